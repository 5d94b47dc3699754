import sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openmm_b200 as b
from openmm_b200 import workloads
from oracle_lib import Oracle, rel_rms
from scipy.spatial import cKDTree

def build(nexc, periodic_safe):
    sysd = workloads.water_box(3000, seed=9)
    L = sysd["box"][0]
    pos = np.mod(sysd["positions"], L)
    pairs = cKDTree(pos, boxsize=L).query_pairs(0.3, output_type="ndarray")
    pairs = pairs[(pairs[:, 0] // 3 != pairs[:, 1] // 3) & (pairs[:, 0] % 3 == 0) & (pairs[:, 1] % 3 != 0)]
    if periodic_safe:
        d = np.linalg.norm(sysd["positions"][pairs[:, 0]] - sysd["positions"][pairs[:, 1]], axis=1)
        pairs = pairs[d < 0.31]
    used, chosen = set(), []
    for i, j in pairs:
        if i // 3 not in used and j // 3 not in used:
            used.add(i // 3); used.add(j // 3); chosen.append((i, j))
        if len(chosen) == nexc:
            break
    chosen = np.array(chosen, dtype=np.int32).reshape(-1, 2)
    sysd["exclusions"] = np.vstack([sysd["exclusions"], chosen]).astype(np.int32)
    return sysd, chosen

for nexc, safe, params in ((300, False, True), (300, True, True), (300, True, False), (5, True, True)):
    sysd, chosen = build(nexc, safe)
    if params:
        qq = 0.5 * sysd["charge"][chosen[:, 0]] * sysd["charge"][chosen[:, 1]]
        sysd["exceptions"] = (chosen, qq, np.full(len(chosen), 0.2), np.full(len(chosen), 0.3))
    ctx = workloads.make_context(sysd)
    ctx.set_positions(sysd["positions"])
    o = Oracle(sysd, 0.9, *ctx.pme)
    out = []
    for g in (1, 2):
        e = ctx.compute(g); f = ctx.get_forces()
        fo, eo = o.nonbonded(g)
        bad = np.argsort(-np.linalg.norm(f - fo, axis=1))[:4]
        out.append((g, rel_rms(f, fo), e, eo, [(int(i), float(np.linalg.norm((f - fo)[i]))) for i in bad]))
    print("nexc", len(chosen), "safe", safe, "params", params, out, flush=True)
    print("   chosen atoms of first pairs", chosen[:4].tolist())
    ctx.close()
