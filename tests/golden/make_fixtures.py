"""Regenerates the fixtures under tests/golden/ from the reference tree.  Runs only in the build container (needs
/root/reference and oracle/_ref built by `make -C oracle reference`); the fixtures are committed, this script documents
exactly how they were made.

  tip3p_895.npy        float32 [895,3,3] nm: wrappers/python/openmm/app/data/tip3p.pdb (equilibrated 30 A TIP3P box)
  nacl_amorph.npy      float32 [894,3] nm: tests/nacl_amorph.dat (testEwaldPME, tests/TestEwald.h:96-218)
  ref_water_<N>.npz    outputs of the reference's own Reference platform (double precision) for water_box(N):
                       positions (float32), energies and forces of the direct and reciprocal force groups, and the
                       hash + count of the pair set of computeNeighborListVoxelHash.  N = 1536 (small box: per-pair
                       minimum image), 6000 (single periodic copy), 23558 (DHFR size; forces stored as float32).
  ref_ljpme_<N>.npz    the same for NonbondedForce::LJPME (method 5): water_box(N) with CHARMM-style hydrogens (sigma 0.04 nm,
                       eps 0.192464 kJ/mol) so that three particle types enter the dispersion grid; dispersion mesh and alpha
                       as NonbondedForceImpl::calcPMEParameters(lj=true).  N = 1538, 6002.
  water125_dpme.npy    float64 [375,3] Angstrom: the 125-water coordinates of tests/TestDispersionPME.h:406-783 (make_waterbox),
                       used by the LJPME known-answer tests testWater125DpmeVsLongCutoff{No,With}Exclusions
  ref_switch_1536.npz  water_box(1536), PME, Lennard-Jones switching function from 0.75 nm to the 0.9 nm cutoff (Reference platform)
  ref_cutoff_methods_768.npz  water_box(768) evaluated by the Reference platform with NoCutoff, CutoffNonPeriodic and CutoffPeriodic
                       (reaction field, dielectric 78.3): energies and forces per method (oracle groundwork for the next scope rows)
  ref_ewald_768.npz    water_box(768) with NonbondedForce::Ewald (explicit k-space sum, kmax from calcEwaldParameters), Reference platform
  ref_ions_triclinic.npz  the 8-ion triclinic case of tests/TestEwald.h:220-269 evaluated by the Reference platform
  ref_large_<N>.npz / ref_large_ljpme_<N>.npz  the BASELINE.json configurations themselves (92 224, 408 609, 1 066 628 atoms; LJPME at
                       92 224 with a 45^3 dispersion mesh) evaluated once by the Reference platform, in compact form: energies, pair
                       count + hash, sha256 of the float32 positions, and the direct / reciprocal forces of every 16th atom (float32)
"""
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference"


def tip3p():
    pos = []
    for line in open(os.path.join(REF, "wrappers/python/openmm/app/data/tip3p.pdb")):
        if line.startswith("ATOM"):
            pos.append([float(line[30:38]), float(line[38:46]), float(line[46:54])])
    a = 0.1 * np.array(pos).reshape(-1, 3, 3)
    # make every molecule whole (hydrogens next to their oxygen) in the 3 nm cell
    for k in (1, 2):
        d = a[:, k, :] - a[:, 0, :]
        a[:, k, :] -= 3.0 * np.round(d / 3.0)
    np.save(os.path.join(HERE, "tip3p_895.npy"), a.astype(np.float32))


def nacl():
    pat = re.compile(r"Vec3\(([-\d.eE+]+),([-\d.eE+]+),([-\d.eE+]+)\)")
    pos = [[float(x) for x in pat.search(l).groups()] for l in open(os.path.join(REF, "tests/nacl_amorph.dat")) if "Vec3" in l]
    np.save(os.path.join(HERE, "nacl_amorph.npy"), np.array(pos, dtype=np.float32))


def ref_eval(sysd, cutoff, alpha, grid, pairs=True, pair_hash_only=False, **kw):
    from openmm_b200 import workloads
    from oracle_lib import REF_DRIVER
    with tempfile.TemporaryDirectory() as td:
        sp, op = os.path.join(td, "s.bin"), os.path.join(td, "o.bin")
        workloads.write_system_file(sp, sysd, cutoff, alpha, grid, **kw)
        subprocess.check_call([REF_DRIVER, "eval", sp, op, "--platform", "Reference"] + ((["--pair-hash"] if pair_hash_only else ["--pairs"]) if pairs else []))
        return workloads.read_eval_file(op, sysd["n"])


LARGE_STRIDE = 16


def large(n, ljpme=False):
    """One Reference-platform evaluation of a BASELINE.json configuration, stored compactly (SURVEY.md 8c gotcha 6)."""
    import hashlib
    import time
    import openmm_b200 as b
    from openmm_b200 import workloads
    t0 = time.time()
    sysd = workloads.ljpme_water_box(n) if ljpme else workloads.water_box(n, seed=n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    kw = {}
    extra = {}
    if ljpme:
        lj_alpha, lj_grid = b.pme_parameters(0.9, sysd["box"], lj=True)
        kw = dict(method=5, lj_alpha=lj_alpha, lj_grid=lj_grid)
        extra = dict(lj_alpha=lj_alpha, lj_grid=np.array(lj_grid))
    print(n, "built in", time.time() - t0, "s; grid", grid, extra, flush=True)
    r = ref_eval(sysd, 0.9, alpha, grid, pairs=True, pair_hash_only=True, **kw)
    sha = hashlib.sha256(sysd["positions"].astype(np.float32).tobytes()).hexdigest()
    name = f"ref_large_ljpme_{n}.npz" if ljpme else f"ref_large_{n}.npz"
    np.savez_compressed(os.path.join(HERE, name), n=n, box=sysd["box"], alpha=alpha, grid=np.array(grid), cutoff=0.9, stride=LARGE_STRIDE,
                        positions_sha256=sha, e_direct=r["e_direct"], e_recip=r["e_recip"],
                        f_direct=r["f_direct"][::LARGE_STRIDE].astype(np.float32), f_recip=r["f_recip"][::LARGE_STRIDE].astype(np.float32),
                        f_direct_norm2=float(np.sum(r["f_direct"] ** 2)), f_recip_norm2=float(np.sum(r["f_recip"] ** 2)),
                        f_total_norm2=float(np.sum((r["f_direct"] + r["f_recip"]) ** 2)),
                        pair_count=r["pair_count"], pair_hash=np.uint64(r["pair_hash"]), **extra)
    print("large", n, "ljpme" if ljpme else "pme", "pairs", r["pair_count"], "E", r["e_direct"], r["e_recip"], "in", time.time() - t0, "s", flush=True)


def water(n, f32=False):
    import openmm_b200 as b
    from openmm_b200 import workloads
    from oracle_lib import pair_hash
    sysd = workloads.water_box(n, seed=n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    r = ref_eval(sysd, 0.9, alpha, grid)
    cnt, h = pair_hash(r["pairs"], n)
    ft = np.float32 if f32 else np.float64
    np.savez_compressed(os.path.join(HERE, f"ref_water_{n}.npz"), positions=sysd["positions"].astype(np.float32), box=sysd["box"],
                        alpha=alpha, grid=np.array(grid), cutoff=0.9, e_direct=r["e_direct"], e_recip=r["e_recip"],
                        f_direct=r["f_direct"].astype(ft), f_recip=r["f_recip"].astype(ft), pair_count=cnt, pair_hash=np.uint64(h))
    print(n, "pairs", cnt, "E", r["e_direct"], r["e_recip"])


def water_ljpme(n):
    import openmm_b200 as b
    from openmm_b200 import workloads
    sysd = workloads.ljpme_water_box(n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    lj_alpha, lj_grid = b.pme_parameters(0.9, sysd["box"], lj=True)
    r = ref_eval(sysd, 0.9, alpha, grid, pairs=False, method=5, lj_alpha=lj_alpha, lj_grid=lj_grid)
    np.savez_compressed(os.path.join(HERE, f"ref_ljpme_{n}.npz"), positions=sysd["positions"].astype(np.float32), box=sysd["box"],
                        alpha=alpha, grid=np.array(grid), lj_alpha=lj_alpha, lj_grid=np.array(lj_grid), cutoff=0.9,
                        e_direct=r["e_direct"], e_recip=r["e_recip"], f_direct=r["f_direct"], f_recip=r["f_recip"])
    print("ljpme", n, "E", r["e_direct"], r["e_recip"])


def water_switch(n=1536, sw=0.75):
    import openmm_b200 as b
    from openmm_b200 import workloads
    sysd = workloads.water_box(n, seed=n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    r = ref_eval(sysd, 0.9, alpha, grid, pairs=False, switch_distance=sw)
    np.savez_compressed(os.path.join(HERE, f"ref_switch_{n}.npz"), positions=sysd["positions"].astype(np.float32), box=sysd["box"], alpha=alpha,
                        grid=np.array(grid), cutoff=0.9, switch_distance=sw, e_direct=r["e_direct"], e_recip=r["e_recip"], f_direct=r["f_direct"], f_recip=r["f_recip"])
    print("switch", n, "E", r["e_direct"], r["e_recip"])


def cutoff_methods(n=768):
    import openmm_b200 as b
    from openmm_b200 import workloads
    sysd = workloads.water_box(n, seed=n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    out = dict(positions=sysd["positions"].astype(np.float32), box=sysd["box"], cutoff=0.9)
    for method, name in ((0, "nocutoff"), (1, "cutoff_nonperiodic"), (2, "cutoff_periodic")):
        r = ref_eval(sysd, 0.9, alpha, grid, pairs=False, method=method)
        out["e_" + name] = r["e_direct"] + r["e_recip"]; out["f_" + name] = r["f_direct"] + r["f_recip"]
        print(name, out["e_" + name])
    np.savez_compressed(os.path.join(HERE, f"ref_cutoff_methods_{n}.npz"), **out)


def ewald(n=768):
    from openmm_b200 import workloads
    from oracle_lib import ewald_parameters
    sysd = workloads.water_box(n, seed=n)
    alpha, kmax = ewald_parameters(0.9, sysd["box"])
    r = ref_eval(sysd, 0.9, alpha, [8, 8, 8], pairs=False, method=3)      # NonbondedForce::Ewald, default tolerance 5e-4
    np.savez_compressed(os.path.join(HERE, f"ref_ewald_{n}.npz"), positions=sysd["positions"].astype(np.float32), box=sysd["box"], alpha=alpha, kmax=np.array(kmax),
                        cutoff=0.9, e_direct=r["e_direct"], e_recip=r["e_recip"], f_direct=r["f_direct"], f_recip=r["f_recip"])
    print("ewald", n, alpha, kmax, r["e_direct"], r["e_recip"])


def water125():
    lines = open(os.path.join(REF, "tests/TestDispersionPME.h")).read().split("\n")
    start = next(i for i, l in enumerate(lines) if "const double coords[375][3]" in l)
    pat = re.compile(r"\{\s*([-\d.]+),\s*([-\d.]+),\s*([-\d.]+)\s*\}")
    pos = []
    for l in lines[start + 1:]:
        m = pat.search(l)
        if not m:
            break
        pos.append([float(x) for x in m.groups()])
    assert len(pos) == 375
    np.save(os.path.join(HERE, "water125_dpme.npy"), np.array(pos))


def ions():
    pos = np.array([[1.744, 2.788, 3.162], [1.048, 0.762, 2.340], [2.489, 1.570, 2.817], [1.027, 1.893, 3.271],
                    [0.937, 0.825, 0.009], [2.290, 1.887, 3.352], [1.266, 1.111, 2.894], [0.933, 1.862, 3.490]])
    sysd = dict(n=8, positions=pos, box=np.array([2.5, 0, 0, 0.5, 3.0, 0, 0.7, 0.9, 3.5]), charge=np.array([-1.0] * 4 + [1.0] * 4),
                sigma=np.array([0.440104] * 4 + [0.332840] * 4), epsilon=np.array([0.4184] * 4 + [0.0115897] * 4), mass=np.ones(8),
                exclusions=np.zeros((0, 2), dtype=np.int32))
    r = ref_eval(sysd, 1.0, 3.45891, [32, 40, 48], pairs=False)
    np.savez(os.path.join(HERE, "ref_ions_triclinic.npz"), e_direct=r["e_direct"], e_recip=r["e_recip"], f_direct=r["f_direct"], f_recip=r["f_recip"])
    print("ions", r["e_direct"] + r["e_recip"])


if __name__ == "__main__":
    only = sys.argv[1:]
    if not only or "ewald" in only:
        ewald()
    if not only or "cutoff" in only:
        cutoff_methods()
    if not only or "switch" in only:
        water_switch()
    if not only or "water125" in only:
        water125()
    if not only or "base" in only:
        tip3p()
        nacl()
        ions()
        for n, f32 in ((1536, False), (6000, False), (23558, True)):
            water(n, f32)
    if not only or "ljpme" in only:
        for n in (1538, 6002):
            water_ljpme(n)
    for key, n, lj in (("large_apoa1", 92224, False), ("large_apoa1_ljpme", 92224, True), ("large_cellulose", 408609, False), ("large_stmv", 1066628, False)):
        if key in only or "large" in only:      # minutes each (the STMV evaluation alone is several minutes): never part of the default run
            large(n, lj)
