import sys, os, tempfile, subprocess, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openmm_b200 as b
from openmm_b200 import workloads
from oracle_lib import REF_DRIVER, rel_rms
for n in (1536, 6000):
    sysd = workloads.water_box(n, seed=n)
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    ctx = workloads.make_context(sysd); ctx.set_positions(sysd["positions"])
    ed = ctx.compute(b.GROUP_DIRECT); fd = ctx.get_forces().copy()
    er = ctx.compute(b.GROUP_RECIPROCAL); fr = ctx.get_forces().copy()
    ctx.close()
    with tempfile.TemporaryDirectory() as td:
        sp, op = os.path.join(td, "s.bin"), os.path.join(td, "o.bin")
        workloads.write_system_file(sp, sysd, 0.9, alpha, grid, dispersion=True)
        out = subprocess.run([REF_DRIVER, "eval", sp, op, "--platform", "B200", "--plugin", b.PLUGIN_PATH], capture_output=True, text=True)
        print(out.stdout[-300:], out.stderr[-300:])
        r = workloads.read_eval_file(op, n)
    print(n, "plugin vs C ABI: direct", rel_rms(r["f_direct"], fd), "recip", rel_rms(r["f_recip"], fr), "E", r["e_direct"], ed, r["e_recip"], er)
    print("  first atoms plugin direct", r["f_direct"][:2], "cabi", fd[:2]); print("  plugin recip", r["f_recip"][:2], "cabi", fr[:2])
