"""Quick GPU probe for kernel variants: STMV-sized box (or --workload), a few warm-up steps, then per-family CUDA-event times
over K steps.  B200_LIB_PATH selects a variant library, B200_SPARSE_MAX / B200_NB_PADDING the list parameters.
   python tools/tile_probe.py [--workload stmv] [--steps 30] [--tag name]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openmm_b200 as b
from openmm_b200 import workloads

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="stmv")
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--tag", default="")
a = ap.parse_args()
sysd = workloads.cached_water_box(workloads.WORKLOADS[a.workload])
ctx = workloads.make_context(sysd)
ctx.set_positions(sysd["positions"]); ctx.set_velocities(sysd["velocities"])
ctx.set_integrator(dt=0.002, seed=1)
ctx.step(20)
total = ctx.timed_steps(a.steps)
ctx.timing_enable(True); ctx.timing_reset()
ctx.step(a.steps)
fam = ctx.timing_read()
st = ctx.nblist_stats()
pairs, _ = ctx.direct_pairs()
dense = int(st["num_tiles"]) - 2 * int(st["num_sparse_tiles"])
print(json.dumps({"tag": a.tag, "lib": os.path.basename(b.LIB_PATH), "sparse_max": os.environ.get("B200_SPARSE_MAX"), "pad": os.environ.get("B200_NB_PADDING"),
                  "ms_per_step": round(total / a.steps, 4), "us": {k: round(1e3 * v[0] / a.steps, 1) for k, v in fam.items() if v[0] > 0},
                  "tile_us_per_launch": round(1e3 * fam["tile"][0] / max(1, fam["tile"][1]), 1), "dense_tiles": dense, "sparse_tiles": int(st["num_sparse_tiles"]),
                  "units": int(st["num_units"]), "pairs": pairs, "rebuild_every": round(st["steps"] / max(1, st["rebuilds"]), 2)}))
ctx.close()
