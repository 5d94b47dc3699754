"""Helper of tests/test_gpu_graph.py: a fixed scenario (steps, box change, charge change, more steps) whose end state is printed as
a hash.  Run once as is (captured CUDA graph replayed for the plain force evaluations) and once with B200_NO_GRAPH=1 (every
kernel launched separately): the integer force sums make the two trajectories identical bit for bit."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openmm_b200 as b
from openmm_b200 import workloads

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
sysd = workloads.water_box(n, seed=n)
ctx = workloads.make_context(sysd)
ctx.set_positions(sysd["positions"]); ctx.set_velocities(sysd["velocities"])
ctx.set_integrator(dt=0.002, seed=11)
l0 = ctx.launch_count()
ctx.step(150)
launches_150 = ctx.launch_count() - l0
h = hashlib.sha256(ctx.get_positions().tobytes())
# a box change (barostat-like) and a parameter change must drop the captured evaluation
box = np.array(sysd["box"], dtype=np.float64) * 1.002
ctx.set_box(box)
ctx.step(40)
h.update(ctx.get_positions().tobytes())
# barostat-like: a box change every 12 steps (each one re-records the evaluation and updates the executable graph in place)
for k in range(6):
    ctx.set_box(box * (1.0 + 0.0005 * (k % 3)))
    ctx.step(12)
h.update(ctx.get_positions().tobytes())
# ... and a change on every third step (short-lived captures: the back-off to launch-by-launch must not change the result either)
for k in range(12):
    ctx.set_box(box * (1.0 + 0.0002 * (k % 2)))
    ctx.step(3)
h.update(ctx.get_positions().tobytes())
q = np.array(sysd["charge"], dtype=np.float64) * 0.98
ctx.update_nonbonded_parameters(q, sysd["sigma"], sysd["epsilon"], len(sysd["exclusions"]))
ctx.step(40)
h.update(ctx.get_positions().tobytes())
e = ctx.compute()
h.update(ctx.get_forces().tobytes())
print(json.dumps({"hash": h.hexdigest(), "energy": e, "launches_150": launches_150}))
ctx.close()
