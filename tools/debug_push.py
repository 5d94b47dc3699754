"""2-GPU debug: forces-only evaluation + get_forces and a chunked trajectory, multi-rank (reciprocal push) against one rank.
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/debug_push.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openmm_b200 as b
from openmm_b200 import workloads

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
wl = os.environ.get("DBG_WORKLOAD")
if wl:
    sysd = workloads.cached_water_box(workloads.WORKLOADS[wl])
else:
    n = int(os.environ.get("DBG_ATOMS", "23558"))
    sysd = workloads.water_box(n, seed=n)


def make(comm):
    ctx = workloads.make_context(sysd, device=local)
    if comm:
        ids = [b.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(rank, world, ids[0])
    ctx.set_positions(sysd["positions"])
    ctx.set_velocities(sysd["velocities"])
    ctx.set_integrator(dt=0.002, seed=7)
    return ctx


single, multi = make(False), make(True)
single.compute(energy=False); multi.compute(energy=False)
if not os.environ.get("DBG_UNCONSUMED"):
    f1, f2 = single.get_forces().copy(), multi.get_forces().copy()
    if rank == 0:
        print("forces-only evaluation identical:", np.array_equal(f1, f2), "max |df|", np.abs(f1 - f2).max(), flush=True)
done = 0
for chunk in [int(x) for x in os.environ.get('DBG_CHUNKS', '1,1,1,1,1,1,1,1,2,2,4,4,8,8,16').split(',')]:
    try:
        single.step(chunk); multi.step(chunk)
    except Exception as e:
        print(f"rank {rank}: exception after {done} steps: {e}", flush=True)
        break
    done += chunk
    p1, p2 = single.get_positions(), multi.get_positions()
    if rank == 0:
        bad = np.nonzero(np.any(p1 != p2, axis=1))[0]
        if len(bad):
            print(f"   differing atoms: {len(bad)} of {len(p1)}, first {bad[:6]}, last {bad[-3:]}, max |dx| at {int(np.abs(p1 - p2).max(axis=1).argmax())}", flush=True)
        print(f"steps {done}: identical {np.array_equal(p1, p2)} max |dx| {np.abs(p1 - p2).max():.3e} nan {int(np.isnan(p2).sum())} stats {multi.nblist_stats()}", flush=True)
single.close(); multi.close()
dist.destroy_process_group()
