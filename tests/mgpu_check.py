"""Multi-GPU parity check, launched with torchrun (one process per GPU, NCCL): forces after the int64 all-reduce must be
bit-identical on every rank and bit-identical to a single-rank evaluation; a short trajectory must stay identical too.
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tests/mgpu_check.py"""
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
sysd = workloads.water_box(23558, seed=23558)


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


single = make(False)
e1 = single.compute()
f1 = single.get_forces().copy()
multi = make(True)
e2 = multi.compute()
f2 = multi.get_forces().copy()
ok_force = np.array_equal(f1, f2)
ok_energy = e1 == e2
# forces-only evaluation (with B200_RECIP_PUSH=1, or four ranks and more, rank 0 pushes the reciprocal forces into the owners'
# landing buffers) read back through the all-reduce path
single.compute(energy=False); multi.compute(energy=False)
ok_force = ok_force and np.array_equal(single.get_forces(), multi.get_forces())
# two b200_step calls with a state read in between (the second call starts on an idle GPU with hosts out of step)
single.step(60); multi.step(60)
p1, p2 = single.get_positions(), multi.get_positions()
ok_traj = np.array_equal(p1, p2)
single.step(40); multi.step(40)
p1, p2 = single.get_positions(), multi.get_positions()
ok_traj = ok_traj and np.array_equal(p1, p2)
flags = torch.tensor([int(ok_force), int(ok_energy), int(ok_traj)], device="cuda")
dist.all_reduce(flags, op=dist.ReduceOp.MIN)
# every rank must hold the same state
h = torch.tensor([float(np.abs(p2).sum())], device="cuda", dtype=torch.float64)
hmax, hmin = h.clone(), h.clone()
dist.all_reduce(hmax, op=dist.ReduceOp.MAX); dist.all_reduce(hmin, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"mgpu_check world={world}: forces bit-identical to 1 rank: {bool(flags[0])}, energy identical: {bool(flags[1])}, "
          f"100-step trajectory identical: {bool(flags[2])}, ranks agree: {hmax.item() == hmin.item()}, e={e2:.6f}")
    assert flags.min().item() == 1 and hmax.item() == hmin.item()
single.close(); multi.close()
dist.destroy_process_group()
