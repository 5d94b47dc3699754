"""World-size-2 test of the multi-rank host logic on CPU (gloo): the row partition covers every block exactly once, and the
fixed-point force reduction is exact, so the all-reduced result is bit-identical to the single-rank one whatever the split."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from openmm_b200 import partition, workloads
from oracle_lib import Oracle


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _pair_forces(sysd, pairs, alpha):
    """direct-space pair forces (same formula as the oracle) for an explicit pair list, numpy"""
    from scipy.special import erfc
    pos, q, sig, eps, L = sysd["positions"], sysd["charge"], sysd["sigma"], sysd["epsilon"], sysd["box"][0]
    i, j = pairs[:, 0], pairs[:, 1]
    d = pos[j] - pos[i]
    d -= L * np.round(d / L)
    r = np.linalg.norm(d, axis=1)
    qq = 138.93545764438198 * q[i] * q[j]
    ar = alpha * r
    s6 = ((0.5 * (sig[i] + sig[j])) / r) ** 6
    dEdR = qq / r ** 3 * (erfc(ar) + 2 * ar * np.exp(-ar * ar) / np.sqrt(np.pi)) + 4 * np.sqrt(eps[i] * eps[j]) * (12 * s6 - 6) * s6 / r ** 2
    return i, j, d * dEdR[:, None]


def _worker(rank, world, port, result_file):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sysd = workloads.water_box(1536, seed=1536)
    n = sysd["n"]
    import openmm_b200 as b
    alpha, grid = b.pme_parameters(0.9, sysd["box"])
    pairs = Oracle(sysd, 0.9, alpha, grid).pairs()
    nblocks = (n + 31) // 32
    mine = np.zeros(nblocks, dtype=bool)
    mine[partition.rows_of_rank(nblocks, rank, world)] = True
    # a pair belongs to the row of its lower-indexed block, as in the tile list (Y >= X)
    row = np.minimum(pairs[:, 0], pairs[:, 1]) // 32
    sel = pairs[mine[row]]
    i, j, f = _pair_forces(sysd, sel, alpha)
    acc = np.zeros((n, 3), dtype=np.int64)
    np.add.at(acc, i, partition.to_fixed(-f))
    np.add.at(acc, j, partition.to_fixed(f))
    t = torch.from_numpy(acc)
    dist.all_reduce(t)                                   # the int64 sum NCCL performs on the GPUs
    cover = torch.from_numpy(mine.astype(np.int64))
    dist.all_reduce(cover)
    lo, hi = partition.exclusion_tile_range(101, rank, world)
    span = torch.tensor([hi - lo])
    dist.all_reduce(span)
    if rank == 0:
        i, j, f = _pair_forces(sysd, pairs, alpha)
        ref = np.zeros((n, 3), dtype=np.int64)
        np.add.at(ref, i, partition.to_fixed(-f))
        np.add.at(ref, j, partition.to_fixed(f))
        np.savez(result_file, same=np.array_equal(ref, t.numpy()), cover=bool((cover.numpy() == 1).all()), span=int(span.item()))
    dist.destroy_process_group()


def test_two_rank_partition_and_integer_reduction(tmp_path):
    out = str(tmp_path / "r.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    assert bool(r["cover"]), "every block row must belong to exactly one rank"
    assert int(r["span"]) == 101, "exclusion tiles must be split without gaps or overlap"
    assert bool(r["same"]), "int64 all-reduce over ranks must equal the single-rank fixed-point sum bit for bit"


def _shard_worker(rank, world, port, result_file):
    """CPU emulation of the peer-memory step: partial int64 forces per rank -> reduce-scatter by integration-unit slice ->
    slice-local position update -> all-gather; must equal the one-rank update bit for bit."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    nmol, extra = 1000, 2                        # 3-site molecules (one unit each) followed by free atoms (one unit each)
    n, units = 3 * nmol + extra, nmol + extra
    pos = rng.random((n, 3))
    partial = [rng.integers(-2 ** 40, 2 ** 40, size=(n, 3), dtype=np.int64) for _ in range(world)]   # every rank's contribution
    total = sum(partial)
    begin, end = partition.unit_shard(units, rank, world)
    atoms = np.concatenate([np.arange(3 * min(begin, nmol), 3 * min(end, nmol)), 3 * nmol + np.arange(max(begin, nmol) - nmol, max(end, nmol) - nmol)]).astype(np.int64)
    gathered = [torch.zeros((n, 3), dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(partial[rank]))      # stands for the NVLink loads out of the peers' force arrays
    mine = sum(g.numpy()[atoms] for g in gathered)                  # exact integer sum for this rank's atoms only
    new = pos.copy()
    new[atoms] += partition.from_fixed(mine) * 1e-6                 # "integration" of the slice
    pub = [torch.zeros((n, 3), dtype=torch.float64) for _ in range(world)]
    dist.all_gather(pub, torch.from_numpy(new))
    owner = np.full(n, -1)
    for r in range(world):
        b, e = partition.unit_shard(units, r, world)
        a = np.concatenate([np.arange(3 * min(b, nmol), 3 * min(e, nmol)), 3 * nmol + np.arange(max(b, nmol) - nmol, max(e, nmol) - nmol)]).astype(np.int64)
        assert (owner[a] == -1).all()
        owner[a] = r
    merged = np.stack([pub[owner[i]].numpy()[i] for i in range(n)])
    if rank == 0:
        ref = pos + partition.from_fixed(total) * 1e-6
        lo, hi = zip(*[partition.excluded_pair_slice(3000, r, world) for r in range(world)])
        np.savez(result_file, same=np.array_equal(ref, merged), cover=bool((owner >= 0).all()), aligned=all(b % 128 == 0 for b in lo),
                 contiguous=all(hi[r] == lo[r + 1] for r in range(world - 1)) and lo[0] == 0 and hi[-1] == 3000)
    dist.destroy_process_group()


def test_two_rank_sharded_integration_matches_one_rank(tmp_path):
    out = str(tmp_path / "s.npz")
    mp.spawn(_shard_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    assert bool(r["cover"]) and bool(r["same"]), "slices must cover every atom once and reproduce the one-rank update exactly"
    assert bool(r["aligned"]) and bool(r["contiguous"]), "excluded-pair slices must be block aligned and gap free"
