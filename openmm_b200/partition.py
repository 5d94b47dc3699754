"""Host-side mirror of how the direct-space work is partitioned over ranks (one process per GPU).

  rows (32-atom blocks X of the tile list)  : weighted round-robin, rank = owner_table[X & 63]             (b200_nblist.cu nbBuild)
  exclusion tiles (static list, sorted)     : contiguous index ranges proportional to the same shares      (b200_tile.cu)
  PME, exclusion corrections, bonded terms  : rank 0
  forces                                    : int64 fixed point (x 2^32); stepping: reduce-scatter over peer memory, sharded
                                              integration, position all-gather; getState: one all-reduce  (b200_comm.cu)

The reference partitions contiguous tile ranges and rebalances them dynamically (CudaNonbondedUtilities::setAtomBlockRange,
CudaNonbondedUtilities.cpp:485-493; CudaParallelKernels.cpp:162-166,247-266); round-robin rows need no balancer because every
rank gets a statistically identical sample of the Hilbert-ordered rows.
"""
import numpy as np

FORCE_SCALE = 4294967296.0


def shares(nranks, pme_cost_ratio=0.38, with_communicator=True):
    """fraction of the direct-space rows each rank takes; rank 0 also runs PME (b200_comm.cu b200_set_partition)"""
    if nranks == 1 or not with_communicator:
        return [1.0 / nranks] * nranks
    s0 = max(0.0, (1.0 - (nranks - 1) * pme_cost_ratio) / nranks)
    return [s0] + [(1.0 - s0) / (nranks - 1)] * (nranks - 1)


def owner_table(nranks, **kw):
    """row X belongs to rank owner_table[X & 63]: weighted round-robin, same largest-credit rule as the C code"""
    share = shares(nranks, **kw)
    credit = [0.0] * nranks
    table = []
    for _ in range(64):
        best = 0
        for r in range(nranks):
            credit[r] += share[r]
            if credit[r] > credit[best] + 1e-12:
                best = r
        table.append(best)
        credit[best] -= 1.0
    return np.array(table)


def rows_of_rank(num_blocks, rank, nranks, **kw):
    table = owner_table(nranks, **kw)
    x = np.arange(num_blocks)
    return x[table[x & 63] == rank]


def exclusion_tile_range(num_tiles, rank, nranks, **kw):
    share = shares(nranks, **kw)
    begin = sum(share[:rank])
    end = 1.0 if rank == nranks - 1 else begin + share[rank]
    return int(round(num_tiles * begin)), int(round(num_tiles * end))


def unit_shard(num_units, rank, nranks, block=128):
    """slice [begin, end) of the integration units (SETTLE clusters, then free atoms) that rank integrates in the peer-memory step:
    cut at multiples of the 128-unit thread block so that block-level partial sums group the same units on any rank count
    (b200_comm.cu b200_peer_plan_shard)"""
    nblk = (num_units + block - 1) // block
    begin = min(num_units, block * ((nblk * rank) // nranks))
    end = min(num_units, block * ((nblk * (rank + 1)) // nranks))
    return begin, end


def excluded_pair_slice(num_pairs, rank, nranks, block=128, **kw):
    """slice of the excluded-pair corrections a rank evaluates: same shares as the rows, aligned to the 128-pair blocks
    (b200_bonded.cu b200_bonded_launch)"""
    share = shares(nranks, **kw)
    begin = sum(share[:rank])
    end = 1.0 if rank == nranks - 1 else begin + share[rank]
    nblk = (num_pairs + block - 1) // block
    return block * int(round(nblk * begin)), min(num_pairs, block * int(round(nblk * end)))


def to_fixed(f):
    """(long long)(f * 2^32): truncation toward zero, platforms/cuda/src/kernels/common.cu:29-31"""
    return np.trunc(np.asarray(f, dtype=np.float64) * FORCE_SCALE).astype(np.int64)


def from_fixed(i):
    return np.asarray(i, dtype=np.float64) / FORCE_SCALE
