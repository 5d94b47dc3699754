"""Host-side mirror of how the direct-space work is partitioned over ranks (one process per GPU).

  rows (32-atom blocks X of the tile list)  : dealt round-robin, rank r owns X with X % R == r            (b200_nblist.cu nbBuild)
  exclusion tiles (static list, sorted)     : contiguous index ranges [n*r/R, n*(r+1)/R)                   (b200_tile.cu)
  PME, exclusion corrections, bonded terms  : rank 0
  forces                                    : int64 fixed point (x 2^32), summed with one all-reduce       (b200_comm.cu)

The reference partitions contiguous tile ranges and rebalances them dynamically (CudaNonbondedUtilities::setAtomBlockRange,
CudaNonbondedUtilities.cpp:485-493; CudaParallelKernels.cpp:162-166,247-266); round-robin rows need no balancer because every
rank gets a statistically identical sample of the Hilbert-ordered rows.
"""
import numpy as np

FORCE_SCALE = 4294967296.0


def rows_of_rank(num_blocks, rank, nranks):
    return np.arange(rank, num_blocks, nranks)


def exclusion_tile_range(num_tiles, rank, nranks):
    return (num_tiles * rank) // nranks, (num_tiles * (rank + 1)) // nranks


def to_fixed(f):
    """(long long)(f * 2^32): truncation toward zero, platforms/cuda/src/kernels/common.cu:29-31"""
    return np.trunc(np.asarray(f, dtype=np.float64) * FORCE_SCALE).astype(np.int64)


def from_fixed(i):
    return np.asarray(i, dtype=np.float64) / FORCE_SCALE
