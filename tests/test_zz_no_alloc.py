"""Allocation-free FFT entry point (b200_fft3d_c2r_noalloc).

Who needs it: a caller that cannot allocate device memory at the point where it wants the transform, for instance because a
memory pool already owns the device or because its loop must never wait on the allocator.  The entry point keeps its
working storage in a static device array, so this test checks (through the C ABI) that it gives the same numbers as the
ordinary entry point b200_fft3d_c2r_host, twice in a row (the second call reuses the scratch the first one dirtied), and
that it refuses grids larger than its built-in storage.  The file name sorts last so that a failure here cannot hide
another test's result in a run that stops at the first failure."""
import numpy as np
import pytest

import openmm_b200 as b

pytestmark = pytest.mark.gpu


def test_noalloc_matches_host_entry_point_twice():
    import torch
    rng = np.random.default_rng(11)
    side = torch.cuda.Stream()      # the last shape runs on a stream of the caller's instead of the default one
    # (25, 25, 27): odd nz, so the generic shared-memory passes run instead of the two-stage register passes
    for call, shape in enumerate([(28, 25, 30), (56, 56, 56), (25, 25, 27), (90, 90, 90)]):
        nx, ny, nz = shape
        real = rng.normal(size=shape).astype(np.float32)
        spec = np.fft.rfftn(real.astype(np.float64)).astype(np.complex64)
        ref = b.fft3d_c2r(spec, nz)                                   # existing entry point (allocates internally)
        d_in = torch.from_numpy(spec.view(np.float32).copy()).cuda()
        d_out = torch.empty(nx * ny * nz, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        for rep in range(2):                                          # two calls in a row share the static scratch
            d_out.zero_()
            torch.cuda.synchronize()
            b.fft3d_c2r_noalloc(d_in.data_ptr(), d_out.data_ptr(), nx, ny, nz, stream=side.cuda_stream if call == 3 else 0)
            torch.cuda.synchronize()
            got = d_out.cpu().numpy().reshape(shape)
            scale = np.abs(ref).max()
            assert np.abs(got - ref).max() / scale < 5e-6, (call, rep)
            # and against the exact answer: the unnormalised inverse of the forward transform is the input times the grid size
            assert np.abs(got / (nx * ny * nz) - real).max() < 5e-6, (call, rep)
        # the input spectrum is left untouched
        assert np.array_equal(d_in.cpu().numpy().ravel(), spec.view(np.float32).ravel())


def test_noalloc_rejects_oversized_grids():
    import torch
    d = torch.zeros(16, dtype=torch.float32, device="cuda")
    with pytest.raises(b.B200Error, match="exceeds the built-in"):
        b.fft3d_c2r_noalloc(d.data_ptr(), d.data_ptr(), 196, 196, 196)
    with pytest.raises(b.B200Error, match="limit of 1024"):
        b.fft3d_c2r_noalloc(d.data_ptr(), d.data_ptr(), 2048, 2, 2)
    with pytest.raises(b.B200Error, match="factor into 2, 3, 5 and 7"):
        b.fft3d_c2r_noalloc(d.data_ptr(), d.data_ptr(), 22, 8, 8)
