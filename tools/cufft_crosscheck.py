"""Cross-check row for profiles/: cuFFT (through torch.fft, which calls cuFFT) on the PME meshes of the BASELINE.json configurations,
R2C + C2R per step as CudaFFT3D::execFFT does (platforms/cuda/src/CudaFFT3D.cpp:66-101), against this repository's hand-written
transform (b200_fft3d_r2c_host / c2r_host time host copies too, so the library side is timed inside a PME step by bench.py: `fft`)."""
import json
import torch

out = {}
for n in (56, 90, 144, 196):
    x = torch.randn(n, n, n, device="cuda", dtype=torch.float32)
    for _ in range(3):
        y = torch.fft.rfftn(x); z = torch.fft.irfftn(y, s=(n, n, n))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 50
    e0.record()
    for _ in range(reps):
        y = torch.fft.rfftn(x); z = torch.fft.irfftn(y, s=(n, n, n))
    e1.record(); torch.cuda.synchronize()
    out[f"{n}^3"] = round(e0.elapsed_time(e1) / reps * 1e3, 1)
print(json.dumps({"cufft_r2c_plus_c2r_us": out, "note": "torch.fft.rfftn + irfftn (cuFFT), out of place, includes torch's normalisation pass on the inverse"}))
