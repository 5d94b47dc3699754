"""Multi-GPU step (peer-memory reduce + sharded integration + position publication) against a one-rank run.  Needs at least two
GPUs on the box; launches tests/mgpu_check.py under torchrun, one process per GPU, exactly as bench.py is launched."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("world,push", [(2, None), (2, "1"), (4, None), (4, "1")])
def test_multi_gpu_step_is_bit_identical_to_one_rank(world, push):
    # push: B200_RECIP_PUSH (reciprocal forces stored straight into the owners' landing buffers); opt-in
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29540 + world), os.path.join(ROOT, "tests", "mgpu_check.py")]
    env = dict(os.environ)
    env.pop("B200_RECIP_PUSH", None)
    if push is not None:
        env["B200_RECIP_PUSH"] = push
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    line = [l for l in out.stdout.splitlines() if l.startswith("mgpu_check")]
    assert out.returncode == 0 and line, out.stdout[-2000:]
    assert "forces bit-identical to 1 rank: True" in line[0] and "energy identical: True" in line[0]
    assert "100-step trajectory identical: True" in line[0] and "ranks agree: True" in line[0]
