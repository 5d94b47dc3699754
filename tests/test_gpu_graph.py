"""the plain force evaluation of an MD step is replayed from a captured CUDA graph (csrc/b200_context.cu,
computeImpl).  The replayed trajectory must equal, bit for bit, the one produced by launching every kernel separately, also
across the events that invalidate the capture (box change, parameter change)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _run(n, no_graph, cond=False):
    env = dict(os.environ)
    env.pop("B200_NO_GRAPH", None); env.pop("B200_GRAPH_COND", None)
    if no_graph:
        env["B200_NO_GRAPH"] = "1"
    if cond:
        env["B200_GRAPH_COND"] = "1"
    out = subprocess.run([sys.executable, os.path.join(HERE, "graph_check.py"), str(n)], capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


@pytest.mark.parametrize("n", [3000, 23558])
def test_graph_replay_is_bit_identical_to_separate_launches(n):
    g, s = _run(n, False), _run(n, True)
    print(g, s)
    assert g["hash"] == s["hash"] and g["energy"] == s["energy"]
    # both count the same kernels; the replayed evaluation is one submission instead of ~15
    assert g["launches_150"] == s["launches_150"]


def test_rebuild_kernels_in_a_conditional_node():
    # B200_GRAPH_COND=1: the three list-rebuild kernels sit in an IF node of the captured evaluation whose condition the first list
    # kernel sets on the device; same trajectory, fewer kernels launched on the steps that keep their list
    g, s = _run(3000, False, cond=True), _run(3000, True)
    print(g, s)
    assert g["hash"] == s["hash"] and g["energy"] == s["energy"]
    assert g["launches_150"] < s["launches_150"]
