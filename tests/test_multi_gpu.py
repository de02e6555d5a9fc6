"""Multi-GPU parity inside `pytest -m gpu`: when at least two GPUs are visible, tests/dp_check.py (data parallel, replicated
table) and tests/shard_check.py (row-sharded table) are launched as 2 ranks (torchrun, rendezvous on 127.0.0.1, NCCL); every
rank compares its step with the float64 ORACLE on the global batch and with the 1-GPU CUDA step.  Skipped on a 1-GPU box; the
host-side logic of both exchanges is covered on the CPU by tests/test_dist_cpu.py (gloo, world size 2)."""
import os
import subprocess

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("script,port", [("dp_check.py", 29561), ("shard_check.py", 29562)])
def test_two_ranks_against_the_oracle(script, port, tmp_path):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    log = str(tmp_path / "ranks.log")
    out = str(tmp_path / "launcher.log")
    with open(out, "w") as fo:          # files, not pipes: nothing the launcher leaves behind can keep this call waiting
        r = subprocess.run(["bash", os.path.join(HERE, "run_ranks.sh"), "300", "2", str(port), log, os.path.join(HERE, script)],
                           stdout=fo, stderr=subprocess.STDOUT, stdin=subprocess.DEVNULL, timeout=400)
    tail = open(log).read()[-3000:] if os.path.exists(log) else ""
    assert r.returncode == 0, open(out).read()[-2000:] + "\n" + tail
    assert "ok: world=2" in tail
