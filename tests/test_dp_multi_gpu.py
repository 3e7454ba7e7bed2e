"""Data-parallel parity over NCCL on >= 2 GPUs of one box (skipped on a single-GPU box): each rank trains
on its shard through the facade; losses and weights must match the single-process oracle."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode", ["fp32", "tc", "tc-nccl", "tc-fit", "fp32-fit", "tc-graph", "tc-peer", "fp32-peer", "fp32-pavg3", "lstm-fp32", "lstm-tc"])
def test_dp_two_ranks_match_oracle(mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "dp_check.py"), mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "OK" in r.stdout


def _final_loss(out):
    return float([l for l in out.splitlines() if l.startswith("Testing set loss:")][-1].split(":")[1])


def test_script_two_gpus_matches_one(tmp_path):
    """The drop-in script with --gpus 2 (it re-launches itself as two ranks; model.fit then runs the library's push/poll
    step) ends with the same test loss as the single-GPU run: same global batches, sums added in a different order."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = os.path.join(ROOT, "deep-learning-at-scale_b200", "scripts", "static_ann.py")
    names = ",".join("f%d" % i for i in range(32)) + ",q"
    types = ",".join(["float"] * 33)
    res = []
    for g in (1, 2):
        store = str(tmp_path / ("g%d" % g)) + "/"
        r = subprocess.run([sys.executable, script, names, types, "q", store, "--synthetic", "65536", "--hidden", "256,256",
                            "--batch-size", "4096", "--epochs", "3", "--loss", "mse", "--mode", "tc", "--gpus", str(g)],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert os.path.exists(os.path.join(store, "models", "static_ann.mdl"))
        res.append(_final_loss(r.stdout))
    assert res[1] == pytest.approx(res[0], rel=2e-3), res
