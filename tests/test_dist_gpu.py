"""Multi-GPU (>= 2 B200s on one node) checks, launched through torchrun; skipped on a single-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script_args, n, timeout=240):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29531"] + script_args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_data_parallel_training_step():
    r = _torchrun([os.path.join(ROOT, "tests", "dist_train_check.py")], 2)
    assert r.returncode == 0 and "DIST_TRAIN_OK" in r.stdout and "DIST_TEARDOWN_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_data_parallel_pose_training_step():
    r = _torchrun([os.path.join(ROOT, "tests", "dist_posetrain_check.py")], 2)
    assert r.returncode == 0 and "DIST_POSETRAIN_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_bench_two_gpus_prints_one_json_line():
    import json
    r = _torchrun([os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "3", "--warmup", "3", "--batch", "64", "--sustain-s", "0.2",
                   "--hires-batch", "16"], 2, timeout=600)
    assert r.returncode == 0, r.stderr[-4000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["n_gpus"] == 2 and j["value"] > 0 and j["config"]["global_batch"] == 128
    # configs[2] / configs[3] ride in the same line: DP training step with a real all-reduce, and the 384x288 variant
    assert j["train"]["value"] > 0 and j["train"]["allreduce_bytes_per_step"] > 100e6
    assert j["hires"]["value"] > 0 and j["e2e"]["h2d_bytes_per_step"] == 64 * 3 * 256 * 192 + 64 * 24
