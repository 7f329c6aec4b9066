"""GPU tier, multi-GPU: the fused gradient reduce + AdamW + weight broadcast kernel over NVLink.
Needs >= 2 GPUs on the box (skipped on a single-GPU box; the CPU tier covers the host logic)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode", ["ddp", "zero", "ddp+vp"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_dp_matches_single_gpu(world, mode):
    """ddp: push-based weight broadcast; zero: sharded matrices, all-gather fused into the forward GEMMs;
    +vp: vocab-parallel LM head + cross-entropy.  All must reproduce the single-GPU oracle."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, DP_MODE=mode.split("+")[0], DP_VP="1" if "vp" in mode else "0")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29700 + world),
                        os.path.join(ROOT, "tools", "dp_check.py")], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and "DP_CHECK_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("mode", ["ddp", "zero", "ddp+vp"])
def test_trainer_paths_with_data_parallelism(mode):
    """Through the Trainer on 2 GPUs: gradient accumulation with the fused reduce+AdamW kernels (replicated modes),
    save_checkpoint with the state spread over the ranks, resume - against a single-GPU oracle
    (tools/dp_trainer_check.py)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, DP_MODE=mode.split("+")[0], DP_VP="1" if "vp" in mode else "0")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29761",
                        os.path.join(ROOT, "tools", "dp_trainer_check.py")], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and "DP_TRAINER_CHECK_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
