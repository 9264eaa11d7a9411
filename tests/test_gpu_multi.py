"""Multi-GPU identity as a test: tools/multi_gpu_check.py under torchrun on 2 GPUs of this box (skipped with fewer).
It asserts that the sharded run -- caller-side all-reduce, the library's own NCCL merge, the native tree loop, a
shard without any active row, and a random forest (forest passes on every shard) -- gives statistics and trees
bit-identical to one GPU over the concatenated rows."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_gpus_bit_identical_to_one():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one box")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tools", "multi_gpu_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "multi-gpu check ok" in out.stdout and "in-library NCCL merge ok" in out.stdout and "empty-shard check ok" in out.stdout
    assert "forest check ok" in out.stdout
