"""-m gpu, needs >= 2 GPUs on the box (skipped otherwise): the NCCL side of configs[4] — tools/fullmap_check.py under
torchrun with 2 ranks asserts that the gathered mosaic is bit-identical to the 1-GPU mosaic and to oracle.stitch_tiles."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
def test_fullmap_gather_bit_identical_2_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tools", "fullmap_check.py"), "4", "30", "8"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    print(r.stdout[-2000:], r.stderr[-2000:])
    assert r.returncode == 0 and "bit-identical to the 1-GPU mosaic: True" in r.stdout
