"""GPU (needs >= 2 GPUs on one NVSwitch box, else skipped): data-parallel discriminator update over NVLS multicast memory
(b2g_adam_nvls + SNDCGANEngine._setup_nvls) — tools/nvls_check.py under torchrun with 2 ranks: the kernel agrees bit for
bit with all_reduce + b2g_adam, the engine's NVLS steps (eager and graph replay) track the NCCL path and the replicas stay
bit-identical."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_nvls_data_parallel_update_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29933", os.path.join(ROOT, "tools", "nvls_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "nvls_check ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
