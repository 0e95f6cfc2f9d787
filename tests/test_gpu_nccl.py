"""The library's native NCCL exchange (glia_b200_build_sharded) on real GPUs: needs >= 2 devices, so it is skipped on the
1-GPU test box; tests/test_gpu_sharded.py covers the same kernels through the thread-emulated torch path there, and
bench.py --gpus N runs the native path at N = 2, 4, 8."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_native_nccl_exchange_equals_single_handle():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (NCCL refuses two ranks on one device)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(min(n, 4)),
                        "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.join(HERE, "nccl_sharded_check.py")],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "native == single: True" in r.stdout and "native == single: False" not in r.stdout
