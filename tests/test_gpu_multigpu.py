"""Multi-GPU parity (needs >= 2 GPUs, skipped otherwise): the database-sharded 2-NN -- local search with
the CUDA kernel on every rank, NCCL all-gather of the packed top-2 keys, merge kernel -- must equal the
single-GPU search of the whole database (tools/multigpu_check.py under torchrun)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_knn2_sharded_nccl_equals_single_gpu():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else 4
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", os.path.join(ROOT, "tools", "multigpu_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert f"knn2 sharded over {world} GPUs == single GPU: True" in r.stdout
