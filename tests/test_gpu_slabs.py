"""z-slab decomposition across GPUs (BASELINE config 5 semantics at test size):
needs >= 2 B200s (`gpurun --gpus 2`); one process per GPU under torchrun, in-kernel halo and
scalar exchange over peer memory.  Checks live in tests/slab_worker.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_slabs_two_ranks_match_oracle_and_single_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29611",
                        os.path.join(ROOT, "tests", "slab_worker.py")], capture_output=True, text=True, timeout=600)
    print(r.stdout[-4000:], r.stderr[-2000:])
    assert r.returncode == 0 and "SLAB_OK" in r.stdout
