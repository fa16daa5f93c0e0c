"""Host-side multi-rank logic on CPU (gloo, world_size 2): slab / ensemble partitioning and the
blob all-gather that wires slab handles together.  No CUDA involved."""
import os
import subprocess
import sys
import textwrap

import pytest

import nls_b200
from nls_b200 import slabs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_slab_partition_tiles_the_grid():
    for nz, g in [(1024, 8), (23, 2), (26, 4), (17, 8), (128, 1)]:
        parts = slabs.slab_partition(nz, g)
        assert parts[0][0] == 0 and sum(n for _, n in parts) == nz
        assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(g - 1))
        assert min(n for _, n in parts) >= 2 and max(n for _, n in parts) - min(n for _, n in parts) <= 1
    with pytest.raises(ValueError):
        slabs.slab_partition(7, 4)


def test_ensemble_partition_covers_all_trajectories():
    for b, g in [(512, 8), (512, 1), (5, 2), (3, 4)]:
        parts = slabs.ensemble_partition(b, g)
        assert parts[0][0] == 0 and sum(c for _, c in parts) == b
        assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(g - 1))


WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r)
    import torch.distributed as dist
    import nls_b200
    from nls_b200 import slabs
    dist.init_process_group("gloo")
    r, w = dist.get_rank(), dist.get_world_size()
    blob = bytes([r + 1]) * 128
    allb = slabs.gather_blobs(blob, dist)
    assert len(allb) == 128 * w and all(allb[128 * q] == q + 1 and allb[128 * q + 127] == q + 1 for q in range(w))
    z0, nz = slabs.slab_partition(23, w)[r]
    first, cnt = slabs.ensemble_partition(5, w)[r]
    import torch
    t = torch.tensor([nz, cnt]); dist.all_reduce(t)
    assert t.tolist() == [23, 5]
    # no GPU here: creating a slab handle must fail loudly, never fall back
    try:
        nls_b200.Engine(nls_b200.Desc.make(8, 8, nz, L=1.0, nz_global=23, z0=z0))
        ok = torch.cuda.is_available()
    except nls_b200.NlsError:
        ok = True
    assert ok
    dist.destroy_process_group()
    print("RANK_OK", r)
""")


def test_blob_gather_world_size_2_gloo(tmp_path):
    nls_b200.build_library()
    script = tmp_path / "w.py"
    script.write_text(WORKER % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29613", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-3000:]
    assert r.stdout.count("RANK_OK") == 2
