"""Host-side wiring of the two multi-GPU partitionings.  One process per GPU.

* ensemble: trajectories are independent -> static block partition, no communication;
* z-slabs: one handle per GPU owns planes [z0, z0+nz) of the global grid; the handles are wired
  together once (128-byte blobs gathered in rank order), after which halo planes and Lanczos
  scalars travel inside the kernels over NVLink peer memory (include/nls.h, nls_slab_*).

torch.distributed is plumbing only (rendezvous + one all-gather of the blobs); works with the
`gloo` backend on CPU for the host-logic tests and `nccl` on the GPU box.
"""
import numpy as np

BLOB_BYTES = 128


def slab_partition(nz_global, nranks):
    """[(z0, nz)] per rank: contiguous plane ranges, sizes differing by at most one, >= 2 planes each."""
    if nranks < 1 or nz_global < 2 * nranks:
        raise ValueError(f"cannot cut {nz_global} planes into {nranks} slabs of >= 2 planes")
    base, rem = divmod(nz_global, nranks)
    out, z = [], 0
    for r in range(nranks):
        nz = base + (1 if r < rem else 0)
        out.append((z, nz))
        z += nz
    return out


def ensemble_partition(ntraj, nranks):
    """[(first, count)] per rank: GPU g advances trajectories [g*B/G, (g+1)*B/G)."""
    return [(r * ntraj // nranks, (r + 1) * ntraj // nranks - r * ntraj // nranks) for r in range(nranks)]


def gather_blobs(blob, dist=None, device=None):
    """all-gather one 128-byte blob per rank, returned concatenated in rank order."""
    import torch
    blob = bytes(blob)
    if len(blob) != BLOB_BYTES:
        raise ValueError("slab blob must be 128 bytes")
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return blob
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
    if device is not None:
        mine = mine.to(device)
    parts = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, mine)
    return b"".join(bytes(p.cpu().numpy().tobytes()) for p in parts)


def make_slab_engine(make_desc, nz_global, dist, rank, world, device=None):
    """Create this rank's slab handle and connect it to its peers.
    make_desc(nz, nz_global, z0) -> Desc.  world == 1 gives the plain single-GPU handle."""
    from .binding import Engine
    z0, nz = slab_partition(nz_global, world)[rank]
    eng = Engine(make_desc(nz, nz_global if world > 1 else 0, z0 if world > 1 else 0))
    if world > 1:
        blobs = gather_blobs(eng.slab_export(), dist, device)
        eng.slab_connect(blobs, rank, world)
        dist.barrier()
    return eng, z0, nz


def local_planes(a, z0, nz):
    """the planes [z0, z0+nz) of a global (..., nz_global, ny, nx) array, contiguous"""
    return np.ascontiguousarray(a[..., z0:z0 + nz, :, :])
