"""Slab sharding across GPUs: one process per GPU, no collective on the data path.

Every (variable, time step) slab -- and inside it every level / ensemble member -- is independent given the read-only
cache tables; this is how the reference's writer splits work between OpenMP threads and MPI ranks
(src/io/netcdf/NetCDF_CDMWriter.cc:599-618: `unLimDimPos % size == rank`, falling back to another axis when there are
at most 3 time steps).  Here: slabs are dealt round-robin; when there are fewer slabs than ranks the level axis of each
slab is cut into contiguous chunks instead.  torch.distributed is only used to hand the finished pieces to the rank that
writes (gather), never between kernels.
"""
from typing import Dict, List, Tuple

WorkItem = Tuple[int, int, int]  # (slab index, first level, one-past-last level)


def assign_slabs(n_slabs: int, rank: int, world: int) -> List[int]:
    """slab t belongs to rank t % world"""
    return [t for t in range(n_slabs) if t % world == rank]


def split_levels(nz: int, part: int, parts: int) -> Tuple[int, int]:
    """contiguous, balanced chunk `part` of `parts` of the level axis"""
    base, rem = divmod(nz, parts)
    z0 = part * base + min(part, rem)
    return z0, z0 + base + (1 if part < rem else 0)


def plan(n_slabs: int, nz: int, world: int) -> List[List[WorkItem]]:
    """work items per rank.  n_slabs >= world: whole slabs round-robin.  Fewer slabs than ranks: each slab's levels are cut
    into world // n_slabs (at least 1) chunks so that every rank has work; left-over ranks idle."""
    out: List[List[WorkItem]] = [[] for _ in range(world)]
    if n_slabs >= world or nz <= 1:
        for r in range(world):
            out[r] = [(t, 0, nz) for t in assign_slabs(n_slabs, r, world)]
        return out
    parts = max(1, min(nz, world // max(n_slabs, 1)))
    r = 0
    for t in range(n_slabs):
        for p in range(parts):
            z0, z1 = split_levels(nz, p, parts)
            if z1 > z0:
                out[r % world].append((t, z0, z1))
                r += 1
    return out


def gather_pieces(local: Dict[WorkItem, object], dst: int = 0):
    """hand the finished pieces to rank `dst` (the writer).  Returns {work item: piece} there, None elsewhere."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(local)
    gathered = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(local, gathered, dst=dst)
    if dist.get_rank() != dst:
        return None
    merged: Dict[WorkItem, object] = {}
    for part in gathered:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError(f"work item {k} produced twice")
            merged[k] = v
    return merged


def assemble(pieces: Dict[WorkItem, object], n_slabs: int, nz: int):
    """stack the level chunks of every slab back together (numpy arrays [z1-z0, ...]); checks complete coverage"""
    import numpy as np
    slabs = []
    for t in range(n_slabs):
        chunks = sorted((k for k in pieces if k[0] == t), key=lambda k: k[1])
        z = 0
        for (_, z0, z1) in chunks:
            if z0 != z:
                raise RuntimeError(f"slab {t}: levels {z}..{z0} missing or duplicated")
            z = z1
        if z != nz:
            raise RuntimeError(f"slab {t}: levels {z}..{nz} missing")
        slabs.append(np.concatenate([pieces[k] for k in chunks], axis=0))
    return slabs
