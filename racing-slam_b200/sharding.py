"""Multi-GPU sharding of the batched workloads.  One process per GPU; shards are
independent, so there is NO data-path collective: each rank matches its own slice and the host
gathers results.  The only optional exchange is replicating the descriptor database / point set
(torch.distributed.broadcast over NCCL), done once before matching.

  all_pairs(n)                      LoopDetector ordered pairs query > candidate (src/LoopDetector.h:33)
  shard_pairs(pairs, rank, world)   block-cyclic split of a pair list, balanced for ragged frames
  shard_keyframes(n, rank, world)   round-robin split of the fuse window (src/Mapper.cpp:208-219)
  replicate(array, src)             the one exchange: the database / point set from one rank to all
"""
from __future__ import annotations

import numpy as np


def all_pairs(n_frames: int) -> np.ndarray:
    """Every (query, candidate) with query > candidate, query-major: [n(n-1)/2, 2] int32."""
    q, c = np.tril_indices(n_frames, -1)
    return np.stack([q, c], 1).astype(np.int32)


def shard_pairs(pairs: np.ndarray, rank: int, world: int, block: int = 64) -> np.ndarray:
    """Block-cyclic: blocks of `block` consecutive pairs go to ranks round-robin.  Consecutive
    pairs share their query frame (L2 reuse on the owning GPU); cycling the blocks balances the
    triangle.  The union over ranks is the input, each pair exactly once."""
    pairs = np.asarray(pairs, np.int32).reshape(-1, 2)
    if world <= 1:
        return pairs
    blk = np.arange(len(pairs)) // block
    return np.ascontiguousarray(pairs[blk % world == rank])


def shard_index(n: int, rank: int, world: int) -> np.ndarray:
    """Round-robin item indices of this rank."""
    return np.arange(rank, n, max(world, 1), dtype=np.int64)


def shard_keyframes(n_keyframes: int, rank: int, world: int) -> np.ndarray:
    return shard_index(n_keyframes, rank, world)


def gather_counts(local: np.ndarray, index: np.ndarray, n_total: int, group=None) -> np.ndarray:
    """Host-side gather of per-item integer results into item order (uses torch.distributed
    when initialised; a single process returns its own)."""
    out = np.zeros(n_total, np.int64)
    out[index] = local
    try:
        import torch
        import torch.distributed as dist
    except ImportError:   # pragma: no cover
        return out
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        t = torch.from_numpy(out)
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)   # shards are disjoint: sum = gather
        out = t.cpu().numpy()
    return out


def replicate(array, src: int = 0, group=None, device=None):
    """Replicate `array` (numpy or a device tensor, read on the source rank only) from rank `src` to
    every rank: shape and dtype travel first, then ONE broadcast of the payload -- over NCCL the payload is a
    device tensor, so the copy runs GPU to GPU through NVLink / NVSwitch and arrives already resident.
    -> (numpy array or None, device tensor or None): NCCL returns the device tensor (host copy left to the caller),
    gloo / single process the numpy array."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        a = np.ascontiguousarray(array)
        return a, (torch.from_numpy(a).to(device) if device is not None else None)
    rank = dist.get_rank(group)
    nccl = dist.get_backend(group) == "nccl"
    dev = torch.device(device) if device is not None else (torch.device("cuda", torch.cuda.current_device()) if nccl
                                                           else torch.device("cpu"))
    meta = [None]
    if rank == src:
        if isinstance(array, torch.Tensor):          # already resident on the source GPU: no staging at all
            resident = array.contiguous()
            meta = [(tuple(resident.shape), np.dtype(str(resident.dtype).replace("torch.", "")).str)]
        else:
            resident = None
            a = np.ascontiguousarray(array)
            meta = [(a.shape, a.dtype.str)]
    dist.broadcast_object_list(meta, src=src, group=group)
    shape, dtype = meta[0]
    nbytes = int(np.prod(shape, dtype=np.int64)) * np.dtype(dtype).itemsize
    if rank == src:
        t = resident.reshape(-1).view(torch.uint8).to(dev) if resident is not None else \
            torch.from_numpy(a.reshape(-1).view(np.uint8)).to(dev)
    else:
        t = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    if nbytes:
        dist.broadcast(t, src=src, group=group)
    if dev.type == "cuda":
        return None, t.view(_TORCH_DTYPES(torch)[np.dtype(dtype).str]).reshape(shape)
    return t.numpy().view(np.dtype(dtype)).reshape(shape), None


def _TORCH_DTYPES(torch):
    return {np.dtype(np.uint8).str: torch.uint8, np.dtype(np.int32).str: torch.int32, np.dtype(np.int64).str: torch.int64,
            np.dtype(np.float32).str: torch.float32, np.dtype(np.float64).str: torch.float64}
