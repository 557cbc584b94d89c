"""Template-bank sharding over ranks and the single exchange of a multi-GPU search.

Templates are independent given the time series, so the bank is cut into contiguous shards, one per rank; every
rank searches its shard with global template indices, and the only communication is ONE all-gather of the per-rank
hit lists (per-bin maxima above threshold, 16 bytes each).  The merge itself is `brp_toplist_from_hits`, which is
order-independent and takes the earliest template on ties.  Works with any torch.distributed backend: NCCL on
GPUs (`bench.py`), gloo on CPU (`tests/test_sharding_cpu.py`).
"""
from __future__ import annotations

import numpy as np

HIT_DTYPE = np.dtype([("power", "<f4"), ("bin", "<u4"), ("tmpl", "<u4"), ("harm_idx", "<u4")])


def shard_range(n_templates: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous range [lo, hi) of global template indices searched by `rank` (sizes differ by at most one)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return rank * n_templates // world, (rank + 1) * n_templates // world


def gather_hits(hits: np.ndarray, dist=None, device=None) -> np.ndarray:
    """All-gather the ranks' hit lists (variable length) and return their concatenation in rank order.

    Two collectives on the wire: a MAX all-reduce of the list length (4 bytes) and one all-gather of the padded
    lists.  `device` is the torch device the backend needs its tensors on (cuda for NCCL, cpu for gloo)."""
    hits = np.ascontiguousarray(hits, HIT_DTYPE)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return hits
    import torch

    world = dist.get_world_size()
    dev = torch.device("cpu") if device is None else device
    n = torch.tensor([len(hits)], dtype=torch.int64, device=dev)
    cap = n.clone()
    dist.all_reduce(cap, op=dist.ReduceOp.MAX)
    rows = int(cap.item())
    buf = torch.zeros((rows + 1, 4), dtype=torch.int32, device=dev)
    buf[0, 0] = len(hits)
    if len(hits):
        buf[1:len(hits) + 1] = torch.from_numpy(hits.view(np.int32).reshape(-1, 4).copy()).to(dev)
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    parts = []
    for o in out:
        o = o.cpu().numpy()
        k = int(o[0, 0])
        parts.append(np.ascontiguousarray(o[1:k + 1]).view(HIT_DTYPE).reshape(-1))
    return np.concatenate(parts) if parts else hits[:0]
