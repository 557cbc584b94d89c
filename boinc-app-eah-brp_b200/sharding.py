"""Template-bank sharding over ranks and the single exchange of a multi-GPU search.

Templates are independent given the time series, so the bank is cut into contiguous shards, one per rank; every
rank searches its shard with global template indices, and the only communication is ONE fixed-size all-gather of the
per-rank toplists: 5 x 100 records {power, bin, global template index, harmonic}, 8 000 bytes per rank
(`local_toplist_records` + `gather_toplists`).  That is exact: a bin of the global top 100 of a harmonic is in the
top 100 of the shard that holds its maximising template, because every bin above it there is also above it globally.
The merge itself is `brp_toplist_from_hits`, which is order-independent and takes the earliest template on ties.
Works with any torch.distributed backend: NCCL on GPUs (`bench.py`), gloo on CPU (`tests/test_sharding_cpu.py`).
`gather_hits` (whole hit lists, variable length, two collectives) is the older form, kept for comparison in tests.
"""
from __future__ import annotations

import numpy as np

HIT_DTYPE = np.dtype([("power", "<f4"), ("bin", "<u4"), ("tmpl", "<u4"), ("harm_idx", "<u4")])


def shard_range(n_templates: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous range [lo, hi) of global template indices searched by `rank` (sizes differ by at most one)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return rank * n_templates // world, (rank + 1) * n_templates // world


N_PER_HARMONIC = 100   # N_CAND_5, demod_binary.c:83


def local_toplist_records(hits: np.ndarray) -> np.ndarray:
    """The 5 x 100 strongest per-bin maxima of one rank's hit list as a fixed-size record array (empty slots: power 0).
    Same total order as brp_toplist_from_hits: descending power, then earliest template, then lowest bin."""
    hits = np.ascontiguousarray(hits, HIT_DTYPE)
    out = np.zeros(5 * N_PER_HARMONIC, HIT_DTYPE)
    for k in range(5):
        h = hits[hits["harm_idx"] == k]
        if len(h) == 0:
            continue
        # one entry per bin (hit lists of a device table already are; concatenated lists need not be)
        o = np.lexsort((h["tmpl"], -h["power"].astype(np.float64), h["bin"]))
        h = h[o]
        h = h[np.concatenate(([True], h["bin"][1:] != h["bin"][:-1]))]
        o = np.lexsort((h["bin"], h["tmpl"], -h["power"].astype(np.float64)))[:N_PER_HARMONIC]
        out[k * N_PER_HARMONIC:k * N_PER_HARMONIC + len(o)] = h[o]
    return out


def gather_toplists(records: np.ndarray, dist=None, device=None) -> np.ndarray:
    """ONE fixed-size all-gather (8 000 bytes per rank) of the ranks' `local_toplist_records`; returns the non-empty
    records of all ranks, in rank order, ready for brp_toplist_from_hits."""
    records = np.ascontiguousarray(records, HIT_DTYPE)
    assert records.shape == (5 * N_PER_HARMONIC,)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return records[records["power"] > 0]
    import torch

    dev = torch.device("cpu") if device is None else device
    mine = torch.from_numpy(records.view(np.int32).reshape(-1, 4).copy()).to(dev)
    out = torch.empty((dist.get_world_size() * mine.shape[0], 4), dtype=torch.int32, device=dev)  # ranks concatenated along dim 0
    dist.all_gather_into_tensor(out, mine)
    allrec = np.ascontiguousarray(out.cpu().numpy()).view(HIT_DTYPE).reshape(-1)
    return allrec[allrec["power"] > 0]


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
