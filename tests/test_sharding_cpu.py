"""CPU-only, world_size 2 over gloo: the N>1 path of the search (contiguous bank shards, one all-gather of the per-rank
hit lists, order-independent merge) gives the toplist of the unsharded search.  The per-rank 'search' is replaced by
synthetic per-bin maxima because there is no GPU here; the device part of the same path is covered on the B200 by
tests/test_gpu_search.py::test_sharded_search_merges_to_the_same_toplist and by `bench.py --gpus N`."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _template_hits(t, fu, w2, rng_seed=1234):
    """deterministic above-threshold maxima of template t (a few hot bins shared by all templates -> cross-shard collisions)"""
    rng = np.random.default_rng(rng_seed + t)
    hot = np.random.default_rng(99).integers(w2, fu, 25)
    rows = []
    for k in range(5):
        bins = np.unique(np.concatenate([rng.choice(hot, 6), rng.integers(w2, fu, 10)]))
        power = (20.0 + 30.0 * rng.random(bins.size)).astype(np.float32)
        rows += [(p, b, t, k) for p, b in zip(power, bins) if b != hot[0]]
        if t % 5 == 2:
            rows.append((np.float32(57.5), hot[0], t, k))   # exact tie on one bin across shards: the earliest template must win
    return rows


def _per_bin_max(rows):
    best = {}
    for p, b, t, k in rows:
        key = (k, int(b))
        if key not in best or p > best[key][0] or (p == best[key][0] and t < best[key][1]):
            best[key] = (p, t)
    from brp_b200.sharding import HIT_DTYPE
    out = np.zeros(len(best), HIT_DTYPE)
    for i, ((k, b), (p, t)) in enumerate(best.items()):
        out[i] = (p, b, t, k)
    return out


def _worker(rank, world, port, n_templates, fu, w2, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, str(ROOT))
    import torch.distributed as dist
    import brp_b200
    from brp_b200 import sharding, synth

    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(n_templates, rank, world)
    rows = [r for t in range(lo, hi) for r in _template_hits(t, fu, w2)]
    mine = _per_bin_max(rows)                                  # what this rank's device table would hold
    if rank == 1:
        mine = mine[::-1].copy()                               # arrival order must not matter
    allhits = sharding.gather_hits(mine, dist)
    bank = synth.make_bank(n_templates)
    cands, tidx = brp_b200.toplist_from_hits(allhits, bank)
    # the production exchange: ONE fixed-size all-gather of the ranks' 5 x 100 strongest per-bin maxima
    records = sharding.gather_toplists(sharding.local_toplist_records(mine), dist)
    cands2, tidx2 = brp_b200.toplist_from_hits(records, bank)
    assert cands2.tobytes() == cands.tobytes() and tidx2.tobytes() == tidx.tobytes()
    assert len(records) <= 2 * 500
    q.put((rank, lo, hi, len(mine), len(allhits), cands.tobytes(), tidx.tobytes()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_shards_merge_to_single_rank_toplist():
    import torch.multiprocessing as mp
    sys.path.insert(0, str(ROOT))
    import brp_b200
    from brp_b200 import sharding, synth

    n_templates, fu, w2 = 137, 6866, 100   # enough hits that the per-rank lists overflow 100 entries per harmonic
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_templates, fu, w2, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, n0, tot0, c0, t0), (r1, lo1, hi1, n1, tot1, c1, t1) = results
    assert (lo0, hi0, lo1, hi1) == (0, 68, 68, 137)             # contiguous, sizes differ by at most one
    assert tot0 == tot1 == n0 + n1
    assert c0 == c1 and t0 == t1                                 # every rank ends with the same merged list
    # single-rank answer
    single = _per_bin_max([r for t in range(n_templates) for r in _template_hits(t, fu, w2)])
    bank = synth.make_bank(n_templates)
    cands, tidx = brp_b200.toplist_from_hits(single, bank)
    assert cands.tobytes() == c0 and tidx.tobytes() == t0
    tie = cands["power"] == np.float32(57.5)
    assert tie.sum() == 5 and np.all(tidx[tie] == 2)             # one per harmonic; earliest of the tied templates (2, 7, 12, ...)
    assert all(np.count_nonzero(cands["power"][100 * k:100 * k + 100]) == 100 for k in range(5))


def test_local_toplist_records_keep_what_the_merge_needs():
    """numpy restatement of the per-rank reduction against brp_toplist_from_hits on random hit lists with ties"""
    sys.path.insert(0, str(ROOT))
    import brp_b200
    from brp_b200 import sharding, synth

    rng = np.random.default_rng(3)
    bank = synth.make_bank(50)
    for trial in range(20):
        n = int(rng.integers(0, 1500))
        hits = np.zeros(n, sharding.HIT_DTYPE)
        hits["power"] = rng.integers(20, 60, n).astype(np.float32)      # few distinct values: many ties
        hits["bin"] = rng.integers(100, 1100, n)
        hits["tmpl"] = rng.integers(0, 50, n)
        hits["harm_idx"] = rng.integers(0, 5, n)
        rec = sharding.local_toplist_records(hits)
        a, ta = brp_b200.toplist_from_hits(hits, bank)
        b, tb = brp_b200.toplist_from_hits(rec[rec["power"] > 0], bank)
        assert a.tobytes() == b.tobytes() and ta.tobytes() == tb.tobytes()
        m = a["power"] > 0
        assert np.array_equal(rec["power"][m], a["power"][m].astype(np.float32)) and np.array_equal(rec["bin"][m], a["f0"][m])


def test_shard_ranges_cover_the_bank():
    sys.path.insert(0, str(ROOT))
    from brp_b200.sharding import shard_range
    for n in (0, 1, 7, 6662):
        for world in (1, 2, 4, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
