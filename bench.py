#!/usr/bin/env python
"""Benchmark of the BRP per-template search loop (BASELINE.json: orbital templates/sec on a 2^22-sample work-unit).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repository's CUDA path
    python bench.py --impl reference [--steps K] [--warmup W]      # the reference's own CPU code (oracle/_ref)

A "step" is one pass of the hot path over the whole template bank (6 662 templates, like the reference's bundled bank)
of one synthetic 2^22-sample work-unit: resample -> R2C FFT + power spectrum -> harmonic summing -> per-bin toplist, for
every template.  With N GPUs the ONE bank is split into N contiguous shards (strong scaling: one process per GPU, no
data-path collective); the only exchange is one fixed-size all-gather of the per-GPU toplists (5 x 100 records each) in the
end-to-end leg.

One JSON line is printed by rank 0.  `value` is device time (CUDA events on the library's stream, series already resident
in HBM, max over ranks); `e2e` is the same metric through the public C ABI with HOST buffers (H2D of the series and the
bank, D2H of the hit list, toplist exchange and host-side merge inside the timed region).

The reference arm imports nothing of the product package: its inputs come from the numpy-only helpers loaded by path
(brp_b200_synth) and its geometry from the oracle.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "orbital templates/sec (2^22-sample WU)"
UNIT = "templates/s"
F0, FA, WINDOW = 400.0, 0.08, 1000


def workload_name(log2n, templates, padding):
    return f"synthetic 2^{log2n}-sample WU, one {templates}-template bank, -P {padding:g} -f 400 -A 0.08 -B 1000"


def make_inputs(args):
    """series + bank from the numpy-only generator (no CUDA library involved)"""
    from brp_b200_synth import synth
    nu = 1 << args.log2n
    series, _ = synth.make_series(nu, seed=synth.SEED)
    bank = synth.make_bank(args.templates, seed=synth.SEED + 1)
    return synth, series, bank


def algorithmic_bytes(g):
    """unique inputs read once + unique outputs written once, float32, per template."""
    nu, n, f, h = g.nsamples_unpadded, g.nsamples, g.fft_size, g.harmonic_idx_hi
    return {"resample": 4 * (nu + n), "fft_power": 4 * (n + f), "harmonic_sum": 4 * h, "total": 4 * (nu + 2 * n + f + h)}


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""

    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self, t0: float, t1: float):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.2] or [r for (_, r) in self.rows]
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------------- reference arm

def _ref_worker(job):
    """One process = one single-threaded reference instance (its state is module-global, like a BOINC task)."""
    series, g_fields, rows, fA = job
    from oracle.oracle import Geometry, Oracle, Reference
    ref, orc = Reference(), Oracle()
    g = Geometry(*g_fields)
    thr = orc.base_thresholds(fA, g.fft_size)
    t0 = time.perf_counter()
    for row in rows:
        t = orc.template(*row, g.step_inv)
        res = ref.resample(series, g, t)                      # demod_binary_resamp_cpu.c run_resampling
        ps = ref.fft_power(res, g.nsamples)                    # demod_binary_fft_fftw.c run_fft (FFT = oracle/shim_fftw.c, NOT FFTW)
        ref.harmonic_summing(ps, g, thr)                       # demod_binary_hs_cpu.c run_harmonic_summing
    return time.perf_counter() - t0


def cpu_reference_rate(series, g, bank, per_core: int, cores: int, fA: float = FA):
    """templates/s of the reference CPU translation units on `cores` host cores, `per_core` templates each."""
    import multiprocessing as mp
    from oracle.oracle import Reference
    if not Reference.available():
        return None
    Reference()  # load oracle/_ref/libbrp_ref.so in this process too, before forking: the workers inherit the mapping
    g_fields = tuple(getattr(g, f[0]) for f in g._fields_)
    jobs = [(series, g_fields, bank[(c * per_core) % len(bank):][:per_core], fA) for c in range(cores)]
    t0 = time.perf_counter()
    if cores == 1:
        _ref_worker(jobs[0])
    else:
        with mp.get_context("fork").Pool(cores) as pool:
            pool.map(_ref_worker, jobs)
    dt = time.perf_counter() - t0
    return per_core * cores / dt, dt


def pocketfft_anchor(n: int):
    """single-thread float32 real FFT of the same length with scipy/pocketfft: a sanity anchor for the FFT behind the FFTW API
    of the CPU baseline (oracle/shim_fftw.c computes in double; FFTW itself is not installable here) -- BASELINE.md section 4"""
    try:
        import scipy.fft
        x = np.random.default_rng(1).standard_normal(n).astype(np.float32)
        scipy.fft.rfft(x, workers=1)
        t0 = time.perf_counter()
        for _ in range(3):
            scipy.fft.rfft(x, workers=1)
        return (time.perf_counter() - t0) / 3 * 1e3
    except Exception:
        return None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.oracle import Oracle
    _, series, bank = make_inputs(args)
    g = Oracle().geometry(1 << args.log2n, 65.47619047619051, args.padding, F0, WINDOW)
    cores = os.cpu_count() or 1
    per_core = args.ref_templates_per_core
    times = []
    for it in range(args.warmup + args.steps):
        r = cpu_reference_rate(series, g, bank, per_core, cores)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"}))
            return
        if it >= args.warmup:
            times.append(r[1])
    total = per_core * cores * len(times)
    value = total / sum(times)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.log2n, args.templates, args.padding),
                   "note": "reference CPU translation units compiled unmodified (oracle/_ref); FFT behind the FFTW API is "
                           "oracle/shim_fftw.c (double-precision mixed radix), FFTW itself is not installable here"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": f"{per_core} templates per core on {cores} cores per step, one single-threaded reference instance per core",
                         "pocketfft_f32_rfft_ms": pocketfft_anchor(g.nsamples)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line, default=float))


# --------------------------------------------------------------------------------------------------- our arm

def stage_report(brp_b200, s, shard, g, nb, peak):
    """CUDA events around each stage of one batch, on the library's stream; fractions = algorithmic bytes / time / HBM peak."""
    alg = algorithmic_bytes(g)
    ms = [float(x) for x in s.profile_batch(shard, nb, repeats=5)]
    front_ms = ms[0] + ms[1] + ms[2]
    stages = {
        "resample": {"ms_per_batch": ms[0], "GBps": alg["resample"] * nb / (ms[0] * 1e-3) / 1e9},
        "fft_power": {"ms_per_batch": ms[1] + ms[2], "pass1_ms": ms[1], "pass2_ms": ms[2],
                      "GBps": alg["fft_power"] * nb / ((ms[1] + ms[2]) * 1e-3) / 1e9},
        "harmonic_sum": {"ms_per_batch": ms[3], "GBps": alg["harmonic_sum"] * nb / (ms[3] * 1e-3) / 1e9},
        # the front end as a whole (what a fusion of the resampler into the FFT would be held against): not a candidate for
        # the dominant stage, the three above are
        "resample_fft_power": {"ms_per_batch": front_ms, "resample_ms": ms[0], "pass1_ms": ms[1], "pass2_ms": ms[2],
                               "GBps": (alg["resample"] + alg["fft_power"]) * nb / (front_ms * 1e-3) / 1e9},
    }
    for v in stages.values():
        v["frac"] = v["GBps"] / peak
    return alg, stages


def other_config(brp_b200, synth, log2n, padding, n_handles, templates, batch, peak, device):
    """One of the other BASELINE.json configurations, device-timed after the headline region (series resident)."""
    import torch
    nu = 1 << log2n
    g = brp_b200.derive_geometry(nu, synth.TSAMPLE_US, padding, F0, WINDOW)
    bank = synth.make_bank(templates, seed=synth.SEED + 1)
    handles = []
    for k in range(n_handles):
        h = brp_b200.Search(g, fA=FA, max_batch=batch, device=device)
        h.set_series(synth.make_series(nu, seed=synth.SEED + 100 + k)[0])
        handles.append(h)
    for h in handles:
        h.run(bank)
    for h in handles:
        h.sync()
    steps = 3
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        for h in handles:                      # all work-units enqueued before any is awaited: they overlap on the device
            h.reset()
            h.run(bank)
    for h in handles:
        h.sync()
    dt = time.perf_counter() - t0
    dev_ms = max(h.last_run_ms() for h in handles)
    alg, stages = stage_report(brp_b200, handles[0], bank, g, min(batch, templates), peak)
    for h in handles:
        h.close()
    value = templates * n_handles * steps / dt
    return {"value": value, "unit": UNIT, "templates": templates, "work_units_in_flight": n_handles, "batch": batch,
            "timing": "wall clock around enqueue of all steps + sync (series resident)", "last_step_device_ms": dev_ms,
            "algorithmic_bytes_per_template": alg["total"], "frac_of_hbm_roofline": alg["total"] * value / 1e9 / peak,
            "stages": {k: {"us_per_template": 1e3 * v["ms_per_batch"] / min(batch, templates), "frac": v["frac"]} for k, v in stages.items()}}


def run_ours(args):
    import torch
    import brp_b200
    from brp_b200 import sharding, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this implementation has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    _, series, bank = make_inputs(args)
    g = brp_b200.derive_geometry(1 << args.log2n, synth.TSAMPLE_US, args.padding, F0, WINDOW)
    n_total = len(bank)
    lo, hi = sharding.shard_range(n_total, rank, world)       # strong scaling: contiguous shards of the ONE bank
    shard = np.ascontiguousarray(bank[lo:hi])

    s = brp_b200.Search(g, fA=FA, max_batch=args.batch, device=local)
    stream = torch.cuda.ExternalStream(s.stream, device=local)
    pinned = torch.from_numpy(series).pin_memory()
    d_series = torch.from_numpy(series).cuda()
    s.set_series_device(d_series.data_ptr())
    s.sync()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def device_timed(templates_bank, first, steps, warmup):
        """(device ms of `steps` steps, max over ranks; kernels this rank launched in them)"""
        for _ in range(warmup):
            s.reset()
            s.run(templates_bank, first)
        s.sync()
        barrier()
        n0 = s.kernel_launches
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for _ in range(steps):
            s.reset()
            s.run(templates_bank, first)
        ev1.record(stream)
        barrier()
        return max_over_ranks(ev0.elapsed_time(ev1)), s.kernel_launches - n0

    # ---- device-resident throughput: K steps, CUDA events on the launching stream, max over ranks
    sampler = ClockSampler(local) if rank == 0 else None
    t_wall0 = time.perf_counter()
    dev_ms, launches = device_timed(shard, lo, args.steps, args.warmup)
    t_wall1 = time.perf_counter()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    value = n_total * args.steps / (dev_ms * 1e-3)

    # ---- end to end through the C ABI with host buffers (series + bank in, hits out, one exchange, host merge).
    # Two handles alternate, as a driver with a queue of work-units would use them: the upload of step k+1's series overlaps the
    # search of step k, and step k's read-back, exchange and host merge overlap the search of step k+1.  Every step still
    # uploads its own series and bank from (pinned) host memory and reads its own hit list back.
    s2 = brp_b200.Search(g, fA=FA, max_batch=args.batch, device=local)
    handles = [s, s2]
    hit_bytes = 0

    def begin(h):
        h.reset()
        h.set_series_ptr(pinned.data_ptr())            # H2D of the work-unit from pinned host memory (asynchronous)

    def launch(h):
        h.run(shard, lo)                               # derives template constants on the host, H2D, enqueues all batches

    def finish(h):
        nonlocal hit_bytes
        hits = h.collect()                             # D2H of the compacted per-bin maxima of this shard
        hit_bytes = hits.nbytes + 4
        # the single exchange of the search: ONE fixed-size all-gather (5 x 100 records per rank) over NCCL
        records = sharding.gather_toplists(sharding.local_toplist_records(hits), dist, torch.device("cuda", local))
        if rank == 0:
            return brp_b200.toplist_from_hits(records, bank)
        return None

    def e2e_steps(n):
        top = None
        begin(handles[0])
        launch(handles[0])
        for k in range(n):
            cur, nxt = handles[k % 2], handles[(k + 1) % 2]
            if k + 1 < n:
                begin(nxt)
            cur.sync()
            if k + 1 < n:
                launch(nxt)
            top = finish(cur)
        return top

    e2e_steps(2)
    barrier()
    t0 = time.perf_counter()
    top = e2e_steps(args.steps)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = n_total * args.steps / e2e_s
    s2.close()

    # ---- secondary: weak scaling (every GPU searches the whole bank), for comparison with round 1's records
    weak = None
    if world > 1 and not args.no_weak:
        wsteps = max(1, min(args.steps, 3))
        weak_ms, _ = device_timed(bank, 0, wsteps, 1)
        weak = {"value": world * n_total * wsteps / (weak_ms * 1e-3), "unit": UNIT, "templates_per_gpu": n_total,
                "ms_per_step": weak_ms / wsteps}

    line = None
    if rank == 0:
        peaks_path = ROOT / "MEASURED_PEAKS.json"
        if peaks_path.exists():
            peak, peak_src = json.loads(peaks_path.read_text())["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        nb = min(args.batch, len(shard))
        alg, stages = stage_report(brp_b200, s, shard, g, nb, peak)
        dom = max(("resample", "fft_power", "harmonic_sum"), key=lambda k: stages[k]["ms_per_batch"])
        # DRAM bytes per launch of the dominant stage, from the committed ncu --set full capture (same batch size only)
        traffic = None
        tpath = ROOT / "profiles" / "r2_traffic.json"
        if tpath.exists() and args.log2n == 22 and args.padding == 1.0:
            t = json.loads(tpath.read_text())
            if t.get("batch") == nb:
                traffic = t["dram_bytes_per_launch"].get(dom)
        roofline = {"bound": "hbm", "kernel": dom, "achieved": stages[dom]["GBps"], "peak": peak, "peak_source": peak_src,
                    "unit": "GB/s", "frac": stages[dom]["GBps"] / peak, "traffic": traffic,
                    "algorithmic_bytes_per_template": alg, "templates_per_launch": nb, "stages": stages,
                    "note": "stage fractions are algorithmic bytes / CUDA-event time / HBM peak; the resampler's 4*Nu input bytes per template "
                            "come from L2 (the series is shared by the batch), so its fraction can exceed 1; 'resample_fft_power' is the front "
                            "end as a whole against the sum of its two stages' bytes",
                    "whole_loop": {"GBps": alg["total"] * value / world / 1e9, "frac": alg["total"] * value / world / 1e9 / peak,
                                   "note": "per GPU"}}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            from oracle.oracle import Oracle
            go = Oracle().geometry(1 << args.log2n, synth.TSAMPLE_US, args.padding, F0, WINDOW)
            r = cpu_reference_rate(series, go, bank, args.cpu_sample, 1)
            if r is not None:
                cpu = {"value": r[0], "unit": UNIT, "cores": 1, "kind": "reference",
                       "sample": f"first {args.cpu_sample} templates of the same bank on 1 core ({r[1]:.1f} s); reference objects from oracle/_ref, "
                                 "FFT = oracle/shim_fftw.c (FFTW not installable)", "host_cores_available": os.cpu_count(),
                       "pocketfft_f32_rfft_ms": pocketfft_anchor(g.nsamples)}
        others = None
        if world == 1 and not args.no_other_configs and args.log2n == 22 and args.padding == 1.0:
            s.close()
            others = {
                "2^22, -P 3.0 (reference bench padding)": other_config(brp_b200, synth, 22, 3.0, 1, 1536, args.batch, peak, local),
                "2^23-sample padded series (-P 2.0)": other_config(brp_b200, synth, 22, 2.0, 1, 2048, args.batch, peak, local),
                "2^20-sample WUs, 4 in flight": other_config(brp_b200, synth, 20, 1.0, 4, 2048, args.batch, peak, local),
            }
        n_cands = int(np.count_nonzero(top[0]["power"])) if top else 0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.log2n, args.templates, args.padding), "templates_per_step": n_total,
                       "templates_per_gpu": len(shard), "batch": args.batch,
                       "parallelism": f"one bank split into {world} contiguous shard(s), one GPU each",
                       "l2": "per-batch working set (FFT intermediate + spectrum of 64 templates) exceeds the 126 MB L2; the 16.8 MB input "
                             "series is shared by all templates and deliberately L2-resident", "candidates_found": n_cands},
            "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(series.nbytes + 16 * len(shard)),
                    "d2h_bytes_per_step": int(hit_bytes), "ms_per_step": 1e3 * e2e_s / args.steps,
                    "pipelining": "two handles alternate: step k+1's upload overlaps step k's search, step k's read-back and merge overlap step k+1's search",
                    "exchange": "one all-gather of 8 000 bytes per rank" if world > 1 else "none (one GPU)"},
            "gpu_launches": int(launches),
        }
        if weak:
            line["weak_scaling"] = weak
        if others:
            line["other_configs"] = others
    s.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line, default=float))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=int(os.environ.get("BRP_BATCH", "64")))
    ap.add_argument("--templates", type=int, default=6662)
    ap.add_argument("--log2n", type=int, default=22)
    ap.add_argument("--padding", type=float, default=1.0)
    ap.add_argument("--cpu-sample", type=int, default=32)
    ap.add_argument("--ref-templates-per-core", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    ap.add_argument("--no-weak", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
