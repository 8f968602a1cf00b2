#!/usr/bin/env python
"""Offer-matching benchmark: job x offer evaluations per second on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one batch of pending jobs matched and ranked against the whole resident catalog
(table build + match + merge kernels; for N>1 also the NCCL all-gather of per-shard top-k and
the shard merge).

  N = 1   config 3 of BASELINE.json: 4096 jobs x 1,048,576 offers, top-16, seed 3.
  N > 1   config 4: 16384 jobs x 67,108,864 offers, top-16, seed 4, the catalog sharded over the
          N GPUs in rank order (strong scaling: total work fixed).

`value` is measured with the job records and the catalog resident in HBM (CUDA events, per
step, L2 flushed between steps); `e2e` goes through the C-ABI host-buffer call: pinned host
records in, results out, copies inside the timed region.  `--impl reference` times the CPU
restatement of the reference path (oracle/reference_path.py: the reference is Python and
cannot travel to the GPU box) on all host cores, on a bounded sample of the same workload.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "job_x_offer_evaluations_per_second"
UNIT = "evals/s"


def workload(n_gpus: int, override: str = ""):
    name = override or ("cfg3" if n_gpus == 1 else "cfg4")
    if name == "cfg3" or name.startswith("cfg3:k"):  # cfg3:k50 etc. = config 3 with another k
        k = int(name.split(":k")[1]) if ":k" in name else 16
        return dict(name=name, workload=f"cfg3: 4096 jobs x 1048576 offers, top-{k}, seed 3", jobs=4096,
                    offers=1 << 20, k=k, seed=3, shapes=4096)
    if name == "cfg4":
        return dict(name="cfg4", workload="cfg4: 16384 jobs x 67108864 offers sharded over the GPUs, top-16, seed 4",
                    jobs=16384, offers=1 << 26, k=16, seed=4, shapes=4096)
    if name.startswith("custom:"):  # tuning: custom:<jobs>x<offers>
        j, n = (int(x) for x in name.split(":", 1)[1].split("x"))
        return dict(name=name, workload=f"{name}, top-16, seed 4", jobs=j, offers=n, k=16, seed=4, shapes=4096)
    if name == "tiny":  # CPU-side plumbing tests
        return dict(name="tiny", workload="tiny: 256 jobs x 65536 offers, top-16, seed 3", jobs=256, offers=1 << 16, k=16,
                    seed=3, shapes=512)
    raise SystemExit(f"unknown workload {name}")


# ------------------------------------------------------------------------------------------ clocks


class ClockSampler:
    """Polls SM clock and throttle reasons through NVML from a thread while the timed regions
    run (the regions last milliseconds, so `nvidia-smi -lms` is too coarse)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self.thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception as e:  # no NVML: report that instead of a number
            self.error = repr(e)

    def _poll(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.0005)

    def stop(self):
        if self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "error", "?")]}
        self._stop.set()
        self.thread.join(timeout=2)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ------------------------------------------------------------------------------------------ CPU baseline

_BASE = {}


def _baseline_worker(job_ids):
    from oracle import reference_path as ref

    backends, jobs = _BASE["backends"], _BASE["jobs"]
    for j in job_ids:
        profile, req = jobs[j]
        res = ref.get_offers_by_requirements(backends, profile, req)
        ref.plan_slice([o for _, o in res], _BASE["k"])
    return len(job_ids)


class CpuBaseline:
    """The object-level restatement of the reference path (a fresh CatalogItem and a `matches`
    call per (job, offer), heapq.merge, stable partition, slice) on all host cores.  The
    reference itself is Python and cannot travel to the GPU box, so this is kind "port"."""

    def __init__(self, wl, cores: int, n_offers: int = 20_000, jobs_per_core: int = 48):
        import multiprocessing as mp

        from workloads import synth
        from dstack_b200.encode import encode_table
        from oracle import reference_path as ref

        self.cores, self.n_offers = cores, n_offers
        cat = synth.make_catalog(n_offers, seed=wl["seed"], n_shapes=min(wl["shapes"], 1024))
        per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
        table = encode_table(cat.columns)
        self.n_jobs = cores * jobs_per_core
        sj = synth.make_jobs(cat, table, self.n_jobs, seed=wl["seed"] + 100)
        _BASE.update(backends=[b for b, _ in per_backend], jobs=synth.job_objects(sj), k=wl["k"])
        self.pool = mp.get_context("fork").Pool(cores)
        self.pool.map(_baseline_worker, [[0]] * cores)  # start and warm the workers

    def step(self, jobs_per_core: int):
        n = min(self.n_jobs, jobs_per_core * self.cores)
        chunks = [list(range(i, n, self.cores)) for i in range(self.cores)]
        t0 = time.perf_counter()
        done = sum(self.pool.map(_baseline_worker, chunks))
        dt = time.perf_counter() - t0
        return done * self.n_offers / dt, dt, done

    def close(self):
        self.pool.close()
        self.pool.join()

    def describe(self, value, done, dt):
        return {"value": value, "unit": UNIT, "cores": self.cores, "kind": "port",
                "sample": f"{done} jobs x {self.n_offers} offers from the workload's generators, {dt:.2f} s wall "
                          f"on {self.cores} processes; oracle/reference_path.py (the Python reference cannot "
                          f"travel to the GPU box)"}


def cpu_baseline(wl, cores: int):
    cb = CpuBaseline(wl, cores)
    try:
        value, dt, done = cb.step(48)
    finally:
        cb.close()
    return cb.describe(value, done, dt)


def run_reference(args, wl, rank: int):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_core = max(2, min(48, 96 // max(1, args.steps + args.warmup)))
    cb = CpuBaseline(wl, cores, jobs_per_core=per_core)
    vals, secs, done = [], [], 0
    try:
        for i in range(args.warmup + args.steps):
            v, dt, done = cb.step(per_core)
            if i >= args.warmup:
                vals.append(v)
                secs.append(dt)
    finally:
        cb.close()
    value = statistics.mean(vals)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * statistics.mean(secs),
        "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "python float/int (f64)", "data": "synthetic",
        "config": {"workload": wl["workload"], "note": "each step is a bounded sample of the workload"},
        "cpu_baseline": cb.describe(value, done, statistics.mean(secs)),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------ ours


def verify(cat_cols, sj, k, index, count, max_price, rng, n_sample=24, order=None):
    """Full-size parity: sampled jobs against the numpy oracle on the whole catalog, plus
    size-independent properties for every job."""
    import dataclasses

    from oracle import columnar

    j = sj.columns.n
    valid = index >= 0
    assert ((valid.sum(axis=1)) == np.minimum(count, k)).all(), "number of returned rows != min(count, k)"
    assert (np.isnan(max_price) == (count == 0)).all(), "max_price / count disagree"
    if order is None:
        order = columnar.rank_order(cat_cols)
    rank_of = np.empty(len(order), np.int64)
    rank_of[order] = np.arange(len(order))
    ranks = np.where(valid, rank_of[np.maximum(index, 0)], np.iinfo(np.int64).max)
    assert (np.sort(ranks, axis=1) == ranks).all(), "top-k not in rank order"
    pick = rng.choice(j, size=min(n_sample, j), replace=False)
    jc = sj.columns
    sub = {}
    for f in dataclasses.fields(jc):
        v = getattr(jc, f.name)
        sub[f.name] = v
    # slice the sampled jobs (CSR fields need re-basing)
    def take_csr(ptr, ids):
        parts = [ids[ptr[i]:ptr[i + 1]] for i in pick]
        return np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64), \
            (np.concatenate(parts) if parts else np.zeros(0, np.int64)).astype(np.int64)
    for name in ("max_price", "spot", "cpu", "memory", "disk", "gpu_count", "gpu_memory", "gpu_total", "gpu_vendor",
                 "cc_min", "gpu_name_any", "src_backend_mask", "off_backend_mask", "exclude_not_available",
                 "region_any", "itype_any"):
        sub[name] = getattr(jc, name)[pick]
    sub["gpu_name_ptr"], sub["gpu_name_ids"] = take_csr(jc.gpu_name_ptr, jc.gpu_name_ids)
    sub["region_ptr"], sub["region_ids"] = take_csr(jc.region_ptr, jc.region_ids)
    sub["itype_ptr"], sub["itype_ids"] = take_csr(jc.itype_ptr, jc.itype_ids)
    sub["n"] = len(pick)
    want = columnar.topk(cat_cols, type(jc)(**sub), k, order=order)
    assert np.array_equal(want.count, count[pick]), "sampled counts differ from the oracle"
    assert np.array_equal(want.index, index[pick]), "sampled top-k differ from the oracle"
    assert np.array_equal(np.nan_to_num(want.max_price, nan=-1), np.nan_to_num(max_price[pick], nan=-1))
    return len(pick)


def run_ours(args, wl, rank: int, world: int, local_rank: int):
    import ctypes as C

    import torch

    from dstack_b200 import native
    from workloads import synth
    from dstack_b200.sharded import ShardedMatcher, rank_order

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    lib = native.lib()
    J, N, k = wl["jobs"], wl["offers"], wl["k"]

    t_setup = time.perf_counter()
    cat = synth.make_catalog(N, seed=wl["seed"], n_shapes=wl["shapes"])
    full = cat.columns
    order = rank_order(full)
    sm = ShardedMatcher(full, rank=rank, world=world, device=local_rank, order=order)
    table = sm.table
    sj = synth.make_jobs(cat, table, J, seed=wl["seed"])  # dictionaries are shared by every shard
    packed = sm.pack(sj.columns)
    rec_words = packed.layout["rec_words"]
    setup_s = time.perf_counter() - t_setup

    stream = sm.stream
    m = sm.matcher
    h = m.handle
    with torch.cuda.stream(stream):
        b = sm.prepare(packed, k)
        pin_rec = native.PinnedArray((J, rec_words), np.uint32)
        pin_rec.array[:] = packed.records
        pin_idx = native.PinnedArray((J, k), np.uint32)
        pin_cnt = native.PinnedArray((J,), np.uint32)
        pin_max = native.PinnedArray((J,), np.uint32)
        t_pin = torch.from_numpy(pin_rec.array.view(np.int32))
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        if world > 1:
            pin_oidx = torch.empty((J, k), dtype=torch.int64).pin_memory()
            pin_ocnt = torch.empty((J,), dtype=torch.int64).pin_memory()
            pin_omp = torch.empty((J,), dtype=torch.float64).pin_memory()

        def step_e2e():
            """Host buffers in, host buffers out."""
            if world == 1:
                native.check(lib.dsm_match_topk(h, sm.device_table.handle, C.byref(b["lay"]), pin_rec.array.ctypes.data,
                                                J, k, pin_idx.array.ctypes.data, pin_cnt.array.ctypes.data,
                                                pin_max.array.ctypes.data))
            else:
                b["rec"].copy_(t_pin, non_blocking=True)
                sm.step()
                pin_oidx.copy_(b["o_idx"], non_blocking=True)
                pin_ocnt.copy_(b["o_cnt"], non_blocking=True)
                pin_omp.copy_(b["o_mp"], non_blocking=True)
                stream.synchronize()

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)

        warm = max(3, args.warmup)
        for _ in range(warm):
            sm.step()
            flush.zero_()
        barrier()

        # ---- timed region 1: resident inputs
        clocks = ClockSampler(local_rank)
        native.check(lib.dsm_timing_enable(h, 1))
        launches0 = m.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        barrier()
        for e0, e1 in evs:
            e0.record(stream)
            sm.step()
            e1.record(stream)
            flush.zero_()  # evict the table from L2; outside the event pair
        barrier()
        launches = m.launch_count() - launches0
        step_ms = [e0.elapsed_time(e1) for e0, e1 in evs]
        km = C.c_double()
        kn = C.c_uint32()
        native.check(lib.dsm_timing_read(h, C.byref(km), C.byref(kn)))
        native.check(lib.dsm_timing_enable(h, 0))
        match_ms = km.value / max(1, kn.value)

        # ---- timed region 2: end to end through host buffers
        for _ in range(2):
            step_e2e()
            flush.zero_()
        barrier()
        e2e_s = []
        for _ in range(args.steps):
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            step_e2e()
            e2e_s.append(time.perf_counter() - t0)
            flush.zero_()
        barrier()
        clock_info = clocks.stop()

        mean_ms = statistics.mean(step_ms)
        e2e_mean = statistics.mean(e2e_s)
        e2e_p50 = statistics.median(e2e_s)
        if world > 1:
            t = torch.tensor([mean_ms, e2e_mean, e2e_p50], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            mean_ms, e2e_mean, e2e_p50 = (float(x) for x in t.cpu())
            per_rank = [torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(world)]
            dist.all_gather(per_rank, torch.tensor([match_ms], dtype=torch.float64, device=dev))
            per_rank_match_ms = [float(x.item()) for x in per_rank]
        else:
            per_rank_match_ms = [match_ms]

        # ---- parity at full size (outside the timed regions)
        sm.step()
        src, cnt, mp_ = sm.result_in_source_space()
        if world == 1:
            assert np.array_equal(b["idx"].cpu().numpy().view(np.uint32), pin_idx.array) and \
                np.array_equal(cnt, pin_cnt.array.astype(np.int64)), "host-buffer path and resident path disagree"
        verified = 0
        if rank == 0:
            verified = verify(full, sj, k, src, cnt, mp_, np.random.default_rng(0),
                              n_sample=24 if N <= (1 << 22) else 4, order=order)

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    evals = float(J) * float(N)
    value = evals / (mean_ms * 1e-3)
    peaks_src = "fallback 6650 GB/s (B200_PROFILING.md)"
    hbm_peak = 6650.0
    try:
        hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peaks_src = "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        pass
    traffic = None
    issue_fraction = None
    try:  # per-launch DRAM bytes of k_match from the committed ncu capture of this workload
        t = json.load(open(os.path.join(ROOT, "profiles", "k_match_traffic.json"))).get(wl["name"])
        if t and world == 1:
            traffic = int(t["dram_read_bytes"]) + int(t["dram_write_bytes"])
            # warp instructions of one launch (same capture) against what the
            # 4 schedulers of every SM can issue in the live launch time at the sampled clock
            sm_mhz = (clock_info or {}).get("sm_mhz")
            if t.get("warp_instructions") and sm_mhz:
                sms = torch.cuda.get_device_properties(dev).multi_processor_count
                issue_fraction = float(t["warp_instructions"]) / (match_ms * 1e-3 * sms * 4 * sm_mhz * 1e6)
    except Exception:
        pass
    groups = (J + 127) // 128
    row_bytes = table.n_cols * 2
    alg_bytes = groups * table.n_rows * row_bytes + J * rec_words * 4 + J * (k + 2) * 4
    achieved = alg_bytes / (match_ms * 1e-3) / 1e9
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": mean_ms, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
        "vs_baseline": None, "dtype": "u16 codes / u32 bit-sliced words", "data": "synthetic",
        "config": {"workload": wl["workload"], "jobs": J, "offers": N, "k": k, "jobs_per_pass": 128,
                   "row_bytes": row_bytes, "rows_per_gpu": int(table.n_rows),
                   "l2": "flushed between timed steps (256 MiB memset outside the event pairs)",
                   "parallelism": f"catalog sharded x{world} (result order dealt out in 1024-row chunks), "
                                  "one NCCL all_gather of per-shard top-k + k-way merge kernel"
                   if world > 1 else "1 GPU"},
        "e2e": {"value": evals / e2e_mean, "unit": UNIT, "h2d_bytes_per_step": int(J * rec_words * 4),
                "d2h_bytes_per_step": int(J * (k + 2) * 4 if world == 1 else J * (k * 8 + 16)),
                "p50_batch_latency_ms": 1e3 * e2e_p50,
                "path": "dsm_match_topk (C ABI, pinned host buffers)" if world == 1 else
                        "pinned H2D + dsm_match_topk_dev + NCCL all_gather + dsm_merge_shards_interleaved_dev + D2H"},
        "gpu_launches": int(launches),
        "clocks": clock_info,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "kernel": "k_match",
                     "hbm_fraction": achieved / hbm_peak, "issue_fraction": issue_fraction, "jobs_per_tile": 128,
                     "kernel_ms": match_ms, "kernel_share_of_step": match_ms / statistics.mean(step_ms),
                     "kernel_ms_per_rank": per_rank_match_ms,
                     "algorithmic_bytes_per_launch": int(alg_bytes), "peak_source": peaks_src,
                     "note": "k_match is bound by the LSU/shared-memory data pipe (ncu: 73 %), not HBM: 128 jobs "
                             "share every 16-byte row read, DRAM sees the table once (DESIGN.md section 3)"},
        "parity": {"verified_jobs_vs_oracle": int(verified), "properties_all_jobs": True},
        "setup_seconds": setup_s,
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(wl, os.cpu_count() or 1)
    print(json.dumps(out))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("for --gpus N > 1 launch with torch.distributed.run --nproc-per-node N")
    wl = workload(args.gpus, args.workload)
    if args.impl == "reference":
        run_reference(args, wl, rank)
    else:
        run_ours(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
