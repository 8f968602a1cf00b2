#!/usr/bin/env python
"""Offer-matching benchmark: job x offer evaluations per second on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload cfg1|cfg2|cfg3|cfg4|cfg5|custom:JxN|tiny] [--no-extras] [--no-cpu-baseline]

One "step" = one batch of pending jobs matched and ranked against the whole resident catalog
(table build + match + extract kernels; for N>1 also the price gather, the NCCL all-gather of the
per-shard top-k and the shard merge).

  N = 1   config 3 of BASELINE.json: 4096 jobs x 1,048,576 offers, top-16, seed 3 (the headline).
          The same JSON line also carries, unless --no-extras:
            strong_scaling_anchor  config 4 (16384 jobs x 67,108,864 offers) on this one GPU -- the
                                   N=1 point of the N=2/4/8 lines below;
            other_configs          configs 1, 2 and 5: p50 latency through the Python API and a
                                   parity verdict against the CPU oracle.
  N > 1   config 4, the catalog sharded over the N GPUs (strong scaling: total work fixed).

`value` is measured with the job records and the catalog resident in HBM (CUDA events per step,
L2 flushed between steps); `e2e` goes through the C-ABI host-buffer call: pinned host records in,
results out, copies inside the timed region; `e2e_api` is the repo's Python API (job columns in,
source positions out: record packing on the host included).  `--impl reference` times the CPU
restatement of the reference path (oracle/reference_path.py: the reference is Python that needs
packages absent here and cannot travel to the GPU box) on all host cores, on a bounded sample.
"""

from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
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
    if name in ("cfg1", "cfg2", "cfg5"):
        return dict(name=name, workload=name)
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


# ------------------------------------------------------------------------------------------ CPU baselines

_BASE = {}


def _baseline_worker(job_ids):
    from oracle import reference_path as ref

    backends, jobs = _BASE["backends"], _BASE["jobs"]
    for j in job_ids:
        profile, req = jobs[j]
        res = ref.get_offers_by_requirements(backends, profile, req)
        ref.plan_slice([o for _, o in res], _BASE["k"])
    return len(job_ids)


def _numpy_worker(job_ids):
    from oracle import columnar

    columnar.topk(_BASE["columns"], _take_jobs(_BASE["job_columns"], np.asarray(job_ids)), _BASE["k"],
                  order=_BASE["order"])
    return len(job_ids)


class CpuBaseline:
    """The object-level restatement of the reference path (a fresh CatalogItem and a `matches`
    call per (job, offer), heapq.merge, stable partition, slice) on all host cores.  The
    reference itself is Python and cannot travel to the GPU box, so this is kind "port"."""

    def __init__(self, wl, cores: int, n_offers: int = 20_000, jobs_per_core: int = 48):
        import multiprocessing as mp

        from dstack_b200.encode import encode_table
        from oracle import columnar
        from oracle import reference_path as ref
        from workloads import synth

        self.cores, self.n_offers = cores, n_offers
        cat = synth.make_catalog(n_offers, seed=wl["seed"], n_shapes=min(wl["shapes"], 1024))
        per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
        table = encode_table(cat.columns)
        self.n_jobs = cores * jobs_per_core
        sj = synth.make_jobs(cat, table, self.n_jobs, seed=wl["seed"] + 100)
        _BASE.update(backends=[b for b, _ in per_backend], jobs=synth.job_objects(sj), k=wl["k"],
                     columns=cat.columns, job_columns=sj.columns, order=columnar.rank_order(cat.columns))
        self.pool = mp.get_context("fork").Pool(cores)
        self.pool.map(_baseline_worker, [[0]] * cores)  # start and warm the workers

    def _chunks(self, jobs_per_core: int):
        n = min(self.n_jobs, jobs_per_core * self.cores)
        return [list(range(i, n, self.cores)) for i in range(self.cores)]

    def step(self, jobs_per_core: int):
        t0 = time.perf_counter()
        done = sum(self.pool.map(_baseline_worker, self._chunks(jobs_per_core)))
        dt = time.perf_counter() - t0
        return done * self.n_offers / dt, dt, done

    def step_numpy(self, jobs_per_core: int):
        """The same sample through the numpy-vectorised restatement (oracle/columnar.py): the
        "fair CPU" line of SURVEY 8(d)(c)."""
        t0 = time.perf_counter()
        done = sum(self.pool.map(_numpy_worker, self._chunks(jobs_per_core)))
        dt = time.perf_counter() - t0
        return done * self.n_offers / dt, dt, done

    def step_single(self, n_jobs: int = 4):
        """(a) of SURVEY 8(d): the verbatim loop in ONE process."""
        t0 = time.perf_counter()
        done = _baseline_worker(list(range(min(n_jobs, self.n_jobs))))
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
        out = cb.describe(value, done, dt)
        v1, dt1, d1 = cb.step_single(4)
        out["single_process"] = {"value": v1, "unit": UNIT, "cores": 1,
                                 "sample": f"{d1} jobs x {cb.n_offers} offers, {dt1:.2f} s, the same loop in one process"}
        vn, dtn, dn = cb.step_numpy(48)
        out["numpy_vectorised"] = {"value": vn, "unit": UNIT, "cores": cb.cores,
                                   "sample": f"{dn} jobs x {cb.n_offers} offers, {dtn:.2f} s on {cb.cores} processes; "
                                             f"oracle/columnar.py (one vectorised pass per job batch: a CPU "
                                             f"implementation written for speed, not the reference's loop)"}
    finally:
        cb.close()
    return out


def run_reference(args, wl, rank: int):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_core = max(2, min(48, 96 // max(1, args.steps + args.warmup)))
    if wl["name"] in ("cfg1", "cfg2", "cfg5"):
        wl = workload(1, "cfg3")
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


# ------------------------------------------------------------------------------------------ parity helpers


def _take_jobs(jc, pick):
    """Rows ``pick`` of a catalog JobColumns (CSR fields re-based)."""
    import dataclasses

    sub = {f.name: getattr(jc, f.name) for f in dataclasses.fields(jc)}

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
    return type(jc)(**sub)


def verify(cat_cols, sj, k, index, count, max_price, rng, n_sample=24, order=None):
    """Full-size parity: sampled jobs against the numpy oracle on the whole catalog, plus
    size-independent properties for every job."""
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
    done = 0
    for lo in range(0, len(pick), 32):  # the oracle materialises a [jobs, offers] matrix
        part = pick[lo:lo + 32]
        want = columnar.topk(cat_cols, _take_jobs(sj.columns, part), k, order=order)
        assert np.array_equal(want.count, count[part]), "sampled counts differ from the oracle"
        assert np.array_equal(want.index, index[part]), "sampled top-k differ from the oracle"
        assert np.array_equal(np.nan_to_num(want.max_price, nan=-1), np.nan_to_num(max_price[part], nan=-1))
        done += len(part)
    return done


def verify_shard(sm, sj, k, rng, n_sample):
    """Config 4 on N GPUs: every rank checks sampled jobs of ITS OWN shard result against the numpy
    oracle run on its shard's offers (the merge is checked on rank 0 against the gathered lists)."""
    from dstack_b200.encode import NO_ROW
    from dstack_b200.sharded import subset_columns
    from oracle import columnar

    b, t = sm._bufs, sm.table
    sm.stream.synchronize()
    idx = b["idx"].cpu().numpy().view(np.uint32)
    cnt = b["cnt"].cpu().numpy().view(np.uint32).astype(np.int64)
    local = np.where(idx == NO_ROW, -1, t.perm[np.minimum(idx, t.n_rows - 1)]).astype(np.int64)
    cols = subset_columns(sm.full, sm.positions)
    pick = rng.choice(sj.columns.n, size=min(n_sample, sj.columns.n), replace=False)
    done = 0
    for lo in range(0, len(pick), 8):
        part = pick[lo:lo + 8]
        want = columnar.topk(cols, _take_jobs(sj.columns, part), k)
        assert np.array_equal(want.count, cnt[part]), "shard counts differ from the oracle"
        assert np.array_equal(want.index, local[part]), "shard top-k differ from the oracle"
        done += len(part)
    return done


# ------------------------------------------------------------------------------------------ ours: configs 3 / 4


def _file_sha16(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()[:16]


def run_topk_workload(args, wl, rank, world, local_rank, dist, *, steps):
    """Build the (sharded) table and the job batch of a config-3/4 style workload and time it."""
    import ctypes as C

    import torch

    from dstack_b200 import native
    from dstack_b200.sharded import ShardedMatcher, rank_order
    from workloads import synth

    dev = torch.device("cuda", local_rank)
    lib = native.lib()
    J, N, k = wl["jobs"], wl["offers"], wl["k"]

    t_setup = time.perf_counter()
    cat = synth.make_catalog(N, seed=wl["seed"], n_shapes=wl["shapes"])
    full = cat.columns
    order = rank_order(full)
    sm = ShardedMatcher(full, rank=rank, world=world, device=local_rank, order=order, exchange=args.exchange)
    table = sm.table
    sj = synth.make_jobs(cat, table, J, seed=wl["seed"])  # dictionaries are shared by every shard
    t_pack = time.perf_counter()
    packed = sm.pack(sj.columns)
    pack_ms = 1e3 * (time.perf_counter() - t_pack)
    rec_words = packed.layout["rec_words"]
    setup_s = time.perf_counter() - t_setup

    stream = sm.stream
    m = sm.matcher
    h = m.handle
    top_api = None
    with torch.cuda.stream(stream):
        b = sm.prepare(packed, k)
        pin_rec = native.PinnedArray((J, rec_words), np.uint32)
        pin_rec.array[:] = packed.records
        pin_idx = native.PinnedArray((J, k), np.uint32)
        pin_cnt = native.PinnedArray((J,), np.uint32)
        pin_max = native.PinnedArray((J,), np.uint32)
        t_pin = torch.from_numpy(pin_rec.array.view(np.int32))
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        host_order = None if packed.order is None else np.ascontiguousarray(packed.order, dtype=np.uint32)
        lay_host = m._layout(packed, native.ptr(host_order))  # host-buffer call: the order is a host array
        if world > 1:
            pin_oidx = torch.empty((J, k), dtype=torch.int64).pin_memory()
            pin_ocnt = torch.empty((J,), dtype=torch.int64).pin_memory()
            pin_omp = torch.empty((J,), dtype=torch.float64).pin_memory()

        def step_e2e():
            """Host buffers in, host buffers out."""
            if world == 1:
                native.check(lib.dsm_match_topk(h, sm.device_table.handle, C.byref(lay_host), pin_rec.array.ctypes.data,
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

        # ---- timed region 1: resident inputs (no events inside the step)
        clocks = ClockSampler(local_rank)
        launches0 = m.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for e0, e1 in evs:
            e0.record(stream)
            sm.step()
            e1.record(stream)
            flush.zero_()  # evict the table from L2; outside the event pair
        barrier()
        launches = m.launch_count() - launches0
        step_ms = [e0.elapsed_time(e1) for e0, e1 in evs]

        # ---- the same steps with events around every kernel and around the exchange: where the
        # step's time goes (the events add a little, so `value` does not come from this loop)
        native.check(lib.dsm_timing_enable(h, 1))
        n_bd = max(3, min(steps, 10))
        marks = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(n_bd)]
        for mk in marks:
            sm.step(marks=mk)
            flush.zero_()
        barrier()
        parts = (C.c_double * 3)()
        kn = C.c_uint32()
        native.check(lib.dsm_timing_read_parts(h, C.byref(parts), C.byref(kn)))
        native.check(lib.dsm_timing_enable(h, 0))
        bd = {"build": parts[0] / kn.value, "match": parts[1] / kn.value, "extract": parts[2] / kn.value}
        if world > 1:
            bd["gather_price"] = statistics.mean(mk[0].elapsed_time(mk[1]) for mk in marks)
            bd["all_gather"] = statistics.mean(mk[1].elapsed_time(mk[2]) for mk in marks)
            bd["merge"] = statistics.mean(mk[2].elapsed_time(mk[3]) for mk in marks)
        match_ms = bd["match"]

        # ---- timed region 2: end to end through host buffers
        for _ in range(2):
            step_e2e()
            flush.zero_()
        barrier()
        e2e_s = []
        for _ in range(steps):
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            step_e2e()
            e2e_s.append(time.perf_counter() - t0)
            flush.zero_()
        barrier()

        # ---- timed region 3 (1 GPU): the Python API, job columns in, source positions out
        api_s = []
        if world == 1:
            for i in range(2 + min(steps, 10)):
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
                top_api = m.match_topk(sm.device_table, sj.columns, k)
                if i >= 2:
                    api_s.append(time.perf_counter() - t0)
                flush.zero_()
        clock_info = clocks.stop()

        mean_ms = statistics.mean(step_ms)
        e2e_mean = statistics.mean(e2e_s)
        e2e_p50 = statistics.median(e2e_s)
        bd_keys = sorted(bd)
        if world > 1:
            t = torch.tensor([mean_ms, e2e_mean, e2e_p50] + [bd[x] for x in bd_keys], dtype=torch.float64, device=dev)
            allr = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allr, t)
            allr = torch.stack(allr).cpu().numpy()
            mean_ms, e2e_mean, e2e_p50 = (float(x) for x in allr[:, :3].max(axis=0))
            bd_max = {x: float(allr[:, 3 + i].max()) for i, x in enumerate(bd_keys)}
            per_rank_match_ms = [float(x) for x in allr[:, 3 + bd_keys.index("match")]]
        else:
            bd_max = dict(bd)
            per_rank_match_ms = [match_ms]

        missing = sm.exchange_status()
        assert missing == 0, f"peer exchange: a wait timed out (rank {missing - 1} never arrived)"

        # ---- parity at full size (outside the timed regions)
        sm.step()
        rng = np.random.default_rng(1000 + rank)
        shard_verified = 0
        if world > 1:
            shard_verified = verify_shard(sm, sj, k, rng, n_sample=32)
            tv = torch.tensor([shard_verified], dtype=torch.int64, device=dev)
            dist.all_reduce(tv)
            shard_verified = int(tv.item())
        src, cnt, mp_ = sm.result_in_source_space()
        if world == 1:
            assert np.array_equal(b["idx"].cpu().numpy().view(np.uint32), pin_idx.array) and \
                np.array_equal(cnt, pin_cnt.array.astype(np.int64)), "host-buffer path and resident path disagree"
            assert np.array_equal(top_api.index, src) and np.array_equal(top_api.count, cnt), \
                "Python API and resident path disagree"
        verified = 0
        if rank == 0:
            n_sample = 256 if N <= (1 << 22) else (8 if world == 1 else 4)
            verified = verify(full, sj, k, src, cnt, mp_, np.random.default_rng(0), n_sample=n_sample, order=order)

    res = dict(
        wl=wl, J=J, N=N, k=k, evals=float(J) * float(N), mean_ms=mean_ms, step_ms=step_ms, e2e_mean=e2e_mean,
        e2e_p50=e2e_p50, api_s=api_s, pack_ms=pack_ms, breakdown=bd_max, match_ms=match_ms,
        per_rank_match_ms=per_rank_match_ms, launches=int(launches), clock_info=clock_info, verified=int(verified),
        shard_verified=int(shard_verified), setup_s=setup_s, rec_words=rec_words, table=table, packed=packed,
        warm=warm, steps=steps, sms=torch.cuda.get_device_properties(dev).multi_processor_count,
        exchange=sm.exchange_used, exchange_note=sm.exchange_note)
    sm.close()
    del flush
    torch.cuda.empty_cache()
    return res


def roofline_block(r, world):
    """algorithmic bytes of the dominant kernel over its live duration, against the
    measured HBM peak -- and the pipe that actually binds it."""
    peaks_src = "fallback 6650 GB/s (B200_PROFILING.md)"
    hbm_peak = 6650.0
    try:
        hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peaks_src = "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        pass
    table, J, k = r["table"], r["J"], r["k"]
    groups = (J + 127) // 128
    row_bytes = table.n_cols * 2
    alg_bytes = groups * table.n_rows * row_bytes + J * r["rec_words"] * 4 + J * (k + 2) * 4
    match_ms = r["match_ms"]
    achieved = alg_bytes / (match_ms * 1e-3) / 1e9
    out = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
           "traffic": None, "kernel": "k_match", "hbm_fraction": achieved / hbm_peak, "jobs_per_tile": 128,
           "kernel_ms": match_ms, "kernel_share_of_step": match_ms / statistics.mean(r["step_ms"]),
           "kernel_ms_per_rank": r["per_rank_match_ms"], "algorithmic_bytes_per_launch": int(alg_bytes),
           "peak_source": peaks_src,
           "binding": "lsu",
           "note": "the north star asks for the HBM fraction of the SURVEY 8(d) algorithmic bytes (every 16-byte row "
                   "counted once per group of 128 jobs); DRAM itself sees the table about once per launch (`traffic`), "
                   "the rest of those reads hit L2.  The pipe that binds k_match is the LSU / shared-memory data pipe "
                   "(`lsu_fraction`)."}
    # per-launch DRAM bytes, warp instructions and LSU wavefronts of k_match come from the committed
    # ncu capture of this workload -- valid only for the kernel source it was taken from
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "k_match_traffic.json")))
        t = cap.get(r["wl"]["name"])
        src_sha = _file_sha16(os.path.join(ROOT, "dstack_b200", "csrc", "kernels.cuh"))
        if t and world == 1:
            if cap.get("kernels_cuh_sha16") != src_sha:
                out["traffic_source"] = (f"profiles/k_match_traffic.json is STALE: captured for kernels.cuh "
                                         f"{cap.get('kernels_cuh_sha16')}, this is {src_sha}; traffic withheld")
            else:
                out["traffic"] = int(t["dram_read_bytes"]) + int(t["dram_write_bytes"])
                out["dram_fraction"] = out["traffic"] / (match_ms * 1e-3) / 1e9 / hbm_peak
                out["traffic_source"] = f"profiles/k_match_traffic.json (ncu --set full, {cap.get('captured', '?')})"
                sm_mhz = (r["clock_info"] or {}).get("sm_mhz")
                if sm_mhz:
                    cyc = match_ms * 1e-3 * sm_mhz * 1e6
                    if t.get("warp_instructions"):
                        out["issue_fraction"] = float(t["warp_instructions"]) / (cyc * r["sms"] * 4)
                    if t.get("lsu_wavefronts"):  # one wavefront per SM and clock
                        out["lsu_fraction"] = float(t["lsu_wavefronts"]) / (cyc * r["sms"])
    except Exception as e:
        out["traffic_source"] = f"no capture: {e!r}"
    return out


EXCHANGES = {
    "p2p": "dsm_xchg_all_gather_dev: own kernels, every rank stores its record into its peers' IPC-mapped buffers "
           "over NVLink and raises a flag",
    "nccl": "NCCL all_gather_into_tensor",
    "none": "no exchange",
}


def topk_line(args, r, world):
    wl, J, N, k = r["wl"], r["J"], r["N"], r["k"]
    table = r["table"]
    evals = r["evals"]
    out = {
        "metric": METRIC, "value": evals / (r["mean_ms"] * 1e-3), "unit": UNIT, "n_gpus": world, "steps": r["steps"],
        "warmup": r["warm"], "ms_per_step": r["mean_ms"], "higher_is_better": True,
        "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
        "dtype": "u16 codes / u32 bit-sliced words", "data": "synthetic",
        "config": {"workload": wl["workload"], "jobs": J, "offers": N, "k": k, "jobs_per_pass": 128,
                   "row_bytes": table.n_cols * 2, "rows_per_gpu": int(table.n_rows),
                   "columns": [c.name for c in table.columns][:r["packed"].layout["n_used"]],
                   "job_order": "grouped by the last column a job filters on" if r["packed"].order is not None else "as given",
                   "l2": "flushed between timed steps (256 MiB memset outside the event pairs)",
                   "parallelism": f"catalog sharded x{world} (result order dealt out in 1024-row chunks), "
                                  f"one all-gather of per-shard top-k ({EXCHANGES[r['exchange']]}) + k-way merge kernel"
                   if world > 1 else "1 GPU"},
        "e2e": {"value": evals / r["e2e_mean"], "unit": UNIT, "h2d_bytes_per_step": int(J * r["rec_words"] * 4),
                "d2h_bytes_per_step": int(J * (k + 2) * 4 if world == 1 else J * (k * 8 + 16)),
                "p50_batch_latency_ms": 1e3 * r["e2e_p50"],
                "path": "dsm_match_topk (C ABI, pinned host buffers)" if world == 1 else
                        f"pinned H2D + dsm_match_topk_dev + {EXCHANGES[r['exchange']]} + "
                        "dsm_merge_shards_interleaved_dev + D2H"},
        "gpu_launches": r["launches"],
        "clocks": r["clock_info"],
        "step_breakdown_ms": {**r["breakdown"],
                              "note": "CUDA-event pairs around every launch of a step (max over ranks), measured in a "
                                      "separate loop from `value`; launch gaps are the remainder"},
        "roofline": roofline_block(r, world),
        "parity": {"verified_jobs_vs_oracle": r["verified"], "properties_all_jobs": True,
                   **({"verified_shard_results_vs_oracle": r["shard_verified"]} if world > 1 else {})},
        "setup_seconds": r["setup_s"],
    }
    if world > 1:
        out["exchange"] = {"kind": r["exchange"], "what": EXCHANGES[r["exchange"]],
                           **({"fell_back_because": r["exchange_note"]} if r["exchange_note"] else {})}
        out["anchor_workload"] = "cfg4 (the N=1 line carries the same workload on one GPU as strong_scaling_anchor)"
    if r["api_s"]:
        out["e2e_api"] = {"value": evals / statistics.mean(r["api_s"]), "unit": UNIT,
                          "p50_batch_latency_ms": 1e3 * statistics.median(r["api_s"]),
                          "path": "Matcher.match_topk(JobColumns) -> TopK through dsm_match_topk_cols: job columns in, "
                                  "records + evaluation order built on the device, source positions / counts / "
                                  "max prices out",
                          "host_numpy_pack_ms_for_comparison": r["pack_ms"]}
    return out


# ------------------------------------------------------------------------------------------ ours: configs 1, 2, 5


def _p50(fn, n=15, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return 1e3 * statistics.median(ts)


def run_cfg1(device: int):
    """Config 1: one task against a 10k-offer catalog -- the full get_offers_by_requirements list
    and the first 50, against the object-level oracle (which is also timed: this config is the
    reference's own CPU-runnable case)."""
    from dstack_b200 import services as S
    from dstack_b200.engine import Matcher
    from oracle import reference_path as ref
    from workloads import configs, synth

    m = Matcher(device)
    cat = synth.make_catalog(10_000, seed=1, n_shapes=1024)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    backends = [b for b, _ in per_backend]
    profile, req = configs.config1_request()
    t0 = time.perf_counter()
    want = ref.get_offers_by_requirements(backends, profile, req)
    first, total, top_price = ref.plan_slice([o for _, o in want], 50)
    ref_ms = 1e3 * (time.perf_counter() - t0)
    catalog = S.OfferCatalog(per_backend, matcher=m)
    try:
        f = [S.job_filter(profile, req)]
        full = catalog.offers_for(f)[0]
        top = catalog.match_topk(f, 50)
        ok = ([id(o) for _, o in full] == [id(o) for _, o in want]
              and [id(o) for _, o in catalog.pairs(top.index[0])] == [id(o) for o in first]
              and int(top.count[0]) == total and top.max_price[0] == top_price)
        topk_ms = _p50(lambda: catalog.match_topk([S.job_filter(profile, req)], 50))
        full_ms = _p50(lambda: catalog.offers_for([S.job_filter(profile, req)]))
    finally:
        catalog.close()
        m.close()
    return {"workload": "cfg1: 1 task 'gpu: 24GB..80GB:1..8' x 10000 offers", "matches": total,
            "p50_ms_top50": topk_ms, "p50_ms_full_list": full_ms, "reference_cpu_ms": ref_ms,
            "parity": "bit-exact (same objects, same order, total, max price)" if ok else "MISMATCH",
            "path": "OfferCatalog.match_topk / offers_for (job filter -> records -> GPU -> objects), table resident"}


def run_cfg2(device: int):
    """Config 2: 1000 pending jobs x 50000 pool rows through PoolIndex (one batched device call)."""
    from dstack_b200 import pools as PP
    from dstack_b200.engine import Matcher
    from dstack_b200.ingest import PoolParseCache
    from oracle import reference_path as ref
    from workloads import configs, synth

    m = Matcher(device)
    rows = synth.make_pool(50_000, seed=2, n_fleets=200)
    calls = configs.config2_jobs(rows, 1000)
    cache = PoolParseCache()
    t0 = time.perf_counter()
    index = PP.PoolIndex(rows, matcher=m, parse_cache=cache)
    cold_s = time.perf_counter() - t0
    index.close()
    t0 = time.perf_counter()
    index = PP.PoolIndex(rows, matcher=m, parse_cache=cache)
    warm_s = time.perf_counter() - t0
    try:
        filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"],
                                   fleet_model=kw.get("fleet_model"), multinode=kw.get("multinode", False),
                                   master_job_provisioning_data=kw.get("master_job_provisioning_data"),
                                   volumes=None, shared=False) for p, kw in calls]
        got = index.match(filters)
        match_ms = _p50(lambda: index.match(filters), n=7, warm=2)
    finally:
        index.close()
        m.close()
    ok = True
    sample = [0, 1, 2] + [j for j, (p, kw) in enumerate(calls) if "fleet_model" in kw][:1] + \
        [j for j, (p, kw) in enumerate(calls) if kw.get("multinode")][:1]
    t0 = time.perf_counter()
    for j in sample:
        want = ref.filter_pool_instances(rows, calls[j][0], **calls[j][1])
        ok = ok and len(want) == len(got[j]) and all(rows[i] is w for i, w in zip(got[j], want))
    ref_ms_per_job = 1e3 * (time.perf_counter() - t0) / len(sample)
    return {"workload": "cfg2: 1000 jobs x 50000 pool rows (30 % same-fleet, 20 % multinode, 10 % reuse)",
            "p50_ms_batched_match": match_ms, "evals_per_s": 1000 * 50000 / (match_ms * 1e-3),
            "table_build_s": {"cold_json_parse": cold_s, "warm_parse_cache": warm_s},
            "hits": int(sum(len(g) for g in got)), "reference_cpu_ms_per_job": ref_ms_per_job,
            "parity": f"bit-exact on {len(sample)} sampled jobs (the oracle parses 50000 JSON documents per job)"
            if ok else "MISMATCH",
            "path": "PoolIndex.match (filters -> records -> dsm_match_bitmap -> per-job row lists)"}


def run_cfg5(device: int, n_offers: int = 10_000_000):
    """Config 5: 8 multinode tasks of 8 nodes `gpu: H100:8` against 10 M offers: masters in the
    first tick, followers pinned to their master's backend and region in the second
    (scheduler.process_submitted_batch); parity on the masters' offers and on a follower's filter
    against the numpy oracle."""
    from dstack_b200 import scheduler
    from dstack_b200 import services as S
    from dstack_b200.engine import Matcher
    from dstack_b200.ingest import job_columns
    from oracle import columnar
    from workloads import configs, synth

    m = Matcher(device)
    cat = synth.make_catalog(n_offers, seed=5, n_shapes=4096)
    backends, materialize = synth.lazy_catalog(cat)
    t0 = time.perf_counter()
    catalog = S.OfferCatalog.from_columns(cat.columns, materialize, matcher=m)
    build_s = time.perf_counter() - t0
    jobs = configs.config5_jobs()
    launch = configs.fake_launch(fail_every=4)
    try:
        outcomes = scheduler.process_submitted_batch([], catalog, jobs, launch, matcher=m)
        pass_ms = _p50(lambda: scheduler.process_submitted_batch([], catalog, jobs, launch, matcher=m), n=7, warm=2)
        masters = [j for j in jobs if j.job_num == 0]
        mf = [S.job_filter(j.profile, j.requirements, exclude_not_available=True, multinode=True) for j in masters]
        top = catalog.match_topk(mf, 16)
        match_ms = _p50(lambda: catalog.match_topk(mf, 16), n=7, warm=2)
        # parity: a master's top-16 and the followers' filter of cluster 0 against the numpy oracle
        jpd = outcomes[0].job_provisioning_data
        ff = [S.job_filter(jobs[1].profile, jobs[1].requirements, exclude_not_available=True, multinode=True,
                           master_job_provisioning_data=jpd)] if jpd is not None else []
        filters = mf[:1] + ff
        got = catalog.match_topk(filters, 16)
        want = columnar.topk(cat.columns, job_columns(catalog.table, filters), 16)
        ok = np.array_equal(got.index, want.index) and np.array_equal(got.count, want.count)
        placed = sum(o.status == "provisioning" for o in outcomes)
        pinned = all(o.offer.backend == outcomes[(i // 8) * 8].offer.backend and
                     o.offer.region == outcomes[(i // 8) * 8].offer.region
                     for i, o in enumerate(outcomes)
                     if o.status == "provisioning" and outcomes[(i // 8) * 8].status == "provisioning")
        groups = scheduler.placement_group_keys([o.offer for o in outcomes if o.status == "provisioning"])
    finally:
        catalog.close()
        m.close()
    return {"workload": f"cfg5: 8 multinode tasks x 8 nodes 'gpu: H100:8' x {n_offers} offers, top-16, two ticks",
            "p50_ms_whole_pass": pass_ms, "p50_ms_masters_top16": match_ms, "nodes_placed": int(placed),
            "master_matches": int(top.count[0]), "followers_in_master_region": bool(pinned),
            "placement_groups": [f"{b.value}/{r}" for b, r in groups], "table_build_s": build_s,
            "parity": "bit-exact (a master's and a follower's top-16 + counts vs the numpy oracle on all rows)"
            if ok else "MISMATCH",
            "path": "scheduler.process_submitted_batch over OfferCatalog.from_columns (two batched catalog matches)"}


# ------------------------------------------------------------------------------------------ main


def run_ours(args, wl, rank: int, world: int, local_rank: int):
    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if wl["name"] in ("cfg1", "cfg2", "cfg5"):
        if rank == 0:
            fn = {"cfg1": run_cfg1, "cfg2": run_cfg2, "cfg5": run_cfg5}[wl["name"]]
            print(json.dumps({"metric": METRIC, "workload": wl["name"], **fn(local_rank)}))
        return

    r = run_topk_workload(args, wl, rank, world, local_rank, dist, steps=args.steps)
    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return
    out = topk_line(args, r, world)
    if world == 1 and wl["name"] == "cfg3" and not args.no_extras:
        # the N=1 point of the strong-scaling curve: config 4 on this one GPU
        t0 = time.perf_counter()
        a = run_topk_workload(args, workload(2, "cfg4"), 0, 1, local_rank, None, steps=max(3, min(args.steps, 5)))
        al = topk_line(args, a, 1)
        out["strong_scaling_anchor"] = {
            "workload": al["config"]["workload"], "n_gpus": 1, "value": al["value"], "ms_per_step": al["ms_per_step"],
            "steps": a["steps"], "kernel_ms": a["match_ms"], "step_breakdown_ms": al["step_breakdown_ms"],
            "e2e": al["e2e"], "clocks": al["clocks"], "roofline_hbm_fraction": al["roofline"]["hbm_fraction"],
            "parity": al["parity"], "wall_seconds": time.perf_counter() - t0,
            "note": "divide the N=2/4/8 `value`s by N times this one for the strong-scaling efficiency"}
        extras = {}
        for name, fn in (("cfg1", run_cfg1), ("cfg2", run_cfg2), ("cfg5", run_cfg5)):
            t0 = time.perf_counter()
            try:
                extras[name] = fn(local_rank)
            except Exception as e:  # an extra must not take the headline down with it
                extras[name] = {"error": repr(e)}
            extras[name]["wall_seconds"] = time.perf_counter() - t0
        out["other_configs"] = extras
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(r["wl"] if r["wl"]["name"] != "tiny" else workload(1, "cfg3"), os.cpu_count() or 1)
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
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N>1: the step's all-gather (own peer-store kernels, or NCCL)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="N=1: skip the config-4 anchor and configs 1, 2, 5")
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
