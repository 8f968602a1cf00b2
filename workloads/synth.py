"""Seeded synthetic catalogs, pools and job batches.

Everything is generated in columnar form first (so 64M-offer catalogs are feasible) and can
be materialised as Python objects for the object-level oracle at small sizes.  Per-backend
lists are price-sorted and contiguous in project-backend order, as the reference's merge
requires (``server/services/backends/__init__.py:449``); prices are rounded to 4 decimals so
ties are frequent and the stable tie-break is exercised.
"""

from __future__ import annotations

import json
import math
import types
import uuid
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from dstack_b200 import models as M
from dstack_b200._gpuhunt import QueryFilter
from dstack_b200.encode import BACKEND_CODES, VENDOR_CODES, JobColumns, OfferColumns, ShapeTable
from dstack_b200.ingest import JobFilter

# provider weights of App. D, in project-backend order
PROVIDERS: List[Tuple[str, float]] = [
    ("aws", .30), ("gcp", .20), ("azure", .20), ("oci", .05), ("runpod", .05), ("vastai", .05),
    ("lambda", .02), ("tensordock", .03), ("datacrunch", .03), ("cudo", .03), ("nebius", .02), ("vultr", .02),
]

# (vendor, name as the provider spells it, memory GiB)
ACCELERATORS: List[Tuple[str, str, float]] = [
    ("nvidia", "T4", 16), ("nvidia", "L4", 24), ("nvidia", "A10G", 24), ("nvidia", "A10", 24),
    ("nvidia", "A100", 40), ("nvidia", "A100", 80), ("nvidia", "H100", 80), ("nvidia", "H200", 141),
    ("nvidia", "V100", 16), ("nvidia", "V100", 32), ("nvidia", "L40S", 48), ("nvidia", "RTX4090", 24),
    ("nvidia", "A6000", 48), ("nvidia", "A5000", 24), ("nvidia", "P100", 16), ("nvidia", "a100", 80),
    ("nvidia", "h100", 80), ("nvidia", "L40", 48), ("nvidia", "A40", 48), ("nvidia", "RTX3090", 24),
    ("nvidia", "A4000", 16), ("nvidia", "RTX6000Ada", 48), ("nvidia", "B200", 180), ("nvidia", "H100NVL", 94),
    ("nvidia", "RTX4080", 16), ("nvidia", "A30", 24), ("nvidia", "K80", 12), ("nvidia", "GH200", 96),
    ("amd", "MI300X", 192), ("amd", "MI250", 128), ("google", "v5litepod-8", 0), ("google", "v3-8", 0),
]

CPU_CHOICES = np.array([1, 2, 4, 8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 224])
FRACTIONAL_MEMORY = np.array([15.25, 30.5, 61.0, 3.75, 7.5, 0.6, 1.7])
DISK_CHOICES = np.array([100.0, 200.0, 500.0, 1000.0, 2000.0])
GPU_COUNT_CHOICES = np.array([1, 2, 4, 8])
AVAILABILITY = [M.InstanceAvailability.UNKNOWN, M.InstanceAvailability.AVAILABLE,
                M.InstanceAvailability.NO_QUOTA, M.InstanceAvailability.NOT_AVAILABLE]
AVAILABILITY_P = [.60, .25, .10, .05]


@dataclass
class SynthCatalog:
    columns: OfferColumns
    availability: np.ndarray          # int8 index into AVAILABILITY
    shape_names: List[str]            # instance type name per shape
    providers: List[str]              # project-backend order
    backend_slices: List[Tuple[int, int]]  # [start, end) of each provider's list


def make_shapes(rng: np.random.Generator, n_shapes: int) -> Tuple[ShapeTable, List[str], np.ndarray]:
    cpus = CPU_CHOICES[rng.integers(0, len(CPU_CHOICES), n_shapes)].astype(np.int64)
    mem = cpus * np.array([2.0, 4.0, 8.0])[rng.integers(0, 3, n_shapes)]
    frac = rng.random(n_shapes) < 0.06
    mem = np.where(frac, FRACTIONAL_MEMORY[rng.integers(0, len(FRACTIONAL_MEMORY), n_shapes)], mem)
    # memory as the offers carry it: whole MiB, then / 1024 (offer_to_catalog_item)
    mem_mib = np.round(mem * 1024).astype(np.int64)
    has_gpu = rng.random(n_shapes) >= 0.40
    gcount = np.where(has_gpu, GPU_COUNT_CHOICES[rng.integers(0, 4, n_shapes)], 0).astype(np.int64)
    acc = rng.integers(0, len(ACCELERATORS), n_shapes)
    vendor = np.array([VENDOR_CODES[ACCELERATORS[a][0]] for a in acc], dtype=np.int8)
    gname: List[Optional[str]] = [ACCELERATORS[a][1] if g else None for a, g in zip(acc, has_gpu)]
    gmem = np.array([ACCELERATORS[a][2] for a in acc], dtype=np.float64)
    disk_known = rng.random(n_shapes) >= 0.70
    disk = np.where(disk_known, DISK_CHOICES[rng.integers(0, len(DISK_CHOICES), n_shapes)], 100.0)
    base_price = (0.012 * cpus + 0.004 * mem + gcount * (0.25 + gmem / 40.0)) * rng.uniform(0.8, 1.25, n_shapes)
    shapes = ShapeTable(
        cpus=cpus, memory_gib=mem_mib / 1024, disk_gib=disk.astype(np.float64), gpu_count=gcount,
        gpu_vendor=np.where(has_gpu, vendor, 0).astype(np.int8), gpu_name=gname,
        gpu_memory_gib=np.where(has_gpu, gmem, np.nan),
    )
    names = [f"shape-{i:04d}" for i in range(n_shapes)]
    return shapes, names, base_price


def make_catalog(n_offers: int, seed: int = 1, n_shapes: int = 4096, n_providers: int = 12) -> SynthCatalog:
    rng = np.random.Generator(np.random.PCG64(seed))
    n_shapes = max(1, min(n_shapes, max(1, n_offers)))
    shapes, shape_names, base_price = make_shapes(rng, n_shapes)
    provs = PROVIDERS[:n_providers]
    w = np.array([p[1] for p in provs])
    w = w / w.sum()
    counts = np.floor(w * n_offers).astype(np.int64)
    counts[0] += n_offers - counts.sum()
    regions: List[str] = []
    cols: Dict[str, List[np.ndarray]] = {k: [] for k in
                                         ("shape", "price", "spot", "avail", "backend", "region")}
    slices = []
    start = 0
    for (pname, _), cnt in zip(provs, counts):
        cnt = int(cnt)
        n_reg = int(rng.integers(8, 41))
        reg_base = len(regions)
        regions.extend(f"{pname}-region-{r}" for r in range(n_reg))
        zipf = 1.0 / np.arange(1, n_reg + 1)
        reg = rng.choice(n_reg, size=cnt, p=zipf / zipf.sum()) + reg_base
        factor = 1.0 + 0.03 * (reg - reg_base)
        shape = rng.integers(0, n_shapes, cnt)
        spot = rng.random(cnt) < 0.5
        price = base_price[shape] * factor * np.where(spot, rng.uniform(0.3, 0.7, cnt), 1.0)
        price = np.maximum(np.round(price, 4), 0.0001)
        avail = rng.choice(len(AVAILABILITY), size=cnt, p=AVAILABILITY_P).astype(np.int8)
        order = np.argsort(price, kind="stable")  # the provider's own price-sorted list
        for key, arr in (("shape", shape), ("price", price), ("spot", spot), ("avail", avail), ("region", reg)):
            cols[key].append(arr[order])
        cols["backend"].append(np.full(cnt, BACKEND_CODES[pname], dtype=np.int8))
        slices.append((start, start + cnt))
        start += cnt
    cat = {k: np.concatenate(v) if v else np.zeros(0) for k, v in cols.items()}
    avail_idx = cat["avail"].astype(np.int8)
    is_avail = np.array([a.is_available() for a in AVAILABILITY])[avail_idx] if n_offers else np.zeros(0, bool)
    oc = OfferColumns(
        shapes=shapes, shape_id=cat["shape"].astype(np.int32), price=cat["price"].astype(np.float64),
        spot=cat["spot"].astype(bool), available=is_avail, backend=cat["backend"].astype(np.int8),
        src_backend=cat["backend"].astype(np.int8).copy(), region_id=cat["region"].astype(np.int32),
        regions=regions, itype_id=cat["shape"].astype(np.int32), itypes=shape_names,
    )
    return SynthCatalog(oc, avail_idx, shape_names, [p[0] for p in provs], slices)


class FakeCompute:
    def __init__(self, offers: List[Any], filter_fn=None):
        self._offers, self._filter = offers, filter_fn

    def get_offers_cached(self, requirements: Any) -> List[Any]:
        if self._filter is None or requirements is None:
            return list(self._offers)
        return self._filter(self._offers, requirements)


class FakeBackend:
    """Object with ``TYPE`` and ``compute()`` like the reference's ``Backend``; tests of the
    reference mock exactly this surface (``test_process_submitted_jobs.py:126-148``)."""

    def __init__(self, backend_type: M.BackendType, offers: List[Any], filter_fn=None):
        self.TYPE = backend_type
        self._compute = FakeCompute(offers, filter_fn)

    def compute(self) -> FakeCompute:
        return self._compute

    def __repr__(self) -> str:
        return f"FakeBackend({self.TYPE.value})"


_VENDOR_OF = {v: k for k, v in VENDOR_CODES.items()}


def offer_object(cat: SynthCatalog, i: int) -> M.InstanceOfferWithAvailability:
    """Offer ``i`` (source position) of a synthetic catalog as an object."""
    oc, sh = cat.columns, cat.columns.shapes
    sid = int(oc.shape_id[i])
    n_gpu = int(sh.gpu_count[sid])
    gpus = []
    if n_gpu:
        gpus = [M.Gpu(name=sh.gpu_name[sid], vendor=_VENDOR_OF[int(sh.gpu_vendor[sid])],
                      memory_mib=int(round(sh.gpu_memory_gib[sid] * 1024)))] * n_gpu
    res = M.Resources(cpus=int(sh.cpus[sid]), memory_mib=int(round(sh.memory_gib[sid] * 1024)),
                      gpus=gpus, spot=bool(oc.spot[i]),
                      disk=M.Disk(size_mib=int(round(sh.disk_gib[sid] * 1024))))
    pname = cat.providers[provider_of(cat, i)]
    return M.InstanceOfferWithAvailability(
        backend=M.BackendType(pname), region=oc.regions[int(oc.region_id[i])], price=float(oc.price[i]),
        instance=M.InstanceType(name=oc.itypes[int(oc.itype_id[i])], resources=res),
        availability=AVAILABILITY[int(cat.availability[i])])


def provider_of(cat: SynthCatalog, i: int) -> int:
    """Index of the provider whose price-sorted list holds source position ``i``."""
    for p, (s, e) in enumerate(cat.backend_slices):
        if s <= i < e:
            return p
    raise IndexError(i)


def catalog_objects(cat: SynthCatalog, filter_fn=None) -> List[Tuple[FakeBackend, List[M.InstanceOfferWithAvailability]]]:
    """Materialise a (small) synthetic catalog as per-backend offer lists."""
    out = []
    for pname, (s, e) in zip(cat.providers, cat.backend_slices):
        offers = [offer_object(cat, i) for i in range(s, e)]
        out.append((FakeBackend(M.BackendType(pname), offers, filter_fn), offers))
    return out


def lazy_catalog(cat: SynthCatalog):
    """``(backends, materialize)`` for :meth:`dstack_b200.services.OfferCatalog.from_columns`: one
    :class:`FakeBackend` per provider (no offer objects) and ``materialize(i) -> (backend, offer)``."""
    backends = [FakeBackend(M.BackendType(p), []) for p in cat.providers]
    return backends, (lambda i: (backends[provider_of(cat, i)], offer_object(cat, i)))


# ---------------------------------------------------------------------------- jobs


@dataclass
class SynthJobs:
    """A job batch in columnar form plus what is needed to rebuild the same jobs as objects."""

    columns: JobColumns
    gpu_names: List[List[str]]                 # [] = no name filter
    backend_lists: List[Optional[List[str]]]
    region_lists: List[Optional[List[str]]]


def make_jobs(cat: SynthCatalog, table, n_jobs: int, seed: int = 3) -> SynthJobs:
    """Config-3 distribution: GPU-name sets of size 0 / 1 / 2-4 with
    p = .4/.3/.3 from the accelerator dictionary, ``max_price`` None 50% else lognormal(ln 4, 1),
    backend filter None 60% else 1-3 providers, region filter None 70% else 1-4 regions, spot
    None/True/False 40/20/40%.  ``table`` is the EncodedTable the ids refer to."""
    rng = np.random.Generator(np.random.PCG64(seed))
    j = n_jobs
    nan2 = lambda: np.full((j, 2), np.nan)
    cpu, mem, disk, gcnt, gmem, gtot = nan2(), nan2(), nan2(), nan2(), nan2(), nan2()
    cpu[:, 0] = rng.choice([1, 2, 2, 2, 4, 8, 16, 32], j)
    mem[:, 0] = rng.choice([4.0, 8.0, 8.0, 8.0, 16.0, 32.0, 64.0, 15.25], j)
    disk[:, 0] = rng.choice([100.0, 100.0, 100.0, 50.0, 200.0, 500.0], j)
    wants_gpu = rng.random(j) < 0.85
    count_kind = rng.integers(0, 4, j)
    lo = np.select([count_kind == 0, count_kind == 1, count_kind == 2], [1, 1, 2], 4).astype(float)
    hi = np.select([count_kind == 0, count_kind == 1, count_kind == 2], [1, 8, 4], 8).astype(float)
    gcnt[:, 0] = np.where(wants_gpu, lo, np.nan)
    gcnt[:, 1] = np.where(wants_gpu, hi, np.nan)
    mem_kind = rng.random(j)
    gmem[:, 0] = np.where(wants_gpu & (mem_kind < 0.35), rng.choice([16.0, 24.0, 40.0, 80.0], j), np.nan)
    gmem[:, 1] = np.where(wants_gpu & (mem_kind < 0.15), 80.0, np.nan)
    gtot[:, 0] = np.where(wants_gpu & (rng.random(j) < 0.08), rng.choice([48.0, 160.0, 320.0], j), np.nan)
    cc = np.where(wants_gpu & (rng.random(j) < 0.10), rng.choice([7000, 7005, 8000, 8006, 9000], j), -1)
    vendor = np.where(wants_gpu & (rng.random(j) < 0.10), rng.choice([1, 1, 2, 3], j), 0).astype(np.int8)

    names_lower = table.gpu_dict["names"]
    all_names = sorted({a[1] for a in ACCELERATORS})
    size_kind = rng.random(j)
    gpu_names: List[List[str]] = []
    name_any = np.ones(j, bool)
    n_ptr, n_ids = [0], []
    for i in range(j):
        if wants_gpu[i] and size_kind[i] >= 0.4:
            k = 1 if size_kind[i] < 0.7 else int(rng.integers(2, 5))
            pick = [all_names[x] for x in rng.choice(len(all_names), size=k, replace=False)]
            # requests are case-insensitive: flip the case of some
            pick = [p.lower() if rng.random() < 0.3 else p for p in pick]
            gpu_names.append(pick)
            name_any[i] = False
            n_ids.extend(names_lower.get(p.lower(), -1) for p in pick)
        else:
            gpu_names.append([])
        n_ptr.append(len(n_ids))

    max_price = np.where(rng.random(j) < 0.5, np.nan, np.round(rng.lognormal(math.log(4.0), 1.0, j), 2))
    spot = rng.choice([-1, 1, 0], size=j, p=[.4, .2, .4]).astype(np.int8)

    reg_lookup = table.columns[table.col("region")].lookup
    backend_lists: List[Optional[List[str]]] = []
    region_lists: List[Optional[List[str]]] = []
    src_mask = np.full(j, 0xFFFFFFFF, np.uint32)
    off_mask = np.full(j, 0xFFFFFFFF, np.uint32)
    r_any = np.ones(j, bool)
    r_ptr, r_ids = [0], []
    for i in range(j):
        if rng.random() < 0.4:
            k = int(rng.integers(1, 4))
            bl = [cat.providers[x] for x in rng.choice(len(cat.providers), size=min(k, len(cat.providers)), replace=False)]
            backend_lists.append(bl)
            m = 0
            for b in bl:
                m |= 1 << BACKEND_CODES[b]
            off_mask[i] = m
            src_mask[i] = m | (1 << BACKEND_CODES["dstack"])
        else:
            backend_lists.append(None)
        if rng.random() < 0.3:
            k = int(rng.integers(1, 5))
            rl = [cat.columns.regions[x] for x in rng.choice(len(cat.columns.regions), size=k, replace=False)]
            region_lists.append(rl)
            r_any[i] = False
            r_ids.extend(reg_lookup.get(r, -1) for r in rl)
        else:
            region_lists.append(None)
        r_ptr.append(len(r_ids))

    jc = JobColumns(
        n=j, max_price=max_price, spot=spot, cpu=cpu, memory=mem, disk=disk, gpu_count=gcnt, gpu_memory=gmem,
        gpu_total=gtot, gpu_vendor=vendor, cc_min=cc.astype(np.int64), gpu_name_any=name_any,
        gpu_name_ptr=np.array(n_ptr, np.int64), gpu_name_ids=np.array(n_ids, np.int64),
        src_backend_mask=src_mask, off_backend_mask=off_mask, exclude_not_available=np.zeros(j, bool),
        region_any=r_any, region_ptr=np.array(r_ptr, np.int64), region_ids=np.array(r_ids, np.int64),
        itype_any=np.ones(j, bool), itype_ptr=np.zeros(j + 1, np.int64), itype_ids=np.zeros(0, np.int64),
    )
    return SynthJobs(jc, gpu_names, backend_lists, region_lists)


def widen_jobs(sj: SynthJobs, table, seed: int = 0) -> SynthJobs:
    """Perturb a job batch in place so that every kind of bound occurs: maxima as well as minima,
    bounds exactly equal to values of the table's dictionaries, bounds outside them, no bounds at
    all, explicit block counts, ``exclude_not_available``.  For tests of the record packers."""
    rng = np.random.Generator(np.random.PCG64(seed))
    jc = sj.columns
    j = jc.n
    gd = table.gpu_dict
    lists = {"cpu": table.cpu_values, "memory": table.memory_values,
             "disk": table.columns[table.col("disk")].values, "gpu_count": gd["count"], "gpu_memory": gd["memory"],
             "gpu_total": gd["total"]}
    for name, vals in lists.items():
        b = getattr(jc, name)
        vals = np.asarray(vals, dtype=np.float64)
        if len(vals) == 0:
            continue
        for side in (0, 1):
            u = rng.random(j)
            exact = vals[rng.integers(0, len(vals), j)]
            b[:, side] = np.where(u < 0.15, exact, b[:, side])                       # exactly a table value
            b[:, side] = np.where((u >= 0.15) & (u < 0.20), exact + 1e-9, b[:, side])  # just above one
            b[:, side] = np.where((u >= 0.20) & (u < 0.25), exact - 1e-9, b[:, side])  # just below one
            b[:, side] = np.where((u >= 0.25) & (u < 0.30), np.nan, b[:, side])       # not set
            b[:, side] = np.where((u >= 0.30) & (u < 0.32), vals.max() * 10 + 1, b[:, side])
            b[:, side] = np.where((u >= 0.32) & (u < 0.34), -1.0, b[:, side])
    prices = np.concatenate([table.price_a, table.price_b])
    if len(prices):
        u = rng.random(j)
        jc.max_price[:] = np.where(u < 0.2, prices[rng.integers(0, len(prices), j)], jc.max_price)
        jc.max_price[:] = np.where((u >= 0.2) & (u < 0.25), 0.0, jc.max_price)
    jc.exclude_not_available[:] = rng.random(j) < 0.3
    return sj


def job_objects(sj: SynthJobs) -> List[Tuple[M.Profile, M.Requirements]]:
    """The same batch as ``(Profile, Requirements)`` pairs for the object-level oracle."""
    jc = sj.columns
    opt = lambda v: None if np.isnan(v) else float(v)
    vendor_of = {v: k for k, v in VENDOR_CODES.items()}
    out = []
    for i in range(jc.n):
        gpu = None
        if not np.isnan(jc.gpu_count[i, 0]):
            spec: Dict[str, Any] = {"count": {"min": int(jc.gpu_count[i, 0]), "max": int(jc.gpu_count[i, 1])}}
            if not np.isnan(jc.gpu_memory[i]).all():
                spec["memory"] = {"min": opt(jc.gpu_memory[i, 0]), "max": opt(jc.gpu_memory[i, 1])}
            if not np.isnan(jc.gpu_total[i]).all():
                spec["total_memory"] = {"min": opt(jc.gpu_total[i, 0]), "max": opt(jc.gpu_total[i, 1])}
            if jc.cc_min[i] >= 0:
                spec["compute_capability"] = (int(jc.cc_min[i]) // 1000, int(jc.cc_min[i]) % 1000)
            if jc.gpu_vendor[i]:
                spec["vendor"] = vendor_of[int(jc.gpu_vendor[i])]
            if sj.gpu_names[i]:
                spec["name"] = list(sj.gpu_names[i])
            gpu = M.GPUSpec.parse_obj(spec)
        res = M.ResourcesSpec(cpu={"min": int(jc.cpu[i, 0]), "max": None}, memory={"min": float(jc.memory[i, 0]), "max": None},
                              disk={"size": {"min": float(jc.disk[i, 0]), "max": None}}, gpu=gpu)
        req = M.Requirements(resources=res, max_price=opt(jc.max_price[i]),
                             spot=None if jc.spot[i] < 0 else bool(jc.spot[i]), reservation=None)
        prof = M.Profile(name="p", backends=sj.backend_lists[i], regions=sj.region_lists[i])
        out.append((prof, req))
    return out


# ---------------------------------------------------------------------------- pool rows (config 2)


def make_pool(n_instances: int, seed: int = 2, n_fleets: int = 200) -> List[Any]:
    """Config-2 pool: fleets x instances; status IDLE 70 / BUSY 20 / other 10 %; unreachable 2 %;
    total_blocks 1 (80 %), 2/4/8 (18 %), None (2 %); busy_blocks uniform < total."""
    rng = np.random.Generator(np.random.PCG64(seed))
    shapes, shape_names, base_price = make_shapes(rng, min(512, max(1, n_instances)))
    vendor_of = {v: k for k, v in VENDOR_CODES.items()}
    fleets = [uuid.UUID(int=int(x)) for x in rng.integers(1, 2 ** 62, n_fleets)]
    provs = ["aws", "gcp", "azure", "oci", "runpod", "vastai"]
    rows = []
    for i in range(n_instances):
        sid = int(rng.integers(0, len(shapes)))
        n_gpu = int(shapes.gpu_count[sid])
        tb_draw = rng.random()
        total_blocks: Optional[int] = 1
        if tb_draw < 0.02:
            total_blocks = None
        elif tb_draw < 0.20:
            total_blocks = int(rng.choice([2, 4, 8]))
        st_draw = rng.random()
        other = [M.InstanceStatus.PENDING, M.InstanceStatus.PROVISIONING,
                 M.InstanceStatus.TERMINATING, M.InstanceStatus.TERMINATED]
        status = (M.InstanceStatus.IDLE if st_draw < 0.7 else M.InstanceStatus.BUSY if st_draw < 0.9
                  else other[int(rng.integers(0, len(other)))])
        backend = M.BackendType(provs[int(rng.integers(0, len(provs)))])
        region = f"{backend.value}-region-{int(rng.integers(0, 6))}"
        gpus = []
        if n_gpu:
            gpus = [M.Gpu(name=shapes.gpu_name[sid], vendor=vendor_of[int(shapes.gpu_vendor[sid])],
                          memory_mib=int(round(shapes.gpu_memory_gib[sid] * 1024)))] * n_gpu
        res = M.Resources(cpus=int(shapes.cpus[sid]), memory_mib=int(round(shapes.memory_gib[sid] * 1024)),
                          gpus=gpus, spot=bool(rng.random() < 0.3),
                          disk=M.Disk(size_mib=int(round(shapes.disk_gib[sid] * 1024))))
        price = float(np.round(base_price[sid] * rng.uniform(0.5, 1.5), 4))
        itype = M.InstanceType(name=shape_names[sid], resources=res)
        offer = M.InstanceOfferWithAvailability(backend=backend, instance=itype, region=region, price=price,
                                                availability=M.InstanceAvailability.AVAILABLE,
                                                total_blocks=total_blocks or 1)
        has_offer = rng.random() >= 0.01
        has_jpd = rng.random() >= 0.05
        jpd = M.JobProvisioningData(backend=backend, instance_type=itype, instance_id=f"i-{i}", region=region,
                                    availability_zone=f"{region}-{'abc'[int(rng.integers(0, 3))]}" if rng.random() < 0.8 else None,
                                    price=price, username="ubuntu", dockerized=True)
        rows.append(types.SimpleNamespace(
            id=uuid.UUID(int=i + 1), name=f"instance-{i}", fleet_id=fleets[int(rng.integers(0, n_fleets))] if rng.random() < 0.95 else None,
            status=M.InstanceStatus(status), unreachable=bool(rng.random() < 0.02), backend=backend, region=region,
            price=price, offer=offer.json() if has_offer else None,
            job_provisioning_data=jpd.json() if has_jpd else None,
            total_blocks=total_blocks,
            busy_blocks=int(rng.integers(0, total_blocks)) if total_blocks else 0,
        ))
    return rows
