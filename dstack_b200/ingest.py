"""Python objects -> columnar inputs of the encoder.

* offers (``InstanceOffer`` / ``InstanceOfferWithAvailability``, duck-typed) -> ``OfferColumns``
  -- the one-time replacement of the reference's per-(job, offer) ``offer_to_catalog_item``
  walk (``core/backends/base/offers.py:86-109``);
* pool rows (``InstanceModel``-like) -> ``OfferColumns`` with the admin attributes
  ``filter_pool_instances`` tests (``server/services/pools.py:457-488``); the stored offer
  JSON is parsed ONCE per row here instead of once per (job, row) (``pools.py:499``);
* job filters (``JobFilter``) -> ``JobColumns``.
"""

from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

from . import models as M
from ._gpuhunt import QueryFilter
from .encode import (BACKEND_CODES, FLAG_AUTO_DIVISIBLE, FLAG_BLOCKS_UNKNOWN, FLAG_NO_OFFER, FLAG_SHARED,
                     FLAG_UNREACHABLE, NO_BACKEND_CODE, NO_STATUS_CODE, STATUS_CODES, EncodedTable,
                     JobColumns, OfferColumns, ShapeTable, backend_code, vendor_code)


class _Interner:
    def __init__(self):
        self.ids: Dict[Any, int] = {}
        self.values: List[Any] = []

    def __call__(self, v: Any) -> int:
        i = self.ids.get(v)
        if i is None:
            i = self.ids[v] = len(self.values)
            self.values.append(v)
        return i


class _Shapes:
    """Dedup of resource shapes: (cpus, memory_mib, disk_mib, gpu count, vendor, name, gpu mib)."""

    def __init__(self):
        self.intern = _Interner()

    def add(self, cpus, memory_gib, disk_gib, gpu_count, vendor, name, gpu_gib, disk_conf=-1) -> int:
        return self.intern((cpus, memory_gib, disk_gib, gpu_count, vendor, name, gpu_gib, disk_conf))

    def table(self) -> ShapeTable:
        v = self.intern.values
        return ShapeTable(
            cpus=np.array([s[0] for s in v], dtype=np.int64),
            memory_gib=np.array([s[1] for s in v], dtype=np.float64),
            disk_gib=np.array([math.nan if s[2] is None else s[2] for s in v], dtype=np.float64),
            gpu_count=np.array([s[3] for s in v], dtype=np.int64),
            gpu_vendor=np.array([s[4] for s in v], dtype=np.int8),
            gpu_name=[s[5] for s in v],
            gpu_memory_gib=np.array([math.nan if s[6] is None else s[6] for s in v], dtype=np.float64),
            disk_conf=np.array([s[7] if len(s) > 7 else -1 for s in v], dtype=np.int64),
        )


def _shape_tuple(res: Any) -> Tuple:
    n = len(res.gpus)
    if n > 0:
        g = res.gpus[0]
        return (res.cpus, res.memory_mib / 1024, res.disk.size_mib / 1024, n,
                vendor_code(g.vendor), g.name, g.memory_mib / 1024)
    return (res.cpus, res.memory_mib / 1024, res.disk.size_mib / 1024, 0, 0, None, None)


def _shape_of_resources(shapes: _Shapes, res: Any) -> int:
    return shapes.intern(_shape_tuple(res))


class PoolParseCache:
    """What the matcher needs from an instance's stored JSON, keyed by the JSON text itself.

    The reference parses ``instance.offer`` for every (job, instance) pair
    (``server/services/pools.py:499``); a table build parses it once per row, and with this cache
    once per distinct text across scheduler ticks: an instance whose stored offer and provisioning
    data did not change costs two dictionary look-ups when the pool is indexed again (SURVEY §8(f)
    row 2, "removes the 100 µs/row JSON parse").  Status, reachability, blocks, fleet, backend and
    region are plain columns of the row and are always read fresh."""

    def __init__(self, max_entries: int = 2_000_000):
        self.max_entries = max_entries
        self._offers: Dict[str, Tuple[Tuple, float, bool]] = {}
        self._jpds: Dict[str, Tuple[str, Optional[str]]] = {}
        self.parsed = 0  # JSON documents actually parsed (for tests and logs)

    def offer_facts(self, raw: str) -> Tuple[Tuple, float, bool]:
        hit = self._offers.get(raw)
        if hit is None:
            offer = M.InstanceOffer.parse_raw(raw)  # a malformed document raises, as in the reference
            hit = (_shape_tuple(offer.instance.resources), offer.price, bool(offer.instance.resources.spot))
            self.parsed += 1
            if len(self._offers) >= self.max_entries:
                self._offers.clear()
            self._offers[raw] = hit
        return hit

    def jpd_facts(self, raw: str) -> Tuple[str, Optional[str]]:
        hit = self._jpds.get(raw)
        if hit is None:
            jpd = M.JobProvisioningData.parse_raw(raw)
            hit = (jpd.instance_type.name, jpd.availability_zone)
            self.parsed += 1
            if len(self._jpds) >= self.max_entries:
                self._jpds.clear()
            self._jpds[raw] = hit
        return hit


def columns_from_offers(offers: Sequence[Any], src_backends: Optional[Sequence[Any]] = None) -> OfferColumns:
    """``offers`` in source order; ``src_backends[i]`` is the ``BackendType`` of the Backend
    object that served offer i (defaults to the offer's own backend)."""
    shapes, regions, itypes = _Shapes(), _Interner(), _Interner()
    n = len(offers)
    shape_id = np.empty(n, np.int32)
    price = np.empty(n, np.float64)
    spot = np.empty(n, bool)
    avail = np.ones(n, bool)
    backend = np.empty(n, np.int8)
    src = np.empty(n, np.int8)
    region_id = np.empty(n, np.int32)
    itype_id = np.empty(n, np.int32)
    for i, o in enumerate(offers):
        res = o.instance.resources
        shape_id[i] = _shape_of_resources(shapes, res)
        price[i] = o.price
        spot[i] = res.spot
        a = getattr(o, "availability", None)
        avail[i] = True if a is None else a.is_available()
        backend[i] = backend_code(o.backend)
        src[i] = backend[i] if src_backends is None else backend_code(src_backends[i])
        region_id[i] = regions(o.region)
        itype_id[i] = itypes(o.instance.name)
    return OfferColumns(shapes=shapes.table(), shape_id=shape_id, price=price, spot=spot, available=avail,
                        backend=backend, src_backend=src, region_id=region_id, regions=regions.values,
                        itype_id=itype_id, itypes=itypes.values)


def columns_from_catalog_items(items: Sequence[Any]) -> OfferColumns:
    """``gpuhunt.CatalogItem`` rows of ONE provider (already price-sorted), for
    ``get_catalog_offers`` (``core/backends/base/offers.py:18-43``)."""
    shapes, regions, itypes = _Shapes(), _Interner(), _Interner()
    n = len(items)
    shape_id = np.empty(n, np.int32)
    price = np.empty(n, np.float64)
    spot = np.empty(n, bool)
    region_id = np.empty(n, np.int32)
    itype_id = np.empty(n, np.int32)
    for i, it in enumerate(items):
        shape_id[i] = shapes.add(it.cpu, float(it.memory), None if it.disk_size is None else float(it.disk_size),
                                 it.gpu_count, vendor_code(it.gpu_vendor), it.gpu_name,
                                 None if it.gpu_memory is None else float(it.gpu_memory))
        price[i] = it.price
        spot[i] = it.spot
        region_id[i] = regions(it.location)
        itype_id[i] = itypes(it.instance_name)
    z = np.zeros(n, np.int8)
    return OfferColumns(shapes=shapes.table(), shape_id=shape_id, price=price, spot=spot,
                        available=np.ones(n, bool), backend=z, src_backend=z, region_id=region_id,
                        regions=regions.values, itype_id=itype_id, itypes=itypes.values)


def _range_gib(rng: Any) -> Tuple[float, float]:
    return (math.nan if rng is None or rng.min is None else float(rng.min),
            math.nan if rng is None or rng.max is None else float(rng.max))


def columns_from_backend_items(items_by_backend: Sequence[Tuple[Any, Sequence[Any], Any]]) -> OfferColumns:
    """``[(backend_type, CatalogItem rows, configurable_disk_size), ...]`` in project-backend
    order, each list price-sorted: what every backend's ``get_catalog_offers`` reads from gpuhunt
    BEFORE a job's requirements are applied (``core/backends/base/offers.py:18-43``).  An item
    without a disk size keeps none: whether a job can have it -- its disk range must overlap the
    backend's configurable range (``:161-175``) -- is part of the device predicate, and its size
    is chosen when a result row is turned into an offer."""
    shapes, regions, itypes = _Shapes(), _Interner(), _Interner()
    ranges = _Interner()
    n = sum(len(items) for _, items, _ in items_by_backend)
    shape_id = np.empty(n, np.int32)
    price = np.empty(n, np.float64)
    spot = np.empty(n, bool)
    backend = np.empty(n, np.int8)
    region_id = np.empty(n, np.int32)
    itype_id = np.empty(n, np.int32)
    i = 0
    for btype, items, conf in items_by_backend:
        code = backend_code(btype)
        conf_id = ranges(_range_gib(conf))
        for it in items:
            no_disk = not it.disk_size  # choose_disk_size_mib tests truthiness (:166)
            shape_id[i] = shapes.add(it.cpu, float(it.memory), None if no_disk else float(it.disk_size),
                                     it.gpu_count, vendor_code(it.gpu_vendor), it.gpu_name,
                                     None if it.gpu_memory is None else float(it.gpu_memory),
                                     conf_id if no_disk else -1)
            price[i] = it.price
            spot[i] = it.spot
            backend[i] = code
            region_id[i] = regions(it.location)
            itype_id[i] = itypes(it.instance_name)
            i += 1
    return OfferColumns(shapes=shapes.table(), shape_id=shape_id, price=price, spot=spot,
                        available=np.ones(n, bool), backend=backend, src_backend=backend.copy(),
                        region_id=region_id, regions=regions.values, itype_id=itype_id, itypes=itypes.values,
                        disk_conf_ranges=list(ranges.values))


def columns_from_pool_instances(instances: Sequence[Any], *, blocks_of: Optional[Sequence[int]] = None,
                                offers: Optional[Sequence[Any]] = None,
                                parse_cache: Optional[PoolParseCache] = None) -> OfferColumns:
    """Pool rows in input order.  ``offers[i]`` may carry a pre-parsed (or block-sliced)
    offer for row i; otherwise ``instance.offer`` JSON is parsed here, once -- or not at all when
    ``parse_cache`` has seen the same text on an earlier call."""
    cache = parse_cache if parse_cache is not None else PoolParseCache()
    shapes, regions, itypes = _Shapes(), _Interner(), _Interner()
    fleets, names, zones = _Interner(), _Interner(), _Interner()
    n = len(instances)
    shape_id = np.zeros(n, np.int32)
    price = np.zeros(n, np.float64)
    spot = np.zeros(n, bool)
    backend = np.empty(n, np.int8)
    region_id = np.empty(n, np.int32)
    itype_id = np.full(n, -1, np.int32)
    status = np.empty(n, np.int8)
    flags = np.zeros(n, np.int16)
    fleet_id = np.empty(n, np.int32)
    name_id = np.empty(n, np.int32)
    zone_id = np.full(n, -1, np.int32)
    dummy = shapes.add(0, 0.0, 0.0, 0, 0, None, None)
    for i, inst in enumerate(instances):
        offer = offers[i] if offers is not None else None
        if offer is not None:
            shape_id[i] = _shape_of_resources(shapes, offer.instance.resources)
            price[i] = offer.price
            spot[i] = offer.instance.resources.spot
        elif inst.offer is not None:
            shape, price[i], spot[i] = cache.offer_facts(inst.offer)
            shape_id[i] = shapes.intern(shape)
        else:
            flags[i] |= FLAG_NO_OFFER
            shape_id[i] = dummy
        backend[i] = backend_code(inst.backend)
        region_id[i] = -1 if inst.region is None else regions(inst.region)
        jpd_raw = inst.job_provisioning_data
        if jpd_raw is not None:
            itype_name, zone = cache.jpd_facts(jpd_raw)
            itype_id[i] = itypes(itype_name)
            if zone is not None:
                zone_id[i] = zones(zone)
        st = inst.status
        status[i] = STATUS_CODES[st.value if hasattr(st, "value") else str(st)]
        if inst.unreachable:
            flags[i] |= FLAG_UNREACHABLE
        if inst.total_blocks is None:
            flags[i] |= FLAG_BLOCKS_UNKNOWN
        elif inst.total_blocks > 1:
            flags[i] |= FLAG_SHARED
        fleet_id[i] = -1 if inst.fleet_id is None else fleets(inst.fleet_id)
        name_id[i] = names(inst.name)
    return OfferColumns(shapes=shapes.table(), shape_id=shape_id, price=price, spot=spot,
                        available=np.ones(n, bool), backend=backend, src_backend=backend.copy(),
                        region_id=region_id, regions=regions.values, itype_id=itype_id, itypes=itypes.values,
                        status=status, flags=flags, fleet_id=fleet_id, fleets=fleets.values,
                        name_id=name_id, names=names.values, zone_id=zone_id, zones=zones.values)


# job filters -----------------------------------------------------------------------------


@dataclass
class JobFilter:
    """Everything one job asks of a row, after the reference's derivations
    (``server/services/offers.py:46-82`` / ``pools.py:425-456``) have been applied.  Immutable
    once it has been matched: ``job_columns`` keeps the table-independent numbers on the object."""

    query: QueryFilter = field(default_factory=QueryFilter)
    backend_types: Optional[Sequence[Any]] = None   # None = unconstrained
    keep_dstack_backend: bool = False               # the pre-filter keeps BackendType.DSTACK (offers.py:82)
    regions: Optional[Sequence[str]] = None
    instance_types: Optional[Sequence[str]] = None
    exclude_not_available: bool = False
    require_auto_divisible: bool = False            # blocks="auto" (offers.py:102-118)
    blocks: int = 0                                 # explicit block count: 0 none, b > 1, -1 invalid
    # pool rows
    status: Optional[Any] = None
    fleet_id: Any = None
    any_fleet: bool = True
    instance_name: Optional[str] = None
    zones: Optional[Sequence[str]] = None
    shared: Optional[bool] = None                   # None = not a pool filter
    require_offer: bool = False


def _pair(lo, hi) -> Tuple[float, float]:
    return (math.nan if lo is None else float(lo), math.nan if hi is None else float(hi))


_POOL_FORBID = FLAG_UNREACHABLE | FLAG_BLOCKS_UNKNOWN


def _filter_scalars(f: JobFilter) -> Tuple[Tuple[float, ...], Optional[List[str]]]:
    """The part of a filter's columns that does not depend on any table: 23 numbers (all exactly
    representable as float64, ``nan`` = no bound) and the lower-cased GPU names.  Kept on the filter
    object after the first use -- a pending job is matched again every scheduler pass -- so a
    filter and its query are to be treated as immutable once matched (``dataclasses.replace``
    gives a fresh filter)."""
    q = f.query
    if q.min_price is not None:
        raise ValueError("min_price is not part of the dstack path (offers.py:117 sets max_price only)")
    if q.max_compute_capability is not None:
        raise ValueError("max_compute_capability is never set by dstack (offers.py:143-144)")
    nan = math.nan
    cc = q.min_compute_capability
    if f.backend_types is not None:
        off = 0
        for b in f.backend_types:
            off |= 1 << backend_code(b)
        src = off | ((1 << BACKEND_CODES["dstack"]) if f.keep_dstack_backend else 0)
    else:
        off = src = 0xFFFFFFFF
    req = FLAG_AUTO_DIVISIBLE if f.require_auto_divisible else 0
    forbid, st_mask = 0, 0xFF
    if f.shared is not None:  # pool filter (pools.py:460,484-488)
        forbid |= _POOL_FORBID
        if f.shared:
            req |= FLAG_SHARED
        else:
            forbid |= FLAG_SHARED
        if f.require_offer:
            forbid |= FLAG_NO_OFFER
        if f.status is not None:
            st = f.status
            st_mask = 1 << STATUS_CODES[st.value if hasattr(st, "value") else str(st)]
    bounds = (q.min_cpu, q.max_cpu, q.min_memory, q.max_memory, q.min_disk_size, q.max_disk_size,
              q.min_gpu_count, q.max_gpu_count, q.min_gpu_memory, q.max_gpu_memory,
              q.min_total_gpu_memory, q.max_total_gpu_memory)
    scalars = (nan if q.max_price is None else float(q.max_price), -1 if q.spot is None else (1 if q.spot else 0),
               *[nan if v is None else float(v) for v in bounds],
               vendor_code(q.gpu_vendor), -1 if cc is None else int(cc[0]) * 1000 + int(cc[1]),
               off, src, 1 if f.exclude_not_available else 0, req, forbid, f.blocks, st_mask)
    return scalars, (None if q.gpu_name is None else [n.lower() for n in q.gpu_name])


def job_columns(table: EncodedTable, filters: Sequence[JobFilter]) -> JobColumns:
    """Job filters -> columns.  One pass over the filter objects: the table-independent numbers
    come from the filter (computed once per filter, see ``_filter_scalars``) and go into one
    float64 array at the end; names, regions, instance types, zones, fleets and instance names are
    looked up in this table's dictionaries."""
    j = len(filters)
    cols = {c.name: c for c in table.columns}

    def cached(f: JobFilter):
        c = f.__dict__.get("_scalars")
        if c is None:
            c = f.__dict__["_scalars"] = _filter_scalars(f)
        return c

    per_filter = [cached(f) for f in filters]
    rows = [c[0] for c in per_filter]

    def csr(values: List[Optional[Sequence[Any]]], lookup: Dict[Any, int]):
        """Per job a set of dictionary ids (-1: a value this table does not hold), or "any"."""
        any_ = np.array([v is None for v in values], bool)
        if any_.all():
            return any_, np.zeros(j + 1, np.int64), np.zeros(0, np.int64)
        ptr, ids, get = [0], [], lookup.get
        for v in values:
            if v is not None:
                ids.extend([get(x, -1) for x in v])
            ptr.append(len(ids))
        return any_, np.array(ptr, np.int64), np.array(ids, np.int64)

    gn_any, gn_ptr, gn_ids = csr([c[1] for c in per_filter], table.gpu_dict["names"])
    r_any, r_ptr, r_ids = csr([f.regions for f in filters], cols["region"].lookup)
    t_any, t_ptr, t_ids = csr([f.instance_types for f in filters], cols["itype"].lookup)
    if table.is_pool:
        pool = [f.shared is not None for f in filters]
        z_any, z_ptr, z_ids = csr([f.zones if p else None for f, p in zip(filters, pool)], cols["zone"].lookup)
        fleet_get, name_get = cols["fleet"].lookup.get, cols["name"].lookup.get
        fleet_any = [not p or f.any_fleet for f, p in zip(filters, pool)]
        fleet_ids = [-1 if a else fleet_get(f.fleet_id, -1) for f, a in zip(filters, fleet_any)]
        inst_any = [not p or f.instance_name is None for f, p in zip(filters, pool)]
        inst_ids = [-1 if a else name_get(f.instance_name, -1) for f, a in zip(filters, inst_any)]

    a = np.array(rows, dtype=np.float64).reshape(j, 23)
    pair = lambda k: np.ascontiguousarray(a[:, 2 + 2 * k:4 + 2 * k])
    blocks_a = a[:, 21].astype(np.int64)
    jc = JobColumns(
        n=j, max_price=np.ascontiguousarray(a[:, 0]), spot=a[:, 1].astype(np.int8), cpu=pair(0), memory=pair(1),
        disk=pair(2), gpu_count=pair(3), gpu_memory=pair(4), gpu_total=pair(5), gpu_vendor=a[:, 14].astype(np.int8),
        cc_min=a[:, 15].astype(np.int64), gpu_name_any=gn_any, gpu_name_ptr=gn_ptr,
        gpu_name_ids=gn_ids, src_backend_mask=a[:, 17].astype(np.uint32), off_backend_mask=a[:, 16].astype(np.uint32),
        exclude_not_available=a[:, 18] != 0, region_any=r_any, region_ptr=r_ptr, region_ids=r_ids,
        itype_any=t_any, itype_ptr=t_ptr, itype_ids=t_ids, flag_req=a[:, 19].astype(np.uint16),
        flag_forbid=a[:, 20].astype(np.uint16), status_mask=a[:, 22].astype(np.uint8),
        blocks=blocks_a if blocks_a.any() else None,
    )
    if table.is_pool:
        jc.fleet_any, jc.fleet_ids = np.array(fleet_any, bool), np.array(fleet_ids, np.int64)
        jc.name_any, jc.name_ids = np.array(inst_any, bool), np.array(inst_ids, np.int64)
        jc.zone_any, jc.zone_ptr, jc.zone_ids = z_any, z_ptr, z_ids
    return jc
