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

    def add(self, cpus, memory_gib, disk_gib, gpu_count, vendor, name, gpu_gib) -> int:
        return self.intern((cpus, memory_gib, disk_gib, gpu_count, vendor, name, gpu_gib))

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
    (``server/services/offers.py:46-82`` / ``pools.py:425-456``) have been applied."""

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


def job_columns(table: EncodedTable, filters: Sequence[JobFilter]) -> JobColumns:
    j = len(filters)
    cols = {c.name: c for c in table.columns}
    names = table.gpu_dict["names"]

    f64 = lambda: np.full((j, 2), np.nan)
    cpu, mem, disk, gcnt, gmem, gtot = f64(), f64(), f64(), f64(), f64(), f64()
    max_price = np.full(j, np.nan)
    spot = np.full(j, -1, np.int8)
    vendor = np.zeros(j, np.int8)
    cc_min = np.full(j, -1, np.int64)
    name_any = np.ones(j, bool)
    src_mask = np.full(j, 0xFFFFFFFF, np.uint32)
    off_mask = np.full(j, 0xFFFFFFFF, np.uint32)
    excl = np.zeros(j, bool)
    flag_req = np.zeros(j, np.uint16)
    flag_forbid = np.zeros(j, np.uint16)
    blocks = np.zeros(j, np.int64)
    status_mask = np.full(j, 0xFF, np.uint8)
    fleet_any = np.ones(j, bool)
    fleet_ids = np.full(j, -1, np.int64)
    inst_any = np.ones(j, bool)
    inst_ids = np.full(j, -1, np.int64)

    class Csr:
        def __init__(self, lookup: Dict[Any, int]):
            self.lookup, self.any, self.ptr, self.ids = lookup, np.ones(j, bool), [0], []

        def put(self, i: int, values: Optional[Sequence[Any]]) -> None:
            if values is not None:
                self.any[i] = False
                self.ids.extend(self.lookup.get(v, -1) for v in values)
            self.ptr.append(len(self.ids))

        def arrays(self):
            return self.any, np.array(self.ptr, np.int64), np.array(self.ids, np.int64)

    gname = Csr(names)
    region = Csr(cols["region"].lookup)
    itype = Csr(cols["itype"].lookup)
    zone = Csr(cols["zone"].lookup) if table.is_pool else None

    for i, f in enumerate(filters):
        q = f.query
        if q.min_price is not None:
            raise ValueError("min_price is not part of the dstack path (offers.py:117 sets max_price only)")
        if q.max_compute_capability is not None:
            raise ValueError("max_compute_capability is never set by dstack (offers.py:143-144)")
        if q.max_price is not None:
            max_price[i] = q.max_price
        if q.spot is not None:
            spot[i] = 1 if q.spot else 0
        cpu[i] = _pair(q.min_cpu, q.max_cpu)
        mem[i] = _pair(q.min_memory, q.max_memory)
        disk[i] = _pair(q.min_disk_size, q.max_disk_size)
        gcnt[i] = _pair(q.min_gpu_count, q.max_gpu_count)
        gmem[i] = _pair(q.min_gpu_memory, q.max_gpu_memory)
        gtot[i] = _pair(q.min_total_gpu_memory, q.max_total_gpu_memory)
        vendor[i] = vendor_code(q.gpu_vendor)
        if q.min_compute_capability is not None:
            cc_min[i] = int(q.min_compute_capability[0]) * 1000 + int(q.min_compute_capability[1])
        if q.gpu_name is not None:
            name_any[i] = False
        gname.put(i, None if q.gpu_name is None else [n.lower() for n in q.gpu_name])
        if f.backend_types is not None:
            m = 0
            for b in f.backend_types:
                m |= 1 << backend_code(b)
            off_mask[i] = m
            src_mask[i] = m | ((1 << BACKEND_CODES["dstack"]) if f.keep_dstack_backend else 0)
        excl[i] = f.exclude_not_available
        region.put(i, f.regions)
        itype.put(i, f.instance_types)
        if f.require_auto_divisible:
            flag_req[i] |= FLAG_AUTO_DIVISIBLE
        blocks[i] = f.blocks
        if f.shared is not None:  # pool filter (pools.py:460,484-488)
            flag_forbid[i] |= FLAG_UNREACHABLE | FLAG_BLOCKS_UNKNOWN
            if f.shared:
                flag_req[i] |= FLAG_SHARED
            else:
                flag_forbid[i] |= FLAG_SHARED
            if f.require_offer:
                flag_forbid[i] |= FLAG_NO_OFFER
            if f.status is not None:
                st = f.status
                status_mask[i] = 1 << STATUS_CODES[st.value if hasattr(st, "value") else str(st)]
            if not f.any_fleet:
                fleet_any[i] = False
                fleet_ids[i] = cols["fleet"].lookup.get(f.fleet_id, -1)
            if f.instance_name is not None:
                inst_any[i] = False
                inst_ids[i] = cols["name"].lookup.get(f.instance_name, -1)
            zone.put(i, f.zones)
        elif zone is not None:
            zone.put(i, None)

    gn_any, gn_ptr, gn_ids = gname.arrays()
    r_any, r_ptr, r_ids = region.arrays()
    t_any, t_ptr, t_ids = itype.arrays()
    jc = JobColumns(
        n=j, max_price=max_price, spot=spot, cpu=cpu, memory=mem, disk=disk, gpu_count=gcnt, gpu_memory=gmem,
        gpu_total=gtot, gpu_vendor=vendor, cc_min=cc_min, gpu_name_any=name_any, gpu_name_ptr=gn_ptr,
        gpu_name_ids=gn_ids, src_backend_mask=src_mask, off_backend_mask=off_mask,
        exclude_not_available=excl, region_any=r_any, region_ptr=r_ptr, region_ids=r_ids,
        itype_any=t_any, itype_ptr=t_ptr, itype_ids=t_ids, flag_req=flag_req, flag_forbid=flag_forbid,
        status_mask=status_mask, blocks=blocks if blocks.any() else None,
    )
    if table.is_pool:
        z_any, z_ptr, z_ids = zone.arrays()
        jc.fleet_any, jc.fleet_ids = fleet_any, fleet_ids
        jc.name_any, jc.name_ids = inst_any, inst_ids
        jc.zone_any, jc.zone_ptr, jc.zone_ids = z_any, z_ptr, z_ids
    return jc
