"""Host-side encoding: offers -> dictionary-encoded table, job filters -> predicate records.

Everything the GPU compares is a small integer.  Each attribute of the offers is replaced
by its ordinal among the sorted distinct values of that attribute in the table, and each
job bound is translated to an ordinal with the same IEEE-754 double comparisons the
reference performs in Python (``gpuhunt.matches``), so the device
never compares a float and results are bit-exact by construction.

Row order: the table is stored in the order the reference returns offers --
``heapq.merge`` of price-sorted per-backend lists in project-backend order followed by a
stable partition on ``availability.is_available()``
(``server/services/backends/__init__.py:449-451``).  That is a stable sort of the source
list by ``(not available, price)``; a job's ``max_price`` then is a row cut-off in each of
the two segments and "first k by price" is "first k matching rows".

Nothing here evaluates a (job, offer) pair: this module only builds inputs for
``libdsmatch.so``.
"""

from __future__ import annotations

import dataclasses
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

from ._gpuhunt import AcceleratorVendor, compute_capability_of
from .models import BackendType, InstanceStatus
from .native import (RAW_JOB, RAWF_EXCLUDE_NOT_AVAILABLE, RAWF_FLEET_ANY, RAWF_GPU_NAME_ANY, RAWF_ITYPE_ANY,
                     RAWF_NAME_ANY, RAWF_REGION_ANY, RAWF_ZONE_ANY, ROLES)

# codes -------------------------------------------------------------------------------------

BACKEND_CODES: Dict[str, int] = {b.value: i for i, b in enumerate(BackendType)}
NO_BACKEND_CODE = len(BACKEND_CODES)  # pool rows may have backend = None
VENDOR_CODES: Dict[Optional[str], int] = {None: 0, "nvidia": 1, "amd": 2, "google": 3, "intel": 4}
STATUS_CODES: Dict[str, int] = {s.value: i for i, s in enumerate(InstanceStatus)}
NO_STATUS_CODE = len(STATUS_CODES)

# flag bits of the backend/flags dictionary
FLAG_UNREACHABLE = 1
FLAG_BLOCKS_UNKNOWN = 2   # total_blocks is None (still provisioning)
FLAG_SHARED = 4           # total_blocks > 1
FLAG_NO_OFFER = 8         # instance.offer is None
FLAG_AUTO_DIVISIBLE = 16  # is_divisible_into_blocks(cpu, gpu, "auto")[0]
# Pool tables keep what changes between scheduler passes -- status and these four flags -- in a
# column of its own ("state", code = status * 16 + flags, every combination enumerated), so an
# instance that turns BUSY or unreachable is a one-row patch and never a new dictionary entry.
DYNAMIC_FLAGS = FLAG_UNREACHABLE | FLAG_BLOCKS_UNKNOWN | FLAG_SHARED | FLAG_NO_OFFER

KIND_RANGE, KIND_BITSET, KIND_GPU, KIND_BK, KIND_PAIR = 1, 2, 3, 4, 5
COLF_NULL_PASSES = 1
COLF_PAIR_A_BITSET = 2
GPUF_MEM_BOUNDED, GPUF_CC_BOUNDED, GPUF_NAME_ANY = 1, 2, 8
NO_ROW = 0xFFFFFFFF
MAX_CODES = 65535
LUT_THRESHOLD = 1024  # set columns with more dictionary entries go through a per-call remap
# Shared memory one of three resident CTAs per SM has for a job group's acceptance tables
# (dsmatch.cu:make_plan).  Every entry is stored 8 times for conflict-free gathers; when that does
# not fit the library halves the copies of the largest column.  cpu and memory are merged into one
# pair-coded column only if the tables then still fit with at least 4 copies of every column.
TABLE_AREA_BUDGET = 220 * 1024 // 3 - 1024 - 1024


def table_copies(cards: Sequence[int]) -> List[int]:
    """log2 of the copies per column the library will keep in shared memory for these dictionary
    sizes (mirror of dsmatch.cu:make_plan): 8 of everything, else the largest column 4 / 2 / 1,
    else one copy of everything."""
    big = max(range(len(cards)), key=lambda i: (cards[i], -i))
    pad = lambda n: (n + 127) // 128 * 128
    area = lambda bshift: sum(pad((c * 16) << (bshift if i == big else 3)) for i, c in enumerate(cards))
    bshift = 3
    while bshift > 0 and area(bshift) > TABLE_AREA_BUDGET:
        bshift -= 1
    if area(bshift) > TABLE_AREA_BUDGET:
        return [0] * len(cards)
    return [bshift if i == big else 3 for i in range(len(cards))]


def backend_code(b: Any) -> int:
    if b is None:
        return NO_BACKEND_CODE
    return BACKEND_CODES[b.value if hasattr(b, "value") else str(b)]


def vendor_code(v: Any) -> int:
    if v is None:
        return 0
    return VENDOR_CODES[v.value if hasattr(v, "value") else str(v).lower()]


def _cc_key(cc: Optional[Tuple[int, int]]) -> int:
    return -1 if cc is None else int(cc[0]) * 1000 + int(cc[1])


# offers in columnar form -------------------------------------------------------------------


@dataclass
class ShapeTable:
    """Distinct resource shapes (instance types).  All resource predicates depend on the
    shape only, so dictionaries are built over shapes, not offers."""

    cpus: np.ndarray            # int64
    memory_gib: np.ndarray      # float64: memory_mib / 1024 (offers) or CatalogItem.memory
    disk_gib: np.ndarray        # float64, NaN = no disk size on the item (always passes)
    gpu_count: np.ndarray       # int64
    gpu_vendor: np.ndarray      # int8 vendor code, 0 = none
    gpu_name: List[Optional[str]]  # original casing; None without GPUs
    gpu_memory_gib: np.ndarray  # float64, NaN without GPUs
    # catalog items without a disk size whose backend lets the job choose one: index into
    # OfferColumns.disk_conf_ranges (-1 / None: the item has its own size, or passes as it is)
    disk_conf: Optional[np.ndarray] = None

    def __len__(self) -> int:
        return len(self.cpus)


@dataclass
class OfferColumns:
    """Offers in SOURCE order (per-backend lists concatenated in project-backend order)."""

    shapes: ShapeTable
    shape_id: np.ndarray        # int32 -> shapes
    price: np.ndarray           # float64
    spot: np.ndarray            # bool
    available: np.ndarray       # bool: availability.is_available()
    backend: np.ndarray         # int8: code of offer.backend
    src_backend: np.ndarray     # int8: code of the Backend object that served the offer
    region_id: np.ndarray       # int32 -> regions
    regions: List[str]
    itype_id: np.ndarray        # int32 -> itypes; -1 = no instance type known (passes)
    itypes: List[str]
    # pool-only admin attributes (None for catalog tables)
    status: Optional[np.ndarray] = None      # int8 status code
    flags: Optional[np.ndarray] = None       # int16 FLAG_* bits
    fleet_id: Optional[np.ndarray] = None    # int32 -> fleets, -1 = no fleet
    fleets: List[Any] = field(default_factory=list)
    name_id: Optional[np.ndarray] = None     # int32 -> names
    names: List[str] = field(default_factory=list)
    zone_id: Optional[np.ndarray] = None     # int32 -> zones, -1 = unknown zone (passes)
    zones: List[str] = field(default_factory=list)
    # (min, max) GiB of every configurable-disk range shapes refer to (NaN = open end); such an
    # item matches a job iff the job's disk range and this one overlap (choose_disk_size_mib,
    # core/backends/base/offers.py:161-175)
    disk_conf_ranges: List[Tuple[float, float]] = field(default_factory=list)

    def __len__(self) -> int:
        return len(self.price)

    @property
    def is_pool(self) -> bool:
        return self.fleet_id is not None


def _auto_divisible(cpus: np.ndarray, gpu_eff: np.ndarray) -> np.ndarray:
    """``is_divisible_into_blocks(cpu, gpu, "auto")[0]`` (server/services/offers.py:121-136)."""
    b = np.where(gpu_eff == 0, cpus, np.minimum(cpus, gpu_eff))
    ok = b >= 1
    bs = np.where(ok, b, 1)
    return ok & (cpus % bs == 0) & (gpu_eff % bs == 0)


# encoded table ---------------------------------------------------------------------------


@dataclass
class Column:
    name: str
    kind: int
    card: int                      # dictionary entries incl. null / dead entries
    flags: int = 0
    values: Optional[np.ndarray] = None   # RANGE: sorted distinct values (float64)
    lookup: Optional[Dict[Any, int]] = None  # BITSET: value -> code


@dataclass
class EncodedTable:
    """Device-ready table plus the host-side facts needed to pack jobs and map rows back."""

    n_src: int
    n_rows: int               # physical rows (padded)
    split_row: int
    n_avail: int
    n_cols: int               # 8 or 16
    rows: np.ndarray          # uint16 [n_rows, n_cols]
    price: np.ndarray         # float64 [n_rows] (NaN on padding rows)
    perm: np.ndarray          # int64 [n_rows]: physical row -> source index, -1 on padding
    columns: List[Column]
    gpu_entries: np.ndarray   # uint32 [n_gpu, 3]
    bk_entries: np.ndarray    # uint32 [n_bk, 2]
    gpu_dict: Dict[str, Any]  # count / memory / total / cc value arrays, name -> id
    price_a: np.ndarray       # prices of the first segment (ascending)
    price_b: np.ndarray       # prices of the second segment (ascending)
    is_pool: bool
    # cpu and memory as ONE column (kind PAIR) when the dictionaries are small enough: entry =
    # cpu ordinal | memory ordinal << 16; cpu_values / memory_values are then kept here
    pair_entries: Optional[np.ndarray] = None   # uint32 [n_pairs]
    cpu_values: Optional[np.ndarray] = None
    memory_values: Optional[np.ndarray] = None
    disk_conf_ranges: List[Tuple[float, float]] = field(default_factory=list)

    @property
    def n_rest(self) -> int:
        """Live rows of the second segment."""
        return self.n_src - self.n_avail

    def col(self, name: str) -> int:
        for i, c in enumerate(self.columns):
            if c.name == name:
                return i
        raise KeyError(name)


def _ordinals(values: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    uniq, inv = np.unique(values, return_inverse=True)
    return uniq.astype(np.float64), inv.astype(np.int64)


def _check_card(name: str, n: int) -> None:
    if n > MAX_CODES:
        raise ValueError(f"column {name!r} has {n} distinct values; the u16 row format allows {MAX_CODES}")


def encode_table(oc: OfferColumns, *, check_sorted_backends: bool = True,
                 pair_columns: Optional[bool] = None) -> EncodedTable:
    """``pair_columns``: merge cpu and memory into one pair-coded column (one table gather per row
    instead of two).  ``None`` = when the tables then still fit with at least 4 copies of every column (``table_copies``)."""
    n = len(oc)
    sh = oc.shapes
    price = np.asarray(oc.price, dtype=np.float64)
    avail = np.asarray(oc.available, dtype=bool)
    if np.isnan(price).any():
        raise ValueError("NaN price")

    if check_sorted_backends and n > 1 and not oc.is_pool:
        # heapq.merge equals a sort only if every backend's list is price-sorted (gpuhunt's
        # contract, contributing/GPUHUNT.md:33-35).  Refuse anything else loudly.
        src = np.asarray(oc.src_backend)
        same = src[1:] == src[:-1]
        if (same & (price[1:] < price[:-1])).any():
            raise ValueError("offers of a backend are not sorted by price; heapq.merge order would be "
                             "job-dependent (server/services/backends/__init__.py:449)")
        # lists must be contiguous per backend for source order to mean (backend_pos, position)
        starts = np.concatenate(([True], ~same))
        if len(np.unique(src[starts])) != int(starts.sum()):
            raise ValueError("offers of one backend must be contiguous in the source list")

    # ---- resource dictionaries over shapes
    cpu_vals, cpu_ord = _ordinals(np.asarray(sh.cpus, dtype=np.float64))
    mem_vals, mem_ord = _ordinals(np.asarray(sh.memory_gib, dtype=np.float64))
    disk = np.asarray(sh.disk_gib, dtype=np.float64)
    disk_null = np.isnan(disk)
    disk_vals, disk_ord_nn = _ordinals(disk[~disk_null]) if (~disk_null).any() else (np.zeros(0), np.zeros(0, np.int64))
    n_conf = len(oc.disk_conf_ranges)
    disk_ord = np.full(len(sh), len(disk_vals) + n_conf, dtype=np.int64)  # null entry = last
    disk_ord[~disk_null] = disk_ord_nn
    if n_conf and sh.disk_conf is not None:  # entries [n values, n values + n_conf): one per configurable range
        conf = np.asarray(sh.disk_conf, dtype=np.int64)
        pick = disk_null & (conf >= 0)
        disk_ord[pick] = len(disk_vals) + conf[pick]
    for nm, vals in (("cpu", cpu_vals), ("memory", mem_vals), ("disk", disk_vals)):
        _check_card(nm, len(vals) + 1 + (n_conf if nm == "disk" else 0))

    # ---- accelerator configuration dictionary
    gcount = np.asarray(sh.gpu_count, dtype=np.int64)
    has_gpu = gcount > 0
    gmem = np.where(has_gpu, np.asarray(sh.gpu_memory_gib, dtype=np.float64), 0.0)
    gmem = np.where(np.isnan(gmem), 0.0, gmem)
    # vendor / name are taken as the item states them, also on items without GPUs: the
    # reference tests them independently of the count (App. A.1 items 7, 8, 10)
    gvend = np.asarray(sh.gpu_vendor, dtype=np.int64)
    lower_names: Dict[str, int] = {}
    name_id = np.full(len(sh), -1, dtype=np.int64)
    for i, nm in enumerate(sh.gpu_name):
        if nm is not None:
            name_id[i] = lower_names.setdefault(nm.lower(), len(lower_names))
    cnt_vals, cnt_ord = _ordinals(gcount.astype(np.float64))
    gm_vals, gm_ord = _ordinals(gmem)
    total = np.where(has_gpu, gcount.astype(np.float64) * gmem, 0.0)  # int * float, as Python does
    tot_vals, tot_ord = _ordinals(total)
    cc_key = np.array([_cc_key(compute_capability_of(nm)) for nm in sh.gpu_name], dtype=np.int64)
    cc_vals = np.unique(cc_key[cc_key >= 0]).astype(np.float64)
    cc_ord = np.where(cc_key >= 0, np.searchsorted(cc_vals, cc_key.astype(np.float64)), 0xFFFF)
    gkey = np.stack([cnt_ord, gm_ord, gvend, name_id, has_gpu.astype(np.int64)], axis=1)
    guniq, ginv = np.unique(gkey, axis=0, return_inverse=True)
    ginv = np.asarray(ginv).reshape(-1)
    _check_card("gpu", len(guniq))
    # representative shape of each entry, to read total / cc ordinals
    rep = np.zeros(len(guniq), dtype=np.int64)
    rep[ginv] = np.arange(len(sh))
    gpu_entries = np.zeros((len(guniq), 3), dtype=np.uint32)
    e_name = np.where(guniq[:, 3] >= 0, guniq[:, 3], 0xFFFF)
    gpu_entries[:, 0] = (guniq[:, 0] | (guniq[:, 1] << 16)).astype(np.uint32)
    gpu_entries[:, 1] = (tot_ord[rep] | (cc_ord[rep] << 16)).astype(np.uint32)
    # bits 25..31: GPUs that count for block splitting (a TPU slice counts as one,
    # server/services/offers.py:110-112), saturated at 127
    e_count = cnt_vals[guniq[:, 0]].astype(np.int64)
    e_eff = np.minimum(np.where(guniq[:, 2] == VENDOR_CODES["google"], np.minimum(e_count, 1), e_count), 127)
    gpu_entries[:, 2] = (e_name | (guniq[:, 2] << 16) | (guniq[:, 4] << 24) | (e_eff << 25)).astype(np.uint32)

    # ---- backend / spot / status / flags dictionary (entry 0 = dead)
    sid = np.asarray(oc.shape_id, dtype=np.int64)
    gpu_eff = np.where(np.asarray(sh.gpu_vendor)[sid] == VENDOR_CODES["google"],
                       np.minimum(gcount[sid], 1), gcount[sid])  # TPUs count as one (offers.py:110-112)
    flags = np.zeros(n, dtype=np.int64) if oc.flags is None else np.asarray(oc.flags, dtype=np.int64).copy()
    flags |= np.where(_auto_divisible(np.asarray(sh.cpus, dtype=np.int64)[sid], gpu_eff), FLAG_AUTO_DIVISIBLE, 0)
    status = np.full(n, NO_STATUS_CODE, dtype=np.int64) if oc.status is None else np.asarray(oc.status, dtype=np.int64)
    state_code = None
    if oc.is_pool:  # status and the dynamic flags live in the "state" column
        state_code = status * 16 + (flags & DYNAMIC_FLAGS)
        status = np.full(n, NO_STATUS_CODE, dtype=np.int64)
        flags = flags & ~np.int64(DYNAMIC_FLAGS)
    bkey = (np.asarray(oc.src_backend, dtype=np.int64) | (np.asarray(oc.backend, dtype=np.int64) << 5)
            | (np.asarray(oc.spot, dtype=np.int64) << 10) | (status << 11) | (flags << 14))
    present = np.unique(bkey)
    _check_card("bk", len(present) + 1)
    bk_code = np.searchsorted(present, bkey) + 1
    bk_entries = np.zeros((len(present) + 1, 2), dtype=np.uint32)
    bk_entries[1:, 0] = ((present & 31) | (((present >> 5) & 31) << 8) | (((present >> 10) & 1) << 16)
                         | (((present >> 11) & 7) << 24)).astype(np.uint32)
    bk_entries[1:, 1] = (present >> 14).astype(np.uint32)

    # ---- set columns: code = id, plus one entry for "unknown" (-1) where that passes
    def set_codes(ids: np.ndarray, card: int, name: str) -> Tuple[np.ndarray, int]:
        ids = np.asarray(ids, dtype=np.int64)
        _check_card(name, card + 1)
        return np.where(ids < 0, card, ids), card + 1  # last entry = unknown

    # Column order = gather order.  k_match gathers, per group of 128 jobs, only the columns up to
    # the last one that rejects anything for that group, so the columns jobs usually leave open
    # (default disk, no region / instance-type filter ...) come last.
    reg_codes, reg_card = set_codes(oc.region_id, len(oc.regions), "region")
    it_codes, it_card = set_codes(oc.itype_id, len(oc.itypes), "itype")
    tail_cols: List[Column] = [
        Column("gpu", KIND_GPU, len(guniq)),
        Column("bk", KIND_BK, len(present) + 1),
        Column("disk", KIND_BITSET if n_conf else KIND_RANGE, len(disk_vals) + n_conf + 1,
               flags=0 if n_conf else COLF_NULL_PASSES, values=disk_vals),
        Column("region", KIND_BITSET, reg_card, lookup={r: i for i, r in enumerate(oc.regions)}),
    ]
    tail_rows: List[np.ndarray] = [ginv[sid], bk_code, disk_ord[sid], reg_codes]
    if oc.is_pool:
        tail_cols.insert(2, Column("state", KIND_BITSET, (NO_STATUS_CODE + 1) * 16))
        tail_rows.insert(2, state_code)
        fl_codes, fl_card = set_codes(oc.fleet_id, len(oc.fleets), "fleet")
        nm_codes, nm_card = set_codes(oc.name_id, len(oc.names), "name")
        zn_codes, zn_card = set_codes(oc.zone_id, len(oc.zones), "zone")
        tail_cols += [
            Column("fleet", KIND_BITSET, fl_card, lookup={r: i for i, r in enumerate(oc.fleets)}),
            Column("itype", KIND_BITSET, it_card, lookup={r: i for i, r in enumerate(oc.itypes)}),
            Column("name", KIND_BITSET, nm_card, lookup={r: i for i, r in enumerate(oc.names)}),
            Column("zone", KIND_BITSET, zn_card, lookup={r: i for i, r in enumerate(oc.zones)}),
        ]
        tail_rows += [fl_codes, it_codes, nm_codes, zn_codes]
    else:
        tail_cols.append(Column("itype", KIND_BITSET, it_card, lookup={r: i for i, r in enumerate(oc.itypes)}))
        tail_rows.append(it_codes)

    # cpu + memory: one pair-coded column when the tables stay small enough for eight copies
    pkey = cpu_ord.astype(np.int64) | (mem_ord.astype(np.int64) << 16)
    puniq, pinv = np.unique(pkey, return_inverse=True)
    resident = [c.card for c in tail_cols if c.card <= LUT_THRESHOLD]
    if pair_columns is None:
        pair_columns = min(table_copies([len(puniq)] + resident)) >= 2 and len(puniq) <= MAX_CODES
    pair_entries = None
    if pair_columns:
        _check_card("cpu_memory", len(puniq))
        pair_entries = puniq.astype(np.uint32)
        columns = [Column("cpu_memory", KIND_PAIR, len(puniq))] + tail_cols
        per_row = [np.asarray(pinv).reshape(-1)[sid]] + tail_rows
    else:
        columns = [Column("cpu", KIND_RANGE, len(cpu_vals), values=cpu_vals),
                   Column("memory", KIND_RANGE, len(mem_vals), values=mem_vals)] + tail_cols
        per_row = [cpu_ord[sid], mem_ord[sid]] + tail_rows
    n_cols = 8 if len(columns) <= 8 else 16

    # ---- row order and padding
    order = np.lexsort((price, ~avail))  # stable: ties keep source order
    n_avail = int(avail.sum())
    split_row = (n_avail + 31) // 32 * 32
    n_b = n - n_avail
    n_rows = split_row + (n_b + 31) // 32 * 32
    if n_rows == 0:
        n_rows = 32
        split_row = 32
    perm = np.full(n_rows, -1, dtype=np.int64)
    perm[:n_avail] = order[:n_avail]
    perm[split_row:split_row + n_b] = order[n_avail:]
    live = perm >= 0
    rows = np.zeros((n_rows, n_cols), dtype=np.uint16)  # padding rows: bk code 0 = dead entry
    src = perm[live]
    for c, codes in enumerate(per_row):
        rows[live, c] = codes[src].astype(np.uint16)
    price_phys = np.full(n_rows, np.nan)
    price_phys[live] = price[src]

    return EncodedTable(
        n_src=n, n_rows=n_rows, split_row=split_row, n_avail=n_avail, n_cols=n_cols, rows=rows,
        price=price_phys, perm=perm, columns=columns, gpu_entries=gpu_entries, bk_entries=bk_entries,
        gpu_dict=dict(count=cnt_vals, memory=gm_vals, total=tot_vals, cc=cc_vals, names=lower_names),
        price_a=price[order[:n_avail]], price_b=price[order[n_avail:]], is_pool=oc.is_pool,
        pair_entries=pair_entries, cpu_values=cpu_vals, memory_values=mem_vals,
        disk_conf_ranges=list(oc.disk_conf_ranges),
    )


# jobs -------------------------------------------------------------------------------------


@dataclass
class JobColumns:
    """J job filters in columnar form.  NaN = bound not set.  Sets are CSR over the TABLE's
    dictionary ids (``*_ptr`` length J+1; an id of -1 stands for a value the table does not
    contain and matches nothing); ``*_any`` marks "no filter"."""

    n: int
    max_price: np.ndarray
    spot: np.ndarray                 # int8: -1 either, 0 on-demand, 1 spot
    cpu: np.ndarray                  # float64 [J,2] (min, max)
    memory: np.ndarray
    disk: np.ndarray
    gpu_count: np.ndarray
    gpu_memory: np.ndarray
    gpu_total: np.ndarray
    gpu_vendor: np.ndarray           # int8 vendor code, 0 = no filter
    cc_min: np.ndarray               # int64 key major*1000+minor, -1 = no bound
    gpu_name_any: np.ndarray         # bool
    gpu_name_ptr: np.ndarray
    gpu_name_ids: np.ndarray
    src_backend_mask: np.ndarray     # uint32
    off_backend_mask: np.ndarray     # uint32
    exclude_not_available: np.ndarray  # bool
    region_any: np.ndarray
    region_ptr: np.ndarray
    region_ids: np.ndarray
    itype_any: np.ndarray
    itype_ptr: np.ndarray
    itype_ids: np.ndarray
    # flags that rows must / must not carry, and the statuses accepted (pool filters, blocks=auto)
    flag_req: Optional[np.ndarray] = None     # uint16
    flag_forbid: Optional[np.ndarray] = None  # uint16
    status_mask: Optional[np.ndarray] = None  # uint8 bit per status code; 0xFF any
    fleet_any: Optional[np.ndarray] = None
    fleet_ids: Optional[np.ndarray] = None    # int64 [J] single id (or -1)
    name_any: Optional[np.ndarray] = None
    name_ids: Optional[np.ndarray] = None
    zone_any: Optional[np.ndarray] = None
    zone_ptr: Optional[np.ndarray] = None
    zone_ids: Optional[np.ndarray] = None
    # explicit block count per job: 0 = none, b > 1 = rows must split into b blocks, -1 = invalid count
    blocks: Optional[np.ndarray] = None


def _range_words(vals: np.ndarray, bounds: np.ndarray) -> np.ndarray:
    """(min,max) float bounds -> packed inclusive ordinal range ``lo | hi<<16``.

    lo = first i with vals[i] >= min, hi = last i with vals[i] <= max: exactly the set the
    reference's ``not (v < min) and not (v > max)`` keeps.  Empty -> (1, 0)."""
    lo_b, hi_b = bounds[:, 0], bounds[:, 1]
    no_lo, no_hi = np.isnan(lo_b), np.isnan(hi_b)  # NaN = bound not set
    lo = np.where(no_lo, 0, np.searchsorted(vals, np.where(no_lo, 0.0, lo_b), side="left"))
    hi = np.where(no_hi, len(vals) - 1, np.searchsorted(vals, np.where(no_hi, 0.0, hi_b), side="right") - 1)
    empty = (lo > hi) | (len(vals) == 0)
    lo = np.where(empty, 1, lo).astype(np.uint32)
    hi = np.where(empty, 0, hi).astype(np.uint32)
    return lo | (hi << np.uint32(16))


def _pack_bool_rows(ok: np.ndarray) -> np.ndarray:
    """bool [J, card] -> uint32 [J, ceil(card / 32)], bit e of row j = ok[j, e]."""
    j, card = ok.shape
    nw = (card + 31) // 32
    padded = np.zeros((j, nw * 32), dtype=bool)
    padded[:, :card] = ok
    return np.packbits(padded, axis=1, bitorder="little").view("<u4").reshape(j, nw)


def _bitset_from_csr(any_mask: np.ndarray, ptr: np.ndarray, ids: np.ndarray, n_words: int,
                     pass_bits: Sequence[int] = ()) -> np.ndarray:
    """[J, n_words] uint32 with one bit per accepted code; ``any`` rows are all ones."""
    j = len(any_mask)
    out = np.zeros((j, n_words), dtype=np.uint32)
    if len(ids):
        owner = np.repeat(np.arange(j), np.diff(ptr))
        ok = ids >= 0
        o, i = owner[ok], ids[ok].astype(np.int64)
        np.bitwise_or.at(out, (o, i >> 5), (np.uint32(1) << (i & 31).astype(np.uint32)))
    for b in pass_bits:
        out[:, b >> 5] |= np.uint32(1) << np.uint32(b & 31)
    out[any_mask] = np.uint32(0xFFFFFFFF)
    return out


@dataclass
class PackedJobs:
    records: np.ndarray            # uint32 [J, rec_words], in the caller's job order
    layout: Dict[str, Any]         # n_used, kind[], card[], flags[], aux[], rec_off[], rec_words
    luts: Dict[int, np.ndarray]    # column -> uint16 remap, uploaded per call
    order: Optional[np.ndarray] = None  # uint32 [J] evaluation order (dsm_layout.job_order), None = as given


DEFAULT_DISK_MIN_GIB = 100.0  # DEFAULT_DISK = DiskSpec(size=Range(min=100GB)), core/models/resources.py:250


def disk_request_bounds(disk: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """(lo, hi) of the disk range a job brings to ``choose_disk_size_mib``: its own range, or
    ``DEFAULT_DISK.size`` when it states none (offers.py:170); open ends as -inf / +inf."""
    d = np.asarray(disk, dtype=np.float64)
    unset = np.isnan(d).all(axis=1)
    lo = np.where(np.isnan(d[:, 0]), -np.inf, d[:, 0])
    hi = np.where(np.isnan(d[:, 1]), np.inf, d[:, 1])
    return np.where(unset, DEFAULT_DISK_MIN_GIB, lo), np.where(unset, np.inf, hi)


def _range_trivial(words: np.ndarray, n_values: int) -> np.ndarray:
    """Jobs whose packed range accepts every one of ``n_values`` ordinals."""
    if n_values == 0:
        return np.ones(len(words), bool)
    return ((words & 0xFFFF) == 0) & ((words >> 16) >= n_values - 1)


def pack_jobs(t: EncodedTable, jc: JobColumns, n_used: Optional[int] = None) -> PackedJobs:
    """``n_used``: keep exactly that many leading columns (at least the ones some job filters on);
    default: up to the last column any job filters on."""
    j = jc.n
    luts: Dict[int, np.ndarray] = {}
    # per table column: (record words [J, w], kind, card, flags, aux, jobs that accept every entry)
    preds: Dict[str, Tuple[np.ndarray, int, int, int, int, np.ndarray]] = {}

    # words 0,1: row cut-offs (max_price and exclude_not_available); padding rows lie behind them
    mp = np.asarray(jc.max_price, dtype=np.float64)
    no_mp = np.isnan(mp)
    mp0 = np.where(no_mp, 0.0, mp)
    cut_a = np.where(no_mp, t.n_avail, np.searchsorted(t.price_a, mp0, side="right"))
    cut_b = np.where(no_mp, t.split_row + t.n_rest, t.split_row + np.searchsorted(t.price_b, mp0, side="right"))
    cut_b = np.where(np.asarray(jc.exclude_not_available, dtype=bool), t.split_row, cut_b)
    cuts = np.stack([cut_a, cut_b], axis=1).astype(np.uint32)

    def put(name: str, words: np.ndarray, kind: int, card: int, trivial: np.ndarray, flags: int = 0,
            ax: int = 0) -> None:
        words = np.ascontiguousarray(words, dtype=np.uint32).reshape(j, -1)
        preds[name] = (words, kind, card, flags, ax, np.asarray(trivial, dtype=bool))

    def put_set(name: str, any_mask, ptr, ids, unknown_passes: bool) -> None:
        col_idx = t.col(name)
        card = t.columns[col_idx].card  # ids 0..card-2, last = unknown
        any_mask = np.asarray(any_mask, dtype=bool)
        ids = np.asarray(ids, dtype=np.int64)
        if card > LUT_THRESHOLD:
            # per-call remap: class 0 = values no job mentions, classes 1.. = mentioned values,
            # last class = unknown
            mentioned = np.unique(ids[ids >= 0])
            lut = np.zeros(card, dtype=np.uint16)
            lut[mentioned] = np.arange(1, len(mentioned) + 1, dtype=np.uint16)
            unk = len(mentioned) + 1
            lut[card - 1] = unk
            ids = np.where(ids >= 0, np.searchsorted(mentioned, np.maximum(ids, 0)) + 1, -1)
            card = unk + 1
            luts[col_idx] = lut
        nw = (card + 31) // 32
        bits = _bitset_from_csr(any_mask, ptr, ids, nw, pass_bits=[card - 1] if unknown_passes else [])
        put(name, bits, KIND_BITSET, card, any_mask)

    # explicit block counts (offers.py:102-118): b > 1 keeps rows with cpu % b == 0 and gpus % b == 0,
    # b < 0 stands for an invalid count (nothing is divisible); 0 = no block filter
    blocks = np.zeros(j, dtype=np.int64) if jc.blocks is None else np.asarray(jc.blocks, dtype=np.int64)
    if (blocks > 0xFFFF).any():
        raise ValueError("block counts above 65535 are not supported")
    if (blocks > 1).any() and len(t.gpu_entries) and int((t.gpu_entries[:, 2] >> 25).max()) >= 127:
        raise ValueError("an instance with more than 126 GPUs cannot be matched against an explicit block count")
    cpu_vals, mem_vals = t.cpu_values, t.memory_values
    cpu_w = _range_words(cpu_vals, jc.cpu)
    mem_w = _range_words(mem_vals, jc.memory)
    cpu_bits = None
    if (blocks != 0).any():
        # the cpu test becomes "in range and divisible": an irregular set, sent as a bitset
        e = np.arange(len(cpu_vals), dtype=np.uint32)[None, :]
        ok = (e >= (cpu_w & 0xFFFF)[:, None]) & (e <= (cpu_w >> 16)[:, None])
        vals = cpu_vals.astype(np.int64)[None, :]
        ok &= np.where(blocks[:, None] > 1, vals % np.maximum(blocks, 1)[:, None] == 0, blocks[:, None] == 0)
        cpu_bits = _pack_bool_rows(ok)
    cpu_triv = _range_trivial(cpu_w, len(cpu_vals)) & (blocks == 0)
    mem_triv = _range_trivial(mem_w, len(mem_vals))
    if t.pair_entries is not None:
        a_words = cpu_w[:, None] if cpu_bits is None else cpu_bits
        put("cpu_memory", np.concatenate([a_words, mem_w[:, None]], axis=1), KIND_PAIR, len(t.pair_entries),
            cpu_triv & mem_triv, flags=0 if cpu_bits is None else COLF_PAIR_A_BITSET, ax=a_words.shape[1])
    else:
        if cpu_bits is None:
            put("cpu", cpu_w, KIND_RANGE, len(cpu_vals), cpu_triv)
        else:
            put("cpu", cpu_bits, KIND_BITSET, len(cpu_vals), cpu_triv)
        put("memory", mem_w, KIND_RANGE, len(mem_vals), mem_triv)
    dcol = t.columns[t.col("disk")]
    disk_w = _range_words(dcol.values, jc.disk)
    if not t.disk_conf_ranges:
        put("disk", disk_w, KIND_RANGE, dcol.card, _range_trivial(disk_w, len(dcol.values)), COLF_NULL_PASSES)
    else:
        # entries: the real sizes (by range), one per configurable range (the job's disk range --
        # DEFAULT_DISK when it states none -- must overlap it), the always-passing null entry
        nv = len(dcol.values)
        e = np.arange(nv, dtype=np.uint32)[None, :]
        ok = np.zeros((j, dcol.card), dtype=bool)
        ok[:, :nv] = (e >= (disk_w & 0xFFFF)[:, None]) & (e <= (disk_w >> 16)[:, None])
        lo, hi = disk_request_bounds(jc.disk)
        for ci, (cmin, cmax) in enumerate(t.disk_conf_ranges):
            start = lo if np.isnan(cmin) else np.maximum(lo, cmin)
            end = hi if np.isnan(cmax) else np.minimum(hi, cmax)
            ok[:, nv + ci] = ~(start > end)
        ok[:, dcol.card - 1] = True
        put("disk", _pack_bool_rows(ok), KIND_BITSET, dcol.card, ok.all(axis=1))

    # accelerator sub-predicate
    gd = t.gpu_dict
    n_names = max(1, len(gd["names"]))
    nwn = (n_names + 31) // 32
    gflags = np.zeros(j, dtype=np.uint32)
    gm = np.asarray(jc.gpu_memory, dtype=np.float64)
    gflags |= np.where(~np.isnan(gm).all(axis=1), GPUF_MEM_BOUNDED, 0).astype(np.uint32)
    ccm = np.asarray(jc.cc_min, dtype=np.int64)
    gflags |= np.where(ccm >= 0, GPUF_CC_BOUNDED, 0).astype(np.uint32)
    gflags |= np.where(np.asarray(jc.gpu_name_any, dtype=bool), GPUF_NAME_ANY, 0).astype(np.uint32)
    cc_ord = np.searchsorted(gd["cc"], np.maximum(ccm, 0).astype(np.float64), side="left").astype(np.uint32)
    gv = np.asarray(jc.gpu_vendor, dtype=np.int64)
    vmask = np.where(gv == 0, 0xFF, 1 << gv).astype(np.uint32)
    name_bits = _bitset_from_csr(np.zeros(j, dtype=bool), jc.gpu_name_ptr, jc.gpu_name_ids, nwn)
    cnt_w = _range_words(gd["count"], jc.gpu_count)
    tot_w = _range_words(gd["total"], jc.gpu_total)
    gwords = np.concatenate([
        cnt_w[:, None],
        _range_words(gd["memory"], gm)[:, None],
        tot_w[:, None],
        (np.minimum(cc_ord, 0xFFFE) | (gflags << np.uint32(16)))[:, None],
        (vmask | (np.where(blocks > 1, blocks, 0).astype(np.uint32) << np.uint32(16)))[:, None],
        name_bits,
    ], axis=1)
    gpu_triv = (_range_trivial(cnt_w, len(gd["count"])) & _range_trivial(tot_w, len(gd["total"]))
                & (gflags == GPUF_NAME_ANY) & (gv == 0) & (blocks <= 1))
    put("gpu", gwords, KIND_GPU, t.columns[t.col("gpu")].card, gpu_triv, ax=nwn)

    # backend / spot / status / flags
    spot = np.asarray(jc.spot, dtype=np.int64)
    spot_mask = np.where(spot < 0, 3, np.where(spot == 0, 1, 2)).astype(np.uint32)
    st_mask = np.full(j, 0xFF, dtype=np.uint32) if jc.status_mask is None else np.asarray(jc.status_mask, dtype=np.uint32)
    req = np.zeros(j, dtype=np.uint32) if jc.flag_req is None else np.asarray(jc.flag_req, dtype=np.uint32)
    forbid = np.zeros(j, dtype=np.uint32) if jc.flag_forbid is None else np.asarray(jc.flag_forbid, dtype=np.uint32)
    src_m = np.asarray(jc.src_backend_mask, dtype=np.uint32)
    off_m = np.asarray(jc.off_backend_mask, dtype=np.uint32)
    if t.is_pool:
        # status and the dynamic flags are tested on the "state" column: entry = status * 16 + flags
        e = np.arange((NO_STATUS_CODE + 1) * 16, dtype=np.uint32)[None, :]
        dyn_req, dyn_forbid = (req & DYNAMIC_FLAGS)[:, None], (forbid & DYNAMIC_FLAGS)[:, None]
        ok = (((st_mask[:, None] >> (e >> 4)) & 1) != 0) & (((e & 15) & dyn_req) == dyn_req) & (((e & 15) & dyn_forbid) == 0)
        put("state", _pack_bool_rows(ok), KIND_BITSET, ok.shape[1], ok.all(axis=1))
        st_mask = np.full(j, 0xFF, dtype=np.uint32)
        req, forbid = req & ~np.uint32(DYNAMIC_FLAGS), forbid & ~np.uint32(DYNAMIC_FLAGS)
    bwords = np.stack([src_m, off_m, spot_mask | (st_mask << np.uint32(8)), req | (forbid << np.uint32(16))], axis=1)
    bk_triv = ((src_m == 0xFFFFFFFF) & (off_m == 0xFFFFFFFF) & (spot_mask == 3) & (st_mask == 0xFF)
               & (req == 0) & (forbid == 0))
    put("bk", bwords, KIND_BK, t.columns[t.col("bk")].card, bk_triv)

    put_set("region", jc.region_any, jc.region_ptr, jc.region_ids, unknown_passes=False)
    put_set("itype", jc.itype_any, jc.itype_ptr, jc.itype_ids, unknown_passes=True)
    if t.is_pool:
        ar = np.arange(j + 1)
        fany = np.ones(j, bool) if jc.fleet_any is None else jc.fleet_any
        fids = np.full(j, -1, np.int64) if jc.fleet_ids is None else jc.fleet_ids
        put_set("fleet", fany, ar, fids, unknown_passes=False)
        nany = np.ones(j, bool) if jc.name_any is None else jc.name_any
        nids = np.full(j, -1, np.int64) if jc.name_ids is None else jc.name_ids
        put_set("name", nany, ar, nids, unknown_passes=False)
        zany = np.ones(j, bool) if jc.zone_any is None else jc.zone_any
        zptr = np.zeros(j + 1, np.int64) if jc.zone_ptr is None else jc.zone_ptr
        zids = np.zeros(0, np.int64) if jc.zone_ids is None else jc.zone_ids
        put_set("zone", zany, zptr, zids, unknown_passes=True)

    # ---- records in table-column order; trailing columns on which no job of this call filters
    # are not part of the call at all
    names = [c.name for c in t.columns]
    last_needed = np.zeros(j, dtype=np.int64)  # per job: 1 + index of the last column that rejects anything
    for ci, name in enumerate(names):
        last_needed = np.where(preds[name][5], last_needed, ci + 1)
    needed = max(1, int(last_needed.max()) if j else 1)
    if n_used is None:
        n_used = needed
    elif not needed <= n_used <= len(names):
        raise ValueError(f"n_used={n_used}: the batch filters on {needed} columns, the table has {len(names)}")
    for ci in list(luts):
        if ci >= n_used:
            luts.pop(ci)
    kinds, cards, cflags, aux, offs, parts = [], [], [], [], [], [cuts]
    pos = 2
    for name in names[:n_used]:
        words, kind, card, flags, ax, _ = preds[name]
        kinds.append(kind); cards.append(card); cflags.append(flags); aux.append(ax); offs.append(pos)
        parts.append(words)
        pos += words.shape[1]
    rec_words = (pos + 3) // 4 * 4
    rec = np.zeros((j, rec_words), dtype=np.uint32)
    o = 0
    for part in parts:
        rec[:, o:o + part.shape[1]] = part
        o += part.shape[1]
    # ---- evaluation order: jobs that need the same leading columns side by side, so that whole
    # groups of 128 skip the gathers of the columns behind (only worth it beyond one group)
    order = None
    if j > 128 and (last_needed != last_needed[0]).any():
        order = np.argsort(last_needed, kind="stable").astype(np.uint32)
    layout = dict(n_used=len(kinds), kind=kinds, card=cards, flags=cflags, aux=aux, rec_off=offs,
                  rec_words=rec_words, role=[ROLES[n] for n in names[:n_used]])
    return PackedJobs(records=rec, layout=layout, luts=luts, order=order)


# raw jobs: the same translation on the device ------------------------------------------------


def _column_shapes(t: EncodedTable, cpu_as_set: bool):
    """Per table column: (kind, card, flags, aux, record words) of its predicate."""
    out = []
    n_cpu_words = (len(t.cpu_values) + 31) // 32
    for col in t.columns:
        if col.name == "cpu_memory":
            aux = n_cpu_words if cpu_as_set else 1
            out.append((KIND_PAIR, col.card, COLF_PAIR_A_BITSET if cpu_as_set else 0, aux, aux + 1))
        elif col.name == "cpu":
            out.append((KIND_BITSET, col.card, 0, 0, n_cpu_words) if cpu_as_set else (KIND_RANGE, col.card, 0, 0, 1))
        elif col.name == "memory":
            out.append((KIND_RANGE, col.card, 0, 0, 1))
        elif col.name == "disk":  # (a table with configurable-disk entries is packed on the host: raw_layout)
            out.append((KIND_RANGE, col.card, COLF_NULL_PASSES, 0, 1))
        elif col.name == "gpu":
            nwn = (max(1, len(t.gpu_dict["names"])) + 31) // 32
            out.append((KIND_GPU, col.card, 0, nwn, 5 + nwn))
        elif col.name == "bk":
            out.append((KIND_BK, col.card, 0, 0, 4))
        else:
            out.append((KIND_BITSET, col.card, 0, 0, (col.card + 31) // 32))
    return out


def raw_layout(t: EncodedTable, jc: JobColumns):
    """The record layout of a batch whose records the device builds (``dsm_match_topk_raw`` /
    ``dsm_match_topk_cols``), or ``None`` when the batch filters on a column that needs a per-call
    remap (more than ``LUT_THRESHOLD`` dictionary entries: instance types or instance names) --
    those batches are packed on the host.  Trailing set columns no job filters on are left out;
    whether the remaining columns reject anything is found out on the device, per job group."""
    j = jc.n
    any_of = {"region": jc.region_any, "itype": jc.itype_any, "fleet": jc.fleet_any, "name": jc.name_any,
              "zone": jc.zone_any}
    names = [c.name for c in t.columns]
    n_used = len(names)
    while n_used > 1 and names[n_used - 1] in any_of and \
            (any_of[names[n_used - 1]] is None or bool(np.asarray(any_of[names[n_used - 1]]).all())):
        n_used -= 1
    if any(c.card > LUT_THRESHOLD for c in t.columns[:n_used]) or t.disk_conf_ranges:
        return None
    cpu_as_set = False
    if jc.blocks is not None:
        blocks = np.asarray(jc.blocks)
        cpu_as_set = bool((blocks != 0).any())
        if (blocks > 0xFFFF).any():
            raise ValueError("block counts above 65535 are not supported")
        if (blocks > 1).any() and len(t.gpu_entries) and int((t.gpu_entries[:, 2] >> 25).max()) >= 127:
            raise ValueError("an instance with more than 126 GPUs cannot be matched against an explicit block count")
    shapes = _column_shapes(t, cpu_as_set)[:n_used]
    offs, pos = [], 2
    for s in shapes:
        offs.append(pos)
        pos += s[4]
    return dict(n_used=n_used, kind=[s[0] for s in shapes], card=[s[1] for s in shapes],
                flags=[s[2] for s in shapes], aux=[s[3] for s in shapes], rec_off=offs,
                rec_words=(pos + 3) // 4 * 4, role=[ROLES[n] for n in names[:n_used]])


def raw_jobs(t: EncodedTable, jc: JobColumns):
    """Job filters as ``dsm_raw_job`` records for ``dsm_match_topk_raw`` (the device does what
    :func:`pack_jobs` does: bounds -> ordinals, max_price -> cut-off rows, sets -> bitsets, and the
    evaluation order).  Returns ``(raw, ids, layout)`` or ``None`` (see :func:`raw_layout`).  The
    product path hands the columns to ``dsm_match_topk_cols`` instead, which lays these records out
    in C; this function is the same layout in numpy (tests, callers of the raw entry point)."""
    layout = raw_layout(t, jc)
    if layout is None:
        return None
    j = jc.n
    ones = np.ones(j, bool)
    any_of = {"region": np.asarray(jc.region_any, bool), "itype": np.asarray(jc.itype_any, bool),
              "fleet": ones if jc.fleet_any is None else np.asarray(jc.fleet_any, bool),
              "name": ones if jc.name_any is None else np.asarray(jc.name_any, bool),
              "zone": ones if jc.zone_any is None else np.asarray(jc.zone_any, bool)}
    blocks = np.zeros(j, dtype=np.int64) if jc.blocks is None else np.asarray(jc.blocks, dtype=np.int64)

    # dsm_raw_job is 13 doubles followed by 16 32-bit words; both blocks are filled column-wise in
    # contiguous scratch arrays and transposed into the records in two copies (assigning the
    # fields of a structured array one by one costs 4x as much for a few thousand jobs)
    f = np.empty((13, j), dtype=np.float64)
    f[0] = jc.max_price
    for i, name in enumerate(("cpu", "memory", "disk", "gpu_count", "gpu_memory", "gpu_total")):
        f[1 + 2 * i:3 + 2 * i] = getattr(jc, name).T
    u = np.empty((16, j), dtype=np.uint32)
    i32 = u.view(np.int32)
    i32[0] = jc.cc_min
    i32[1] = blocks
    u[2] = jc.src_backend_mask
    u[3] = jc.off_backend_mask
    if jc.flag_req is None and jc.flag_forbid is None:
        u[4] = 0
    else:
        req = 0 if jc.flag_req is None else np.asarray(jc.flag_req, dtype=np.uint32)
        forbid = 0 if jc.flag_forbid is None else np.asarray(jc.flag_forbid, dtype=np.uint32)
        u[4] = req | (forbid << np.uint32(16))
    # byte 0 status mask, 1 spot, 2 vendor, 3 flags
    b4 = u[5].view(np.uint8).reshape(j, 4)
    b4[:, 0] = 0xFF if jc.status_mask is None else jc.status_mask
    b4[:, 1] = np.asarray(jc.spot).view(np.uint8) if np.asarray(jc.spot).dtype == np.int8 else np.asarray(jc.spot, np.int8).view(np.uint8)
    b4[:, 2] = jc.gpu_vendor
    fl = np.asarray(jc.exclude_not_available, bool).view(np.uint8) * np.uint8(RAWF_EXCLUDE_NOT_AVAILABLE)
    fl = fl | np.asarray(jc.gpu_name_any, bool).view(np.uint8) * np.uint8(RAWF_GPU_NAME_ANY)
    for name, bit in (("region", RAWF_REGION_ANY), ("itype", RAWF_ITYPE_ANY), ("fleet", RAWF_FLEET_ANY),
                      ("name", RAWF_NAME_ANY), ("zone", RAWF_ZONE_ANY)):
        a = any_of[name]
        fl = fl | (np.uint8(bit) if a is ones else a.view(np.uint8) * np.uint8(bit))
    b4[:, 3] = fl
    i32[6] = -1 if jc.fleet_ids is None else jc.fleet_ids
    i32[7] = -1 if jc.name_ids is None else jc.name_ids
    parts, base = [], 0
    zptr = None if jc.zone_ptr is None else np.asarray(jc.zone_ptr)
    for i, (ptr, ids) in enumerate(((jc.gpu_name_ptr, jc.gpu_name_ids), (jc.region_ptr, jc.region_ids),
                                    (jc.itype_ptr, jc.itype_ids), (zptr, jc.zone_ids))):
        if ptr is None or len(ids) == 0:
            u[8 + 2 * i] = base
            u[9 + 2 * i] = 0
            continue
        u[8 + 2 * i] = ptr[:-1]
        if base:
            u[8 + 2 * i] += np.uint32(base)
        u[9 + 2 * i] = ptr[1:] - ptr[:-1]
        parts.append(ids)
        base += len(ids)
    words = np.empty((j, RAW_JOB.itemsize // 4), dtype=np.uint32)
    words[:, :26].view(np.float64)[:] = f.T  # 8-byte aligned: rows are 168 bytes
    words[:, 26:] = u.T
    raw = words.view(RAW_JOB).reshape(j)
    ids_all = np.concatenate(parts).astype(np.int32) if base else np.zeros(0, np.int32)
    return raw, ids_all, layout


def pack_dicts(t: EncodedTable) -> Tuple[np.ndarray, List[int]]:
    """The table's sorted value lists in ``dsm_pack_dict`` order, for ``dsm_table_set_pack_dicts``."""
    gd = t.gpu_dict
    lists = [t.cpu_values, t.memory_values, t.columns[t.col("disk")].values, gd["count"], gd["memory"], gd["total"],
             gd["cc"]]
    lists = [np.asarray(x, dtype=np.float64) for x in lists]
    return (np.concatenate(lists) if sum(len(x) for x in lists) else np.zeros(0)), [len(x) for x in lists]


def state_code(status_code: int, flags: int) -> int:
    """Code of a pool table's "state" column for an instance status and its FLAG_* bits."""
    return int(status_code) * 16 + (int(flags) & DYNAMIC_FLAGS)
