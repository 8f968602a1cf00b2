"""numpy restatement of the path on columnar inputs.  TEST INFRASTRUCTURE (see oracle/__init__).

Works on the raw attribute VALUES of ``OfferColumns`` / ``JobColumns`` with float64
comparisons -- it does not use the encoder's ordinals, dictionaries' codes or cut-off rows,
so it is independent of ``dstack_b200.encode.encode_table`` / ``pack_jobs``.  It follows:

  * the predicate: ``oracle.gpuhunt_restated.matches`` applied to
    ``offer_to_catalog_item`` values (``core/backends/base/offers.py:86-109``);
  * backend / region / instance-type filters: ``server/services/offers.py:81-100``;
  * order: ``heapq.merge`` by price over price-sorted per-backend lists, then stable
    partition on availability (``server/services/backends/__init__.py:449-451``)
    == stable sort of the source list by ``(not available, price)``;
  * slicing: ``offers[:k]``, ``len``, ``max(price)`` (``server/services/runs.py:341-348``).

``tests/test_oracle_columnar.py`` checks it against the object-level ``oracle.reference_path``.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np

from .gpuhunt_restated import compute_capability_of  # the oracle's own table
from dstack_b200.encode import BACKEND_CODES, FLAG_AUTO_DIVISIBLE, JobColumns, OfferColumns


@dataclass
class TopK:
    index: np.ndarray      # int64 [J, k] source positions, -1 = none
    count: np.ndarray      # int64 [J]
    max_price: np.ndarray  # float64 [J], NaN = no match


def _between(v: np.ndarray, bounds: np.ndarray) -> np.ndarray:
    """[J, N]: not (v < lo) and not (v > hi), NaN bound = unbounded."""
    lo, hi = bounds[:, 0][:, None], bounds[:, 1][:, None]
    ok = np.ones((len(bounds), len(v)), bool)
    ok &= ~(~np.isnan(lo) & (v[None, :] < np.nan_to_num(lo)))
    ok &= ~(~np.isnan(hi) & (v[None, :] > np.nan_to_num(hi)))
    return ok


def _in_csr(any_mask, ptr, ids, values: np.ndarray, unknown_passes: bool = False) -> np.ndarray:
    j = len(any_mask)
    ok = np.zeros((j, len(values)), bool)
    for i in range(j):
        if any_mask[i]:
            ok[i] = True
        else:
            ok[i] = np.isin(values, ids[ptr[i]:ptr[i + 1]])
            if unknown_passes:
                ok[i] |= values < 0
    return ok


def lower_name_ids(oc: OfferColumns) -> Dict[str, int]:
    """lower-cased GPU name -> id in first-appearance order over the shapes (the id space
    ``JobColumns.gpu_name_ids`` refers to)."""
    ids: Dict[str, int] = {}
    for nm in oc.shapes.gpu_name:
        if nm is not None:
            ids.setdefault(nm.lower(), len(ids))
    return ids


def match_matrix(oc: OfferColumns, jc: JobColumns) -> np.ndarray:
    """bool [J, N] in SOURCE order."""
    sh, sid = oc.shapes, np.asarray(oc.shape_id, dtype=np.int64)
    n = len(oc)
    price = np.asarray(oc.price, dtype=np.float64)
    mp = np.asarray(jc.max_price, dtype=np.float64)[:, None]
    ok = ~(~np.isnan(mp) & (price[None, :] > np.nan_to_num(mp)))                      # 2
    spot = np.asarray(jc.spot)[:, None]
    ok &= (spot < 0) | (np.asarray(oc.spot, dtype=np.int64)[None, :] == spot)          # 3
    ok &= _between(np.asarray(sh.cpus, dtype=np.float64)[sid], jc.cpu)                 # 4
    ok &= _between(np.asarray(sh.memory_gib, dtype=np.float64)[sid], jc.memory)        # 5
    gcount = np.asarray(sh.gpu_count, dtype=np.int64)[sid]
    ok &= _between(gcount.astype(np.float64), jc.gpu_count)                            # 6
    vend = np.asarray(sh.gpu_vendor, dtype=np.int64)[sid]
    jv = np.asarray(jc.gpu_vendor, dtype=np.int64)[:, None]
    ok &= (jv == 0) | ((vend[None, :] != 0) & (vend[None, :] == jv))                   # 7
    names = lower_name_ids(oc)
    nid_shape = np.array([-1 if nm is None else names[nm.lower()] for nm in sh.gpu_name], dtype=np.int64)
    nid = nid_shape[sid]
    name_ok = _in_csr(np.zeros(jc.n, bool), jc.gpu_name_ptr, jc.gpu_name_ids, nid) & (nid >= 0)[None, :]
    ok &= np.asarray(jc.gpu_name_any, dtype=bool)[:, None] | name_ok                   # 8
    gmem = np.asarray(sh.gpu_memory_gib, dtype=np.float64)[sid]
    gm_bounded = ~np.isnan(jc.gpu_memory).all(axis=1)[:, None]
    gm_ok = (gcount > 0)[None, :] & ~np.isnan(gmem)[None, :] & _between(np.nan_to_num(gmem), jc.gpu_memory)
    ok &= ~gm_bounded | gm_ok                                                          # 9
    cc_shape = np.array([-1 if (c := compute_capability_of(nm)) is None else c[0] * 1000 + c[1]
                         for nm in sh.gpu_name], dtype=np.int64)
    cc = cc_shape[sid]
    ccm = np.asarray(jc.cc_min, dtype=np.int64)[:, None]
    ok &= (ccm < 0) | ((cc >= 0)[None, :] & (cc[None, :] >= ccm))                      # 10
    total = np.where((gcount > 0) & ~np.isnan(gmem), gcount.astype(np.float64) * np.nan_to_num(gmem), 0.0)
    ok &= _between(total, jc.gpu_total)                                                # 11
    disk = np.asarray(sh.disk_gib, dtype=np.float64)[sid]
    disk_ok = np.isnan(disk)[None, :] | _between(np.nan_to_num(disk), jc.disk)        # 12
    if oc.disk_conf_ranges and sh.disk_conf is not None:
        # an item without a disk size on a backend with configurable disks survives
        # catalog_item_to_offer only if the job's disk range (DEFAULT_DISK = 100GB.. when it states
        # none) overlaps the backend's range: core/backends/base/offers.py:63-66, 161-175
        d = np.asarray(jc.disk, dtype=np.float64)
        unset = np.isnan(d).all(axis=1)
        lo = np.where(unset, 100.0, np.where(np.isnan(d[:, 0]), -np.inf, d[:, 0]))
        hi = np.where(unset, np.inf, np.where(np.isnan(d[:, 1]), np.inf, d[:, 1]))
        conf = np.asarray(sh.disk_conf, dtype=np.int64)[sid]
        for ci, (cmin, cmax) in enumerate(oc.disk_conf_ranges):
            start = np.maximum(lo, -np.inf if np.isnan(cmin) else cmin)
            end = np.minimum(hi, np.inf if np.isnan(cmax) else cmax)
            rows = np.isnan(disk) & (conf == ci)
            disk_ok[:, rows] = ~(start > end)[:, None]
    ok &= disk_ok
    # backend / region / instance type (server/services/offers.py:81-100)
    src = np.asarray(oc.src_backend, dtype=np.int64)
    off = np.asarray(oc.backend, dtype=np.int64)
    ok &= ((np.asarray(jc.src_backend_mask, dtype=np.uint64)[:, None] >> src[None, :].astype(np.uint64)) & 1).astype(bool)
    ok &= ((np.asarray(jc.off_backend_mask, dtype=np.uint64)[:, None] >> off[None, :].astype(np.uint64)) & 1).astype(bool)
    ok &= _in_csr(jc.region_any, jc.region_ptr, jc.region_ids, np.asarray(oc.region_id, dtype=np.int64))
    ok &= _in_csr(jc.itype_any, jc.itype_ptr, jc.itype_ids, np.asarray(oc.itype_id, dtype=np.int64),
                  unknown_passes=True)
    ok &= ~(np.asarray(jc.exclude_not_available, dtype=bool)[:, None] & ~np.asarray(oc.available, dtype=bool)[None, :])
    return ok


def rank_order(oc: OfferColumns) -> np.ndarray:
    """Source positions in the order the reference returns offers."""
    return np.lexsort((np.asarray(oc.price, dtype=np.float64), ~np.asarray(oc.available, dtype=bool)))


def topk(oc: OfferColumns, jc: JobColumns, k: int, order: Optional[np.ndarray] = None) -> TopK:
    if order is None:
        order = rank_order(oc)
    price_ranked = np.asarray(oc.price, dtype=np.float64)[order]
    j = jc.n
    idx = np.full((j, k), -1, dtype=np.int64)
    cnt = np.zeros(j, dtype=np.int64)
    mx = np.full(j, np.nan)
    m = match_matrix(oc, jc)[:, order]
    for i in range(j):
        rows = np.flatnonzero(m[i])
        cnt[i] = len(rows)
        idx[i, :min(k, len(rows))] = order[rows[:k]]
        if len(rows):
            mx[i] = price_ranked[rows].max()
    return TopK(idx, cnt, mx)
