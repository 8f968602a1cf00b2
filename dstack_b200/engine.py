"""Device-resident offer tables and the batched matcher on top of ``libdsmatch.so``.

``DeviceTable``  an :class:`~dstack_b200.encode.EncodedTable` uploaded to one GPU.
``Matcher``      a context (CUDA stream + scratch) that runs job batches against tables and
                 maps the rows that come back to positions in the caller's source list.
"""

from __future__ import annotations

import ctypes as C
import threading
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import native
from .encode import (EncodedTable, JobColumns, OfferColumns, PackedJobs, encode_table, pack_dicts, pack_jobs,
                     raw_jobs, raw_layout)

NO_ROW = native.NO_ROW


@dataclass
class TopK:
    """Per job: source positions of the first ``k`` matches in reference order (-1 = none),
    number of matches, and the highest price among all matches (NaN if none)."""

    index: np.ndarray      # int64 [J, k]
    count: np.ndarray      # int64 [J]
    max_price: np.ndarray  # float64 [J]


class Matcher:
    def __init__(self, device: int = 0, stream: Optional[int] = None):
        self._lib = native.lib()
        h = C.c_void_p()
        native.check(self._lib.dsm_ctx_create(device, C.c_void_p(stream or 0), C.byref(h)))
        self._h = h
        self.device = device
        # One shared Matcher serves every drop-in function of a process (default_matcher), and the
        # reference calls this path from a thread pool: the per-call remap upload and the match that
        # reads it must not interleave with another thread's.
        self._call_lock = threading.Lock()

    # -- lifecycle
    def close(self) -> None:
        if self._h is not None:
            self._lib.dsm_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self) -> C.c_void_p:
        if self._h is None:
            raise RuntimeError("matcher is closed")
        return self._h

    def launch_count(self) -> int:
        return int(self._lib.dsm_launch_count(self.handle))

    def sync(self) -> None:
        native.check(self._lib.dsm_sync(self.handle))

    # -- tables
    def upload(self, table: EncodedTable) -> "DeviceTable":
        return DeviceTable(self, table)

    # -- helpers
    def _layout(self, packed: PackedJobs, job_order: Optional[int] = None) -> native.Layout:
        """Upload the per-call remaps and describe the records.  ``job_order``: address of the
        evaluation order where the records of the call will live (see ``native.make_layout``)."""
        lut_ptrs: Dict[int, int] = {}
        for col, lut in packed.luts.items():
            lut = np.ascontiguousarray(lut, dtype=np.uint16)
            out = C.c_void_p()
            native.check(self._lib.dsm_lut_upload(self.handle, col, native.ptr(lut), len(lut), C.byref(out)))
            lut_ptrs[col] = out.value
        return native.make_layout(packed.layout, lut_ptrs, job_order)

    def plan(self, dt: "DeviceTable", packed: PackedJobs, k: int = 16) -> Dict[str, int]:
        """Launch geometry the library would pick for this batch (``k=0``: bitmap mode)."""
        out = (C.c_uint32 * 8)()
        lay = native.make_layout(packed.layout)
        native.check(self._lib.dsm_plan_query(self.handle, dt.handle, C.byref(lay), len(packed.records), k,
                                              C.byref(out)))
        keys = ["grid", "units", "pieces_a", "pieces_b", "slots", "smem", "entries", "groups"]
        return dict(zip(keys, [int(x) for x in out]))

    # -- matching with host buffers (H2D + kernels + D2H inside the call)
    def topk_rows(self, dt: "DeviceTable", packed: PackedJobs, k: int,
                  out: Optional[Tuple[np.ndarray, np.ndarray, np.ndarray]] = None):
        """Raw result in physical-row space: (idx [J,k] u32, cnt [J] u32, max_row [J] u32)."""
        j = len(packed.records)
        rec = np.ascontiguousarray(packed.records, dtype=np.uint32)
        if out is None:
            out = (np.empty((j, k), np.uint32), np.empty(j, np.uint32), np.empty(j, np.uint32))
        order = None if packed.order is None else np.ascontiguousarray(packed.order, dtype=np.uint32)
        with self._call_lock:  # the remaps live in per-context buffers until the match has run
            lay = self._layout(packed, native.ptr(order))
            native.check(self._lib.dsm_match_topk(self.handle, dt.handle, C.byref(lay), native.ptr(rec), j, k,
                                                  native.ptr(out[0]), native.ptr(out[1]), native.ptr(out[2])))
        return out

    def topk_rows_raw(self, dt: "DeviceTable", raw, k: int):
        """``topk_rows`` from raw jobs (``encode.raw_jobs``): records and evaluation order are built
        on the device (``dsm_match_topk_raw``)."""
        jobs, ids, layout = raw
        j = len(jobs)
        out = (np.empty((j, k), np.uint32), np.empty(j, np.uint32), np.empty(j, np.uint32))
        lay = native.make_layout(layout)
        native.check(self._lib.dsm_match_topk_raw(self.handle, dt.handle, C.byref(lay), native.ptr(jobs), j,
                                                  native.ptr(ids), len(ids), k,
                                                  native.ptr(out[0]), native.ptr(out[1]), native.ptr(out[2])))
        return out

    def match_lists(self, dt: "DeviceTable", packed: PackedJobs, by_source: bool) -> List[np.ndarray]:
        """Per job the source positions of every match, ascending (``by_source``) or in table
        order; the bitmaps are turned into lists on the device (``dsm_match_indices``)."""
        j = len(packed.records)
        rec = np.ascontiguousarray(packed.records, dtype=np.uint32)
        order = None if packed.order is None else np.ascontiguousarray(packed.order, dtype=np.uint32)
        ptr = np.empty(j + 1, np.uint64)
        out = C.c_void_p()
        with self._call_lock:
            lay = self._layout(packed, native.ptr(order))
            native.check(self._lib.dsm_match_indices(self.handle, dt.handle, C.byref(lay), native.ptr(rec), j,
                                                     1 if by_source else 0, native.ptr(ptr), C.byref(out)))
            total = int(ptr[j])
            if total:
                buf = (C.c_uint32 * total).from_address(out.value)
                ids = np.frombuffer(buf, dtype=np.int32).copy()  # copy: the buffer is the context's
            else:
                ids = np.zeros(0, np.int32)
        return np.split(ids, ptr[1:-1].astype(np.int64))

    def bitmap_rows(self, dt: "DeviceTable", packed: PackedJobs) -> np.ndarray:
        j = len(packed.records)
        rec = np.ascontiguousarray(packed.records, dtype=np.uint32)
        bits = np.empty((j, dt.table.n_rows // 32), np.uint32)
        order = None if packed.order is None else np.ascontiguousarray(packed.order, dtype=np.uint32)
        with self._call_lock:
            lay = self._layout(packed, native.ptr(order))
            native.check(self._lib.dsm_match_bitmap(self.handle, dt.handle, C.byref(lay), native.ptr(rec), j,
                                                    native.ptr(bits)))
        return bits

    # -- results in source space
    def match_topk(self, dt: "DeviceTable", jobs: JobColumns, k: int) -> TopK:
        t = dt.table
        if t.n_src == 0:
            return TopK(np.full((jobs.n, k), -1, np.int64), np.zeros(jobs.n, np.int64), np.full(jobs.n, np.nan))
        layout = raw_layout(t, jobs)
        if layout is not None:  # records, evaluation order and the answer in source terms: all in the library
            src, cnt, price = np.empty((jobs.n, k), np.int32), np.empty(jobs.n, np.uint32), np.empty(jobs.n, np.float64)
            lay = native.make_layout(layout)
            cols, keep = native.job_columns_c(jobs)
            native.check(self._lib.dsm_match_topk_cols(self.handle, dt.handle, C.byref(lay), C.byref(cols), jobs.n, k,
                                                       native.ptr(src), native.ptr(cnt), native.ptr(price)))
            del keep
            return TopK(src.astype(np.int64), cnt.astype(np.int64), price)
        # a column needs a per-call remap: packed here
        idx, cnt, mx = self.topk_rows(dt, pack_jobs(t, jobs), k)
        src = np.where(idx == NO_ROW, -1, t.perm[np.minimum(idx, t.n_rows - 1)])
        price = np.where(mx == NO_ROW, np.nan, t.price[np.minimum(mx, t.n_rows - 1)])
        return TopK(src.astype(np.int64), cnt.astype(np.int64), price)

    def match_all(self, dt: "DeviceTable", jobs: JobColumns, order: str = "source") -> List[np.ndarray]:
        """Every match per job as source positions, in ``"source"`` (input) order or in
        ``"rank"`` (reference result) order."""
        t = dt.table
        if t.n_src == 0:
            return [np.zeros(0, np.int64) for _ in range(jobs.n)]
        return self.match_lists(dt, pack_jobs(t, jobs), by_source=(order == "source"))


class DeviceTable:
    def __init__(self, matcher: Matcher, table: EncodedTable):
        self.table = table
        self._lib = native.lib()
        self._matcher = matcher
        n_pairs = 0 if table.pair_entries is None else len(table.pair_entries)
        desc = native.TableDesc(table.n_rows, table.split_row, table.n_cols,
                                len(table.gpu_entries), len(table.bk_entries), n_pairs,
                                table.n_avail, table.n_rest)
        rows = np.ascontiguousarray(table.rows, dtype=np.uint16)
        price = np.ascontiguousarray(table.price, dtype=np.float64)
        ge = np.ascontiguousarray(table.gpu_entries, dtype=np.uint32)
        be = np.ascontiguousarray(table.bk_entries, dtype=np.uint32)
        pe = None if table.pair_entries is None else np.ascontiguousarray(table.pair_entries, dtype=np.uint32)
        h = C.c_void_p()
        native.check(self._lib.dsm_table_create(matcher.handle, C.byref(desc), native.ptr(rows), native.ptr(price),
                                                native.ptr(ge), native.ptr(be), native.ptr(pe), C.byref(h)))
        self._h = h
        # what the device needs to build job records itself and to answer in source positions
        values, counts = pack_dicts(table)
        values = np.ascontiguousarray(values, dtype=np.float64)
        native.check(self._lib.dsm_table_set_pack_dicts(matcher.handle, h, native.ptr(values),
                                                        C.byref((C.c_uint32 * native.N_PACK_DICTS)(*counts))))
        to_source = np.ascontiguousarray(table.perm, dtype=np.int32)
        native.check(self._lib.dsm_table_set_source_map(matcher.handle, h, native.ptr(to_source), max(1, table.n_src)))

    @property
    def handle(self) -> C.c_void_p:
        if self._h is None:
            raise RuntimeError("table is closed")
        return self._h

    def patch_rows(self, rows: np.ndarray) -> None:
        """Copy the host table's current codes of physical rows ``rows`` to the device (after they
        were changed in ``self.table.rows``): an in-place update, no re-upload of the table."""
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        if len(rows) == 0:
            return
        codes = np.ascontiguousarray(self.table.rows[rows], dtype=np.uint16)
        native.check(self._lib.dsm_table_patch_rows(self._matcher.handle, self.handle, native.ptr(rows), len(rows),
                                                    native.ptr(codes)))

    def close(self) -> None:
        if self._h is not None:
            self._lib.dsm_table_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default: Dict[int, Matcher] = {}
_default_lock = threading.Lock()


def default_matcher(device: int = 0) -> Matcher:
    """One shared matcher per device for the drop-in functions."""
    with _default_lock:
        m = _default.get(device)
        if m is None:
            m = _default[device] = Matcher(device)
        return m
