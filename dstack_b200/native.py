"""ctypes binding of ``libdsmatch.so`` (C ABI in ``include/dsmatch.h``).

There is no CPU fallback: if the shared library is missing or a call fails, an exception is
raised.  ctypes releases the GIL for the duration of every foreign call, so several host
threads can drive their own contexts concurrently (the reference calls this path from a
thread pool and from interleaved asyncio tasks, SURVEY.md section 8b "Threading").
"""

from __future__ import annotations

import ctypes as C
import os
from typing import Any, Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DSM_LIB") or os.path.join(_HERE, "libdsmatch.so")
MAX_COLS = 16
NO_ROW = 0xFFFFFFFF

EXPORTS = [
    "dsm_ctx_create", "dsm_ctx_destroy", "dsm_last_error", "dsm_build_info", "dsm_table_create",
    "dsm_table_destroy", "dsm_lut_upload", "dsm_match_topk", "dsm_match_bitmap",
    "dsm_match_topk_dev", "dsm_match_bitmap_dev", "dsm_merge_shards_dev", "dsm_gather_price_dev",
    "dsm_sync", "dsm_host_alloc", "dsm_host_free", "dsm_launch_count", "dsm_plan_query",
    "dsm_timing_enable", "dsm_timing_read", "dsm_match_bitmap_static", "dsm_merge_shards_strided_dev",
    "dsm_merge_shards_interleaved_dev", "dsm_timing_read_parts", "dsm_table_set_pack_dicts",
    "dsm_table_set_source_map", "dsm_match_topk_raw", "dsm_jobs_pack_dev", "dsm_match_indices",
    "dsm_match_topk_raw_src", "dsm_match_topk_cols", "dsm_reserve",
    "dsm_table_patch_rows",
    "dsm_xchg_create", "dsm_xchg_local_base", "dsm_xchg_connect_ipc", "dsm_xchg_connect_ptrs",
    "dsm_xchg_all_gather_dev", "dsm_xchg_status", "dsm_xchg_destroy",
]
IPC_HANDLE_BYTES = 64  # DSM_IPC_HANDLE_BYTES
N_PACK_DICTS = 7  # cpu, memory, disk, gpu count, gpu memory, gpu total, compute capability

# dsm_raw_job (include/dsmatch.h): little-endian, 168 bytes
RAW_JOB = np.dtype([
    ("max_price", "<f8"), ("cpu", "<f8", 2), ("memory", "<f8", 2), ("disk", "<f8", 2), ("gpu_count", "<f8", 2),
    ("gpu_memory", "<f8", 2), ("gpu_total", "<f8", 2), ("cc_min", "<i4"), ("blocks", "<i4"),
    ("src_backend_mask", "<u4"), ("off_backend_mask", "<u4"), ("flag_req", "<u2"), ("flag_forbid", "<u2"),
    ("status_mask", "u1"), ("spot", "i1"), ("gpu_vendor", "i1"), ("flags", "u1"), ("fleet_id", "<i4"),
    ("name_id", "<i4"), ("gpu_name_off", "<u4"), ("gpu_name_cnt", "<u4"), ("region_off", "<u4"),
    ("region_cnt", "<u4"), ("itype_off", "<u4"), ("itype_cnt", "<u4"), ("zone_off", "<u4"), ("zone_cnt", "<u4"),
])
RAWF_EXCLUDE_NOT_AVAILABLE, RAWF_GPU_NAME_ANY, RAWF_REGION_ANY, RAWF_ITYPE_ANY = 1, 2, 4, 8
RAWF_FLEET_ANY, RAWF_NAME_ANY, RAWF_ZONE_ANY = 16, 32, 64
ROLES = {"cpu": 1, "memory": 2, "cpu_memory": 3, "disk": 4, "gpu": 5, "bk": 6, "region": 7, "itype": 8,
         "fleet": 9, "name": 10, "zone": 11, "state": 12}
STATIC_SCRATCH_BYTES = 2 << 20  # DSM_STATIC_SCRATCH_BYTES


class DsmError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libdsmatch error {code}: {message}")
        self.code = code


class TableDesc(C.Structure):
    _fields_ = [("n_rows", C.c_uint32), ("split_row", C.c_uint32), ("n_cols", C.c_uint32),
                ("n_gpu_entries", C.c_uint32), ("n_bk_entries", C.c_uint32), ("n_pair_entries", C.c_uint32),
                ("n_live_a", C.c_uint32), ("n_live_b", C.c_uint32)]


class Layout(C.Structure):
    _fields_ = [("n_used", C.c_uint32),
                ("kind", C.c_uint32 * MAX_COLS), ("card", C.c_uint32 * MAX_COLS),
                ("flags", C.c_uint32 * MAX_COLS), ("aux", C.c_uint32 * MAX_COLS),
                ("rec_off", C.c_uint32 * MAX_COLS), ("rec_words", C.c_uint32),
                ("lut", C.c_void_p * MAX_COLS), ("job_order", C.c_void_p), ("role", C.c_uint32 * MAX_COLS)]


class JobColumnsC(C.Structure):
    """``dsm_job_columns``: one host pointer per attribute of a job batch."""

    FIELDS = [("max_price", "f8"), ("cpu", "f8"), ("memory", "f8"), ("disk", "f8"), ("gpu_count", "f8"),
              ("gpu_memory", "f8"), ("gpu_total", "f8"), ("cc_min", "i8"), ("blocks", "i8"),
              ("src_backend_mask", "u4"), ("off_backend_mask", "u4"), ("flag_req", "u2"), ("flag_forbid", "u2"),
              ("status_mask", "u1"), ("spot", "i1"), ("gpu_vendor", "i1"), ("exclude_not_available", "u1"),
              ("gpu_name_any", "u1"), ("region_any", "u1"), ("itype_any", "u1"), ("fleet_any", "u1"),
              ("name_any", "u1"), ("zone_any", "u1"), ("fleet_ids", "i8"), ("name_ids", "i8"),
              ("gpu_name_ptr", "i8"), ("gpu_name_ids", "i8"), ("region_ptr", "i8"), ("region_ids", "i8"),
              ("itype_ptr", "i8"), ("itype_ids", "i8"), ("zone_ptr", "i8"), ("zone_ids", "i8")]
    _fields_ = [(name, C.c_void_p) for name, _ in FIELDS]


_FIELD_DTYPES = None  # [(name, dtype)] of JobColumnsC.FIELDS, built on first use


def job_columns_c(jc: Any):
    """A ``dsm_job_columns`` over the arrays of a ``JobColumns`` (converted only where the dtype
    differs; bool arrays are passed as bytes).  Returns the struct and the arrays it points into,
    which must stay alive for the call."""
    global _FIELD_DTYPES
    if _FIELD_DTYPES is None:
        _FIELD_DTYPES = [(name, np.dtype(code)) for name, code in JobColumnsC.FIELDS]
    st, keep = JobColumnsC(), []
    u1, b1 = np.dtype("u1"), np.dtype(bool)
    for name, dt in _FIELD_DTYPES:
        a = getattr(jc, name, None)
        if a is None:
            continue
        if not isinstance(a, np.ndarray):
            a = np.asarray(a)
        if a.dtype == b1 and dt == u1:
            a = a.view(u1) if a.flags.c_contiguous else np.ascontiguousarray(a).view(u1)
        elif a.dtype != dt or not a.flags.c_contiguous:
            a = np.ascontiguousarray(a, dtype=dt)
        keep.append(a)
        if a.size:
            try:  # the address without numpy's ctypes helper object (a third of the cost; 33 fields a call)
                addr = C.addressof(C.c_char.from_buffer(a))
            except (TypeError, ValueError, BufferError):  # read-only arrays have no writable buffer
                addr = a.ctypes.data
            setattr(st, name, addr)
    return st, keep


_lib: Optional[C.CDLL] = None


def lib() -> C.CDLL:
    """Load the library once.  Raises if it has not been built (``python -c 'import
    __graft_entry__ as g; g.build()'``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not found: build the CUDA extension first "
                          "(__graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    L.dsm_last_error.restype = C.c_char_p
    L.dsm_build_info.restype = C.c_char_p
    L.dsm_launch_count.restype = u64
    L.dsm_launch_count.argtypes = [vp]
    L.dsm_ctx_create.argtypes = [C.c_int, vp, C.POINTER(vp)]
    L.dsm_ctx_destroy.argtypes = [vp]
    L.dsm_table_create.argtypes = [vp, C.POINTER(TableDesc), vp, vp, vp, vp, vp, C.POINTER(vp)]
    L.dsm_table_destroy.argtypes = [vp]
    L.dsm_lut_upload.argtypes = [vp, u32, vp, u32, C.POINTER(vp)]
    L.dsm_match_topk.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, u32, vp, vp, vp]
    L.dsm_match_bitmap.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, vp]
    L.dsm_match_topk_dev.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, u32, vp, vp, vp]
    L.dsm_match_bitmap_dev.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, vp]
    L.dsm_match_bitmap_static.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, vp]
    L.dsm_merge_shards_dev.argtypes = [vp, u32, u32, u32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.dsm_merge_shards_strided_dev.argtypes = [vp, u32, u32, u32, vp, vp, vp, vp, C.c_size_t, vp, vp, vp, vp, vp]
    L.dsm_merge_shards_interleaved_dev.argtypes = [vp, u32, u32, u32, vp, vp, vp, vp, C.c_size_t, vp, u32,
                                                   vp, vp, vp, vp]
    L.dsm_gather_price_dev.argtypes = [vp, vp, vp, u32, vp]
    L.dsm_xchg_create.argtypes = [vp, u32, u32, u64, C.POINTER(vp), vp]
    L.dsm_xchg_local_base.argtypes = [vp, C.POINTER(vp)]
    L.dsm_xchg_connect_ipc.argtypes = [vp, vp]
    L.dsm_xchg_connect_ptrs.argtypes = [vp, C.POINTER(vp)]
    L.dsm_xchg_all_gather_dev.argtypes = [vp, vp, u64, C.POINTER(vp), C.POINTER(u64)]
    L.dsm_xchg_status.argtypes = [vp, C.POINTER(u32)]
    L.dsm_xchg_destroy.argtypes = [vp]
    L.dsm_sync.argtypes = [vp]
    L.dsm_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    L.dsm_host_free.argtypes = [vp]
    L.dsm_timing_enable.argtypes = [vp, C.c_int]
    L.dsm_timing_read.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(u32)]
    L.dsm_timing_read_parts.argtypes = [vp, C.POINTER(C.c_double * 3), C.POINTER(u32)]
    L.dsm_table_set_pack_dicts.argtypes = [vp, vp, vp, C.POINTER(u32 * N_PACK_DICTS)]
    L.dsm_table_set_source_map.argtypes = [vp, vp, vp, u32]
    L.dsm_match_topk_raw.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, vp, u32, u32, vp, vp, vp]
    L.dsm_match_topk_raw_src.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, vp, u32, u32, vp, vp, vp]
    L.dsm_match_topk_cols.argtypes = [vp, vp, C.POINTER(Layout), C.POINTER(JobColumnsC), u32, u32, vp, vp, vp]
    L.dsm_table_patch_rows.argtypes = [vp, vp, vp, u32, vp]
    L.dsm_reserve.argtypes = [vp, vp, C.POINTER(Layout), u32, u32]
    L.dsm_jobs_pack_dev.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, vp, vp, vp]
    L.dsm_match_indices.argtypes = [vp, vp, C.POINTER(Layout), vp, u32, C.c_int, vp, C.POINTER(vp)]
    L.dsm_plan_query.argtypes = [vp, vp, C.POINTER(Layout), u32, u32, C.POINTER(u32 * 8)]
    for name in EXPORTS:
        fn = getattr(L, name)
        if name not in ("dsm_last_error", "dsm_build_info", "dsm_launch_count"):
            fn.restype = C.c_int
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise DsmError(rc, lib().dsm_last_error().decode(errors="replace"))


def ptr(a: Optional[np.ndarray]) -> Optional[int]:
    return None if a is None else a.ctypes.data


def make_layout(d: Dict[str, Any], lut_ptrs: Optional[Dict[int, int]] = None,
                job_order: Optional[int] = None) -> Layout:
    """``job_order``: address of the evaluation order (u32 [n_jobs]) in the memory space of the
    call's job records (host for the host-buffer entry points, device for ``*_dev``), or None."""
    L = Layout()
    L.job_order = job_order
    L.n_used = d["n_used"]
    L.rec_words = d["rec_words"]
    roles = d.get("role")
    for i in range(d["n_used"]):
        if roles is not None:
            L.role[i] = roles[i]
        L.kind[i] = d["kind"][i]
        L.card[i] = d["card"][i]
        L.flags[i] = d["flags"][i]
        L.aux[i] = d["aux"][i]
        L.rec_off[i] = d["rec_off"][i]
    for i, p in (lut_ptrs or {}).items():
        L.lut[i] = p
    return L


class PinnedArray:
    """A numpy view over ``cudaHostAlloc`` memory (for the host-buffer entry points)."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        check(lib().dsm_host_alloc(max(self.nbytes, 16), C.byref(p)))
        self._p = p
        buf = (C.c_char * max(self.nbytes, 16)).from_address(p.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self) -> None:
        if self._p is not None:
            self.array = None
            lib().dsm_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
