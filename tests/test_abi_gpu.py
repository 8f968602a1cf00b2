"""C-ABI behaviour on a GPU: error codes instead of crashes, per-context re-entrancy, k limits."""

import ctypes as C
import threading

import numpy as np
import pytest

from dstack_b200 import native
from workloads import synth
from dstack_b200.encode import encode_table, pack_jobs
from dstack_b200.engine import Matcher

pytestmark = pytest.mark.gpu

DSM_E_ARG, DSM_E_DICT_TOO_LARGE = -2, -3


def _setup(matcher, n_offers=5000, n_jobs=150, seed=51):
    cat = synth.make_catalog(n_offers, seed=seed, n_shapes=256)
    table = encode_table(cat.columns)
    packed = pack_jobs(table, synth.make_jobs(cat, table, n_jobs, seed=seed + 1).columns)
    return cat, table, packed


def test_bad_arguments_return_error_codes(matcher):
    lib = native.lib()
    cat, table, packed = _setup(matcher)
    dt = matcher.upload(table)
    try:
        lay = matcher._layout(packed)
        rec = np.ascontiguousarray(packed.records)
        j = len(rec)
        idx, cnt, mx = np.empty((j, 16), np.uint32), np.empty(j, np.uint32), np.empty(j, np.uint32)
        args = (matcher.handle, dt.handle, C.byref(lay), rec.ctypes.data)
        assert lib.dsm_match_topk(*args, 0, 16, idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        assert b"n_jobs" in lib.dsm_last_error()
        assert lib.dsm_match_topk(*args, j, 0, idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        assert lib.dsm_match_topk(*args, j, 5000, idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        assert lib.dsm_match_topk(matcher.handle, dt.handle, C.byref(lay), None, j, 16, idx.ctypes.data,
                                  cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        # a dictionary that cannot fit in shared memory is refused, not truncated
        huge = matcher._layout(packed)
        ranges = [c for c in range(huge.n_used) if huge.kind[c] == 1]  # a range predicate is one word whatever the card
        assert ranges, "the batch is expected to filter on disk"
        huge.card[ranges[0]] = 60000
        assert lib.dsm_match_topk(matcher.handle, dt.handle, C.byref(huge), rec.ctypes.data, j, 16, idx.ctypes.data,
                                  cnt.ctypes.data, mx.ctypes.data) == DSM_E_DICT_TOO_LARGE
        bad_kind = matcher._layout(packed)
        bad_kind.kind[0] = 9
        assert lib.dsm_match_topk(matcher.handle, dt.handle, C.byref(bad_kind), rec.ctypes.data, j, 16,
                                  idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        # and the context still works afterwards
        native.check(lib.dsm_match_topk(*args, j, 16, idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data))
        assert cnt.sum() > 0
        # a predicate that would reach past the end of the record is refused (not read out of bounds)
        bad_off = matcher._layout(packed)
        bad_off.rec_off[bad_off.n_used - 1] = bad_off.rec_words
        assert lib.dsm_match_topk(matcher.handle, dt.handle, C.byref(bad_off), rec.ctypes.data, j, 16,
                                  idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        # so is an evaluation order that is not a permutation
        bad_order = np.zeros(j, np.uint32)
        bad_lay = matcher._layout(packed, bad_order.ctypes.data)
        assert lib.dsm_match_topk(matcher.handle, dt.handle, C.byref(bad_lay), rec.ctypes.data, j, 16,
                                  idx.ctypes.data, cnt.ctypes.data, mx.ctypes.data) == DSM_E_ARG
        desc = native.TableDesc(33, 32, 8, 0, 0, 0, 32, 1)  # rows not a multiple of 32
        h = C.c_void_p()
        assert lib.dsm_table_create(matcher.handle, C.byref(desc), rec.ctypes.data, None, None, None, None,
                                    C.byref(h)) == DSM_E_ARG
    finally:
        dt.close()


def test_two_contexts_from_two_threads_agree(matcher):
    """The reference calls this path from a thread pool; each thread drives its own context."""
    cat, table, packed = _setup(matcher, n_offers=40_000, n_jobs=600, seed=61)
    want = None
    dt0 = matcher.upload(table)
    try:
        want = matcher.topk_rows(dt0, packed, 15)
    finally:
        dt0.close()
    results, errors = {}, []

    def worker(name):
        try:
            m = Matcher(0)
            dt = m.upload(table)
            for _ in range(5):
                results[name] = m.topk_rows(dt, packed, 15)
            dt.close()
            m.close()
        except Exception as e:  # pragma: no cover
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(3)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors
    for got in results.values():
        for a, b in zip(got, want):
            assert np.array_equal(a, b)


@pytest.mark.parametrize("k", [1, 7, 64, 300])
def test_any_k(matcher, k):
    from oracle import columnar

    cat = synth.make_catalog(6000, seed=71, n_shapes=128)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, 70, seed=72)
    dt = matcher.upload(table)
    try:
        got = matcher.match_topk(dt, sj.columns, k)
    finally:
        dt.close()
    want = columnar.topk(cat.columns, sj.columns, k)
    assert np.array_equal(got.index, want.index) and np.array_equal(got.count, want.count)


def test_reserve_sizes_the_scratch_ahead_of_the_first_call(matcher):
    lib = native.lib()
    cat, table, packed = _setup(matcher, n_offers=60_000, n_jobs=900, seed=91)
    m = Matcher(0)  # a fresh context: nothing allocated yet
    dt = m.upload(table)
    try:
        lay = m._layout(packed)
        native.check(lib.dsm_reserve(m.handle, dt.handle, C.byref(lay), 900, 50))
        native.check(lib.dsm_reserve(m.handle, dt.handle, C.byref(lay), 900, 0))
        assert lib.dsm_reserve(m.handle, dt.handle, C.byref(lay), 0, 16) == DSM_E_ARG
        got = m.topk_rows(dt, packed, 50)
        want = matcher.topk_rows(matcher.upload(table), packed, 50)
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
    finally:
        dt.close()
        m.close()
