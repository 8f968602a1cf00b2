"""`dsm_match_bitmap_static` -- the match that allocates nothing -- against `dsm_match_bitmap_dev`.

Who needs it: a host that may not call cudaMalloc while matching -- a scheduler thread that shares
the GPU with serving streams (an allocation can stall every stream on the device), or a process
whose device memory is already owned in full by a pool allocator.  For those callers the scratch
of the match (acceptance tables + work counter) sits in a fixed 2 MiB array inside the library
instead of being allocated on demand; inputs and outputs stay the caller's buffers.

The test goes through the C ABI with the caller's device buffers (torch tensors, only as memory),
runs the new entry point twice in a row -- the second time with a different, smaller batch, so
whatever the first call left in the shared array must not leak into the second result -- and
requires bit-identical bitmaps to the allocating entry point.  It also checks that a batch too
large for the array is refused with DSM_E_TOO_LARGE rather than served.

The file name sorts last on purpose: a `pytest -x` run has executed every other test before a
failure here could stop it.
"""

import ctypes as C

import numpy as np
import pytest

from dstack_b200 import native
from workloads import synth
from dstack_b200.encode import encode_table, pack_jobs
from oracle import columnar

pytestmark = pytest.mark.gpu

DSM_E_TOO_LARGE = -5


def _bitmaps(lib, matcher, dt, packed, entry):
    """Run `entry` (a C function taking device pointers) and return the bitmap as numpy."""
    import torch

    dev = torch.device("cuda", matcher.device)
    rec = np.ascontiguousarray(packed.records, dtype=np.uint32)
    n_jobs = len(rec)
    words = dt.table.n_rows // 32
    d_rec = torch.from_numpy(rec.view(np.int32)).to(dev)
    d_bits = torch.full((n_jobs, words), -1, dtype=torch.int32, device=dev)  # poison
    # the evaluation order (if the packer chose one) lives where the records live: on the device
    d_order = None if packed.order is None else \
        torch.from_numpy(np.ascontiguousarray(packed.order, dtype=np.uint32).view(np.int32)).to(dev)
    lay = matcher._layout(packed, None if d_order is None else d_order.data_ptr())
    torch.cuda.synchronize(dev)
    rc = entry(matcher.handle, dt.handle, C.byref(lay), d_rec.data_ptr(), n_jobs, d_bits.data_ptr())
    native.check(rc)
    matcher.sync()
    return d_bits.cpu().numpy().view(np.uint32)


def test_static_scratch_matches_allocating_entry_point_twice_in_a_row(matcher):
    lib = native.lib()
    cat = synth.make_catalog(30_000, seed=41, n_shapes=512)
    table = encode_table(cat.columns)
    dt = matcher.upload(table)
    try:
        jobs = [synth.make_jobs(cat, table, 300, seed=42), synth.make_jobs(cat, table, 70, seed=43)]
        for sj in jobs:  # two calls in a row on the shared array
            packed = pack_jobs(table, sj.columns)
            got = _bitmaps(lib, matcher, dt, packed, lib.dsm_match_bitmap_static)
            want = _bitmaps(lib, matcher, dt, packed, lib.dsm_match_bitmap_dev)
            assert got.shape == want.shape
            assert np.array_equal(got, want)  # integer work: bit-exact
            assert want.any(), "vacuous comparison: no job matched anything"
            # and against the CPU oracle directly, so this file stands on its own: bit l of word b
            # of job j = physical row 32 b + l matches; the oracle answers in source positions
            oracle = columnar.match_matrix(cat.columns, sj.columns)
            rows = np.unpackbits(got.view(np.uint8), axis=1, bitorder="little").astype(bool)
            live = table.perm >= 0
            assert not rows[:, ~live].any(), "a padding row matched"
            assert np.array_equal(rows[:, live], oracle[:, table.perm[live]])
    finally:
        dt.close()


def test_static_scratch_refuses_batches_beyond_its_capacity(matcher):
    import torch

    lib = native.lib()
    cat = synth.make_catalog(2_000, seed=44, n_shapes=256)
    table = encode_table(cat.columns)
    dt = matcher.upload(table)
    try:
        probe = pack_jobs(table, synth.make_jobs(cat, table, 128, seed=45).columns)
        entries = sum(probe.layout["card"])
        per_group = entries * 16
        n_jobs = 128 * (native.STATIC_SCRATCH_BYTES // per_group + 1)  # one group too many
        rec = np.ascontiguousarray(np.tile(probe.records, (n_jobs // 128, 1)), dtype=np.uint32)
        big = type(probe)(records=rec, layout=probe.layout, luts=probe.luts)
        dev = torch.device("cuda", matcher.device)
        d_rec = torch.from_numpy(rec.view(np.int32)).to(dev)
        d_bits = torch.zeros((n_jobs, table.n_rows // 32), dtype=torch.int32, device=dev)
        lay = matcher._layout(big)
        rc = lib.dsm_match_bitmap_static(matcher.handle, dt.handle, C.byref(lay), d_rec.data_ptr(), n_jobs,
                                         d_bits.data_ptr())
        assert rc == DSM_E_TOO_LARGE
        message = lib.dsm_last_error().decode()
        assert "scratch" in message and str(native.STATIC_SCRATCH_BYTES) in message
        # the same batch is fine through the allocating entry point
        native.check(lib.dsm_match_bitmap_dev(matcher.handle, dt.handle, C.byref(lay), d_rec.data_ptr(), n_jobs,
                                              d_bits.data_ptr()))
        matcher.sync()
    finally:
        dt.close()
