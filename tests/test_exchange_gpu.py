"""The peer exchange (``dsm_xchg_*``, include/dsmatch.h) on one GPU.

The exchange replaces the NCCL all-gather of the sharded step; here its ranks share ``cuda:0``:
as two contexts of one process wired with base pointers, and as two processes wired with CUDA IPC
handles (the form ``ShardedMatcher`` uses, one process per GPU).  Every rank must end up with all
records in rank order, epoch after epoch, and a rank whose peer never arrives must give up instead
of spinning for ever."""

import ctypes as C
import multiprocessing as mp
import os
import sys

import numpy as np
import pytest

from dstack_b200 import native
from dstack_b200.engine import Matcher

pytestmark = pytest.mark.gpu
DSM_E_ARG = -2


def _record(rank: int, epoch: int, n_words: int) -> np.ndarray:
    return (np.arange(n_words, dtype=np.uint32) * np.uint32(2654435761) + np.uint32(rank * 7919 + epoch * 104729)).astype(np.uint32)


def _gather(lib, x, src_ptr, nbytes):
    out, stride = C.c_void_p(), C.c_uint64()
    native.check(lib.dsm_xchg_all_gather_dev(x, src_ptr, nbytes, C.byref(out), C.byref(stride)))
    return out.value, stride.value


def _read(torch, ptr, stride, world, n_words):
    """Copy [world][n_words] u32 out of a raw device pointer."""
    got = np.empty((world, n_words), dtype=np.uint32)
    for r in range(world):
        buf = torch.empty((n_words,), dtype=torch.int32, device="cuda:0")
        # wrap the raw pointer with the CUDA array interface, let torch copy
        src = torch.as_tensor(_RawCuda(ptr + r * stride, n_words), device="cuda:0")
        buf.copy_(src)
        got[r] = buf.cpu().numpy().view(np.uint32)
    return got


class _RawCuda:
    def __init__(self, ptr, n_words):
        self.__cuda_array_interface__ = {"shape": (n_words,), "typestr": "<i4", "data": (ptr, False), "version": 3}


@pytest.mark.parametrize("world,n_words", [(2, 4096), (3, 1 << 18), (1, 64)])
def test_exchange_between_contexts_of_one_process(world, n_words):
    import torch

    lib = native.lib()
    ms = [Matcher(0) for _ in range(world)]
    xs = []
    try:
        for r in range(world):
            x = C.c_void_p()
            native.check(lib.dsm_xchg_create(ms[r].handle, world, r, 4 * n_words, C.byref(x), None))
            xs.append(x)
        bases = (C.c_void_p * world)()
        for r in range(world):
            b = C.c_void_p()
            native.check(lib.dsm_xchg_local_base(xs[r], C.byref(b)))
            bases[r] = b.value
        for r in range(world):
            if world > 1:
                native.check(lib.dsm_xchg_connect_ptrs(xs[r], bases))
        for epoch in range(1, 6):
            srcs = [torch.from_numpy(_record(r, epoch, n_words).view(np.int32)).to("cuda:0") for r in range(world)]
            torch.cuda.synchronize()
            outs = [_gather(lib, xs[r], srcs[r].data_ptr(), 4 * n_words) for r in range(world)]
            for m in ms:
                m.sync()
            want = np.stack([_record(r, epoch, n_words) for r in range(world)])
            for r in range(world):
                assert np.array_equal(_read(torch, outs[r][0], outs[r][1], world, n_words), want), (r, epoch)
        for r in range(world):
            bad = C.c_uint32(9)
            native.check(lib.dsm_xchg_status(xs[r], C.byref(bad)))
            assert bad.value == 0
    finally:
        for x in xs:
            lib.dsm_xchg_destroy(x)
        for m in ms:
            m.close()


def test_exchange_rejects_bad_arguments():
    import torch

    lib = native.lib()
    m = Matcher(0)
    x = C.c_void_p()
    try:
        assert lib.dsm_xchg_create(m.handle, 0, 0, 64, C.byref(x), None) == DSM_E_ARG
        assert lib.dsm_xchg_create(m.handle, 2, 2, 64, C.byref(x), None) == DSM_E_ARG
        assert lib.dsm_xchg_create(m.handle, 2, 0, 24, C.byref(x), None) == DSM_E_ARG
        native.check(lib.dsm_xchg_create(m.handle, 2, 0, 64, C.byref(x), None))
        src = torch.zeros((64,), dtype=torch.int32, device="cuda:0")
        out = C.c_void_p()
        # not connected yet; too large; misaligned
        assert lib.dsm_xchg_all_gather_dev(x, src.data_ptr(), 64, C.byref(out), None) == DSM_E_ARG
        bases = (C.c_void_p * 2)()
        b = C.c_void_p()
        native.check(lib.dsm_xchg_local_base(x, C.byref(b)))
        bases[0], bases[1] = b.value, b.value
        native.check(lib.dsm_xchg_connect_ptrs(x, bases))
        assert lib.dsm_xchg_all_gather_dev(x, src.data_ptr(), 512, C.byref(out), None) == DSM_E_ARG
        assert lib.dsm_xchg_all_gather_dev(x, src.data_ptr() + 4, 64, C.byref(out), None) == DSM_E_ARG
    finally:
        lib.dsm_xchg_destroy(x)
        m.close()


def test_a_missing_peer_times_out_instead_of_hanging():
    import torch

    lib = native.lib()
    m0, m1 = Matcher(0), Matcher(0)
    x0, x1 = C.c_void_p(), C.c_void_p()
    try:
        native.check(lib.dsm_xchg_create(m0.handle, 2, 0, 256, C.byref(x0), None))
        native.check(lib.dsm_xchg_create(m1.handle, 2, 1, 256, C.byref(x1), None))
        bases = (C.c_void_p * 2)()
        for i, x in enumerate((x0, x1)):
            b = C.c_void_p()
            native.check(lib.dsm_xchg_local_base(x, C.byref(b)))
            bases[i] = b.value
        native.check(lib.dsm_xchg_connect_ptrs(x0, bases))
        native.check(lib.dsm_xchg_connect_ptrs(x1, bases))
        src = torch.zeros((64,), dtype=torch.int32, device="cuda:0")
        _gather(lib, x0, src.data_ptr(), 256)  # rank 1 never calls
        bad = C.c_uint32()
        native.check(lib.dsm_xchg_status(x0, C.byref(bad)))  # returns after the 5 s limit
        assert bad.value == 2  # 1 + the rank that was missing
    finally:
        lib.dsm_xchg_destroy(x0)
        lib.dsm_xchg_destroy(x1)
        m0.close()
        m1.close()


def _ipc_worker(rank, world, n_words, epochs, inbox, barrier, result_q):
    try:
        import torch

        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from dstack_b200 import native as nat
        from dstack_b200.engine import Matcher as M

        lib = nat.lib()
        m = M(0)
        x = C.c_void_p()
        handle = (C.c_ubyte * nat.IPC_HANDLE_BYTES)()
        nat.check(lib.dsm_xchg_create(m.handle, world, rank, 4 * n_words, C.byref(x), handle))
        handles = {rank: bytes(handle)}
        for r in range(world):
            if r != rank:
                inbox[r].put((rank, bytes(handle)))
        while len(handles) < world:
            r, h = inbox[rank].get(timeout=120)
            handles[r] = h
        nat.check(lib.dsm_xchg_connect_ipc(x, b"".join(handles[r] for r in range(world))))
        ok = True
        for epoch in range(1, epochs + 1):
            src = torch.from_numpy(_record(rank, epoch, n_words).view(np.int32)).to("cuda:0")
            torch.cuda.synchronize()
            ptr, stride = _gather(lib, x, src.data_ptr(), 4 * n_words)
            m.sync()
            want = np.stack([_record(r, epoch, n_words) for r in range(world)])
            ok = ok and np.array_equal(_read(torch, ptr, stride, world, n_words), want)
        bad = C.c_uint32(9)
        nat.check(lib.dsm_xchg_status(x, C.byref(bad)))
        barrier.wait(timeout=120)  # nobody unmaps or frees while a peer may still store
        lib.dsm_xchg_destroy(x)
        m.close()
        result_q.put((rank, bool(ok), int(bad.value), ""))
    except Exception as e:  # noqa: BLE001
        result_q.put((rank, False, -1, repr(e)))


def test_exchange_between_processes_over_cuda_ipc():
    """Two processes on cuda:0 map each other's buffer with the IPC handles, as two ranks of
    ``ShardedMatcher`` do with theirs on two GPUs."""
    ctx = mp.get_context("spawn")
    world, n_words, epochs = 2, 1 << 16, 4
    inbox = [ctx.Queue() for _ in range(world)]
    barrier = ctx.Barrier(world)
    result_q = ctx.Queue()
    procs = [ctx.Process(target=_ipc_worker, args=(r, world, n_words, epochs, inbox, barrier, result_q))
             for r in range(world)]
    for p in procs:
        p.start()
    try:
        results = sorted(result_q.get(timeout=300) for _ in range(world))
    finally:
        for p in procs:
            p.join(timeout=60)
            if p.is_alive():
                p.kill()
    assert results == [(0, True, 0, ""), (1, True, 0, "")], results
