"""The sharded step's all-gather alone: NCCL against the library's peer exchange.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29517 tools/xchg_bench.py [--jobs 16384] [--k 16] [--iters 200]

Every rank contributes one per-shard record of a J-job, top-k batch (config 4: 1.25 MiB).  Times
are CUDA-event pairs around one exchange on the stream, ranks released together by a barrier
before every batch of 20; mean and max over ranks of the per-iteration mean.  The result of both
is compared word for word."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import native  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--jobs", type=int, default=16384)
ap.add_argument("--k", type=int, default=16)
ap.add_argument("--iters", type=int, default=200)
a = ap.parse_args()

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = native.lib()
words = (a.jobs * a.k + 4 * a.jobs + 3) // 4 * 4
stream = torch.cuda.Stream(device=dev)
with torch.cuda.stream(stream):
    m = Matcher(local, stream=stream.cuda_stream)
    src = torch.from_numpy(np.random.default_rng(rank).integers(0, 1 << 31, words, dtype=np.int64).astype(np.int32)).to(dev)
    g_nccl = torch.empty((world, words), dtype=torch.int32, device=dev)
    # set-up of the peer exchange: handles through NCCL
    handle = (C.c_ubyte * native.IPC_HANDLE_BYTES)()
    x = C.c_void_p()
    native.check(lib.dsm_xchg_create(m.handle, world, rank, 4 * words, C.byref(x), handle))
    mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=dev)
    every = torch.empty((world, native.IPC_HANDLE_BYTES), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(every, mine)
    native.check(lib.dsm_xchg_connect_ipc(x, every.cpu().numpy().tobytes()))

    def nccl():
        dist.all_gather_into_tensor(g_nccl, src)

    out, stride = C.c_void_p(), C.c_uint64()

    def p2p():
        native.check(lib.dsm_xchg_all_gather_dev(x, src.data_ptr(), 4 * words, C.byref(out), C.byref(stride)))

    def timed(fn):
        for _ in range(10):
            fn()
        ms = []
        for i in range(a.iters):
            if i % 20 == 0:
                stream.synchronize()
                dist.barrier()
                stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            ms.append((e0, e1))
        stream.synchronize()
        return float(np.mean([e0.elapsed_time(e1) for e0, e1 in ms]))

    t_nccl = timed(nccl)
    t_p2p = timed(p2p)
    bad = C.c_uint32()
    native.check(lib.dsm_xchg_status(x, C.byref(bad)))

    class Raw:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 3}

    same = True
    for r in range(world):
        got = torch.as_tensor(Raw(out.value + r * stride.value, words), device=dev)
        same = same and bool(torch.equal(got, g_nccl[r]))
    t = torch.tensor([t_nccl, t_p2p, float(same), float(bad.value)], dtype=torch.float64, device=dev)
    allr = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allr, t)
    allr = torch.stack(allr).cpu().numpy()
    if rank == 0:
        print(json.dumps({"world": world, "bytes_per_rank": 4 * words,
                          "nccl_all_gather_ms": {"mean": allr[:, 0].mean(), "max": allr[:, 0].max()},
                          "peer_exchange_ms": {"mean": allr[:, 1].mean(), "max": allr[:, 1].max()},
                          "identical": bool(allr[:, 2].min() == 1.0), "timeouts": int(allr[:, 3].max())}))
    stream.synchronize()
    dist.barrier()
    lib.dsm_xchg_destroy(x)
    m.close()
dist.destroy_process_group()
