"""Time the interleaved shard merge on one GPU: synthetic per-shard results of a J-job, top-k batch
from G shards.  python tools/merge_bench.py [G] [J] [k]

Measured on a B200 (G=8, J=16384, k=16): 25 us for the thread-per-job kernel.  A warp-per-job
variant that ranked all G*k candidates at once with binary searches took 62 us -- 25 times the
instructions (position arithmetic and searches for 128 candidates where the serial selection
touches 24) -- and was dropped."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import native  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
J = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
k = int(sys.argv[3]) if len(sys.argv) > 3 else 16
rows_per_shard, chunk = 8 << 20, 1024
rng = np.random.default_rng(5)
w_idx, w_cnt = J * k, J
w_head = (w_idx + 2 * w_cnt + 1) // 2 * 2
words = (w_head + 2 * J + 3) // 4 * 4
g = np.zeros((G, words), dtype=np.uint32)
for s in range(G):
    n = rng.integers(0, 3 * k, J)
    idx = np.sort(rng.integers(0, rows_per_shard, (J, k)).astype(np.uint32), axis=1)
    idx[np.arange(k)[None, :] >= n[:, None]] = native.NO_ROW
    g[s, :w_idx] = idx.ravel()
    g[s, w_idx:w_idx + J] = n
    mx = np.where(n > 0, rng.integers(0, rows_per_shard, J), native.NO_ROW).astype(np.uint32)
    g[s, w_idx + J:w_idx + 2 * J] = mx
    g[s, w_head:w_head + 2 * J] = rng.random(J).view(np.uint32)
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
with torch.cuda.stream(stream):
    m = Matcher(0, stream=stream.cuda_stream)
    lib = native.lib()
    gd = torch.from_numpy(g.view(np.int32)).to(dev)
    geom = torch.tensor([[rows_per_shard, rows_per_shard]] * G, dtype=torch.int32, device=dev)
    o_idx = torch.empty((J, k), dtype=torch.int64, device=dev)
    o_cnt = torch.empty((J,), dtype=torch.int64, device=dev)
    o_max = torch.empty((J,), dtype=torch.int64, device=dev)
    o_mp = torch.empty((J,), dtype=torch.float64, device=dev)
    p0 = gd.data_ptr()

    def run():
        native.check(lib.dsm_merge_shards_interleaved_dev(
            m.handle, G, J, k, p0, p0 + 4 * w_idx, p0 + 4 * (w_idx + w_cnt), p0 + 4 * w_head, 4 * words,
            geom.data_ptr(), chunk, o_idx.data_ptr(), o_cnt.data_ptr(), o_max.data_ptr(), o_mp.data_ptr()))

    for _ in range(5):
        run()
    evs = []
    for _ in range(50):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        run()
        e1.record(stream)
        evs.append((e0, e1))
    stream.synchronize()
    t = np.mean([a.elapsed_time(b) for a, b in evs])
    chk = (int(o_idx.sum().item()) & 0xFFFFFFFFFFFF, int(o_cnt.sum().item()), int(o_max.sum().item()) & 0xFFFFFFFFFFFF,
           float(o_mp.nan_to_num().sum().item()))
    print(f"G={G} J={J} k={k}: {1e3 * t:.1f} us  checksum {chk}")
