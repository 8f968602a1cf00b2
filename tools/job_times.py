"""Debug tool: per-job duration of k_extract on the bench workload (instrumented build, see cta_times.py)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import native  # noqa: E402
from workloads import synth  # noqa: E402
from dstack_b200.encode import encode_table, pack_jobs  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402

n_offers = int(os.environ.get("N_OFFERS", 1 << 20))
n_jobs = int(os.environ.get("N_JOBS", 4096))
cat = synth.make_catalog(n_offers, seed=3, n_shapes=4096)
table = encode_table(cat.columns)
sj = synth.make_jobs(cat, table, n_jobs, seed=3)
packed = pack_jobs(table, sj.columns)
m = Matcher(0)
dt = m.upload(table)
print("plan", m.plan(dt, packed))
for _ in range(3):
    idx, cnt, mx = m.topk_rows(dt, packed, 16)
lib = native.lib()
buf = np.zeros((n_jobs, 2), np.uint64)
lib.dsm_debug_job_times.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
native.check(lib.dsm_debug_job_times(m.handle, buf.ctypes.data, n_jobs))
t0, t1 = buf[:, 0].astype(np.int64), buf[:, 1].astype(np.int64)
dur = (t1 - t0) / 1e3
print("kernel span us", (t1.max() - t0.min()) / 1e3, "job dur us: mean %.1f p50 %.1f p90 %.1f p99 %.1f max %.1f" %
      (dur.mean(), np.median(dur), np.percentile(dur, 90), np.percentile(dur, 99), dur.max()))
print("start spread us", (t0.max() - t0.min()) / 1e3)
order = np.argsort(-dur)
first = np.where(idx[:, 0] == native.NO_ROW, -1, idx[:, 0].astype(np.int64))
last = np.where(idx[:, 15] == native.NO_ROW, -1, idx[:, 15].astype(np.int64))
for j in order[:12]:
    print("job %5d dur %.1f us start+%.1f count %8d first %8d 16th %8d max_row %d cuts %s" %
          (j, dur[j], (t0[j] - t0.min()) / 1e3, cnt[j], first[j], last[j], int(mx[j]), packed.records[j, :2]))
for lo, hi in [(0, 1), (1, 16), (16, 1000), (1000, 100000), (100000, 10 ** 9)]:
    sel = (cnt >= lo) & (cnt < hi)
    if sel.any():
        print("count in [%d,%d): %d jobs, mean dur %.1f us, max %.1f" % (lo, hi, sel.sum(), dur[sel].mean(), dur[sel].max()))
