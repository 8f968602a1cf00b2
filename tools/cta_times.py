"""Debug tool: per-CTA duration / SM placement of k_match on the bench workload.

Build the instrumented library first (not shipped, *.so is git-ignored):
    nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DDSM_CTA_TIMING -shared \
         -Xcompiler -fPIC -o dstack_b200/libdsmatch_dbg.so dstack_b200/csrc/dsmatch.cu
then run on the GPU box:  DSM_LIB=dstack_b200/libdsmatch_dbg.so python tools/cta_times.py
"""
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
plan = m.plan(dt, packed)
print("plan", plan)
for _ in range(3):
    m.topk_rows(dt, packed, 16)
lib = native.lib()
n = plan["grid"]
buf = np.zeros((n, 3), np.uint64)
lib.dsm_debug_cta_times.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
native.check(lib.dsm_debug_cta_times(m.handle, buf.ctypes.data, n))
t0 = buf[:, 0].astype(np.int64)
t1 = buf[:, 1].astype(np.int64)
sm = buf[:, 2].astype(np.int64)
base = t0.min()
dur = (t1 - t0) / 1e3
print(f"kernel span {(t1.max() - base) / 1e3:.1f} us; CTA duration us: min {dur.min():.1f} "
      f"mean {dur.mean():.1f} max {dur.max():.1f}; start spread {(t0.max() - base) / 1e3:.2f} us")
per_sm_end = {}
for s, e in zip(sm, t1):
    per_sm_end[s] = max(per_sm_end.get(s, 0), e)
ends = np.array(sorted((v - base) / 1e3 for v in per_sm_end.values()))
print("per-SM finish us: min %.1f median %.1f max %.1f; SMs used %d" % (ends.min(), np.median(ends), ends.max(), len(ends)))
cnt = np.bincount(sm)
print("CTAs per SM histogram:", np.bincount(cnt[cnt > 0]))
