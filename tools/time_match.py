"""Kernel times of one top-k call on a bench-shaped workload (for kernel experiments).

    python tools/time_match.py JOBS OFFERS K [--no-pair] [--no-order] [--seed S] [--check N]

Prints the launch plan and the mean time of k_build_tables / k_match / k_extract over 20 calls
(CUDA events inside the library, table resident, L2 not flushed).  ``--check N`` compares N
sampled jobs with the numpy oracle first.
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import native  # noqa: E402
from workloads import synth  # noqa: E402
from dstack_b200.encode import encode_table, pack_jobs  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("jobs", type=int)
ap.add_argument("offers", type=int)
ap.add_argument("k", type=int)
ap.add_argument("--no-pair", action="store_true")
ap.add_argument("--no-order", action="store_true")
ap.add_argument("--seed", type=int, default=3)
ap.add_argument("--check", type=int, default=0)
a = ap.parse_args()

m = Matcher(0)
lib = native.lib()
cat = synth.make_catalog(a.offers, seed=a.seed, n_shapes=4096)
table = encode_table(cat.columns, pair_columns=False if a.no_pair else None)
dt = m.upload(table)
sj = synth.make_jobs(cat, table, a.jobs, seed=a.seed)
packed = pack_jobs(table, sj.columns)
if a.no_order:
    packed.order = None
print("columns", [(c.name, c.card) for c in table.columns][:packed.layout["n_used"]], "ordered", packed.order is not None)
print(m.plan(dt, packed, a.k))
for _ in range(3):
    out = m.topk_rows(dt, packed, a.k)
if a.check:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench

    t = table
    idx, cnt, mx = out
    src = np.where(idx == native.NO_ROW, -1, t.perm[np.minimum(idx, t.n_rows - 1)]).astype(np.int64)
    price = np.where(mx == native.NO_ROW, np.nan, t.price[np.minimum(mx, t.n_rows - 1)])
    n = bench.verify(cat.columns, sj, a.k, src, cnt.astype(np.int64), price, np.random.default_rng(0), n_sample=a.check)
    print(f"verified {n} sampled jobs against the oracle, properties for all")
native.check(lib.dsm_timing_enable(m.handle, 1))
for _ in range(20):
    out = m.topk_rows(dt, packed, a.k)
parts, n = (C.c_double * 3)(), C.c_uint32()
native.check(lib.dsm_timing_read_parts(m.handle, C.byref(parts), C.byref(n)))
native.check(lib.dsm_timing_enable(m.handle, 0))
b, k_, e = (x / n.value for x in parts)
print(f"build {b:.4f} ms  k_match {k_:.4f} ms  k_extract {e:.4f} ms  over {n.value} calls; "
      f"total matches {int(out[1].astype('int64').sum())}")
