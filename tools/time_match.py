"""k_match time alone on a bench workload (results are NOT checked: for kernel experiments)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import native, synth  # noqa: E402
from dstack_b200.encode import encode_table, pack_jobs  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402

J, N, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
m = Matcher(0)
lib = native.lib()
cat = synth.make_catalog(N, seed=3, n_shapes=4096)
table = encode_table(cat.columns)
dt = m.upload(table)
packed = pack_jobs(table, synth.make_jobs(cat, table, J, seed=3).columns)
print(m.plan(dt, packed, k))
for _ in range(3):
    out = m.topk_rows(dt, packed, k)
native.check(lib.dsm_timing_enable(m.handle, 1))
for _ in range(20):
    out = m.topk_rows(dt, packed, k)
ms, n = C.c_double(), C.c_uint32()
native.check(lib.dsm_timing_read(m.handle, C.byref(ms), C.byref(n)))
print(f"k_match {ms.value / n.value:.4f} ms over {n.value} launches; total matches {int(out[1].astype('int64').sum())}")
