"""Latency of small batches against a resident catalog (config 1: one task, 10k offers)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import models as M  # noqa: E402
from workloads import synth  # noqa: E402
from dstack_b200.encode import encode_table, pack_jobs  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402
from dstack_b200.ingest import job_columns  # noqa: E402
from dstack_b200.services import job_filter  # noqa: E402

m = Matcher(0)
for n_offers in (10_000, 1_000_000):
    cat = synth.make_catalog(n_offers, seed=1, n_shapes=1024)
    table = encode_table(cat.columns)
    dt = m.upload(table)
    req = M.Requirements(resources=M.ResourcesSpec(gpu="24GB..80GB:1..8"), max_price=None, spot=None)
    prof = M.Profile(name="p")
    for n_jobs in (1, 5, 128, 1024):
        def once():
            t0 = time.perf_counter()
            f = [job_filter(prof, req)] * n_jobs
            jc = job_columns(table, f)
            t1 = time.perf_counter()
            packed = pack_jobs(table, jc)
            t2 = time.perf_counter()
            out = m.topk_rows(dt, packed, 50)
            t3 = time.perf_counter()
            return t1 - t0, t2 - t1, t3 - t2, out
        for _ in range(3):
            once()
        runs = [once() for _ in range(20)]
        med = lambda i: 1e3 * float(np.median([r[i] for r in runs]))
        print(f"offers {n_offers:8d} jobs {n_jobs:5d}: filter->columns {med(0):.3f} ms, pack {med(1):.3f} ms, "
              f"dsm_match_topk (H2D+kernels+D2H) {med(2):.3f} ms; matches of job 0: {int(runs[0][3][1][0])}")
    dt.close()
