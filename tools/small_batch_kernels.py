"""One small batch against a resident table; run under ncu for the per-kernel launch list."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dstack_b200 import models as M  # noqa: E402
from workloads import synth  # noqa: E402
from dstack_b200.encode import encode_table, pack_jobs  # noqa: E402
from dstack_b200.engine import Matcher  # noqa: E402
from dstack_b200.ingest import job_columns  # noqa: E402
from dstack_b200.services import job_filter  # noqa: E402

n_offers, n_jobs = int(sys.argv[1]), int(sys.argv[2])
m = Matcher(0)
table = encode_table(synth.make_catalog(n_offers, seed=1, n_shapes=1024).columns)
dt = m.upload(table)
req = M.Requirements(resources=M.ResourcesSpec(gpu="24GB..80GB:1..8"), max_price=None, spot=None)
packed = pack_jobs(table, job_columns(table, [job_filter(M.Profile(name="p"), req)] * n_jobs))
print(m.plan(dt, packed))
for _ in range(3):
    out = m.topk_rows(dt, packed, 50)
print(int(out[1][0]))
