import sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from dstack_b200 import pools as PP
from workloads import synth
from dstack_b200.engine import Matcher
from dstack_b200.ingest import PoolParseCache
import test_configs as T
m = Matcher(0)
rows = synth.make_pool(50_000, seed=2, n_fleets=200)
calls = T._config2_jobs(rows, 1000)
cache = PoolParseCache()
t0 = time.perf_counter(); index = PP.PoolIndex(rows, matcher=m, parse_cache=cache); t1 = time.perf_counter()
index.close()
t2 = time.perf_counter(); index = PP.PoolIndex(rows, matcher=m, parse_cache=cache); t3 = time.perf_counter()
filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"], fleet_model=kw.get("fleet_model"),
                           multinode=kw.get("multinode", False), master_job_provisioning_data=kw.get("master_job_provisioning_data"),
                           volumes=None, shared=False) for p, kw in calls]
for _ in range(2): got = index.match(filters)
ts = []
for _ in range(5):
    t = time.perf_counter(); got = index.match(filters); ts.append(time.perf_counter() - t)
print(f"pool index build cold {t1-t0:.2f} s, warm (parse cache) {t3-t2:.2f} s; 1000 jobs x 50000 rows: match {1e3*np.median(ts):.2f} ms "
      f"({1000*50000/np.median(ts)/1e9:.2f} G evals/s incl. host packing and unpacking); hits {sum(len(g) for g in got)}")
