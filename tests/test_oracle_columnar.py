"""The numpy oracle against the object-level oracle (both CPU, both test infrastructure)."""

import numpy as np
import pytest

from workloads import synth
from dstack_b200.encode import encode_table
from oracle import columnar
from oracle import reference_path as ref


@pytest.mark.parametrize("n_offers,n_jobs,seed", [(1500, 48, 1), (400, 32, 7)])
def test_columnar_oracle_equals_object_oracle(n_offers, n_jobs, seed):
    cat = synth.make_catalog(n_offers, seed=seed, n_shapes=200)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    table = encode_table(cat.columns)  # only for the dictionaries the job ids refer to
    sj = synth.make_jobs(cat, table, n_jobs, seed=seed + 2)
    got = columnar.topk(cat.columns, sj.columns, 15)
    backends = [b for b, _ in per_backend]
    offers = [o for _, lst in per_backend for o in lst]
    pos = {id(o): i for i, o in enumerate(offers)}
    nonempty = 0
    for j, (profile, req) in enumerate(synth.job_objects(sj)):
        want = ref.get_offers_by_requirements(backends, profile, req)
        first, total, top_price = ref.plan_slice([o for _, o in want], 15)
        assert got.count[j] == total
        assert [i for i in got.index[j] if i >= 0] == [pos[id(o)] for o in first]
        if total:
            nonempty += 1
            assert got.max_price[j] == top_price
        else:
            assert np.isnan(got.max_price[j])
    assert nonempty >= n_jobs // 8  # the batch must not be vacuous
