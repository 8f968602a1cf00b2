"""BASELINE.json `configs` as parity cases (config 3 / 4 are the bench workloads).

  config 1  one task `gpu: 24GB..80GB:1..8, spot: auto` against a 10k-offer catalog: full ordered
            list and first 50, against the object-level oracle.
  config 2  1 000 pending jobs x 50 000 pool rows with reuse / same-fleet / multinode constraints:
            one batched PoolIndex query; sampled jobs against the object-level oracle (which
            parses the offer JSON per (job, row), ~5 s per job) plus properties for every job.
  config 5  multinode placement: 8 cluster requests x 8 nodes `gpu: H100:8`, masters first, then
            followers pinned to their master's backend and region (offers.py:71-79), top-16.
"""

import types

import numpy as np
import pytest

from dstack_b200 import models as M
from dstack_b200 import pools as PP
from dstack_b200 import services as S
from workloads import synth
from oracle import reference_path as ref

pytestmark = pytest.mark.gpu
BT = M.BackendType


def test_config1_single_task_10k_offers(matcher):
    cat = synth.make_catalog(10_000, seed=1, n_shapes=1024)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    req = M.Requirements(resources=M.ResourcesSpec(gpu="24GB..80GB:1..8"), max_price=None, spot=None)
    profile = M.Profile(name="p")
    want = ref.get_offers_by_requirements([b for b, _ in per_backend], profile, req)
    catalog = S.OfferCatalog(per_backend, matcher=matcher)
    try:
        f = [S.job_filter(profile, req)]
        full = catalog.offers_for(f)[0]
        top = catalog.match_topk(f, 50)
    finally:
        catalog.close()
    assert len(want) > 500
    assert [id(o) for _, o in full] == [id(o) for _, o in want]
    first, total, top_price = ref.plan_slice([o for _, o in want], 50)
    assert [id(o) for _, o in catalog.pairs(top.index[0])] == [id(o) for o in first]
    assert top.count[0] == total and top.max_price[0] == top_price


from workloads.configs import config2_jobs as _config2_jobs  # noqa: E402


def test_config2_pool_1k_jobs_50k_instances(matcher):
    rows = synth.make_pool(50_000, seed=2, n_fleets=200)
    calls = _config2_jobs(rows, 1000)
    index = PP.PoolIndex(rows, matcher=matcher)
    try:
        filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"],
                                   fleet_model=kw.get("fleet_model"), multinode=kw.get("multinode", False),
                                   master_job_provisioning_data=kw.get("master_job_provisioning_data"),
                                   volumes=None, shared=False) for p, kw in calls]
        got = index.match(filters)  # ONE batched device call for the 1000 jobs
    finally:
        index.close()
    # properties for every job: ascending input order, only IDLE / reachable / unshared rows with an offer
    for g in got:
        assert (np.diff(g) > 0).all()
        for i in g[:50]:
            r = rows[i]
            assert r.status == M.InstanceStatus.IDLE and not r.unreachable and r.total_blocks == 1 and r.offer is not None
    assert sum(len(g) > 0 for g in got) > 300
    # sampled jobs against the object-level oracle (JSON parse per (job, row), as pools.py:499 does)
    sample = [0, 1, 2, 3] + [j for j, (p, kw) in enumerate(calls) if "fleet_model" in kw][:1] + \
        [j for j, (p, kw) in enumerate(calls) if kw.get("multinode")][:1]
    for j in sample:
        profile, kw = calls[j]
        want = ref.filter_pool_instances(rows, profile, **kw)
        assert [rows[i] is w for i, w in zip(got[j], want)] == [True] * len(want) and len(got[j]) == len(want), j
    # the scheduler takes the cheapest candidate (process_submitted_jobs.py:389-390)
    j = sample[0]
    want = ref.filter_pool_instances(rows, calls[j][0], **calls[j][1])
    if want:
        ranked = ref.rank_pool_candidates(want, [])
        best = min(got[j], key=lambda i: (rows[i].price or 0, i))
        assert ranked[0][0] is rows[best]


def test_config5_multinode_two_phase_placement(matcher):
    cat = synth.make_catalog(40_000, seed=5, n_shapes=512)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    backends = [b for b, _ in per_backend]
    req = M.Requirements(resources=M.ResourcesSpec(gpu="H100:8"), max_price=None, spot=None)
    catalog = S.OfferCatalog(per_backend, matcher=matcher)
    try:
        # phase 1: the 8 masters (multinode => only BACKENDS_WITH_MULTINODE_SUPPORT)
        masters = [S.job_filter(M.Profile(name=f"c{i}", regions=None), req, multinode=True, exclude_not_available=True)
                   for i in range(8)]
        top = catalog.match_topk(masters, 16)
        want = ref.get_offers_by_requirements(backends, M.Profile(name="c"), req, multinode=True, exclude_not_available=True)
        assert len(want) >= 8, "synthetic catalog has too few H100:8 multinode offers"
        for i in range(8):
            assert [id(o) for _, o in catalog.pairs(top.index[i])] == [id(o) for _, o in want[:16]]
            assert top.count[i] == len(want)
        # each cluster takes a different offer of the ranked list as its master
        chosen = [want[i] for i in range(8)]
        assert all(b.TYPE in M.BACKENDS_WITH_MULTINODE_SUPPORT for b, _ in chosen)
        # phase 2: 7 followers per cluster, pinned to the master's backend and region
        followers, expect = [], []
        for b, o in chosen:
            jpd = M.JobProvisioningData(backend=o.backend, instance_type=o.instance, region=o.region, price=o.price)
            f = S.job_filter(M.Profile(name="c"), req, multinode=True, master_job_provisioning_data=jpd,
                             exclude_not_available=True)
            w = ref.get_offers_by_requirements(backends, M.Profile(name="c"), req, multinode=True,
                                               master_job_provisioning_data=jpd, exclude_not_available=True)
            followers += [f] * 7
            expect += [w] * 7
        top2 = catalog.match_topk(followers, 16)
        for j, w in enumerate(expect):
            assert [id(o) for _, o in catalog.pairs(top2.index[j])] == [id(o) for _, o in w[:16]]
            assert top2.count[j] == len(w)
            assert all(o.backend == chosen[j // 7][1].backend and o.region == chosen[j // 7][1].region for _, o in w)
    finally:
        catalog.close()
