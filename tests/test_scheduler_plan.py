"""the batched scheduler step and the plan slice against the
reference's sequential logic (restated in oracle/reference_path.py)."""

import copy

import numpy as np
import pytest

from dstack_b200 import models as M
from dstack_b200 import scheduler, services as S
from workloads import synth
from oracle import reference_path as ref

gpu = pytest.mark.gpu


def _requests(rows, n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    fleets = sorted({r.fleet_id for r in rows if r.fleet_id is not None}, key=str)
    specs = [None, None, "16GB..:1..8", "A100", "1..2", "T4,L4:1..4", "24GB..80GB:1..8"]
    out = []
    import types
    for _ in range(n):
        gpu = specs[int(rng.integers(0, len(specs)))]
        req = M.Requirements(resources=M.ResourcesSpec(gpu=gpu, cpu=f"{int(rng.choice([1, 2, 4, 8]))}.."),
                             max_price=None if rng.random() < 0.6 else float(np.round(rng.lognormal(1.0, 1.0), 2)), spot=None)
        kw = {}
        if rng.random() < 0.25:
            kw["fleet_model"] = types.SimpleNamespace(id=fleets[int(rng.integers(0, len(fleets)))])
        if rng.random() < 0.15:
            kw["multinode"] = True
        out.append(scheduler.JobRequest(profile=M.Profile(name="p"), requirements=req, **kw))
    return out


@gpu
def test_batched_assignment_equals_sequential_reference(matcher):
    rows = synth.make_pool(4000, seed=31, n_fleets=25)
    jobs = _requests(rows, 300, seed=32)
    seq_rows = copy.deepcopy(rows)
    want = []
    for j in jobs:
        res = ref.assign_job_to_pool_instance(seq_rows, j.profile, j.requirements, fleet_model=j.fleet_model,
                                              multinode=j.multinode)
        want.append(None if res is None else (seq_rows.index(res[0]), res[1].blocks, res[1].json()))
    got = scheduler.assign_jobs_to_pool(rows, jobs, matcher=matcher)
    assert len(got) == len(want)
    assigned = shared_hits = 0
    for g, w in zip(got, want):
        if w is None:
            assert g.instance is None
            continue
        assigned += 1
        assert rows.index(g.instance) == w[0] and g.blocks == w[1]
        assert g.offer.json() == w[2]
        shared_hits += (g.instance.total_blocks or 1) > 1
    assert assigned > 150 and shared_hits > 5
    # the pool was mutated exactly like the sequential loop mutates it
    assert [(r.status, r.busy_blocks) for r in rows] == [(r.status, r.busy_blocks) for r in seq_rows]


@gpu
def test_plan_slice_equals_reference_get_plan(matcher):
    cat = synth.make_catalog(3000, seed=33, n_shapes=200)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    backends = [b for b, _ in per_backend]
    pool = synth.make_pool(400, seed=34, n_fleets=5)
    catalog = S.OfferCatalog(per_backend, matcher=matcher)
    reqs = [M.Requirements(resources=M.ResourcesSpec(), max_price=None, spot=None),
            M.Requirements(resources=M.ResourcesSpec(gpu="H200:8"), max_price=None, spot=None),       # few / no matches
            M.Requirements(resources=M.ResourcesSpec(gpu="16GB..:1..8"), max_price=2.0, spot=None),
            M.Requirements(resources=M.ResourcesSpec(cpu="128.."), max_price=None, spot=True),
            M.Requirements(resources=M.ResourcesSpec(gpu="v5litepod-8:tpu"), max_price=None, spot=None)]
    try:
        profile = M.Profile(name="p")
        filters = [S.job_filter(profile, r) for r in reqs]
        top = catalog.match_topk(filters, 50)
        kinds = set()
        for j, r in enumerate(reqs):
            pool_offers = ref.get_pool_offers(pool, profile, r)
            offers = ref.get_offers_by_requirements(backends, profile, r)
            want_first, want_total, want_max = ref.get_plan_offers(pool_offers, offers, 50)
            first, total, top_price = scheduler.plan_slice(pool_offers, catalog, top, j, 50)
            assert [o.json() for o in first] == [o.json() for o in want_first]
            assert (total, top_price) == (want_total, want_max)
            kinds.add((len(offers) >= 50, bool(pool_offers)))
        assert len(kinds) >= 2
    finally:
        catalog.close()


@gpu
def test_process_submitted_batch_equals_sequential_reference(matcher):
    """The whole pass -- waiting rules, pool step, offers[:15] of new instances with launch
    failures, REUSE-only jobs, followers pinned to their master in the second tick -- against the
    sequential restatement of process_submitted_jobs.py:115-260, 418-490."""
    from workloads import configs

    rows = synth.make_pool(1500, seed=61, n_fleets=12)
    cat = synth.make_catalog(6000, seed=62, n_shapes=300)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    backends = [b for b, _ in per_backend]
    jobs = configs.scheduler_jobs(rows, 240, seed=63)
    launch = configs.fake_launch(fail_every=3)
    seq_rows = copy.deepcopy(rows)
    want = ref.process_submitted_jobs(seq_rows, backends, jobs, launch)
    catalog = S.OfferCatalog(per_backend, matcher=matcher)
    try:
        got = scheduler.process_submitted_batch(rows, catalog, jobs, launch, matcher=matcher)
    finally:
        catalog.close()
    kinds = {"pool": 0, "new": 0, "terminating": 0, "waiting": 0, "follower_new": 0, "reuse_failed": 0}
    for j, (g, w) in enumerate(zip(got, want)):
        status, inst, offer, jpd = w
        assert g.status == status, j
        if status != "provisioning":
            kinds[status] += 1
            kinds["reuse_failed"] += jobs[j].profile.creation_policy == M.CreationPolicy.REUSE
            continue
        if inst is not None:
            assert g.instance is rows[seq_rows.index(inst)], j
            assert g.offer.json() == offer.json(), j
            kinds["pool"] += 1
        else:
            assert g.instance is None and g.offer.json() == offer.json(), j
            assert g.job_provisioning_data.json() == jpd.json(), j
            kinds["new"] += 1
            kinds["follower_new"] += jobs[j].job_num != 0
    assert [(r.status, r.busy_blocks) for r in rows] == [(r.status, r.busy_blocks) for r in seq_rows]
    assert kinds["pool"] > 40 and kinds["new"] > 40 and kinds["reuse_failed"] > 3 and kinds["follower_new"] > 5, kinds


def test_placement_group_keys_first_offer_of_each_backend_region():
    mk = lambda b, r: __import__("types").SimpleNamespace(backend=b, region=r)
    offers = [mk(M.BackendType.AWS, "us-east-1"), mk(M.BackendType.GCP, "us-central1"), mk(M.BackendType.AWS, "us-east-1"),
              mk(M.BackendType.AWS, "eu-west-1"), mk(M.BackendType.AWS, "us-west-2")]
    existing = [(M.BackendType.AWS, "us-west-2")]
    want = [(M.BackendType.AWS, "us-east-1"), (M.BackendType.AWS, "eu-west-1")]
    assert scheduler.placement_group_keys(offers, existing) == want == ref.placement_group_keys(offers, existing)
