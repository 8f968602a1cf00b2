"""filter_pool_instances / shared-block search / get_offers_by_requirements.

CPU: the cases the reference's own tests pin, on the oracle and on the
product's reference-signature host logic.  GPU: `PoolIndex`, `SharedPoolIndex`, `OfferCatalog`
against the oracle on seeded synthetic pools / catalogs with every filter combination.
"""

import asyncio
import types
import uuid

import numpy as np
import pytest

from dstack_b200 import models as M
from dstack_b200 import pools as PP
from dstack_b200 import services as S
from workloads import synth
from oracle import reference_path as ref

BT = M.BackendType


def make_offer(backend=BT.AWS, region="eu-west", gpu_count=0, cpu_count=2, memory_gib=12, disk_gib=100.0, spot=False,
               price=1.0, availability=M.InstanceAvailability.AVAILABLE, name="instance", total_blocks=1):
    """The reference's fixture, server/testing/common.py:580-610."""
    gpus = [M.Gpu(name="T4", memory_mib=16384, vendor="nvidia")] * gpu_count
    return M.InstanceOfferWithAvailability(
        backend=backend, region=region, price=price, availability=availability, total_blocks=total_blocks,
        instance=M.InstanceType(name=name, resources=M.Resources(
            cpus=cpu_count, memory_mib=int(memory_gib * 1024), gpus=gpus, spot=spot,
            disk=M.Disk(size_mib=int(disk_gib * 1024)), description="")))


def make_instance(i=0, offer=None, status=M.InstanceStatus.IDLE, backend=BT.AWS, region="eu-west", fleet_id=None,
                  unreachable=False, total_blocks=1, busy_blocks=0, zone=None, has_jpd=True, name=None):
    offer = offer if offer is not None else make_offer(backend=backend, region=region)
    jpd = M.JobProvisioningData(backend=backend, instance_type=offer.instance, instance_id=f"i-{i}", region=region,
                                availability_zone=zone, price=offer.price, username="ubuntu", dockerized=True)
    return types.SimpleNamespace(
        id=uuid.UUID(int=i + 1), name=name or f"instance-{i}", fleet_id=fleet_id, status=status, unreachable=unreachable,
        backend=backend, region=region, price=offer.price, offer=offer.json(),
        job_provisioning_data=jpd.json() if has_jpd else None, total_blocks=total_blocks, busy_blocks=busy_blocks)


def volume(backend, region, zone=None):
    return types.SimpleNamespace(configuration=types.SimpleNamespace(backend=backend, region=region),
                                 provisioning_data=None if zone is None else types.SimpleNamespace(availability_zone=zone))


DEFAULT_REQ = M.Requirements(resources=M.ResourcesSpec(), max_price=None, spot=None)


# ------------------------------------------------------------------------------------------ CPU: reference-pinned cases


def test_default_requirements_match_idle_fixture_instance():
    # test_process_submitted_jobs.py:422-455 (cpu >= 2, mem >= 8GB, disk >= 100GB vs 2 / 12GiB / 100GiB)
    inst = make_instance()
    prof = M.Profile(name="p")
    assert ref.filter_pool_instances([inst], prof, requirements=DEFAULT_REQ, status=M.InstanceStatus.IDLE) == [inst]
    busy = make_instance(1, status=M.InstanceStatus.BUSY)
    assert ref.filter_pool_instances([busy], prof, requirements=DEFAULT_REQ, status=M.InstanceStatus.IDLE) == []


def test_volume_restricts_backend_region_zone():
    # test_process_submitted_jobs.py:459-526, pools.py:428-438,477-483
    a = make_instance(0, backend=BT.AWS, region="us-east-1", zone="us-east-1a")
    b = make_instance(1, backend=BT.AWS, region="us-east-1", zone="us-east-1b")
    c = make_instance(2, backend=BT.GCP, region="us-east-1")
    prof = M.Profile(name="p")
    vols = [[volume(BT.AWS, "us-east-1", zone="us-east-1a")]]
    assert ref.filter_pool_instances([a, b, c], prof, requirements=DEFAULT_REQ, volumes=vols) == [a]
    assert ref.filter_pool_instances([a, b, c], prof, requirements=DEFAULT_REQ, volumes=[[volume(BT.AWS, "us-east-1")]]) == [a, b]


def test_shared_instance_smallest_block():
    # test_process_submitted_jobs.py:530-570: 8 GPU / 64 CPU / 128 GiB, total_blocks 4, busy 1 -> blocks 1
    offer = make_offer(gpu_count=8, cpu_count=64, memory_gib=128, total_blocks=4)
    inst = make_instance(0, offer=offer, total_blocks=4, busy_blocks=1)
    req = M.Requirements(resources=M.ResourcesSpec(gpu="1..2"), max_price=None, spot=None)
    [(got_inst, got_offer)] = ref.get_shared_pool_instances_with_offers([inst], M.Profile(name="p"), req, idle_only=True)
    assert got_inst is inst
    assert (got_offer.blocks, got_offer.total_blocks) == (1, 4)
    assert got_offer.availability == M.InstanceAvailability.IDLE
    assert (got_offer.instance.resources.cpus, len(got_offer.instance.resources.gpus)) == (16, 2)
    full = make_instance(1, offer=offer, total_blocks=4, busy_blocks=4)
    assert ref.get_shared_pool_instances_with_offers([full], M.Profile(name="p"), req, idle_only=True) == []
    [(_, busy_offer)] = ref.get_shared_pool_instances_with_offers([full], M.Profile(name="p"), req, idle_only=False)
    assert busy_offer.availability == M.InstanceAvailability.BUSY


class CountingBackend(synth.FakeBackend):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.calls = 0

    def compute(self):
        self.calls += 1
        return super().compute()


def run(coro):
    return asyncio.run(coro)


def _plan(get_offers, backends, profile, **kw):
    offers = get_offers(backends, profile, DEFAULT_REQ, **kw)
    return ref.plan_slice([o for _, o in offers], 50), offers


def product_get_offers(backends, profile, requirements, **kw):
    project = types.SimpleNamespace(backends=list(backends))
    return run(S.get_offers_by_requirements(project, profile, requirements, **kw))


@pytest.mark.parametrize("get_offers", [ref.get_offers_by_requirements, product_get_offers], ids=["oracle", "product"])
def test_run_plan_order_total_and_max_price(get_offers):
    # test_runs.py:695-877 (TestGetRunPlan): aws $1 before runpod $2, total_offers, max_price
    aws = CountingBackend(BT.AWS, [make_offer(BT.AWS, price=1.0)])
    runpod = CountingBackend(BT.RUNPOD, [make_offer(BT.RUNPOD, price=2.0)])
    (first, total, top), offers = _plan(get_offers, [runpod, aws], M.Profile(name="p"))
    assert [o.backend for o in first] == [BT.AWS, BT.RUNPOD] and total == 2 and top == 2.0
    assert [b.TYPE for b, _ in offers] == [BT.AWS, BT.RUNPOD]
    # privileged: runpod has no create-instance support -> only aws is even asked (offers.py:59-62,81-82)
    aws.calls = runpod.calls = 0
    (first, total, top), _ = _plan(get_offers, [aws, runpod], M.Profile(name="p"), privileged=True)
    assert [o.backend for o in first] == [BT.AWS] and (aws.calls, runpod.calls) == (1, 0)
    # no backend survives -> nobody is asked, empty plan
    aws.calls = runpod.calls = 0
    (first, total, top), _ = _plan(get_offers, [runpod], M.Profile(name="p"), instance_mounts=True)
    assert (first, total, top, runpod.calls) == ([], 0, None, 0)


@pytest.mark.parametrize("get_offers", [ref.get_offers_by_requirements, product_get_offers], ids=["oracle", "product"])
def test_ordering_ties_availability_and_filters(get_offers):
    na = M.InstanceAvailability.NOT_AVAILABLE
    a = [make_offer(BT.AWS, price=1.0, name="a1"), make_offer(BT.AWS, price=1.0, name="a2", availability=na),
         make_offer(BT.AWS, price=3.0, name="a3", region="us")]
    g = [make_offer(BT.GCP, price=0.5, name="g1", availability=M.InstanceAvailability.NO_QUOTA),
         make_offer(BT.GCP, price=1.0, name="g2"), make_offer(BT.GCP, price=3.0, name="g3", region="us")]
    aws, gcp = synth.FakeBackend(BT.AWS, a), synth.FakeBackend(BT.GCP, g)
    names = lambda res: [o.instance.name for _, o in res]
    # price ties go to the earlier backend, unavailable offers go last regardless of price
    assert names(get_offers([aws, gcp], M.Profile(name="p"), DEFAULT_REQ)) == ["a1", "g2", "a3", "g3", "g1", "a2"]
    assert names(get_offers([gcp, aws], M.Profile(name="p"), DEFAULT_REQ)) == ["g2", "a1", "g3", "a3", "g1", "a2"]
    assert names(get_offers([aws, gcp], M.Profile(name="p"), DEFAULT_REQ, exclude_not_available=True)) == ["a1", "g2", "a3", "g3"]
    assert names(get_offers([aws, gcp], M.Profile(name="p", regions=["us"]), DEFAULT_REQ)) == ["a3", "g3"]
    assert names(get_offers([aws, gcp], M.Profile(name="p", instance_types=["g2", "a2"]), DEFAULT_REQ)) == ["g2", "a2"]
    assert names(get_offers([aws, gcp], M.Profile(name="p", backends=[BT.GCP]), DEFAULT_REQ)) == ["g2", "g3", "g1"]
    # `backends: [dstack]` is reset to None (offers.py:39-44) -- mutation is part of the contract
    prof = M.Profile(name="p", backends=[BT.DSTACK])
    assert len(get_offers([aws, gcp], prof, DEFAULT_REQ)) == 6 and prof.backends is None
    # multinode keeps only backends with multinode support; master node pins backend and region
    rp = synth.FakeBackend(BT.RUNPOD, [make_offer(BT.RUNPOD, price=0.1, name="r1")])
    assert names(get_offers([rp, aws], M.Profile(name="p"), DEFAULT_REQ, multinode=True)) == ["a1", "a3", "a2"]
    jpd = M.JobProvisioningData(backend=BT.GCP, instance_type=g[0].instance, region="us", price=1.0)
    assert names(get_offers([aws, gcp], M.Profile(name="p"), DEFAULT_REQ, multinode=True,
                            master_job_provisioning_data=jpd)) == ["g3"]
    # a DSTACK backend may serve offers of other backend types: kept by the pre-filter, checked per offer
    sky = synth.FakeBackend(BT.DSTACK, [make_offer(BT.AWS, price=0.2, name="s-aws"), make_offer(BT.GCP, price=0.3, name="s-gcp")])
    assert names(get_offers([aws, sky], M.Profile(name="p", backends=[BT.AWS]), DEFAULT_REQ)) == ["s-aws", "a1", "a3", "a2"]
    # blocks: 4-way split needs cpu % 4 == 0; `auto` = as many as possible; total_blocks is written back
    big = synth.FakeBackend(BT.AWS, [make_offer(BT.AWS, cpu_count=32, gpu_count=8, memory_gib=64, name="b8"),
                                      make_offer(BT.AWS, cpu_count=6, gpu_count=0, name="c6")])
    res = get_offers([big], M.Profile(name="p"), DEFAULT_REQ, blocks=4)
    assert [(o.instance.name, o.total_blocks) for _, o in res] == [("b8", 4)]
    res = get_offers([big], M.Profile(name="p"), DEFAULT_REQ, blocks="auto")
    assert [(o.instance.name, o.total_blocks) for _, o in res] == [("b8", 8), ("c6", 6)]


def test_backend_errors_are_skipped():
    # backends/__init__.py:427-440: BackendError -> warning, any other exception -> error; both skipped
    def raising(exc):
        b = synth.FakeBackend(BT.GCP, [])
        b.compute().get_offers_cached = lambda requirements: (_ for _ in ()).throw(exc)
        return b

    ok = synth.FakeBackend(BT.AWS, [make_offer(BT.AWS)])
    for exc in (S.BackendError("quota api down"), RuntimeError("boom")):
        assert len(product_get_offers([raising(exc), ok], M.Profile(name="p"), DEFAULT_REQ)) == 1


# ------------------------------------------------------------------------------------------ GPU


def _pool_filter_cases(rows):
    fleets = [r.fleet_id for r in rows if r.fleet_id is not None]
    some = rows[len(rows) // 3]
    jpd = M.JobProvisioningData.parse_raw(next(r.job_provisioning_data for r in rows if r.job_provisioning_data))
    zone = next(M.JobProvisioningData.parse_raw(r.job_provisioning_data).availability_zone for r in rows
                if r.job_provisioning_data and M.JobProvisioningData.parse_raw(r.job_provisioning_data).availability_zone)
    gpu_req = M.Requirements(resources=M.ResourcesSpec(gpu="16GB..:1..8"), max_price=3.0, spot=None)
    cases = []
    for shared in (False, True):
        cases += [
            dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, status=M.InstanceStatus.IDLE, shared=shared),
            dict(profile=M.Profile(name="p"), requirements=None, shared=shared),
            dict(profile=M.Profile(name="p"), requirements=gpu_req, shared=shared),
            dict(profile=M.Profile(name="p", backends=[BT.AWS, BT.GCP], regions=[some.region]), requirements=DEFAULT_REQ, shared=shared),
            dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, fleet_model=types.SimpleNamespace(id=fleets[0]), shared=shared),
            dict(profile=M.Profile(name="p", instance_name=some.name), requirements=None, shared=shared),
            dict(profile=M.Profile(name="p", instance_types=[jpd.instance_type.name, "nope"]), requirements=DEFAULT_REQ, shared=shared),
            dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, multinode=True, shared=shared),
            dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, multinode=True, master_job_provisioning_data=jpd, shared=shared),
            dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ,
                 volumes=[[volume(jpd.backend, jpd.region, zone=zone)]], shared=shared),
            dict(profile=M.Profile(name="p"), requirements=M.Requirements(
                resources=M.ResourcesSpec(cpu="64.."), max_price=0.01, spot=True), shared=shared),
        ]
    return cases


@pytest.mark.gpu
def test_gpu_pool_index_equals_oracle(matcher):
    rows = synth.make_pool(3000, seed=2, n_fleets=40)
    index = PP.PoolIndex(rows, matcher=matcher)
    try:
        nonempty = 0
        for kw in _pool_filter_cases(rows):
            profile = kw.pop("profile")
            want = ref.filter_pool_instances(rows, profile, **kw)
            got = index.filter(profile, **kw)
            assert [id(x) for x in got] == [id(x) for x in want], kw  # same objects, input order
            nonempty += bool(want)
        assert nonempty >= 12
    finally:
        index.close()
    kw = _pool_filter_cases(rows)[0]
    profile = kw.pop("profile")
    assert PP.filter_pool_instances(rows[:200], profile, matcher=matcher, **kw) == ref.filter_pool_instances(rows[:200], profile, **kw)
    assert PP.filter_pool_instances([], profile, matcher=matcher, **kw) == []


@pytest.mark.gpu
def test_gpu_shared_pool_offers_equal_oracle(matcher):
    rows = synth.make_pool(1500, seed=5, n_fleets=20)
    reqs = [DEFAULT_REQ, M.Requirements(resources=M.ResourcesSpec(gpu="1..2"), max_price=None, spot=None),
            M.Requirements(resources=M.ResourcesSpec(cpu="8..", memory="32GB.."), max_price=5.0, spot=None)]
    seen = 0
    for req in reqs:
        for idle_only in (False, True):
            want = ref.get_shared_pool_instances_with_offers(rows, M.Profile(name="p"), req, idle_only=idle_only)
            got = PP.get_shared_pool_instances_with_offers(rows, M.Profile(name="p"), req, idle_only=idle_only, matcher=matcher)
            assert [(id(i), o.json()) for i, o in got] == [(id(i), o.json()) for i, o in want]
            seen += len(want)
    assert seen > 50


@pytest.mark.gpu
def test_gpu_offer_catalog_equals_oracle(matcher):
    cat = synth.make_catalog(4000, seed=13, n_shapes=300)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    backends = [b for b, _ in per_backend]
    catalog = S.OfferCatalog(per_backend, matcher=matcher)
    gpu_req = M.Requirements(resources=M.ResourcesSpec(gpu="16GB..:1..8"), max_price=6.0, spot=None)
    master = M.JobProvisioningData(backend=BT.GCP, instance_type=per_backend[1][1][0].instance,
                                   region=per_backend[1][1][5].region, price=1.0)
    calls = [
        dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ),
        dict(profile=M.Profile(name="p"), requirements=gpu_req, exclude_not_available=True),
        dict(profile=M.Profile(name="p", backends=[BT.AWS, BT.OCI], regions=[per_backend[0][1][3].region]), requirements=DEFAULT_REQ),
        dict(profile=M.Profile(name="p", instance_types=["shape-0007", "shape-0100", "shape-0200"]), requirements=DEFAULT_REQ),
        dict(profile=M.Profile(name="p"), requirements=gpu_req, multinode=True),
        dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, multinode=True, master_job_provisioning_data=master),
        dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, privileged=True),
        dict(profile=M.Profile(name="p", reservation="cr-1"), requirements=DEFAULT_REQ),
        dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, volumes=[[volume(BT.AWS, per_backend[0][1][0].region)]]),
        dict(profile=M.Profile(name="p"), requirements=gpu_req, blocks="auto"),
        dict(profile=M.Profile(name="p"), requirements=gpu_req, blocks=2),
        dict(profile=M.Profile(name="p"), requirements=DEFAULT_REQ, blocks=4),
        dict(profile=M.Profile(name="p", backends=[BT.DSTACK]), requirements=DEFAULT_REQ),
    ]
    try:
        filters = []
        wants = []
        for kw in calls:
            kw = dict(kw)
            profile, req = kw.pop("profile"), kw.pop("requirements")
            filters.append(S.job_filter(profile, req, **kw))
            wants.append(ref.get_offers_by_requirements(backends, profile.copy(deep=True), req, **kw))
        full = catalog.offers_for(filters)
        top = catalog.match_topk(filters, 15)
        for j, want in enumerate(wants):
            assert [(b.TYPE, id(o)) for b, o in full[j]] == [(b.TYPE, id(o)) for b, o in want], calls[j]
            first, total, top_price = ref.plan_slice([o for _, o in want], 15)
            assert [id(o) for _, o in catalog.pairs(top.index[j])] == [id(o) for o in first]
            assert top.count[j] == total
            assert (np.isnan(top.max_price[j]) and top_price is None) or top.max_price[j] == top_price
        assert sum(bool(w) for w in wants) >= 9
    finally:
        catalog.close()


@pytest.mark.gpu
def test_pool_index_refresh_patches_rows_in_place(matcher):
    """A scheduler pass changes statuses, reachability and fleets; PoolIndex.refresh rewrites those
    rows on the device and every job gets what the reference's filter gives on the changed pool."""
    import test_pool_parse_cache as T
    from workloads import configs

    rng = np.random.default_rng(11)
    rows = synth.make_pool(4000, seed=12, n_fleets=8)
    calls = configs.config2_jobs(rows, 40, seed=13)
    index = PP.PoolIndex(rows, matcher=matcher)
    try:
        for tick in range(3):
            changed = T._status_tick(rows, rng, share=0.02)
            assert index.refresh(changed) is True and index.rebuilds == 1
            filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"],
                                       fleet_model=kw.get("fleet_model"), multinode=kw.get("multinode", False),
                                       master_job_provisioning_data=kw.get("master_job_provisioning_data"),
                                       volumes=None, shared=False) for p, kw in calls]
            got = index.match(filters)
            for j in (0, 1, 2, 3, 4, 5 + tick):
                want = ref.filter_pool_instances(rows, calls[j][0], **calls[j][1])
                assert len(want) == len(got[j]) and all(rows[i] is w for i, w in zip(got[j], want)), (tick, j)
        # an instance moving to a fleet the pool has never seen cannot be patched: rebuilt instead
        rows[7].fleet_id = uuid.UUID(int=987654321)
        assert index.refresh([7]) is False and index.rebuilds == 2
        want = ref.filter_pool_instances(rows, calls[0][0], **calls[0][1])
        got = index.match([PP._pool_filter(calls[0][0], requirements=calls[0][1]["requirements"],
                                           status=calls[0][1]["status"], fleet_model=calls[0][1].get("fleet_model"),
                                           multinode=calls[0][1].get("multinode", False),
                                           master_job_provisioning_data=calls[0][1].get("master_job_provisioning_data"),
                                           volumes=None, shared=False)])[0]
        assert len(want) == len(got) and all(rows[i] is w for i, w in zip(got, want))
    finally:
        index.close()


def test_job_columns_keeps_the_table_independent_numbers_on_the_filter():
    """``job_columns`` computes a filter's table-independent numbers once and keeps them on the
    object (a pending job is matched again every pass): the second use must give the same columns
    as the first, and ``dataclasses.replace`` must give a filter that is computed afresh."""
    import dataclasses

    from dstack_b200.encode import encode_table
    from dstack_b200.ingest import job_columns
    from workloads import synth

    cat = synth.make_catalog(2_000, seed=5, n_shapes=64)
    table = encode_table(cat.columns)
    req = M.Requirements(resources=M.ResourcesSpec(cpu="4..", memory="16GB..", gpu="24GB..:1..2"), max_price=3.5,
                         spot=False)
    f = S.job_filter(M.Profile(name="p", regions=["eu-west-1"]), req, exclude_not_available=True)
    first = job_columns(table, [f])
    assert "_scalars" in f.__dict__
    again = job_columns(table, [f])
    for name, a in vars(first).items():
        b = getattr(again, name)
        if isinstance(a, np.ndarray):
            assert np.array_equal(a, b, equal_nan=a.dtype.kind == "f"), name
        else:
            assert a == b or (a is None and b is None), name
    g = dataclasses.replace(f, exclude_not_available=False)
    assert "_scalars" not in g.__dict__
    assert not job_columns(table, [g]).exclude_not_available[0] and first.exclude_not_available[0]
