"""Items without a disk size on backends with configurable disks (AWS, GCP, Azure ...).

In the reference such an item becomes an offer whose disk is sized to the job
(core/backends/base/offers.py:46-83, 161-175): it matches a job iff the job's disk range
(DEFAULT_DISK = 100GB.. when the job states none) overlaps the backend's configurable range.
``OfferCatalog.from_catalog_items`` evaluates that overlap on the device for every (job, item) and
sizes the disk when a row is turned into an offer; the expected lists come from running the
reference flow -- get_catalog_offers per backend and job, merge by price -- on the oracle."""

import heapq

import numpy as np
import pytest

import device_emulator as emu
from dstack_b200 import models as M
from dstack_b200 import services as S
from dstack_b200._gpuhunt import CatalogItem
from dstack_b200.encode import NO_ROW, encode_table, pack_jobs
from dstack_b200.ingest import columns_from_backend_items, job_columns
from oracle import columnar
from oracle import reference_path as ref
from oracle.gpuhunt_restated import Catalog as RefCatalog

BT = M.BackendType
GB = M.Memory.parse


def _items(provider, n, seed, disks):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        gpus = int(rng.choice([0, 1, 4]))
        out.append(CatalogItem(
            instance_name=f"{provider}-type-{int(rng.integers(0, 12))}", location=f"{provider}-r{int(rng.integers(0, 3))}",
            price=float(np.round(rng.uniform(0.1, 9.0), 2)), cpu=int(rng.choice([2, 8, 32])), memory=float(rng.choice([8, 64, 256])),
            gpu_count=gpus, gpu_name="A100" if gpus else None, gpu_memory=80.0 if gpus else None, spot=bool(rng.random() < 0.4),
            disk_size=disks[int(rng.integers(0, len(disks)))], gpu_vendor="nvidia" if gpus else None, provider=provider))
    return sorted(out, key=lambda it: it.price)


class _Backend:
    def __init__(self, btype):
        self.TYPE = btype


CASES = [
    (_Backend(BT.AWS), _items("aws", 60, 1, [None]), M.Range[M.Memory](min=GB("1GB"), max=None)),
    (_Backend(BT.GCP), _items("gcp", 50, 2, [None, 100.0, 1000.0]), M.Range[M.Memory](min=GB("10GB"), max=GB("300GB"))),
    (_Backend(BT.RUNPOD), _items("runpod", 40, 3, [50.0, 200.0, 2000.0]), M.Range[M.Memory](min=GB("1GB"), max=None)),
]
DISKS = [None, "500GB..", "50GB..80GB", "..5GB", "200GB..250GB", "100GB..", "400GB..1000GB"]


def _requirements(disk, gpu=None):
    res = M.ResourcesSpec(gpu=gpu) if disk is None else M.ResourcesSpec(gpu=gpu, disk=disk)
    return M.Requirements(resources=res, max_price=None, spot=None)


def _reference_offers(req):
    """get_catalog_offers per backend (disk sized per job, infeasible items dropped), merged by price."""
    lists = []
    for backend, items, conf in CASES:
        cat = RefCatalog()
        cat.add_items("lambdalabs" if backend.TYPE == BT.LAMBDA else backend.TYPE.value, items)
        lists.append(ref.get_catalog_offers(backend.TYPE, requirements=req, configurable_disk_size=conf, catalog=cat))
    return list(heapq.merge(*lists, key=lambda o: o.price))


def _filters():
    reqs = [_requirements(d, gpu) for d in DISKS for gpu in (None, "A100:1..4")]
    return reqs, [S.job_filter(M.Profile(name="p"), r) for r in reqs]


def test_overlap_predicate_on_the_emulator_equals_reference_flow():
    """CPU: encoder + documented device semantics (tests/device_emulator.py) against the reference flow."""
    cols = columns_from_backend_items([(b.TYPE, items, conf) for b, items, conf in CASES])
    table = encode_table(cols)
    assert table.disk_conf_ranges and table.columns[table.col("disk")].kind == 2  # a bitset column
    reqs, filters = _filters()
    jc = job_columns(table, filters)
    packed = pack_jobs(table, jc)
    idx, cnt, _ = emu.topk(table, packed, 200)
    oracle = columnar.topk(cols, jc, 200)
    assert np.array_equal(cnt.astype(np.int64), oracle.count)
    flat = [(b, it) for b, items, _ in CASES for it in items]
    some = 0
    for j, req in enumerate(reqs):
        want = _reference_offers(req)
        rows = idx[j][idx[j] != NO_ROW]
        got = [flat[table.perm[r]] for r in rows]
        assert int(cnt[j]) == len(want), (DISKS[j // 2], int(cnt[j]), len(want))
        assert [(b.TYPE, it.instance_name, it.location, it.price) for b, it in got] == \
               [(o.backend, o.instance.name, o.region, o.price) for o in want]
        some += len(want) > 0
    assert some >= 8
    # the advisor's case: only disk-less AWS items, a job that wants 500 GB -> all of them, sized 500 GB
    aws_only = _reference_offers(_requirements("500GB.."))
    assert any(o.backend == BT.AWS and o.instance.resources.disk.size_mib == 500 * 1024 for o in aws_only)


@pytest.mark.gpu
def test_catalog_of_items_sizes_disks_like_the_reference(matcher):
    catalog = S.OfferCatalog.from_catalog_items(CASES, matcher=matcher)
    try:
        reqs, filters = _filters()
        top = catalog.match_topk(filters, 200)
        full = catalog.offers_for(filters)
        for j, req in enumerate(reqs):
            want = _reference_offers(req)
            assert int(top.count[j]) == len(want)
            for got in (catalog.pairs(top.index[j], filters[j]), full[j]):
                assert [(o.backend, o.instance.name, o.region, o.price, o.instance.resources.disk.size_mib,
                         o.instance.resources.description) for _, o in got] == \
                       [(o.backend, o.instance.name, o.region, o.price, o.instance.resources.disk.size_mib,
                         o.instance.resources.description) for o in want]
    finally:
        catalog.close()
