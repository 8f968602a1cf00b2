"""Golden vectors produced by the REFERENCE's own functions (tests/golden/make_golden.py).

CPU part: the oracle restatement (oracle/reference_path.py) and the product's host-side mirror
(dstack_b200/models.py, offers.py, services.py) reproduce them.  GPU part: the CUDA path
(`match_requirements`, `GpuCatalog.query` / `get_catalog_offers`) reproduces them bit for bit.
"""

import json
import os
from dataclasses import asdict

import pytest

from dstack_b200 import models as M
from dstack_b200 import offers as P
from dstack_b200 import services as S
from dstack_b200._gpuhunt import AcceleratorVendor, CatalogItem
from oracle import gpuhunt_restated
from oracle import reference_path as ref

try:
    from pydantic.v1 import ValidationError, parse_obj_as
except ImportError:  # pragma: no cover
    from pydantic import ValidationError, parse_obj_as

G = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.json")))


def enc(v):
    if v is None or isinstance(v, (bool, int, str)):
        return v
    if isinstance(v, float):
        return float(v)
    if hasattr(v, "value"):
        return v.value
    if isinstance(v, (list, tuple)):
        return [enc(x) for x in v]
    raise TypeError(type(v))


def dc(obj):
    return {k: enc(v) for k, v in asdict(obj).items()}


def requirements(spec):
    return None if spec is None else M.Requirements.parse_obj(spec)


OFFERS = [M.InstanceOfferWithAvailability.parse_obj(o) for o in G["offers"]]


def catalog_items():
    items = []
    for d in G["catalog_items"]:
        d = dict(d)
        if d["gpu_vendor"] is not None:
            d["gpu_vendor"] = AcceleratorVendor(d["gpu_vendor"])
        items.append(CatalogItem(**d))
    return items


# ------------------------------------------------------------------------------------------ CPU


@pytest.mark.parametrize("fn", [ref.requirements_to_query_filter, P.requirements_to_query_filter], ids=["oracle", "product"])
def test_requirements_to_query_filter(fn):
    for case in G["requirements_to_query_filter"]:
        assert dc(fn(requirements(case["requirements"]))) == case["query_filter"], case["requirements"]


@pytest.mark.parametrize("fn", [ref.offer_to_catalog_item, P.offer_to_catalog_item], ids=["oracle", "product"])
def test_offer_to_catalog_item(fn):
    assert len(OFFERS) == len(G["offer_to_catalog_item"])
    for offer, want in zip(OFFERS, G["offer_to_catalog_item"]):
        assert dc(fn(offer)) == want


def test_offers_round_trip_json():
    for offer, raw in zip(OFFERS, G["offers"]):
        assert json.loads(offer.json()) == raw


def test_oracle_match_requirements():
    hit_something = 0
    for case in G["match_requirements"]:
        kept = ref.match_requirements(OFFERS, requirements(case["requirements"]))
        pos = {id(o): i for i, o in enumerate(OFFERS)}
        assert [pos[id(o)] for o in kept] == case["matched"], case["requirements"]
        hit_something += bool(case["matched"])
    assert hit_something >= 15


def _oracle_catalog():
    cat = gpuhunt_restated.Catalog(balance_resources=False, auto_reload=False)
    items = catalog_items()
    cat.add_items("aws", [it for i, it in enumerate(items) if i % 2 == 0])
    cat.add_items("lambdalabs", [it for i, it in enumerate(items) if i % 2 == 1])
    return cat


def _check_catalog_offers(get_catalog_offers, cat):
    for case in G["get_catalog_offers"]:
        kw = {}
        if case["disk"] is not None:
            kw["configurable_disk_size"] = M.Range[M.Memory](**case["disk"])
        got = get_catalog_offers(M.BackendType(case["backend"]), locations=case["locations"],
                                 requirements=requirements(case["requirements"]), catalog=cat, **kw)
        assert [json.loads(o.json()) for o in got] == case["offers"], case


def test_oracle_get_catalog_offers():
    _check_catalog_offers(ref.get_catalog_offers, _oracle_catalog())


@pytest.mark.parametrize("fn", [ref.choose_disk_size_mib, P.choose_disk_size_mib], ids=["oracle", "product"])
def test_choose_disk_size_mib(fn):
    default = M.Range[M.Memory](min=M.Memory.parse("1GB"), max=None)
    for case in G["choose_disk_size_mib"]:
        rd = None if case["req_disk"] is None else M.Range[M.Memory](**case["req_disk"])
        cd = default if case["configurable"] is None else M.Range[M.Memory](**case["configurable"])
        assert fn(case["item_disk"], rd, cd) == case["result"], case


def test_range_intersect():
    for case in G["range_intersect"]:
        a = M.Range[int](min=case["a"][0], max=case["a"][1])
        b = M.Range[int](min=case["b"][0], max=case["b"][1])
        for x, y in ((a, b), (b, a)):
            r = x.intersect(y)
            assert (None if r is None else {"min": r.min, "max": r.max}) == case["result"]


def test_model_parsing():
    for case in G["gpu_spec_parse"]:
        assert json.loads(parse_obj_as(M.GPUSpec, case["text"]).json()) == case["spec"], case["text"]
    for text in G["gpu_spec_errors"]:
        with pytest.raises(ValidationError):
            parse_obj_as(M.GPUSpec, text)
    for case in G["memory_parse"]:
        assert float(M.Memory.parse(case["text"])) == case["gb"]
    for case in G["range_parse"]:
        typ = M.Range[M.Memory] if case["memory"] else M.Range[int]
        r = parse_obj_as(typ, case["text"])
        assert {"min": r.min, "max": r.max} == case["range"]
    for case in G["gpu_vendor_default"]:
        g = M.Gpu(**case["kwargs"])
        assert (g.name, g.vendor.value) == (case["name"], case["vendor"])
    assert json.loads(M.ResourcesSpec().json()) == G["resources_defaults"]


@pytest.mark.parametrize("fn", [ref.is_divisible_into_blocks, S.is_divisible_into_blocks], ids=["oracle", "product"])
def test_is_divisible_into_blocks(fn):
    for case in G["is_divisible_into_blocks"]:
        assert fn(case["cpus"], case["gpus"], case["blocks"]) == (True, case["expected"])
    assert fn(32, 8, 3) == (False, 0)
    assert fn(6, 4, 4) == (False, 0)
    assert fn(8, 0, 0) == (False, 0)


# ------------------------------------------------------------------------------------------ GPU


@pytest.mark.gpu
def test_gpu_match_requirements_reproduces_reference(matcher):
    for case in G["match_requirements"]:
        kept = P.match_requirements(OFFERS, requirements(case["requirements"]), matcher=matcher)
        pos = {id(o): i for i, o in enumerate(OFFERS)}
        assert [pos[id(o)] for o in kept] == case["matched"], case["requirements"]  # same objects, input order


@pytest.mark.gpu
def test_gpu_catalog_reproduces_reference(matcher):
    cat = P.GpuCatalog(balance_resources=False, auto_reload=False, matcher=matcher)
    items = catalog_items()
    cat.add_items("aws", [it for i, it in enumerate(items) if i % 2 == 0])
    cat.add_items("lambdalabs", [it for i, it in enumerate(items) if i % 2 == 1])
    _check_catalog_offers(P.get_catalog_offers, cat)
