"""Generate the golden vectors in this directory from the REFERENCE's own code.

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py

It imports the reference modules unmodified through oracle/ref_shim.py (pydantic -> pydantic.v1
alias, pydantic_duality stub, gpuhunt = restated types + predicate) and records, as JSON,
inputs in a neutral form (spec strings / `.json()` dumps) and the outputs of:

  requirements_to_query_filter   core/backends/base/offers.py:112-146
  offer_to_catalog_item          core/backends/base/offers.py:86-109
  match_requirements             core/backends/base/offers.py:149-158
  get_catalog_offers             core/backends/base/offers.py:18-43 (+ catalog_item_to_offer :46-83)
  choose_disk_size_mib           core/backends/base/offers.py:161-175
  Range.intersect                core/models/resources.py:59-73   (table of tests/_internal/core/models/test_resources.py:192-206)
  GPUSpec / Range / Memory parsing  core/models/resources.py       (cases of test_resources.py:17-189)
  Gpu vendor defaulting          core/models/instances.py:21-36   (test_instances.py:6-36)

What this pins: the reference's OWN half of the path.  `gpuhunt.matches` inside
match_requirements / Catalog.query is the restatement in oracle/gpuhunt_restated.py (the real
package is not available here), so the predicate itself stays "parity unpinned" (DESIGN.md §5).
The `is_divisible_into_blocks` table is lifted from the reference's test
(tests/_internal/server/background/tasks/test_process_instances.py:486-497) because its module
imports sqlalchemy and cannot be imported here.
"""

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from oracle import ref_shim  # noqa: E402

ns = ref_shim.load()
R = ns.resources
I = ns.instances
RUNS = ns.runs
OFF = ns.offers
BT = ns.backends_base.BackendType
parse_obj_as = ns.pydantic.parse_obj_as


def enc(v):
    """QueryFilter / CatalogItem field -> JSON value."""
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
    from dataclasses import asdict

    return {k: enc(v) for k, v in asdict(obj).items()}


# ---------------------------------------------------------------------------------- requirement specs

REQUIREMENTS = [
    # docs/docs/concepts/tasks.md:161-173 and BASELINE config 1
    {"resources": {"gpu": "24GB..80GB:1..8"}, "max_price": None, "spot": None},
    {"resources": {"gpu": "40GB..80GB:4"}, "max_price": None, "spot": None},
    {"resources": {"gpu": "A100:80GB"}, "max_price": None, "spot": None},
    {"resources": {"gpu": "24GB..40GB:2"}, "max_price": None, "spot": None},
    {"resources": {}, "max_price": None, "spot": None},
    {"resources": {"cpu": "4..", "memory": "16GB..64GB"}, "max_price": 2.5, "spot": False},
    {"resources": {"cpu": 8, "memory": "32GB", "disk": "200GB.."}, "max_price": None, "spot": True},
    {"resources": {"cpu": "..16", "memory": "..128GB", "disk": None}, "max_price": 0.5, "spot": None},
    {"resources": {"gpu": "A10,A10G:2"}, "max_price": None, "spot": None},
    {"resources": {"gpu": "a100"}, "max_price": 10, "spot": None},
    {"resources": {"gpu": {"name": ["H100", "h200"], "count": "1..8"}}, "max_price": None, "spot": None},
    {"resources": {"gpu": "nvidia:16GB..:1.."}, "max_price": None, "spot": None},
    {"resources": {"gpu": "MI300X:AMD:8"}, "max_price": None, "spot": None},
    {"resources": {"gpu": "tpu:v5litepod-8"}, "max_price": None, "spot": None},
    {"resources": {"gpu": {"count": "0..2"}}, "max_price": None, "spot": None},
    {"resources": {"gpu": {"total_memory": "160GB..", "count": "1..8"}}, "max_price": None, "spot": None},
    {"resources": {"gpu": {"compute_capability": "8.0", "count": "1..8"}}, "max_price": None, "spot": None},
    {"resources": {"gpu": {"compute_capability": 7.5, "memory": "..24GB", "count": "1..4"}}, "max_price": 3, "spot": True},
    {"resources": {"gpu": {"name": "T4", "memory": "16GB"}, "cpu": "2..8", "memory": "8GB..32GB"}, "max_price": 1.0, "spot": None},
    {"resources": {"gpu": 2}, "max_price": None, "spot": None},
    {"resources": {"gpu": "16GB..32"}, "max_price": None, "spot": None},
    {"resources": {"memory": "15.25GB..", "cpu": "1.."}, "max_price": None, "spot": None},
    {"resources": {"gpu": {"vendor": "amd"}, "disk": "50GB..500GB"}, "max_price": None, "spot": None},
    {"resources": {"gpu": {"name": []}}, "max_price": None, "spot": None},
]


def make_offers():
    """A hand-built list of offers that exercises every bound incl. equality."""
    rng = np.random.Generator(np.random.PCG64(2024))
    gpus = [("T4", 16384, "nvidia"), ("A10G", 24576, "nvidia"), ("A100", 40960, "nvidia"), ("A100", 81920, "nvidia"),
            ("a100", 81920, "nvidia"), ("H100", 81920, "nvidia"), ("H200", 144384, "nvidia"), ("L4", 24576, "nvidia"),
            ("V100", 16384, "nvidia"), ("V100", 32768, "nvidia"), ("MI300X", 196608, "amd"), ("v5litepod-8", 0, "google"),
            ("GH200", 98304, "nvidia"), ("tpu-v3-8", 0, None)]
    offers = []
    backends = [BT.AWS, BT.GCP, BT.AZURE, BT.RUNPOD, BT.LAMBDA]
    for i in range(90):
        cpus = int(rng.choice([1, 2, 4, 8, 16, 32, 64, 96]))
        mem_mib = int(rng.choice([2048, 8192, 12288, 15616, 16384, 32768, 65536, 131072, 614, 1741]))
        n_gpu = int(rng.choice([0, 0, 1, 1, 2, 4, 8]))
        g = gpus[int(rng.integers(0, len(gpus)))]
        disk = int(rng.choice([51200, 102400, 102400, 204800, 512000]))
        res = I.Resources(cpus=cpus, memory_mib=mem_mib, spot=bool(rng.random() < 0.4),
                          gpus=[I.Gpu(name=g[0], memory_mib=g[1], vendor=g[2])] * n_gpu,
                          disk=I.Disk(size_mib=disk))
        offers.append(I.InstanceOfferWithAvailability(
            backend=backends[i % len(backends)], region=f"region-{i % 7}",
            price=float(np.round(rng.uniform(0.05, 12.0), 2)) if i % 9 else 2.5,
            instance=I.InstanceType(name=f"type-{i % 23}", resources=res),
            availability=I.InstanceAvailability.AVAILABLE))
    # the reference's own fixture: server/testing/common.py:580-610 (T4 16GiB, cpu 2, 12GiB, 100GiB, $1)
    for n_gpu in (0, 1, 8):
        offers.append(I.InstanceOfferWithAvailability(
            backend=BT.AWS, region="eu-west", price=1,
            instance=I.InstanceType(name="instance", resources=I.Resources(
                cpus=2 if n_gpu < 8 else 64, memory_mib=int((12 if n_gpu < 8 else 128) * 1024),
                gpus=[I.Gpu(name="T4", memory_mib=16384, vendor="nvidia")] * n_gpu, spot=False,
                disk=I.Disk(size_mib=int(100.0 * 1024)), description="")),
            availability=I.InstanceAvailability.AVAILABLE))
    return offers


def make_catalog_items():
    CI = ns.gpuhunt.CatalogItem
    V = ns.gpuhunt.AcceleratorVendor
    rng = np.random.Generator(np.random.PCG64(7))
    items = []
    for i in range(60):
        n_gpu = int(rng.choice([0, 1, 2, 8]))
        name, mem = [("A100", 80.0), ("T4", 16.0), ("L4", 24.0), ("H100", 80.0)][int(rng.integers(0, 4))]
        items.append(CI(
            instance_name=f"ci-{i % 17}", location=f"loc-{i % 5}", price=float(np.round(0.1 + 0.07 * i, 4)),
            cpu=int(rng.choice([2, 4, 8, 48])), memory=float(rng.choice([7.5, 15.25, 16.0, 0.6, 61.0, 192.0])),
            gpu_count=n_gpu, gpu_name=name if n_gpu else None, gpu_memory=mem if n_gpu else None,
            spot=bool(i % 3 == 0), disk_size=[None, None, 100.0, 500.0, 30.0][i % 5],
            gpu_vendor=V.NVIDIA if n_gpu else None, provider=None))
    return items


def main():
    out = {}

    # (a) requirements -> QueryFilter
    out["requirements_to_query_filter"] = []
    reqs = []
    for spec in REQUIREMENTS:
        req = RUNS.Requirements.parse_obj(spec)
        reqs.append(req)
        out["requirements_to_query_filter"].append({"requirements": spec, "query_filter": dc(OFF.requirements_to_query_filter(req))})
    out["requirements_to_query_filter"].append({"requirements": None, "query_filter": dc(OFF.requirements_to_query_filter(None))})

    # (b) offers -> CatalogItem, (c) match_requirements
    offers = make_offers()
    out["offers"] = [json.loads(o.json()) for o in offers]
    out["offer_to_catalog_item"] = [dc(OFF.offer_to_catalog_item(o)) for o in offers]
    out["match_requirements"] = []
    for spec, req in zip(REQUIREMENTS, reqs):
        kept = OFF.match_requirements(offers, req)
        assert all(any(k is o for o in offers) for k in kept)
        pos = {id(o): i for i, o in enumerate(offers)}
        out["match_requirements"].append({"requirements": spec, "matched": [pos[id(o)] for o in kept]})
    out["match_requirements"].append({"requirements": None, "matched": list(range(len(offers)))})
    assert OFF.match_requirements(offers, None) == offers

    # (d) get_catalog_offers over a restated Catalog
    items = make_catalog_items()
    cat = ns.gpuhunt.Catalog(balance_resources=False, auto_reload=False)
    cat.add_items("aws", [it for i, it in enumerate(items) if i % 2 == 0])
    cat.add_items("lambdalabs", [it for i, it in enumerate(items) if i % 2 == 1])
    out["catalog_items"] = [dc(it) for it in items]
    out["get_catalog_offers"] = []
    cases = [
        dict(backend="aws", locations=None, req=0, disk=None),
        dict(backend="aws", locations=["loc-0", "loc-2"], req=4, disk=None),
        dict(backend="lambda", locations=None, req=4, disk=None),
        dict(backend="lambda", locations=None, req=5, disk={"min": 50, "max": 150}),
        dict(backend="aws", locations=None, req=6, disk={"min": 1, "max": 100}),
        dict(backend="aws", locations=None, req=7, disk=None),
        dict(backend="aws", locations=None, req=None, disk=None),
    ]
    for c in cases:
        req = None if c["req"] is None else reqs[c["req"]]
        kw = {}
        if c["disk"] is not None:
            kw["configurable_disk_size"] = R.Range[R.Memory](**c["disk"])
        res = OFF.get_catalog_offers(BT(c["backend"]), locations=c["locations"], requirements=req, catalog=cat, **kw)
        out["get_catalog_offers"].append({**c, "requirements": None if c["req"] is None else REQUIREMENTS[c["req"]],
                                          "offers": [json.loads(o.json()) for o in res]})

    # (e) small tables
    cfg = R.Range[R.Memory](min=R.Memory.parse("1GB"), max=None)
    out["choose_disk_size_mib"] = []
    for item_disk, req_disk, conf in [
        (100.0, None, None), (None, None, None), (0.0, {"min": 200, "max": None}, None),
        (None, {"min": 50, "max": 80}, None), (None, {"min": 50, "max": 80}, {"min": 100, "max": 200}),
        (None, {"min": None, "max": 80}, {"min": 30, "max": None}), (30.5, {"min": 500, "max": None}, None),
        (None, {"min": 0.3, "max": None}, {"min": 0.25, "max": None}), (None, {"min": 10.0005, "max": None}, None),
    ]:
        rd = None if req_disk is None else R.Range[R.Memory](**req_disk)
        cd = cfg if conf is None else R.Range[R.Memory](**conf)
        out["choose_disk_size_mib"].append({"item_disk": item_disk, "req_disk": req_disk, "configurable": conf,
                                            "result": OFF.choose_disk_size_mib(item_disk, rd, cd)})

    def rj(r):
        return None if r is None else {"min": r.min, "max": r.max}

    out["range_intersect"] = []
    for a, b in [((1, 2), (3, 4)), ((1, 2), (2, 3)), ((1, 2), (1, 2)), ((1, 3), (2, 4)), ((1, 4), (2, 3)),
                 ((None, 1), (2, None)), ((None, 1), (1, None)), ((None, 2), (1, None)), ((None, 1), (None, 2)),
                 ((1, None), (2, None)), ((1, None), (None, 2))]:
        ra, rb = R.Range[int](min=a[0], max=a[1]), R.Range[int](min=b[0], max=b[1])
        assert rj(ra.intersect(rb)) == rj(rb.intersect(ra))
        out["range_intersect"].append({"a": list(a), "b": list(b), "result": rj(ra.intersect(rb))})

    out["gpu_spec_parse"] = []
    for text in ["1", "Nvidia", "google:v3-64", "tpu:v5p-1024", "v5litepod-64:TPU", "MI300X:AMD", "A100", "tpu-v3-2048",
                 "16GB", "A10,A10G:2", "16GB..32", "24GB..80GB:1..8", "A100:80GB", "40GB..80GB:4", "H100:8", "a100,h100:2..4:40GB.."]:
        out["gpu_spec_parse"].append({"text": text, "spec": json.loads(parse_obj_as(R.GPUSpec, text).json())})
    out["gpu_spec_errors"] = []
    for text in ["A100,:2", "A100:", "Nvidia:A100:2:AMD", "A100:2:3"]:
        try:
            parse_obj_as(R.GPUSpec, text)
            raise SystemExit(f"{text} should not parse")
        except ns.pydantic.ValidationError:
            out["gpu_spec_errors"].append(text)
    out["memory_parse"] = [{"text": t, "gb": float(R.Memory.parse(t))} for t in ["512MB", "1TB", "8GB", "16 gb", 4, 1.5, "2"]]
    out["range_parse"] = []
    for t in ["1..2", "4..", "..8", 3, "512MB..1TB"]:
        typ = R.Range[R.Memory] if isinstance(t, str) and "B" in t else R.Range[int]
        out["range_parse"].append({"text": t, "memory": typ is not R.Range[int], "range": rj(parse_obj_as(typ, t))})
    out["gpu_vendor_default"] = []
    for kw in [{"name": "A100", "memory_mib": 1}, {"name": "tpu-v3-8", "memory_mib": 0}, {"name": "MI300X", "memory_mib": 1, "vendor": "AMD"},
               {"name": "v5p", "memory_mib": 0, "vendor": "google"}]:
        g = I.Gpu(**kw)
        out["gpu_vendor_default"].append({"kwargs": kw, "name": g.name, "vendor": g.vendor.value})
    out["resources_defaults"] = json.loads(R.ResourcesSpec().json())

    # (f) lifted from the reference's test table (module not importable here)
    out["is_divisible_into_blocks"] = [
        {"cpus": c, "gpus": g, "blocks": b, "expected": e} for c, g, b, e in
        [(32, 8, 1, 1), (32, 8, 2, 2), (32, 8, 4, 4), (32, 8, "auto", 8), (4, 8, "auto", 4), (8, 8, "auto", 8),
         (32, 0, 1, 1), (32, 0, 2, 2), (32, 0, 4, 4), (32, 0, "auto", 32)]
    ]

    path = os.path.join(HERE, "reference_vectors.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", path, {k: len(v) if hasattr(v, "__len__") else 1 for k, v in out.items()})


if __name__ == "__main__":
    main()
