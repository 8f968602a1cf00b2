"""Known-answer table for the predicate.

The reference's tests hold NO vector for `gpuhunt.matches` (third-party, absent here), so this
hand-derived table is the de-facto pin of its semantics -- labelled as such: "parity unpinned"
in DESIGN.md §5 refers to exactly this.  Every case is checked on the CPU oracle, and (marked
gpu) on the CUDA path through `match_requirements`, which must agree with the oracle.
"""

import pytest

from dstack_b200 import models as M
from dstack_b200 import offers as P
from oracle import reference_path as ref


def offer(cpus=4, mem_gib=16.0, gpus=0, gpu="A100", gpu_gib=80.0, vendor=None, disk_gib=100.0, spot=False, price=1.0,
          backend="aws", region="us-east-1", name="i1", availability="available"):
    g = [M.Gpu(name=gpu, memory_mib=int(gpu_gib * 1024), vendor=vendor)] * gpus
    return M.InstanceOfferWithAvailability(
        backend=backend, region=region, price=price, availability=availability,
        instance=M.InstanceType(name=name, resources=M.Resources(
            cpus=cpus, memory_mib=int(mem_gib * 1024), gpus=g, spot=spot, disk=M.Disk(size_mib=int(disk_gib * 1024)))))


def req(max_price=None, spot=None, **resources):
    return M.Requirements(resources=M.ResourcesSpec(**resources), max_price=max_price, spot=spot)


# (requirements, offer, expected match)
CASES = [
    # defaults: cpu >= 2, memory >= 8GB, disk >= 100GB (resources.py:125-127,250)
    ("defaults vs fixture", req(), offer(cpus=2, mem_gib=12, disk_gib=100), True),
    ("cpu below default", req(), offer(cpus=1), False),
    ("memory below default", req(), offer(mem_gib=7.99), False),
    ("disk below default", req(), offer(disk_gib=99.0), False),
    # boundary equality on every bound
    ("cpu == min", req(cpu="4.."), offer(cpus=4), True),
    ("cpu == max", req(cpu="..4"), offer(cpus=4), True),
    ("cpu above max", req(cpu="2..3"), offer(cpus=4), False),
    ("memory == min", req(memory="16GB.."), offer(mem_gib=16), True),
    ("memory == max", req(memory="8GB..16GB"), offer(mem_gib=16), True),
    ("memory fractional below", req(memory="15.25GB.."), offer(mem_gib=15.0), False),
    ("memory fractional equal", req(memory="15.25GB.."), offer(mem_gib=15.25), True),
    ("disk == max", req(disk="50GB..100GB"), offer(disk_gib=100), True),
    ("disk above max", req(disk="50GB..99GB"), offer(disk_gib=100), False),
    ("disk unconstrained", req(disk=None), offer(disk_gib=1), True),
    ("max_price == price", req(max_price=1.0), offer(price=1.0), True),
    ("max_price below", req(max_price=0.9999), offer(price=1.0), False),
    # spot
    ("spot None / spot offer", req(spot=None), offer(spot=True), True),
    ("spot True / on-demand offer", req(spot=True), offer(spot=False), False),
    ("spot False / on-demand offer", req(spot=False), offer(spot=False), True),
    ("spot False / spot offer", req(spot=False), offer(spot=True), False),
    # no GPU requirement matches GPU offers too (res.gpu is None sets no bound)
    ("cpu job vs gpu offer", req(), offer(gpus=8), True),
    # default count 1..1
    ("A100 vs 1xA100", req(gpu="A100"), offer(gpus=1), True),
    ("A100 vs 2xA100", req(gpu="A100"), offer(gpus=2), False),
    ("A100 vs cpu offer", req(gpu="A100"), offer(gpus=0), False),
    ("count range hi", req(gpu="A100:1..8"), offer(gpus=8), True),
    ("count range above", req(gpu="A100:1..4"), offer(gpus=8), False),
    ("count 0..2 allows cpu offer", req(gpu={"count": "0..2"}), offer(gpus=0), True),
    # names: case-insensitive, set membership
    ("lower-case request", req(gpu="a100"), offer(gpus=1, gpu="A100"), True),
    ("lower-case offer", req(gpu="A100"), offer(gpus=1, gpu="a100"), True),
    ("name set", req(gpu="A10,A10G:2"), offer(gpus=2, gpu="A10G", gpu_gib=24), True),
    ("name not in set", req(gpu="A10,A10G:2"), offer(gpus=2, gpu="A100"), False),
    ("empty name list", req(gpu={"name": []}), offer(gpus=1), False),
    # vendor
    ("vendor match", req(gpu="nvidia"), offer(gpus=1), True),
    ("vendor mismatch", req(gpu="amd"), offer(gpus=1), False),
    ("vendor amd", req(gpu="MI300X:AMD:8"), offer(gpus=8, gpu="MI300X", gpu_gib=192, vendor="amd"), True),
    ("tpu prefix", req(gpu="tpu:v5litepod-8"), offer(gpus=1, gpu="tpu-v5litepod-8", gpu_gib=0), True),
    ("vendor filter vs cpu offer (count 0..1)", req(gpu={"vendor": "nvidia", "count": "0..1"}), offer(gpus=0), False),
    # gpu memory
    ("gpu mem == min", req(gpu="80GB"), offer(gpus=1, gpu_gib=80), True),
    ("gpu mem below", req(gpu="80GB.."), offer(gpus=1, gpu_gib=40), False),
    ("gpu mem range hi", req(gpu="24GB..80GB:1..8"), offer(gpus=4, gpu_gib=80), True),
    ("gpu mem bound, no gpu (count 0..)", req(gpu={"memory": "16GB..", "count": "0.."}), offer(gpus=0), False),
    # total memory
    ("total == min", req(gpu={"total_memory": "160GB..", "count": "1..8"}), offer(gpus=2, gpu_gib=80), True),
    ("total below", req(gpu={"total_memory": "161GB..", "count": "1..8"}), offer(gpus=2, gpu_gib=80), False),
    ("total with no gpus is 0", req(gpu={"total_memory": "..100GB", "count": "0.."}), offer(gpus=0), True),
    # compute capability
    ("cc equal", req(gpu={"compute_capability": "8.0"}), offer(gpus=1, gpu="A100"), True),
    ("cc below", req(gpu={"compute_capability": "9.0"}), offer(gpus=1, gpu="A100"), False),
    ("cc minor compare", req(gpu={"compute_capability": "8.6"}), offer(gpus=1, gpu="L4", gpu_gib=24), True),
    ("cc unknown gpu fails", req(gpu={"compute_capability": "7.0"}), offer(gpus=1, gpu="GH200", gpu_gib=96), False),
    ("cc vs cpu offer", req(gpu={"compute_capability": "7.0", "count": "0.."}), offer(gpus=0), False),
    # ---- gpu_count == 0 against jobs whose count range includes 0 (`gpu: 0..2` style).  gpuhunt is
    # absent here, so each row states the READING of upstream `matches` it assumes (all unpinned,
    # SURVEY App. A.1 / B) and the call-site evidence for it (core/backends/base/offers.py:128-144:
    # every one of these bounds is only ever set inside `if res.gpu:`).
    # reading: any per-GPU memory bound, a maximum alone too, requires gpu_count > 0
    ("max gpu mem only, no gpu", req(gpu={"memory": "..16GB", "count": "0.."}), offer(gpus=0), False),
    ("max gpu mem only, small gpu", req(gpu={"memory": "..16GB", "count": "0.."}), offer(gpus=1, gpu="T4", gpu_gib=16), True),
    # reading: total memory of an item without GPUs is 0, compared like any other value
    ("max total only, no gpu", req(gpu={"total_memory": "..100GB", "count": "0.."}), offer(gpus=0), True),
    ("min total only, no gpu", req(gpu={"total_memory": "1GB..", "count": "0.."}), offer(gpus=0), False),
    # reading: a vendor filter needs the item to name that vendor; an item without GPUs names none
    ("vendor set, no gpu", req(gpu={"vendor": "amd", "count": "0.."}), offer(gpus=0), False),
    # reading: a name filter needs the item to name a GPU
    ("name set, no gpu", req(gpu={"name": ["A100"], "count": "0.."}), offer(gpus=0), False),
    # reading: the compute-capability test is NOT skipped for items without GPUs (unknown capability fails)
    ("min cc, no gpu, count 0..0", req(gpu={"compute_capability": "3.5", "count": "0..0"}), offer(gpus=0), False),
    # and the count range itself still admits the CPU-only item when nothing else is asked
    ("count 0.. alone, no gpu", req(gpu={"count": "0.."}), offer(gpus=0), True),
]


@pytest.mark.parametrize("name,requirements,o,expected", CASES, ids=[c[0] for c in CASES])
def test_oracle_known_answer(name, requirements, o, expected):
    assert (ref.match_requirements([o], requirements) == [o]) is expected


@pytest.mark.gpu
def test_gpu_known_answers(matcher):
    # all offers in one table, every case a job
    offers = [c[2] for c in CASES]
    for i, (name, requirements, o, expected) in enumerate(CASES):
        kept = P.match_requirements(offers, requirements, matcher=matcher)
        want = ref.match_requirements(offers, requirements)
        assert [id(x) for x in kept] == [id(x) for x in want], name
        assert (any(x is o for x in kept)) is expected, name
    assert P.match_requirements([], CASES[0][1], matcher=matcher) == []
    assert P.match_requirements(offers, None, matcher=matcher) == offers


def test_oracle_and_product_compute_capability_tables_agree():
    """The oracle carries its own table (typed from NVIDIA's public list); an entry that is wrong
    on one side only would otherwise be invisible to every parity test."""
    from dstack_b200 import _gpuhunt
    from oracle import gpuhunt_restated as G

    names = {g.name for g in _gpuhunt.KNOWN_NVIDIA_GPUS} | set(G.COMPUTE_CAPABILITY)
    assert len(names) >= 35
    for name in sorted(names):
        assert G.compute_capability_of(name) == _gpuhunt.compute_capability_of(name), name
        assert G.compute_capability_of(name.lower()) == G.compute_capability_of(name.upper())
    assert G.compute_capability_of("GH200") is None and G.compute_capability_of(None) is None
