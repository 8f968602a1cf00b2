"""Restatement of the third-party ``gpuhunt`` predicate.  TEST INFRASTRUCTURE (see oracle/__init__).

PARITY UNPINNED: ``gpuhunt`` (pinned ``>=0.0.19,<0.1.0`` at reference ``setup.py:37``) is not in
``/root/reference`` and cannot be installed here, and the reference's tests hold no vector
for it.  This file restates the published algorithm of ``gpuhunt.matches`` and the parts
of ``gpuhunt.Catalog.query`` the reference relies on; parity
is anchored on the reference's call sites:

  * ``gpuhunt.matches(item, q)``  <- ``core/backends/base/offers.py:156``,
    ``server/services/pools.py:501,536``
  * ``catalog.query(**asdict(q))`` <- ``core/backends/base/offers.py:34``
"""

from __future__ import annotations

import heapq
from typing import Dict, Iterable, List, Optional, Sequence

from dstack_b200._gpuhunt import CatalogItem, QueryFilter  # plain dataclasses: field names only

# Compute capability by GPU name -- the oracle's OWN table, typed from NVIDIA's public list
# ("CUDA GPUs - Compute Capability", developer.nvidia.com/cuda-gpus; architecture in the comment),
# so that a wrong entry in the product's table (dstack_b200/_gpuhunt.py) shows up as a parity
# failure instead of being shared by both sides.  tests/test_oracle_known_answers.py compares the two.
COMPUTE_CAPABILITY = {
    # Kepler / Maxwell
    "K80": (3, 7),         # Tesla K80, GK210
    "M60": (5, 2),         # Tesla M60, GM204
    # Pascal
    "P100": (6, 0),        # Tesla P100, GP100
    "P4": (6, 1),          # Tesla P4, GP104
    "P40": (6, 1),         # Tesla P40, GP102
    # Volta
    "V100": (7, 0),        # Tesla V100, GV100
    # Turing
    "T4": (7, 5),          # Tesla T4, TU104
    "RTX6000": (7, 5),     # Quadro RTX 6000, TU102
    "RTX8000": (7, 5),     # Quadro RTX 8000, TU102
    # Ampere
    "A100": (8, 0),        # GA100
    "A30": (8, 0),         # GA100
    "A2": (8, 6),          # GA107
    "A10": (8, 6),         # GA102
    "A10G": (8, 6),        # GA102 (AWS G5)
    "A16": (8, 6),         # GA107 x4
    "A40": (8, 6),         # GA102
    "A4000": (8, 6),       # RTX A4000, GA104
    "A4500": (8, 6),       # RTX A4500, GA102
    "A5000": (8, 6),       # RTX A5000, GA102
    "A6000": (8, 6),       # RTX A6000, GA102
    "RTX3060": (8, 6), "RTX3070": (8, 6), "RTX3080": (8, 6), "RTX3090": (8, 6),  # GeForce RTX 30, GA10x
    # Ada Lovelace
    "L4": (8, 9),          # AD104
    "L40": (8, 9),         # AD102
    "L40S": (8, 9),        # AD102
    "RTX2000Ada": (8, 9), "RTX4000Ada": (8, 9), "RTX5000Ada": (8, 9), "RTX6000Ada": (8, 9),  # RTX Ada generation
    "RTX4070": (8, 9), "RTX4080": (8, 9), "RTX4090": (8, 9),                                  # GeForce RTX 40, AD10x
    # Hopper
    "H100": (9, 0),        # GH100
    "H100NVL": (9, 0),     # GH100
    "H200": (9, 0),        # GH100 + HBM3e
    # Blackwell
    "B200": (10, 0),       # GB100
}
_CC_LOWER = {name.lower(): cc for name, cc in COMPUTE_CAPABILITY.items()}


def compute_capability_of(gpu_name: Optional[str]):
    """Case-insensitive; ``None`` for a name the table does not hold."""
    return None if gpu_name is None else _CC_LOWER.get(gpu_name.lower())


def _within(value, low, high) -> bool:
    """Closed interval test; a ``None`` bound is unbounded (App. A.1 ``between``)."""
    if low is not None and value < low:
        return False
    if high is not None and value > high:
        return False
    return True


def matches(item: CatalogItem, q: QueryFilter) -> bool:
    """Conjunction of the twelve range/set predicates of SURVEY.md App. A.1."""
    # 1. provider allow-list (case-insensitive)
    if q.provider is not None:
        allowed = {p.lower() for p in q.provider}
        if item.provider is None or item.provider.lower() not in allowed:
            return False
    # 2. price, 3. spot, 4. cpu, 5. memory (GiB, f64)
    if not _within(item.price, q.min_price, q.max_price):
        return False
    if q.spot is not None and item.spot != q.spot:
        return False
    if not _within(item.cpu, q.min_cpu, q.max_cpu):
        return False
    if not _within(item.memory, q.min_memory, q.max_memory):
        return False
    # 6. gpu count
    if not _within(item.gpu_count, q.min_gpu_count, q.max_gpu_count):
        return False
    # 7. vendor: a filter needs the item to carry the same vendor
    if q.gpu_vendor is not None:
        if item.gpu_vendor is None or item.gpu_vendor != q.gpu_vendor:
            return False
    # 8. gpu name: case-insensitive membership
    if q.gpu_name is not None:
        if item.gpu_name is None:
            return False
        if item.gpu_name.lower() not in {n.lower() for n in q.gpu_name}:
            return False
    # 9. per-GPU memory: any bound -- a maximum alone too -- needs at least one GPU.  Reading
    # assumed (unpinned): upstream returns False for gpu_count == 0 before comparing; the other
    # reading (compare gpu_memory = 0 / None with the bounds) would let a CPU-only item pass
    # "memory ..16GB, count 0..".  Call-site evidence: offers.py:128-136 only sets these bounds
    # inside `if res.gpu:`, i.e. for jobs that speak about GPUs at all.
    if q.min_gpu_memory is not None or q.max_gpu_memory is not None:
        if item.gpu_count <= 0 or item.gpu_memory is None:
            return False
        if not _within(item.gpu_memory, q.min_gpu_memory, q.max_gpu_memory):
            return False
    # 10. compute capability: an unknown capability -- which includes an item without any GPU
    # (gpu_name None) -- fails a bound.  Reading assumed (unpinned): the test is not skipped for
    # gpu_count == 0; dstack only sets the bound together with a GPU spec (offers.py:143-144).
    if q.min_compute_capability is not None or q.max_compute_capability is not None:
        cc = compute_capability_of(item.gpu_name)
        if cc is None:
            return False
        if not _within(tuple(cc),
                       None if q.min_compute_capability is None else tuple(q.min_compute_capability),
                       None if q.max_compute_capability is None else tuple(q.max_compute_capability)):
            return False
    # 11. total GPU memory = count * per-GPU memory (0 without GPUs)
    if q.min_total_gpu_memory is not None or q.max_total_gpu_memory is not None:
        total = item.gpu_count * item.gpu_memory if item.gpu_count > 0 and item.gpu_memory is not None else 0
        if not _within(total, q.min_total_gpu_memory, q.max_total_gpu_memory):
            return False
    # 12. disk: an item without a disk size is unconstrained
    if item.disk_size is not None:
        if not _within(item.disk_size, q.min_disk_size, q.max_disk_size):
            return False
    return True


class Catalog:
    """The slice of ``gpuhunt.Catalog`` the reference drives: per-provider item lists that are
    already price-sorted (``contributing/GPUHUNT.md:33-35``), filtered with :func:`matches`
    and merged by lowest price."""

    def __init__(self, balance_resources: bool = True, auto_reload: bool = False):
        self.balance_resources = balance_resources
        self._items: Dict[str, List[CatalogItem]] = {}

    def add_items(self, provider: str, items: Iterable[CatalogItem]) -> None:
        bucket = self._items.setdefault(provider, [])
        for it in items:
            if it.provider is None:
                it.provider = provider
            bucket.append(it)

    def query(self, *, provider: Optional[Sequence[str]] = None, **bounds) -> List[CatalogItem]:
        q = QueryFilter(provider=list(provider) if provider is not None else None, **bounds)
        wanted = None if provider is None else {p.lower() for p in provider}
        per_provider = []
        for name, items in self._items.items():
            if wanted is not None and name.lower() not in wanted:
                continue
            per_provider.append([it for it in items if matches(it, q)])
        return list(heapq.merge(*per_provider, key=lambda it: it.price))
