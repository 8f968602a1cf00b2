"""Drop-in mirror of ``dstack._internal.core.backends.base.offers`` with the predicate loop on the GPU.

Same names, arguments, ordering and object-identity behaviour as the reference module
(``core/backends/base/offers.py``); what changes is where the per-offer predicate runs:

* ``match_requirements(offers, requirements)`` (ref ``:149-158``) encodes the offers once and
  evaluates the job on the GPU (``dsm_match_bitmap``), returning the same objects in input order.
* ``GpuCatalog.query(**asdict(q))`` stands in for ``gpuhunt.Catalog.query`` (ref ``:33-34``):
  the catalog rows live in HBM, the filter runs on the GPU, matching items come back in the
  price-merged order gpuhunt returns them in.
* ``requirements_to_query_filter`` / ``offer_to_catalog_item`` / ``catalog_item_to_offer`` /
  ``choose_disk_size_mib`` are per-job or per-result conversions and stay on the host.

There is no CPU fallback for the predicate: without ``libdsmatch.so`` these functions raise.
"""

from __future__ import annotations

from dataclasses import asdict
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from . import models as M
from ._gpuhunt import CatalogItem, QueryFilter
from .encode import encode_table
from .engine import DeviceTable, Matcher, default_matcher
from .ingest import JobFilter, columns_from_catalog_items, columns_from_offers, job_columns

# ---------------------------------------------------------------------------- conversions (host)


def requirements_to_query_filter(req: Optional[Any]) -> QueryFilter:
    """``Requirements`` -> ``QueryFilter`` (ref ``:112-146``).  ``resources.gpu is None`` sets no
    GPU bound at all, so CPU-only jobs also match GPU offers."""
    q = QueryFilter()
    if req is None:
        return q
    q.max_price = req.max_price
    q.spot = req.spot
    spec = req.resources
    for rng, lo, hi in ((spec.cpu, "min_cpu", "max_cpu"), (spec.memory, "min_memory", "max_memory")):
        if rng:
            setattr(q, lo, rng.min)
            setattr(q, hi, rng.max)
    if spec.disk:
        q.min_disk_size, q.max_disk_size = spec.disk.size.min, spec.disk.size.max
    g = spec.gpu
    if g:
        q.gpu_vendor = g.vendor
        q.gpu_name = g.name
        for rng, lo, hi in ((g.memory, "min_gpu_memory", "max_gpu_memory"),
                            (g.count, "min_gpu_count", "max_gpu_count"),
                            (g.total_memory, "min_total_gpu_memory", "max_total_gpu_memory")):
            if rng:
                setattr(q, lo, rng.min)
                setattr(q, hi, rng.max)
        if g.compute_capability:
            q.min_compute_capability = g.compute_capability
    return q


def offer_to_catalog_item(offer: Any) -> CatalogItem:
    """``InstanceOffer`` -> ``CatalogItem`` (ref ``:86-109``).  Kept for API compatibility; the
    matcher does not call it per pair -- :func:`dstack_b200.ingest.columns_from_offers`
    performs the same conversion once per offer."""
    res = offer.instance.resources
    gpus = res.gpus
    head = gpus[0] if gpus else None
    return CatalogItem(
        provider=offer.backend.value, instance_name=offer.instance.name, location=offer.region,
        price=offer.price, cpu=res.cpus, memory=res.memory_mib / 1024, gpu_count=len(gpus),
        gpu_vendor=head.vendor if head else None, gpu_name=head.name if head else None,
        gpu_memory=head.memory_mib / 1024 if head else None, spot=res.spot,
        disk_size=res.disk.size_mib / 1024,
    )


def choose_disk_size_mib(catalog_item_disk_size_gib: Optional[float],
                         requirements_disk_size: Optional[Any],
                         configurable_disk_size: Any) -> Optional[int]:
    """Disk of a converted offer (ref ``:161-175``): the item's own size when it has one, else the
    smallest size both the job and the backend allow; ``None`` drops the offer."""
    if catalog_item_disk_size_gib:
        return round(catalog_item_disk_size_gib * 1024)
    asked = requirements_disk_size or M.DEFAULT_DISK.size
    feasible = asked.intersect(configurable_disk_size)
    return None if feasible is None else round(feasible.min * 1024)


def catalog_item_to_offer(backend: Any, item: Any, requirements: Optional[Any],
                          configurable_disk_size: Any) -> Optional[M.InstanceOffer]:
    """ref ``:46-83``."""
    asked_disk = requirements.resources.disk.size if requirements and requirements.resources.disk else None
    disk_mib = choose_disk_size_mib(item.disk_size, asked_disk, configurable_disk_size)
    if disk_mib is None:
        return None
    gpus: List[M.Gpu] = []
    if item.gpu_count > 0:
        gpus = [M.Gpu(vendor=item.gpu_vendor, name=item.gpu_name,
                      memory_mib=round(item.gpu_memory * 1024))] * item.gpu_count
    res = M.Resources(cpus=item.cpu, memory_mib=round(item.memory * 1024), gpus=gpus, spot=item.spot,
                      disk=M.Disk(size_mib=disk_mib))
    res.description = res.pretty_format()
    return M.InstanceOffer(backend=backend, region=item.location, price=item.price,
                           instance=M.InstanceType(name=item.instance_name, resources=res))


# ---------------------------------------------------------------------------- GPU-backed


def match_requirements(offers: List[Any], requirements: Optional[Any], *,
                       matcher: Optional[Matcher] = None) -> List[Any]:
    """Offers that satisfy ``requirements`` -- the same objects, in input order (ref ``:149-158``)."""
    if not offers:
        return []
    m = matcher or default_matcher()
    table = encode_table(columns_from_offers(offers), check_sorted_backends=False)
    dt = m.upload(table)
    try:
        jobs = job_columns(table, [JobFilter(query=requirements_to_query_filter(requirements))])
        keep = m.match_all(dt, jobs, order="source")[0]
    finally:
        dt.close()
    return [offers[i] for i in keep]


class GpuCatalog:
    """Device-resident stand-in for ``gpuhunt.Catalog`` as the reference uses it.

    ``add_items(provider, items)`` takes one provider's ``CatalogItem`` rows, price-sorted as
    gpuhunt guarantees (``contributing/GPUHUNT.md:33-35``); ``query(**asdict(q))`` filters on the
    GPU and returns the matching items merged by lowest price across providers, ties going to the
    provider added first -- what ``heapq.merge`` over the per-provider lists yields."""

    def __init__(self, balance_resources: bool = True, auto_reload: bool = False,
                 matcher: Optional[Matcher] = None):
        self.balance_resources = balance_resources
        self._matcher = matcher
        self._providers: List[str] = []
        self._items: List[Any] = []
        self._owner: List[int] = []
        self._dt: Optional[DeviceTable] = None

    def add_items(self, provider: str, items: Sequence[Any]) -> None:
        if provider in self._providers:
            raise ValueError(f"provider {provider!r} already added")
        pid = len(self._providers)
        self._providers.append(provider)
        for it in items:
            if it.provider is None:
                it.provider = provider
            self._items.append(it)
            self._owner.append(pid)
        self._invalidate()

    def add_csv(self, provider: str, text: str) -> None:
        """One provider's CSV from gpuhunt's ``catalog.zip`` (see :mod:`dstack_b200.catalog_csv`)."""
        from .catalog_csv import catalog_items_from_csv

        self.add_items(provider, catalog_items_from_csv(text, provider))

    def _invalidate(self) -> None:
        if self._dt is not None:
            self._dt.close()
            self._dt = None

    def _table(self) -> DeviceTable:
        if self._dt is None:
            m = self._matcher or default_matcher()
            cols = columns_from_catalog_items(self._items)
            owner = np.asarray(self._owner, dtype=np.int8)
            cols.src_backend = owner  # provider position: breaks price ties, selects providers
            cols.backend = owner.copy()
            if len(self._providers) > 16:
                raise ValueError("at most 16 providers per catalog")
            self._dt = m.upload(encode_table(cols))
            self._m = m
        return self._dt

    def query(self, *, provider: Optional[Sequence[str]] = None, **bounds: Any) -> List[Any]:
        if not self._items:
            return []
        dt = self._table()
        q = QueryFilter(provider=None, **bounds)
        f = JobFilter(query=q)
        jobs = job_columns(dt.table, [f])
        if provider is not None:  # case-insensitive allow-list (App. A.1 item 1)
            wanted = {p.lower() for p in provider}
            mask = 0
            for pid, name in enumerate(self._providers):
                if name.lower() in wanted:
                    mask |= 1 << pid
            jobs.src_backend_mask[:] = mask
            jobs.off_backend_mask[:] = mask
        rows = self._m.match_all(dt, jobs, order="rank")[0]
        return [self._items[i] for i in rows]


_DEFAULT_CONFIGURABLE_DISK = M.Range[M.Memory](min=M.Memory.parse("1GB"), max=None)


def get_catalog_offers(backend: Any, locations: Optional[List[str]] = None,
                       requirements: Optional[Any] = None,
                       configurable_disk_size: Any = _DEFAULT_CONFIGURABLE_DISK,
                       extra_filter: Optional[Callable[[Any], bool]] = None,
                       catalog: Optional[Any] = None) -> List[M.InstanceOffer]:
    """ref ``:18-43``.  ``catalog`` must be given (a :class:`GpuCatalog`): there is no network to
    fetch gpuhunt's default catalog from, and no CPU catalog to fall back to."""
    if catalog is None:
        raise ValueError("get_catalog_offers needs a GpuCatalog (no default catalog offline)")
    provider = "lambdalabs" if backend == M.BackendType.LAMBDA else backend.value
    q = requirements_to_query_filter(requirements)
    q.provider = [provider]
    result = []
    for item in catalog.query(**asdict(q)):
        if locations is not None and item.location not in locations:
            continue
        offer = catalog_item_to_offer(backend, item, requirements, configurable_disk_size)
        if offer is None or (extra_filter is not None and not extra_filter(offer)):
            continue
        result.append(offer)
    return result
