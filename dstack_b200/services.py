"""Mirror of ``dstack._internal.server.services.offers`` (+ ``backends.get_instance_offers``)
and the batched, device-resident form of the same path.

Two layers:

1. Reference-signature functions -- ``get_offers_by_requirements``, ``get_instance_offers``,
   ``is_divisible_into_blocks``, ``generate_shared_offer`` -- with the reference's semantics
   (``server/services/offers.py:24-162``, ``server/services/backends/__init__.py:417-451``).
   They ask each backend for offers exactly as the reference does; a backend whose
   ``Compute.get_offers`` is built on :class:`dstack_b200.offers.GpuCatalog` or
   :func:`dstack_b200.offers.match_requirements` runs its predicate loop on the GPU.

2. :class:`OfferCatalog` -- every backend's offers resident in HBM in result order; a whole
   batch of pending jobs is matched and ranked in one call (``match_topk`` = the reference's
   ``offers[:k]``, ``len(offers)`` and ``max(price)`` per job; ``offers_for`` = the full list).
   This is what the scheduler loop (``process_submitted_jobs``) would call once per tick.
"""

from __future__ import annotations

import asyncio
import heapq
import logging
from dataclasses import dataclass
from typing import Any, List, Literal, Optional, Sequence, Tuple, Union

import numpy as np

from . import models as M
from ._gpuhunt import AcceleratorVendor
from .encode import EncodedTable, JobColumns, encode_table
from .engine import DeviceTable, Matcher, TopK, default_matcher
from .ingest import JobFilter, columns_from_offers, job_columns
from .offers import requirements_to_query_filter

logger = logging.getLogger(__name__)


class BackendError(Exception):
    """Stand-in for ``dstack._internal.core.errors.BackendError``: a backend that raises it is
    skipped with a warning (``backends/__init__.py:427-433``)."""


async def get_project_backends(project: Any) -> List[Any]:
    """Hook the reference resolves through ``backends_services.get_project_backends``; tests
    patch it the same way the reference's tests do.  Default: ``project.backends``."""
    return list(project.backends)


# ---------------------------------------------------------------------------- derivations


def derive_backend_types_and_regions(profile: Any, *, multinode: bool = False,
                                     master_job_provisioning_data: Optional[Any] = None,
                                     volumes: Optional[List[List[Any]]] = None,
                                     privileged: bool = False, instance_mounts: bool = False):
    """Backend / region restrictions a job implies (``server/services/offers.py:46-79``)."""
    backend_types, regions = profile.backends, profile.regions
    if volumes:
        first_mount = volumes[0]
        backend_types = [v.configuration.backend for v in first_mount]
        regions = [v.configuration.region for v in first_mount]

    def restrict(current, capable):
        return [b for b in (current or capable) if b in capable]

    if multinode:
        backend_types = restrict(backend_types, M.BACKENDS_WITH_MULTINODE_SUPPORT)
    if privileged or instance_mounts:
        backend_types = restrict(backend_types, M.BACKENDS_WITH_CREATE_INSTANCE_SUPPORT)
    if profile.reservation is not None:
        backend_types = restrict(backend_types, M.BACKENDS_WITH_RESERVATION_SUPPORT)
    jpd = master_job_provisioning_data
    if jpd is not None:  # all nodes of a cluster go to the master's backend and region
        base = jpd.get_base_backend()
        backend_types = [b for b in (backend_types or [base]) if b == base]
        regions = [r for r in (regions or [jpd.region]) if r == jpd.region]
    return backend_types, regions


def is_divisible_into_blocks(cpu_count: int, gpu_count: int,
                             blocks: Union[int, Literal["auto"]]) -> Tuple[bool, int]:
    """``(True, n)`` if an instance splits into ``n`` equal blocks, else ``(False, 0)``
    (``server/services/offers.py:121-136``)."""
    n = blocks
    if n == "auto":
        n = min(cpu_count, gpu_count) if gpu_count else cpu_count
    if n < 1 or cpu_count % n != 0 or gpu_count % n != 0:
        return False, 0
    return True, n


def generate_shared_offer(offer: Any, blocks: int, total_blocks: int) -> M.InstanceOfferWithAvailability:
    """The ``blocks``-of-``total_blocks`` slice of an instance (``server/services/offers.py:139-162``)."""
    whole = offer.instance.resources
    share = M.Resources(
        cpus=whole.cpus // total_blocks * blocks,
        memory_mib=whole.memory_mib // total_blocks * blocks,
        gpus=whole.gpus[: len(whole.gpus) // total_blocks * blocks],
        spot=whole.spot, disk=whole.disk, description=whole.description,
    )
    return M.InstanceOfferWithAvailability(
        backend=offer.backend, region=offer.region, price=offer.price, availability=offer.availability,
        instance=M.InstanceType(name=offer.instance.name, resources=share),
        blocks=blocks, total_blocks=total_blocks,
    )


def _blocks_of_offer(offer: Any, blocks: Union[int, Literal["auto"]]) -> Tuple[bool, int]:
    res = offer.instance.resources
    gpu_count = len(res.gpus)
    if gpu_count > 0 and res.gpus[0].vendor == AcceleratorVendor.GOOGLE:
        gpu_count = 1  # TPUs cannot be shared
    return is_divisible_into_blocks(res.cpus, gpu_count, blocks)


# ---------------------------------------------------------------------------- reference-signature layer


async def get_instance_offers(backends: List[Any], requirements: Any,
                              exclude_not_available: bool = False) -> List[Tuple[Any, Any]]:
    """Offers of every backend merged by price, not-available ones last
    (``server/services/backends/__init__.py:417-451``)."""
    logger.info("Requesting instance offers from backends: %s", [b.TYPE.value for b in backends])
    loop = asyncio.get_running_loop()
    calls = [loop.run_in_executor(None, b.compute().get_offers_cached, requirements) for b in backends]
    per_backend = []
    for backend, result in zip(backends, await asyncio.gather(*calls, return_exceptions=True)):
        if isinstance(result, BackendError):
            logger.warning("Failed to get offers from backend %s: %r", backend.TYPE, result)
            continue
        if isinstance(result, BaseException):
            logger.error("Got exception when requesting offers from backend %s", backend.TYPE, exc_info=result)
            continue
        per_backend.append([(backend, o) for o in result
                            if not exclude_not_available or o.availability.is_available()])
    merged = heapq.merge(*per_backend, key=lambda bo: bo[1].price)
    return sorted(merged, key=lambda bo: not bo[1].availability.is_available())


async def get_offers_by_requirements(
    project: Any,
    profile: Any,
    requirements: Any,
    exclude_not_available: bool = False,
    multinode: bool = False,
    master_job_provisioning_data: Optional[Any] = None,
    volumes: Optional[List[List[Any]]] = None,
    privileged: bool = False,
    instance_mounts: bool = False,
    blocks: Union[int, Literal["auto"]] = 1,
) -> List[Tuple[Any, Any]]:
    """``server/services/offers.py:24-118`` (same mutations of ``profile.backends`` and
    ``offer.total_blocks``)."""
    backends = await get_project_backends(project=project)
    if profile.backends is not None and len(profile.backends) == 1 and M.BackendType.DSTACK in profile.backends:
        profile.backends = None
    backend_types, regions = derive_backend_types_and_regions(
        profile, multinode=multinode, master_job_provisioning_data=master_job_provisioning_data,
        volumes=volumes, privileged=privileged, instance_mounts=instance_mounts)
    if backend_types is not None:
        backends = [b for b in backends if b.TYPE in backend_types or b.TYPE == M.BackendType.DSTACK]
    offers = await get_instance_offers(backends=backends, requirements=requirements,
                                       exclude_not_available=exclude_not_available)
    if backend_types is not None:
        offers = [(b, o) for b, o in offers if o.backend in backend_types]
    if regions is not None:
        offers = [(b, o) for b, o in offers if o.region in regions]
    if profile.instance_types is not None:
        offers = [(b, o) for b, o in offers if o.instance.name in profile.instance_types]
    if blocks == 1:
        return offers
    shareable = []
    for backend, offer in offers:
        ok, n = _blocks_of_offer(offer, blocks)
        if ok:
            offer.total_blocks = n
            shareable.append((backend, offer))
    return shareable


# ---------------------------------------------------------------------------- batched, device-resident layer


class _ProfileView:
    """The three profile fields ``derive_backend_types_and_regions`` reads, without touching the profile."""

    __slots__ = ("backends", "regions", "reservation")

    def __init__(self, backends: Any, regions: Any, reservation: Any):
        self.backends, self.regions, self.reservation = backends, regions, reservation


def job_filter(profile: Any, requirements: Any, *, backend_types_of_project: Optional[Sequence[Any]] = None,
               exclude_not_available: bool = False, multinode: bool = False,
               master_job_provisioning_data: Optional[Any] = None, volumes: Optional[List[List[Any]]] = None,
               privileged: bool = False, instance_mounts: bool = False,
               blocks: Union[int, Literal["auto"]] = 1) -> JobFilter:
    """Fold one ``get_offers_by_requirements`` call into a :class:`JobFilter`."""
    backends = profile.backends
    if backends is not None and len(backends) == 1 and M.BackendType.DSTACK in backends:
        backends = None

    p = _ProfileView(backends, profile.regions, profile.reservation)
    backend_types, regions = derive_backend_types_and_regions(
        p, multinode=multinode, master_job_provisioning_data=master_job_provisioning_data, volumes=volumes,
        privileged=privileged, instance_mounts=instance_mounts)
    # blocks == 1 filters nothing; "auto" needs min(cpu, gpu) to divide both (a per-row flag);
    # an explicit count b needs cpu % b == 0 and gpu % b == 0; b < 1 is never divisible
    explicit = 0
    if blocks != "auto" and blocks != 1:
        explicit = int(blocks) if int(blocks) > 1 else -1
    return JobFilter(query=requirements_to_query_filter(requirements), backend_types=backend_types,
                     keep_dstack_backend=True, regions=regions, instance_types=profile.instance_types,
                     exclude_not_available=exclude_not_available, require_auto_divisible=(blocks == "auto"),
                     blocks=explicit)


class OfferCatalog:
    """All offers of a project's backends, resident on one GPU in result order.

    ``offers_by_backend``: ``[(backend, [InstanceOfferWithAvailability, ...]), ...]`` in
    project-backend order, each list price-sorted -- i.e. what every backend's
    ``get_offers_cached`` returns for an unconstrained request.  The catalog lives until
    ``close()``; rebuild it when the backends' 30 s offer cache expires
    (``core/backends/base/compute.py:47-48``)."""

    def __init__(self, offers_by_backend: Sequence[Tuple[Any, Sequence[Any]]], *,
                 matcher: Optional[Matcher] = None):
        self.matcher = matcher or default_matcher()
        self.backends: List[Any] = []
        self.offers: List[Any] = []
        self._materialize = None
        src = []
        for backend, offers in offers_by_backend:
            for o in offers:
                self.backends.append(backend)
                self.offers.append(o)
                src.append(backend.TYPE)
        self.table: EncodedTable = encode_table(columns_from_offers(self.offers, src))
        self.device_table: DeviceTable = self.matcher.upload(self.table)

    @classmethod
    def from_columns(cls, columns: Any, materialize, *, matcher: Optional[Matcher] = None) -> "OfferCatalog":
        """A catalog that never existed as Python objects: ``columns`` is an
        :class:`~dstack_b200.encode.OfferColumns` in source order (per-backend lists concatenated in
        project-backend order, each price-sorted) and ``materialize(i) -> (backend, offer)`` builds
        the objects of source position ``i`` on demand -- only the few rows a job's top-k names are
        ever turned into objects (tens of millions of offers stay columnar)."""
        self = cls.__new__(cls)
        self.matcher = matcher or default_matcher()
        self.backends, self.offers = None, None
        self._materialize = materialize
        self.table = encode_table(columns)
        self.device_table = self.matcher.upload(self.table)
        return self

    @classmethod
    def from_catalog_items(cls, items_by_backend: Sequence[Tuple[Any, Sequence[Any], Any]], *,
                           matcher: Optional[Matcher] = None) -> "OfferCatalog":
        """A catalog of ``gpuhunt.CatalogItem`` rows: ``[(backend, items, configurable_disk_size),
        ...]`` in project-backend order, each list price-sorted -- what the backends'
        ``get_catalog_offers`` reads BEFORE a job's requirements are applied.

        This is the form to use for backends with configurable disks (AWS, GCP, Azure ...).  In the
        reference an item without a disk size becomes an offer whose disk is sized to the job
        (``catalog_item_to_offer`` / ``choose_disk_size_mib``, ``core/backends/base/offers.py:46-83,
        161-175``), so whether it matches depends on the job's disk range overlapping the backend's
        configurable range, not on a fixed size.  Here that overlap test is part of the device
        predicate, and a result row is turned into an offer -- with the disk the job gets -- by
        :meth:`pairs` / :meth:`offers_for`.  (A catalog built from ready-made offers,
        ``OfferCatalog(offers_by_backend)``, takes every offer's disk as a fixed size.)"""
        from .ingest import columns_from_backend_items

        self = cls.__new__(cls)
        self.matcher = matcher or default_matcher()
        self.backends, self.offers, self._materialize = None, None, None
        self._items = [(b, it, conf) for b, items, conf in items_by_backend for it in items]
        self.table = encode_table(columns_from_backend_items([(b.TYPE, items, conf) for b, items, conf in items_by_backend]))
        self.device_table = self.matcher.upload(self.table)
        return self

    def pair_at(self, i: int, disk_range: Optional[Any] = None) -> Tuple[Any, Any]:
        """``(backend, offer)`` of source position ``i``; for a catalog of items the offer is made
        here, its disk sized for a job whose disk range is ``disk_range`` (None: DEFAULT_DISK)."""
        if getattr(self, "_items", None) is not None:
            from .offers import choose_disk_size_mib

            backend, item, conf = self._items[int(i)]
            disk_mib = choose_disk_size_mib(item.disk_size, disk_range, conf)
            if disk_mib is None:
                raise RuntimeError("a matched catalog item has no feasible disk size: the device predicate "
                                   "and choose_disk_size_mib disagree")
            gpus = []
            if item.gpu_count > 0:
                gpus = [M.Gpu(vendor=item.gpu_vendor, name=item.gpu_name,
                              memory_mib=round(item.gpu_memory * 1024))] * item.gpu_count
            res = M.Resources(cpus=item.cpu, memory_mib=round(item.memory * 1024), gpus=gpus, spot=item.spot,
                              disk=M.Disk(size_mib=disk_mib))
            res.description = res.pretty_format()
            return backend, M.InstanceOfferWithAvailability(
                backend=backend.TYPE, region=item.location, price=item.price,
                instance=M.InstanceType(name=item.instance_name, resources=res),
                availability=M.InstanceAvailability.UNKNOWN)
        if self._materialize is not None:
            # building an offer object costs ~50 us (pydantic); a pass asks for the same cheap rows
            # for many jobs, so the objects made so far are kept (bounded)
            cache = self.__dict__.setdefault("_made", {})
            pair = cache.get(int(i))
            if pair is None:
                if len(cache) >= 200_000:
                    cache.clear()
                pair = cache[int(i)] = self._materialize(int(i))
            return pair
        return self.backends[i], self.offers[i]

    @staticmethod
    def _disk_range_of(f: JobFilter) -> Optional[Any]:
        q = f.query
        if q.min_disk_size is None and q.max_disk_size is None:
            return None
        return M.Range[M.Memory](min=q.min_disk_size, max=q.max_disk_size)

    def close(self) -> None:
        self.device_table.close()

    def jobs(self, filters: Sequence[JobFilter]) -> JobColumns:
        return job_columns(self.table, filters)

    def match_topk(self, filters: Sequence[JobFilter], k: int) -> TopK:
        return self.matcher.match_topk(self.device_table, self.jobs(filters), k)

    def offers_for(self, filters: Sequence[JobFilter]) -> List[List[Tuple[Any, Any]]]:
        """The complete ``get_offers_by_requirements`` result of every job."""
        hits = self.matcher.match_all(self.device_table, self.jobs(filters), order="rank")
        out = []
        for f, rows in zip(filters, hits):
            disk = self._disk_range_of(f)
            pairs = [self.pair_at(i, disk) for i in rows]
            if f.require_auto_divisible or f.blocks > 1:  # offers.py:116 writes the count back
                for _, o in pairs:
                    o.total_blocks = _blocks_of_offer(o, "auto" if f.require_auto_divisible else f.blocks)[1]
            out.append(pairs)
        return out

    def pairs(self, index_row: np.ndarray, job: Optional[JobFilter] = None) -> List[Tuple[Any, Any]]:
        """``(backend, offer)`` of a ``match_topk`` row; ``job``: the filter the row answers (a
        catalog of items sizes the offers' disks for it)."""
        disk = None if job is None else self._disk_range_of(job)
        return [self.pair_at(i, disk) for i in index_row if i >= 0]


class OfferCatalogCache:
    """An :class:`OfferCatalog` with the life time of the reference's offers cache.

    The reference keeps each backend's offers for 30 s (``TTLCache(maxsize=10, ttl=30)`` around
    ``get_offers_cached``, ``core/backends/base/compute.py:46-48,201-209``); the resident table
    follows the same clock: ``get()`` returns the current catalog and rebuilds it from
    ``fetch()`` -- ``[(backend, price-sorted offers), ...]`` -- on the first use after ``ttl``
    seconds.  The previous table is freed once the new one is up.  Calls may come from several
    threads (the reference fetches offers from a thread pool): one of them rebuilds, outside the
    lock, the others keep matching against the previous table until the swap."""

    def __init__(self, fetch, *, ttl: float = 30.0, matcher: Optional[Matcher] = None,
                 clock=None, build=None):
        import threading
        import time

        self._fetch = fetch
        self.ttl = ttl
        self._matcher = matcher
        self._clock = clock or time.monotonic
        self._build = build or (lambda offers_by_backend: OfferCatalog(offers_by_backend, matcher=self._matcher))
        self._lock = threading.Lock()
        self._ready = threading.Condition(self._lock)
        self._building = False
        self._catalog: Optional[Any] = None
        self._born = 0.0
        self.rebuilds = 0

    def get(self) -> Any:
        """The current catalog.  When it has expired, the caller that notices builds the replacement
        -- fetch, encode, upload -- WITHOUT holding the lock, and swaps it in when it is up; matcher
        threads that call meanwhile keep getting the previous catalog (only the very first build
        makes them wait).  The replaced table is freed after the swap."""
        with self._lock:
            now = self._clock()
            fresh_enough = self._catalog is not None and now - self._born < self.ttl
            if fresh_enough or (self._catalog is not None and self._building):
                return self._catalog
            if self._catalog is None and self._building:
                while self._building:  # first build in progress on another thread
                    self._ready.wait()
                return self._catalog
            self._building = True
        try:
            fresh = self._build(self._fetch())
        except BaseException:
            with self._lock:
                self._building = False
                self._ready.notify_all()
            raise
        with self._lock:
            old, self._catalog, self._born = self._catalog, fresh, self._clock()
            self._building = False
            self.rebuilds += 1
            self._ready.notify_all()
        if old is not None:
            old.close()
        return fresh

    def invalidate(self) -> None:
        """Force a rebuild on the next ``get()`` (e.g. after a backend was added or removed)."""
        with self._lock:
            self._born = -float("inf")

    def close(self) -> None:
        with self._lock:
            if self._catalog is not None:
                self._catalog.close()
                self._catalog = None
