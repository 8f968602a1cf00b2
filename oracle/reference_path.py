"""Object-level CPU restatement of the reference's offer-matching path.

TEST INFRASTRUCTURE (see oracle/__init__): the checker for the CUDA path and the timed CPU
baseline.  Every function names the reference lines it follows (paths relative to
``/root/reference/src/dstack/_internal``).  Inputs are duck-typed: the objects of
``dstack_b200.models`` and the reference's own pydantic objects both work, and pool rows
only need the attributes of ``InstanceModel`` (``server/models.py:476-561``) that are read.

The cost structure is kept on purpose (a fresh ``CatalogItem`` per (job, offer), a JSON
parse per (job, pool row)): this is what ``bench.py --impl reference`` times.
"""

from __future__ import annotations

import heapq
from typing import Any, Callable, Iterable, List, Optional, Sequence, Tuple

from dstack_b200 import models as M
from dstack_b200._gpuhunt import AcceleratorVendor, CatalogItem, QueryFilter

from .gpuhunt_restated import matches

# --------------------------------------------------------------------------- a1 / a2 / a4


def requirements_to_query_filter(req) -> QueryFilter:
    """core/backends/base/offers.py:112-146."""
    q = QueryFilter()
    if req is None:
        return q
    q.max_price, q.spot = req.max_price, req.spot
    res = req.resources
    if res.cpu:
        q.min_cpu, q.max_cpu = res.cpu.min, res.cpu.max
    if res.memory:
        q.min_memory, q.max_memory = res.memory.min, res.memory.max
    if res.disk:
        q.min_disk_size, q.max_disk_size = res.disk.size.min, res.disk.size.max
    gpu = res.gpu
    if gpu:
        q.gpu_vendor, q.gpu_name = gpu.vendor, gpu.name
        if gpu.memory:
            q.min_gpu_memory, q.max_gpu_memory = gpu.memory.min, gpu.memory.max
        if gpu.count:
            q.min_gpu_count, q.max_gpu_count = gpu.count.min, gpu.count.max
        if gpu.total_memory:
            q.min_total_gpu_memory, q.max_total_gpu_memory = gpu.total_memory.min, gpu.total_memory.max
        if gpu.compute_capability:
            q.min_compute_capability = gpu.compute_capability
    return q


def offer_to_catalog_item(offer) -> CatalogItem:
    """core/backends/base/offers.py:86-109: MiB -> GiB by /1024; GPU facts from gpus[0]."""
    res = offer.instance.resources
    n = len(res.gpus)
    first = res.gpus[0] if n > 0 else None
    return CatalogItem(
        provider=offer.backend.value,
        instance_name=offer.instance.name,
        location=offer.region,
        price=offer.price,
        cpu=res.cpus,
        memory=res.memory_mib / 1024,
        gpu_count=n,
        gpu_vendor=None if first is None else first.vendor,
        gpu_name=None if first is None else first.name,
        gpu_memory=None if first is None else first.memory_mib / 1024,
        spot=res.spot,
        disk_size=res.disk.size_mib / 1024,
    )


def match_requirements(offers: Sequence[Any], requirements) -> List[Any]:
    """core/backends/base/offers.py:149-158: same objects, input order."""
    q = requirements_to_query_filter(requirements)
    return [o for o in offers if matches(offer_to_catalog_item(o), q)]


# --------------------------------------------------------------------------- a5 / a6


def choose_disk_size_mib(catalog_item_disk_size_gib, requirements_disk_size, configurable_disk_size):
    """core/backends/base/offers.py:161-175."""
    if catalog_item_disk_size_gib:
        gib = catalog_item_disk_size_gib
    else:
        wanted = requirements_disk_size or M.DEFAULT_DISK.size
        common = wanted.intersect(configurable_disk_size)
        if common is None:
            return None
        gib = common.min
    return round(gib * 1024)


def catalog_item_to_offer(backend, item, requirements, configurable_disk_size):
    """core/backends/base/offers.py:46-83."""
    gpus = []
    if item.gpu_count > 0:
        gpus = [M.Gpu(vendor=item.gpu_vendor, name=item.gpu_name,
                      memory_mib=round(item.gpu_memory * 1024))] * item.gpu_count
    wanted_disk = None
    if requirements and requirements.resources.disk:
        wanted_disk = requirements.resources.disk.size
    disk_mib = choose_disk_size_mib(item.disk_size, wanted_disk, configurable_disk_size)
    if disk_mib is None:
        return None
    res = M.Resources(cpus=item.cpu, memory_mib=round(item.memory * 1024), gpus=gpus,
                      spot=item.spot, disk=M.Disk(size_mib=disk_mib))
    res.description = res.pretty_format()
    return M.InstanceOffer(backend=backend, instance=M.InstanceType(name=item.instance_name, resources=res),
                           region=item.location, price=item.price)


_DEFAULT_CONFIGURABLE_DISK = M.Range[M.Memory](min=M.Memory.parse("1GB"), max=None)


def get_catalog_offers(backend, locations=None, requirements=None,
                       configurable_disk_size=_DEFAULT_CONFIGURABLE_DISK,
                       extra_filter: Optional[Callable[[Any], bool]] = None, catalog=None):
    """core/backends/base/offers.py:18-43 (``catalog`` is required here: no network)."""
    from dataclasses import asdict

    provider = "lambdalabs" if backend == M.BackendType.LAMBDA else backend.value
    q = requirements_to_query_filter(requirements)
    q.provider = [provider]
    out = []
    for item in catalog.query(**asdict(q)):
        if locations is not None and item.location not in locations:
            continue
        offer = catalog_item_to_offer(backend, item, requirements, configurable_disk_size)
        if offer is None:
            continue
        if extra_filter is not None and not extra_filter(offer):
            continue
        out.append(offer)
    return out


# --------------------------------------------------------------------------- a8 / a9 / a10


def get_instance_offers(backends: Sequence[Any], requirements, exclude_not_available: bool = False):
    """server/services/backends/__init__.py:417-451, synchronous.

    ``backends`` are objects with ``TYPE`` and ``compute().get_offers_cached(requirements)``.
    A backend that raises contributes nothing (``:427-440``)."""
    lists = []
    for backend in backends:
        try:
            result = backend.compute().get_offers_cached(requirements)
        except Exception:
            continue
        lists.append([(backend, o) for o in result
                      if not exclude_not_available or o.availability.is_available()])
    merged = heapq.merge(*lists, key=lambda pair: pair[1].price)
    return sorted(merged, key=lambda pair: not pair[1].availability.is_available())


def is_divisible_into_blocks(cpu_count: int, gpu_count: int, blocks) -> Tuple[bool, int]:
    """server/services/offers.py:121-136."""
    if blocks == "auto":
        blocks = cpu_count if gpu_count == 0 else min(cpu_count, gpu_count)
    if blocks < 1 or cpu_count % blocks or gpu_count % blocks:
        return False, 0
    return True, blocks


def generate_shared_offer(offer, blocks: int, total_blocks: int):
    """server/services/offers.py:139-162: a ``blocks/total_blocks`` slice of an instance."""
    full = offer.instance.resources
    part = M.Resources(
        cpus=full.cpus // total_blocks * blocks,
        memory_mib=full.memory_mib // total_blocks * blocks,
        gpus=full.gpus[: len(full.gpus) // total_blocks * blocks],
        spot=full.spot,
        disk=full.disk,
        description=full.description,
    )
    return M.InstanceOfferWithAvailability(
        backend=offer.backend,
        instance=M.InstanceType(name=offer.instance.name, resources=part),
        region=offer.region,
        price=offer.price,
        availability=offer.availability,
        blocks=blocks,
        total_blocks=total_blocks,
    )


def _derive_backend_and_region_filters(profile, volumes, multinode, master_jpd, *,
                                       privileged=False, instance_mounts=False, for_pool=False):
    """Shared derivation: server/services/offers.py:46-79 and server/services/pools.py:425-456.

    Returns ``(backend_types, regions, zones)``; each ``None`` means unconstrained."""
    backend_types, regions, zones = profile.backends, profile.regions, None
    if volumes:
        first = volumes[0]
        backend_types = [v.configuration.backend for v in first]
        regions = [v.configuration.region for v in first]
        if for_pool:
            vz = [v.provisioning_data.availability_zone for v in first if v.provisioning_data is not None]
            if vz:
                zones = vz
    if multinode:
        if not backend_types:
            backend_types = M.BACKENDS_WITH_MULTINODE_SUPPORT
        backend_types = [b for b in backend_types if b in M.BACKENDS_WITH_MULTINODE_SUPPORT]
    if not for_pool:
        if privileged or instance_mounts:
            if not backend_types:
                backend_types = M.BACKENDS_WITH_CREATE_INSTANCE_SUPPORT
            backend_types = [b for b in backend_types if b in M.BACKENDS_WITH_CREATE_INSTANCE_SUPPORT]
        if profile.reservation is not None:
            if not backend_types:
                backend_types = M.BACKENDS_WITH_RESERVATION_SUPPORT
            backend_types = [b for b in backend_types if b in M.BACKENDS_WITH_RESERVATION_SUPPORT]
    if master_jpd is not None:
        base = master_jpd.get_base_backend()
        if not backend_types:
            backend_types = [base]
        if not regions:
            regions = [master_jpd.region]
        backend_types = [b for b in backend_types if b == base]
        regions = [r for r in regions if r == master_jpd.region]
    return backend_types, regions, zones


def get_offers_by_requirements(backends: Sequence[Any], profile, requirements, exclude_not_available=False,
                               multinode=False, master_job_provisioning_data=None, volumes=None,
                               privileged=False, instance_mounts=False, blocks=1):
    """server/services/offers.py:24-118, synchronous; ``backends`` stands for
    ``get_project_backends(project)`` (project-backend order matters for ties)."""
    if (profile.backends is not None and len(profile.backends) == 1
            and M.BackendType.DSTACK in profile.backends):
        profile.backends = None  # :39-44, mutation is part of the contract
    backend_types, regions, _ = _derive_backend_and_region_filters(
        profile, volumes, multinode, master_job_provisioning_data,
        privileged=privileged, instance_mounts=instance_mounts)
    if backend_types is not None:
        backends = [b for b in backends if b.TYPE in backend_types or b.TYPE == M.BackendType.DSTACK]
    offers = get_instance_offers(backends, requirements, exclude_not_available)
    if backend_types is not None:
        offers = [(b, o) for b, o in offers if o.backend in backend_types]
    if regions is not None:
        offers = [(b, o) for b, o in offers if o.region in regions]
    if profile.instance_types is not None:
        offers = [(b, o) for b, o in offers if o.instance.name in profile.instance_types]
    if blocks == 1:
        return offers
    shareable = []
    for b, o in offers:
        res = o.instance.resources
        gpu_count = len(res.gpus)
        if gpu_count > 0 and res.gpus[0].vendor == AcceleratorVendor.GOOGLE:
            gpu_count = 1  # TPUs cannot be shared (:110-112)
        ok, n = is_divisible_into_blocks(res.cpus, gpu_count, blocks)
        if not ok:
            continue
        o.total_blocks = n
        shareable.append((b, o))
    return shareable


# --------------------------------------------------------------------------- a11 / a12 / a13


def _jpd_of(instance):
    raw = instance.job_provisioning_data
    return None if raw is None else M.JobProvisioningData.parse_raw(raw)


def filter_pool_instances(pool_instances: Sequence[Any], profile, *, requirements=None, status=None,
                          fleet_model=None, multinode=False, master_job_provisioning_data=None,
                          volumes=None, shared=False) -> List[Any]:
    """server/services/pools.py:409-503: admin predicates in the reference's order, then the
    resource predicate on the stored offer JSON; survivors in input order."""
    backend_types, regions, zones = _derive_backend_and_region_filters(
        profile, volumes, multinode, master_job_provisioning_data, for_pool=True)
    candidates = []
    for inst in pool_instances:
        if fleet_model is not None and inst.fleet_id != fleet_model.id:
            continue
        if inst.unreachable:
            continue
        if profile.instance_name is not None and inst.name != profile.instance_name:
            continue
        if status is not None and inst.status != status:
            continue
        if backend_types is not None and inst.backend not in backend_types:
            continue
        if regions is not None and inst.region not in regions:
            continue
        jpd = _jpd_of(inst)
        if jpd is not None:
            if profile.instance_types is not None and jpd.instance_type.name not in profile.instance_types:
                continue
            if jpd.availability_zone is not None and zones is not None and jpd.availability_zone not in zones:
                continue
        if inst.total_blocks is None:
            continue  # still provisioning (:484-486)
        if (inst.total_blocks > 1) != shared:
            continue
        candidates.append(inst)
    if requirements is None:
        return candidates
    q = requirements_to_query_filter(requirements)
    kept = []
    for inst in candidates:
        if inst.offer is None:
            continue
        offer = M.InstanceOffer.parse_raw(inst.offer)  # the per-(job,row) JSON parse, :499
        if matches(offer_to_catalog_item(offer), q):
            kept.append(inst)
    return kept


def get_shared_pool_instances_with_offers(pool_instances: Sequence[Any], profile, requirements, *,
                                          idle_only=False, fleet_model=None, volumes=None):
    """server/services/pools.py:506-544: smallest matching block count per shared instance."""
    q = requirements_to_query_filter(requirements)
    out = []
    for inst in filter_pool_instances(pool_instances, profile, fleet_model=fleet_model,
                                      multinode=False, volumes=volumes, shared=True):
        if idle_only and inst.status not in (M.InstanceStatus.IDLE, M.InstanceStatus.BUSY):
            continue
        if inst.offer is None:
            continue
        offer = M.InstanceOfferWithAvailability.parse_raw(inst.offer)
        total = inst.total_blocks
        idle = total - inst.busy_blocks
        for blocks in range(1, total + 1):
            part = generate_shared_offer(offer, blocks, total)
            if matches(offer_to_catalog_item(part), q):
                part.availability = (M.InstanceAvailability.IDLE if blocks <= idle
                                     else M.InstanceAvailability.BUSY)
                if part.availability == M.InstanceAvailability.IDLE or not idle_only:
                    out.append((inst, part))
                break
    return out


def rank_pool_candidates(nonshared: Sequence[Any], shared_with_offers: Sequence[Tuple[Any, Any]]):
    """server/background/tasks/process_submitted_jobs.py:371-390: ``(instance, offer)`` pairs,
    non-shared first then shared, stably sorted by ``instance.price or 0``."""
    pairs = []
    for inst in nonshared:
        offer = M.InstanceOfferWithAvailability.parse_raw(inst.offer)
        offer.availability = M.InstanceAvailability.IDLE
        pairs.append((inst, offer))
    pairs.extend(shared_with_offers)
    pairs.sort(key=lambda p: p[0].price or 0)
    return pairs


# --------------------------------------------------------------------------- a14


def plan_slice(offers: Sequence[Any], k: int):
    """``offers[:k]``, ``len(offers)``, ``max(price)`` as consumed at
    server/services/runs.py:341-348, server/services/fleets.py:268-270 and
    server/background/tasks/process_submitted_jobs.py:450."""
    prices = [o.price for o in offers]
    return list(offers[:k]), len(offers), (max(prices) if prices else None)


# --------------------------------------------------------------------------- callers


def assign_job_to_pool_instance(pool_instances: Sequence[Any], profile, requirements, *, fleet_model=None,
                                multinode=False, master_job_provisioning_data=None, volumes=None):
    """server/background/tasks/process_submitted_jobs.py:347-400 without the DB session: pick the
    cheapest candidate (unshared IDLE instances first, then shared ones with enough idle blocks),
    mark it BUSY and grow its busy_blocks.  Returns ``(instance, offer)`` or ``None``."""
    unshared = filter_pool_instances(pool_instances, profile, requirements=requirements, status=M.InstanceStatus.IDLE,
                                     fleet_model=fleet_model, multinode=multinode,
                                     master_job_provisioning_data=master_job_provisioning_data, volumes=volumes,
                                     shared=False)
    candidates = [(inst, M.InstanceOfferWithAvailability.parse_raw(inst.offer)) for inst in unshared]
    if not multinode:
        candidates.extend(get_shared_pool_instances_with_offers(pool_instances, profile, requirements, idle_only=True,
                                                                fleet_model=fleet_model, volumes=volumes))
    if not candidates:
        return None
    candidates.sort(key=lambda pair: pair[0].price or 0)
    inst, offer = candidates[0]
    inst.status = M.InstanceStatus.BUSY
    inst.busy_blocks += offer.blocks
    return inst, offer


def get_plan_offers(pool_offers: Sequence[Any], offers: Sequence[Tuple[Any, Any]], k: int = 50):
    """server/services/runs.py:336-349: pool offers, then backend offers, stable sort on
    availability, first ``k``, total, max price."""
    job_offers = list(pool_offers)
    job_offers.extend(o for _, o in offers)
    job_offers.sort(key=lambda o: not o.availability.is_available())
    return job_offers[:k], len(job_offers), max((o.price for o in job_offers), default=None)


def get_pool_offers(pool_instances: Sequence[Any], profile, requirements, *, multinode=False, volumes=None):
    """server/services/runs.py:674-715 without the DB: shared offers first, then unshared
    instances (IDLE / BUSY by status), sorted by price."""
    pool_offers = []
    if not multinode:
        for _, offer in get_shared_pool_instances_with_offers(pool_instances, profile, requirements, volumes=volumes):
            pool_offers.append(offer)
    for inst in filter_pool_instances(pool_instances, profile, requirements=requirements, multinode=multinode,
                                      volumes=volumes, shared=False):
        if inst.offer is None:
            continue
        offer = M.InstanceOfferWithAvailability.parse_raw(inst.offer)
        offer.availability = M.InstanceAvailability.IDLE if inst.status == M.InstanceStatus.IDLE else M.InstanceAvailability.BUSY
        pool_offers.append(offer)
    pool_offers.sort(key=lambda o: o.price)
    return pool_offers


# --------------------------------------------------------------------------- the scheduler pass


def pool_rows_open_to_jobs(pool_instances: Sequence[Any]) -> List[Any]:
    """The rows ``_process_submitted_job`` fetches for a job:
    ``deleted == False and total_blocks > busy_blocks``
    (server/background/tasks/process_submitted_jobs.py:186-193; a NULL ``total_blocks`` fails the
    SQL comparison, so such rows are not fetched at all)."""
    return [r for r in pool_instances
            if not getattr(r, "deleted", False) and r.total_blocks is not None
            and r.total_blocks > (r.busy_blocks or 0)]


def process_submitted_jobs(pool_instances: Sequence[Any], backends: Sequence[Any], jobs: Sequence[Any],
                           launch: Callable[[Any, Any, Any], Any], *, max_ticks: int = 2, k: int = 15):
    """The reference's one-job-at-a-time pass over SUBMITTED jobs
    (server/background/tasks/process_submitted_jobs.py:115-260, 347-400, 418-490) without the DB:

    * a follower (``job_num != 0``) waits until its master has provisioning data (:138-143), any
      non-first job waits until the run has a fleet (:148-153);
    * step 1 tries the pool (``assign_job_to_pool_instance``); the chosen row turns BUSY;
    * step 2: no instance and ``creation_policy == REUSE`` -> terminating (:244-250); otherwise
      ``get_offers_by_requirements(exclude_not_available=True, ...)[:15]`` are tried in order with
      ``launch(job, backend, offer)``, a ``BackendError``-like exception moves on to the next offer
      (:436-488); none left -> terminating (no capacity).

    The reference spreads the two steps of a job over two passes of its background task; here a
    *tick* runs both.  Whether a job waits is decided from the state at the START of a tick (what a
    pass that fetched the job rows up front would see); jobs that waited are tried again in the
    next tick, in their original order, for ``max_ticks`` ticks.  ``jobs`` carry the attributes of
    ``dstack_b200.scheduler.SubmittedJob``.  Returns one ``(status, instance, offer, jpd)`` tuple
    per job: status "provisioning" / "terminating" / "waiting"; ``instance`` is the pool row
    (None for a new instance)."""
    out: List[Any] = [("waiting", None, None, None)] * len(jobs)
    master_jpd = {}
    pending = list(range(len(jobs)))
    for _ in range(max_ticks):
        known = dict(master_jpd)  # state at the start of the tick
        waiting = []
        for i in pending:
            job = jobs[i]
            jpd = job.master_job_provisioning_data
            if job.job_num != 0:
                if jpd is None:
                    jpd = known.get((job.run_id, job.replica_num))
                if jpd is None:
                    waiting.append(i)
                    continue
            if (job.job_num != 0 or job.replica_num != 0) and job.fleet_model is None:
                waiting.append(i)
                continue
            multinode = job.jobs_per_replica > 1
            res = assign_job_to_pool_instance(pool_rows_open_to_jobs(pool_instances), job.profile, job.requirements,
                                              fleet_model=job.fleet_model, multinode=multinode,
                                              master_job_provisioning_data=jpd, volumes=job.volumes)
            if res is not None:
                inst, offer = res
                out[i] = ("provisioning", inst, offer, _jpd_of(inst))
            elif job.profile.creation_policy == M.CreationPolicy.REUSE:
                out[i] = ("terminating", None, None, None)
            else:
                offers = get_offers_by_requirements(
                    backends, job.profile, job.requirements, exclude_not_available=True,
                    multinode=multinode or job.cluster_placement, master_job_provisioning_data=jpd,
                    volumes=job.volumes, privileged=job.privileged, instance_mounts=job.instance_mounts)
                out[i] = ("terminating", None, None, None)
                for backend, offer in offers[:k]:
                    try:
                        new_jpd = launch(job, backend, offer)
                    except Exception:  # BackendError and anything else: next offer (:463-488)
                        continue
                    out[i] = ("provisioning", None, offer, new_jpd)
                    break
            if job.job_num == 0 and out[i][0] == "provisioning":
                master_jpd[(job.run_id, job.replica_num)] = out[i][3]
        pending = waiting
        if not pending:
            break
    return out


def placement_group_keys(offers: Sequence[Any], existing: Sequence[Tuple[Any, str]] = ()):
    """``(backend, region)`` pairs for which trying ``offers`` in order creates a placement group
    (server/background/tasks/process_instances.py:523-544, 910-932): the first offer of every
    pair not seen before, for backends that support placement groups."""
    seen = set(existing)
    new = []
    for o in offers:
        if o.backend not in M.BACKENDS_WITH_PLACEMENT_GROUPS_SUPPORT:
            continue
        key = (o.backend, o.region)
        if key not in seen:
            seen.add(key)
            new.append(key)
    return new
