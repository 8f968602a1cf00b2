"""Batched form of the scheduler's pool step and of the plan slice (row 3).

The reference assigns SUBMITTED jobs to pool instances one at a time
(``server/background/tasks/process_submitted_jobs.py:347-400``): per job it filters the whole
pool twice (``filter_pool_instances`` for unshared IDLE instances, ``get_shared_pool_instances_with_offers``
for shared ones), sorts the candidates by ``instance.price or 0`` and takes the first; the chosen
instance becomes BUSY and its ``busy_blocks`` grows, which the next job sees.  Here the two
filters of ALL pending jobs are answered by one GPU call each (the predicates do not depend on
the mutable state), and only the cheap part -- walking the jobs in order and skipping instances
that an earlier job took -- stays sequential on the host.  The result is exactly what the
sequential reference loop produces for the same job order.

``process_submitted_batch`` is the whole pass over SUBMITTED jobs
(``process_submitted_jobs.py:115-260, 418-490``): waiting rules for followers, the pool step, and
for the jobs the pool could not take the ``offers[:15]`` of ``_run_job_on_new_instance`` from ONE
batched catalog match, tried in order through the caller's ``launch``; followers of a master
placed in this pass are matched in a second phase against the master's backend and region.

``plan_slice`` is the ``get_plan`` aggregation (``server/services/runs.py:336-349``): pool offers
first, stable re-partition by availability, first ``k`` / total / max price, computed from the
catalog's device top-k instead of the full offer list.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import models as M
from .engine import Matcher, TopK
from .pools import PoolIndex, SharedPoolIndex, _pool_filter


@dataclass
class JobRequest:
    """What ``_assign_job_to_pool_instance`` reads of a job."""

    profile: Any
    requirements: Any
    fleet_model: Optional[Any] = None
    multinode: bool = False
    master_job_provisioning_data: Optional[Any] = None
    volumes: Optional[List[List[Any]]] = None


@dataclass
class Assignment:
    instance: Optional[Any]      # the chosen pool row (None: no candidate -> provision a new instance)
    blocks: int = 0              # blocks taken on it (1 for an unshared instance)
    offer: Optional[Any] = None  # the (possibly block-sliced) offer, availability IDLE


def assign_jobs_to_pool(pool_instances: Sequence[Any], jobs: Sequence[JobRequest], *,
                        matcher: Optional[Matcher] = None, mutate: bool = True,
                        parse_cache: Optional[Any] = None) -> List[Assignment]:
    """Assign ``jobs`` (in order) to pool instances as the reference's per-job loop would.

    With ``mutate`` the chosen rows get ``status = BUSY`` and ``busy_blocks += blocks`` like
    ``process_submitted_jobs.py:398-399``.  ``parse_cache`` (a ``PoolParseCache`` kept by the
    scheduler between ticks) spares re-parsing the stored JSON of unchanged instances."""
    rows = list(pool_instances)
    if not jobs:
        return []
    unshared = PoolIndex(rows, matcher=matcher, parse_cache=parse_cache)
    shared = SharedPoolIndex(rows, matcher=matcher)
    try:
        f_unshared = [_pool_filter(j.profile, requirements=j.requirements, status=M.InstanceStatus.IDLE,
                                   fleet_model=j.fleet_model, multinode=j.multinode,
                                   master_job_provisioning_data=j.master_job_provisioning_data,
                                   volumes=j.volumes, shared=False) for j in jobs]
        hits_unshared = unshared.match(f_unshared)
        # the reference skips the shared search for multinode jobs (:374)
        single = [i for i, j in enumerate(jobs) if not j.multinode]
        f_shared = [_pool_filter(jobs[i].profile, requirements=jobs[i].requirements, status=None,
                                 fleet_model=jobs[i].fleet_model, multinode=False, master_job_provisioning_data=None,
                                 volumes=jobs[i].volumes, shared=True) for i in single]
        picks_shared = dict(zip(single, shared.smallest_blocks(f_shared))) if single else {}
    finally:
        unshared.close()
        shared.close()

    price = np.array([(r.price or 0) for r in rows], dtype=np.float64)
    taken = np.zeros(len(rows), dtype=bool)                      # unshared row already given away
    busy = np.array([r.busy_blocks or 0 for r in rows], dtype=np.int64)
    total = np.array([r.total_blocks or 0 for r in rows], dtype=np.int64)
    usable = np.array([r.status in (M.InstanceStatus.IDLE, M.InstanceStatus.BUSY) for r in rows], dtype=bool)
    vblocks = np.asarray(shared.blocks, dtype=np.int64)

    out: List[Assignment] = []
    for i, job in enumerate(jobs):
        cand = hits_unshared[i]
        cand = cand[~taken[cand]]
        keys = [price[cand]]
        sh_pos = sh_row = None
        if i in picks_shared and picks_shared[i]:
            sh = np.asarray(picks_shared[i], dtype=np.int64).reshape(-1, 2)
            ok = usable[sh[:, 0]] & (vblocks[sh[:, 1]] <= total[sh[:, 0]] - busy[sh[:, 0]])  # idle_only
            sh_pos, sh_row = sh[ok, 0], sh[ok, 1]
            keys.append(price[sh_pos])
        key = np.concatenate(keys)
        if len(key) == 0:
            out.append(Assignment(None))
            continue
        best = int(np.argmin(key))  # first minimum: unshared before shared, each in pool order
        if best < len(cand):
            pos = int(cand[best])
            taken[pos] = True
            offer = M.InstanceOfferWithAvailability.parse_raw(rows[pos].offer)
            blocks = offer.blocks
        else:
            pos, vrow = int(sh_pos[best - len(cand)]), int(sh_row[best - len(cand)])
            offer = shared.block_offers[vrow].copy()
            offer.availability = M.InstanceAvailability.IDLE
            blocks = int(vblocks[vrow])
        busy[pos] += blocks
        if mutate:
            rows[pos].status = M.InstanceStatus.BUSY
            rows[pos].busy_blocks = int(busy[pos])
        out.append(Assignment(rows[pos], blocks, offer))
    return out


def plan_slice(pool_offers: Sequence[Any], catalog: Any, top: TopK, job: int, k: int = 50):
    """``offers[:k]``, ``total_offers`` and ``max_price`` of a job plan (``runs.py:336-349``).

    ``pool_offers``: the job's pool offers, already price-sorted (``runs.py:714``);
    ``catalog`` / ``top``: an :class:`~dstack_b200.services.OfferCatalog` and its
    ``match_topk(filters, k)`` result (same ``k``).  The plan list is pool offers followed by
    catalog offers, stably re-partitioned so that unavailable ones come last; the catalog part
    is already in that order on the device, so its first ``k`` rows are all that is needed."""
    if top.index.shape[1] < k:
        raise ValueError("the catalog top-k must be at least as long as the plan slice")
    table = catalog.table
    idx = top.index[job]
    idx = idx[idx >= 0]
    cloud = [catalog.offers[i] for i in idx]
    cloud_avail = [o for o in cloud if o.availability.is_available()]
    cloud_rest = [o for o in cloud if not o.availability.is_available()]
    pool_avail = [o for o in pool_offers if o.availability.is_available()]
    pool_rest = [o for o in pool_offers if not o.availability.is_available()]
    first = (pool_avail + cloud_avail + pool_rest + cloud_rest)[:k]
    total = len(pool_offers) + int(top.count[job])
    prices = [o.price for o in pool_offers]
    if not np.isnan(top.max_price[job]):
        prices.append(float(top.max_price[job]))
    return first, total, (max(prices) if prices else None)


# ---------------------------------------------------------------------------- the whole pass


@dataclass
class SubmittedJob:
    """What ``_process_submitted_job`` reads of a SUBMITTED job and its run
    (``process_submitted_jobs.py:115-260``)."""

    run_id: Any
    profile: Any
    requirements: Any
    job_num: int = 0
    replica_num: int = 0
    jobs_per_replica: int = 1               # > 1: a multinode task
    fleet_model: Optional[Any] = None       # run_model.fleet (its ``id`` is read); None: the run has no fleet yet
    cluster_placement: bool = False         # the fleet asks for InstanceGroupPlacement.CLUSTER
    master_job_provisioning_data: Optional[Any] = None  # of a master provisioned in an earlier pass
    volumes: Optional[List[List[Any]]] = None
    privileged: bool = False
    instance_mounts: bool = False


@dataclass
class JobOutcome:
    status: str                       # "provisioning" | "terminating" | "waiting"
    instance: Optional[Any] = None    # the pool row taken (None: a new instance, or nothing)
    offer: Optional[Any] = None
    job_provisioning_data: Optional[Any] = None


def _open_to_jobs(pool_instances: Sequence[Any]) -> List[Any]:
    # ``deleted == False and total_blocks > busy_blocks`` (process_submitted_jobs.py:186-193)
    return [r for r in pool_instances
            if not getattr(r, "deleted", False) and r.total_blocks is not None
            and r.total_blocks > (r.busy_blocks or 0)]


def process_submitted_batch(pool_instances: Sequence[Any], catalog: Any, jobs: Sequence[SubmittedJob],
                            launch: Callable[[SubmittedJob, Any, Any], Any], *, k: int = 15,
                            max_ticks: int = 2, matcher: Optional[Matcher] = None,
                            parse_cache: Optional[Any] = None) -> List[JobOutcome]:
    """One scheduler pass over ``jobs`` (in order) with both matching steps batched on the GPU.

    Per tick: the jobs that need not wait (followers need their master's provisioning data, known
    at the start of the tick; non-first jobs need the run's fleet) go through
    :func:`assign_jobs_to_pool`; those left without an instance -- unless their profile says
    REUSE, which ends them -- get ``offers[:k]`` from one ``catalog.match_topk`` call
    (``exclude_not_available=True`` and each job's multinode / master / volumes / privileged /
    instance-mount restrictions) and ``launch(job, backend, offer)`` is tried on them in order; an
    exception means "next offer".  Jobs that waited run in the next tick (``max_ticks`` in all),
    so the followers of a master placed in tick 1 are matched in tick 2 against the master's
    backend and region.  Outcomes equal those of the sequential reference loop for the same job
    order (``oracle/reference_path.py:process_submitted_jobs``)."""
    from . import services as S

    out = [JobOutcome("waiting") for _ in jobs]
    master_jpd: Dict[Tuple[Any, int], Any] = {}
    pending = list(range(len(jobs)))
    for _ in range(max_ticks):
        ready: List[int] = []
        jpds: Dict[int, Any] = {}
        waiting: List[int] = []
        for i in pending:
            job = jobs[i]
            jpd = job.master_job_provisioning_data
            if job.job_num != 0:
                if jpd is None:
                    jpd = master_jpd.get((job.run_id, job.replica_num))
                if jpd is None:
                    waiting.append(i)
                    continue
            if (job.job_num != 0 or job.replica_num != 0) and job.fleet_model is None:
                waiting.append(i)
                continue
            ready.append(i)
            jpds[i] = jpd
        if ready:
            # step 1: the pool, for every ready job at once
            requests = [JobRequest(profile=jobs[i].profile, requirements=jobs[i].requirements,
                                   fleet_model=jobs[i].fleet_model, multinode=jobs[i].jobs_per_replica > 1,
                                   master_job_provisioning_data=jpds[i], volumes=jobs[i].volumes) for i in ready]
            assigned = assign_jobs_to_pool(_open_to_jobs(pool_instances), requests, matcher=matcher,
                                           parse_cache=parse_cache)
            need_new: List[int] = []
            for i, a in zip(ready, assigned):
                if a.instance is not None:
                    raw = a.instance.job_provisioning_data
                    out[i] = JobOutcome("provisioning", a.instance, a.offer,
                                        None if raw is None else M.JobProvisioningData.parse_raw(raw))
                elif jobs[i].profile.creation_policy == M.CreationPolicy.REUSE:
                    out[i] = JobOutcome("terminating")
                else:
                    need_new.append(i)
            # step 2: a new instance, offers[:k] of every such job from one catalog match
            if need_new:
                filters = [S.job_filter(jobs[i].profile, jobs[i].requirements, exclude_not_available=True,
                                        multinode=jobs[i].jobs_per_replica > 1 or jobs[i].cluster_placement,
                                        master_job_provisioning_data=jpds[i], volumes=jobs[i].volumes,
                                        privileged=jobs[i].privileged, instance_mounts=jobs[i].instance_mounts)
                           for i in need_new]
                top = catalog.match_topk(filters, k)
                for row, i in enumerate(need_new):
                    out[i] = JobOutcome("terminating")
                    for backend, offer in catalog.pairs(top.index[row]):
                        try:
                            new_jpd = launch(jobs[i], backend, offer)
                        except Exception:  # BackendError or anything else: next offer
                            continue
                        out[i] = JobOutcome("provisioning", None, offer, new_jpd)
                        break
            for i in ready:
                if jobs[i].job_num == 0 and out[i].status == "provisioning":
                    master_jpd[(jobs[i].run_id, jobs[i].replica_num)] = out[i].job_provisioning_data
        pending = waiting
        if not pending:
            break
    return out


def placement_group_keys(offers: Sequence[Any], existing: Sequence[Tuple[Any, str]] = ()) -> List[Tuple[Any, str]]:
    """The ``(backend, region)`` pairs for which trying ``offers`` in order creates a placement group
    (``process_instances.py:523-544, 910-932``): one per pair not seen before, only for backends
    with placement-group support.  The cloud call that creates the group is the caller's."""
    seen = set(existing)
    new: List[Tuple[Any, str]] = []
    for o in offers:
        if o.backend not in M.BACKENDS_WITH_PLACEMENT_GROUPS_SUPPORT:
            continue
        key = (o.backend, o.region)
        if key not in seen:
            seen.add(key)
            new.append(key)
    return new
