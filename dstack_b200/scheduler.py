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

``plan_slice`` is the ``get_plan`` aggregation (``server/services/runs.py:336-349``): pool offers
first, stable re-partition by availability, first ``k`` / total / max price, computed from the
catalog's device top-k instead of the full offer list.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Any, List, Optional, Sequence, Tuple

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
