"""Mirror of the matching part of ``dstack._internal.server.services.pools`` (``:409-544``).

``filter_pool_instances`` and ``get_shared_pool_instances_with_offers`` keep the reference's
signatures and return the same objects in the same order.  The admin predicates (fleet,
reachability, name, status, backend, region, instance type, zone, blocks) and the resource
predicate are all evaluated on the GPU against a :class:`PoolIndex`: the pool rows are
encoded once -- including the ONE parse of each row's offer JSON that the reference repeats
for every (job, row) pair (``pools.py:499``) -- and every job is a record in a batch.
"""

from __future__ import annotations

from typing import Any, List, Optional, Sequence, Tuple

import numpy as np

from . import models as M
from .encode import (FLAG_BLOCKS_UNKNOWN, FLAG_NO_OFFER, FLAG_SHARED, FLAG_UNREACHABLE, STATUS_CODES, EncodedTable,
                     encode_table, state_code)
from .engine import DeviceTable, Matcher, default_matcher
from .ingest import JobFilter, PoolParseCache, columns_from_pool_instances, job_columns
from .offers import requirements_to_query_filter
from .services import generate_shared_offer


def _pool_filter(profile: Any, *, requirements: Optional[Any], status: Optional[Any], fleet_model: Optional[Any],
                 multinode: bool, master_job_provisioning_data: Optional[Any],
                 volumes: Optional[List[List[Any]]], shared: bool) -> JobFilter:
    """The derivations of ``pools.py:425-456`` folded into a :class:`JobFilter`."""
    backend_types, regions, zones = profile.backends, profile.regions, None
    if volumes:
        first_mount = volumes[0]
        backend_types = [v.configuration.backend for v in first_mount]
        regions = [v.configuration.region for v in first_mount]
        known = [v.provisioning_data.availability_zone for v in first_mount if v.provisioning_data is not None]
        zones = known or None
    if multinode:
        backend_types = [b for b in (backend_types or M.BACKENDS_WITH_MULTINODE_SUPPORT)
                         if b in M.BACKENDS_WITH_MULTINODE_SUPPORT]
    jpd = master_job_provisioning_data
    if jpd is not None:
        base = jpd.get_base_backend()
        backend_types = [b for b in (backend_types or [base]) if b == base]
        regions = [r for r in (regions or [jpd.region]) if r == jpd.region]
    return JobFilter(
        query=requirements_to_query_filter(requirements), backend_types=backend_types, regions=regions,
        instance_types=profile.instance_types, status=status,
        any_fleet=fleet_model is None, fleet_id=None if fleet_model is None else fleet_model.id,
        instance_name=profile.instance_name, zones=zones, shared=shared,
        require_offer=requirements is not None,
    )


def patch_pool_rows(t: EncodedTable, instances: Sequence[Any], changed: Sequence[int], row_of: np.ndarray,
                    offers: Optional[Sequence[Any]] = None) -> Optional[np.ndarray]:
    """Host half of :meth:`PoolIndex.refresh`: re-encode status / reachability / block state / fleet
    of the instances at ``changed`` into ``t.rows``.  Returns the physical rows that changed, or
    ``None`` -- nothing written -- when an instance moved to a fleet the table's dictionary does not
    hold (status and flags are a closed dictionary: every combination has its code)."""
    state_col, fleet_col = t.col("state"), t.col("fleet")
    fleet_lookup = t.columns[fleet_col].lookup
    unknown_fleet = t.columns[fleet_col].card - 1
    rows = row_of[np.asarray(changed, dtype=np.int64)]
    new_codes = []
    for pos in changed:
        inst = instances[pos]
        flags = 0
        if inst.unreachable:
            flags |= FLAG_UNREACHABLE
        if inst.total_blocks is None:
            flags |= FLAG_BLOCKS_UNKNOWN
        elif inst.total_blocks > 1:
            flags |= FLAG_SHARED
        if (offers is None or offers[pos] is None) and inst.offer is None:
            flags |= FLAG_NO_OFFER
        st = inst.status
        code = state_code(STATUS_CODES[st.value if hasattr(st, "value") else str(st)], flags)
        fleet = unknown_fleet if inst.fleet_id is None else fleet_lookup.get(inst.fleet_id, -1)
        if fleet < 0:
            return None
        new_codes.append((code, fleet))
    for r, (code, fleet) in zip(rows, new_codes):
        t.rows[r, state_col] = code
        t.rows[r, fleet_col] = fleet
    return rows


class PoolIndex:
    """Pool rows resident on the GPU.  Build once per scheduler tick, query for many jobs."""

    def __init__(self, pool_instances: Sequence[Any], *, matcher: Optional[Matcher] = None,
                 offers: Optional[Sequence[Any]] = None, parse_cache: Optional[PoolParseCache] = None):
        """``parse_cache``: a :class:`~dstack_b200.ingest.PoolParseCache` kept by the caller across
        ticks, so that only rows whose stored JSON changed are parsed again."""
        self.matcher = matcher or default_matcher()
        self.instances = list(pool_instances)
        self._offers, self._parse_cache = offers, parse_cache
        self.table: Optional[EncodedTable] = None
        self.device_table: Optional[DeviceTable] = None
        self.rebuilds = 0
        self._build()

    def _build(self) -> None:
        if self.device_table is not None:
            self.device_table.close()
            self.device_table = None
        if self.instances:
            self.table = encode_table(columns_from_pool_instances(self.instances, offers=self._offers,
                                                                  parse_cache=self._parse_cache))
            self.device_table = self.matcher.upload(self.table)
            self._row_of = np.empty(self.table.n_src, np.int64)  # source position -> physical row
            live = self.table.perm >= 0
            self._row_of[self.table.perm[live]] = np.flatnonzero(live)
        self.rebuilds += 1

    def refresh(self, changed: Sequence[int]) -> bool:
        """The instances at positions ``changed`` (of the list this index was built from) have a new
        status, reachability, block state or fleet -- what a scheduler pass changes
        (``process_submitted_jobs.py:398-399``) and what ``filter_pool_instances`` reads fresh for
        every job (``pools.py:458-488``).  Their rows are rewritten in place on the device.  Returns
        True when that sufficed; when an instance moved to a fleet never seen in this pool the table
        is rebuilt and False is returned.  Changes to an instance's offer, backend, region, name or
        provisioning data, and added or removed instances, need a new index."""
        if self.table is None or len(changed) == 0:
            return True
        rows = patch_pool_rows(self.table, self.instances, changed, self._row_of, self._offers)
        if rows is None:
            self._build()
            return False
        self.device_table.patch_rows(rows)
        return True

    def close(self) -> None:
        if self.device_table is not None:
            self.device_table.close()
            self.device_table = None

    def match(self, filters: Sequence[JobFilter]) -> List[np.ndarray]:
        """Per job: positions (input order) of the rows that pass."""
        if not self.instances:
            return [np.zeros(0, np.int64) for _ in filters]
        jobs = job_columns(self.table, filters)
        return self.matcher.match_all(self.device_table, jobs, order="source")

    def filter(self, profile: Any, *, requirements: Optional[Any] = None, status: Optional[Any] = None,
               fleet_model: Optional[Any] = None, multinode: bool = False,
               master_job_provisioning_data: Optional[Any] = None,
               volumes: Optional[List[List[Any]]] = None, shared: bool = False) -> List[Any]:
        f = _pool_filter(profile, requirements=requirements, status=status, fleet_model=fleet_model,
                         multinode=multinode, master_job_provisioning_data=master_job_provisioning_data,
                         volumes=volumes, shared=shared)
        return [self.instances[i] for i in self.match([f])[0]]


def filter_pool_instances(
    pool_instances: List[Any],
    profile: Any,
    *,
    requirements: Optional[Any] = None,
    status: Optional[Any] = None,
    fleet_model: Optional[Any] = None,
    multinode: bool = False,
    master_job_provisioning_data: Optional[Any] = None,
    volumes: Optional[List[List[Any]]] = None,
    shared: bool = False,
    matcher: Optional[Matcher] = None,
    parse_cache: Optional[PoolParseCache] = None,
) -> List[Any]:
    """``server/services/pools.py:409-503``: matching rows, same objects, input order."""
    index = PoolIndex(pool_instances, matcher=matcher, parse_cache=parse_cache)
    try:
        return index.filter(profile, requirements=requirements, status=status, fleet_model=fleet_model,
                            multinode=multinode, master_job_provisioning_data=master_job_provisioning_data,
                            volumes=volumes, shared=shared)
    finally:
        index.close()


class SharedPoolIndex:
    """Shared (multi-block) instances expanded to one row per block count ``1..total_blocks``
    (``pools.py:533-543``): row ``(i, b)`` carries the resources of ``generate_shared_offer(offer_i,
    b, total_i)``.  A job's answer for instance ``i`` is its smallest matching ``b``."""

    def __init__(self, pool_instances: Sequence[Any], *, matcher: Optional[Matcher] = None):
        self.instances = list(pool_instances)
        self.owner: List[int] = []
        self.blocks: List[int] = []
        self.block_offers: List[Any] = []
        rows = []
        for i, inst in enumerate(self.instances):
            if inst.offer is None or inst.total_blocks is None:
                continue
            offer = M.InstanceOfferWithAvailability.parse_raw(inst.offer)
            for b in range(1, inst.total_blocks + 1):
                self.owner.append(i)
                self.blocks.append(b)
                self.block_offers.append(generate_shared_offer(offer, b, inst.total_blocks))
                rows.append(inst)
        self.index = PoolIndex(rows, matcher=matcher, offers=self.block_offers)

    def close(self) -> None:
        self.index.close()

    def smallest_blocks(self, filters: Sequence[JobFilter]) -> List[List[Tuple[int, int]]]:
        """Per job: ``[(instance position, virtual row)]`` with the smallest matching block
        count of each instance, in instance order."""
        out = []
        for rows in self.index.match(filters):
            best: List[Tuple[int, int]] = []
            seen = -1
            for r in rows:  # rows ascend; block counts ascend within an instance
                i = self.owner[r]
                if i != seen:
                    best.append((i, int(r)))
                    seen = i
            out.append(best)
        return out


def get_shared_pool_instances_with_offers(
    pool_instances: List[Any],
    profile: Any,
    requirements: Any,
    *,
    idle_only: bool = False,
    fleet_model: Optional[Any] = None,
    volumes: Optional[List[List[Any]]] = None,
    matcher: Optional[Matcher] = None,
) -> List[Tuple[Any, Any]]:
    """``server/services/pools.py:506-544``."""
    shared = SharedPoolIndex(pool_instances, matcher=matcher)
    try:
        f = _pool_filter(profile, requirements=requirements, status=None, fleet_model=fleet_model,
                         multinode=False, master_job_provisioning_data=None, volumes=volumes, shared=True)
        picks = shared.smallest_blocks([f])[0]
    finally:
        shared.close()
    result = []
    for i, row in picks:
        inst = shared.instances[i]
        if idle_only and inst.status not in (M.InstanceStatus.IDLE, M.InstanceStatus.BUSY):
            continue
        offer = shared.block_offers[row]
        idle_blocks = inst.total_blocks - inst.busy_blocks
        offer.availability = (M.InstanceAvailability.IDLE if shared.blocks[row] <= idle_blocks
                              else M.InstanceAvailability.BUSY)
        if offer.availability == M.InstanceAvailability.IDLE or not idle_only:
            result.append((inst, offer))
    return result
