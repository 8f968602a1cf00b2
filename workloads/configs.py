"""The job side of BASELINE.json's configs 1, 2 and 5, and a mixed scheduler pass.

The catalogs / pools come from :mod:`workloads.synth`; everything is seeded.
"""

from __future__ import annotations

import types
from typing import Any, Dict, List, Tuple

import numpy as np

from dstack_b200 import models as M
from dstack_b200.scheduler import SubmittedJob

GPU_SPECS = ["", "16GB..:1..8", "A100", "H100:8", "24GB..80GB:1..8", "T4,L4:1..4"]


def config1_request() -> Tuple[M.Profile, M.Requirements]:
    """One task ``gpu: 24GB..80GB:1..8, spot: auto``."""
    return M.Profile(name="p"), M.Requirements(resources=M.ResourcesSpec(gpu="24GB..80GB:1..8"), max_price=None, spot=None)


def config2_jobs(rows: List[Any], n_jobs: int, seed: int = 2) -> List[Tuple[M.Profile, Dict[str, Any]]]:
    """``filter_pool_instances`` calls of config 2: 30 % same-fleet, 20 % multinode with a master
    node, 10 % reuse by instance name.  Returns ``(profile, keyword arguments)`` per job."""
    rng = np.random.Generator(np.random.PCG64(seed))
    fleets = sorted({r.fleet_id for r in rows if r.fleet_id is not None}, key=str)
    with_jpd = [r for r in rows if r.job_provisioning_data is not None]
    calls = []
    for _ in range(n_jobs):
        gpu = GPU_SPECS[int(rng.integers(0, len(GPU_SPECS)))]
        req = M.Requirements(resources=M.ResourcesSpec(gpu=gpu or None, cpu=f"{int(rng.choice([1, 2, 4, 8]))}.."),
                             max_price=None if rng.random() < 0.5 else float(np.round(rng.lognormal(1.0, 1.0), 2)),
                             spot=None)
        kw: Dict[str, Any] = dict(requirements=req, status=M.InstanceStatus.IDLE, shared=False)
        profile = M.Profile(name="p")
        u = rng.random()
        if u < 0.3:
            kw["fleet_model"] = types.SimpleNamespace(id=fleets[int(rng.integers(0, len(fleets)))])
        elif u < 0.5:
            jpd = M.JobProvisioningData.parse_raw(with_jpd[int(rng.integers(0, len(with_jpd)))].job_provisioning_data)
            kw.update(multinode=True, master_job_provisioning_data=jpd)
        elif u < 0.6:
            profile = M.Profile(name="p", instance_name=rows[int(rng.integers(0, len(rows)))].name,
                                creation_policy=M.CreationPolicy.REUSE)
        calls.append((profile, kw))
    return calls


def config5_jobs(n_clusters: int = 8, n_nodes: int = 8, gpu: str = "H100:8") -> List[SubmittedJob]:
    """Config 5: ``n_clusters`` multinode tasks of ``n_nodes`` nodes ``gpu: H100:8`` each; node 0 of
    a task is its master, nodes 1.. follow it to its backend and region (offers.py:71-79)."""
    req = M.Requirements(resources=M.ResourcesSpec(gpu=gpu), max_price=None, spot=None)
    jobs = []
    for c in range(n_clusters):
        fleet = types.SimpleNamespace(id=f"cluster-fleet-{c}")
        for n in range(n_nodes):
            jobs.append(SubmittedJob(run_id=f"cluster-{c}", profile=M.Profile(name=f"cluster-{c}"), requirements=req,
                                     job_num=n, jobs_per_replica=n_nodes, fleet_model=fleet, cluster_placement=True))
    return jobs


def scheduler_jobs(rows: List[Any], n_jobs: int, seed: int = 6) -> List[SubmittedJob]:
    """A mixed pass: single-node jobs (some pinned to a fleet, some REUSE-only, some with a
    max_price), and every sixth run a 3-node multinode task whose followers come right behind
    their master."""
    rng = np.random.Generator(np.random.PCG64(seed))
    fleets = sorted({r.fleet_id for r in rows if r.fleet_id is not None}, key=str)
    specs = [None, None, "16GB..:1..8", "A100", "1..2", "T4,L4:1..4", "24GB..80GB:1..8", "H100:8"]
    jobs: List[SubmittedJob] = []
    run = 0
    while len(jobs) < n_jobs:
        gpu = specs[int(rng.integers(0, len(specs)))]
        req = M.Requirements(resources=M.ResourcesSpec(gpu=gpu, cpu=f"{int(rng.choice([1, 2, 4, 8]))}.."),
                             max_price=None if rng.random() < 0.6 else float(np.round(rng.lognormal(1.0, 1.0), 2)),
                             spot=None)
        fleet = None
        if rng.random() < 0.3 and fleets:
            fleet = types.SimpleNamespace(id=fleets[int(rng.integers(0, len(fleets)))])
        policy = M.CreationPolicy.REUSE if rng.random() < 0.15 else None
        profile = M.Profile(name=f"run-{run}", creation_policy=policy)
        if run % 6 == 5:
            cluster_fleet = fleet or types.SimpleNamespace(id=f"new-fleet-{run}")
            for n in range(3):
                jobs.append(SubmittedJob(run_id=run, profile=profile, requirements=req, job_num=n, jobs_per_replica=3,
                                         fleet_model=cluster_fleet, cluster_placement=True))
        else:
            jobs.append(SubmittedJob(run_id=run, profile=profile, requirements=req, fleet_model=fleet))
        run += 1
    return jobs[:n_jobs]


def fake_launch(fail_every: int = 4):
    """A ``launch(job, backend, offer)`` for tests and the bench: deterministic "no capacity" for
    every ``fail_every``-th (instance type, region) pair, otherwise provisioning data naming the
    offer's backend and region (what followers are pinned to)."""
    import zlib

    def launch(job, backend, offer):
        h = zlib.crc32(f"{offer.instance.name}/{offer.region}".encode())
        if fail_every and h % fail_every == 0:
            raise RuntimeError("no capacity")
        return M.JobProvisioningData(backend=offer.backend, instance_type=offer.instance, instance_id=f"i-{h:08x}",
                                     region=offer.region, price=offer.price, username="ubuntu", dockerized=True)

    return launch
