"""Pool ingest across scheduler ticks: the stored offer / provisioning JSON
of an instance is parsed once per distinct text, not once per tick and never per (job, row) as at
``server/services/pools.py:499``."""

import dataclasses

import numpy as np
import pytest

from dstack_b200 import models as M
from workloads import synth
from dstack_b200.ingest import PoolParseCache, columns_from_pool_instances


def _same_columns(a, b):
    for f in dataclasses.fields(a):
        x, y = getattr(a, f.name), getattr(b, f.name)
        if f.name == "shapes":
            for g in dataclasses.fields(x):
                u, v = getattr(x, g.name), getattr(y, g.name)
                if isinstance(u, np.ndarray):
                    assert np.array_equal(u, v, equal_nan=u.dtype.kind == "f"), g.name
                else:
                    assert u == v, g.name
        elif isinstance(x, np.ndarray):
            assert np.array_equal(x, y, equal_nan=x.dtype.kind == "f"), f.name
        else:
            assert x == y, f.name


def test_cached_ingest_equals_plain_ingest_and_parses_once(monkeypatch):
    pool = synth.make_pool(400, seed=2, n_fleets=8)
    plain = columns_from_pool_instances(pool)
    cache = PoolParseCache()
    first = columns_from_pool_instances(pool, parse_cache=cache)
    _same_columns(plain, first)
    parsed = cache.parsed
    assert 0 < parsed <= 2 * len(pool)

    # second tick: nothing may be parsed again
    def boom(*a, **k):
        raise AssertionError("unchanged JSON was parsed again")

    monkeypatch.setattr(M.InstanceOffer, "parse_raw", classmethod(boom))
    monkeypatch.setattr(M.JobProvisioningData, "parse_raw", classmethod(boom))
    second = columns_from_pool_instances(pool, parse_cache=cache)
    _same_columns(plain, second)
    assert cache.parsed == parsed


def test_changed_rows_are_parsed_again_and_fresh_columns_are_read_every_tick():
    pool = synth.make_pool(200, seed=5, n_fleets=4)
    cache = PoolParseCache()
    columns_from_pool_instances(pool, parse_cache=cache)
    parsed = cache.parsed
    victim = next(i for i, inst in enumerate(pool) if inst.offer is not None)
    offer = M.InstanceOffer.parse_raw(pool[victim].offer)
    offer.price = offer.price + 123.0
    pool[victim].offer = offer.json()
    other = (victim + 1) % len(pool)
    pool[other].status = M.InstanceStatus.TERMINATED  # a plain column: no JSON involved
    pool[other].unreachable = True
    again = columns_from_pool_instances(pool, parse_cache=cache)
    assert cache.parsed == parsed + 1
    _same_columns(columns_from_pool_instances(pool), again)
    assert again.price[victim] == offer.price


def test_malformed_stored_offer_raises_like_the_reference():
    pool = synth.make_pool(10, seed=7, n_fleets=2)
    victim = next(i for i, inst in enumerate(pool) if inst.offer is not None)
    pool[victim].offer = '{"backend": "aws"'
    with pytest.raises(Exception) as e:
        columns_from_pool_instances(pool, parse_cache=PoolParseCache())
    assert "ValidationError" in type(e.value).__name__ or "JSON" in type(e.value).__name__ or "Error" in type(e.value).__name__


def test_cache_is_bounded():
    cache = PoolParseCache(max_entries=16)
    pool = synth.make_pool(100, seed=9, n_fleets=2)
    a = columns_from_pool_instances(pool, parse_cache=cache)
    assert len(cache._offers) <= 16 and len(cache._jpds) <= 16
    _same_columns(columns_from_pool_instances(pool), a)


def test_offer_catalog_cache_follows_the_30_s_clock():
    from dstack_b200.services import OfferCatalogCache

    now = [100.0]
    built, closed = [], []

    class FakeCatalog:
        def __init__(self, offers_by_backend):
            built.append(offers_by_backend)

        def close(self):
            closed.append(self)

    fetches = []

    def fetch():
        fetches.append(now[0])
        return [("backend", [len(fetches)])]

    cache = OfferCatalogCache(fetch, ttl=30.0, clock=lambda: now[0], build=FakeCatalog)
    a = cache.get()
    now[0] += 29.9
    assert cache.get() is a and len(fetches) == 1
    now[0] += 0.2  # 30.1 s after the build
    b = cache.get()
    assert b is not a and closed == [a] and cache.rebuilds == 2
    cache.invalidate()
    c = cache.get()
    assert c is not b and closed == [a, b]
    cache.close()
    assert closed == [a, b, c]


def test_offer_catalog_cache_serves_the_old_table_while_the_new_one_is_built():
    """Matcher threads must not stall for a rebuild (0.5 s per Mi rows): the thread that notices
    the expiry builds outside the lock, the others keep the previous catalog until the swap."""
    import threading

    from dstack_b200.services import OfferCatalogCache

    now = [0.0]
    gate, started = threading.Event(), threading.Event()
    built = []

    class SlowCatalog:
        def __init__(self, offers_by_backend):
            if built:  # the first build is immediate, the rebuild waits for the test
                started.set()
                assert gate.wait(10)
            built.append(self)

        def close(self):
            pass

    cache = OfferCatalogCache(lambda: [], ttl=30.0, clock=lambda: now[0], build=SlowCatalog)
    first = cache.get()
    now[0] = 31.0
    result = []
    t = threading.Thread(target=lambda: result.append(cache.get()))
    t.start()
    assert started.wait(10)          # the rebuild is running ...
    assert cache.get() is first      # ... and other callers are served at once, from the old table
    gate.set()
    t.join(10)
    assert result and result[0] is not first and cache.get() is result[0] and cache.rebuilds == 2


def _status_tick(rows, rng, share=0.01):
    """What a scheduler pass changes on a pool: some idle instances turn busy (and back), one
    becomes unreachable, one moves to another fleet that the pool already knows."""
    n = max(4, int(len(rows) * share))
    changed = sorted(int(i) for i in rng.choice(len(rows), size=n, replace=False))
    fleets = [r.fleet_id for r in rows if r.fleet_id is not None]
    for c, i in enumerate(changed):
        r = rows[i]
        if c % 4 == 0:
            r.unreachable = not r.unreachable
        elif c % 4 == 1:
            r.fleet_id = fleets[int(rng.integers(0, len(fleets)))]
        elif r.status == M.InstanceStatus.IDLE:
            r.status, r.busy_blocks = M.InstanceStatus.BUSY, (r.total_blocks or 1)
        elif r.status == M.InstanceStatus.BUSY:
            r.status, r.busy_blocks = M.InstanceStatus.IDLE, 0
        else:
            r.status = M.InstanceStatus.IDLE
    return changed


def test_rows_patched_in_place_answer_like_a_rebuilt_table():
    """status / reachability / fleet changes rewrite the affected rows of the
    resident table instead of rebuilding it; every job must get the rows a rebuilt table gives."""
    import device_emulator as emu
    from dstack_b200 import pools as PP
    from dstack_b200.encode import encode_table, pack_jobs
    from dstack_b200.ingest import job_columns
    from workloads import configs

    rng = np.random.default_rng(7)
    rows = synth.make_pool(3000, seed=71, n_fleets=10)
    table = encode_table(columns_from_pool_instances(rows))
    row_of = np.empty(table.n_src, np.int64)
    live = table.perm >= 0
    row_of[table.perm[live]] = np.flatnonzero(live)
    calls = configs.config2_jobs(rows, 60, seed=72)
    filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"], fleet_model=kw.get("fleet_model"),
                               multinode=kw.get("multinode", False),
                               master_job_provisioning_data=kw.get("master_job_provisioning_data"), volumes=None,
                               shared=False) for p, kw in calls]
    patched_ticks = 0
    for tick in range(4):
        changed = _status_tick(rows, rng)
        got_rows = PP.patch_pool_rows(table, rows, changed, row_of)
        fresh = encode_table(columns_from_pool_instances(rows))
        if got_rows is None:  # a combination the dictionaries do not hold: the caller rebuilds
            table = fresh
            continue
        patched_ticks += 1
        assert len(got_rows) == len(changed)
        a = emu.match_matrix(table, pack_jobs(table, job_columns(table, filters)))
        b = emu.match_matrix(fresh, pack_jobs(fresh, job_columns(fresh, filters)))
        # same source rows for every job (the two tables may number their dictionaries differently)
        src_a = [np.sort(table.perm[np.flatnonzero(m)]) for m in a]
        src_b = [np.sort(fresh.perm[np.flatnonzero(m)]) for m in b]
        assert all(np.array_equal(x, y) for x, y in zip(src_a, src_b))
        assert sum(len(x) for x in src_a) > 100
    assert patched_ticks >= 2
