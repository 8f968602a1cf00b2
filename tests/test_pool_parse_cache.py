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
