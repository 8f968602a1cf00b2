"""Parity of the CUDA path (through the C ABI) with the oracle.  Bit-exact: integer/index work."""

import numpy as np
import pytest

from workloads import synth
from dstack_b200.encode import NO_ROW, encode_table, pack_jobs
from oracle import columnar

pytestmark = pytest.mark.gpu


def _check(matcher, cat, sj, k):
    table = encode_table(cat.columns)
    dt = matcher.upload(table)
    try:
        got = matcher.match_topk(dt, sj.columns, k)
        want = columnar.topk(cat.columns, sj.columns, k)
        assert np.array_equal(got.count, want.count)
        assert np.array_equal(got.index, want.index)
        assert np.array_equal(np.nan_to_num(got.max_price, nan=-1.0), np.nan_to_num(want.max_price, nan=-1.0))
        return got
    finally:
        dt.close()


@pytest.mark.parametrize("k", [15, 16, 50])
def test_topk_64_jobs_100k_offers(matcher, k):
    cat = synth.make_catalog(100_000, seed=3, n_shapes=2048)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, 64, seed=4)
    got = _check(matcher, cat, sj, k)
    assert (got.count > 0).sum() >= 16


@pytest.mark.parametrize("n_offers,n_jobs", [(1, 1), (31, 3), (32, 129), (33, 128), (1000, 257), (4097, 1)])
def test_topk_ragged_sizes(matcher, n_offers, n_jobs):
    cat = synth.make_catalog(n_offers, seed=21, n_shapes=64)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, n_jobs, seed=22)
    _check(matcher, cat, sj, 16)


@pytest.mark.parametrize("n_jobs", [1, 5, 130])
def test_small_batch_over_many_pieces(matcher, n_jobs):
    """A scheduler tick: a handful of jobs against a large table is cut into more than a thousand
    one-window slots, whose counts k_extract then sums with a whole CTA per job."""
    cat = synth.make_catalog(600_000, seed=31, n_shapes=1024)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, n_jobs, seed=32)
    dt = matcher.upload(table)
    try:
        assert matcher.plan(dt, pack_jobs(table, sj.columns), 50)["slots"] >= 1024
    finally:
        dt.close()
    got = _check(matcher, cat, sj, 50)
    if n_jobs >= 5:
        assert ((got.count > 0) & (got.count < 50)).any() or (got.count >= 50).any()


def test_large_k_stays_exact(matcher):
    """k = 2048 over 1024 jobs: the scratch of a call (counts per slot and window) does not grow
    with k, so the launch geometry is the same as for k = 16; results stay those of the oracle."""
    cat = synth.make_catalog(60_000, seed=41, n_shapes=256)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, 1024, seed=42)
    dt = matcher.upload(table)
    try:
        packed = pack_jobs(table, sj.columns)
        assert matcher.plan(dt, packed, 16) == matcher.plan(dt, packed, 2048)
    finally:
        dt.close()
    got = _check(matcher, cat, sj, 2048)
    assert (got.count > 2048).any()


def test_long_slots_with_many_windows(monkeypatch):
    """Few, long row pieces (what a 16k-job batch over tens of millions of offers gets): every
    slot holds more than 32 windows, so k_extract pages through a slot's window records.
    Checked like the benchmark does at full size: sampled jobs against the oracle, properties
    for every job."""
    import bench

    from dstack_b200.engine import Matcher

    monkeypatch.setenv("DSM_UNITS_PER_CTA", "1")  # tuning knobs are read when a context is created
    monkeypatch.setenv("DSM_FINE_PCT", "0")       # no short units at the end of the row axis: long slots only
    matcher = Matcher(0)
    cat = synth.make_catalog(2_000_000, seed=71, n_shapes=1024)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, 4096, seed=72)
    dt = matcher.upload(table)
    try:
        plan = matcher.plan(dt, pack_jobs(table, sj.columns), 16)
        blocks_per_warp = table.n_rows // 32 / (plan["pieces_a"] + plan["pieces_b"]) / 8
        assert blocks_per_warp > 32 * 15
        got = matcher.match_topk(dt, sj.columns, 16)
    finally:
        dt.close()
        matcher.close()
    n = bench.verify(cat.columns, sj, 16, got.index, got.count, got.max_price, np.random.default_rng(5), n_sample=12)
    assert n == 12 and (got.count > 16).any() and ((got.count > 0) & (got.count < 16)).any()


def test_dense_and_single_match_jobs(matcher):
    """The extremes of k_extract's candidate search in one batch: jobs without any filter (every
    live row matches: all 32 row lanes of every window, 480 candidates per window), jobs whose
    matches are a price prefix ending at a chosen row (the last-match search must land exactly on
    it: first rows, middle, end of the available segment), and jobs that ask for one instance
    type (a handful of rows scattered over 300k: one candidate lane per window)."""
    cat = synth.make_catalog(300_000, seed=51, n_shapes=256)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, 40, seed=52)
    jc = sj.columns
    nan2 = np.full((jc.n, 2), np.nan)
    jc.max_price[:] = np.nan
    jc.spot[:] = -1
    for name in ("cpu", "memory", "disk", "gpu_count", "gpu_memory", "gpu_total"):
        getattr(jc, name)[:] = nan2
    jc.gpu_vendor[:] = 0
    jc.cc_min[:] = -1
    jc.gpu_name_any[:] = True
    jc.src_backend_mask[:] = 0xFFFFFFFF
    jc.off_backend_mask[:] = 0xFFFFFFFF
    jc.exclude_not_available[:] = False
    jc.region_any[:] = True
    jc.itype_any[:] = True
    # jobs 8.. : a price window around one row of the result order (prices are not unique, so
    # take rows whose price occurs once among the available offers)
    price = np.asarray(cat.columns.price)
    order = columnar.rank_order(cat.columns)
    n_avail = int(np.asarray(cat.columns.available).sum())
    uniq, first, counts = np.unique(price[order[:n_avail]], return_index=True, return_counts=True)
    singles = np.sort(first[counts == 1])  # positions in the result order
    assert len(singles) > 16
    picks = [singles[0], singles[len(singles) // 3], singles[len(singles) // 2], singles[-2], singles[-1]]
    for j, pos in enumerate(picks, start=8):
        jc.max_price[j] = price[order[pos]]
        jc.exclude_not_available[j] = True
    # jobs 16..23: one instance type each (CSR over the table's instance-type ids)
    rng = np.random.default_rng(53)
    want_types = rng.choice(len(cat.columns.itypes), size=8, replace=False)
    counts_per_job = np.zeros(jc.n, np.int64)
    counts_per_job[16:24] = 1
    jc.itype_any[16:24] = False
    jc.itype_ptr = np.concatenate(([0], np.cumsum(counts_per_job))).astype(np.int64)
    jc.itype_ids = want_types.astype(np.int64)
    got = _check(matcher, cat, sj, 50)
    assert (got.count[:8] == len(price)).all()
    for j, pos in enumerate(picks, start=8):
        assert got.count[j] == pos + 1 and got.max_price[j] == price[order[pos]]


def test_random_shapes_against_oracle(matcher):
    """Forty random (offers, jobs, k, share of not-available rows) combinations: segment
    boundaries, partial last windows and slots of a block or two fall wherever they fall."""
    rng = np.random.default_rng(2024)
    for case in range(40):
        n_offers = int(rng.choice([1, 5, 31, 32, 33, 479, 480, 481, 960, 1500, 4096, 9999, 20011, 50000]))
        n_jobs = int(rng.choice([1, 2, 31, 33, 127, 128, 129, 300]))
        k = int(rng.choice([1, 2, 15, 16, 17, 50, 64, 200]))
        cat = synth.make_catalog(n_offers, seed=1000 + case, n_shapes=int(rng.choice([1, 8, 64, 300])))
        share = float(rng.choice([0.0, 0.1, 0.5, 1.0]))
        flip = rng.random(n_offers) < share
        cat.columns.available[:] = ~flip
        table = encode_table(cat.columns)
        sj = synth.make_jobs(cat, table, n_jobs, seed=2000 + case)
        if case % 3 == 0:
            sj.columns.exclude_not_available[:] = rng.random(n_jobs) < 0.5
        try:
            _check(matcher, cat, sj, k)
        except AssertionError as e:  # name the case that failed
            raise AssertionError(f"case {case}: offers={n_offers} jobs={n_jobs} k={k} not-available={share}") from e


def test_bitmap_equals_oracle_matrix(matcher):
    cat = synth.make_catalog(20_000, seed=8, n_shapes=512)
    table = encode_table(cat.columns)
    sj = synth.make_jobs(cat, table, 200, seed=9)
    dt = matcher.upload(table)
    try:
        rows = matcher.match_all(dt, sj.columns, order="source")
    finally:
        dt.close()
    want = columnar.match_matrix(cat.columns, sj.columns)
    for j in range(sj.columns.n):
        assert np.array_equal(rows[j], np.flatnonzero(want[j]))


def test_all_not_available_and_all_available(matcher):
    for avail in (0, 3):  # UNKNOWN (available) / NOT_AVAILABLE
        cat = synth.make_catalog(3000, seed=2, n_shapes=64)
        cat.availability[:] = avail
        cat.columns.available[:] = synth.AVAILABILITY[avail].is_available()
        table = encode_table(cat.columns)
        sj = synth.make_jobs(cat, table, 40, seed=5)
        sj.columns.exclude_not_available[::2] = True
        _check(matcher, cat, sj, 16)


def test_topk_on_a_pool_table_agrees_with_its_bitmap(matcher):
    """32-byte rows (pool tables, up to 16 code columns) through the count-mode kernel: first k,
    count and max price must be what the bitmap of the same call implies."""
    import test_configs as T
    from dstack_b200 import pools as PP
    from dstack_b200.ingest import job_columns

    rows = synth.make_pool(6000, seed=11, n_fleets=12)
    calls = T._config2_jobs(rows, 90)
    index = PP.PoolIndex(rows, matcher=matcher)
    try:
        assert index.table.n_cols == 16
        filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"],
                                   fleet_model=kw.get("fleet_model"), multinode=kw.get("multinode", False),
                                   master_job_provisioning_data=kw.get("master_job_provisioning_data"),
                                   volumes=None, shared=False) for p, kw in calls]
        jobs = job_columns(index.table, filters)
        top = matcher.match_topk(index.device_table, jobs, 16)
        every = matcher.match_all(index.device_table, jobs, order="rank")
    finally:
        index.close()
    price = np.array([PP.M.InstanceOffer.parse_raw(r.offer).price if r.offer is not None else np.nan for r in rows])
    assert sum(len(e) > 16 for e in every) > 10 and sum(0 < len(e) <= 16 for e in every) > 0
    for j, e in enumerate(every):
        assert top.count[j] == len(e)
        n = min(16, len(e))
        assert list(top.index[j][:n]) == list(e[:n]) and (top.index[j][n:] == -1).all()
        if len(e):
            assert top.max_price[j] == price[e].max()
        else:
            assert np.isnan(top.max_price[j])


def test_tables_that_fit_eight_times_only_without_a_staging_copy(matcher):
    """Dictionaries of 467..570 entries: eight shared-memory copies fit beside three CTAs; whether
    a ninth (the TMA staging copy) does too decides between the TMA path and the threads filling
    the copies straight from global memory.  The plan must say which, and results stay exact."""
    cat = synth.make_catalog(200_000, seed=4, n_shapes=4096)
    table = encode_table(cat.columns, pair_columns=False)
    sj = synth.make_jobs(cat, table, 300, seed=4)
    dt = matcher.upload(table)
    try:
        packed = pack_jobs(table, sj.columns)
        plan = matcher.plan(dt, packed, 16)
        got = matcher.topk_rows(dt, packed, 16)
    finally:
        dt.close()
    from dstack_b200.encode import TABLE_AREA_BUDGET

    e = plan["entries"]
    assert 466 < e <= 570
    staged = e * 144 <= TABLE_AREA_BUDGET  # room for the TMA staging copy beside the eight copies?
    assert plan["smem"] == e * (144 if staged else 128) + 1024
    want = columnar.topk(cat.columns, sj.columns, 16)
    assert np.array_equal(got[1].astype(np.int64), want.count)
    src = np.where(got[0] == NO_ROW, -1, table.perm[np.minimum(got[0], table.n_rows - 1)])
    assert np.array_equal(src, want.index)


def test_largest_column_with_fewer_copies(matcher):
    """cpu and memory paired into one column push the tables past what fits eight times beside
    three CTAs; the library then keeps the largest column (regions) four times and the rest
    eight times.  The plan must show that layout, and results stay exact."""
    from dstack_b200.encode import table_copies

    cat = synth.make_catalog(200_000, seed=4, n_shapes=4096)
    table = encode_table(cat.columns)
    assert table.pair_entries is not None
    sj = synth.make_jobs(cat, table, 300, seed=4)
    packed = pack_jobs(table, sj.columns)
    cards = packed.layout["card"]
    copies = table_copies(cards)
    assert min(copies) == 2 and copies.count(2) == 1 and cards[copies.index(2)] == max(cards)
    dt = matcher.upload(table)
    try:
        plan = matcher.plan(dt, packed, 16)
    finally:
        dt.close()
    pad = lambda n: (n + 127) // 128 * 128
    assert plan["entries"] > 570
    from dstack_b200.encode import TABLE_AREA_BUDGET

    area = sum(pad((c * 16) << s) for c, s in zip(cards, copies))
    staging = plan["entries"] * 16 if area + plan["entries"] * 16 <= TABLE_AREA_BUDGET else 0  # TMA staging copy
    assert plan["smem"] == area + staging + 1024
    _check(matcher, cat, sj, 16)




def test_every_layout_of_one_batch_gives_the_same_answer(matcher):
    """One catalog, one batch of jobs, four layouts: cpu and memory as one paired column or as two
    range columns, jobs evaluated in the grouped order (whole groups skip trailing columns) or as
    given.  All four must return the oracle's rows, counts and max-price rows."""
    cat = synth.make_catalog(60_000, seed=95, n_shapes=1024)
    sj = synth.make_jobs(cat, encode_table(cat.columns), 700, seed=96)
    want = columnar.topk(cat.columns, sj.columns, 16)
    seen = []
    for pair in (None, False):
        table = encode_table(cat.columns, pair_columns=pair)
        assert (table.pair_entries is not None) == (pair is None)
        dt = matcher.upload(table)
        try:
            for ordered in (True, False):
                packed = pack_jobs(table, sj.columns)
                assert packed.order is not None
                if not ordered:
                    packed.order = None
                idx, cnt, mx = matcher.topk_rows(dt, packed, 16)
                src = np.where(idx == NO_ROW, -1, table.perm[np.minimum(idx, table.n_rows - 1)]).astype(np.int64)
                price = np.where(mx == NO_ROW, np.nan, table.price[np.minimum(mx, table.n_rows - 1)])
                assert np.array_equal(src, want.index), (pair, ordered)
                assert np.array_equal(cnt.astype(np.int64), want.count), (pair, ordered)
                assert np.array_equal(np.nan_to_num(price, nan=-1.0), np.nan_to_num(want.max_price, nan=-1.0))
                seen.append((src, cnt.copy()))
        finally:
            dt.close()
    assert len(seen) == 4 and all(np.array_equal(seen[0][0], s) and np.array_equal(seen[0][1], c) for s, c in seen)
