"""Host encoder (dictionaries, ordinals, records, cut-offs) checked on the CPU by running the
documented table/record semantics in numpy (tests/device_emulator.py) against the oracle."""

import numpy as np
import pytest

import device_emulator as emu
from workloads import synth
from dstack_b200.encode import NO_ROW, encode_table, pack_jobs
from dstack_b200.ingest import columns_from_offers, job_columns
from dstack_b200.services import job_filter
from oracle import columnar


def _to_source(table, idx):
    return np.where(idx == NO_ROW, -1, table.perm[np.minimum(idx, table.n_rows - 1)])


@pytest.mark.parametrize("n_offers,n_jobs,k", [(5000, 96, 16), (777, 33, 50), (31, 5, 15)])
def test_encoded_batch_equals_oracle(n_offers, n_jobs, k):
    cat = synth.make_catalog(n_offers, seed=5, n_shapes=256)
    table = encode_table(cat.columns)
    assert table.n_rows % 32 == 0 and table.split_row % 32 == 0
    sj = synth.make_jobs(cat, table, n_jobs, seed=6)
    packed = pack_jobs(table, sj.columns)
    idx, cnt, mx = emu.topk(table, packed, k)
    want = columnar.topk(cat.columns, sj.columns, k)
    assert np.array_equal(cnt.astype(np.int64), want.count)
    assert np.array_equal(_to_source(table, idx), want.index)
    got_price = np.where(mx == NO_ROW, np.nan, table.price[np.minimum(mx, table.n_rows - 1)])
    assert np.array_equal(np.nan_to_num(got_price, nan=-1), np.nan_to_num(want.max_price, nan=-1))


def test_object_ingest_equals_columnar_ingest():
    cat = synth.make_catalog(1200, seed=9, n_shapes=128)
    per_backend = synth.catalog_objects(cat)
    offers = [o for _, lst in per_backend for o in lst]
    src = [b.TYPE for b, lst in per_backend for _ in lst]
    t_col = encode_table(cat.columns)
    t_obj = encode_table(columns_from_offers(offers, src))
    sj = synth.make_jobs(cat, t_col, 40, seed=10)
    filters = [job_filter(p, r) for p, r in synth.job_objects(sj)]
    a = emu.topk(t_col, pack_jobs(t_col, sj.columns), 16)
    b = emu.topk(t_obj, pack_jobs(t_obj, job_columns(t_obj, filters)), 16)
    assert np.array_equal(a[1], b[1])
    assert np.array_equal(_to_source(t_col, a[0]), _to_source(t_obj, b[0]))


def test_unsorted_backend_list_is_rejected():
    cat = synth.make_catalog(200, seed=3, n_shapes=32)
    cat.columns.price[:2] = cat.columns.price[:2][::-1] + np.array([1.0, 0.0])
    with pytest.raises(ValueError, match="not sorted by price"):
        encode_table(cat.columns)


@pytest.mark.parametrize("blocks", [2, 4, 3, 8, 0, "auto"])
def test_block_counts_equal_reference(blocks):
    """`blocks` of get_offers_by_requirements (server/services/offers.py:102-118): explicit counts
    become a cpu bitset + a GPU-count divisibility test, "auto" a per-row flag."""
    from dstack_b200 import models as M
    from oracle import reference_path as ref

    cat = synth.make_catalog(1500, seed=23, n_shapes=160)
    per_backend = synth.catalog_objects(cat, filter_fn=ref.match_requirements)
    backends = [b for b, _ in per_backend]
    offers = [o for _, lst in per_backend for o in lst]
    src = [b.TYPE for b, lst in per_backend for _ in lst]
    table = encode_table(columns_from_offers(offers, src))
    reqs = [M.Requirements(resources=M.ResourcesSpec(), max_price=None, spot=None),
            M.Requirements(resources=M.ResourcesSpec(gpu="1..8"), max_price=None, spot=None),
            M.Requirements(resources=M.ResourcesSpec(gpu="tpu:0..8", cpu="1.."), max_price=None, spot=None)]
    filters = [job_filter(M.Profile(name="p"), r, blocks=blocks) for r in reqs]
    m = emu.match_matrix(table, pack_jobs(table, job_columns(table, filters)))
    pos = {id(o): i for i, o in enumerate(offers)}
    seen = 0
    for j, r in enumerate(reqs):
        want = ref.get_offers_by_requirements(backends, M.Profile(name="p"), r, blocks=blocks)
        got = [int(table.perm[row]) for row in np.flatnonzero(m[j])]
        assert got == [pos[id(o)] for _, o in want], (blocks, j)
        seen += len(want)
    assert (seen > 0) == (blocks != 0)
