"""The shard-merge kernels on one GPU: G shard tables of one catalog live side by side on the
device, each is matched with ``dsm_match_topk_dev``, the per-shard results are laid out as the
all-gathered record ``ShardedMatcher`` uses, and ``dsm_merge_shards_strided_dev`` /
``dsm_merge_shards_interleaved_dev`` must reproduce the oracle's answer on the unsharded catalog
(bit-exact: indices, counts, max price)."""

import ctypes as C

import numpy as np
import pytest

from dstack_b200 import native, sharded
from workloads import synth
from dstack_b200.encode import encode_table, pack_jobs
from oracle import columnar

pytestmark = pytest.mark.gpu


def _run(matcher, n_offers, n_jobs, k, world, chunk_rows, seed):
    import torch

    lib = native.lib()
    dev = torch.device("cuda", 0)
    cat = synth.make_catalog(n_offers, seed=seed, n_shapes=128)
    full = cat.columns
    order = sharded.rank_order(full)
    tables = [encode_table(sharded.subset_columns(full, sharded.shard_positions(order, r, world, chunk_rows)))
              for r in range(world)]
    sj = synth.make_jobs(cat, tables[0], n_jobs, seed=seed + 1)
    j = n_jobs
    w_idx, w_cnt = j * k, j
    w_head = (w_idx + 2 * w_cnt + 1) // 2 * 2
    words = w_head + 2 * j
    g = torch.zeros((world, words), dtype=torch.int32, device=dev)
    dts = []
    h = matcher.handle
    try:
        for r, t in enumerate(tables):
            dt = matcher.upload(t)
            dts.append(dt)
            packed = pack_jobs(t, sj.columns)
            rec = torch.from_numpy(np.ascontiguousarray(packed.records).view(np.int32)).to(dev)
            lay = matcher._layout(packed)
            base = g[r].data_ptr()
            native.check(lib.dsm_match_topk_dev(h, dt.handle, C.byref(lay), rec.data_ptr(), j, k,
                                                base, base + 4 * w_idx, base + 4 * (w_idx + w_cnt)))
            native.check(lib.dsm_gather_price_dev(h, dt.handle, base + 4 * (w_idx + w_cnt), j, base + 4 * w_head))
            matcher.sync()
        o_idx = torch.empty((j, k), dtype=torch.int64, device=dev)
        o_cnt = torch.empty((j,), dtype=torch.int64, device=dev)
        o_max = torch.empty((j,), dtype=torch.int64, device=dev)
        o_mp = torch.empty((j,), dtype=torch.float64, device=dev)
        p0 = g.data_ptr()
        fields = (p0, p0 + 4 * w_idx, p0 + 4 * (w_idx + w_cnt), p0 + 4 * w_head, 4 * words)
        outs = (o_idx.data_ptr(), o_cnt.data_ptr(), o_max.data_ptr(), o_mp.data_ptr())
        if chunk_rows:
            geom = torch.tensor([[t.split_row, t.n_avail] for t in tables], dtype=torch.int32, device=dev)
            native.check(lib.dsm_merge_shards_interleaved_dev(h, world, j, k, *fields, geom.data_ptr(), chunk_rows,
                                                              *outs))
            matcher.sync()
            pos = o_idx.cpu().numpy()
            src = np.where(pos < 0, -1, order[np.clip(pos, 0, n_offers - 1)])
        else:
            bases = sharded.shard_bases([t.n_rows for t in tables])
            d_base = torch.from_numpy(bases.astype(np.int64)).to(dev)
            native.check(lib.dsm_merge_shards_strided_dev(h, world, j, k, *fields, d_base.data_ptr(), *outs))
            matcher.sync()
            src = sharded.global_rows_to_source(
                o_idx.cpu().numpy(), bases, [t.perm for t in tables],
                [sharded.shard_positions(order, r, world, 0) for r in range(world)])
        want = columnar.topk(full, sj.columns, k, order=order)
        assert np.array_equal(o_cnt.cpu().numpy(), want.count)
        assert np.array_equal(src, want.index)
        assert np.array_equal(np.nan_to_num(o_mp.cpu().numpy(), nan=-1.0), np.nan_to_num(want.max_price, nan=-1.0))
        return want
    finally:
        for dt in dts:
            dt.close()


@pytest.mark.parametrize("world,chunk_rows", [(2, 0), (3, 0), (2, 64), (4, 1024), (8, 37), (5, 1)])
def test_merge_kernels_equal_unsharded_oracle(matcher, world, chunk_rows):
    want = _run(matcher, 20_000, 150, 16, world, chunk_rows, seed=61)
    assert (want.count > 16).any() and (want.count < 16).any()


def test_interleaved_merge_with_not_available_segment_and_k50(matcher):
    _run(matcher, 5_000, 40, 50, 4, 128, seed=67)


@pytest.mark.parametrize("world,k", [(8, 64), (3, 200), (32, 17)])
def test_interleaved_merge_many_shards_or_long_lists(matcher, world, k):
    _run(matcher, 6_000, 70, k, world, 16, seed=71)


def test_interleaved_merge_rejects_bad_arguments(matcher):
    lib = native.lib()
    h = matcher.handle
    one = C.c_void_p(16)
    assert lib.dsm_merge_shards_interleaved_dev(h, 33, 1, 1, one, one, one, one, 8, one, 1, one, one, one, one) == -2
    assert lib.dsm_merge_shards_interleaved_dev(h, 2, 1, 1, one, one, one, one, 8, one, 0, one, one, one, one) == -2
    assert lib.dsm_merge_shards_interleaved_dev(h, 2, 1, 256, one, one, one, one, 8, one, 4, one, one, one, one) == -2
    assert lib.dsm_merge_shards_interleaved_dev(h, 2, 1, 1, one, one, one, one, 12, one, 4, one, one, one, one) == -2
    assert b"" != lib.dsm_last_error()
