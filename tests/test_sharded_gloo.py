"""Host-side logic of the sharded catalog under a real 2-process `gloo` group (CPU).

The device steps (dsm_match_topk_dev, dsm_merge_shards_dev) are replaced by their numpy
emulation from tests/device_emulator.py; what is exercised for real is everything around them:
the split of the result order into contiguous or chunk-interleaved shards, per-shard tables with their own
dictionaries, job packing per shard, the all-gather, shard bases and the mapping of global rows
back to source positions.  The merged answer must equal the oracle's on the unsharded catalog.
"""

import os
import socket

import numpy as np
import pytest


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_offers, n_jobs, k, chunk_rows, out):
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    import device_emulator as emu
    from dstack_b200 import sharded
    from workloads import synth
    from dstack_b200.encode import encode_table, pack_jobs
    from oracle import columnar

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cat = synth.make_catalog(n_offers, seed=17, n_shapes=128)
        full = cat.columns
        order = sharded.rank_order(full)
        mine = sharded.shard_positions(order, rank, world, chunk_rows)
        table = encode_table(sharded.subset_columns(full, mine))
        sj = synth.make_jobs(cat, table, n_jobs, seed=18)      # ids are valid for every shard
        idx, cnt, mx = emu.topk(table, pack_jobs(table, sj.columns), k)
        mp = np.where(mx == emu.NO_ROW, np.nan, table.price[np.minimum(mx, table.n_rows - 1)])

        def gather(a):
            t = torch.from_numpy(np.ascontiguousarray(a))
            parts = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(parts, t)
            return np.stack([p.numpy() for p in parts])

        g_idx, g_cnt, g_max, g_mp = gather(idx.astype(np.int64)), gather(cnt.astype(np.int64)), \
            gather(mx.astype(np.int64)), gather(mp)
        sizes = gather(np.array([table.n_rows], dtype=np.int64)).reshape(-1)
        bases = sharded.shard_bases(sizes)
        meta = [None] * world
        dist.all_gather_object(meta, (mine, table.perm))
        if chunk_rows:
            geom = gather(np.array([table.split_row, table.n_avail], dtype=np.int64))
            o_idx, o_cnt, o_max, o_mp = emu.merge_shards_interleaved(g_idx, g_cnt, g_max, g_mp, geom, chunk_rows, k)
            src = np.where(o_idx < 0, -1, order[np.maximum(o_idx, 0)])
        else:
            o_idx, o_cnt, o_max, o_mp = emu.merge_shards(g_idx, g_cnt, g_max, g_mp, bases, k)
            src = sharded.global_rows_to_source(o_idx, bases, [m[1] for m in meta], [m[0] for m in meta])

        want = columnar.topk(full, sj.columns, k, order=order)
        assert np.array_equal(o_cnt, want.count)
        assert np.array_equal(src, want.index)
        assert np.array_equal(np.nan_to_num(o_mp, nan=-1.0), np.nan_to_num(want.max_price, nan=-1.0))
        # shards partition the catalog and are in result order
        assert sum(len(m[0]) for m in meta) == n_offers
        rank_of = np.empty(n_offers, np.int64)
        rank_of[order] = np.arange(n_offers)
        if chunk_rows:  # chunk c of the result order belongs to rank c % world
            assert ((rank_of[mine] // chunk_rows) % world == rank).all()
        elif rank + 1 < world:
            assert rank_of[meta[rank][0]].max() < rank_of[meta[rank + 1][0]].min()
        out[rank] = int(want.count.sum())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_offers,n_jobs,k,chunk_rows", [(3000, 40, 16, 0), (257, 9, 50, 0),
                                                          (3000, 40, 16, 64), (1000, 9, 50, 7), (257, 5, 16, 1024)])
def test_two_rank_sharded_merge_equals_unsharded_oracle(n_offers, n_jobs, k, chunk_rows):
    import torch.multiprocessing as mp

    world = 2
    port = _free_port()
    with mp.Manager() as manager:
        out = manager.dict()
        mp.spawn(_worker, args=(world, port, n_offers, n_jobs, k, chunk_rows, out), nprocs=world, join=True)
        assert len(out) == world and out[0] == out[1] and out[0] > 0
