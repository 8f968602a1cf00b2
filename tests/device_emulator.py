"""numpy emulation of what libdsmatch's kernels compute, for CPU-only tests of the host encoder.

TEST INFRASTRUCTURE: lives under tests/, is never imported by ``dstack_b200`` and is not a
fallback.  It re-implements the record/table formats documented in ``include/dsmatch.h``
so that ``encode.py`` / ``ingest.py`` can be checked against the oracle without a GPU; the
``-m gpu`` tests then check the real kernels against the same oracle.
"""

from __future__ import annotations

import numpy as np

from dstack_b200.encode import (COLF_NULL_PASSES, COLF_PAIR_A_BITSET, GPUF_CC_BOUNDED, GPUF_MEM_BOUNDED,
                                GPUF_NAME_ANY, KIND_BITSET, KIND_BK, KIND_GPU, KIND_PAIR, KIND_RANGE, NO_ROW,
                                EncodedTable, PackedJobs)


def acceptance(table: EncodedTable, packed: PackedJobs):
    """Per used column: bool [J, card] -- bit (j, e) of the acceptance tables."""
    rec = packed.records.astype(np.uint64)
    lay = packed.layout
    out = []
    for c in range(lay["n_used"]):
        kind, card, off = lay["kind"][c], lay["card"][c], lay["rec_off"][c]
        e = np.arange(card, dtype=np.uint64)[None, :]
        if kind == KIND_RANGE:
            w = rec[:, off][:, None]
            ok = (e >= (w & 0xFFFF)) & (e <= (w >> 16))
            if lay["flags"][c] & COLF_NULL_PASSES:
                ok[:, card - 1] = True
        elif kind == KIND_BITSET:
            nw = (card + 31) // 32
            words = rec[:, off:off + nw]
            ok = ((words[:, (e[0] >> 5).astype(np.int64)] >> (e & 31)) & 1).astype(bool)
        elif kind == KIND_PAIR:
            ent = table.pair_entries.astype(np.uint64)
            a, b = (ent & 0xFFFF)[None, :], (ent >> 16)[None, :]
            aw = lay["aux"][c]
            if lay["flags"][c] & COLF_PAIR_A_BITSET:
                words = rec[:, off:off + aw]
                ai = a[0].astype(np.int64)
                ok = ((words[:, ai >> 5] >> (a & 31)) & 1).astype(bool)
            else:
                w = rec[:, off][:, None]
                ok = (a >= (w & 0xFFFF)) & (a <= (w >> 16))
            wb = rec[:, off + aw][:, None]
            ok = ok & (b >= (wb & 0xFFFF)) & (b <= (wb >> 16))
        elif kind == KIND_GPU:
            ent = table.gpu_entries.astype(np.uint64)
            cnt, mem = (ent[:, 0] & 0xFFFF)[None, :], (ent[:, 0] >> 16)[None, :]
            tot, cc = (ent[:, 1] & 0xFFFF)[None, :], (ent[:, 1] >> 16)[None, :]
            name, vendor = (ent[:, 2] & 0xFFFF)[None, :], ((ent[:, 2] >> 16) & 0xFF)[None, :]
            has_gpu = (((ent[:, 2] >> 24) & 1) == 1)[None, :]
            r0, r1, r2, r3, r4 = (rec[:, off + i][:, None] for i in range(5))
            gflags, cc_min = r3 >> 16, r3 & 0xFFFF
            ok = (cnt >= (r0 & 0xFFFF)) & (cnt <= (r0 >> 16))
            ok &= ((r4 >> vendor) & 1).astype(bool)
            nwn = lay["aux"][c]
            bits = rec[:, off + 5:off + 5 + nwn]
            safe = np.where(name == 0xFFFF, 0, name).astype(np.int64)
            name_ok = ((bits[:, (safe[0] >> 5)] >> (safe & 31).astype(np.uint64)) & 1).astype(bool) & (name != 0xFFFF)
            ok &= np.where((gflags & GPUF_NAME_ANY) != 0, True, name_ok)
            mem_ok = has_gpu & (mem >= (r1 & 0xFFFF)) & (mem <= (r1 >> 16))
            ok &= np.where((gflags & GPUF_MEM_BOUNDED) != 0, mem_ok, True)
            cc_ok = (cc != 0xFFFF) & (cc >= cc_min)
            ok &= np.where((gflags & GPUF_CC_BOUNDED) != 0, cc_ok, True)
            ok &= (tot >= (r2 & 0xFFFF)) & (tot <= (r2 >> 16))
            blocks, eff = (r4 >> 16).astype(np.int64), (ent[:, 2] >> 25).astype(np.int64)[None, :]
            ok &= np.where(blocks > 0, eff % np.maximum(blocks, 1) == 0, True)
        elif kind == KIND_BK:
            ent = table.bk_entries.astype(np.uint64)
            src, offb = (ent[:, 0] & 0xFF)[None, :], ((ent[:, 0] >> 8) & 0xFF)[None, :]
            spot, status = ((ent[:, 0] >> 16) & 0xFF)[None, :], (ent[:, 0] >> 24)[None, :]
            fl = ent[:, 1][None, :]
            r0, r1, r2, r3 = (rec[:, off + i][:, None] for i in range(4))
            ok = np.ones((len(rec), card), bool)
            ok[:, 0] = False
            ok &= ((r0 >> src) & 1).astype(bool) & ((r1 >> offb) & 1).astype(bool)
            ok &= ((r2 >> spot) & 1).astype(bool) & ((r2 >> (8 + status)) & 1).astype(bool)
            req, forbid = r3 & 0xFFFF, r3 >> 16
            ok &= ((fl & req) == req) & ((fl & forbid) == 0)
        else:
            raise ValueError(kind)
        out.append(ok)
    return out


def match_matrix(table: EncodedTable, packed: PackedJobs) -> np.ndarray:
    """bool [J, n_rows]: what k_match's bitmap mode produces."""
    acc = acceptance(table, packed)
    j = len(packed.records)
    rows = np.arange(table.n_rows)
    # padding rows never pass, whatever the cut-offs say (k_match clamps them to the live rows)
    cut_a = np.minimum(packed.records[:, 0].astype(np.int64), table.n_avail)[:, None]
    cut_b = np.minimum(packed.records[:, 1].astype(np.int64), table.split_row + table.n_rest)[:, None]
    m = np.where(rows[None, :] < table.split_row, rows[None, :] < cut_a, rows[None, :] < cut_b)
    for c, ok in enumerate(acc):
        codes = table.rows[:, c].astype(np.int64)
        if c in packed.luts:
            codes = packed.luts[c][codes].astype(np.int64)
        m &= ok[:, codes]
    return m


def topk(table: EncodedTable, packed: PackedJobs, k: int):
    m = match_matrix(table, packed)
    j = len(m)
    idx = np.full((j, k), NO_ROW, dtype=np.uint32)
    cnt = m.sum(axis=1).astype(np.uint32)
    mx = np.full(j, NO_ROW, dtype=np.uint32)
    for i in range(j):
        rows = np.flatnonzero(m[i])
        idx[i, :min(k, len(rows))] = rows[:k]
        if len(rows):
            a = rows[rows < table.split_row]
            b = rows[rows >= table.split_row]
            best = a[-1] if len(a) else b[-1]
            if len(a) and len(b) and table.price[b[-1]] > table.price[a[-1]]:
                best = b[-1]
            mx[i] = best
    return idx, cnt, mx


def merge_shards(idx_all, cnt_all, max_all, maxprice_all, bases, k):
    """What k_merge_shards computes: shards are in table order, so the first k matches are the
    concatenation of the per-shard first-k lists; counts add up; the max row is the one with the
    highest price (earlier shard wins ties).  Rows come out as global physical rows, -1 = none."""
    n_shards, j = cnt_all.shape
    idx_out = np.full((j, k), -1, dtype=np.int64)
    cnt_out = np.zeros(j, dtype=np.int64)
    max_out = np.full(j, -1, dtype=np.int64)
    mp_out = np.full(j, np.nan)
    for job in range(j):
        have = 0
        for s in range(n_shards):
            c = int(cnt_all[s, job])
            if c == 0:
                continue
            take = min(c, k, k - have)
            idx_out[job, have:have + take] = bases[s] + idx_all[s, job, :take].astype(np.int64)
            have += take
            cnt_out[job] += c
            m = int(max_all[s, job])
            if m != NO_ROW:
                pr = float(maxprice_all[s, job])
                if max_out[job] < 0 or pr > mp_out[job]:
                    max_out[job], mp_out[job] = bases[s] + m, pr
    return idx_out, cnt_out, max_out, mp_out


def merge_shards_interleaved(idx_all, cnt_all, max_all, maxprice_all, geom, chunk_rows, k):
    """What k_merge_shards_interleaved computes: shard s was dealt chunks s, s+G, ... of the
    result order; geom[s] = (split_row, n_avail) turns its physical rows into global positions;
    the k smallest positions over all shards are kept.  Positions out, -1 = none."""
    from dstack_b200.sharded import local_rows_to_positions

    n_shards, j = cnt_all.shape
    idx_out = np.full((j, k), -1, dtype=np.int64)
    cnt_out = cnt_all.astype(np.int64).sum(axis=0)
    max_out = np.full(j, -1, dtype=np.int64)
    mp_out = np.full(j, np.nan)
    pos = np.stack([local_rows_to_positions(idx_all[s], int(geom[s][0]), int(geom[s][1]), s, n_shards, chunk_rows)
                    for s in range(n_shards)])  # [G, J, k]
    for job in range(j):
        cand = pos[:, job, :].reshape(-1)
        cand = np.sort(cand[cand >= 0])[:k]
        idx_out[job, :len(cand)] = cand
        for s in range(n_shards):
            m = int(max_all[s, job])
            if m != NO_ROW and m >= 0:
                pr = float(maxprice_all[s, job])
                if max_out[job] < 0 or pr > mp_out[job]:
                    max_out[job] = local_rows_to_positions(np.array([m]), int(geom[s][0]), int(geom[s][1]), s,
                                                           n_shards, chunk_rows)[0]
                    mp_out[job] = pr
    return idx_out, cnt_out, max_out, mp_out
