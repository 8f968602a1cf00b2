"""Catalog sharded over the GPUs of one node: one process per GPU, one exchange per step.

Rank r holds a contiguous range of the global result order (stable sort of the source list by
``(not available, price)``, the order ``server/services/backends/__init__.py:449-451``
produces) as its own device table with its own dictionaries; job records are replicated.  A
step is: local ``dsm_match_topk_dev`` -> ``dsm_gather_price_dev`` -> ONE all-gather of a per-shard
record (first-k rows, counts, max rows and their prices) -> a merge kernel.  The all-gather is the
library's own peer exchange (``dsm_xchg_*``: every rank stores its record straight into its peers'
IPC-mapped buffers over NVLink and raises a flag; ``exchange="p2p"``) with NCCL's
``all_gather_into_tensor`` as the fallback when the ranks cannot map each other (``"nccl"``);
``torch.distributed`` carries the set-up (handles, sizes) either way.  By default the result
order is dealt out to the ranks in chunks of 1024 positions (``dsm_merge_shards_interleaved_dev`` keeps
the k smallest global positions), which balances the ranks; with ``chunk_rows=0`` every rank holds
one contiguous range, the merge is a concatenation (``dsm_merge_shards_strided_dev``) and the total
order is the *global physical row* ``shard_base[s] + local row``.

The helpers at the top are pure numpy (used by the CPU ``gloo`` tests as well); the device path
is :class:`ShardedMatcher`.
"""

from __future__ import annotations

import ctypes as C
import dataclasses
from typing import Any, List, Optional, Sequence, Tuple

import numpy as np

from .encode import NO_ROW, EncodedTable, JobColumns, OfferColumns, PackedJobs, encode_table, pack_jobs


def rank_order(oc: OfferColumns) -> np.ndarray:
    """Source positions in result order."""
    return np.lexsort((np.asarray(oc.price, dtype=np.float64), ~np.asarray(oc.available, dtype=bool)))


def shard_positions(order: np.ndarray, rank: int, world: int, chunk_rows: int = 0) -> np.ndarray:
    """Source positions (ascending) of the offers rank ``rank`` owns.

    ``chunk_rows == 0``: the ``rank``-th of ``world`` equal contiguous ranges of the result order.
    ``chunk_rows > 0``: the result order is dealt out in chunks of that many positions, chunk ``c``
    to rank ``c % world``, so every shard sees the same mix of cheap and expensive rows and the
    ranks finish together (a contiguous split left one rank of four 7 % behind the others)."""
    n = len(order)
    if chunk_rows > 0:
        p = np.arange(n)
        return np.sort(order[(p // chunk_rows) % world == rank])
    lo, hi = rank * n // world, (rank + 1) * n // world
    return np.sort(order[lo:hi])


def local_rows_to_positions(rows: np.ndarray, split_row: int, n_avail: int, rank: int, world: int,
                            chunk_rows: int) -> np.ndarray:
    """Physical rows of an interleaved shard's table -> positions in the global result order
    (``NO_ROW`` / negative stay "none" = -1).  A shard's table keeps its rows in result order with
    the first segment padded to ``split_row``; its ``q``-th row is the global position
    ``((q // chunk) * world + rank) * chunk + q % chunk``."""
    r = np.asarray(rows).astype(np.int64)
    none = (r < 0) | (r == NO_ROW)
    q = np.where(r < split_row, r, n_avail + (r - split_row))
    p = ((q // chunk_rows) * world + rank) * chunk_rows + q % chunk_rows
    return np.where(none, -1, p)


def subset_columns(oc: OfferColumns, idx: np.ndarray) -> OfferColumns:
    """Rows ``idx`` of ``oc``; dictionaries (shapes, regions, instance types) stay shared, so job
    filters packed against one shard's id space are valid for every shard."""
    return dataclasses.replace(
        oc, shape_id=oc.shape_id[idx], price=oc.price[idx], spot=oc.spot[idx], available=oc.available[idx],
        backend=oc.backend[idx], src_backend=oc.src_backend[idx], region_id=oc.region_id[idx],
        itype_id=oc.itype_id[idx])


def shard_bases(n_rows_per_shard: Sequence[int]) -> np.ndarray:
    sizes = np.asarray(n_rows_per_shard, dtype=np.int64)
    return np.cumsum(sizes) - sizes


def global_rows_to_source(global_rows: np.ndarray, bases: np.ndarray, perms: Sequence[np.ndarray],
                          positions: Sequence[np.ndarray]) -> np.ndarray:
    """Global physical rows (``shard_base + local row``, ``-1``/``2^64-1`` = none) -> positions in
    the full source list, through the owning shard's ``perm`` and its source positions."""
    g = np.asarray(global_rows).astype(np.int64)
    out = np.full(g.shape, -1, dtype=np.int64)
    for s, (perm, pos) in enumerate(zip(perms, positions)):
        sel = (g >= bases[s]) & (g < bases[s] + len(perm))
        local = perm[g[sel] - bases[s]]
        out[sel] = np.asarray(pos)[local]
    return out


class ShardedMatcher:
    """Device path.  ``torch`` supplies device memory, the stream and the process group only."""

    def __init__(self, full: OfferColumns, *, rank: int, world: int, device: int, order: Optional[np.ndarray] = None,
                 chunk_rows: int = 1024, exchange: str = "p2p"):
        """``chunk_rows > 0`` deals the result order out in chunks (balanced shards, the default);
        ``0`` gives every rank one contiguous range.  ``exchange``: ``"p2p"`` (peer stores over
        NVLink, NCCL if the ranks cannot map each other's memory) or ``"nccl"``; what a batch
        shape ended up with is ``self.exchange_used`` after :meth:`prepare`."""
        import torch
        import torch.distributed as dist

        from . import native
        from .engine import Matcher

        self.torch, self.dist, self.native = torch, dist, native
        self.rank, self.world = rank, world
        self.dev = torch.device("cuda", device)
        self.order = rank_order(full) if order is None else order
        self.full = full
        self.chunk_rows = chunk_rows
        if exchange not in ("p2p", "nccl"):
            raise ValueError(f"exchange must be 'p2p' or 'nccl', not {exchange!r}")
        self.exchange = exchange
        self.exchange_used = "none" if world == 1 else "nccl"
        self.exchange_note = ""
        self._xchg = None
        self.positions = shard_positions(self.order, rank, world, chunk_rows) if world > 1 else None
        cols = subset_columns(full, self.positions) if world > 1 else full
        self.table: EncodedTable = encode_table(cols)
        self.stream = torch.cuda.Stream(device=self.dev)
        with torch.cuda.stream(self.stream):
            self.matcher = Matcher(device, stream=self.stream.cuda_stream)
            self.device_table = self.matcher.upload(self.table)
            if world > 1:
                sizes = torch.tensor([self.table.n_rows], dtype=torch.int64, device=self.dev)
                all_sizes = torch.empty((world,), dtype=torch.int64, device=self.dev)
                dist.all_gather_into_tensor(all_sizes, sizes)
                self.base = torch.cumsum(all_sizes, 0) - all_sizes
                geom = torch.tensor([self.table.split_row, self.table.n_avail], dtype=torch.int32, device=self.dev)
                self.geom = torch.empty((world, 2), dtype=torch.int32, device=self.dev)
                dist.all_gather_into_tensor(self.geom, geom)
        self._bufs = None

    def pack(self, jobs: JobColumns) -> PackedJobs:
        return pack_jobs(self.table, jobs)

    def prepare(self, packed: PackedJobs, k: int):
        """Allocate the device buffers of a batch shape once (records stay resident).  The cached
        layout points at this matcher's per-context remap buffers (columns with more than 1024
        entries): ``self.matcher`` is private to this object and must not serve other calls while
        a batch is prepared -- a later ``_layout`` on the same column would replace the remap."""
        torch = self.torch
        j = len(packed.records)
        dev = self.dev
        b = dict(j=j, k=k)
        b["rec"] = torch.from_numpy(np.ascontiguousarray(packed.records).view(np.int32)).to(dev)
        # the evaluation order lives where the records live: on the device
        b["order"] = None if packed.order is None else \
            torch.from_numpy(np.ascontiguousarray(packed.order, dtype=np.uint32).view(np.int32)).to(dev)
        b["lay"] = self.matcher._layout(packed, None if b["order"] is None else b["order"].data_ptr())
        # one record per shard: [idx J*k | cnt J | max J | (pad) | max price J f64], so that a
        # single all_gather moves everything
        w_idx, w_cnt = j * k, j
        w_head = (w_idx + 2 * w_cnt + 1) // 2 * 2          # keep the f64 part 8-byte aligned
        words = w_head + 2 * j
        b["words"] = words
        rec_out = torch.empty(((words + 3) // 4 * 4,), dtype=torch.int32, device=dev)  # whole 16-byte units
        b["out"] = rec_out
        b["idx"] = rec_out[:w_idx].view(j, k)
        b["cnt"] = rec_out[w_idx:w_idx + w_cnt]
        b["max"] = rec_out[w_idx + w_cnt:w_idx + 2 * w_cnt]
        b["mp"] = rec_out[w_head:w_head + 2 * j].view(torch.float64)
        if self.world > 1:
            w = self.world
            b["g_out"] = torch.empty((w, rec_out.numel()), dtype=torch.int32, device=dev)
            b["o_idx"] = torch.empty((j, k), dtype=torch.int64, device=dev)
            b["o_cnt"] = torch.empty((j,), dtype=torch.int64, device=dev)
            b["o_max"] = torch.empty((j,), dtype=torch.int64, device=dev)
            b["o_mp"] = torch.empty((j,), dtype=torch.float64, device=dev)
            self._close_exchange()
            if self.exchange == "p2p":
                self._open_exchange(4 * rec_out.numel())
        self._bufs = b
        return b

    def _open_exchange(self, nbytes: int) -> None:
        """Collective: every rank creates its exchange buffer, the IPC handles travel through
        ``torch.distributed``, every rank maps its peers; all or none use it."""
        torch, dist, native = self.torch, self.dist, self.native
        lib = native.lib()
        handle = (C.c_ubyte * native.IPC_HANDLE_BYTES)()
        x = C.c_void_p()
        rc = lib.dsm_xchg_create(self.matcher.handle, self.world, self.rank, nbytes, C.byref(x), handle)
        note = "" if rc == 0 else lib.dsm_last_error().decode(errors="replace")
        mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=self.dev)
        every = torch.empty((self.world, native.IPC_HANDLE_BYTES), dtype=torch.uint8, device=self.dev)
        dist.all_gather_into_tensor(every, mine)
        if rc == 0:
            blob = every.cpu().numpy().tobytes()
            rc = lib.dsm_xchg_connect_ipc(x, blob)
            if rc != 0:
                note = lib.dsm_last_error().decode(errors="replace")
        ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=self.dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 1:
            self._xchg = x
            self.exchange_used = "p2p"
            return
        # some rank could not map its peers: nobody uses the exchange (nobody has pushed yet)
        if x.value:
            dist.barrier()
            lib.dsm_xchg_destroy(x)
        else:
            dist.barrier()
        self.exchange_used = "nccl"
        self.exchange_note = note or "a peer could not map this rank's buffer"

    def _close_exchange(self) -> None:
        if self._xchg is not None:
            self.stream.synchronize()
            self.dist.barrier()  # peers may still be storing into this rank's buffer
            self.native.lib().dsm_xchg_destroy(self._xchg)
            self._xchg = None
            self.exchange_used = "nccl"

    def exchange_status(self) -> int:
        """0 = every wait of the peer exchange saw all ranks; else 1 + a rank that was missing."""
        if self._xchg is None:
            return 0
        bad = C.c_uint32()
        self.native.check(self.native.lib().dsm_xchg_status(self._xchg, C.byref(bad)))
        return int(bad.value)

    def step(self, marks=None) -> None:
        """One batch on the resident buffers; asynchronous on ``self.stream`` (call under it).
        ``marks``: four CUDA events recorded after the local match, the price gather, the
        all-gather and the merge (bench.py's step breakdown)."""
        b, lib, h = self._bufs, self.native.lib(), self.matcher.handle
        check = self.native.check
        mark = (lambda i: marks[i].record(self.stream)) if marks is not None else (lambda i: None)
        check(lib.dsm_match_topk_dev(h, self.device_table.handle, C.byref(b["lay"]), b["rec"].data_ptr(), b["j"], b["k"],
                                     b["idx"].data_ptr(), b["cnt"].data_ptr(), b["max"].data_ptr()))
        mark(0)
        if self.world > 1:
            dist = self.dist
            check(lib.dsm_gather_price_dev(h, self.device_table.handle, b["max"].data_ptr(), b["j"], b["mp"].data_ptr()))
            mark(1)
            # the one exchange of the step
            if self._xchg is not None:
                gathered, stride = C.c_void_p(), C.c_uint64()
                check(lib.dsm_xchg_all_gather_dev(self._xchg, b["out"].data_ptr(), 4 * b["out"].numel(),
                                                  C.byref(gathered), C.byref(stride)))
                base_ptr, stride_bytes = gathered.value, stride.value
            else:
                dist.all_gather_into_tensor(b["g_out"], b["out"])
                base_ptr, stride_bytes = b["g_out"].data_ptr(), 4 * b["out"].numel()
            mark(2)
            w_idx, w_cnt = b["j"] * b["k"], b["j"]
            fields = (base_ptr, base_ptr + 4 * w_idx, base_ptr + 4 * (w_idx + w_cnt),
                      base_ptr + 4 * (b["words"] - 2 * b["j"]), stride_bytes)
            outs = (b["o_idx"].data_ptr(), b["o_cnt"].data_ptr(), b["o_max"].data_ptr(), b["o_mp"].data_ptr())
            if self.chunk_rows > 0:
                check(lib.dsm_merge_shards_interleaved_dev(h, self.world, b["j"], b["k"], *fields,
                                                           self.geom.data_ptr(), self.chunk_rows, *outs))
            else:
                check(lib.dsm_merge_shards_strided_dev(h, self.world, b["j"], b["k"], *fields,
                                                       self.base.data_ptr(), *outs))
            mark(3)

    def result_in_source_space(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        """(index [J,k] source positions, count [J], max_price [J]) of the last step, on every rank."""
        b, t = self._bufs, self.table
        self.stream.synchronize()
        if self.world == 1:
            idx = b["idx"].cpu().numpy().view(np.uint32)
            src = np.where(idx == NO_ROW, -1, t.perm[np.minimum(idx, t.n_rows - 1)]).astype(np.int64)
            cnt = b["cnt"].cpu().numpy().view(np.uint32).astype(np.int64)
            mx = b["max"].cpu().numpy().view(np.uint32)
            price = np.where(mx == NO_ROW, np.nan, t.price[np.minimum(mx, t.n_rows - 1)])
            return src, cnt, price
        if self.chunk_rows > 0:  # merged indices are positions in the global result order
            pos = b["o_idx"].cpu().numpy()
            src = np.where(pos < 0, -1, self.order[np.clip(pos, 0, len(self.order) - 1)]).astype(np.int64)
            return src, b["o_cnt"].cpu().numpy(), b["o_mp"].cpu().numpy()
        gathered: List[Any] = [None] * self.world
        self.dist.all_gather_object(gathered, (self.positions, t.perm))
        bases = self.base.cpu().numpy()
        src = global_rows_to_source(b["o_idx"].cpu().numpy(), bases, [g[1] for g in gathered], [g[0] for g in gathered])
        return src, b["o_cnt"].cpu().numpy(), b["o_mp"].cpu().numpy()

    def close(self) -> None:
        self._close_exchange()
        self.device_table.close()
        self.matcher.close()
