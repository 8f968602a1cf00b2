// dsmatch.cu -- host side and C ABI of libdsmatch (see include/dsmatch.h); kernels in kernels.cuh.
//
// Data flow of one call (J jobs against a table of N rows, J_t = 128 jobs per group):
//
//   records [J][rec_words] u32 --k_build_tables--> tabs [groups][E][4] u32        (scratch)
//   rows [N][n_cols] u16 + tabs --k_match--> counts                                (scratch)
//        part_cnt [J][S] matches per slot, win_rows [groups][S][W][128] matching row lanes per
//        15-block window (S slots per job, W windows per slot), group_nact [groups]
//   counts + rows + tabs + price --k_extract--> idx [J][k], cnt [J], max [J]
//   (bitmap calls: k_match writes every 32-row match word, job-major, and nothing follows)
//
// Jobs are evaluated in the order dsm_layout.job_order gives (128 per group); scratch is indexed
// by evaluation slot, results by job.
//
// k_match is the hot kernel.  One thread owns one row (a single 128-bit load of its 8 codes,
// or two for 16-code rows) and, for each column, gathers a 128-bit word group from the
// shared-memory acceptance table: bit j of word t says whether job 32t+j of the group accepts
// this dictionary entry.  ANDing the gathered groups evaluates 128 jobs against the row.  In a
// top-k call the match bits are summed in per-lane bit-sliced counters and transposed across
// the warp once per 15-block window; in a bitmap call every block is transposed so that lane j
// holds the 32-row match word of job j.  No tensor cores: there is no contraction here, the
// work is integer gathers and bit logic.
//
// Reference being replaced: gpuhunt.matches per (job, offer) at
// core/backends/base/offers.py:154-157 and server/services/pools.py:496-502,533-543, and the
// ranker at server/services/backends/__init__.py:449-451 (see include/dsmatch.h).

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "kernels.cuh"
#include "exchange.cuh"

#define MAX_SMEM_BYTES (200 * 1024)

// ------------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------------

static thread_local std::string g_err;

static int fail(int code, const char* fmt, const char* a = "", const char* b = "") {
  char buf[512];
  snprintf(buf, sizeof buf, fmt, a, b);
  g_err = buf;
  return code;
}

#define CK(call)                                                                  \
  do {                                                                            \
    cudaError_t e_ = (call);                                                      \
    if (e_ != cudaSuccess) return fail(DSM_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

// ------------------------------------------------------------------------------------------
// objects
// ------------------------------------------------------------------------------------------

struct dsm_table {
  int device;
  dsm_table_desc desc;
  uint4* rows;
  double* price;
  uint32_t* gpu_entries;
  uint32_t* bk_entries;
  uint32_t* pair_entries;
  double* dicts;                          // the DSM_N_PACK_DICTS value lists, one after the other
  uint32_t dict_n[DSM_N_PACK_DICTS];
  uint32_t dict_off[DSM_N_PACK_DICTS];
  int32_t* row_to_source;                 // [n_rows], -1 on padding rows
  uint32_t n_source;
};

// events around the launches of one call: build | match | extract (n = events used)
struct StepEvents {
  cudaEvent_t e[4];
  int n = 0;
};

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct dsm_ctx {
  int device;
  cudaStream_t stream;
  bool own_stream;
  int sm_count;
  // tuning knobs, read from the environment once when the context is created (0 = default)
  int force_rep = -1;        // DSM_REP_SHIFT: 0 single table copy, 3 eight copies
  bool no_direct = false;    // DSM_NO_DIRECT_FILL
  bool force_direct = false; // DSM_FORCE_DIRECT_FILL
  uint32_t units_per_cta = 0;// DSM_UNITS_PER_CTA
  uint32_t force_wpw = 0;    // DSM_WPW: windows per warp and unit
  uint32_t force_wpw_f = 0;  // DSM_WPW_FINE: windows per warp in the fine units at the end of the row axis
  int fine_pct = -1;         // DSM_FINE_PCT: share of the windows that goes out in fine units (default: two fine units per CTA)
  std::mutex mu;
  uint64_t launches;
  // optional timing of the match kernel alone (bench.py's roofline): event pairs recorded
  // around each k_match launch while enabled, read back by dsm_timing_read
  bool timing = false;
  std::vector<StepEvents> timing_events;
  size_t timing_used = 0;
  DevBuf tabs, counter, part_cnt, win_rows, slot_start, group_nact, d_jobs, d_order, d_idx, d_cnt, d_max, d_bits;
  DevBuf lut[DSM_MAX_COLS];
  DevBuf d_raw, d_ids, d_keys, d_sbits, d_cnt64, d_ptr64, d_out_ids;
  void* h_out_ids = nullptr;  // pinned, dsm_match_indices' result
  size_t h_out_cap = 0;
  void* h_stage = nullptr;    // pinned, raw jobs + ids laid out by dsm_match_topk_cols
  size_t h_stage_cap = 0;
};

// Every entry point makes the context's device current for its own work and puts the caller's
// device back on return, so a host framework that keeps "its" device current is not disturbed.
struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    ok = (prev == device) || cudaSetDevice(device) == cudaSuccess;
    if (prev == device) prev = -1;  // nothing to restore
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};
#define USE_DEVICE(dev)   \
  DeviceGuard guard_(dev); \
  if (!guard_.ok) return fail(DSM_E_CUDA, "cudaSetDevice failed")

static int ensure(DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return 0;
  if (b.p) CK(cudaFree(b.p));
  b.p = nullptr;
  b.cap = 0;
  size_t want = bytes + bytes / 4 + 256;
  CK(cudaMalloc(&b.p, want));
  b.cap = want;
  return 0;
}

// ------------------------------------------------------------------------------------------
// launch plan
// ------------------------------------------------------------------------------------------

struct Plan {
  LayoutDev lay;
  Pieces pc;
  uint32_t n_groups, n_blocks, split_block, slots, smem, grid, wps, dead_entry;
  uint32_t col_off[DSM_MAX_COLS], col_shift[DSM_MAX_COLS], tab_bytes, stage;  // see MatchParams
  uint32_t rep_shift, big_col, big_shift;
};

// Words of a record a column's predicate occupies.
static uint32_t pred_words(const dsm_layout* L, uint32_t c) {
  switch (L->kind[c]) {
    case DSM_COL_RANGE: return 1;
    case DSM_COL_BITSET: return (L->card[c] + 31u) / 32u;
    case DSM_COL_GPU: return 5u + L->aux[c];
    case DSM_COL_BK: return 4;
    case DSM_COL_PAIR: return L->aux[c] + 1u;
    default: return 0;
  }
}

// Geometry of a call; pure arithmetic.  Does not depend on k.
static int make_plan(const dsm_ctx* ctx, const dsm_table* t, const dsm_layout* L, uint32_t n_jobs, Plan& pl) {
  if (!L || L->n_used < 1 || L->n_used > t->desc.n_cols)
    return fail(DSM_E_ARG, "layout.n_used out of range");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  if (L->rec_words < 3 || L->rec_words > 65536) return fail(DSM_E_ARG, "layout.rec_words out of range");
  LayoutDev& d = pl.lay;
  memset(&d, 0, sizeof d);
  d.n_used = L->n_used;
  d.rec_words = L->rec_words;
  uint32_t base = 0;
  pl.dead_entry = 0xFFFFFFFFu;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    if (L->kind[c] < DSM_COL_RANGE || L->kind[c] > DSM_COL_PAIR)
      return fail(DSM_E_ARG, "layout.kind invalid");
    if (L->card[c] == 0 || L->card[c] > 65536) return fail(DSM_E_ARG, "layout.card invalid");
    if (L->kind[c] == DSM_COL_GPU && L->card[c] != t->desc.n_gpu_entries)
      return fail(DSM_E_ARG, "GPU column card != table GPU dictionary size");
    if (L->kind[c] == DSM_COL_BK && L->card[c] != t->desc.n_bk_entries)
      return fail(DSM_E_ARG, "BK column card != table BK dictionary size");
    if (L->kind[c] == DSM_COL_PAIR && (L->card[c] != t->desc.n_pair_entries || L->aux[c] == 0))
      return fail(DSM_E_ARG, "PAIR column card != table pair dictionary size (or aux = 0)");
    if (L->kind[c] == DSM_COL_GPU && L->aux[c] > 2048) return fail(DSM_E_ARG, "GPU-name bitset too long");
    // the predicate must lie inside the record, behind the two cut-off words
    const uint32_t words = pred_words(L, c);
    if (L->rec_off[c] < 2 || (unsigned long long)L->rec_off[c] + words > L->rec_words)
      return fail(DSM_E_ARG, "layout.rec_off: a column's predicate does not fit in the record");
    if (L->kind[c] == DSM_COL_BK) pl.dead_entry = base;  // entry 0 of that column
    d.kind[c] = L->kind[c];
    d.card[c] = L->card[c];
    d.flags[c] = L->flags[c];
    d.aux[c] = L->aux[c];
    d.rec_off[c] = L->rec_off[c];
    d.base[c] = base;
    base += L->card[c];
  }
  d.n_entries = base;
  // Shared-memory layout of a group's tables.  Eight copies of an entry make its gathers bank-
  // conflict free; every column gets eight while MATCH_CTAS_PER_SM CTAs' tables fit in an SM,
  // otherwise the copies of the column with the largest area are halved until they do (a table
  // of 312 regions gathered from 4 copies costs about 7 wavefronts instead of 4 when it is
  // gathered at all, which buys the room to pair cpu with memory: one gather less on every row).
  const uint32_t cuts = 2u * JPG * 4u;
  if (base * 16u + cuts > MAX_SMEM_BYTES)
    return fail(DSM_E_DICT_TOO_LARGE, "acceptance tables need more shared memory than one CTA has");
  const uint32_t budget = 220u * 1024u / MATCH_CTAS_PER_SM - 1024u - cuts;  // table area of one of 3 CTAs
  // every column 8 copies; if that does not fit, the largest column 4, 2 or 1; if not even that,
  // one copy of everything
  uint32_t big = 0;
  for (uint32_t c = 1; c < L->n_used; ++c) if (L->card[c] > L->card[big]) big = c;
  auto area_with = [&](uint32_t rep, uint32_t bshift) {
    uint32_t a = 0;
    for (uint32_t c = 0; c < L->n_used; ++c) {
      const uint32_t bytes = (L->card[c] * 16u) << (c == big ? bshift : rep);
      a += (bytes + 127u) / 128u * 128u;  // column areas start on a 128-byte boundary
    }
    return a;
  };
  pl.rep_shift = 3;
  pl.big_shift = 3;
  if (ctx->force_rep >= 0) {
    pl.rep_shift = pl.big_shift = ctx->force_rep >= 3 ? 3u : 0u;
  } else {
    while (pl.big_shift > 0 && area_with(3, pl.big_shift) > budget) --pl.big_shift;
    if (area_with(3, pl.big_shift) > budget) pl.rep_shift = pl.big_shift = 0;
  }
  pl.big_col = pl.big_shift < pl.rep_shift ? big : 0xFFFFFFFFu;
  bool single = pl.rep_shift == 0;
  uint32_t off = 0;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    pl.col_off[c] = off;
    pl.col_shift[c] = c == pl.big_col ? pl.big_shift : pl.rep_shift;
    const uint32_t bytes = (L->card[c] * 16u) << pl.col_shift[c];
    off += single ? bytes : (bytes + 127u) / 128u * 128u;
  }
  pl.tab_bytes = off;
  // TMA into a staging copy when there is room for one beside the table area (or straight in
  // place when the area is one copy of everything), else the threads read from global memory
  pl.stage = single ? 1u : ((pl.tab_bytes + base * 16u <= budget && !ctx->force_direct) || ctx->no_direct ? 1u : 0u);
  pl.smem = pl.tab_bytes + ((pl.stage && !single) ? base * 16u : 0u) + cuts;
  if (pl.smem > MAX_SMEM_BYTES) return fail(DSM_E_DICT_TOO_LARGE, "acceptance tables need more shared memory than one CTA has");
  pl.n_groups = (n_jobs + JPG - 1) / JPG;
  pl.n_blocks = t->desc.n_rows / 32u;
  pl.split_block = t->desc.split_row / 32u;

  // Resident CTAs: a multiple of the SM count.  Work units (piece, group): about units_per_cta
  // per resident CTA so that the dynamic draw balances the SMs; a piece is a whole number of
  // 15-block windows and gives every warp at least one.
  uint32_t per_sm = MATCH_CTAS_PER_SM;
  const uint32_t smem_cap = 220u * 1024u;
  while (per_sm > 1 && per_sm * (pl.smem + 1024u) > smem_cap) --per_sm;
  const uint32_t resident = (uint32_t)ctx->sm_count * per_sm;
  // Units per resident CTA: many short units balance the SMs (8 per CTA measured best on a
  // 128-group batch over 8 M rows: 3.73 ms against 3.84 with 4 and 4.10 with 2), but every unit
  // stages its group's table first, so a warp's share of a unit is kept at MIN_WPW windows or more
  // (4096 jobs x 1 M rows: 0.146 ms with 10 windows per warp, 0.154 with 5, 0.168 with 2)
  // -- unless that would leave CTAs without any unit.
  const uint32_t MIN_WPW = 8;  // k_extract reads the records of slots of up to 8 windows a chunk of slots at a time
  const uint32_t units_per_cta = ctx->units_per_cta ? ctx->units_per_cta : 8u;
  const uint32_t a_blocks = pl.split_block, b_blocks = pl.n_blocks - pl.split_block;
  const unsigned long long n_win = (unsigned long long)Pieces::windows(a_blocks) + Pieces::windows(b_blocks);
  auto wpw_for = [&](unsigned long long units) -> unsigned long long {  // windows per warp for about `units` units
    unsigned long long pieces = (units + pl.n_groups - 1) / pl.n_groups;
    if (pieces < 1) pieces = 1;
    const unsigned long long w = (n_win + pieces * MATCH_WARPS / 2) / (pieces * MATCH_WARPS);
    return w < 1 ? 1 : w;
  };
  unsigned long long wpw = wpw_for((unsigned long long)resident * units_per_cta);
  if (!ctx->units_per_cta && wpw < MIN_WPW) {
    const unsigned long long fill = wpw_for(resident);  // the share that gives every CTA one unit
    wpw = fill < MIN_WPW ? fill : MIN_WPW;
  }
  if (ctx->force_wpw) wpw = ctx->force_wpw;
  pl.pc.n_groups = pl.n_groups;
  pl.pc.n_blocks = pl.n_blocks;
  pl.pc.split_block = pl.split_block;
  pl.pc.wpw = (uint32_t)wpw;
  // The last windows (in draw order: the end of the second segment, before it the end of the
  // first) go out in units a quarter as long, about two of them per resident CTA and at most 30 %
  // of the table, so that the CTAs finish nearly together while most rows are covered by long
  // units: 4096 jobs x 1 M rows 0.141 -> 0.134 ms, 16384 jobs x 8 M rows 3.47 -> 3.33 ms.  Fine
  // units get longer if they alone would push a job past the slot count of its k_extract variant.
  pl.pc.wpw_f = ctx->force_wpw_f ? ctx->force_wpw_f : (pl.pc.wpw >= 4u ? std::max(2u, pl.pc.wpw / 4u) : pl.pc.wpw);
  if (pl.pc.wpw_f > pl.pc.wpw) pl.pc.wpw_f = pl.pc.wpw;
  const uint32_t wins_b = Pieces::windows(b_blocks);
  // k_extract: counts in registers up to 512 slots, a warp per job below 1024, else a CTA per job
  const uint32_t coarse_slots = (Pieces::pieces_for_windows(Pieces::windows(a_blocks), pl.pc.wpw) +
                                 Pieces::pieces_for_windows(wins_b, pl.pc.wpw)) * MATCH_WARPS;
  const uint32_t slot_limit = coarse_slots <= 512u ? 512u : (coarse_slots < 1024u ? 1023u : 0xFFFFFFFFu);
  for (;;) {
    unsigned long long fine_total;
    if (ctx->fine_pct >= 0) {
      fine_total = n_win * (unsigned long long)ctx->fine_pct / 100ull;
    } else {
      const unsigned long long fine_pieces = (2ull * resident + pl.n_groups - 1) / pl.n_groups;
      fine_total = std::min(fine_pieces * MATCH_WARPS * pl.pc.wpw_f, n_win * 30ull / 100ull);
    }
    const uint32_t fine_b = (uint32_t)std::min<unsigned long long>(wins_b, fine_total);
    const uint32_t fine_a = (uint32_t)(fine_total - fine_b);
    pl.pc.pa = pl.pc.cut(a_blocks, fine_a, pl.pc.ca);
    pl.pc.pb = pl.pc.cut(b_blocks, fine_b, pl.pc.cb);
    if (pl.pc.per_group() * MATCH_WARPS <= slot_limit || pl.pc.wpw_f >= pl.pc.wpw || ctx->force_wpw_f) break;
    ++pl.pc.wpw_f;
  }
  pl.slots = pl.pc.per_group() * MATCH_WARPS;
  pl.wps = pl.pc.wpw;  // windows per slot
  const uint32_t units = pl.pc.count();
  pl.grid = units < resident ? units : resident;
  return 0;
}

// k_build_tables: x = 32-job tiles, y = columns, z = chunks of BUILD_CHUNK dictionary entries
static dim3 build_grid(const Plan& pl) {
  uint32_t longest = 1;
  for (uint32_t c = 0; c < pl.lay.n_used; ++c) longest = pl.lay.card[c] > longest ? pl.lay.card[c] : longest;
  return dim3(pl.n_groups * TPG, pl.lay.n_used, (longest + BUILD_CHUNK - 1) / BUILD_CHUNK);
}

template <int NCMAX, bool ROW32, int MODE, bool HAS_LUT, bool EVFAST, bool MIXED>
static cudaError_t launch_match_v(const MatchParams& mp, uint32_t grid, uint32_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(k_match<NCMAX, ROW32, MODE, HAS_LUT, EVFAST, MIXED>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  k_match<NCMAX, ROW32, MODE, HAS_LUT, EVFAST, MIXED><<<grid, MATCH_THREADS, smem, st>>>(mp);
  return cudaGetLastError();
}

template <int NCMAX, bool ROW32, int MODE, bool HAS_LUT, bool EVFAST>
static cudaError_t launch_match_m(const MatchParams& mp, uint32_t grid, uint32_t smem, cudaStream_t st) {
  return mp.big_col != 0xFFFFFFFFu ? launch_match_v<NCMAX, ROW32, MODE, HAS_LUT, EVFAST, true>(mp, grid, smem, st)
                                   : launch_match_v<NCMAX, ROW32, MODE, HAS_LUT, EVFAST, false>(mp, grid, smem, st);
}

// evfast: see k_match; only 16-byte rows in count mode have the unrolled windows it concerns
template <int MODE>
static cudaError_t launch_match(bool row32, const MatchParams& mp, bool has_lut, bool evfast, uint32_t grid,
                                uint32_t smem, cudaStream_t st) {
  if (MODE != MODE_COUNT || row32) evfast = false;
#ifdef DSM_FAST_BUILD  // experiments: only the shape the benchmark uses
  if (row32 || has_lut) return cudaErrorInvalidValue;
  return evfast ? launch_match_m<8, false, MODE, false, true>(mp, grid, smem, st)
                : launch_match_m<8, false, MODE, false, false>(mp, grid, smem, st);
#else
  if (!row32) {
    if (evfast) return has_lut ? launch_match_m<8, false, MODE, true, true>(mp, grid, smem, st)
                               : launch_match_m<8, false, MODE, false, true>(mp, grid, smem, st);
    return has_lut ? launch_match_m<8, false, MODE, true, false>(mp, grid, smem, st)
                   : launch_match_m<8, false, MODE, false, false>(mp, grid, smem, st);
  }
  return has_lut ? launch_match_m<16, true, MODE, true, false>(mp, grid, smem, st)
                 : launch_match_m<16, true, MODE, false, false>(mp, grid, smem, st);
#endif
}

// Fill the kernel parameters shared by every way of running k_build_tables + k_match.
static void fill_params(const dsm_table* t, const dsm_layout* L, const Plan& pl, const uint32_t* d_jobs,
                        const uint32_t* d_order, uint32_t n_jobs, uint4* tabs, uint32_t* counter,
                        uint32_t* group_nact, BuildParams& bp, MatchParams& mp, bool& has_lut) {
  bp.lay = pl.lay;
  bp.recs = d_jobs;
  bp.job_order = d_order;
  bp.n_jobs = n_jobs;
  bp.gpu_entries = t->gpu_entries;
  bp.bk_entries = t->bk_entries;
  bp.pair_entries = t->pair_entries;
  bp.tabs = tabs;
  bp.work_counter = counter;

  memset(&mp, 0, sizeof mp);
  mp.rows = t->rows;
  mp.tabs = tabs;
  mp.recs = d_jobs;
  mp.job_order = d_order;
  has_lut = false;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    mp.lut[c] = L->lut[c];
    has_lut = has_lut || L->lut[c] != nullptr;
    mp.base[c] = pl.lay.base[c];
  }
  mp.pc = pl.pc;
  mp.n_used = L->n_used;
  mp.rec_words = L->rec_words;
  mp.n_entries = pl.lay.n_entries;
  mp.dead_entry = pl.dead_entry;
  mp.live_a = t->desc.n_live_a;
  mp.live_b = t->desc.split_row + t->desc.n_live_b;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    mp.col_off[c] = pl.col_off[c];
    mp.col_shift[c] = pl.col_shift[c];
    mp.gbase[c] = pl.col_off[c] / (16u << pl.rep_shift);
  }
  mp.rep_shift = pl.rep_shift;
  mp.big_col = pl.big_col;
  mp.big_shift = pl.big_shift;
  mp.tab_bytes = pl.tab_bytes;
  mp.stage = pl.stage;
  mp.n_jobs = n_jobs;
  mp.slots = pl.slots;
  mp.work_counter = counter;
  mp.group_nact = group_nact;
  mp.wps = pl.wps;
}

// The scratch of one call: acceptance tables, work counter, per-group column counts and, for a
// top-k call, the counts per slot and the records per window.  Grown on demand (cudaFree +
// cudaMalloc: a device-wide synchronisation) -- dsm_reserve does it ahead of time.
static int ensure_match_scratch(dsm_ctx* ctx, const Plan& pl, bool bitmap) {
  int rc;
  const size_t jobs_pad = (size_t)pl.n_groups * JPG;
  if ((rc = ensure(ctx->tabs, (size_t)pl.n_groups * pl.lay.n_entries * 16))) return rc;
  if ((rc = ensure(ctx->counter, 256))) return rc;
  if ((rc = ensure(ctx->group_nact, (size_t)pl.n_groups * 4))) return rc;
  if (!bitmap) {
    if ((unsigned long long)jobs_pad * pl.slots * pl.wps * 4ull > (64ull << 30))
      return fail(DSM_E_TOO_LARGE, "per-window match records of this call exceed 64 GiB; split the jobs over several calls");
    if ((rc = ensure(ctx->part_cnt, jobs_pad * pl.slots * 4))) return rc;
    if ((rc = ensure(ctx->win_rows, jobs_pad * pl.slots * pl.wps * 4))) return rc;
    if ((rc = ensure(ctx->slot_start, ((size_t)pl.slots + 1) * 4))) return rc;
  }
  return 0;
}

// Issue build -> match -> extract on the context's stream.  d_jobs, d_order (may be NULL) and the
// outputs are device pointers.  Caller holds ctx->mu and has made the context's device current.
static int run_match(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* L, const uint32_t* d_jobs,
                     const uint32_t* d_order, uint32_t n_jobs, uint32_t k, bool bitmap, uint32_t* d_idx,
                     uint32_t* d_cnt, uint32_t* d_max, uint32_t* d_bits) {
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  if (!bitmap && (k == 0 || k > 4096)) return fail(DSM_E_ARG, "k must be in 1..4096");
  Plan pl;
  int rc = make_plan(ctx, t, L, n_jobs, pl);
  if (rc) return rc;
  const bool row32 = t->desc.n_cols == 16;

  if ((rc = ensure_match_scratch(ctx, pl, bitmap))) return rc;

  BuildParams bp;
  MatchParams mp;
  bool has_lut;
  fill_params(t, L, pl, d_jobs, d_order, n_jobs, (uint4*)ctx->tabs.p, (uint32_t*)ctx->counter.p,
              (uint32_t*)ctx->group_nact.p, bp, mp, has_lut);
  mp.part_cnt = (uint32_t*)ctx->part_cnt.p;
  mp.win_rows = (uint32_t*)ctx->win_rows.p;
  mp.slot_start = (uint32_t*)ctx->slot_start.p;
  mp.bits = d_bits;

  StepEvents* tev = nullptr;
  if (ctx->timing) {
    if (ctx->timing_used == ctx->timing_events.size()) {
      StepEvents ev;
      for (cudaEvent_t& e : ev.e) CK(cudaEventCreate(&e));
      ctx->timing_events.push_back(ev);
    }
    tev = &ctx->timing_events[ctx->timing_used++];
    tev->n = bitmap ? 3 : 4;
    CK(cudaEventRecord(tev->e[0], ctx->stream));
  }
  k_build_tables<<<build_grid(pl), 256, 0, ctx->stream>>>(bp);
  CK(cudaGetLastError());
  ctx->launches++;
  if (tev) CK(cudaEventRecord(tev->e[1], ctx->stream));

  // windows with a cut-off inside: up to 2 x 128 cut-offs per job group over the table's windows
  const unsigned long long n_win_all = (unsigned long long)Pieces::windows(pl.split_block) +
                                       Pieces::windows(pl.n_blocks - pl.split_block);
  const bool evfast = 2ull * JPG * 20ull > n_win_all;  // more than ~5 % of the windows
  cudaError_t e = bitmap ? launch_match<MODE_BITMAP>(row32, mp, has_lut, false, pl.grid, pl.smem, ctx->stream)
                         : launch_match<MODE_COUNT>(row32, mp, has_lut, evfast, pl.grid, pl.smem, ctx->stream);
  if (e != cudaSuccess) return fail(DSM_E_CUDA, "k_match launch: %s", cudaGetErrorString(e));
  if (tev) CK(cudaEventRecord(tev->e[2], ctx->stream));
  ctx->launches++;

  if (!bitmap) {
    ExtractParams gp;
    memset(&gp, 0, sizeof gp);
    gp.pc = pl.pc;
    gp.win_rows = mp.win_rows;
    gp.slot_start = mp.slot_start;
    gp.wps = mp.wps;
    gp.group_nact = mp.group_nact;
    gp.part_cnt = mp.part_cnt;
    gp.price = t->price;
    gp.job_order = d_order;
    gp.n_jobs = n_jobs;
    gp.k = k;
    gp.slots = pl.slots;
    gp.slots_a = pl.pc.pa * MATCH_WARPS;
    gp.rows = t->rows;
    gp.tabs = mp.tabs;
    gp.recs = d_jobs;
    for (uint32_t c = 0; c < DSM_MAX_COLS; ++c) { gp.lut[c] = mp.lut[c]; gp.base[c] = mp.base[c]; }
    gp.n_cols = t->desc.n_cols;
    gp.n_used = L->n_used;
    gp.n_entries = pl.lay.n_entries;
    gp.rec_words = L->rec_words;
    gp.live_a = mp.live_a;
    gp.live_b = mp.live_b;
    gp.out_idx = d_idx;
    gp.out_cnt = d_cnt;
    gp.out_max = d_max;
    // few slots per job: a warp per job; thousands (small batches): a CTA per job
    const uint32_t chunk_bytes = (pl.slots + 31u) / 32u * 4u;
    const bool wide = pl.slots >= 1024u;
    const dim3 grid(wide ? n_jobs : (n_jobs + 7) / 8);
    const uint32_t sm_bytes = wide ? chunk_bytes : 8u * chunk_bytes;
#define EXTRACT(W, R, L, S) k_extract<W, R, L, S><<<grid, 256, sm_bytes, ctx->stream>>>(gp)
    const bool shortw = !wide && pl.slots <= 512u;  // counts in registers, needed slots known before any further load
    if (wide) {
      if (row32) { if (has_lut) EXTRACT(8, true, true, 0); else EXTRACT(8, true, false, 0); }
      else       { if (has_lut) EXTRACT(8, false, true, 0); else EXTRACT(8, false, false, 0); }
    } else if (shortw && pl.wps <= 8u) {  // ... and their records in one batch
      if (row32) { if (has_lut) EXTRACT(1, true, true, 1); else EXTRACT(1, true, false, 1); }
      else       { if (has_lut) EXTRACT(1, false, true, 1); else EXTRACT(1, false, false, 1); }
    } else if (shortw) {
      if (row32) { if (has_lut) EXTRACT(1, true, true, 2); else EXTRACT(1, true, false, 2); }
      else       { if (has_lut) EXTRACT(1, false, true, 2); else EXTRACT(1, false, false, 2); }
    } else {
      if (row32) { if (has_lut) EXTRACT(1, true, true, 0); else EXTRACT(1, true, false, 0); }
      else       { if (has_lut) EXTRACT(1, false, true, 0); else EXTRACT(1, false, false, 0); }
    }
#undef EXTRACT
    CK(cudaGetLastError());
    ctx->launches++;
    if (tev) CK(cudaEventRecord(tev->e[3], ctx->stream));
  }
  return 0;
}

// Host-buffer entry points: layout->job_order is a host array; validate that it is a permutation
// (a bad one would make the kernels write outside the result buffers) and copy it to the device.
static int upload_order(dsm_ctx* ctx, const dsm_layout* L, uint32_t n_jobs, const uint32_t** d_order) {
  *d_order = nullptr;
  if (!L->job_order) return 0;
  std::vector<bool> seen(n_jobs, false);
  for (uint32_t i = 0; i < n_jobs; ++i) {
    const uint32_t j = L->job_order[i];
    if (j >= n_jobs || seen[j]) return fail(DSM_E_ARG, "layout.job_order is not a permutation of 0..n_jobs-1");
    seen[j] = true;
  }
  int rc = ensure(ctx->d_order, (size_t)n_jobs * 4);
  if (rc) return rc;
  CK(cudaMemcpyAsync(ctx->d_order.p, L->job_order, (size_t)n_jobs * 4, cudaMemcpyHostToDevice, ctx->stream));
  *d_order = (const uint32_t*)ctx->d_order.p;
  return 0;
}

extern "C" {

const char* dsm_last_error(void) { return g_err.c_str(); }

const char* dsm_build_info(void) { return "libdsmatch 0.2 sm_100a jobs_per_group=128"; }

int dsm_ctx_create(int device, void* stream, dsm_ctx** out) {
  if (!out) return fail(DSM_E_ARG, "out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(DSM_E_NO_DEVICE, "no CUDA device: %s", cudaGetErrorString(e));
  if (device < 0 || device >= n) return fail(DSM_E_ARG, "device ordinal out of range");
  USE_DEVICE(device);
  dsm_ctx* c = new (std::nothrow) dsm_ctx();
  if (!c) return fail(DSM_E_ARG, "out of host memory");
  c->device = device;
  c->launches = 0;
  c->own_stream = stream == nullptr;
  if (c->own_stream) {
    e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete c; return fail(DSM_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
  } else {
    c->stream = (cudaStream_t)stream;
  }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) { delete c; return fail(DSM_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e)); }
  c->sm_count = prop.multiProcessorCount;
  // tuning knobs are read here, once: a stray variable cannot change a running scheduler
  if (const char* v = getenv("DSM_REP_SHIFT")) c->force_rep = atoi(v);
  c->no_direct = getenv("DSM_NO_DIRECT_FILL") != nullptr;
  c->force_direct = getenv("DSM_FORCE_DIRECT_FILL") != nullptr;
  if (const char* v = getenv("DSM_UNITS_PER_CTA")) c->units_per_cta = (uint32_t)atoi(v);
  if (const char* v = getenv("DSM_WPW")) c->force_wpw = (uint32_t)atoi(v);
  if (const char* v = getenv("DSM_WPW_FINE")) c->force_wpw_f = (uint32_t)atoi(v);
  if (const char* v = getenv("DSM_FINE_PCT")) c->fine_pct = std::min(100, std::max(0, atoi(v)));
  *out = c;
  return 0;
}

int dsm_ctx_destroy(dsm_ctx* ctx) {
  if (!ctx) return 0;
  DeviceGuard guard_(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->tabs, &ctx->counter, &ctx->part_cnt, &ctx->win_rows, &ctx->slot_start, &ctx->group_nact,
                    &ctx->d_jobs, &ctx->d_order, &ctx->d_idx, &ctx->d_cnt, &ctx->d_max, &ctx->d_bits,
                    &ctx->d_raw, &ctx->d_ids, &ctx->d_keys, &ctx->d_sbits, &ctx->d_cnt64, &ctx->d_ptr64, &ctx->d_out_ids};
  if (ctx->h_out_ids) cudaFreeHost(ctx->h_out_ids);
  if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
  for (DevBuf* b : bufs) if (b->p) cudaFree(b->p);
  for (int c = 0; c < DSM_MAX_COLS; ++c) if (ctx->lut[c].p) cudaFree(ctx->lut[c].p);
  for (auto& ev : ctx->timing_events) for (cudaEvent_t e : ev.e) cudaEventDestroy(e);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return 0;
}

int dsm_table_create(dsm_ctx* ctx, const dsm_table_desc* desc, const uint16_t* rows,
                     const double* price, const uint32_t* gpu_entries, const uint32_t* bk_entries,
                     const uint32_t* pair_entries, dsm_table** out) {
  if (!ctx || !desc || !rows || !out) return fail(DSM_E_ARG, "NULL argument");
  if (desc->n_cols != 8 && desc->n_cols != 16) return fail(DSM_E_ARG, "n_cols must be 8 or 16");
  if (desc->n_rows == 0 || desc->n_rows % 32 || desc->split_row % 32 || desc->split_row > desc->n_rows)
    return fail(DSM_E_ARG, "n_rows / split_row must be multiples of 32, split_row <= n_rows");
  if (desc->n_live_a > desc->split_row || desc->n_live_b > desc->n_rows - desc->split_row)
    return fail(DSM_E_ARG, "n_live_a / n_live_b exceed their segments");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  dsm_table* t = new (std::nothrow) dsm_table();
  if (!t) return fail(DSM_E_ARG, "out of host memory");
  memset(t, 0, sizeof *t);
  t->device = ctx->device;
  t->desc = *desc;
  const size_t row_bytes = (size_t)desc->n_rows * desc->n_cols * 2;
  cudaError_t e = cudaMalloc((void**)&t->rows, row_bytes);
  if (e == cudaSuccess) e = cudaMemcpyAsync(t->rows, rows, row_bytes, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess && price) {
    e = cudaMalloc((void**)&t->price, (size_t)desc->n_rows * 8);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->price, price, (size_t)desc->n_rows * 8, cudaMemcpyHostToDevice, ctx->stream);
  }
  if (e == cudaSuccess && desc->n_gpu_entries) {
    if (!gpu_entries) { e = cudaErrorInvalidValue; }
    else {
      const size_t b = (size_t)desc->n_gpu_entries * DSM_GPU_ENTRY_WORDS * 4;
      e = cudaMalloc((void**)&t->gpu_entries, b);
      if (e == cudaSuccess) e = cudaMemcpyAsync(t->gpu_entries, gpu_entries, b, cudaMemcpyHostToDevice, ctx->stream);
    }
  }
  if (e == cudaSuccess && desc->n_bk_entries) {
    if (!bk_entries) { e = cudaErrorInvalidValue; }
    else {
      const size_t b = (size_t)desc->n_bk_entries * DSM_BK_ENTRY_WORDS * 4;
      e = cudaMalloc((void**)&t->bk_entries, b);
      if (e == cudaSuccess) e = cudaMemcpyAsync(t->bk_entries, bk_entries, b, cudaMemcpyHostToDevice, ctx->stream);
    }
  }
  if (e == cudaSuccess && desc->n_pair_entries) {
    if (!pair_entries) { e = cudaErrorInvalidValue; }
    else {
      const size_t b = (size_t)desc->n_pair_entries * 4;
      e = cudaMalloc((void**)&t->pair_entries, b);
      if (e == cudaSuccess) e = cudaMemcpyAsync(t->pair_entries, pair_entries, b, cudaMemcpyHostToDevice, ctx->stream);
    }
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    if (t->pair_entries) cudaFree(t->pair_entries);
    if (t->rows) cudaFree(t->rows);
    if (t->price) cudaFree(t->price);
    if (t->gpu_entries) cudaFree(t->gpu_entries);
    if (t->bk_entries) cudaFree(t->bk_entries);
    delete t;
    return fail(DSM_E_CUDA, "dsm_table_create: %s", cudaGetErrorString(e));
  }
  *out = t;
  return 0;
}

int dsm_table_destroy(dsm_table* t) {
  if (!t) return 0;
  DeviceGuard guard_(t->device);
  if (t->pair_entries) cudaFree(t->pair_entries);
  if (t->dicts) cudaFree(t->dicts);
  if (t->row_to_source) cudaFree(t->row_to_source);
  if (t->rows) cudaFree(t->rows);
  if (t->price) cudaFree(t->price);
  if (t->gpu_entries) cudaFree(t->gpu_entries);
  if (t->bk_entries) cudaFree(t->bk_entries);
  delete t;
  return 0;
}

int dsm_table_patch_rows(dsm_ctx* ctx, dsm_table* t, const uint32_t* rows, uint32_t n, const uint16_t* codes) {
  if (!ctx || !t || (n && (!rows || !codes))) return fail(DSM_E_ARG, "NULL argument");
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  if (n == 0) return 0;
  for (uint32_t i = 0; i < n; ++i)
    if (rows[i] >= t->desc.n_rows) return fail(DSM_E_ARG, "row index out of range");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  const size_t row_bytes = (size_t)t->desc.n_cols * 2;
  int rc;
  if ((rc = ensure(ctx->d_ids, (size_t)n * 4 + 16))) return rc;
  if ((rc = ensure(ctx->d_raw, (size_t)n * row_bytes + 16))) return rc;
  CK(cudaMemcpyAsync(ctx->d_ids.p, rows, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->d_raw.p, codes, (size_t)n * row_bytes, cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t row_u4 = t->desc.n_cols / 8;
  k_patch_rows<<<(n * row_u4 + 255) / 256, 256, 0, ctx->stream>>>(t->rows, (const uint32_t*)ctx->d_ids.p, n,
                                                                   (const uint4*)ctx->d_raw.p, row_u4);
  CK(cudaGetLastError());
  ctx->launches++;
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_table_set_pack_dicts(dsm_ctx* ctx, dsm_table* t, const double* values,
                             const uint32_t counts[DSM_N_PACK_DICTS]) {
  if (!ctx || !t || !values || !counts) return fail(DSM_E_ARG, "NULL argument");
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  size_t total = 0;
  for (int i = 0; i < DSM_N_PACK_DICTS; ++i) {
    if (counts[i] > 65536) return fail(DSM_E_ARG, "a value list has more than 65536 entries");
    t->dict_off[i] = (uint32_t)total;
    t->dict_n[i] = counts[i];
    total += counts[i];
  }
  if (t->dicts) { CK(cudaFree(t->dicts)); t->dicts = nullptr; }
  CK(cudaMalloc((void**)&t->dicts, (total + 1) * 8));
  if (total) CK(cudaMemcpyAsync(t->dicts, values, total * 8, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_table_set_source_map(dsm_ctx* ctx, dsm_table* t, const int32_t* row_to_source, uint32_t n_source) {
  if (!ctx || !t || !row_to_source) return fail(DSM_E_ARG, "NULL argument");
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  for (uint32_t r = 0; r < t->desc.n_rows; ++r)
    if (row_to_source[r] >= (int64_t)n_source) return fail(DSM_E_ARG, "row_to_source entry >= n_source");
  if (t->row_to_source) { CK(cudaFree(t->row_to_source)); t->row_to_source = nullptr; }
  CK(cudaMalloc((void**)&t->row_to_source, (size_t)t->desc.n_rows * 4));
  CK(cudaMemcpyAsync(t->row_to_source, row_to_source, (size_t)t->desc.n_rows * 4, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  t->n_source = n_source;
  return 0;
}

int dsm_lut_upload(dsm_ctx* ctx, uint32_t column, const uint16_t* lut, uint32_t n,
                   const uint16_t** dev_out) {
  if (!ctx || !lut || !dev_out || column >= DSM_MAX_COLS || n == 0) return fail(DSM_E_ARG, "bad argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  int rc = ensure(ctx->lut[column], (size_t)n * 2);
  if (rc) return rc;
  CK(cudaMemcpyAsync(ctx->lut[column].p, lut, (size_t)n * 2, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  *dev_out = (const uint16_t*)ctx->lut[column].p;
  return 0;
}

// k_pack_jobs (+ k_order_jobs) on the context's stream.  Caller holds ctx->mu, device current.
static int run_pack(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* L, const dsm_raw_job* d_raw, uint32_t n_jobs,
                    const int32_t* d_ids, uint32_t* d_records, uint32_t* d_order) {
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  if (!t->dicts) return fail(DSM_E_ARG, "the table has no value lists: call dsm_table_set_pack_dicts first");
  if (L->job_order) return fail(DSM_E_ARG, "layout.job_order must be NULL when the records are built on the device");
  Plan pl;
  int rc = make_plan(ctx, t, L, n_jobs, pl);  // validates the layout
  if (rc) return rc;
  PackParams pp;
  memset(&pp, 0, sizeof pp);
  pp.raw = d_raw;
  pp.ids = d_ids;
  pp.n_jobs = n_jobs;
  pp.lay = pl.lay;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    if (L->lut[c]) return fail(DSM_E_ARG, "columns with a per-call remap cannot be packed on the device");
    const uint32_t role = L->role[c], kind = L->kind[c];
    const bool ok = (role == DSM_ROLE_CPU && (kind == DSM_COL_RANGE || kind == DSM_COL_BITSET)) ||
                    ((role == DSM_ROLE_MEMORY || role == DSM_ROLE_DISK) && kind == DSM_COL_RANGE) ||
                    (role == DSM_ROLE_CPU_MEMORY && kind == DSM_COL_PAIR) || (role == DSM_ROLE_GPU && kind == DSM_COL_GPU) ||
                    (role == DSM_ROLE_BK && kind == DSM_COL_BK) ||
                    (role >= DSM_ROLE_REGION && role <= DSM_ROLE_ZONE && kind == DSM_COL_BITSET) ||
                    (role == DSM_ROLE_STATE && kind == DSM_COL_BITSET && L->card[c] <= 256 && L->card[c] % 16 == 0);
    if (role == DSM_ROLE_STATE) pp.has_state = 1;
    if (!ok) return fail(DSM_E_ARG, "layout.role does not fit the column's kind");
    if (role == DSM_ROLE_CPU && kind == DSM_COL_BITSET && L->card[c] != t->dict_n[DSM_DICT_CPU])
      return fail(DSM_E_ARG, "cpu bitset column: card != number of cpu values");
    pp.role[c] = role;
  }
  for (int i = 0; i < DSM_N_PACK_DICTS; ++i) {
    pp.dict[i] = t->dicts + t->dict_off[i];
    pp.dict_n[i] = t->dict_n[i];
  }
  if (!t->price) return fail(DSM_E_ARG, "the table has no prices (needed for max_price cut-offs)");
  pp.price = t->price;
  pp.n_live_a = t->desc.n_live_a;
  pp.split_row = t->desc.split_row;
  pp.n_live_b = t->desc.n_live_b;
  pp.recs = d_records;
  pp.keys = nullptr;
  if (d_order) {
    if ((rc = ensure(ctx->d_keys, (size_t)n_jobs * 4))) return rc;
    pp.keys = (uint32_t*)ctx->d_keys.p;
  }
  k_pack_jobs<<<(n_jobs + 127) / 128, 128, 0, ctx->stream>>>(pp);
  CK(cudaGetLastError());
  ctx->launches++;
  if (d_order) {
    k_order_jobs<<<1, 1024, 0, ctx->stream>>>(pp.keys, n_jobs, d_order);
    CK(cudaGetLastError());
    ctx->launches++;
  }
  return 0;
}

int dsm_jobs_pack_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_raw_job* d_jobs,
                      uint32_t n_jobs, const int32_t* d_ids, uint32_t* d_records, uint32_t* d_order) {
  if (!ctx || !t || !layout || !d_jobs || !d_records) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  return run_pack(ctx, t, layout, d_jobs, n_jobs, d_ids, d_records, d_order);
}

static int match_topk_raw_locked(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_raw_job* jobs,
                                 uint32_t n_jobs, const int32_t* ids, uint32_t n_ids, uint32_t k, uint32_t* out_idx,
                                 uint32_t* out_cnt, uint32_t* out_max, int32_t* out_src, double* out_max_price);

// Shared by the two raw-job entry points.  src != NULL: answer in source positions / prices.
static int match_topk_raw(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_raw_job* jobs,
                          uint32_t n_jobs, const int32_t* ids, uint32_t n_ids, uint32_t k, uint32_t* out_idx,
                          uint32_t* out_cnt, uint32_t* out_max, int32_t* out_src, double* out_max_price) {
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  if (k == 0 || k > 4096) return fail(DSM_E_ARG, "k must be in 1..4096");
  if (layout->rec_words < 3 || layout->rec_words > 65536) return fail(DSM_E_ARG, "layout.rec_words out of range");
  if (out_src && !t->row_to_source)
    return fail(DSM_E_ARG, "the table has no source map: call dsm_table_set_source_map first");
  // the id ranges of every job must lie inside the id array (the kernel would read outside it)
  for (uint32_t j = 0; j < n_jobs; ++j) {
    const dsm_raw_job& r = jobs[j];
    const unsigned long long ends[4] = {(unsigned long long)r.gpu_name_off + r.gpu_name_cnt,
                                        (unsigned long long)r.region_off + r.region_cnt,
                                        (unsigned long long)r.itype_off + r.itype_cnt,
                                        (unsigned long long)r.zone_off + r.zone_cnt};
    for (unsigned long long e : ends)
      if (e > n_ids) return fail(DSM_E_ARG, "a raw job's id range reaches past the id array");
  }
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  return match_topk_raw_locked(ctx, t, layout, jobs, n_jobs, ids, n_ids, k, out_idx, out_cnt, out_max, out_src,
                               out_max_price);
}

// Caller holds ctx->mu, device current, arguments validated.
static int match_topk_raw_locked(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_raw_job* jobs,
                                 uint32_t n_jobs, const int32_t* ids, uint32_t n_ids, uint32_t k, uint32_t* out_idx,
                                 uint32_t* out_cnt, uint32_t* out_max, int32_t* out_src, double* out_max_price) {
  int rc;
  if ((rc = ensure(ctx->d_raw, (size_t)n_jobs * sizeof(dsm_raw_job)))) return rc;
  if ((rc = ensure(ctx->d_ids, (size_t)(n_ids + 1) * 4))) return rc;
  if ((rc = ensure(ctx->d_jobs, (size_t)n_jobs * layout->rec_words * 4))) return rc;
  if ((rc = ensure(ctx->d_order, (size_t)n_jobs * 4))) return rc;
  if ((rc = ensure(ctx->d_idx, (size_t)n_jobs * k * 4))) return rc;
  if ((rc = ensure(ctx->d_cnt, (size_t)n_jobs * 4))) return rc;
  if ((rc = ensure(ctx->d_max, (size_t)n_jobs * 4))) return rc;
  CK(cudaMemcpyAsync(ctx->d_raw.p, jobs, (size_t)n_jobs * sizeof(dsm_raw_job), cudaMemcpyHostToDevice, ctx->stream));
  if (n_ids) CK(cudaMemcpyAsync(ctx->d_ids.p, ids, (size_t)n_ids * 4, cudaMemcpyHostToDevice, ctx->stream));
  uint32_t* d_order = n_jobs > JPG ? (uint32_t*)ctx->d_order.p : nullptr;  // one group: nothing to order
  rc = run_pack(ctx, t, layout, (const dsm_raw_job*)ctx->d_raw.p, n_jobs, (const int32_t*)ctx->d_ids.p,
                (uint32_t*)ctx->d_jobs.p, d_order);
  if (rc) return rc;
  rc = run_match(ctx, t, layout, (const uint32_t*)ctx->d_jobs.p, d_order, n_jobs, k, false,
                 (uint32_t*)ctx->d_idx.p, (uint32_t*)ctx->d_cnt.p, (uint32_t*)ctx->d_max.p, nullptr);
  if (rc) return rc;
  if (out_src) {
    // rows -> source positions and the max row's price, on the device: [src J*k | price J]
    const size_t n = (size_t)n_jobs * k;
    if ((rc = ensure(ctx->d_out_ids, n * 4 + (size_t)n_jobs * 8 + 8))) return rc;
    int32_t* d_src = (int32_t*)ctx->d_out_ids.p;
    double* d_price = (double*)((char*)ctx->d_out_ids.p + (n * 4 + 7) / 8 * 8);
    k_rows_to_source<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>((const uint32_t*)ctx->d_idx.p, (uint32_t)n,
                                                                          t->row_to_source, d_src);
    k_gather_price<<<(n_jobs + 255) / 256, 256, 0, ctx->stream>>>(t->price, (const uint32_t*)ctx->d_max.p, n_jobs, d_price);
    CK(cudaGetLastError());
    ctx->launches += 2;
    CK(cudaMemcpyAsync(out_src, d_src, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(out_max_price, d_price, (size_t)n_jobs * 8, cudaMemcpyDeviceToHost, ctx->stream));
  } else {
    CK(cudaMemcpyAsync(out_idx, ctx->d_idx.p, (size_t)n_jobs * k * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(out_max, ctx->d_max.p, (size_t)n_jobs * 4, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaMemcpyAsync(out_cnt, ctx->d_cnt.p, (size_t)n_jobs * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_match_topk_raw(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_raw_job* jobs,
                       uint32_t n_jobs, const int32_t* ids, uint32_t n_ids, uint32_t k, uint32_t* out_idx,
                       uint32_t* out_cnt, uint32_t* out_max) {
  if (!ctx || !t || !layout || !jobs || !out_idx || !out_cnt || !out_max || (n_ids && !ids))
    return fail(DSM_E_ARG, "NULL argument");
  return match_topk_raw(ctx, t, layout, jobs, n_jobs, ids, n_ids, k, out_idx, out_cnt, out_max, nullptr, nullptr);
}

int dsm_match_topk_raw_src(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_raw_job* jobs,
                           uint32_t n_jobs, const int32_t* ids, uint32_t n_ids, uint32_t k, int32_t* out_src,
                           uint32_t* out_cnt, double* out_max_price) {
  if (!ctx || !t || !layout || !jobs || !out_src || !out_cnt || !out_max_price || (n_ids && !ids))
    return fail(DSM_E_ARG, "NULL argument");
  return match_topk_raw(ctx, t, layout, jobs, n_jobs, ids, n_ids, k, nullptr, out_cnt, nullptr, out_src, out_max_price);
}

int dsm_match_topk_cols(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const dsm_job_columns* c,
                        uint32_t n_jobs, uint32_t k, int32_t* out_src, uint32_t* out_cnt, double* out_max_price) {
  if (!ctx || !t || !layout || !c || !out_src || !out_cnt || !out_max_price) return fail(DSM_E_ARG, "NULL argument");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  if (k == 0 || k > 4096) return fail(DSM_E_ARG, "k must be in 1..4096");
  if (layout->rec_words < 3 || layout->rec_words > 65536) return fail(DSM_E_ARG, "layout.rec_words out of range");
  if (!t->row_to_source) return fail(DSM_E_ARG, "the table has no source map: call dsm_table_set_source_map first");
  if (!c->max_price || !c->cpu || !c->memory || !c->disk || !c->gpu_count || !c->gpu_memory || !c->gpu_total ||
      !c->cc_min || !c->src_backend_mask || !c->off_backend_mask || !c->spot || !c->gpu_vendor)
    return fail(DSM_E_ARG, "dsm_job_columns: a required column is NULL");
  struct Csr { const int64_t* ptr; const int64_t* ids; } sets[4] = {
      {c->gpu_name_ptr, c->gpu_name_ids}, {c->region_ptr, c->region_ids}, {c->itype_ptr, c->itype_ids},
      {c->zone_ptr, c->zone_ids}};
  unsigned long long n_ids = 0;
  for (const Csr& s : sets) {
    if (!s.ptr) continue;
    if (s.ptr[0] != 0 || s.ptr[n_jobs] < 0 || (s.ptr[n_jobs] > 0 && !s.ids))
      return fail(DSM_E_ARG, "dsm_job_columns: malformed CSR set");
    for (uint32_t j = 0; j < n_jobs; ++j)
      if (s.ptr[j + 1] < s.ptr[j]) return fail(DSM_E_ARG, "dsm_job_columns: CSR pointers must not decrease");
    n_ids += (unsigned long long)s.ptr[n_jobs];
  }
  if (n_ids > 0x7FFFFFFFull) return fail(DSM_E_TOO_LARGE, "more than 2^31 set members in one call");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  const size_t raw_bytes = (size_t)n_jobs * sizeof(dsm_raw_job);
  const size_t need = raw_bytes + (size_t)(n_ids + 1) * 4;
  if (need > ctx->h_stage_cap) {
    if (ctx->h_stage) CK(cudaFreeHost(ctx->h_stage));
    ctx->h_stage = nullptr;
    ctx->h_stage_cap = 0;
    CK(cudaHostAlloc(&ctx->h_stage, need + need / 4 + 4096, cudaHostAllocDefault));
    ctx->h_stage_cap = need + need / 4 + 4096;
  }
  dsm_raw_job* raw = (dsm_raw_job*)ctx->h_stage;
  int32_t* ids = (int32_t*)((char*)ctx->h_stage + raw_bytes);
  uint32_t base[4], at = 0;
  for (int s = 0; s < 4; ++s) {
    base[s] = at;
    if (!sets[s].ptr) continue;
    const int64_t n = sets[s].ptr[n_jobs];
    for (int64_t i = 0; i < n; ++i) ids[at + i] = (int32_t)sets[s].ids[i];
    at += (uint32_t)n;
  }
  for (uint32_t j = 0; j < n_jobs; ++j) {
    dsm_raw_job& r = raw[j];
    r.max_price = c->max_price[j];
    r.cpu[0] = c->cpu[2 * j]; r.cpu[1] = c->cpu[2 * j + 1];
    r.memory[0] = c->memory[2 * j]; r.memory[1] = c->memory[2 * j + 1];
    r.disk[0] = c->disk[2 * j]; r.disk[1] = c->disk[2 * j + 1];
    r.gpu_count[0] = c->gpu_count[2 * j]; r.gpu_count[1] = c->gpu_count[2 * j + 1];
    r.gpu_memory[0] = c->gpu_memory[2 * j]; r.gpu_memory[1] = c->gpu_memory[2 * j + 1];
    r.gpu_total[0] = c->gpu_total[2 * j]; r.gpu_total[1] = c->gpu_total[2 * j + 1];
    r.cc_min = (int32_t)c->cc_min[j];
    const int64_t b = c->blocks ? c->blocks[j] : 0;
    if (b > 0xFFFF) return fail(DSM_E_ARG, "block counts above 65535 are not supported");
    r.blocks = (int32_t)b;
    r.src_backend_mask = c->src_backend_mask[j];
    r.off_backend_mask = c->off_backend_mask[j];
    r.flag_req = c->flag_req ? c->flag_req[j] : 0;
    r.flag_forbid = c->flag_forbid ? c->flag_forbid[j] : 0;
    r.status_mask = c->status_mask ? c->status_mask[j] : 0xFF;
    r.spot = c->spot[j];
    r.gpu_vendor = c->gpu_vendor[j];
    uint8_t fl = 0;
    if (c->exclude_not_available && c->exclude_not_available[j]) fl |= DSM_RAWF_EXCLUDE_NOT_AVAILABLE;
    if (!c->gpu_name_any || c->gpu_name_any[j]) fl |= DSM_RAWF_GPU_NAME_ANY;
    if (!c->region_any || c->region_any[j]) fl |= DSM_RAWF_REGION_ANY;
    if (!c->itype_any || c->itype_any[j]) fl |= DSM_RAWF_ITYPE_ANY;
    if (!c->fleet_any || c->fleet_any[j]) fl |= DSM_RAWF_FLEET_ANY;
    if (!c->name_any || c->name_any[j]) fl |= DSM_RAWF_NAME_ANY;
    if (!c->zone_any || c->zone_any[j]) fl |= DSM_RAWF_ZONE_ANY;
    r.flags = fl;
    r.fleet_id = c->fleet_ids ? (int32_t)c->fleet_ids[j] : -1;
    r.name_id = c->name_ids ? (int32_t)c->name_ids[j] : -1;
    uint32_t* oc[4][2] = {{&r.gpu_name_off, &r.gpu_name_cnt}, {&r.region_off, &r.region_cnt},
                          {&r.itype_off, &r.itype_cnt}, {&r.zone_off, &r.zone_cnt}};
    for (int s = 0; s < 4; ++s) {
      *oc[s][0] = base[s] + (sets[s].ptr ? (uint32_t)sets[s].ptr[j] : 0u);
      *oc[s][1] = sets[s].ptr ? (uint32_t)(sets[s].ptr[j + 1] - sets[s].ptr[j]) : 0u;
    }
  }
  return match_topk_raw_locked(ctx, t, layout, raw, n_jobs, ids, (uint32_t)n_ids, k, nullptr, out_cnt, nullptr, out_src,
                               out_max_price);
}

int dsm_match_indices(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const uint32_t* jobs,
                      uint32_t n_jobs, int by_source, uint64_t* out_ptr, const uint32_t** out_ids) {
  if (!ctx || !t || !layout || !jobs || !out_ptr || !out_ids) return fail(DSM_E_ARG, "NULL argument");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  if (layout->rec_words < 3 || layout->rec_words > 65536) return fail(DSM_E_ARG, "layout.rec_words out of range");
  if (!t->row_to_source) return fail(DSM_E_ARG, "the table has no source map: call dsm_table_set_source_map first");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  int rc;
  const uint32_t n_words = t->desc.n_rows / 32u;
  const uint32_t s_words = (t->n_source + 31u) / 32u;
  const size_t jb = (size_t)n_jobs * layout->rec_words * 4;
  if ((rc = ensure(ctx->d_jobs, jb))) return rc;
  if ((rc = ensure(ctx->d_bits, (size_t)n_jobs * n_words * 4))) return rc;
  if ((rc = ensure(ctx->d_cnt64, (size_t)n_jobs * 8))) return rc;
  if ((rc = ensure(ctx->d_ptr64, ((size_t)n_jobs + 1) * 8))) return rc;
  CK(cudaMemcpyAsync(ctx->d_jobs.p, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t* d_order = nullptr;
  if ((rc = upload_order(ctx, layout, n_jobs, &d_order))) return rc;
  rc = run_match(ctx, t, layout, (const uint32_t*)ctx->d_jobs.p, d_order, n_jobs, 1, true, nullptr, nullptr, nullptr,
                 (uint32_t*)ctx->d_bits.p);
  if (rc) return rc;
  const uint32_t* bits = (const uint32_t*)ctx->d_bits.p;
  uint32_t words = n_words;
  const int32_t* map = t->row_to_source;
  if (by_source) {  // re-file every match under its source position, then list the set bits in order
    if ((rc = ensure(ctx->d_sbits, (size_t)n_jobs * s_words * 4 + 4))) return rc;
    CK(cudaMemsetAsync(ctx->d_sbits.p, 0, (size_t)n_jobs * s_words * 4, ctx->stream));
    const size_t total = (size_t)n_jobs * n_words;
    k_scatter_to_source<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(
        bits, n_jobs, n_words, t->row_to_source, (uint32_t*)ctx->d_sbits.p, s_words);
    CK(cudaGetLastError());
    ctx->launches++;
    bits = (const uint32_t*)ctx->d_sbits.p;
    words = s_words;
    map = nullptr;
  }
  k_count_bits<<<(n_jobs + 7) / 8, 256, 0, ctx->stream>>>(bits, n_jobs, words, (unsigned long long*)ctx->d_cnt64.p);
  k_prefix_u64<<<1, 1024, 0, ctx->stream>>>((const unsigned long long*)ctx->d_cnt64.p, n_jobs,
                                             (unsigned long long*)ctx->d_ptr64.p);
  CK(cudaGetLastError());
  ctx->launches += 2;
  CK(cudaMemcpyAsync(out_ptr, ctx->d_ptr64.p, ((size_t)n_jobs + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  const size_t total_ids = (size_t)out_ptr[n_jobs];
  if ((rc = ensure(ctx->d_out_ids, (total_ids + 1) * 4))) return rc;
  if (total_ids * 4 > ctx->h_out_cap) {
    if (ctx->h_out_ids) CK(cudaFreeHost(ctx->h_out_ids));
    ctx->h_out_ids = nullptr;
    ctx->h_out_cap = 0;
    const size_t want = total_ids * 4 + total_ids + 4096;
    CK(cudaHostAlloc(&ctx->h_out_ids, want, cudaHostAllocDefault));
    ctx->h_out_cap = want;
  }
  if (total_ids) {
    k_fill_indices<<<(n_jobs + 7) / 8, 256, 0, ctx->stream>>>(bits, n_jobs, words, map,
                                                              (const unsigned long long*)ctx->d_ptr64.p,
                                                              (uint32_t*)ctx->d_out_ids.p);
    CK(cudaGetLastError());
    ctx->launches++;
    CK(cudaMemcpyAsync(ctx->h_out_ids, ctx->d_out_ids.p, total_ids * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  *out_ids = (const uint32_t*)ctx->h_out_ids;
  return 0;
}

int dsm_match_topk_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                       const uint32_t* d_jobs, uint32_t n_jobs, uint32_t k, uint32_t* d_idx,
                       uint32_t* d_cnt, uint32_t* d_max) {
  if (!ctx || !t || !layout || !d_jobs || !d_idx || !d_cnt || !d_max) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  return run_match(ctx, t, layout, d_jobs, layout->job_order, n_jobs, k, false, d_idx, d_cnt, d_max, nullptr);
}

int dsm_match_bitmap_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                         const uint32_t* d_jobs, uint32_t n_jobs, uint32_t* d_bits) {
  if (!ctx || !t || !layout || !d_jobs || !d_bits) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  return run_match(ctx, t, layout, d_jobs, layout->job_order, n_jobs, 1, true, nullptr, nullptr, nullptr, d_bits);
}

int dsm_match_topk(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const uint32_t* jobs,
                   uint32_t n_jobs, uint32_t k, uint32_t* out_idx, uint32_t* out_cnt,
                   uint32_t* out_max) {
  if (!ctx || !t || !layout || !jobs || !out_idx || !out_cnt || !out_max) return fail(DSM_E_ARG, "NULL argument");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  if (k == 0 || k > 4096) return fail(DSM_E_ARG, "k must be in 1..4096");
  if (layout->rec_words < 3 || layout->rec_words > 65536) return fail(DSM_E_ARG, "layout.rec_words out of range");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  int rc;
  const size_t jb = (size_t)n_jobs * layout->rec_words * 4;
  if ((rc = ensure(ctx->d_jobs, jb))) return rc;
  if ((rc = ensure(ctx->d_idx, (size_t)n_jobs * k * 4))) return rc;
  if ((rc = ensure(ctx->d_cnt, (size_t)n_jobs * 4))) return rc;
  if ((rc = ensure(ctx->d_max, (size_t)n_jobs * 4))) return rc;
  CK(cudaMemcpyAsync(ctx->d_jobs.p, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t* d_order = nullptr;
  if ((rc = upload_order(ctx, layout, n_jobs, &d_order))) return rc;
  rc = run_match(ctx, t, layout, (const uint32_t*)ctx->d_jobs.p, d_order, n_jobs, k, false,
                 (uint32_t*)ctx->d_idx.p, (uint32_t*)ctx->d_cnt.p, (uint32_t*)ctx->d_max.p, nullptr);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_idx, ctx->d_idx.p, (size_t)n_jobs * k * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(out_cnt, ctx->d_cnt.p, (size_t)n_jobs * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(out_max, ctx->d_max.p, (size_t)n_jobs * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_match_bitmap(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const uint32_t* jobs,
                     uint32_t n_jobs, uint32_t* out_bits) {
  if (!ctx || !t || !layout || !jobs || !out_bits) return fail(DSM_E_ARG, "NULL argument");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  if (layout->rec_words < 3 || layout->rec_words > 65536) return fail(DSM_E_ARG, "layout.rec_words out of range");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  int rc;
  const size_t jb = (size_t)n_jobs * layout->rec_words * 4;
  const size_t bb = (size_t)n_jobs * (t->desc.n_rows / 32) * 4;
  if ((rc = ensure(ctx->d_jobs, jb))) return rc;
  if ((rc = ensure(ctx->d_bits, bb))) return rc;
  CK(cudaMemcpyAsync(ctx->d_jobs.p, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t* d_order = nullptr;
  if ((rc = upload_order(ctx, layout, n_jobs, &d_order))) return rc;
  rc = run_match(ctx, t, layout, (const uint32_t*)ctx->d_jobs.p, d_order, n_jobs, 1, true, nullptr, nullptr,
                 nullptr, (uint32_t*)ctx->d_bits.p);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_bits, ctx->d_bits.p, bb, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static int merge_shards(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                        const uint32_t* d_idx_all, const uint32_t* d_cnt_all, const uint32_t* d_max_all,
                        const double* d_maxprice_all, size_t shard_stride_bytes,
                        const uint64_t* d_shard_base, uint64_t* d_idx_out, uint64_t* d_cnt_out,
                        uint64_t* d_max_out, double* d_maxprice_out) {
  if (!ctx || !d_idx_all || !d_cnt_all || !d_max_all || !d_maxprice_all || !d_shard_base ||
      !d_idx_out || !d_cnt_out || !d_max_out || !d_maxprice_out)
    return fail(DSM_E_ARG, "NULL argument");
  if (n_shards == 0 || n_jobs == 0 || k == 0) return fail(DSM_E_ARG, "empty merge");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  k_merge_shards<<<(n_jobs + 127) / 128, 128, 0, ctx->stream>>>(
      n_shards, n_jobs, k, d_idx_all, d_cnt_all, d_max_all, d_maxprice_all, shard_stride_bytes,
      d_shard_base, d_idx_out, d_cnt_out, d_max_out, d_maxprice_out);
  CK(cudaGetLastError());
  ctx->launches++;
  return 0;
}

int dsm_merge_shards_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                         const uint32_t* d_idx_all, const uint32_t* d_cnt_all,
                         const uint32_t* d_max_all, const double* d_maxprice_all,
                         const uint64_t* d_shard_base, uint64_t* d_idx_out, uint64_t* d_cnt_out,
                         uint64_t* d_max_out, double* d_maxprice_out) {
  return merge_shards(ctx, n_shards, n_jobs, k, d_idx_all, d_cnt_all, d_max_all, d_maxprice_all, 0,
                      d_shard_base, d_idx_out, d_cnt_out, d_max_out, d_maxprice_out);
}

int dsm_merge_shards_strided_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                                 const uint32_t* d_idx0, const uint32_t* d_cnt0, const uint32_t* d_max0,
                                 const double* d_maxprice0, size_t shard_stride_bytes,
                                 const uint64_t* d_shard_base, uint64_t* d_idx_out, uint64_t* d_cnt_out,
                                 uint64_t* d_max_out, double* d_maxprice_out) {
  if (shard_stride_bytes == 0 || shard_stride_bytes % 8) return fail(DSM_E_ARG, "shard stride must be a non-zero multiple of 8");
  return merge_shards(ctx, n_shards, n_jobs, k, d_idx0, d_cnt0, d_max0, d_maxprice0, shard_stride_bytes,
                      d_shard_base, d_idx_out, d_cnt_out, d_max_out, d_maxprice_out);
}

int dsm_merge_shards_interleaved_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                                     const uint32_t* d_idx0, const uint32_t* d_cnt0, const uint32_t* d_max0,
                                     const double* d_maxprice0, size_t shard_stride_bytes,
                                     const uint32_t* d_shard_geom, uint32_t chunk_rows,
                                     uint64_t* d_idx_out, uint64_t* d_cnt_out, uint64_t* d_max_out,
                                     double* d_maxprice_out) {
  if (!ctx || !d_idx0 || !d_cnt0 || !d_max0 || !d_maxprice0 || !d_shard_geom || !d_idx_out || !d_cnt_out ||
      !d_max_out || !d_maxprice_out)
    return fail(DSM_E_ARG, "NULL argument");
  if (n_shards == 0 || n_shards > 32) return fail(DSM_E_ARG, "interleaved merge takes 1..32 shards");
  if (n_jobs == 0 || k == 0 || k > 255) return fail(DSM_E_ARG, "interleaved merge needs n_jobs > 0 and k in 1..255");
  if (chunk_rows == 0) return fail(DSM_E_ARG, "chunk_rows must be > 0");
  if (shard_stride_bytes == 0 || shard_stride_bytes % 8) return fail(DSM_E_ARG, "shard stride must be a non-zero multiple of 8");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  k_merge_shards_interleaved<<<(n_jobs + 127) / 128, 128, 0, ctx->stream>>>(
      n_shards, n_jobs, k, d_idx0, d_cnt0, d_max0, d_maxprice0, shard_stride_bytes,
      reinterpret_cast<const uint2*>(d_shard_geom), chunk_rows, d_idx_out, d_cnt_out, d_max_out, d_maxprice_out);
  CK(cudaGetLastError());
  ctx->launches++;
  return 0;
}

int dsm_gather_price_dev(dsm_ctx* ctx, const dsm_table* t, const uint32_t* d_rows, uint32_t n,
                         double* d_out) {
  if (!ctx || !t || !d_rows || !d_out || n == 0) return fail(DSM_E_ARG, "bad argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  k_gather_price<<<(n + 255) / 256, 256, 0, ctx->stream>>>(t->price, d_rows, n, d_out);
  CK(cudaGetLastError());
  ctx->launches++;
  return 0;
}

int dsm_reserve(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, uint32_t n_jobs, uint32_t k) {
  if (!ctx || !t || !layout) return fail(DSM_E_ARG, "NULL argument");
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  if (k > 4096) return fail(DSM_E_ARG, "k must be in 0..4096");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  Plan pl;
  int rc = make_plan(ctx, t, layout, n_jobs, pl);
  if (rc) return rc;
  if ((rc = ensure_match_scratch(ctx, pl, k == 0))) return rc;
  // buffers of the host-buffer and raw-job entry points
  if ((rc = ensure(ctx->d_jobs, (size_t)n_jobs * layout->rec_words * 4))) return rc;
  if ((rc = ensure(ctx->d_order, (size_t)n_jobs * 4))) return rc;
  if ((rc = ensure(ctx->d_raw, (size_t)n_jobs * sizeof(dsm_raw_job)))) return rc;
  if ((rc = ensure(ctx->d_keys, (size_t)n_jobs * 4))) return rc;
  if (k) {
    if ((rc = ensure(ctx->d_idx, (size_t)n_jobs * k * 4))) return rc;
    if ((rc = ensure(ctx->d_cnt, (size_t)n_jobs * 4))) return rc;
    if ((rc = ensure(ctx->d_max, (size_t)n_jobs * 4))) return rc;
    if ((rc = ensure(ctx->d_out_ids, (size_t)n_jobs * k * 4 + (size_t)n_jobs * 8 + 8))) return rc;
  } else {
    if ((rc = ensure(ctx->d_bits, (size_t)n_jobs * (t->desc.n_rows / 32) * 4))) return rc;
  }
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_sync(dsm_ctx* ctx) {
  if (!ctx) return fail(DSM_E_ARG, "NULL ctx");
  USE_DEVICE(ctx->device);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_host_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) return fail(DSM_E_ARG, "bad argument");
  CK(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
  return 0;
}

int dsm_host_free(void* p) {
  if (p) CK(cudaFreeHost(p));
  return 0;
}

uint64_t dsm_launch_count(const dsm_ctx* ctx) { return ctx ? ctx->launches : 0; }

int dsm_timing_enable(dsm_ctx* ctx, int on) {
  if (!ctx) return fail(DSM_E_ARG, "NULL ctx");
  std::lock_guard<std::mutex> g(ctx->mu);
  ctx->timing = on != 0;
  ctx->timing_used = 0;
  return 0;
}

int dsm_timing_read(dsm_ctx* ctx, double* total_ms, uint32_t* n_launches) {
  if (!ctx || !total_ms || !n_launches) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  CK(cudaStreamSynchronize(ctx->stream));
  double sum = 0.0;
  for (size_t i = 0; i < ctx->timing_used; ++i) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, ctx->timing_events[i].e[1], ctx->timing_events[i].e[2]));
    sum += ms;
  }
  *total_ms = sum;
  *n_launches = (uint32_t)ctx->timing_used;
  ctx->timing_used = 0;
  return 0;
}

int dsm_timing_read_parts(dsm_ctx* ctx, double parts_ms[3], uint32_t* n_calls) {
  if (!ctx || !parts_ms || !n_calls) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  CK(cudaStreamSynchronize(ctx->stream));
  parts_ms[0] = parts_ms[1] = parts_ms[2] = 0.0;
  for (size_t i = 0; i < ctx->timing_used; ++i) {
    const StepEvents& ev = ctx->timing_events[i];
    for (int q = 0; q + 1 < ev.n; ++q) {
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, ev.e[q], ev.e[q + 1]));
      parts_ms[q] += ms;
    }
  }
  *n_calls = (uint32_t)ctx->timing_used;
  ctx->timing_used = 0;
  return 0;
}

// ------------------------------------------------------------------------------------------
// dsm_match_bitmap_static: the bitmap match without any allocation
// ------------------------------------------------------------------------------------------
//
// The other entry points grow their scratch with cudaMalloc on first use.  Some hosts of this
// library cannot do that at call time: a scheduler thread that shares the GPU with serving or
// training work must not run into the device-wide synchronisation and driver locks that
// cudaMalloc / cudaFree can take while other streams are busy, and a process whose device memory
// is already handed out in full to a pool allocator (a framework's caching allocator, a fixed
// MIG slice) may simply be refused a new allocation.  For them the scratch of this variant --
// the acceptance tables k_build_tables writes and k_match reads, and the work counter -- lives
// in a fixed array in the library's device image.  Every GPU that loads the library carries the
// array until the process exits; DSM_STATIC_SCRATCH_BYTES is its size and the hard limit.
//
// g_static_scratch names the array for DEVICE code.  In host code the same name is only the
// host-side handle of the symbol, not a device address -- a kernel handed that value would fault
// or touch unrelated memory -- so the address the kernels get is looked up once per device with
// cudaGetSymbolAddress while that device is current.
static __device__ __align__(256) unsigned char g_static_scratch[DSM_STATIC_SCRATCH_BYTES];

#define DSM_STATIC_MAX_DEVICES 64
static std::mutex g_static_mu;
static void* g_static_dev_ptr[DSM_STATIC_MAX_DEVICES];  // device address of the array, per GPU

// Caller has made `device` current.
static int static_scratch_address(int device, unsigned char** out) {
  if (device < 0 || device >= DSM_STATIC_MAX_DEVICES)
    return fail(DSM_E_ARG, "device ordinal beyond the static scratch table");
  std::lock_guard<std::mutex> g(g_static_mu);
  if (g_static_dev_ptr[device] == nullptr) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_static_scratch));
    g_static_dev_ptr[device] = p;
  }
  *out = static_cast<unsigned char*>(g_static_dev_ptr[device]);
  return 0;
}

int dsm_match_bitmap_static(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                            const uint32_t* d_jobs, uint32_t n_jobs, uint32_t* d_bits) {
  if (!ctx || !t || !layout || !d_jobs || !d_bits) return fail(DSM_E_ARG, "NULL argument");
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  std::lock_guard<std::mutex> g(ctx->mu);
  Plan pl;
  int rc = make_plan(ctx, t, layout, n_jobs, pl);  // pure arithmetic, allocates nothing
  if (rc) return rc;
  const bool row32 = t->desc.n_cols == 16;

  // layout of the array: [0,256) work counter, then one word per job group (columns the group
  // needs, written and read by k_match), then the tables [groups][entries] x 16 bytes
  const size_t nact_bytes = ((size_t)pl.n_groups * 4 + 255) / 256 * 256;
  const size_t table_bytes = (size_t)pl.n_groups * pl.lay.n_entries * 16;
  if (256 + nact_bytes + table_bytes > DSM_STATIC_SCRATCH_BYTES) {
    char msg[256];
    snprintf(msg, sizeof msg,
             "dsm_match_bitmap_static: %u job groups x %u dictionary entries need %zu bytes of "
             "scratch, the built-in array holds %u; split the batch or use dsm_match_bitmap_dev",
             pl.n_groups, pl.lay.n_entries, 256 + nact_bytes + table_bytes, (unsigned)DSM_STATIC_SCRATCH_BYTES);
    return fail(DSM_E_TOO_LARGE, "%s", msg);
  }
  USE_DEVICE(ctx->device);
  unsigned char* scratch = nullptr;
  if ((rc = static_scratch_address(ctx->device, &scratch))) return rc;

  BuildParams bp;
  MatchParams mp;
  bool has_lut;
  fill_params(t, layout, pl, d_jobs, layout->job_order, n_jobs, reinterpret_cast<uint4*>(scratch + 256 + nact_bytes),
              reinterpret_cast<uint32_t*>(scratch) /* k_build_tables zeroes it */,
              reinterpret_cast<uint32_t*>(scratch + 256), bp, mp, has_lut);
  mp.bits = d_bits;
  k_build_tables<<<build_grid(pl), 256, 0, ctx->stream>>>(bp);
  CK(cudaGetLastError());
  ctx->launches++;
  cudaError_t e = launch_match<MODE_BITMAP>(row32, mp, has_lut, false, pl.grid, pl.smem, ctx->stream);
  if (e != cudaSuccess) return fail(DSM_E_CUDA, "k_match launch: %s", cudaGetErrorString(e));
  ctx->launches++;
  return 0;
}

#ifdef DSM_CTA_TIMING
// debug builds only: start / end globaltimer of every job of the last k_extract
int dsm_debug_job_times(dsm_ctx* ctx, unsigned long long* out, uint32_t n_jobs) {
  USE_DEVICE(ctx->device);
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaMemcpyFromSymbol(out, g_job_times, (size_t)n_jobs * 2 * 8));
  return 0;
}
// debug builds only: copy the per-CTA (start ns, end ns, SM id) triples of the last k_match
int dsm_debug_cta_times(dsm_ctx* ctx, unsigned long long* out, uint32_t n_ctas) {
  USE_DEVICE(ctx->device);
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaMemcpyFromSymbol(out, g_cta_times, (size_t)n_ctas * 3 * 8));
  return 0;
}
#endif

int dsm_plan_query(const dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, uint32_t n_jobs,
                   uint32_t k, uint32_t out[8]) {
  if (!ctx || !t || !layout || !out) return fail(DSM_E_ARG, "NULL argument");
  (void)k;  // the geometry of a call does not depend on k
  Plan pl;
  int rc = make_plan(ctx, t, layout, n_jobs, pl);
  if (rc) return rc;
  out[0] = pl.grid;
  out[1] = pl.pc.count();
  out[2] = pl.pc.pa;
  out[3] = pl.pc.pb;
  out[4] = pl.slots;
  out[5] = pl.smem;
  out[6] = pl.lay.n_entries;
  out[7] = pl.n_groups;
  return 0;
}

// ------------------------------------------------------------------------------------------
// peer exchange (exchange.cuh): the sharded step's all-gather over IPC-mapped peer memory
// ------------------------------------------------------------------------------------------

}  // extern "C"

struct dsm_xchg {
  dsm_ctx* ctx;
  uint32_t world, rank;
  uint64_t slot_bytes;          // multiple of 256
  char* local = nullptr;        // [2][world][slot_bytes] | flags [2][world] | ticket | err
  size_t flags_off, ticket_off, err_off, total;
  XchgPeers peers;
  bool opened[DSM_XCHG_MAX_WORLD];  // mapped with cudaIpcOpenMemHandle (to be closed)
  bool connected = false;
  uint32_t epoch = 0;
};

static void xchg_wire(dsm_xchg* x, uint32_t r, char* base) {
  x->peers.data[r] = base;
  x->peers.flags[r] = reinterpret_cast<uint32_t*>(base + x->flags_off);
}

extern "C" {

int dsm_xchg_create(dsm_ctx* ctx, uint32_t world, uint32_t rank, uint64_t bytes_per_rank, dsm_xchg** out,
                    void* ipc_handle_out) {
  if (!ctx || !out) return fail(DSM_E_ARG, "NULL argument");
  if (world == 0 || world > DSM_XCHG_MAX_WORLD || rank >= world) return fail(DSM_E_ARG, "exchange takes 1..32 ranks");
  if (bytes_per_rank == 0 || bytes_per_rank % 16) return fail(DSM_E_ARG, "bytes_per_rank must be a non-zero multiple of 16");
  static_assert(sizeof(cudaIpcMemHandle_t) == DSM_IPC_HANDLE_BYTES, "IPC handle size");
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  dsm_xchg* x = new (std::nothrow) dsm_xchg();
  if (!x) return fail(DSM_E_ARG, "out of host memory");
  x->ctx = ctx;
  x->world = world;
  x->rank = rank;
  x->slot_bytes = (bytes_per_rank + 255) / 256 * 256;
  x->flags_off = 2 * (size_t)world * x->slot_bytes;
  x->ticket_off = x->flags_off + 256;  // 2 * 32 flag words fit in 256 bytes
  x->err_off = x->ticket_off + 256;
  x->total = x->err_off + 256;
  memset(&x->peers, 0, sizeof x->peers);
  memset(x->opened, 0, sizeof x->opened);
  // a plain cudaMalloc of its own: IPC handles cover whole allocations
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&x->local), x->total);
  if (e != cudaSuccess) {
    delete x;
    return fail(DSM_E_CUDA, "cudaMalloc (exchange): %s", cudaGetErrorString(e));
  }
  e = cudaMemset(x->local + x->flags_off, 0, x->total - x->flags_off);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e == cudaSuccess && ipc_handle_out)
    e = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(ipc_handle_out), x->local);
  if (e != cudaSuccess) {
    cudaFree(x->local);
    delete x;
    return fail(DSM_E_CUDA, "exchange setup: %s", cudaGetErrorString(e));
  }
  xchg_wire(x, rank, x->local);
  *out = x;
  return 0;
}

int dsm_xchg_local_base(dsm_xchg* x, void** base) {
  if (!x || !base) return fail(DSM_E_ARG, "NULL argument");
  *base = x->local;
  return 0;
}

int dsm_xchg_connect_ipc(dsm_xchg* x, const void* handles) {
  if (!x || !handles) return fail(DSM_E_ARG, "NULL argument");
  if (x->connected) return fail(DSM_E_ARG, "exchange already connected");
  std::lock_guard<std::mutex> g(x->ctx->mu);
  USE_DEVICE(x->ctx->device);
  const cudaIpcMemHandle_t* h = static_cast<const cudaIpcMemHandle_t*>(handles);
  for (uint32_t r = 0; r < x->world; ++r) {
    if (r == x->rank) continue;
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (uint32_t q = 0; q < r; ++q)
        if (x->opened[q]) { cudaIpcCloseMemHandle(x->peers.data[q]); x->opened[q] = false; }
      (void)cudaGetLastError();
      return fail(DSM_E_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
    }
    x->opened[r] = true;
    xchg_wire(x, r, static_cast<char*>(p));
  }
  x->connected = true;
  return 0;
}

int dsm_xchg_connect_ptrs(dsm_xchg* x, void* const* bases) {
  if (!x || !bases) return fail(DSM_E_ARG, "NULL argument");
  if (x->connected) return fail(DSM_E_ARG, "exchange already connected");
  for (uint32_t r = 0; r < x->world; ++r) {
    if (r == x->rank) continue;
    if (!bases[r]) return fail(DSM_E_ARG, "NULL peer base");
    xchg_wire(x, r, static_cast<char*>(bases[r]));
  }
  x->connected = true;
  return 0;
}

int dsm_xchg_all_gather_dev(dsm_xchg* x, const void* d_src, uint64_t bytes, const void** d_gathered,
                            uint64_t* slot_stride_bytes) {
  if (!x || !d_src || !d_gathered) return fail(DSM_E_ARG, "NULL argument");
  if (!x->connected && x->world > 1) return fail(DSM_E_ARG, "exchange not connected");
  if (bytes == 0 || bytes % 16 || bytes > x->slot_bytes) return fail(DSM_E_ARG, "bytes must be a multiple of 16 within the slot");
  if (reinterpret_cast<uintptr_t>(d_src) % 16) return fail(DSM_E_ARG, "source must be 16-byte aligned");
  dsm_ctx* ctx = x->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  USE_DEVICE(ctx->device);
  const uint32_t epoch = ++x->epoch, parity = epoch & 1;
  XchgPushParams p;
  p.peers = x->peers;
  p.src = static_cast<const uint4*>(d_src);
  p.n_vec = (uint32_t)(bytes / 16);
  p.world = x->world;
  p.rank = x->rank;
  p.parity = parity;
  p.epoch = epoch;
  p.slot_bytes = x->slot_bytes;
  p.ticket = reinterpret_cast<uint32_t*>(x->local + x->ticket_off);
  uint32_t chunks = (p.n_vec + 1023) / 1024;
  const uint32_t most = std::max(1u, 2u * (uint32_t)ctx->sm_count / x->world);
  if (chunks > most) chunks = most;
  k_xchg_push<<<dim3(chunks, x->world), 256, 0, ctx->stream>>>(p);
  CK(cudaGetLastError());
  k_xchg_wait<<<1, 32, 0, ctx->stream>>>(reinterpret_cast<const uint32_t*>(x->local + x->flags_off), x->world, parity,
                                          epoch, 5ull * 1000 * 1000 * 1000,
                                          reinterpret_cast<uint32_t*>(x->local + x->err_off));
  CK(cudaGetLastError());
  ctx->launches += 2;
  *d_gathered = x->local + (size_t)parity * x->world * x->slot_bytes;
  if (slot_stride_bytes) *slot_stride_bytes = x->slot_bytes;
  return 0;
}

int dsm_xchg_status(dsm_xchg* x, uint32_t* waited_for_rank_plus_1) {
  if (!x || !waited_for_rank_plus_1) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(x->ctx->mu);
  USE_DEVICE(x->ctx->device);
  CK(cudaMemcpyAsync(waited_for_rank_plus_1, x->local + x->err_off, 4, cudaMemcpyDeviceToHost, x->ctx->stream));
  CK(cudaStreamSynchronize(x->ctx->stream));
  return 0;
}

int dsm_xchg_destroy(dsm_xchg* x) {
  if (!x) return 0;
  DeviceGuard guard_(x->ctx->device);
  cudaStreamSynchronize(x->ctx->stream);
  for (uint32_t r = 0; r < x->world; ++r)
    if (x->opened[r]) cudaIpcCloseMemHandle(x->peers.data[r]);
  cudaFree(x->local);
  delete x;
  return 0;
}

}  // extern "C"
