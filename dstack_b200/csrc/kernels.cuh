// kernels.cuh -- device code of libdsmatch (sm_100a).  Included by dsmatch.cu only.
//
//   k_build_tables  job records -> bit-sliced acceptance tables (one u32 per 32 jobs per entry)
//   k_match         rows x job groups -> match counts per slot and per 15-block window (top-k
//                   calls), or every match word job-major (bitmap calls)
//   k_extract       per job: first k matching rows, total, last match, from those counts
//   k_merge_shards* per-job merge of all-gathered per-GPU results
//
// Work decomposition of k_match.  The rows of the table are cut into P pieces per job group
// (pa pieces in the "available" segment, pb in the rest; a piece never straddles the segment
// boundary).  The unit of work is (piece position, job group); CTAs draw units from an atomic
// counter in position-major order, so every CTA sees cheap and expensive rows and dense and
// sparse job groups alike and the SMs finish together.  Inside a unit the piece is split evenly
// over the 8 warps; warp w's results go to slot position*8 + w of each of the group's 128 jobs,
// so a job's slots are in row order and k_extract finds "the first k matches" by walking them.
//
// Top-k calls do not transpose every block.  Each thread (= row lane) adds its match bits into
// 4-bit vertical counters, one per job bit; after 15 blocks (a window) the four counter planes
// are transposed once, which gives every job its match count in the window and the set of row
// lanes (row mod 32) that matched.  k_extract then re-evaluates, for one job at a time, only
// the candidate rows of the few windows that hold its first k matches and its last match.

#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dsmatch.h"

#define JPG DSM_JOBS_PER_GROUP  // 128 jobs per group
#define TPG 4                   // 32-job tiles per group
#ifndef MATCH_WARPS
#define MATCH_WARPS 8
#endif
#define MATCH_THREADS (32 * MATCH_WARPS)
#ifndef MATCH_CTAS_PER_SM
#define MATCH_CTAS_PER_SM 3
#endif

// ------------------------------------------------------------------------------------------
// small device helpers
// ------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// 128-bit read-only row load.  Rows are loaded into registers two blocks ahead of their use,
// and every warp asks for the lines of its next window (15 blocks) to be brought into L2 while
// it works on the current one: with three CTAs' tables in shared memory only ~24 KB of L1 is left
// per SM, so the point of the prefetch is to pull rows out of DRAM early (measured in round 1 on
// config 3 with the table cold in DRAM: 0.226 ms without any prefetch, 0.203 ms with).
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr)
               : "memory");
  return r;
}

// 32x32 bit-matrix transpose across a warp (used interleaved, four at a time, in k_match): five
// butterfly steps; at step i a lane exchanges its word with lane^s (s = 16 >> i), rotates the
// partner's word left by s (lanes with bit s clear) or right by s (the others) and keeps its
// own bits under `keep`.  Afterwards lane j holds in bit l what lane l held in bit j.
struct XposeCtl {
  uint32_t sel[2];  // steps 0, 1 move half-words / bytes: one byte permute each
  uint32_t rot[5];
  uint32_t keep[5];
};
__device__ __forceinline__ void transpose_setup(XposeCtl& c, uint32_t lane) {
  const uint32_t lowmask[5] = {0x0000FFFFu, 0x00FF00FFu, 0x0F0F0F0Fu, 0x33333333u, 0x55555555u};
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const uint32_t s = 16u >> i;
    const bool upper = (lane & s) != 0;
    c.rot[i] = upper ? 32u - s : s;
    c.keep[i] = upper ? ~lowmask[i] : lowmask[i];
  }
  // __byte_perm(own, partner, sel): lanes with bit 16 (8) clear keep their low half-word (even
  // bytes) and take the partner's low half-word (even bytes) into the high one (odd bytes); the
  // other lanes keep the high part and take the partner's high part into the low one
  c.sel[0] = (lane & 16u) ? 0x3276u : 0x5410u;
  c.sel[1] = (lane & 8u) ? 0x3715u : 0x6240u;
}

// ------------------------------------------------------------------------------------------
// parameters
// ------------------------------------------------------------------------------------------

struct LayoutDev {
  uint32_t n_used;
  uint32_t kind[DSM_MAX_COLS];
  uint32_t card[DSM_MAX_COLS];
  uint32_t flags[DSM_MAX_COLS];
  uint32_t aux[DSM_MAX_COLS];
  uint32_t rec_off[DSM_MAX_COLS];
  uint32_t base[DSM_MAX_COLS];  // first table entry of the column inside a group's table
  uint32_t rec_words;
  uint32_t n_entries;           // entries per group table (sum of card)
};

struct BuildParams {
  LayoutDev lay;
  const uint32_t* recs;
  const uint32_t* job_order;  // evaluation slot -> job (nullptr: identity); see dsm_layout.job_order
  uint32_t n_jobs;
  const uint32_t* gpu_entries;
  const uint32_t* bk_entries;
  const uint32_t* pair_entries;
  uint4* tabs;             // [groups][n_entries] x 4 words
  uint32_t* work_counter;  // reset here for the k_match that follows in the stream
};

// count mode: a window is WIN_BLOCKS 32-row blocks; a 4-bit counter per (row lane, job) holds it
#define WIN_PLANES 4
#define WIN_BLOCKS 15
#define EXTRACT_PENDING 16  // windows k_extract queues before it evaluates their candidate rows
#ifndef EXTRACT_LANE_ROWS
#define EXTRACT_LANE_ROWS 2  // candidate rows per lane and evaluation step (their loads overlap; 4 measured slower)
#endif

// Pieces of the row axis (see the header comment); shared by k_match and k_extract.  Every warp
// of a piece gets the same whole number of windows, counted from the start of the piece's
// segment, so the eight warps of a unit reach its closing barrier together and only the last
// piece of a segment has short or empty shares.  A segment is cut into coarse pieces (wpw windows
// per warp) followed by fine ones (wpw_f): units are drawn in row order, so the last units of a
// launch are short and the CTAs run out of work nearly together, while most rows are covered by
// long units that stage their table once for many windows.
struct Pieces {
  uint32_t n_groups;
  uint32_t n_blocks;     // 32-row blocks of the table
  uint32_t split_block;  // first block of the second segment
  uint32_t pa, pb;       // pieces in the first / second segment
  uint32_t ca, cb;       // of which coarse (the first ca / cb of the segment)
  uint32_t wpw;          // windows per warp in a coarse piece
  uint32_t wpw_f;        // ... in a fine piece (<= wpw)
  __host__ __device__ uint32_t per_group() const { return pa + pb; }
  __host__ __device__ uint32_t count() const { return n_groups * (pa + pb); }
  __host__ __device__ static uint32_t windows(uint32_t blocks) { return (blocks + WIN_BLOCKS - 1) / WIN_BLOCKS; }
  // pieces needed for `wins` windows when every warp takes wpw of them
  __host__ __device__ static uint32_t pieces_for_windows(uint32_t wins, uint32_t wpw) {
    const uint32_t per_piece = wpw * MATCH_WARPS;
    return (wins + per_piece - 1) / per_piece;
  }
  // Cut a segment of `blocks` blocks: the last `fine_windows` windows (rounded up to whole coarse
  // pieces' worth of remainder) in fine pieces.  Returns the piece count, sets `coarse`.
  __host__ uint32_t cut(uint32_t blocks, uint32_t fine_windows, uint32_t& coarse) const {
    const uint32_t wins = windows(blocks);
    if (fine_windows == 0u || wpw_f == wpw) return coarse = pieces_for_windows(wins, wpw);  // coarse pieces only
    const uint32_t fine = fine_windows < wins ? fine_windows : wins;
    coarse = (wins - fine) / (wpw * MATCH_WARPS);
    return coarse + pieces_for_windows(wins - coarse * wpw * MATCH_WARPS, wpw_f);
  }
  // blocks [b0, b1) of warp `warp` of piece `pos`
  __host__ __device__ void warp_blocks(uint32_t pos, uint32_t warp, uint32_t& b0, uint32_t& b1) const {
    uint32_t seg_lo = 0, seg_hi = split_block, i = pos, c = ca;
    if (pos >= pa) { seg_lo = split_block; seg_hi = n_blocks; i = pos - pa; c = cb; }
    unsigned long long w0, len = wpw;  // first window, windows
    if (i < c) {
      w0 = ((unsigned long long)i * MATCH_WARPS + warp) * wpw;
    } else {
      w0 = (unsigned long long)c * MATCH_WARPS * wpw + ((unsigned long long)(i - c) * MATCH_WARPS + warp) * wpw_f;
      len = wpw_f;
    }
    const unsigned long long x0 = seg_lo + w0 * WIN_BLOCKS;
    const unsigned long long x1 = x0 + len * WIN_BLOCKS;
    b0 = x0 < seg_hi ? (uint32_t)x0 : seg_hi;
    b1 = x1 < seg_hi ? (uint32_t)x1 : seg_hi;
  }
};

struct MatchParams {
  const uint4* rows;
  const uint4* tabs;
  const uint32_t* recs;
  const uint32_t* job_order;  // evaluation slot -> job (nullptr: identity)
  const uint16_t* lut[DSM_MAX_COLS];
  uint32_t base[DSM_MAX_COLS];
  Pieces pc;
  uint32_t n_used;       // columns of the call
  uint32_t rec_words;
  uint32_t n_entries;
  uint32_t dead_entry;   // table entry every job rejects by construction (padding rows), or ~0
  uint32_t live_a, live_b;  // rows >= live_a in the first segment / >= live_b in the second are padding
  // shared-memory layout of a group's acceptance tables: column c's entries are stored
  // 2^col_shift[c] times each (copy r of entry e at col_off[c] + ((e << shift) + r) * 16), see k_match
  uint32_t col_off[DSM_MAX_COLS];    // bytes from the start of the table area
  uint32_t col_shift[DSM_MAX_COLS];  // log2 copies: rep_shift for every column but big_col (big_shift)
  uint32_t gbase[DSM_MAX_COLS];      // col_off / (16 << rep_shift): entry offset of a column in gather units
  uint32_t rep_shift;    // log2 copies of the columns: 3 = conflict-free gathers, 0 = one copy
  uint32_t big_col;      // column stored with fewer copies (big_shift < rep_shift) to make room, or ~0
  uint32_t big_shift;
  uint32_t tab_bytes;    // size of the table area
  uint32_t stage;        // 1: the tables arrive by TMA in a staging copy in front of the table area (or,
                         // with every col_shift 0, straight in place); 0: the threads read them from global memory
  uint32_t n_jobs;
  uint32_t slots;        // slots per job = pieces per group * 8
  uint32_t* work_counter;
  uint32_t* part_cnt;    // count mode: [jobs_pad][slots] matches per slot (indexed by evaluation slot)
  uint32_t* bits;        // bitmap mode: [n_jobs][n_blocks]
  uint32_t* win_rows;    // count mode: [groups][slots][wps][128]: per window of WIN_BLOCKS blocks and job,
                         // the row lanes (row mod 32) with at least one match
  uint32_t wps;          // count mode: windows reserved per slot
  uint32_t* slot_start;  // count mode: [slots + 1] first block of every slot, then n_blocks
  uint32_t* group_nact;  // [groups] leading columns k_match gathered for the group (k_extract re-uses it)
};

// k_match modes
#define MODE_BITMAP 1  // every match word, job-major
#define MODE_COUNT 2   // counts per slot and window only; k_extract re-evaluates the few rows it needs

struct ExtractParams {
  const uint32_t* part_cnt;
  const double* price;
  const uint32_t* job_order;
  uint32_t n_jobs, k, slots, slots_a;  // slots below slots_a belong to the first segment
  // candidate rows are evaluated for one job from the global copies of rows and tables
  const uint4* rows;
  const uint4* tabs;
  const uint32_t* recs;
  const uint16_t* lut[DSM_MAX_COLS];
  uint32_t base[DSM_MAX_COLS];
  uint32_t n_cols, n_used, n_entries, rec_words;
  uint32_t live_a, live_b;
  uint32_t* out_idx;
  uint32_t* out_cnt;
  uint32_t* out_max;
  // k_extract: the pieces k_match worked on (slot -> block range) and its per-window counts
  Pieces pc;
  const uint32_t* win_rows;
  uint32_t wps;
  const uint32_t* slot_start;
  const uint32_t* group_nact;
};

// ------------------------------------------------------------------------------------------
// k_build_tables: grid (groups*4, n_used, entry chunks), block 256 = 8 warps; lane = job of the tile,
// warps stride over the column's dictionary entries; one ballot = one table word.
// ------------------------------------------------------------------------------------------

#ifndef BUILD_CHUNK
#define BUILD_CHUNK 64u
#endif

__device__ __forceinline__ bool range_accepts(uint32_t word, uint32_t e) {
  return e >= (word & 0xFFFFu) && e <= (word >> 16);
}

__global__ void __launch_bounds__(256) k_build_tables(const BuildParams p) {
  const uint32_t tile = blockIdx.x;
  const uint32_t col = blockIdx.y;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = threadIdx.x >> 5;
  if (tile == 0 && col == 0 && blockIdx.z == 0 && threadIdx.x == 0) *p.work_counter = 0u;
  const uint32_t slot = tile * 32u + lane;
  const bool live = slot < p.n_jobs;
  const uint32_t job = live ? (p.job_order ? p.job_order[slot] : slot) : 0u;
  const uint32_t* rec = p.recs + (size_t)job * p.lay.rec_words + p.lay.rec_off[col];
  const uint32_t kind = p.lay.kind[col];
  const uint32_t card = p.lay.card[col];
  uint4* out = p.tabs + (size_t)(tile / TPG) * p.lay.n_entries + p.lay.base[col];
  const uint32_t sub = tile % TPG;

  uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
  if (live) {
    r0 = rec[0];
    if (kind == DSM_COL_GPU) { r1 = rec[1]; r2 = rec[2]; r3 = rec[3]; r4 = rec[4]; }
    if (kind == DSM_COL_BK) { r1 = rec[1]; r2 = rec[2]; r3 = rec[3]; }
  }

  // blockIdx.z cuts long dictionaries into chunks of BUILD_CHUNK entries so that the largest
  // column does not set the kernel's duration (8 entries are in flight per CTA)
  const uint32_t e_end = min(card, (blockIdx.z + 1u) * BUILD_CHUNK);
  for (uint32_t e = blockIdx.z * BUILD_CHUNK + warp; e < e_end; e += 8u) {
    bool ok = false;
    if (live) {
      if (kind == DSM_COL_RANGE) {
        ok = range_accepts(r0, e);
        if ((p.lay.flags[col] & DSM_COLF_NULL_PASSES) && e == card - 1u) ok = true;
      } else if (kind == DSM_COL_BITSET) {
        ok = (rec[e >> 5] >> (e & 31u)) & 1u;
      } else if (kind == DSM_COL_PAIR) {
        // entry = (ordinal of member A | ordinal of member B << 16); member A's predicate is a
        // range word or (DSM_COLF_PAIR_A_BITSET) a bitset of aux words, member B's a range word
        const uint32_t ent = p.pair_entries[e];
        const uint32_t a = ent & 0xFFFFu, b = ent >> 16;
        const uint32_t aw = p.lay.aux[col];  // words of member A's predicate
        if (p.lay.flags[col] & DSM_COLF_PAIR_A_BITSET) ok = (rec[a >> 5] >> (a & 31u)) & 1u;
        else ok = range_accepts(r0, a);
        ok = ok && range_accepts(rec[aw], b);
      } else if (kind == DSM_COL_GPU) {
        const uint32_t* ent = p.gpu_entries + (size_t)e * DSM_GPU_ENTRY_WORDS;
        const uint32_t e0 = ent[0], e1 = ent[1], e2 = ent[2];
        const uint32_t cnt = e0 & 0xFFFFu, mem = e0 >> 16;
        const uint32_t tot = e1 & 0xFFFFu, cc = e1 >> 16;
        const uint32_t name = e2 & 0xFFFFu, vendor = (e2 >> 16) & 0xFFu, has_gpu = (e2 >> 24) & 1u;
        const uint32_t gflags = r3 >> 16, cc_min = r3 & 0xFFFFu;
        ok = cnt >= (r0 & 0xFFFFu) && cnt <= (r0 >> 16);                       // item 6
        ok = ok && ((r4 >> vendor) & 1u);                                      // item 7
        if (!(gflags & DSM_GPUF_NAME_ANY))                                     // item 8
          ok = ok && name != 0xFFFFu && ((rec[5 + (name >> 5)] >> (name & 31u)) & 1u);
        if (gflags & DSM_GPUF_MEM_BOUNDED)                                     // item 9
          ok = ok && has_gpu && mem >= (r1 & 0xFFFFu) && mem <= (r1 >> 16);
        if (gflags & DSM_GPUF_CC_BOUNDED)                                      // item 10
          ok = ok && cc != 0xFFFFu && cc >= cc_min;
        ok = ok && tot >= (r2 & 0xFFFFu) && tot <= (r2 >> 16);                 // item 11
        const uint32_t blocks = r4 >> 16, gpus_for_blocks = e2 >> 25;          // offers.py:102-118
        if (blocks) ok = ok && (gpus_for_blocks % blocks == 0u);
      } else if (kind == DSM_COL_BK) {
        const uint32_t* ent = p.bk_entries + (size_t)e * DSM_BK_ENTRY_WORDS;
        const uint32_t e0 = ent[0], fl = ent[1];
        const uint32_t src = e0 & 0xFFu, off = (e0 >> 8) & 0xFFu, spot = (e0 >> 16) & 0xFFu,
                       status = e0 >> 24;
        ok = e != 0u;  // entry 0 is the dead entry
        ok = ok && ((r0 >> src) & 1u) && ((r1 >> off) & 1u);
        ok = ok && ((r2 >> spot) & 1u) && ((r2 >> (8u + status)) & 1u);
        const uint32_t req = r3 & 0xFFFFu, forbid = r3 >> 16;
        ok = ok && ((fl & req) == req) && ((fl & forbid) == 0u);
      }
    }
    const uint32_t word = __ballot_sync(0xFFFFFFFFu, ok);
    if (lane == 0) reinterpret_cast<uint32_t*>(out + e)[sub] = word;
  }
}

// ------------------------------------------------------------------------------------------
// k_match
// ------------------------------------------------------------------------------------------

#ifdef DSM_CTA_TIMING
// debug builds only (tools/cta_times.py): start / end globaltimer and SM id of every CTA
__device__ unsigned long long g_cta_times[4096 * 3];
__device__ unsigned long long g_job_times[65536 * 2];  // k_extract: start / end per job
#endif

template <int NC>
struct RowCodes {
  uint32_t w[(NC + 1) / 2];
};

// `rows` already points at this lane's row of the block.
template <int NC, bool ROW32>
__device__ __forceinline__ void load_row(const uint4* rows, RowCodes<NC>& rc) {
  if constexpr (!ROW32) {
    const uint4 v = ldg_stream(rows);
    rc.w[0] = v.x; rc.w[1] = v.y; rc.w[2] = v.z; rc.w[3] = v.w;
  } else {
    const uint4 v0 = ldg_stream(rows);
    const uint4 v1 = ldg_stream(rows + 1);
    rc.w[0] = v0.x; rc.w[1] = v0.y; rc.w[2] = v0.z; rc.w[3] = v0.w;
    rc.w[4] = v1.x; rc.w[5] = v1.y; rc.w[6] = v1.z; rc.w[7] = v1.w;
  }
}

template <int V>
struct IntC {
  static constexpr int value = V;
};

__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (c & (a | b)); }

// Cut-off classification of one block for the 128 jobs of a group (lane j of tile t holds job
// 32t+j's cut-off row in cut_lane[32 t]): full = jobs whose cut-off lies behind the block (they
// pass on all its rows, and on the following blocks up to `next_evt`), pw = per row lane, the jobs
// whose cut-off falls inside the block and lies behind this lane's row.  Out of line on purpose:
// it runs once per cut-off, but is reachable from every block of the unrolled windows.
struct BlockClass {
  uint32_t full[TPG];
  uint32_t pw[TPG];
  uint32_t next_evt;
  uint32_t any_part;
};
__device__ __noinline__ BlockClass classify_block(const uint32_t* cut_lane, uint32_t blk, uint32_t lane) {
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  BlockClass r;
  const uint32_t row = blk * 32u + lane;
  const uint32_t row_end = (blk + 1u) * 32u;
  uint32_t nxt = 0xFFFFFFFFu;
  r.any_part = 0;
#pragma unroll
  for (int t = 0; t < TPG; ++t) {
    const uint32_t cutv = cut_lane[t * 32];
    r.full[t] = __ballot_sync(FULL, cutv >= row_end);
    uint32_t part = __ballot_sync(FULL, cutv > blk * 32u && cutv < row_end);
    r.any_part |= part;
    r.pw[t] = 0;
    while (part) {
      const uint32_t j = __ffs(part) - 1;
      part &= part - 1;
      r.pw[t] |= (row < __shfl_sync(FULL, cutv, j)) ? (1u << j) : 0u;
    }
    if (cutv >= row_end) nxt = min(nxt, cutv >> 5);  // block where this job leaves `full`
  }
  r.next_evt = __reduce_min_sync(FULL, nxt);
  return r;
}

// NCMAX = 8 (16-byte rows) or 16 (32-byte rows); the number of columns a call uses and the number
// a job group needs are run-time values (p.n_used, n_act).
// EVFAST (count mode): the unrolled windows also take windows with a cut-off inside (a per-block
// test and an out-of-line call) instead of leaving them to the generic path.  Pays when such
// windows are frequent (few rows per job group: config 3, 0.141 against 0.145 ms) and costs when
// they are rare (8 M rows x 128 groups: 3.88 against 3.66 ms); the host picks per call.
// MIXED: one column (p.big_col) is stored with fewer copies than the others, so the gather
// address takes its stride and lane offset per column (one more instruction per gather than the
// uniform formula: 6 % on a batch that does not need it, hence a variant of its own).
template <int NCMAX, bool ROW32, int MODE, bool HAS_LUT, bool EVFAST, bool MIXED>
__global__ void __launch_bounds__(MATCH_THREADS, MATCH_CTAS_PER_SM) k_match(const MatchParams p) {
  static_assert(MODE == MODE_BITMAP || MODE == MODE_COUNT, "unknown k_match mode");
  constexpr bool COUNT = MODE == MODE_COUNT;
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  constexpr uint32_t ROW_U4 = ROW32 ? 2u : 1u;  // uint4 per row
  // shared memory: [staging copy: E entries, if p.stage][table area][cut-offs: 2 x 128]
  // In the table area every entry of column c is stored R_c = 2^col_shift[c] times, copy r of the
  // column's entry e at uint4 index (e << shift) + r.  With R = 8 copy r lies in bank group r
  // (4 banks), lane l reads copy l%8, and the 8 lanes of every quarter-warp phase of a 128-bit
  // gather hit 8 different bank groups: no bank conflicts whatever the codes are.  The host gives 8
  // copies to every column while three CTAs' tables fit in an SM's shared memory and halves the
  // copies of the largest column(s) otherwise (those gathers then conflict now and then).  The
  // tables arrive by TMA into the staging copy when there is room for one, else the threads read
  // them from global memory; with one copy of everything the TMA lands them in place.
  extern __shared__ __align__(128) unsigned char smem_raw[];
  bool in_place = p.stage != 0u;  // p.stage and a single copy of every column: layout == the global one
#pragma unroll
  for (int c = 0; c < NCMAX; ++c) in_place = in_place && (c >= (int)p.n_used || p.col_shift[c] == 0u);
  const bool staged = p.stage != 0u && !in_place;
  uint4* s_stage = reinterpret_cast<uint4*>(smem_raw);
  unsigned char* s_tab = smem_raw + (staged ? (size_t)p.n_entries * 16u : 0u);
  uint32_t* s_cut = reinterpret_cast<uint32_t*>(s_tab + p.tab_bytes);  // [2][128]
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_unit, s_nact;

  const uint32_t tid = threadIdx.x;
  const uint32_t lane = tid & 31u;
  const uint32_t warp = __shfl_sync(FULL, tid >> 5, 0);  // provably warp-uniform for the compiler
  XposeCtl xc;
  transpose_setup(xc, lane);

  // Gather address of column c for a row: lane_base (table area + this lane's copy)
  //                                      + (code + gbase[c]) * stride;
  // for the one column with fewer copies its own stride and lane offset
  const uint32_t stride = 16u << p.rep_shift;
  const uint32_t lane_base = smem_u32(s_tab) + ((lane & ((1u << p.rep_shift) - 1u)) << 4);
  const uint32_t tab_u32 = smem_u32(s_tab);
  const uint32_t lane16 = lane << 4;
  auto gather_addr = [&](const int c, const uint32_t code) -> uint32_t {
    if constexpr (MIXED) {  // per-column stride and lane offset, branch-free (measured faster than a
                            // predicated special case for the one column: 3.44 against 3.60 ms)
      const uint32_t st = 16u << p.col_shift[c];
      return tab_u32 + p.col_off[c] + code * st + (lane16 & (st - 16u));
    }
    return lane_base + (code + p.gbase[c]) * stride;
  };

  if (tid == 0) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }

  const Pieces pc = p.pc;
  const uint32_t n_units = pc.count();
  const uint32_t n_blocks = pc.n_blocks;
  const uint4* rows_lane = p.rows + ROW_U4 * lane;  // this lane's row inside block 0
  uint32_t parity = 0;
  uint32_t cur_group = 0xFFFFFFFFu;
  uint32_t n_act = p.n_used;
#ifdef DSM_CTA_TIMING
  unsigned long long dbg_t0 = 0;
  if (tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
#endif

  for (;;) {
    __syncthreads();  // the previous unit is finished: table and s_unit may be overwritten
    if (tid == 0) {
      s_unit = atomicAdd(p.work_counter, 1u);
      s_nact = 1u;
    }
    __syncthreads();
    const uint32_t unit = s_unit;
    if (unit >= n_units) break;
    const uint32_t pos = unit / pc.n_groups;  // position-major: all groups' piece 0 first
    const uint32_t group = unit - pos * pc.n_groups;

    if (group != cur_group) {
      // ---- stage this group's acceptance table (TMA bulk copy) and cut-off rows
      if (tid == 0 && p.stage) {
        const uint32_t bytes = p.n_entries * 16u;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&s_bar, bytes);
        bulk_g2s(s_stage, p.tabs + (size_t)group * p.n_entries, bytes, &s_bar);
      }
      if (tid < JPG) {
        const uint32_t slot_j = group * JPG + tid;
        uint32_t ca = 0, cb = 0;  // padding jobs accept nothing
        if (slot_j < p.n_jobs) {
          const uint32_t job = p.job_order ? p.job_order[slot_j] : slot_j;
          ca = min(p.recs[(size_t)job * p.rec_words + 0], p.live_a);  // padding rows never pass
          cb = min(p.recs[(size_t)job * p.rec_words + 1], p.live_b);
        }
        s_cut[tid] = ca;
        s_cut[JPG + tid] = cb;
      }
      // columns are ordered so that the ones jobs most often leave open come last; a column whose
      // every entry is accepted by every job of this group changes nothing, so only the columns
      // up to the last one that rejects something are gathered for this group (n_act)
      const uint32_t nj = min(JPG, p.n_jobs - group * JPG);
      uint32_t live[TPG];
#pragma unroll
      for (int t = 0; t < TPG; ++t)
        live[t] = nj >= 32u * (t + 1) ? FULL : (nj > 32u * t ? (1u << (nj - 32u * t)) - 1u : 0u);
      if (p.stage) {
        mbar_wait(&s_bar, parity);
        parity ^= 1u;
      }
      // one thread per entry: read it once (from global memory, or from the TMA-staged copy) and
      // write its copies, lane l starting at copy l so that the lanes of a store phase hit
      // different bank groups
      {
        const uint4* src = p.stage ? s_stage : p.tabs + (size_t)group * p.n_entries;
        uint32_t need = 0;
        for (uint32_t e = tid; e < p.n_entries; e += MATCH_THREADS) {
          const uint4 v = p.stage ? src[e] : __ldg(src + e);
          uint32_t c = 0;  // the entry's column
#pragma unroll
          for (int i = 1; i < NCMAX; ++i) c += (i < (int)p.n_used && e >= p.base[i]) ? 1u : 0u;
          if (!in_place) {
            const uint32_t sh = p.col_shift[c], R = 1u << sh;
            uint4* dst = reinterpret_cast<uint4*>(s_tab + p.col_off[c]) + ((size_t)(e - p.base[c]) << sh);
#pragma unroll
            for (uint32_t r = 0; r < 8u; ++r)
              if (r < R) dst[(r + lane) & (R - 1u)] = v;
          }
          const bool all = ((v.x & live[0]) == live[0]) && ((v.y & live[1]) == live[1]) &&
                           ((v.z & live[2]) == live[2]) && ((v.w & live[3]) == live[3]);
          if (!all && e != p.dead_entry) need = max(need, c + 1u);
        }
        if (need > 1u) atomicMax(&s_nact, need);
      }
      __syncthreads();
      n_act = s_nact;
      if (tid == 0) p.group_nact[group] = n_act;
      cur_group = group;
    }

    // ---- this warp's rows of the piece
    uint32_t b0, b1;
    pc.warp_blocks(pos, warp, b0, b1);
    const uint32_t slot = pos * MATCH_WARPS + warp;
    if constexpr (COUNT) {
      // k_extract reads the slots' block ranges instead of redoing this arithmetic per job:
      // slot s covers blocks [slot_start[s], slot_start[s + 1])
      if (group == 0 && lane == 0) {
        p.slot_start[slot] = b0;
        if (slot + 1 == p.slots) p.slot_start[slot + 1] = b1;
      }
    }

    // Rows below a job's cut-off pass its price / availability test.  full[t] holds the jobs
    // of tile t that pass on every row of the current block; it only changes at an "event"
    // block, the block a cut-off falls into, where the per-row comparison is done.
    const uint32_t* cut_lane = s_cut + (pos >= pc.pa ? JPG : 0u) + lane;  // + 32t: job 32t+lane's cut-off
    uint32_t full[TPG] = {0, 0, 0, 0};
    uint32_t next_evt = b0;

    // classify block `blk`: full[] for it and the blocks after it, the rows of jobs whose cut-off
    // lies inside it (pw), and the next block at which a job leaves full[]
    auto classify = [&](const uint32_t blk, uint32_t (&pw)[TPG]) -> bool {
      const BlockClass bc = classify_block(cut_lane, blk, lane);
#pragma unroll
      for (int t = 0; t < TPG; ++t) { full[t] = bc.full[t]; pw[t] = bc.pw[t]; }
      next_evt = bc.next_evt;
      return bc.any_part != 0;
    };
    auto block_mask = [&](const uint32_t blk) -> uint4 {
      if (blk == next_evt) {  // warp-uniform, once per cut-off inside this warp's rows
        uint32_t pw[TPG];
        classify(blk, pw);
        return make_uint4(full[0] | pw[0], full[1] | pw[1], full[2] | pw[2], full[3] | pw[3]);
      }
      return make_uint4(full[0], full[1], full[2], full[3]);
    };

    // gather one 128-bit word group per needed column and AND them
    auto gather_generic = [&](const RowCodes<NCMAX>& cur, uint4 m) -> uint4 {
#pragma unroll
      for (int c = 0; c < NCMAX; ++c) {
        if (c < (int)n_act) {
          uint32_t code = (c & 1) ? (cur.w[c >> 1] >> 16) : (cur.w[c >> 1] & 0xFFFFu);
          if constexpr (HAS_LUT) {
            if (p.lut[c] != nullptr) code = __ldg(p.lut[c] + code);
          }
          const uint4 a = lds128(gather_addr(c, code));
          m.x &= a.x; m.y &= a.y; m.z &= a.z; m.w &= a.w;
        }
      }
      return m;
    };

    // four independent 32x32 bit transposes, interleaved step by step so the shuffles overlap
    auto xpose4 = [&](uint32_t (&T)[TPG]) {
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        uint32_t y[TPG];
#pragma unroll
        for (int t = 0; t < TPG; ++t) y[t] = __shfl_xor_sync(FULL, T[t], 16u >> i);
#pragma unroll
        for (int t = 0; t < TPG; ++t) {
          if (i < 2) {  // half-word / byte exchange: one byte permute
            T[t] = __byte_perm(T[t], y[t], xc.sel[i]);
          } else {
            const uint32_t moved = __funnelshift_l(y[t], y[t], xc.rot[i]);
            T[t] = (T[t] & xc.keep[i]) | (moved & ~xc.keep[i]);
          }
        }
      }
    };

    const uint4* rp = rows_lane + (size_t)b0 * (32u * ROW_U4);  // this lane's row of block b0
    // ra: the current block, rb: the one after it; rc: third buffer of the unrolled windows, so
    // that a block's codes need not be copied out of the buffer its successor's load overwrites
    RowCodes<NCMAX> ra, rb, rc;
    if (b0 < b1) load_row<NCMAX, ROW32>(rp, ra);
    if (b0 + 1 < b1) load_row<NCMAX, ROW32>(rp + 32u * ROW_U4, rb);

    if constexpr (COUNT) {
      uint32_t cnt[TPG] = {0, 0, 0, 0};  // per job (lane j, tile t): matches in this slot
      uint32_t blk = b0, widx = 0;
      if (!EVFAST && b0 < b1) {
        uint32_t pw[TPG];
        if (classify(b0, pw)) next_evt = b0;  // a cut-off inside the first block: take the generic path
      }
      while (blk < b1) {
        const uint32_t nb = min((uint32_t)WIN_BLOCKS, b1 - blk);
        // ask for the next window's rows (60 lines of 128 bytes with 16-byte rows) while this one
        // is processed; lines past the end of the table are left alone
        if (blk + WIN_BLOCKS < b1) {
          constexpr uint32_t LINES = WIN_BLOCKS * 4u * ROW_U4;  // lines per window
          const uint32_t first = (blk + WIN_BLOCKS) * 4u * ROW_U4;  // line index of the next window
          const uint32_t last = n_blocks * 4u * ROW_U4;             // lines in the table
#pragma unroll
          for (uint32_t l = 0; l < LINES; l += 32u) {
            const uint32_t line = first + l + lane;
            if (l + lane < LINES && line < last)
              prefetch_l2(reinterpret_cast<const char*>(p.rows) + (size_t)line * 128u);
          }
        }
        uint32_t P[TPG][WIN_PLANES];  // bit b of the number of matches per (row lane, job) in this window
        bool fast_done = false;
        if constexpr (!ROW32) {
          if (nb == WIN_BLOCKS && (EVFAST || next_evt >= blk + WIN_BLOCKS)) {
            // ---- a whole window: the 15 blocks are unrolled and their match words summed with a
            // carry-save adder tree (11 full adders per tile and window instead of a ripple-carry
            // add per block).  full[] only changes at a block a cut-off falls into (next_evt)
            auto window = [&](auto na_tag) {
              constexpr int NA = decltype(na_tag)::value;
              uint32_t s1[TPG], p1[TPG], s2[TPG], p2[TPG], d1[TPG], d2[TPG];
#pragma unroll
              for (int i = 0; i < WIN_BLOCKS; ++i) {
                const RowCodes<NCMAX>& cur = (i % 3 == 0) ? ra : (i % 3 == 1) ? rb : rc;
                RowCodes<NCMAX>& nxt = (i % 3 == 0) ? rc : (i % 3 == 1) ? ra : rb;  // block i + 2
                if (blk + i + 2 < b1) load_row<NCMAX, ROW32>(rp + (size_t)(i + 2) * 32u * ROW_U4, nxt);
                uint4 m = EVFAST ? block_mask(blk + i) : make_uint4(full[0], full[1], full[2], full[3]);
#pragma unroll
                for (int c = 0; c < NA; ++c) {
                  uint32_t code = (c & 1) ? (cur.w[c >> 1] >> 16) : (cur.w[c >> 1] & 0xFFFFu);
                  if constexpr (HAS_LUT) {
                    if (p.lut[c] != nullptr) code = __ldg(p.lut[c] + code);
                  }
                  const uint4 a = lds128(gather_addr(c, code));
                  m.x &= a.x; m.y &= a.y; m.z &= a.z; m.w &= a.w;
                }
                const uint32_t mm[TPG] = {m.x, m.y, m.z, m.w};
#pragma unroll
                for (int t = 0; t < TPG; ++t) {
                  if (i == 0) {
                    s1[t] = mm[t];
                  } else if (i & 1) {
                    p1[t] = mm[t];
                  } else {  // full adder of the running sum, the held word and this one
                    const uint32_t c2 = maj3(s1[t], p1[t], mm[t]);
                    s1[t] ^= p1[t] ^ mm[t];
                    const int j = i / 2;  // carries 1..7 into the twos (a constant once unrolled)
                    if (j == 1) {
                      s2[t] = c2;
                    } else if ((j & 1) == 0) {
                      p2[t] = c2;
                    } else {
                      const uint32_t c4 = maj3(s2[t], p2[t], c2);
                      s2[t] ^= p2[t] ^ c2;
                      if (j == 3) d1[t] = c4;
                      else if (j == 5) d2[t] = c4;
                      else {
                        P[t][2] = d1[t] ^ d2[t] ^ c4;
                        P[t][3] = maj3(d1[t], d2[t], c4);
                      }
                    }
                  }
                }
              }
#pragma unroll
              for (int t = 0; t < TPG; ++t) { P[t][0] = s1[t]; P[t][1] = s2[t]; }
            };
            switch (n_act) {
              case 1: window(IntC<1>{}); break;
              case 2: window(IntC<2>{}); break;
              case 3: window(IntC<3>{}); break;
              case 4: window(IntC<4>{}); break;
              case 5: window(IntC<5>{}); break;
              case 6: window(IntC<6>{}); break;
              case 7: window(IntC<7>{}); break;
              default: window(IntC<8>{}); break;
            }
            fast_done = true;  // 15 = 5 x 3 blocks: ra / rb again hold the next two blocks
          }
        }
        if (!fast_done) {
          // ---- the short last window of a segment, or 32-byte rows: block by block, ripple-carry
          // add into the counter planes
#pragma unroll
          for (int t = 0; t < TPG; ++t)
#pragma unroll
            for (int b = 0; b < WIN_PLANES; ++b) P[t][b] = 0;
#pragma unroll 1
          for (uint32_t i = 0; i < nb; ++i) {
            const RowCodes<NCMAX> cur = ra;
            ra = rb;
            if (blk + i + 2 < b1) load_row<NCMAX, ROW32>(rp + (size_t)(i + 2) * 32u * ROW_U4, rb);
            const uint4 m = gather_generic(cur, block_mask(blk + i));
            const uint32_t mm[TPG] = {m.x, m.y, m.z, m.w};
#pragma unroll
            for (int t = 0; t < TPG; ++t) {
              uint32_t carry = mm[t];
#pragma unroll
              for (int b = 0; b < WIN_PLANES; ++b) {
                const uint32_t up = P[t][b] & carry;
                P[t][b] ^= carry;
                carry = up;
              }
            }
          }
        }
        // ---- the window is complete: after the transposes lane j holds, per plane, which row
        // lanes have that counter bit set for job 32t+j: popcounts give the window's matches, the
        // OR of the planes the row lanes (row mod 32) that matched at all -- k_extract looks only
        // at those rows
        {
          uint32_t c[TPG] = {0, 0, 0, 0}, any[TPG] = {0, 0, 0, 0};
#pragma unroll
          for (int b = 0; b < WIN_PLANES; ++b) {
            uint32_t T[TPG];
#pragma unroll
            for (int t = 0; t < TPG; ++t) T[t] = P[t][b];
            // the high planes are empty unless some job matched 4 (8) or more of a lane's 15 rows:
            // with selective jobs they usually are, and an empty plane needs no transpose
            if (b >= 2 && !__any_sync(FULL, (T[0] | T[1] | T[2] | T[3]) != 0u)) continue;
            xpose4(T);
#pragma unroll
            for (int t = 0; t < TPG; ++t) {
              c[t] += (uint32_t)__popc(T[t]) << b;
              any[t] |= T[t];
            }
          }
          uint32_t* w = p.win_rows + (((size_t)group * p.slots + slot) * p.wps + widx) * JPG + lane;
#pragma unroll
          for (int t = 0; t < TPG; ++t) {
            w[t * 32] = any[t];
            cnt[t] += c[t];
          }
        }
        blk += nb;
        rp += (size_t)nb * 32u * ROW_U4;
        ++widx;
      }
#pragma unroll
      for (int t = 0; t < TPG; ++t)
        p.part_cnt[(size_t)(group * JPG + t * 32 + lane) * p.slots + slot] = cnt[t];
    } else {
      // ---- bitmap mode: every block is transposed and lane j stores the 32-row match word of
      // the job evaluated in slot 32t+j
      uint32_t out_job[TPG];
#pragma unroll
      for (int t = 0; t < TPG; ++t) {
        const uint32_t s = group * JPG + t * 32 + lane;
        out_job[t] = s < p.n_jobs ? (p.job_order ? p.job_order[s] : s) : 0xFFFFFFFFu;
      }
#pragma unroll 1
      for (uint32_t blk = b0; blk < b1; ++blk) {
        const RowCodes<NCMAX> cur = ra;
        ra = rb;
        if (blk + 2 < b1) load_row<NCMAX, ROW32>(rp + 2u * 32u * ROW_U4, rb);
        rp += 32u * ROW_U4;
        const uint4 m = gather_generic(cur, block_mask(blk));
        uint32_t T[TPG] = {m.x, m.y, m.z, m.w};
        xpose4(T);
#pragma unroll
        for (int t = 0; t < TPG; ++t)
          if (out_job[t] != 0xFFFFFFFFu) p.bits[(size_t)out_job[t] * n_blocks + blk] = T[t];
      }
    }
  }
#ifdef DSM_CTA_TIMING
  if (tid == 0 && blockIdx.x < 4096) {
    unsigned long long t1;
    uint32_t smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_cta_times[blockIdx.x * 3 + 0] = dbg_t0;
    g_cta_times[blockIdx.x * 3 + 1] = t1;
    g_cta_times[blockIdx.x * 3 + 2] = smid;
  }
#endif
}

// ------------------------------------------------------------------------------------------
// k_extract: finishes a count-mode k_match.  Per job the per-slot counts and the per-window
// row-lane records say where the first k matches and the last match of each segment lie; only
// the candidate rows of those windows are evaluated again for this one job, lane = row, straight
// from the global rows and tables.  A job needs at most k + 2 windows: one to three when its
// matches are dense, k when they are sparse.  WPJ warps per job: 1 when a job has few slots,
// 8 when a small batch was cut into thousands of pieces (pass 1 then sums the slot counts with
// all 8 warps; the rest is done by the first of them).
// ------------------------------------------------------------------------------------------

// SHORT != 0 (one warp per job, at most 512 slots; 1: of at most 8 windows, 2: longer): lane l
// keeps the counts of slots 16l .. 16l+15 in registers (four 128-bit loads), so one sequential
// pass per lane and two warp scans give the total, the last non-empty slot of each segment and
// -- exactly -- the slots that hold the first k matches: the non-empty ones whose prefix count is
// below k.  Their bounds and, when a slot has at most 8 windows, their window records are then
// fetched in ONE batch (lane i = i-th such slot) instead of chunk by chunk with a dependent round
// trip each.
#ifndef EXTRACT_CTAS_PER_SM
#define EXTRACT_CTAS_PER_SM 4
#endif
template <int WPJ, bool ROW32, bool HAS_LUT, int SHORT>
__global__ void __launch_bounds__(256, EXTRACT_CTAS_PER_SM) k_extract(const ExtractParams p) {
  static_assert(!SHORT || WPJ == 1, "the batched walk is for one warp per job");
  constexpr uint32_t JPC = 8u / WPJ;  // jobs per CTA
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  extern __shared__ uint32_t chunk_sums[];  // [JPC][n_chunks]
  __shared__ uint32_t red_total[8];
  __shared__ int red_last_a[8], red_last_b[8];
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t jl = warp / WPJ, wj = warp % WPJ;
  // k_match evaluated job `job` in evaluation slot `es` (bit es % 32 of tile es / 32 % 4 of group
  // es / 128); its counts and window records are filed under the slot, the results under the job
  const uint32_t es = blockIdx.x * JPC + jl;
  const bool live = es < p.n_jobs;
  const uint32_t job = live ? (p.job_order ? p.job_order[es] : es) : 0u;
  const uint32_t n_slots = p.slots;
  const uint32_t n_chunks = (n_slots + 31u) / 32u;
  const uint32_t* cnt = p.part_cnt + (size_t)(live ? es : 0u) * p.slots;
#ifdef DSM_CTA_TIMING
  unsigned long long dbg_j0 = 0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_j0));
#endif
  uint32_t* csum = chunk_sums + jl * n_chunks;

  uint32_t creg[SHORT ? 16 : 1] = {0};  // SHORT: the counts of slots 16 * lane .. 16 * lane + 15
  uint32_t need_mask = 0, need_off = 0, n_need = 0;  // SHORT: this lane's needed slots, their rank, all of them
  if constexpr (SHORT) {
    // n_slots is a multiple of 8 (8 warps per piece), so a job's counts are 16-byte aligned and
    // every uint4 lies wholly inside or outside them
    const uint4* c4 = reinterpret_cast<const uint4*>(cnt) + lane * 4u;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (live && lane * 16u + q * 4u < n_slots) v = c4[q];
      creg[q * 4 + 0] = v.x; creg[q * 4 + 1] = v.y; creg[q * 4 + 2] = v.z; creg[q * 4 + 3] = v.w;
    }
    uint32_t lane_tot = 0;
    int la = -1, lb = -1;  // slots ascend with u: the last assignment is the lane's last non-empty slot
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      const uint32_t sl = lane * 16u + u;
      lane_tot += creg[u];
      if (creg[u] != 0u) {
        if (sl < p.slots_a) la = (int)sl;
        else lb = (int)sl;
      }
    }
    uint32_t incl = lane_tot;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(FULL, incl, d);
      if (lane >= (uint32_t)d) incl += y;
    }
    const uint32_t tot = __shfl_sync(FULL, incl, 31);
    const int ls_a = __reduce_max_sync(FULL, la), ls_b = __reduce_max_sync(FULL, lb);
    uint32_t run = incl - lane_tot;  // matches in the slots before this lane's
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      if (creg[u] != 0u && run < p.k) need_mask |= 1u << u;
      run += creg[u];
    }
    uint32_t nincl = (uint32_t)__popc(need_mask);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(FULL, nincl, d);
      if (lane >= (uint32_t)d) nincl += y;
    }
    need_off = nincl - (uint32_t)__popc(need_mask);
    n_need = __shfl_sync(FULL, nincl, 31);
    if (lane == 0) { red_total[warp] = tot; red_last_a[warp] = ls_a; red_last_b[warp] = ls_b; }
  } else {  // pass 1: totals of 32 slots at a time, last slot with a match per segment
    uint32_t tot = 0;
    int ls_a = -1, ls_b = -1;
    for (uint32_t ch0 = wj; ch0 < n_chunks; ch0 += 8u * WPJ) {
      uint32_t c[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t s = (ch0 + u * WPJ) * 32u + lane;
        c[u] = (live && s < n_slots) ? cnt[s] : 0u;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t ch = ch0 + u * WPJ;
        if (ch >= n_chunks) break;  // warp-uniform
        const uint32_t s = ch * 32u + lane;
        const uint32_t sum = __reduce_add_sync(FULL, c[u]);
        const uint32_t hit_a = __ballot_sync(FULL, c[u] != 0 && s < p.slots_a);
        const uint32_t hit_b = __ballot_sync(FULL, c[u] != 0 && s >= p.slots_a);
        if (hit_a) ls_a = (int)(ch * 32u + 31u - __clz(hit_a));
        if (hit_b) ls_b = (int)(ch * 32u + 31u - __clz(hit_b));
        if (lane == 0) csum[ch] = sum;
        tot += sum;
      }
    }
    if (lane == 0) { red_total[warp] = tot; red_last_a[warp] = ls_a; red_last_b[warp] = ls_b; }
  }
  __syncthreads();
  if (!live || wj != 0) return;

  uint32_t total = 0;
  int ls_a = -1, ls_b = -1;
#pragma unroll
  for (int w = 0; w < WPJ; ++w) {
    total += red_total[jl * WPJ + w];
    ls_a = max(ls_a, red_last_a[jl * WPJ + w]);
    ls_b = max(ls_b, red_last_b[jl * WPJ + w]);
  }

  // ---- what it takes to evaluate this job on one row per lane
  const uint32_t group = es / JPG, tile = (es % JPG) / 32u, bit = es & 31u;
  const uint32_t* tab = reinterpret_cast<const uint32_t*>(p.tabs + (size_t)group * p.n_entries) + tile;
  const uint32_t cut_a = min(p.recs[(size_t)job * p.rec_words + 0], p.live_a);
  const uint32_t cut_b = min(p.recs[(size_t)job * p.rec_words + 1], p.live_b);
  const uint32_t n_act = min(p.n_used, p.group_nact[group]);  // the columns k_match gathered for this group
  auto eval_row = [&](const uint32_t row, const uint32_t cut) -> uint32_t {
    // the row's codes come with one (16-byte rows) or two 128-bit loads, then the table words of
    // the group's n_act columns are fetched side by side: two dependent round trips per
    // candidate.  n_act is the same for the whole warp, so the unrolled column loop leaves at
    // the first column the group does not need
    const uint4* rp = p.rows + (size_t)row * (ROW32 ? 2u : 1u);
    const uint4 v0 = __ldg(rp);
    uint32_t w[ROW32 ? 8 : 4] = {v0.x, v0.y, v0.z, v0.w};
    if constexpr (ROW32) {
      const uint4 v1 = __ldg(rp + 1);
      w[4] = v1.x; w[5] = v1.y; w[6] = v1.z; w[7] = v1.w;
    }
    uint32_t ok = row < cut ? 1u : 0u;
#pragma unroll
    for (uint32_t c = 0; c < (ROW32 ? 16u : 8u); ++c) {
      if (c >= n_act) break;
      uint32_t code = (c & 1u) ? (w[c >> 1] >> 16) : (w[c >> 1] & 0xFFFFu);
      if constexpr (HAS_LUT) {
        if (p.lut[c] != nullptr) code = __ldg(p.lut[c] + code);
      }
      ok &= __ldg(tab + (size_t)(p.base[c] + code) * 4u) >> bit;
    }
    return ok & 1u;
  };
  // A window's record is the set R of row lanes (row mod 32) with a match somewhere in its
  // blocks, so its candidate rows are {32 * blk + l : blk in window, l in R}; candidate i (in
  // row order) is block i / |R|, the (i mod |R|)-th lane of R.  A sparse job has |R| = 1 or 2.
  auto candidate = [&](const uint32_t i, const uint32_t bs, const uint32_t R) -> uint32_t {
    const uint32_t nR = __popc(R);
    if (nR == 1u) return (bs + i) * 32u + (__ffs(R) - 1);  // the usual case of a sparse job
    if (nR == 32u) return bs * 32u + i;                    // and of a dense one
    const uint32_t b = (i * (65535u / nR + 1u)) >> 16;  // i / nR, exact for i < 512
    uint32_t r = i - b * nR, m = R, l = 0;              // l = position of the r-th set bit of R
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
      const uint32_t c = __popc(m & ((1u << s) - 1u));
      if (r >= c) { r -= c; m >>= s; l += s; }
    }
    return (bs + b) * 32u + l;
  };
  const uint32_t* wr_job = p.win_rows + (size_t)group * p.slots * p.wps * JPG + (es % JPG);

  // ---- windows waiting to be evaluated (warp-private): first block, row-lane set, cut-off,
  // index of the window's first candidate among all queued candidates.  Candidates are
  // evaluated 64 at a time, so a sparse job's one-lane windows share evaluations.
  uint32_t* out = p.out_idx + (size_t)job * p.k;
  const uint32_t k = p.k;
  const uint32_t lt_mask = (1u << lane) - 1u;
  __shared__ uint32_t s_pend[8][EXTRACT_PENDING][4];
  uint32_t (*pend)[4] = s_pend[warp];
  __shared__ uint32_t s_wlist[8][256][2];  // long slots: (window, row-lane set) of non-empty windows
  uint32_t (*wlist)[2] = s_wlist[warp];
  uint32_t n_pend = 0, q_total = 0;
  uint32_t have = 0;             // matches written so far
  auto push = [&](const uint32_t bs, const uint32_t nb, const uint32_t R, const uint32_t cut) {
    if (lane == 0) {
      pend[n_pend][0] = bs;
      pend[n_pend][1] = R;
      pend[n_pend][2] = cut;
      pend[n_pend][3] = q_total;
    }
    q_total += nb * (uint32_t)__popc(R);
    ++n_pend;
  };
  // append the matching candidates of the queued windows to out[] until k are there
  auto flush = [&]() {  // warp-uniform
    __syncwarp();
    for (uint32_t q0 = 0; q0 < q_total; q0 += 32u * EXTRACT_LANE_ROWS) {
      if (have >= k) break;
      uint32_t row[EXTRACT_LANE_ROWS], ok[EXTRACT_LANE_ROWS];
#pragma unroll
      for (int u = 0; u < EXTRACT_LANE_ROWS; ++u) {
        const uint32_t q = q0 + u * 32 + lane;
        row[u] = 0;
        ok[u] = 0;
        if (q < q_total) {
          uint32_t e = 0;  // the queued window that holds candidate q: starts ascend, at most 16 entries
#pragma unroll
          for (uint32_t s = EXTRACT_PENDING / 2; s >= 1u; s >>= 1)
            if (e + s < n_pend && q >= pend[e + s][3]) e += s;
          row[u] = candidate(q - pend[e][3], pend[e][0], pend[e][1]);
          ok[u] = eval_row(row[u], pend[e][2]);
        }
      }
#pragma unroll
      for (int u = 0; u < EXTRACT_LANE_ROWS; ++u) {
        const uint32_t hits = __ballot_sync(FULL, ok[u]);
        const uint32_t rank = have + __popc(hits & lt_mask);
        if (ok[u] && rank < k) out[rank] = row[u];
        have = min(k, have + (uint32_t)__popc(hits));
      }
    }
    n_pend = 0;
    q_total = 0;
    __syncwarp();
  };

  // ---- the last match of each segment (the max-price row) lies in the last non-empty window of
  // the segment's last non-empty slot.  Lanes 0-15 look after the first segment, lanes 16-31 after
  // the second; the slot bounds and window records are requested here, before the walk for the
  // first k matches, so that their latency is hidden behind it.
  const uint32_t half = lane >> 4, hl = lane & 15u;
  const int t_slot = half ? ls_b : ls_a;
  uint32_t t_b0 = 0, t_b1 = 0;
  if (t_slot >= 0) { t_b0 = p.slot_start[t_slot]; t_b1 = p.slot_start[t_slot + 1]; }
  const uint32_t t_nwin = (t_b1 - t_b0 + WIN_BLOCKS - 1) / WIN_BLOCKS;
  // the last (up to) 16 windows of the slot, window t_w = t_nwin - 1 - hl for this lane
  uint32_t t_rec = 0;
  if (t_slot >= 0 && hl < t_nwin)
    t_rec = wr_job[((size_t)t_slot * p.wps + (t_nwin - 1u - hl)) * JPG];

  // ---- first k matches, SHORT: the needed slots are known from the counts in registers; lane i
  // fetches the bounds (and, for slots of at most 8 windows, the window records) of the i-th of
  // them, 32 slots a round; a longer slot's records are fetched 256 windows at a time and the
  // non-empty ones compacted into a list.  Either way the windows reach the queue 8 at a time, so
  // that there is one flush site in the loop.
  if constexpr (SHORT) {
    __shared__ uint32_t s_need[8][32];
    uint32_t* need = s_need[warp];
    constexpr bool short_slots = SHORT == 1;  // SHORT == 2: slots of more than 8 windows
    for (uint32_t round_lo = 0; round_lo < n_need && have < k; round_lo += 32u) {
      {
        uint32_t m = need_mask, pos = need_off - round_lo;  // wraps for slots of earlier rounds: >= 32
        while (m != 0u) {
          const uint32_t u = (uint32_t)__ffs(m) - 1u;
          m &= m - 1u;
          if (pos < 32u) need[pos] = lane * 16u + u;
          ++pos;
        }
      }
      __syncwarp();
      const uint32_t n_here = min(32u, n_need - round_lo);
      uint32_t my_s = 0, my_b0 = 0, my_b1 = 0;
      uint32_t mk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (lane < n_here) {
        my_s = need[lane];
        my_b0 = p.slot_start[my_s];
        my_b1 = p.slot_start[my_s + 1];
        if constexpr (short_slots) {
          const uint32_t* wr = wr_job + (size_t)my_s * p.wps * JPG;
#pragma unroll
          for (int w = 0; w < 8; ++w) mk[w] = (uint32_t)w < p.wps ? wr[(size_t)w * JPG] : 0u;
          const uint32_t my_nwin = (my_b1 - my_b0 + WIN_BLOCKS - 1) / WIN_BLOCKS;
#pragma unroll
          for (int w = 0; w < 8; ++w) mk[w] = (uint32_t)w < my_nwin ? mk[w] : 0u;  // records past the slot's end are stale
        }
      }
      __syncwarp();
      for (uint32_t i = 0; i < n_here && have < k; ++i) {
        const uint32_t sl = __shfl_sync(FULL, my_s, i);
        const uint32_t b0 = __shfl_sync(FULL, my_b0, i), b1 = __shfl_sync(FULL, my_b1, i);
        const uint32_t cut = sl >= p.slots_a ? cut_b : cut_a;
        const uint32_t n_win = (b1 - b0 + WIN_BLOCKS - 1) / WIN_BLOCKS;
        const uint32_t* wr = wr_job + (size_t)sl * p.wps * JPG;
        // w0: first window of the batch of records at hand (8 in mk, or up to 256 in wlist)
        for (uint32_t w0 = 0; w0 < (short_slots ? 1u : n_win) && have < k; w0 += 256u) {
          uint32_t n_list = 8u;
          if constexpr (!short_slots) {
            uint32_t wv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const uint32_t w = w0 + u * 32 + lane;
              wv[u] = w < n_win ? wr[(size_t)w * JPG] : 0u;
            }
            n_list = 0;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const uint32_t nz = __ballot_sync(FULL, wv[u] != 0);
              if (wv[u] != 0) {
                const uint32_t at = n_list + __popc(nz & lt_mask);
                wlist[at][0] = w0 + u * 32 + lane;
                wlist[at][1] = wv[u];
              }
              n_list += __popc(nz);
            }
            __syncwarp();
          }
          for (uint32_t l0 = 0; l0 < n_list && have < k; l0 += 8u) {
#pragma unroll
            for (int w = 0; w < 8; ++w) {
              uint32_t win = (uint32_t)w, R = 0;
              if constexpr (short_slots) {
                R = __shfl_sync(FULL, mk[w], i);
              } else {
                const bool in = l0 + w < n_list;
                win = in ? wlist[l0 + w][0] : 0u;
                R = in ? wlist[l0 + w][1] : 0u;
              }
              if (R != 0) {
                const uint32_t bs = b0 + win * WIN_BLOCKS;
                push(bs, min(bs + WIN_BLOCKS, b1) - bs, R, cut);
              }
            }
            if (n_pend > EXTRACT_PENDING - 8u || q_total >= 32u * EXTRACT_LANE_ROWS) flush();
          }
          __syncwarp();  // before wlist is filled again
        }
      }
      flush();
      __syncwarp();  // before need[] is filled again
    }
  }
  // ---- first k matches, general: walk slots and windows in row order, skipping empty ones
  for (uint32_t c0 = 0; SHORT == 0 && c0 < n_chunks && have < k; c0 += 32) {
    const uint32_t cs = c0 + lane < n_chunks ? csum[c0 + lane] : 0u;
    uint32_t open_ch = __ballot_sync(FULL, cs != 0);
    while (open_ch && have < k) {  // all loops below are warp-uniform
      const uint32_t ch = c0 + __ffs(open_ch) - 1;
      open_ch &= open_ch - 1;
      const uint32_t s = ch * 32u + lane;
      const uint32_t c = s < n_slots ? cnt[s] : 0u;
      const uint32_t my_b0 = s <= n_slots ? p.slot_start[s] : 0u;  // lane 31 may read slot_start[n_slots]
      uint32_t open_sl = __ballot_sync(FULL, c != 0);
      // slots of at most 8 windows: every lane fetches the window records of its own slot, so
      // the whole chunk costs one memory round trip instead of one per non-empty slot
      uint32_t mk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      const bool short_slots = p.wps <= 8u;
      if (short_slots && c != 0) {
        const uint32_t my_b1 = p.slot_start[s + 1];
        const uint32_t my_nwin = (my_b1 - my_b0 + WIN_BLOCKS - 1) / WIN_BLOCKS;
        const uint32_t* wr = wr_job + (size_t)s * p.wps * JPG;
#pragma unroll
        for (int w = 0; w < 8; ++w) mk[w] = (uint32_t)w < my_nwin ? wr[(size_t)w * JPG] : 0u;
      }
      while (open_sl && have < k) {
        const uint32_t sl_lane = __ffs(open_sl) - 1;
        open_sl &= open_sl - 1;
        const uint32_t sl = ch * 32u + sl_lane;
        const uint32_t b0 = __shfl_sync(FULL, my_b0, sl_lane);
        const uint32_t b1 = sl_lane < 31u ? __shfl_sync(FULL, my_b0, sl_lane + 1) : p.slot_start[sl + 1];
        const uint32_t cut = sl >= p.slots_a ? cut_b : cut_a;
        if (short_slots) {
#pragma unroll
          for (int w = 0; w < 8; ++w) {
            const uint32_t R = __shfl_sync(FULL, mk[w], sl_lane);
            if (R != 0) {
              const uint32_t bs = b0 + (uint32_t)w * WIN_BLOCKS;
              push(bs, min(bs + WIN_BLOCKS, b1) - bs, R, cut);
            }
          }
          // one flush site per slot (the queue holds two slots' worth of windows)
          if (n_pend > EXTRACT_PENDING - 8u || q_total >= 32u * EXTRACT_LANE_ROWS) flush();
          continue;
        }
        const uint32_t n_win = (b1 - b0 + WIN_BLOCKS - 1) / WIN_BLOCKS;
        const uint32_t* wr = wr_job + (size_t)sl * p.wps * JPG;
        // long slots: fetch 256 window records at once, compact the non-empty ones into a
        // short list, then queue them from a rolled loop (one flush site)
        for (uint32_t w0 = 0; w0 < n_win && have < k; w0 += 256) {
          uint32_t wv[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t w = w0 + u * 32 + lane;
            wv[u] = w < n_win ? wr[(size_t)w * JPG] : 0u;
          }
          uint32_t n_list = 0;
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t nz = __ballot_sync(FULL, wv[u] != 0);
            if (wv[u] != 0) {
              const uint32_t at = n_list + __popc(nz & lt_mask);
              wlist[at][0] = w0 + u * 32 + lane;
              wlist[at][1] = wv[u];
            }
            n_list += __popc(nz);
          }
          __syncwarp();
          for (uint32_t i = 0; i < n_list && have < k; ++i) {
            const uint32_t bs = b0 + wlist[i][0] * WIN_BLOCKS;
            push(bs, min(bs + WIN_BLOCKS, b1) - bs, wlist[i][1], cut);
            if (n_pend == EXTRACT_PENDING || q_total >= 32u * EXTRACT_LANE_ROWS) flush();
          }
          __syncwarp();
        }
      }
    }
  }
  flush();
  for (uint32_t i = have + lane; i < k; i += 32) out[i] = DSM_NO_ROW;

  // ---- last match of each segment.  Per half-warp: the last non-empty window among the 16
  // requested above (a slot's last match is almost always there; otherwise the earlier windows are
  // read 16 at a time), then its candidate rows from the end, 16 per round, both segments at once.
  uint32_t last_ab[2] = {DSM_NO_ROW, DSM_NO_ROW};
  {
    uint32_t w_off = 0;  // windows before the 16 in t_rec (counted from the slot's end)
    uint32_t nz = (__ballot_sync(FULL, t_rec != 0) >> (16u * half)) & 0xFFFFu;
    // warp-uniform loop; a half that has found its window (or has none) idles
    while (__any_sync(FULL, t_slot >= 0 && nz == 0u && w_off + 16u < t_nwin)) {
      const bool more = t_slot >= 0 && nz == 0u && w_off + 16u < t_nwin;
      if (more) w_off += 16u;
      uint32_t rec = 0;
      if (more && w_off + hl < t_nwin) rec = wr_job[((size_t)t_slot * p.wps + (t_nwin - 1u - w_off - hl)) * JPG];
      const uint32_t z = (__ballot_sync(FULL, rec != 0) >> (16u * half)) & 0xFFFFu;
      if (more) { t_rec = rec; nz = z; }
    }
    // lane with the latest non-empty window of its half = lowest set bit of nz
    const uint32_t src = (nz ? (uint32_t)__ffs(nz) - 1u : 0u) + 16u * half;
    const uint32_t R = __shfl_sync(FULL, t_rec, src);
    const bool has = t_slot >= 0 && nz != 0u;
    const uint32_t w = t_nwin - 1u - w_off - (src & 15u);
    const uint32_t bs = t_b0 + w * WIN_BLOCKS;
    const uint32_t nb = has ? min(bs + WIN_BLOCKS, t_b1) - bs : 0u;
    const uint32_t n_cand = nb * (uint32_t)__popc(R);
    const uint32_t cut = half ? cut_b : cut_a;
    uint32_t found = DSM_NO_ROW;
    for (uint32_t r0 = 0; __any_sync(FULL, has && found == DSM_NO_ROW && r0 < n_cand); r0 += 16u) {
      const uint32_t q = r0 + hl;
      uint32_t row = 0, ok = 0;
      if (has && found == DSM_NO_ROW && q < n_cand) {
        row = candidate(n_cand - 1u - q, bs, R);
        ok = eval_row(row, cut);
      }
      const uint32_t hits = (__ballot_sync(FULL, ok) >> (16u * half)) & 0xFFFFu;
      const uint32_t first = (hits ? (uint32_t)__ffs(hits) - 1u : 0u) + 16u * half;
      const uint32_t r = __shfl_sync(FULL, row, first);
      if (hits && found == DSM_NO_ROW) found = r;
    }
    last_ab[0] = __shfl_sync(FULL, found, 0);
    last_ab[1] = __shfl_sync(FULL, found, 16);
  }
  const uint32_t last_a = last_ab[0], last_b = last_ab[1];
  if (lane == 0) {
    p.out_cnt[job] = total;
    uint32_t best = last_a;
    if (last_b != DSM_NO_ROW) {
      if (best == DSM_NO_ROW) best = last_b;
      else if (p.price != nullptr && p.price[last_b] > p.price[last_a]) best = last_b;
    }
    p.out_max[job] = best;
#ifdef DSM_CTA_TIMING
    if (job < 65536u) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      g_job_times[job * 2 + 0] = dbg_j0;
      g_job_times[job * 2 + 1] = t1;
    }
#endif
  }
}

// ------------------------------------------------------------------------------------------
// utility kernels
// ------------------------------------------------------------------------------------------

__global__ void k_gather_price(const double* price, const uint32_t* rows, uint32_t n, double* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rows[i];
  out[i] = (r == DSM_NO_ROW || price == nullptr) ? __longlong_as_double(0x7FF8000000000000LL) : price[r];
}

// Shards are in table order: concatenate per-shard first-k lists, sum counts, keep the
// highest-priced max row.  One thread per job.  With shard_stride_bytes == 0 the four inputs are
// dense [n_shards][...] arrays; otherwise every pointer addresses shard 0's field and shard s's
// field lies s * shard_stride_bytes further on (all fields of a shard in one all-gathered record).
__global__ void k_merge_shards(uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                               const uint32_t* idx_all, const uint32_t* cnt_all,
                               const uint32_t* max_all, const double* maxprice_all,
                               size_t shard_stride_bytes,
                               const uint64_t* shard_base, uint64_t* idx_out, uint64_t* cnt_out,
                               uint64_t* max_out, double* maxprice_out) {
  const uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= n_jobs) return;
  uint64_t total = 0;
  uint32_t have = 0;
  uint64_t best = ~0ull;
  double best_price = 0.0;
  for (uint32_t s = 0; s < n_shards; ++s) {
    const uint32_t *idx_s, *cnt_s, *max_s;
    const double* mp_s;
    if (shard_stride_bytes) {
      const size_t off = (size_t)s * shard_stride_bytes;
      idx_s = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(idx_all) + off);
      cnt_s = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(cnt_all) + off);
      max_s = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(max_all) + off);
      mp_s = reinterpret_cast<const double*>(reinterpret_cast<const char*>(maxprice_all) + off);
    } else {
      idx_s = idx_all + (size_t)s * n_jobs * k;
      cnt_s = cnt_all + (size_t)s * n_jobs;
      max_s = max_all + (size_t)s * n_jobs;
      mp_s = maxprice_all + (size_t)s * n_jobs;
    }
    const uint32_t c = cnt_s[job];
    if (c == 0) continue;
    const uint32_t take = min(min(c, k), k - have);
    for (uint32_t i = 0; i < take; ++i)
      idx_out[(size_t)job * k + have + i] = shard_base[s] + idx_s[(size_t)job * k + i];
    have += take;
    total += c;
    const uint32_t m = max_s[job];
    if (m != DSM_NO_ROW) {
      const double pr = mp_s[job];
      if (best == ~0ull || pr > best_price) { best = shard_base[s] + m; best_price = pr; }
    }
  }
  for (uint32_t i = have; i < k; ++i) idx_out[(size_t)job * k + i] = ~0ull;
  cnt_out[job] = total;
  max_out[job] = best;
  maxprice_out[job] = (best == ~0ull) ? __longlong_as_double(0x7FF8000000000000LL) : best_price;
}

// Interleaved shards: the result order was dealt out to the shards in chunks of chunk_rows
// positions (chunk c to shard c % n_shards), so a shard's first-k list is ascending in the
// global order but the lists interleave.  geom[s] = (split_row, n_avail) of shard s turns a
// physical row of that shard's table into its position in the shard's own order, and from there
// into the global position; the merge keeps the k smallest positions.  Always strided input
// (every pointer addresses shard 0's field).  One thread per job; n_shards <= 32.
__global__ void k_merge_shards_interleaved(uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                                           const uint32_t* idx_all, const uint32_t* cnt_all,
                                           const uint32_t* max_all, const double* maxprice_all,
                                           size_t shard_stride_bytes, const uint2* geom,
                                           uint32_t chunk_rows, uint64_t* idx_out, uint64_t* cnt_out,
                                           uint64_t* max_out, double* maxprice_out) {
  const uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= n_jobs) return;
  auto field = [&](const void* base, uint32_t s) {
    return reinterpret_cast<const char*>(base) + (size_t)s * shard_stride_bytes;
  };
  auto position = [&](uint32_t s, uint32_t row) -> uint64_t {
    const uint2 g = geom[s];
    const uint32_t q = row < g.x ? row : g.y + (row - g.x);  // < the shard's row count: 32-bit division
    return ((uint64_t)(q / chunk_rows) * n_shards + s) * chunk_rows + q % chunk_rows;
  };
  uint64_t head[32];  // global position of each shard's next unread entry, ~0 = exhausted
  uint8_t used[32];
  uint64_t total = 0, best = ~0ull;
  double best_price = 0.0;
  for (uint32_t s = 0; s < n_shards; ++s) {
    const uint32_t* idx_s = reinterpret_cast<const uint32_t*>(field(idx_all, s)) + (size_t)job * k;
    const uint32_t c = reinterpret_cast<const uint32_t*>(field(cnt_all, s))[job];
    total += c;
    used[s] = 0;
    head[s] = (c != 0 && idx_s[0] != DSM_NO_ROW) ? position(s, idx_s[0]) : ~0ull;
    const uint32_t m = reinterpret_cast<const uint32_t*>(field(max_all, s))[job];
    if (m != DSM_NO_ROW) {
      const double pr = reinterpret_cast<const double*>(field(maxprice_all, s))[job];
      if (best == ~0ull || pr > best_price) { best = position(s, m); best_price = pr; }
    }
  }
  for (uint32_t i = 0; i < k; ++i) {
    uint32_t win = 0;
    uint64_t lo = ~0ull;
    for (uint32_t s = 0; s < n_shards; ++s)
      if (head[s] < lo) { lo = head[s]; win = s; }
    idx_out[(size_t)job * k + i] = lo;
    if (lo == ~0ull) continue;
    const uint32_t* idx_w = reinterpret_cast<const uint32_t*>(field(idx_all, win)) + (size_t)job * k;
    const uint32_t nxt = ++used[win];
    head[win] = (nxt < k && idx_w[nxt] != DSM_NO_ROW) ? position(win, idx_w[nxt]) : ~0ull;
  }
  cnt_out[job] = total;
  max_out[job] = best;
  maxprice_out[job] = (best == ~0ull) ? __longlong_as_double(0x7FF8000000000000LL) : best_price;
}

// ------------------------------------------------------------------------------------------
// k_pack_jobs: raw jobs -> predicate records (one thread per job).  The translation
// dstack_b200/encode.py:pack_jobs does with numpy, with the same double comparisons:
// bound -> ordinal by binary search in the table's sorted value lists (searchsorted left for a
// minimum, right for a maximum), max_price -> cut-off rows by binary search in the prices.
// ------------------------------------------------------------------------------------------

struct PackParams {
  const dsm_raw_job* raw;
  const int32_t* ids;
  uint32_t n_jobs;
  LayoutDev lay;
  uint32_t role[DSM_MAX_COLS];
  const double* dict[DSM_N_PACK_DICTS];
  uint32_t dict_n[DSM_N_PACK_DICTS];
  const double* price;  // physical rows: [0, n_live_a) and [split_row, split_row + n_live_b) ascending
  uint32_t n_live_a, split_row, n_live_b;
  uint32_t has_state;  // a DSM_ROLE_STATE column is part of the layout
  uint32_t* recs;
  uint32_t* keys;  // per job: 1 + index of the last column that rejects anything (for the evaluation order)
};

__device__ __forceinline__ uint32_t lower_bound_f64(const double* v, uint32_t n, double x) {  // first i: v[i] >= x
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (v[mid] < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ uint32_t upper_bound_f64(const double* v, uint32_t n, double x) {  // first i: v[i] > x
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (x < v[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}
// inclusive ordinal range lo | hi<<16 of the values inside [b[0], b[1]] (NaN = open); empty -> (1, 0)
__device__ __forceinline__ uint32_t range_word_f64(const double* v, uint32_t n, const double* b) {
  const long long lo = isnan(b[0]) ? 0ll : (long long)lower_bound_f64(v, n, b[0]);
  const long long hi = isnan(b[1]) ? (long long)n - 1ll : (long long)upper_bound_f64(v, n, b[1]) - 1ll;
  if (lo > hi || n == 0) return 1u;
  return (uint32_t)lo | ((uint32_t)hi << 16);
}
__device__ __forceinline__ bool range_word_trivial(uint32_t w, uint32_t n) {
  return n == 0 || ((w & 0xFFFFu) == 0u && (w >> 16) >= n - 1u);
}

__global__ void __launch_bounds__(128) k_pack_jobs(const PackParams p) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= p.n_jobs) return;
  const dsm_raw_job& r = p.raw[j];
  uint32_t* rec = p.recs + (size_t)j * p.lay.rec_words;
  for (uint32_t w = 0; w < p.lay.rec_words; ++w) rec[w] = 0u;

  // words 0, 1: cut-off rows
  const bool no_mp = isnan(r.max_price);
  uint32_t cut_a = p.n_live_a, cut_b = p.split_row + p.n_live_b;
  if (!no_mp) {
    cut_a = upper_bound_f64(p.price, p.n_live_a, r.max_price);
    cut_b = p.split_row + upper_bound_f64(p.price + p.split_row, p.n_live_b, r.max_price);
  }
  if (r.flags & DSM_RAWF_EXCLUDE_NOT_AVAILABLE) cut_b = p.split_row;
  rec[0] = cut_a;
  rec[1] = cut_b;

  const double* cpu_v = p.dict[DSM_DICT_CPU];
  const uint32_t cpu_n = p.dict_n[DSM_DICT_CPU];
  const uint32_t cpu_w = range_word_f64(cpu_v, cpu_n, r.cpu);
  const uint32_t mem_w = range_word_f64(p.dict[DSM_DICT_MEMORY], p.dict_n[DSM_DICT_MEMORY], r.memory);
  const bool cpu_triv = range_word_trivial(cpu_w, cpu_n) && r.blocks == 0;
  const bool mem_triv = range_word_trivial(mem_w, p.dict_n[DSM_DICT_MEMORY]);
  // cpu as a set: in range and (blocks > 1: divisible by blocks; blocks == 0: any; else nothing)
  auto cpu_bitset = [&](uint32_t* out) {
    for (uint32_t e = 0; e < cpu_n; ++e) {
      bool ok = e >= (cpu_w & 0xFFFFu) && e <= (cpu_w >> 16);
      if (r.blocks > 1) ok = ok && ((long long)cpu_v[e] % (long long)r.blocks == 0ll);
      else ok = ok && r.blocks == 0;
      if (ok) out[e >> 5] |= 1u << (e & 31u);
    }
  };
  auto put_ids = [&](uint32_t* out, uint32_t off, uint32_t cnt) {
    for (uint32_t i = 0; i < cnt; ++i) {
      const int32_t id = p.ids[off + i];
      if (id >= 0) out[(uint32_t)id >> 5] |= 1u << ((uint32_t)id & 31u);
    }
  };

  uint32_t last_needed = 0;
  for (uint32_t c = 0; c < p.lay.n_used; ++c) {
    uint32_t* w = rec + p.lay.rec_off[c];
    const uint32_t card = p.lay.card[c];
    bool triv = false;
    switch (p.role[c]) {
      case DSM_ROLE_CPU:
        if (p.lay.kind[c] == DSM_COL_BITSET) cpu_bitset(w); else w[0] = cpu_w;
        triv = cpu_triv;
        break;
      case DSM_ROLE_MEMORY:
        w[0] = mem_w;
        triv = mem_triv;
        break;
      case DSM_ROLE_CPU_MEMORY:
        if (p.lay.flags[c] & DSM_COLF_PAIR_A_BITSET) cpu_bitset(w); else w[0] = cpu_w;
        w[p.lay.aux[c]] = mem_w;
        triv = cpu_triv && mem_triv;
        break;
      case DSM_ROLE_DISK: {
        const uint32_t n = p.dict_n[DSM_DICT_DISK];
        w[0] = range_word_f64(p.dict[DSM_DICT_DISK], n, r.disk);
        triv = range_word_trivial(w[0], n);
        break;
      }
      case DSM_ROLE_GPU: {
        const uint32_t n_cnt = p.dict_n[DSM_DICT_GPU_COUNT], n_tot = p.dict_n[DSM_DICT_GPU_TOTAL];
        w[0] = range_word_f64(p.dict[DSM_DICT_GPU_COUNT], n_cnt, r.gpu_count);
        w[1] = range_word_f64(p.dict[DSM_DICT_GPU_MEMORY], p.dict_n[DSM_DICT_GPU_MEMORY], r.gpu_memory);
        w[2] = range_word_f64(p.dict[DSM_DICT_GPU_TOTAL], n_tot, r.gpu_total);
        uint32_t gflags = 0;
        if (!(isnan(r.gpu_memory[0]) && isnan(r.gpu_memory[1]))) gflags |= DSM_GPUF_MEM_BOUNDED;
        if (r.cc_min >= 0) gflags |= DSM_GPUF_CC_BOUNDED;
        if (r.flags & DSM_RAWF_GPU_NAME_ANY) gflags |= DSM_GPUF_NAME_ANY;
        const uint32_t cc_ord = lower_bound_f64(p.dict[DSM_DICT_CC], p.dict_n[DSM_DICT_CC], (double)max(r.cc_min, 0));
        w[3] = min(cc_ord, 0xFFFEu) | (gflags << 16);
        const uint32_t vmask = r.gpu_vendor == 0 ? 0xFFu : (1u << r.gpu_vendor);
        w[4] = vmask | ((r.blocks > 1 ? (uint32_t)r.blocks : 0u) << 16);
        put_ids(w + 5, r.gpu_name_off, r.gpu_name_cnt);
        triv = range_word_trivial(w[0], n_cnt) && range_word_trivial(w[2], n_tot) &&
               gflags == DSM_GPUF_NAME_ANY && r.gpu_vendor == 0 && r.blocks <= 1;
        break;
      }
      case DSM_ROLE_BK: {
        const uint32_t spot_mask = r.spot < 0 ? 3u : (r.spot == 0 ? 1u : 2u);
        w[0] = r.src_backend_mask;
        w[1] = r.off_backend_mask;
        // with a state column (pool tables) status and the four dynamic flags are tested there
        const uint32_t st = p.has_state ? 0xFFu : (uint32_t)r.status_mask;
        const uint32_t req = p.has_state ? (r.flag_req & ~15u) : (uint32_t)r.flag_req;
        const uint32_t forbid = p.has_state ? (r.flag_forbid & ~15u) : (uint32_t)r.flag_forbid;
        w[2] = spot_mask | (st << 8);
        w[3] = req | (forbid << 16);
        triv = r.src_backend_mask == 0xFFFFFFFFu && r.off_backend_mask == 0xFFFFFFFFu && spot_mask == 3u &&
               st == 0xFFu && req == 0 && forbid == 0;
        break;
      }
      case DSM_ROLE_REGION:
      case DSM_ROLE_ITYPE:
      case DSM_ROLE_FLEET:
      case DSM_ROLE_NAME:
      case DSM_ROLE_ZONE: {
        const uint32_t role = p.role[c];
        const uint32_t any_bit = role == DSM_ROLE_REGION ? DSM_RAWF_REGION_ANY
                               : role == DSM_ROLE_ITYPE ? DSM_RAWF_ITYPE_ANY
                               : role == DSM_ROLE_FLEET ? DSM_RAWF_FLEET_ANY
                               : role == DSM_ROLE_NAME  ? DSM_RAWF_NAME_ANY : DSM_RAWF_ZONE_ANY;
        const uint32_t nw = (card + 31u) / 32u;
        triv = (r.flags & any_bit) != 0;
        if (triv) {
          for (uint32_t i = 0; i < nw; ++i) w[i] = 0xFFFFFFFFu;
        } else {
          if (role == DSM_ROLE_REGION) put_ids(w, r.region_off, r.region_cnt);
          else if (role == DSM_ROLE_ITYPE) put_ids(w, r.itype_off, r.itype_cnt);
          else if (role == DSM_ROLE_ZONE) put_ids(w, r.zone_off, r.zone_cnt);
          else {
            const int32_t id = role == DSM_ROLE_FLEET ? r.fleet_id : r.name_id;
            if (id >= 0) w[(uint32_t)id >> 5] |= 1u << ((uint32_t)id & 31u);
          }
          // the entry of rows whose value is unknown (last) passes an instance-type / zone filter
          if (role == DSM_ROLE_ITYPE || role == DSM_ROLE_ZONE) w[(card - 1u) >> 5] |= 1u << ((card - 1u) & 31u);
        }
        break;
      }
      case DSM_ROLE_STATE: {
        triv = true;
        const uint32_t dyn_req = r.flag_req & 15u, dyn_forbid = r.flag_forbid & 15u;
        for (uint32_t e = 0; e < card; ++e) {
          const bool ok = ((r.status_mask >> (e >> 4)) & 1u) && ((e & 15u) & dyn_req) == dyn_req && ((e & 15u) & dyn_forbid) == 0u;
          if (ok) w[e >> 5] |= 1u << (e & 31u); else triv = false;
        }
        break;
      }
      default:
        break;
    }
    if (!triv) last_needed = c + 1u;
  }
  if (p.keys) p.keys[j] = last_needed;
}

// Evaluation order: jobs grouped by key (counting sort, keys <= DSM_MAX_COLS; the order inside a
// key is whatever the atomics give -- results do not depend on it).  One CTA.
__global__ void __launch_bounds__(1024) k_order_jobs(const uint32_t* keys, uint32_t n_jobs, uint32_t* order) {
  __shared__ uint32_t hist[DSM_MAX_COLS + 2];
  if (threadIdx.x < DSM_MAX_COLS + 2) hist[threadIdx.x] = 0u;
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < n_jobs; j += blockDim.x) atomicAdd(&hist[min(keys[j], (uint32_t)DSM_MAX_COLS) + 1u], 1u);
  __syncthreads();
  if (threadIdx.x == 0)
    for (uint32_t i = 1; i < DSM_MAX_COLS + 2; ++i) hist[i] += hist[i - 1];  // hist[k] = first slot of key k
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < n_jobs; j += blockDim.x)
    order[atomicAdd(&hist[min(keys[j], (uint32_t)DSM_MAX_COLS)], 1u)] = j;
}

// ------------------------------------------------------------------------------------------
// match bitmaps -> per-job index lists
// ------------------------------------------------------------------------------------------

// bits [n_jobs][n_words] over physical rows -> sbits [n_jobs][s_words] over source positions
__global__ void k_scatter_to_source(const uint32_t* bits, uint32_t n_jobs, uint32_t n_words, const int32_t* row_to_source,
                                    uint32_t* sbits, uint32_t s_words) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_jobs * n_words) return;
  uint32_t m = bits[i];
  const uint32_t job = (uint32_t)(i / n_words), word = (uint32_t)(i % n_words);
  while (m) {
    const uint32_t b = __ffs(m) - 1;
    m &= m - 1;
    const int32_t s = row_to_source[word * 32u + b];
    if (s >= 0) atomicOr(&sbits[(size_t)job * s_words + ((uint32_t)s >> 5)], 1u << ((uint32_t)s & 31u));
  }
}

// one warp per job: matches of the job
__global__ void __launch_bounds__(256) k_count_bits(const uint32_t* bits, uint32_t n_jobs, uint32_t n_words,
                                                    unsigned long long* cnt) {
  const uint32_t job = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (job >= n_jobs) return;
  const uint32_t* b = bits + (size_t)job * n_words;
  uint32_t c = 0;
  for (uint32_t w = lane; w < n_words; w += 32u) c += __popc(b[w]);
  c = __reduce_add_sync(0xFFFFFFFFu, c);
  if (lane == 0) cnt[job] = c;
}

// exclusive prefix of cnt[n] into ptr[n + 1]; one CTA
__global__ void __launch_bounds__(1024) k_prefix_u64(const unsigned long long* cnt, uint32_t n, unsigned long long* ptr) {
  __shared__ unsigned long long part[1024];
  const uint32_t per = (n + blockDim.x - 1) / blockDim.x;
  const uint32_t lo = min(n, threadIdx.x * per), hi = min(n, lo + per);
  unsigned long long s = 0;
  for (uint32_t i = lo; i < hi; ++i) s += cnt[i];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long run = 0;
    for (uint32_t t = 0; t < blockDim.x; ++t) { const unsigned long long v = part[t]; part[t] = run; run += v; }
    ptr[n] = run;
  }
  __syncthreads();
  unsigned long long run = part[threadIdx.x];
  for (uint32_t i = lo; i < hi; ++i) { ptr[i] = run; run += cnt[i]; }
}

// one warp per job: the set bits of its bitmap in ascending order; value = the bit's index, or
// (map != nullptr) map[index]
__global__ void __launch_bounds__(256) k_fill_indices(const uint32_t* bits, uint32_t n_jobs, uint32_t n_words,
                                                      const int32_t* map, const unsigned long long* ptr, uint32_t* out) {
  const uint32_t job = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (job >= n_jobs) return;
  const uint32_t* b = bits + (size_t)job * n_words;
  unsigned long long at = ptr[job];
  for (uint32_t w0 = 0; w0 < n_words; w0 += 32u) {
    const uint32_t w = w0 + lane;
    uint32_t m = w < n_words ? b[w] : 0u;
    const uint32_t c = __popc(m);
    uint32_t incl = c;  // inclusive scan of the lanes' popcounts
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= (uint32_t)d) incl += y;
    }
    unsigned long long o = at + (incl - c);
    while (m) {
      const uint32_t bit = __ffs(m) - 1;
      m &= m - 1;
      const uint32_t idx = w * 32u + bit;
      out[o++] = map ? (uint32_t)map[idx] : idx;
    }
    at += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
}

// rows [n] (DSM_NO_ROW = none) -> source positions (-1 = none)
__global__ void k_rows_to_source(const uint32_t* rows, uint32_t n, const int32_t* row_to_source, int32_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rows[i];
  out[i] = r == DSM_NO_ROW ? -1 : row_to_source[r];
}

// rows[i] <- codes[i] (whole rows of `row_u4` uint4): in-place patch of a resident table
__global__ void k_patch_rows(uint4* table, const uint32_t* rows, uint32_t n, const uint4* codes, uint32_t row_u4) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * row_u4) return;
  table[(size_t)rows[i / row_u4] * row_u4 + i % row_u4] = codes[i];
}
