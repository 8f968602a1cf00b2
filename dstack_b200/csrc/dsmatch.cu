// dsmatch.cu -- sm_100a kernels + C ABI for dstack offer matching (see include/dsmatch.h).
//
// Data flow of one call (J jobs against a table of N rows, J_t = 128 jobs per group):
//
//   records [J][rec_words] u32 --k_build_tables--> tabs [groups][E][4] u32   (scratch)
//   rows [N][n_cols] u16 + tabs --k_match--> per-slot partials                (scratch)
//        part_cnt [S][J], part_last [S][J], part_idx [S][J][k]
//   partials + price --k_merge--> idx [J][k], cnt [J], max [J]
//
// k_match is the hot kernel.  One thread owns one row (a single 128-bit load of its 8 codes,
// or two for 16-code rows) and, for each of the columns, gathers a 128-bit word group from
// the shared-memory acceptance table: bit j of word t says whether job 32t+j of the group
// accepts this dictionary entry.  ANDing the gathered groups evaluates 128 jobs against the
// row; a 32x32 bit transpose across the warp then gives lane j the 32-row match word of job j,
// from which count / first-k / last-match are kept per job in registers.  No tensor cores:
// there is no contraction here, the work is integer gathers and bit logic.
//
// Reference being replaced: gpuhunt.matches per (job, offer) at
// core/backends/base/offers.py:154-157 and server/services/pools.py:496-502,533-543, and the
// ranker at server/services/backends/__init__.py:449-451 (see include/dsmatch.h).

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "../../include/dsmatch.h"

#define JPG DSM_JOBS_PER_GROUP   // 128 jobs per group
#define TPG 4                    // 32-job tiles per group
#define MATCH_THREADS 256
#define MATCH_WARPS 8
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
// device helpers
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
// TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// Streaming 128-bit row load: rows are read once per job group, keep them out of L1.
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// 32x32 bit-matrix transpose across a warp: on return lane j holds, in bit l, bit j of the
// value lane l passed in.  Five butterfly steps; rot[] holds the per-lane rotate amounts and
// keep[] the per-lane keep masks (both precomputed once per thread by transpose_setup).
struct XposeCtl {
  uint32_t rot[5];
  uint32_t keep[5];
};
__device__ __forceinline__ void transpose_setup(XposeCtl& c, uint32_t lane) {
  const uint32_t lowmask[5] = {0x0000FFFFu, 0x00FF00FFu, 0x0F0F0F0Fu, 0x33333333u, 0x55555555u};
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const uint32_t s = 16u >> i;
    const bool upper = (lane & s) != 0;
    c.rot[i] = upper ? 32u - s : s;                  // rotl amount: >>s for upper lanes, <<s for lower
    c.keep[i] = upper ? ~lowmask[i] : lowmask[i];    // bits this lane keeps of its own word
  }
}
__device__ __forceinline__ uint32_t transpose32(uint32_t x, const XposeCtl& c) {
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const uint32_t s = 16u >> i;
    const uint32_t y = __shfl_xor_sync(0xFFFFFFFFu, x, s);
    const uint32_t moved = __funnelshift_l(y, y, c.rot[i]);
    x = (x & c.keep[i]) | (moved & ~c.keep[i]);
  }
  return x;
}

// ------------------------------------------------------------------------------------------
// kernel parameters
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
  uint32_t n_jobs;
  const uint32_t* gpu_entries;
  const uint32_t* bk_entries;
  uint4* tabs;  // [groups][n_entries] x 4 words
};

struct MatchParams {
  const uint4* rows;
  const uint4* tabs;
  const uint32_t* recs;
  const uint16_t* lut[DSM_MAX_COLS];
  uint32_t base[DSM_MAX_COLS];
  uint32_t rec_words;
  uint32_t n_entries;
  uint32_t n_jobs;
  uint32_t n_groups;
  uint32_t n_blocks;      // rows / 32
  uint32_t split_block;   // first block of the second segment
  uint32_t ranges_a;      // row ranges in the first segment
  uint32_t ranges_b;      // row ranges in the second segment
  uint32_t k;
  uint32_t jobs_pad;      // n_groups * 128: leading dimension of the partial arrays
  uint32_t* part_cnt;     // [slots][jobs_pad]
  uint32_t* part_last;    // [slots][jobs_pad]
  uint32_t* part_idx;     // [slots][jobs_pad][k]
  uint32_t* bits;         // bitmap mode: [n_jobs][n_blocks]
};

struct MergeParams {
  const uint32_t* part_cnt;
  const uint32_t* part_last;
  const uint32_t* part_idx;
  const double* price;
  uint32_t n_jobs, jobs_pad, k, slots, slots_a;
  uint32_t* out_idx;
  uint32_t* out_cnt;
  uint32_t* out_max;
};

// ------------------------------------------------------------------------------------------
// k_build_tables: job records -> bit-sliced acceptance tables
//   grid (groups*4, n_used), block 256 = 8 warps; lane = job of the tile, warps stride entries.
// ------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256) k_build_tables(const BuildParams p) {
  const uint32_t tile = blockIdx.x;
  const uint32_t col = blockIdx.y;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t job = tile * 32u + lane;
  const bool live = job < p.n_jobs;
  const uint32_t* rec = p.recs + (size_t)(live ? job : 0) * p.lay.rec_words + p.lay.rec_off[col];
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

  for (uint32_t e = warp; e < card; e += 8u) {
    bool ok = false;
    if (live) {
      if (kind == DSM_COL_RANGE) {
        const uint32_t lo = r0 & 0xFFFFu, hi = r0 >> 16;
        ok = (e >= lo && e <= hi);
        if ((p.lay.flags[col] & DSM_COLF_NULL_PASSES) && e == card - 1u) ok = true;
      } else if (kind == DSM_COL_BITSET) {
        ok = (rec[e >> 5] >> (e & 31u)) & 1u;
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
// k_match: the hot kernel
// ------------------------------------------------------------------------------------------

template <int NC>
struct RowCodes {
  uint32_t w[(NC + 1) / 2];
};

template <int NC>
__device__ __forceinline__ void load_row(const uint4* rows, uint32_t row, RowCodes<NC>& rc) {
  if constexpr (NC <= 8) {
    const uint4 v = ldg_stream(rows + row);
    const uint32_t t[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < (NC + 1) / 2; ++i) rc.w[i] = t[i];
  } else {
    const uint4 v0 = ldg_stream(rows + 2 * (size_t)row);
    const uint4 v1 = ldg_stream(rows + 2 * (size_t)row + 1);
    const uint32_t t[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int i = 0; i < (NC + 1) / 2; ++i) rc.w[i] = t[i];
  }
}

template <int NC, bool BITMAP, bool HAS_LUT>
__global__ void __launch_bounds__(MATCH_THREADS, 2) k_match(const MatchParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint4* s_tab = reinterpret_cast<uint4*>(smem_raw);
  uint32_t* s_cut = reinterpret_cast<uint32_t*>(smem_raw + (size_t)p.n_entries * 16u);  // [2][128]
  __shared__ __align__(8) uint64_t s_bar;

  const uint32_t tid = threadIdx.x;
  const uint32_t lane = tid & 31u;
  const uint32_t warp = tid >> 5;
  XposeCtl xc;
  transpose_setup(xc, lane);

  if (tid == 0) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const uint32_t n_ranges = p.ranges_a + p.ranges_b;
  const uint32_t n_items = p.n_groups * n_ranges;
  uint32_t parity = 0;
  uint32_t cur_group = 0xFFFFFFFFu;

  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t group = item / n_ranges;
    const uint32_t range = item - group * n_ranges;

    if (group != cur_group) {
      __syncthreads();  // everyone is done with the previous group's table
      if (tid == 0) {
        const uint32_t bytes = p.n_entries * 16u;
        // the table was read through the generic proxy; order those reads before the
        // async-proxy overwrite
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&s_bar, bytes);
        bulk_g2s(s_tab, p.tabs + (size_t)group * p.n_entries, bytes, &s_bar);
      }
      if (tid < JPG) {
        const uint32_t job = group * JPG + tid;
        uint32_t ca = 0, cb = 0;  // padding jobs accept nothing
        if (job < p.n_jobs) {
          ca = p.recs[(size_t)job * p.rec_words + 0];
          cb = p.recs[(size_t)job * p.rec_words + 1];
        }
        s_cut[tid] = ca;
        s_cut[JPG + tid] = cb;
      }
      mbar_wait(&s_bar, parity);
      parity ^= 1u;
      __syncthreads();
      cur_group = group;
    }

    // rows of this work item: a range inside one segment, split evenly over the 8 warps
    const bool seg_b = range >= p.ranges_a;
    uint32_t seg_lo, seg_n, ridx, rcnt;
    if (!seg_b) { seg_lo = 0; seg_n = p.split_block; ridx = range; rcnt = p.ranges_a; }
    else { seg_lo = p.split_block; seg_n = p.n_blocks - p.split_block; ridx = range - p.ranges_a; rcnt = p.ranges_b; }
    const uint32_t r0 = seg_lo + (uint32_t)(((uint64_t)seg_n * ridx) / rcnt);
    const uint32_t r1 = seg_lo + (uint32_t)(((uint64_t)seg_n * (ridx + 1)) / rcnt);
    const uint32_t rl = r1 - r0;
    const uint32_t b0 = r0 + (uint32_t)(((uint64_t)rl * warp) / MATCH_WARPS);
    const uint32_t b1 = r0 + (uint32_t)(((uint64_t)rl * (warp + 1)) / MATCH_WARPS);
    const uint32_t* cut = s_cut + (seg_b ? JPG : 0);

    // price / availability cut-offs: classify each job of the group for this warp's rows.
    // full: every row passes; part: the cut-off lies inside; neither: every row fails.
    uint32_t full[TPG], part[TPG];
#pragma unroll
    for (int t = 0; t < TPG; ++t) {
      const uint32_t c = cut[t * 32 + lane];
      full[t] = __ballot_sync(0xFFFFFFFFu, c >= b1 * 32u);
      part[t] = __ballot_sync(0xFFFFFFFFu, c > b0 * 32u && c < b1 * 32u);
    }

    uint32_t cnt[TPG], last[TPG];
#pragma unroll
    for (int t = 0; t < TPG; ++t) { cnt[t] = 0; last[t] = DSM_NO_ROW; }

    const uint32_t slot = range * MATCH_WARPS + warp;

    RowCodes<NC> rc;
    if (b0 < b1) load_row<NC>(p.rows, b0 * 32u + lane, rc);

    for (uint32_t blk = b0; blk < b1; ++blk) {
      RowCodes<NC> cur = rc;
      if (blk + 1 < b1) load_row<NC>(p.rows, (blk + 1) * 32u + lane, rc);  // prefetch
      const uint32_t row = blk * 32u + lane;

      uint4 m = make_uint4(full[0], full[1], full[2], full[3]);
      {
        uint32_t pw[TPG] = {0, 0, 0, 0};
#pragma unroll
        for (int t = 0; t < TPG; ++t) {
          uint32_t rest = part[t];
          while (rest) {  // warp-uniform: usually zero iterations
            const uint32_t j = __ffs(rest) - 1;
            rest &= rest - 1;
            pw[t] |= (row < cut[t * 32 + j]) ? (1u << j) : 0u;
          }
        }
        m.x |= pw[0]; m.y |= pw[1]; m.z |= pw[2]; m.w |= pw[3];
      }

#pragma unroll
      for (int c = 0; c < NC; ++c) {
        uint32_t code = (c & 1) ? (cur.w[c >> 1] >> 16) : (cur.w[c >> 1] & 0xFFFFu);
        if constexpr (HAS_LUT) {
          if (p.lut[c] != nullptr) code = __ldg(p.lut[c] + code);
        }
        const uint4 a = s_tab[p.base[c] + code];
        m.x &= a.x; m.y &= a.y; m.z &= a.z; m.w &= a.w;
      }

      const uint32_t mm[TPG] = {m.x, m.y, m.z, m.w};
#pragma unroll
      for (int t = 0; t < TPG; ++t) {
        const uint32_t T = transpose32(mm[t], xc);  // lane j: rows of this block matching job 32t+j
        if constexpr (BITMAP) {
          const uint32_t job = cur_group * JPG + t * 32 + lane;
          if (job < p.n_jobs) p.bits[(size_t)job * p.n_blocks + blk] = T;
        } else {
          if (T) {
            last[t] = blk * 32u + 31u - __clz(T);
            if (cnt[t] < p.k) {
              uint32_t* dst = p.part_idx +
                              ((size_t)slot * p.jobs_pad + cur_group * JPG + t * 32 + lane) * p.k;
              uint32_t w = T, n = cnt[t];
              while (w && n < p.k) {
                dst[n++] = blk * 32u + (__ffs(w) - 1);
                w &= w - 1;
              }
            }
            cnt[t] += __popc(T);
          }
        }
      }
    }

    if constexpr (!BITMAP) {
#pragma unroll
      for (int t = 0; t < TPG; ++t) {
        const size_t o = (size_t)slot * p.jobs_pad + cur_group * JPG + t * 32 + lane;
        p.part_cnt[o] = cnt[t];
        p.part_last[o] = last[t];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// k_merge: order the per-slot partials of each job (one thread per job)
// ------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(128) k_merge(const MergeParams p) {
  const uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= p.n_jobs) return;
  uint32_t total = 0, have = 0;
  uint32_t last_a = DSM_NO_ROW, last_b = DSM_NO_ROW;
  uint32_t* out = p.out_idx + (size_t)job * p.k;
  for (uint32_t s = 0; s < p.slots; ++s) {
    const size_t o = (size_t)s * p.jobs_pad + job;
    const uint32_t c = p.part_cnt[o];
    if (c == 0) continue;
    const uint32_t l = p.part_last[o];
    if (s < p.slots_a) last_a = l; else last_b = l;  // slots are in row order
    if (have < p.k) {
      const uint32_t take = min(c, p.k - have);
      const uint32_t* src = p.part_idx + o * p.k;
      for (uint32_t i = 0; i < take; ++i) out[have + i] = src[i];
      have += take;
    }
    total += c;
  }
  for (uint32_t i = have; i < p.k; ++i) out[i] = DSM_NO_ROW;
  p.out_cnt[job] = total;
  uint32_t best = last_a;
  if (last_b != DSM_NO_ROW) {
    if (best == DSM_NO_ROW) best = last_b;
    else if (p.price != nullptr && p.price[last_b] > p.price[last_a]) best = last_b;
  }
  p.out_max[job] = best;
}

// ------------------------------------------------------------------------------------------
// small utility kernels
// ------------------------------------------------------------------------------------------

__global__ void k_gather_price(const double* price, const uint32_t* rows, uint32_t n, double* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rows[i];
  out[i] = (r == DSM_NO_ROW || price == nullptr) ? __longlong_as_double(0x7FF8000000000000LL) : price[r];
}

// shards are in table order: concatenate per-shard first-k lists, sum counts, pick the
// highest-priced max row.  One thread per job.
__global__ void k_merge_shards(uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                               const uint32_t* idx_all, const uint32_t* cnt_all,
                               const uint32_t* max_all, const double* maxprice_all,
                               const uint64_t* shard_base, uint64_t* idx_out, uint64_t* cnt_out,
                               uint64_t* max_out, double* maxprice_out) {
  const uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= n_jobs) return;
  uint64_t total = 0;
  uint32_t have = 0;
  uint64_t best = ~0ull;
  double best_price = 0.0;
  for (uint32_t s = 0; s < n_shards; ++s) {
    const size_t o = (size_t)s * n_jobs + job;
    const uint32_t c = cnt_all[o];
    if (c == 0) continue;
    const uint32_t take = min(min(c, k), k - have);
    for (uint32_t i = 0; i < take; ++i)
      idx_out[(size_t)job * k + have + i] = shard_base[s] + idx_all[o * k + i];
    have += take;
    total += c;
    const uint32_t m = max_all[o];
    if (m != DSM_NO_ROW) {
      const double pr = maxprice_all[o];
      if (best == ~0ull || pr > best_price) { best = shard_base[s] + m; best_price = pr; }
    }
  }
  for (uint32_t i = have; i < k; ++i) idx_out[(size_t)job * k + i] = ~0ull;
  cnt_out[job] = total;
  max_out[job] = best;
  maxprice_out[job] = (best == ~0ull) ? __longlong_as_double(0x7FF8000000000000LL) : best_price;
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------

struct dsm_table {
  int device;
  dsm_table_desc desc;
  uint4* rows;
  double* price;
  uint32_t* gpu_entries;
  uint32_t* bk_entries;
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
  std::mutex mu;
  uint64_t launches;
  // optional timing of the match kernel alone (bench.py's roofline): event pairs recorded
  // around each k_match launch while enabled, read back by dsm_timing_read
  bool timing = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timing_events;
  size_t timing_used = 0;
  DevBuf tabs, part_cnt, part_last, part_idx, d_jobs, d_idx, d_cnt, d_max, d_bits;
  DevBuf lut[DSM_MAX_COLS];
};

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

struct Plan {
  LayoutDev lay;
  uint32_t n_groups, n_blocks, split_block, ranges_a, ranges_b, slots, slots_a, grid, smem;
};

static int make_plan(const dsm_ctx* ctx, const dsm_table* t, const dsm_layout* L, uint32_t n_jobs,
                     Plan& pl) {
  if (!L || L->n_used < 1 || L->n_used > t->desc.n_cols)
    return fail(DSM_E_ARG, "layout.n_used out of range");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  LayoutDev& d = pl.lay;
  memset(&d, 0, sizeof d);
  d.n_used = L->n_used;
  d.rec_words = L->rec_words;
  uint32_t base = 0;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    if (L->kind[c] < DSM_COL_RANGE || L->kind[c] > DSM_COL_BK)
      return fail(DSM_E_ARG, "layout.kind invalid");
    if (L->card[c] == 0 || L->card[c] > 65536) return fail(DSM_E_ARG, "layout.card invalid");
    if (L->kind[c] == DSM_COL_GPU && L->card[c] != t->desc.n_gpu_entries)
      return fail(DSM_E_ARG, "GPU column card != table GPU dictionary size");
    if (L->kind[c] == DSM_COL_BK && L->card[c] != t->desc.n_bk_entries)
      return fail(DSM_E_ARG, "BK column card != table BK dictionary size");
    d.kind[c] = L->kind[c];
    d.card[c] = L->card[c];
    d.flags[c] = L->flags[c];
    d.aux[c] = L->aux[c];
    d.rec_off[c] = L->rec_off[c];
    d.base[c] = base;
    base += L->card[c];
  }
  d.n_entries = base;
  pl.smem = base * 16u + 2u * JPG * 4u;
  if (pl.smem > MAX_SMEM_BYTES)
    return fail(DSM_E_DICT_TOO_LARGE, "acceptance tables need more shared memory than one CTA has");
  pl.n_groups = (n_jobs + JPG - 1) / JPG;
  pl.n_blocks = t->desc.n_rows / 32u;
  pl.split_block = t->desc.split_row / 32u;

  // CTAs: two per SM when the tables leave room; work items ~2x the grid so the static
  // round-robin stays balanced; every warp of an item gets at least a few blocks.
  const uint32_t ctas_per_sm = pl.smem <= 100 * 1024 ? 2u : 1u;
  const uint32_t full_grid = (uint32_t)ctx->sm_count * ctas_per_sm;
  uint32_t want_ranges = (2u * full_grid + pl.n_groups - 1) / pl.n_groups;
  const uint32_t max_ranges = (pl.n_blocks + MATCH_WARPS * 4 - 1) / (MATCH_WARPS * 4);
  if (want_ranges > max_ranges) want_ranges = max_ranges;
  if (want_ranges < 1) want_ranges = 1;
  const uint32_t a_blocks = pl.split_block, b_blocks = pl.n_blocks - pl.split_block;
  uint32_t ra = a_blocks ? (uint32_t)(((uint64_t)want_ranges * a_blocks + pl.n_blocks - 1) / pl.n_blocks) : 0;
  uint32_t rb = b_blocks ? (uint32_t)(((uint64_t)want_ranges * b_blocks + pl.n_blocks - 1) / pl.n_blocks) : 0;
  if (a_blocks && ra == 0) ra = 1;
  if (b_blocks && rb == 0) rb = 1;
  pl.ranges_a = ra;
  pl.ranges_b = rb;
  pl.slots = (ra + rb) * MATCH_WARPS;
  pl.slots_a = ra * MATCH_WARPS;
  const uint32_t items = pl.n_groups * (ra + rb);
  pl.grid = items < full_grid ? items : full_grid;
  if (pl.grid == 0) pl.grid = 1;
  return 0;
}

template <int NC, bool BITMAP>
static cudaError_t launch_match_nc(const MatchParams& mp, bool has_lut, uint32_t grid, uint32_t smem,
                                   cudaStream_t st) {
  cudaError_t e;
  if (has_lut) {
    e = cudaFuncSetAttribute(k_match<NC, BITMAP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_match<NC, BITMAP, true><<<grid, MATCH_THREADS, smem, st>>>(mp);
  } else {
    e = cudaFuncSetAttribute(k_match<NC, BITMAP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_match<NC, BITMAP, false><<<grid, MATCH_THREADS, smem, st>>>(mp);
  }
  return cudaGetLastError();
}

template <bool BITMAP>
static cudaError_t launch_match(uint32_t nc, const MatchParams& mp, bool has_lut, uint32_t grid,
                                uint32_t smem, cudaStream_t st) {
  switch (nc) {
#define CASE_NC(n) case n: return launch_match_nc<n, BITMAP>(mp, has_lut, grid, smem, st);
    CASE_NC(1) CASE_NC(2) CASE_NC(3) CASE_NC(4) CASE_NC(5) CASE_NC(6) CASE_NC(7) CASE_NC(8)
    CASE_NC(9) CASE_NC(10) CASE_NC(11) CASE_NC(12) CASE_NC(13) CASE_NC(14) CASE_NC(15) CASE_NC(16)
#undef CASE_NC
  }
  return cudaErrorInvalidValue;
}

// Issue build -> match -> merge on the context's stream.  d_jobs and the outputs are device
// pointers.  Caller holds ctx->mu.
static int run_match(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* L, const uint32_t* d_jobs,
                     uint32_t n_jobs, uint32_t k, bool bitmap, uint32_t* d_idx, uint32_t* d_cnt,
                     uint32_t* d_max, uint32_t* d_bits) {
  if (t->device != ctx->device) return fail(DSM_E_ARG, "table and context are on different devices");
  Plan pl;
  int rc = make_plan(ctx, t, L, n_jobs, pl);
  if (rc) return rc;
  if (!bitmap && (k == 0 || k > 4096)) return fail(DSM_E_ARG, "k must be in 1..4096");
  CK(cudaSetDevice(ctx->device));

  const size_t jobs_pad = (size_t)pl.n_groups * JPG;
  if ((rc = ensure(ctx->tabs, (size_t)pl.n_groups * pl.lay.n_entries * 16))) return rc;
  if (!bitmap) {
    if ((rc = ensure(ctx->part_cnt, (size_t)pl.slots * jobs_pad * 4))) return rc;
    if ((rc = ensure(ctx->part_last, (size_t)pl.slots * jobs_pad * 4))) return rc;
    if ((rc = ensure(ctx->part_idx, (size_t)pl.slots * jobs_pad * k * 4))) return rc;
  }

  BuildParams bp;
  bp.lay = pl.lay;
  bp.recs = d_jobs;
  bp.n_jobs = n_jobs;
  bp.gpu_entries = t->gpu_entries;
  bp.bk_entries = t->bk_entries;
  bp.tabs = (uint4*)ctx->tabs.p;
  k_build_tables<<<dim3(pl.n_groups * TPG, pl.lay.n_used), 256, 0, ctx->stream>>>(bp);
  CK(cudaGetLastError());
  ctx->launches++;

  MatchParams mp;
  memset(&mp, 0, sizeof mp);
  // rows with 16 codes are read as two uint4 per row; the kernel indexes accordingly
  mp.rows = t->rows;
  mp.tabs = (const uint4*)ctx->tabs.p;
  mp.recs = d_jobs;
  bool has_lut = false;
  for (uint32_t c = 0; c < L->n_used; ++c) {
    mp.lut[c] = L->lut[c];
    has_lut = has_lut || L->lut[c] != nullptr;
    mp.base[c] = pl.lay.base[c];
  }
  mp.rec_words = L->rec_words;
  mp.n_entries = pl.lay.n_entries;
  mp.n_jobs = n_jobs;
  mp.n_groups = pl.n_groups;
  mp.n_blocks = pl.n_blocks;
  mp.split_block = pl.split_block;
  mp.ranges_a = pl.ranges_a;
  mp.ranges_b = pl.ranges_b;
  mp.k = k;
  mp.jobs_pad = (uint32_t)jobs_pad;
  mp.part_cnt = (uint32_t*)ctx->part_cnt.p;
  mp.part_last = (uint32_t*)ctx->part_last.p;
  mp.part_idx = (uint32_t*)ctx->part_idx.p;
  mp.bits = d_bits;
  // 16-code tables need every column slot the row has even if fewer are used: NC picks the
  // load width, so round NC up to 9 for 32-byte rows.
  uint32_t nc = L->n_used;
  if (t->desc.n_cols == 16 && nc < 9) return fail(DSM_E_ARG, "16-column tables need n_used >= 9");
  if (t->desc.n_cols == 8 && nc > 8) return fail(DSM_E_ARG, "8-column tables allow n_used <= 8");
  std::pair<cudaEvent_t, cudaEvent_t>* tev = nullptr;
  if (ctx->timing) {
    if (ctx->timing_used == ctx->timing_events.size()) {
      std::pair<cudaEvent_t, cudaEvent_t> pr;
      CK(cudaEventCreate(&pr.first));
      CK(cudaEventCreate(&pr.second));
      ctx->timing_events.push_back(pr);
    }
    tev = &ctx->timing_events[ctx->timing_used++];
    CK(cudaEventRecord(tev->first, ctx->stream));
  }
  cudaError_t e = bitmap ? launch_match<true>(nc, mp, has_lut, pl.grid, pl.smem, ctx->stream)
                         : launch_match<false>(nc, mp, has_lut, pl.grid, pl.smem, ctx->stream);
  if (e != cudaSuccess) return fail(DSM_E_CUDA, "k_match launch: %s", cudaGetErrorString(e));
  if (tev) CK(cudaEventRecord(tev->second, ctx->stream));
  ctx->launches++;

  if (!bitmap) {
    MergeParams gp;
    gp.part_cnt = mp.part_cnt;
    gp.part_last = mp.part_last;
    gp.part_idx = mp.part_idx;
    gp.price = t->price;
    gp.n_jobs = n_jobs;
    gp.jobs_pad = (uint32_t)jobs_pad;
    gp.k = k;
    gp.slots = pl.slots;
    gp.slots_a = pl.slots_a;
    gp.out_idx = d_idx;
    gp.out_cnt = d_cnt;
    gp.out_max = d_max;
    k_merge<<<(n_jobs + 127) / 128, 128, 0, ctx->stream>>>(gp);
    CK(cudaGetLastError());
    ctx->launches++;
  }
  return 0;
}

extern "C" {

const char* dsm_last_error(void) { return g_err.c_str(); }

const char* dsm_build_info(void) { return "libdsmatch 0.1 sm_100a jobs_per_group=128"; }

int dsm_ctx_create(int device, void* stream, dsm_ctx** out) {
  if (!out) return fail(DSM_E_ARG, "out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(DSM_E_NO_DEVICE, "no CUDA device: %s", cudaGetErrorString(e));
  if (device < 0 || device >= n) return fail(DSM_E_ARG, "device ordinal out of range");
  CK(cudaSetDevice(device));
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
  *out = c;
  return 0;
}

int dsm_ctx_destroy(dsm_ctx* ctx) {
  if (!ctx) return 0;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->tabs, &ctx->part_cnt, &ctx->part_last, &ctx->part_idx, &ctx->d_jobs,
                    &ctx->d_idx, &ctx->d_cnt, &ctx->d_max, &ctx->d_bits};
  for (DevBuf* b : bufs) if (b->p) cudaFree(b->p);
  for (int c = 0; c < DSM_MAX_COLS; ++c) if (ctx->lut[c].p) cudaFree(ctx->lut[c].p);
  for (auto& pr : ctx->timing_events) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return 0;
}

int dsm_table_create(dsm_ctx* ctx, const dsm_table_desc* desc, const uint16_t* rows,
                     const double* price, const uint32_t* gpu_entries, const uint32_t* bk_entries,
                     dsm_table** out) {
  if (!ctx || !desc || !rows || !out) return fail(DSM_E_ARG, "NULL argument");
  if (desc->n_cols != 8 && desc->n_cols != 16) return fail(DSM_E_ARG, "n_cols must be 8 or 16");
  if (desc->n_rows == 0 || desc->n_rows % 32 || desc->split_row % 32 || desc->split_row > desc->n_rows)
    return fail(DSM_E_ARG, "n_rows / split_row must be multiples of 32, split_row <= n_rows");
  std::lock_guard<std::mutex> g(ctx->mu);
  CK(cudaSetDevice(ctx->device));
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
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
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
  cudaSetDevice(t->device);
  if (t->rows) cudaFree(t->rows);
  if (t->price) cudaFree(t->price);
  if (t->gpu_entries) cudaFree(t->gpu_entries);
  if (t->bk_entries) cudaFree(t->bk_entries);
  delete t;
  return 0;
}

int dsm_lut_upload(dsm_ctx* ctx, uint32_t column, const uint16_t* lut, uint32_t n,
                   const uint16_t** dev_out) {
  if (!ctx || !lut || !dev_out || column >= DSM_MAX_COLS || n == 0) return fail(DSM_E_ARG, "bad argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  int rc = ensure(ctx->lut[column], (size_t)n * 2);
  if (rc) return rc;
  CK(cudaMemcpyAsync(ctx->lut[column].p, lut, (size_t)n * 2, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  *dev_out = (const uint16_t*)ctx->lut[column].p;
  return 0;
}

int dsm_match_topk_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                       const uint32_t* d_jobs, uint32_t n_jobs, uint32_t k, uint32_t* d_idx,
                       uint32_t* d_cnt, uint32_t* d_max) {
  if (!ctx || !t || !layout || !d_jobs || !d_idx || !d_cnt || !d_max) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  return run_match(ctx, t, layout, d_jobs, n_jobs, k, false, d_idx, d_cnt, d_max, nullptr);
}

int dsm_match_bitmap_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                         const uint32_t* d_jobs, uint32_t n_jobs, uint32_t* d_bits) {
  if (!ctx || !t || !layout || !d_jobs || !d_bits) return fail(DSM_E_ARG, "NULL argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  return run_match(ctx, t, layout, d_jobs, n_jobs, 1, true, nullptr, nullptr, nullptr, d_bits);
}

int dsm_match_topk(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const uint32_t* jobs,
                   uint32_t n_jobs, uint32_t k, uint32_t* out_idx, uint32_t* out_cnt,
                   uint32_t* out_max) {
  if (!ctx || !t || !layout || !jobs || !out_idx || !out_cnt || !out_max) return fail(DSM_E_ARG, "NULL argument");
  if (n_jobs == 0) return fail(DSM_E_ARG, "n_jobs must be > 0");
  std::lock_guard<std::mutex> g(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  int rc;
  const size_t jb = (size_t)n_jobs * layout->rec_words * 4;
  if ((rc = ensure(ctx->d_jobs, jb))) return rc;
  if ((rc = ensure(ctx->d_idx, (size_t)n_jobs * k * 4))) return rc;
  if ((rc = ensure(ctx->d_cnt, (size_t)n_jobs * 4))) return rc;
  if ((rc = ensure(ctx->d_max, (size_t)n_jobs * 4))) return rc;
  CK(cudaMemcpyAsync(ctx->d_jobs.p, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  rc = run_match(ctx, t, layout, (const uint32_t*)ctx->d_jobs.p, n_jobs, k, false,
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
  std::lock_guard<std::mutex> g(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  int rc;
  const size_t jb = (size_t)n_jobs * layout->rec_words * 4;
  const size_t bb = (size_t)n_jobs * (t->desc.n_rows / 32) * 4;
  if ((rc = ensure(ctx->d_jobs, jb))) return rc;
  if ((rc = ensure(ctx->d_bits, bb))) return rc;
  CK(cudaMemcpyAsync(ctx->d_jobs.p, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  rc = run_match(ctx, t, layout, (const uint32_t*)ctx->d_jobs.p, n_jobs, 1, true, nullptr, nullptr,
                 nullptr, (uint32_t*)ctx->d_bits.p);
  if (rc) return rc;
  CK(cudaMemcpyAsync(out_bits, ctx->d_bits.p, bb, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int dsm_merge_shards_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                         const uint32_t* d_idx_all, const uint32_t* d_cnt_all,
                         const uint32_t* d_max_all, const double* d_maxprice_all,
                         const uint64_t* d_shard_base, uint64_t* d_idx_out, uint64_t* d_cnt_out,
                         uint64_t* d_max_out, double* d_maxprice_out) {
  if (!ctx || !d_idx_all || !d_cnt_all || !d_max_all || !d_maxprice_all || !d_shard_base ||
      !d_idx_out || !d_cnt_out || !d_max_out || !d_maxprice_out)
    return fail(DSM_E_ARG, "NULL argument");
  if (n_shards == 0 || n_jobs == 0 || k == 0) return fail(DSM_E_ARG, "empty merge");
  std::lock_guard<std::mutex> g(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  k_merge_shards<<<(n_jobs + 127) / 128, 128, 0, ctx->stream>>>(
      n_shards, n_jobs, k, d_idx_all, d_cnt_all, d_max_all, d_maxprice_all, d_shard_base, d_idx_out,
      d_cnt_out, d_max_out, d_maxprice_out);
  CK(cudaGetLastError());
  ctx->launches++;
  return 0;
}

int dsm_gather_price_dev(dsm_ctx* ctx, const dsm_table* t, const uint32_t* d_rows, uint32_t n,
                         double* d_out) {
  if (!ctx || !t || !d_rows || !d_out || n == 0) return fail(DSM_E_ARG, "bad argument");
  std::lock_guard<std::mutex> g(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  k_gather_price<<<(n + 255) / 256, 256, 0, ctx->stream>>>(t->price, d_rows, n, d_out);
  CK(cudaGetLastError());
  ctx->launches++;
  return 0;
}

int dsm_sync(dsm_ctx* ctx) {
  if (!ctx) return fail(DSM_E_ARG, "NULL ctx");
  CK(cudaSetDevice(ctx->device));
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
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  double sum = 0.0;
  for (size_t i = 0; i < ctx->timing_used; ++i) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, ctx->timing_events[i].first, ctx->timing_events[i].second));
    sum += ms;
  }
  *total_ms = sum;
  *n_launches = (uint32_t)ctx->timing_used;
  ctx->timing_used = 0;
  return 0;
}

int dsm_plan_query(const dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, uint32_t n_jobs,
                   uint32_t out[8]) {
  if (!ctx || !t || !layout || !out) return fail(DSM_E_ARG, "NULL argument");
  Plan pl;
  int rc = make_plan(ctx, t, layout, n_jobs, pl);
  if (rc) return rc;
  out[0] = pl.grid;
  out[1] = pl.n_groups * (pl.ranges_a + pl.ranges_b);
  out[2] = pl.ranges_a;
  out[3] = pl.ranges_b;
  out[4] = pl.slots;
  out[5] = pl.smem;
  out[6] = pl.lay.n_entries;
  out[7] = pl.n_groups;
  return 0;
}

}  // extern "C"
