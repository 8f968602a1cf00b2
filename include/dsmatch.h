/*
 * dsmatch.h -- C ABI of libdsmatch.so: B200 (sm_100a) offer matching for dstack.
 *
 * The reference (solovyevt/dstack) has NO FFI for this path: the hot loops are plain
 * Python.  These entry points are therefore the boundary a
 * maintainer would bind with ctypes; each one names the reference
 * code whose work it takes over (paths relative to src/dstack/_internal).
 *
 * Model
 *   table   a device-resident, dictionary-encoded offer table: one row of n_cols u16 codes
 *           per offer, rows stored in the reference's result order
 *           (not-available last, then price, then project-backend order, then in-backend order:
 *           server/services/backends/__init__.py:449-451), so "top-k by price" is "first k
 *           matches" and a job's max_price is a row-index cut-off.
 *   jobs    one fixed-size predicate record per job (u32 words, format below), built by the
 *           host from Requirements + profile filters
 *           (core/backends/base/offers.py:112-146, server/services/offers.py:46-100).
 *   match   every (job, row) pair is evaluated on the GPU: a table-build kernel turns the
 *           records into bit-sliced acceptance tables (bit j of word = job j accepts this
 *           dictionary entry), the match kernel ANDs one gathered word per column for 32
 *           rows x 128 jobs at a time, and a merge kernel orders the per-range partial
 *           results.  This replaces gpuhunt.matches per (job, offer)
 *           (core/backends/base/offers.py:154-157, server/services/pools.py:496-502,533-543)
 *           and the merge / partition / slice ranker (backends/__init__.py:449-451,
 *           background/tasks/process_submitted_jobs.py:450, services/runs.py:341-348).
 *
 * All functions return 0 on success or a negative DSM_E_* code; dsm_last_error() gives the
 * message for the calling thread.  No exceptions cross the boundary.  A context owns a CUDA
 * stream and its scratch memory; calls on one context serialise, different contexts may be
 * used from different host threads concurrently.  Nothing here falls back to the CPU.
 */
#ifndef DSMATCH_H
#define DSMATCH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSM_MAX_COLS 16
#define DSM_JOBS_PER_GROUP 128u /* J_t: jobs evaluated per pass over the rows */
#define DSM_NO_ROW 0xFFFFFFFFu

enum dsm_error {
  DSM_OK = 0,
  DSM_E_CUDA = -1,         /* a CUDA runtime call failed (message has the cudaError string) */
  DSM_E_ARG = -2,          /* bad argument */
  DSM_E_DICT_TOO_LARGE = -3, /* acceptance tables of one job group exceed shared memory */
  DSM_E_NO_DEVICE = -4,
  DSM_E_TOO_LARGE = -5     /* request exceeds a fixed-capacity entry point's limit */
};

/* Column kinds: how the table-build kernel turns a job's record into acceptance bits. */
enum dsm_col_kind {
  DSM_COL_UNUSED = 0,
  /* 1 word: lo | hi<<16 (inclusive ordinals).  Entry e passes iff lo <= e <= hi; with
   * DSM_COLF_NULL_PASSES the last entry (card-1) passes for every job.
   * cpu / memory / disk bounds: App. A.1 items 4,5,12 of SURVEY.md. */
  DSM_COL_RANGE = 1,
  /* ceil(card/32) words: entry e passes iff bit e is set (region, instance type, fleet,
   * instance name, zone ... sets: server/services/offers.py:93-100, pools.py:458-483). */
  DSM_COL_BITSET = 2,
  /* accelerator configuration dictionary, record = 5 + aux words:
   *   w0 count lo|hi<<16   w1 memory lo|hi<<16   w2 total-memory lo|hi<<16
   *   w3 cc_min | flags<<16 (DSM_GPUF_*)   w4 vendor mask (bit = vendor code) | blocks<<16
   *   w5.. GPU-name bitset (aux words).  App. A.1 items 6-11; blocks != 0 also requires the
   *   entry's GPU count to be divisible by it (server/services/offers.py:102-118). */
  DSM_COL_GPU = 3,
  /* backend/spot/status/flags dictionary, record = 4 words:
   *   w0 serving-backend mask  w1 offer-backend mask  w2 spot mask | status mask<<8
   *   w3 required flags | forbidden flags<<16.  App. A.1 item 3 and
   *   server/services/offers.py:81-94, pools.py:460-488. */
  DSM_COL_BK = 4,
  /* two range attributes stored as ONE code: the id of the (ordinal A, ordinal B) pair the row
   * carries, so that one table gather tests both (cpu and memory: App. A.1 items 4, 5).  The
   * table holds the pair dictionary (dsm_table_create: pair_entries, A | B<<16 per entry).
   * Record = aux words for member A -- one range word, or with DSM_COLF_PAIR_A_BITSET a bitset
   * over A's ordinals -- followed by member B's range word. */
  DSM_COL_PAIR = 5
};

#define DSM_COLF_NULL_PASSES 1u
#define DSM_COLF_PAIR_A_BITSET 2u

#define DSM_GPUF_MEM_BOUNDED 1u
#define DSM_GPUF_CC_BOUNDED 2u
#define DSM_GPUF_NAME_ANY 8u

/* Dictionary entry of a DSM_COL_GPU column (3 words, built by the host encoder):
 *   w0 count ordinal | memory ordinal<<16
 *   w1 total-memory ordinal | compute-capability ordinal<<16 (0xFFFF unknown)
 *   w2 name id (0xFFFF none) | vendor code<<16 | has_gpu<<24 | GPUs counted for block
 *      splitting<<25 (7 bits, saturated)
 * Dictionary entry of a DSM_COL_BK column (2 words):
 *   w0 serving backend | offer backend<<8 | spot<<16 | status<<24      w1 flags
 * Entry 0 of the DSM_COL_BK column is reserved as the dead entry: rows that carry it never
 * match (padding rows). */
#define DSM_GPU_ENTRY_WORDS 3
#define DSM_BK_ENTRY_WORDS 2

typedef struct dsm_table_desc {
  uint32_t n_rows;     /* physical rows incl. padding; multiple of 32 */
  uint32_t split_row;  /* rows [0,split) are the "available" segment, [split,n_rows) the rest;
                          multiple of 32 */
  uint32_t n_cols;     /* codes per row: 8 (16-byte rows) or 16 (32-byte rows) */
  uint32_t n_gpu_entries; /* entries of the DSM_COL_GPU dictionary (0 if none) */
  uint32_t n_bk_entries;  /* entries of the DSM_COL_BK dictionary (0 if none) */
  uint32_t n_pair_entries;/* entries of the DSM_COL_PAIR dictionary (0 if none) */
  uint32_t n_live_a;      /* real rows of the first segment: rows [n_live_a, split_row) are padding */
  uint32_t n_live_b;      /* real rows of the second: rows [split_row + n_live_b, n_rows) are padding.
                             Padding rows never match, whatever a record's cut-offs say. */
} dsm_table_desc;

/* Per-call description of the job records and of the acceptance tables built from them. */
typedef struct dsm_layout {
  uint32_t n_used;               /* columns that take part, 1..n_cols; column i reads code i */
  uint32_t kind[DSM_MAX_COLS];
  uint32_t card[DSM_MAX_COLS];   /* table entries of the column for this call */
  uint32_t flags[DSM_MAX_COLS];  /* DSM_COLF_* */
  uint32_t aux[DSM_MAX_COLS];    /* DSM_COL_GPU: words of the GPU-name bitset */
  uint32_t rec_off[DSM_MAX_COLS];/* word offset of the column's predicate in a record */
  uint32_t rec_words;            /* record stride in u32 words; word 0 = cut-off row inside
                                    the first segment, word 1 = cut-off row inside the second
                                    (rows >= cut-off fail: max_price / exclude_not_available) */
  /* optional per-call remap of a column's codes (device pointer to card_of_table u16
   * values, or NULL): lets a high-cardinality column be looked up through the few classes
   * the jobs of this call mention.  Set by dsm_lut_upload. */
  const uint16_t* lut[DSM_MAX_COLS];
  /* optional evaluation order of the jobs: a permutation of 0..n_jobs-1 (u32), NULL = as given.
   * Jobs are evaluated 128 at a time; a group whose jobs all leave the trailing columns open
   * (no region filter, default disk, ...) skips those columns' gathers, so a host that puts
   * similar jobs next to each other in this order gets a faster match.  Results are always
   * returned in the caller's job order.  The array lives where the call's job records live:
   * host memory for dsm_match_topk / dsm_match_bitmap, device memory for the *_dev and
   * *_static entry points. */
  const uint32_t* job_order;
  /* which attribute of a raw job (dsm_raw_job) feeds the column: DSM_ROLE_*.  Only read by the
   * entry points that build the records on the device (dsm_jobs_pack_dev, dsm_match_topk_raw). */
  uint32_t role[DSM_MAX_COLS];
} dsm_layout;

/* ---- raw jobs: the record building of core/backends/base/offers.py:112-146 on the device ----
 *
 * A raw job carries a job's bounds as the reference states them (IEEE doubles, NaN = not set)
 * and its set filters as ids of the table's dictionaries; the device turns them into the
 * predicate records above with the same double comparisons the reference makes per offer --
 * bound -> ordinal by binary search in the table's sorted value lists, max_price -> cut-off row by
 * binary search in the table's prices.  Little-endian, 8-byte aligned, 168 bytes. */
typedef struct dsm_raw_job {
  double max_price;       /* NaN: none */
  double cpu[2];          /* min, max */
  double memory[2];       /* GiB */
  double disk[2];         /* GiB */
  double gpu_count[2];
  double gpu_memory[2];   /* GiB */
  double gpu_total[2];    /* GiB */
  int32_t cc_min;         /* min compute capability as major*1000+minor, -1: none */
  int32_t blocks;         /* 0: no block filter, b > 1: cpu and GPU count divisible by b, < 0: nothing is */
  uint32_t src_backend_mask, off_backend_mask;
  uint16_t flag_req, flag_forbid;
  uint8_t status_mask;    /* bit per status code, 0xFF any */
  int8_t spot;            /* -1 either, 0 on-demand, 1 spot */
  int8_t gpu_vendor;      /* vendor code, 0 = no filter */
  uint8_t flags;          /* DSM_RAWF_* */
  int32_t fleet_id;       /* dictionary id, -1: a fleet the table does not know */
  int32_t name_id;
  uint32_t gpu_name_off, gpu_name_cnt;   /* ranges of the call's id array (dictionary ids, -1 = */
  uint32_t region_off, region_cnt;       /* a value the table does not contain)                 */
  uint32_t itype_off, itype_cnt;
  uint32_t zone_off, zone_cnt;
} dsm_raw_job;

#define DSM_RAWF_EXCLUDE_NOT_AVAILABLE 1u
#define DSM_RAWF_GPU_NAME_ANY 2u
#define DSM_RAWF_REGION_ANY 4u
#define DSM_RAWF_ITYPE_ANY 8u
#define DSM_RAWF_FLEET_ANY 16u
#define DSM_RAWF_NAME_ANY 32u
#define DSM_RAWF_ZONE_ANY 64u

enum dsm_col_role {
  DSM_ROLE_NONE = 0, DSM_ROLE_CPU = 1, DSM_ROLE_MEMORY = 2, DSM_ROLE_CPU_MEMORY = 3, DSM_ROLE_DISK = 4,
  DSM_ROLE_GPU = 5, DSM_ROLE_BK = 6, DSM_ROLE_REGION = 7, DSM_ROLE_ITYPE = 8, DSM_ROLE_FLEET = 9,
  DSM_ROLE_NAME = 10, DSM_ROLE_ZONE = 11,
  /* pool tables: entry = status * 16 + (unreachable | blocks unknown << 1 | shared << 2 | no offer << 3);
   * accepted iff the status is in status_mask and the flags satisfy flag_req / flag_forbid (their low
   * four bits; the remaining bits are tested on the DSM_ROLE_BK column) */
  DSM_ROLE_STATE = 12
};

/* sorted distinct values of the table's range attributes, in this order */
enum dsm_pack_dict {
  DSM_DICT_CPU = 0, DSM_DICT_MEMORY = 1, DSM_DICT_DISK = 2, DSM_DICT_GPU_COUNT = 3, DSM_DICT_GPU_MEMORY = 4,
  DSM_DICT_GPU_TOTAL = 5, DSM_DICT_CC = 6, DSM_N_PACK_DICTS = 7
};

typedef struct dsm_ctx dsm_ctx;
typedef struct dsm_table dsm_table;

/* ---- lifecycle ------------------------------------------------------------------------- */

/* Create a context on `device` (a CUDA ordinal).  stream = 0 makes the context own a new
 * non-blocking stream; otherwise pass a cudaStream_t to run on (e.g. torch's current stream). */
int dsm_ctx_create(int device, void* stream, dsm_ctx** out);
int dsm_ctx_destroy(dsm_ctx* ctx);
const char* dsm_last_error(void);
/* Compile-time facts: "sm_100a", jobs per group, version. */
const char* dsm_build_info(void);

/* Upload an encoded table.  rows: n_rows*n_cols u16 (host).  price: n_rows f64 (host), used
 * only to pick the max-price match; may be NULL.  gpu_entries / bk_entries: dictionaries of
 * the composite columns (host); pair_entries: n_pair_entries u32 (NULL if none).  Replaces the
 * per-(job,offer) offer_to_catalog_item walk (core/backends/base/offers.py:86-109): attributes
 * are encoded once per table. */
int dsm_table_create(dsm_ctx* ctx, const dsm_table_desc* desc, const uint16_t* rows,
                     const double* price, const uint32_t* gpu_entries,
                     const uint32_t* bk_entries, const uint32_t* pair_entries, dsm_table** out);
int dsm_table_destroy(dsm_table* t);

/* Rewrite n rows of a resident table in place: rows[i] = physical row, codes = n * n_cols u16
 * (host).  For attributes that change between scheduler passes without adding a dictionary
 * value (an instance's status, reachability, busy blocks, fleet: server/services/pools.py:
 * 458-488 reads them fresh for every job).  Row order and prices are not touched.  Serialised
 * with the matches of `ctx` (same stream); matches on other contexts must not overlap it. */
int dsm_table_patch_rows(dsm_ctx* ctx, dsm_table* t, const uint32_t* rows, uint32_t n, const uint16_t* codes);

/* Copy a per-call code->class remap to the device (see dsm_layout.lut).  The returned device
 * pointer stays valid until the next dsm_lut_upload for the same (ctx, column). */
int dsm_lut_upload(dsm_ctx* ctx, uint32_t column, const uint16_t* lut, uint32_t n,
                   const uint16_t** dev_out);

/* ---- matching, host buffers (H2D of the records and D2H of the results inside) ---------- */

/* For each job: the first k matching rows in table order (= the reference's offers[:k]),
 * the number of matching rows (= len(offers)) and the row holding the highest price among
 * the matches (= max(offer.price)), DSM_NO_ROW if none.
 * out_idx: n_jobs*k u32, unused slots DSM_NO_ROW.  out_cnt, out_max: n_jobs u32.  1 <= k <= 4096.
 * Scratch (owned by the context, grown on demand, independent of k): about
 * n_jobs * n_rows / 120 bytes of per-window match records; a call that would need more than
 * 64 GiB of it returns DSM_E_TOO_LARGE -- split the jobs over several calls.
 * Reference: get_instance_offers + slicing, see the header comment. */
int dsm_match_topk(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                   const uint32_t* jobs, uint32_t n_jobs, uint32_t k, uint32_t* out_idx,
                   uint32_t* out_cnt, uint32_t* out_max);

/* For each job the full match bitmap: out_bits[j*(n_rows/32) + b] bit l = row 32*b+l matches.
 * Used where the reference returns every match in input order: match_requirements
 * (core/backends/base/offers.py:149-158) and filter_pool_instances (pools.py:409-503). */
int dsm_match_bitmap(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                     const uint32_t* jobs, uint32_t n_jobs, uint32_t* out_bits);

/* Give a table the value lists raw jobs are translated with: `values` = the DSM_N_PACK_DICTS
 * lists one after the other (host, f64, each ascending), counts[i] = entries of list i. */
int dsm_table_set_pack_dicts(dsm_ctx* ctx, dsm_table* t, const double* values,
                             const uint32_t counts[DSM_N_PACK_DICTS]);

/* dsm_match_topk with the records built on the device: raw jobs and their id array (host) in,
 * results (host) out.  layout.role[] names the attribute behind every column; a column whose
 * kind is DSM_COL_BITSET and role DSM_ROLE_CPU (or a DSM_COL_PAIR with DSM_COLF_PAIR_A_BITSET)
 * gets the "in range and divisible by blocks" set.  layout.job_order must be NULL: the
 * evaluation order (jobs that need the same leading columns side by side) is chosen on the
 * device.  Columns with a per-call remap (layout.lut) are not supported here -- pack those
 * calls on the host.  Reference: requirements_to_query_filter, offers.py:112-146, once per job. */
int dsm_match_topk_raw(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                       const dsm_raw_job* jobs, uint32_t n_jobs, const int32_t* ids, uint32_t n_ids,
                       uint32_t k, uint32_t* out_idx, uint32_t* out_cnt, uint32_t* out_max);

/* The same, answered in the caller's terms: out_src = SOURCE positions of the first k matches
 * (-1 = none; needs dsm_table_set_source_map), out_cnt = number of matches, out_max_price = the
 * highest price among them (NaN = none) -- offers[:k], len(offers) and max(offer.price) of
 * server/services/runs.py:341-348 without a host-side pass over the rows. */
int dsm_match_topk_raw_src(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                           const dsm_raw_job* jobs, uint32_t n_jobs, const int32_t* ids, uint32_t n_ids,
                           uint32_t k, int32_t* out_src, uint32_t* out_cnt, double* out_max_price);

/* Job filters as columns (one array per attribute, the layout of a batch of Requirements +
 * profile filters before any encoding): what dsm_match_topk_cols takes.  Arrays are host memory,
 * n_jobs long unless noted; a NULL pointer means "not set for any job" (any-flags: every job
 * accepts everything; exclude_not_available, blocks, flag_req / flag_forbid: zero; status_mask:
 * 0xFF; fleet_ids / name_ids: -1).  Set filters are CSR: *_ptr has n_jobs + 1 entries into *_ids
 * (dictionary ids of the table, -1 = a value the table does not contain). */
typedef struct dsm_job_columns {
  const double* max_price;      /* NaN: none */
  const double* cpu;            /* [n_jobs][2] min, max; NaN = not set */
  const double* memory;
  const double* disk;
  const double* gpu_count;
  const double* gpu_memory;
  const double* gpu_total;
  const int64_t* cc_min;        /* major*1000+minor, -1: none */
  const int64_t* blocks;
  const uint32_t* src_backend_mask;
  const uint32_t* off_backend_mask;
  const uint16_t* flag_req;
  const uint16_t* flag_forbid;
  const uint8_t* status_mask;
  const int8_t* spot;
  const int8_t* gpu_vendor;
  const uint8_t* exclude_not_available; /* 0 / 1 */
  const uint8_t* gpu_name_any;
  const uint8_t* region_any;
  const uint8_t* itype_any;
  const uint8_t* fleet_any;
  const uint8_t* name_any;
  const uint8_t* zone_any;
  const int64_t* fleet_ids;
  const int64_t* name_ids;
  const int64_t* gpu_name_ptr; const int64_t* gpu_name_ids;
  const int64_t* region_ptr;   const int64_t* region_ids;
  const int64_t* itype_ptr;    const int64_t* itype_ids;
  const int64_t* zone_ptr;     const int64_t* zone_ids;
} dsm_job_columns;

/* dsm_match_topk_raw_src from job columns: the library lays the raw jobs out itself (into a
 * pinned staging buffer of the context), so a Python host passes its arrays as they are. */
int dsm_match_topk_cols(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                        const dsm_job_columns* cols, uint32_t n_jobs, uint32_t k, int32_t* out_src,
                        uint32_t* out_cnt, double* out_max_price);

/* The packing step alone, device buffers: d_jobs/d_ids -> d_records [n_jobs][rec_words] and, if
 * d_order is not NULL, the evaluation order [n_jobs].  Asynchronous on the ctx stream. */
int dsm_jobs_pack_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                      const dsm_raw_job* d_jobs, uint32_t n_jobs, const int32_t* d_ids,
                      uint32_t* d_records, uint32_t* d_order);

/* ---- every match as an index list (filter_pool_instances, match_requirements) ----------- */

/* Give a table the source position of every physical row (host, i32 [n_rows], -1 = padding). */
int dsm_table_set_source_map(dsm_ctx* ctx, dsm_table* t, const int32_t* row_to_source, uint32_t n_source);

/* dsm_match_bitmap, with the bitmaps turned into per-job lists of SOURCE positions on the
 * device: out_ptr[j]..out_ptr[j+1] delimits job j's matches inside *out_ids (n_jobs + 1 u64,
 * host).  by_source != 0: ascending source position (the reference's "input order",
 * pools.py:409-503, offers.py:149-158); 0: table order (price rank).  *out_ids points into a
 * pinned buffer owned by the context, valid until its next call.  Needs
 * dsm_table_set_source_map. */
int dsm_match_indices(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, const uint32_t* jobs,
                      uint32_t n_jobs, int by_source, uint64_t* out_ptr, const uint32_t** out_ids);

/* ---- matching, device buffers (inputs already resident; asynchronous on the ctx stream) - */

int dsm_match_topk_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                       const uint32_t* d_jobs, uint32_t n_jobs, uint32_t k, uint32_t* d_idx,
                       uint32_t* d_cnt, uint32_t* d_max);
int dsm_match_bitmap_dev(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                         const uint32_t* d_jobs, uint32_t n_jobs, uint32_t* d_bits);

/* dsm_match_bitmap_dev for callers that must not allocate.  This entry point has no counterpart
 * in the reference (which never touches a GPU); it exists for two kinds of host:
 *   - a scheduler thread sharing the GPU with serving or training streams, which cannot afford
 *     the device-wide synchronisation and driver locking a cudaMalloc / cudaFree may cause;
 *   - a process whose device memory is already fully owned by a pool (a framework's caching
 *     allocator, a fixed MIG slice), where a fresh allocation can be refused.
 * Same arguments and same result as dsm_match_bitmap_dev: d_jobs (records) and d_bits (output,
 * n_jobs * n_rows/32 words) are the caller's device buffers.  The scratch the other entry
 * points allocate -- the acceptance tables, one word per job group (the columns the group
 * needs) and the work counter -- lives in a fixed array of DSM_STATIC_SCRATCH_BYTES inside the
 * library's device image, so every GPU that loads the library holds that much until the
 * process exits.  No cudaMalloc, cudaFree or other allocating call is made.  layout.job_order,
 * if used, is a device array like d_jobs.
 *   Limit: with g = ceil(n_jobs/128) job groups, 256 + (4 g rounded up to 256) +
 *   g * (sum of layout.card) * 16 bytes must fit in the array (e.g. 562 dictionary entries ->
 *   up to ~29,000 jobs per call); larger requests fail with
 *   DSM_E_TOO_LARGE and a message saying how much was needed.  A layout that uses lut[] needs
 *   its dsm_lut_upload done beforehand (that call does allocate).
 *   First call on a GPU: the array's device address is looked up once per device
 *   (cudaGetSymbolAddress, with that device current); make that first call before starting any
 *   stream capture.
 *   Sharing: all calls on a GPU use the one array and it keeps the previous call's contents.
 *   Everything that has to start from zero (the work counter) is reset by the call itself, the
 *   tables and the per-group words are rewritten for every job group used, but two calls whose kernels overlap on
 *   the same GPU -- different contexts, streams or threads -- race on it: serialise them. */
#define DSM_STATIC_SCRATCH_BYTES (2u << 20)
int dsm_match_bitmap_static(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                            const uint32_t* d_jobs, uint32_t n_jobs, uint32_t* d_bits);

/* Merge per-shard results of a catalog sharded over n_shards GPUs in table order
 * (shard s holds rows that rank before shard s+1).  Inputs are the all-gathered device
 * arrays [n_shards][n_jobs][k], [n_shards][n_jobs], [n_shards][n_jobs] and the per-shard
 * max-price values [n_shards][n_jobs] f64; indices come out as global rows
 * (shard_base[s] + local).  Asynchronous on the ctx stream. */
int dsm_merge_shards_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                         const uint32_t* d_idx_all, const uint32_t* d_cnt_all,
                         const uint32_t* d_max_all, const double* d_maxprice_all,
                         const uint64_t* d_shard_base, uint64_t* d_idx_out, uint64_t* d_cnt_out,
                         uint64_t* d_max_out, double* d_maxprice_out);

/* Same merge when each shard's four result arrays travel in ONE all-gathered record (one
 * collective per step instead of four): the pointers address shard 0's arrays and shard s's
 * lie s * shard_stride_bytes further on. */
int dsm_merge_shards_strided_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                                 const uint32_t* d_idx0, const uint32_t* d_cnt0, const uint32_t* d_max0,
                                 const double* d_maxprice0, size_t shard_stride_bytes,
                                 const uint64_t* d_shard_base, uint64_t* d_idx_out, uint64_t* d_cnt_out,
                                 uint64_t* d_max_out, double* d_maxprice_out);

/* Merge for shards that were dealt the result order in chunks: position p of the global order
 * belongs to shard (p / chunk_rows) % n_shards, so all shards see the same price mix and finish
 * together.  d_shard_geom = [n_shards][2] u32: split_row and number of live first-segment rows
 * (n_avail) of every shard's table.  Outputs are positions in the global result order (not
 * physical rows): the k smallest over all shards, ~0 = none.  Strided input as above;
 * n_shards <= 32, k <= 255.  Like the other merges this has no counterpart in the reference. */
int dsm_merge_shards_interleaved_dev(dsm_ctx* ctx, uint32_t n_shards, uint32_t n_jobs, uint32_t k,
                                     const uint32_t* d_idx0, const uint32_t* d_cnt0, const uint32_t* d_max0,
                                     const double* d_maxprice0, size_t shard_stride_bytes,
                                     const uint32_t* d_shard_geom, uint32_t chunk_rows,
                                     uint64_t* d_idx_out, uint64_t* d_cnt_out, uint64_t* d_max_out,
                                     double* d_maxprice_out);

/* Gather price[row] for rows produced by dsm_match_topk_dev (DSM_NO_ROW -> NaN). */
int dsm_gather_price_dev(dsm_ctx* ctx, const dsm_table* t, const uint32_t* d_rows, uint32_t n,
                         double* d_out);

/* ---- peer exchange: the sharded step's all-gather over NVLink peer memory ---------------------
 *
 * One dsm_xchg per rank (one rank per GPU; the ranks may be processes or threads).  Every rank
 * owns a device buffer [2][world][slot] that its peers map: across processes with the CUDA IPC
 * handle dsm_xchg_create returns (exchange the world handles with whatever the host has --
 * torch.distributed, MPI, a pipe -- and pass all of them to dsm_xchg_connect_ipc), inside one
 * process with the peers' base pointers (dsm_xchg_local_base -> dsm_xchg_connect_ptrs).
 *
 * dsm_xchg_all_gather_dev enqueues, on the context's stream, (1) a kernel that stores the
 * `bytes` at d_src into slot[rank] of every rank's buffer and then raises this rank's flag on
 * every peer, (2) a kernel that waits until the flags of all ranks show this call's epoch.  Work
 * enqueued after it may read *d_gathered = [world][*slot_stride_bytes]: the records of all ranks
 * in rank order, in local memory.  Every rank must make the same sequence of calls; buffers
 * alternate between two parities, so a rank may run at most one call ahead of its peers, which
 * the wait itself guarantees.  The wait gives up after 5 s (a dead peer must not leave this GPU
 * spinning): dsm_xchg_status synchronises the stream and returns 0 if no wait ever timed out,
 * else 1 + the highest rank that was missing -- the gathered data of that call is then undefined.
 * Destroy only after every rank has finished its last call (host barrier).
 * Replaces the NCCL all_gather of the step in dstack_b200/sharded.py; no counterpart in the
 * reference, which runs one scheduler process against one catalog. */
typedef struct dsm_xchg dsm_xchg;
#define DSM_IPC_HANDLE_BYTES 64
int dsm_xchg_create(dsm_ctx* ctx, uint32_t world, uint32_t rank, uint64_t bytes_per_rank, dsm_xchg** out,
                    void* ipc_handle_out /* DSM_IPC_HANDLE_BYTES, may be NULL */);
int dsm_xchg_local_base(dsm_xchg* x, void** base);
int dsm_xchg_connect_ipc(dsm_xchg* x, const void* handles /* [world][DSM_IPC_HANDLE_BYTES] */);
int dsm_xchg_connect_ptrs(dsm_xchg* x, void* const* bases /* [world], own entry ignored */);
int dsm_xchg_all_gather_dev(dsm_xchg* x, const void* d_src, uint64_t bytes, const void** d_gathered,
                            uint64_t* slot_stride_bytes);
int dsm_xchg_status(dsm_xchg* x, uint32_t* waited_for_rank_plus_1);
int dsm_xchg_destroy(dsm_xchg* x);

/* ---- utilities ------------------------------------------------------------------------ */

/* Size the context's scratch for batches of up to n_jobs jobs with this layout against this
 * table (k = 0: bitmap calls, else top-k calls up to k), so that later calls of that shape find
 * their buffers in place: the entry points otherwise grow scratch on first use with cudaFree +
 * cudaMalloc, a device-wide synchronisation a scheduler does not want on its hot path.  The
 * id array of raw-job calls is sized on use. */
int dsm_reserve(dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout, uint32_t n_jobs, uint32_t k);

int dsm_sync(dsm_ctx* ctx);
/* Pinned host memory for record / result buffers of the host-buffer entry points. */
int dsm_host_alloc(size_t bytes, void** out);
int dsm_host_free(void* p);
/* Kernel launches issued by this context since creation (bench.py's gpu_launches). */
uint64_t dsm_launch_count(const dsm_ctx* ctx);
/* Time the kernels of a call: while enabled, CUDA events bracket every launch of a match call
 * on the context's stream; dsm_timing_read synchronises, returns the summed kernel time and
 * the number of launches since the last read, and resets the counter. */
int dsm_timing_enable(dsm_ctx* ctx, int on);
int dsm_timing_read(dsm_ctx* ctx, double* total_ms, uint32_t* n_launches);
/* The same events, per kernel of a call: parts_ms = summed time of k_build_tables, k_match and
 * k_extract (0 for bitmap calls) over the n_calls calls since the last read (bench.py's
 * step_breakdown_ms).  Resets the counter like dsm_timing_read; use one or the other. */
int dsm_timing_read_parts(dsm_ctx* ctx, double parts_ms[3], uint32_t* n_calls);
/* Launch geometry chosen for a call (k = 0: bitmap mode): grid CTAs, work items, row ranges per
 * job group in the two segments, slots, shared memory, dictionary entries, job groups. */
int dsm_plan_query(const dsm_ctx* ctx, const dsm_table* t, const dsm_layout* layout,
                   uint32_t n_jobs, uint32_t k, uint32_t out[8]);

#ifdef __cplusplus
}
#endif
#endif /* DSMATCH_H */
