// Typing sweep of the row pipeline (SURVEY.md §7.1b K3 + K4): what `pd.read_csv` does after
// tokenising — NA set, int64 / uint64 / float64 / bool / str inference per low_memory chunk — fused with
// the measurements `to_string` needs (printed width of every cell under every type its column can still
// become), app/file_parser.py:301-305.
//
// Work item = RFQ2_ROWS consecutive data rows of one document (never across a low_memory chunk
// boundary).  The item's bytes are one contiguous span of the arena: one thread brings it into shared
// memory with a bulk asynchronous copy (cp.async.bulk + mbarrier), then every thread walks ONE row out
// of the stage (csv2_rows.cuh), so the lanes of a warp sit on the same column at the same time and
// their unaligned 8-byte windows are shared-memory reads, not 32 different L1 lines.  Cells the hot
// classifier does not decide (exact float parsing, integers beyond 2^53, text that needs the quote
// state machine) are appended to a list and finished by csv2_type_cold_kernel with the general
// functions of cell_ops.cuh; keeping those out of this kernel keeps its register count small.
// HBM-bound byte work: no tensor cores, by design.
#include <algorithm>
#include <atomic>
#include "csv2.h"
#include "csv2_rows.cuh"
#include "csv2_stage.cuh"
#include "csv_dfa.cuh"
#include "launch.h"
#include "scan.cuh"
#include "table_common.cuh"

namespace rfq {

// What the scan left about a document, completed: data rows used and the failures that show without
// looking at a cell.  Every kernel that needs them before csv2_plan_kernel stores them derives them itself.
__device__ __forceinline__ DocRun doc_view(const Doc2& D, DocRun R) {
  if (R.n_rec == 0) R.flags |= RUN_FAIL;                                                    // EmptyDataError
  if (R.final_state == ST_Q && (unsigned long long)R.n_rec < (unsigned long long)D.max_rows + 1ull) R.flags |= RUN_FAIL;   // EOF inside string
  R.n_data = R.n_rec ? min(R.n_rec - 1u, D.max_rows) : 0u;
  return R;
}

enum { S_FLAGS = 0, S_BAD, S_UINV, S_STR, S_INT, S_UINT, S_BOOL, S_IP, S_TZ, S_OBJ, S_SCI, S_FL, S_NA, S_COUNT };
#define RFQ2_TYPE_STAGE (24u << 10)


// What classify_hot (csv2_rows.cuh) finds goes straight into the item's shared-memory statistics:
// each outcome touches only the entries it can change (a word: flags, text width, first-bad-row marks).
struct TypeSink {
  uint32_t* s_arr; uint32_t W, c, inv_row, extra;
  bool is_cold, is_na;
  __device__ __forceinline__ uint32_t& at(int k) { return s_arr[k * W + c]; }
  __device__ __forceinline__ void smax(int k, uint32_t v) { if (v > at(k)) atomicMax(&at(k), v); }
  __device__ __forceinline__ void flags(uint32_t f) { if ((at(S_FLAGS) & f) != f) atomicOr(&at(S_FLAGS), f); }
  __device__ __forceinline__ void cold() { is_cold = true; }
  __device__ __forceinline__ void na() { is_na = true; }
  __device__ __forceinline__ void word(uint32_t f, uint32_t str_len, uint32_t ex, uint32_t bool_len) {
    flags(f);
    smax(S_BAD, inv_row);
    if (f & SF_UINT_INVALID) smax(S_UINV, inv_row);
    smax(S_STR, str_len);
    if (bool_len) smax(S_BOOL, bool_len);
    extra = ex;
  }
  __device__ __forceinline__ void integer(uint32_t f, uint32_t str_len, uint32_t int_len, bool neg, bool large) {
    flags(f);
    smax(S_STR, str_len); smax(S_INT, int_len);
    if (!neg) smax(S_UINT, int_len);
    smax(S_IP, int_len); smax(S_OBJ, int_len - (neg ? 1u : 0u) + 3u); smax(S_SCI, 12u + (neg ? 1u : 0u));
    const uint32_t fl = MF_FINITE | (large ? MF_LARGE : 0);
    if ((at(S_FL) & fl) != fl) atomicOr(&at(S_FL), fl);
  }
  __device__ __forceinline__ void decimal(uint32_t f, uint32_t str_len, uint32_t ip_len, uint32_t tz, uint32_t obj_len,
                                          bool neg, bool large) {
    flags(f);
    smax(S_BAD, inv_row);
    if (f & SF_UINT_INVALID) smax(S_UINV, inv_row);
    smax(S_STR, str_len);
    smax(S_IP, ip_len); smax(S_TZ, 6u - tz); smax(S_OBJ, obj_len); smax(S_SCI, 12u + (neg ? 1u : 0u));
    const uint32_t fl = MF_FINITE | (large ? MF_LARGE : 0);
    if ((at(S_FL) & fl) != fl) atomicOr(&at(S_FL), fl);
  }
};

__global__ void __launch_bounds__(RFQ2_ROWS) csv2_type_kernel(Csv2Args a, uint32_t max_w) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* s_stage = smem;
  uint32_t* s_arr = reinterpret_cast<uint32_t*>(smem + RFQ2_TYPE_STAGE + 32);
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_doc, s_red[2];
  if (threadIdx.x == 0) mbar_init(&s_bar, 1);
  mbar_init_fence();
  __syncthreads();
  uint32_t phase = 0;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    if (threadIdx.x == 0) { s_doc = ub_search(a.item_base, a.n_docs, item); s_red[0] = 0; s_red[1] = 0; }
    __syncthreads();
    const uint32_t d = s_doc;
    const Doc2 D = a.docs[d];
    const DocRun R = doc_view(D, a.run[d]);
    const uint32_t row0 = (item - D.item_base) * RFQ2_ROWS;
    if ((R.flags & (RUN_FAIL | RUN_UNSUP)) || row0 >= R.n_data) { __syncthreads(); continue; }
    const uint32_t nrows = min(RFQ2_ROWS, R.n_data - row0);
    const uint32_t chunk = row0 / D.bl;
    const uint32_t W = D.W;
    const uint32_t slot0 = D.slot_base + 1u + row0;
    // ---- stage the item's bytes
    const uint32_t start = __ldg(a.rec_start + slot0);
    const uint32_t e_idx = 1u + row0 + nrows;
    const uint32_t end = e_idx < R.n_rec ? __ldg(a.rec_start + D.slot_base + e_idx) : D.len;
    const uint32_t g_lo = (D.off + start) & ~15u;
    const uint32_t g_hi = (D.off + end + 31u) & ~15u;       // 16 bytes of slack for the window loads
    const bool staged = g_hi - g_lo <= RFQ2_TYPE_STAGE;
    if (threadIdx.x == 0) {
      if (staged) bulk_load(s_stage, a.arena + g_lo, g_hi - g_lo, &s_bar);
      if (g_hi - g_lo > a.counters[3]) atomicMax(a.counters + 3, g_hi - g_lo);   // sizes the next run's emit stage
    }
    for (uint32_t c = threadIdx.x; c < W * S_COUNT; c += blockDim.x) s_arr[c] = 0;
    const uint8_t* base = staged ? s_stage - (size_t)(g_lo - D.off) : a.arena + D.off;   // base[pos] = document byte pos
    if (staged) mbar_wait(&s_bar, phase), phase ^= 1u;
    __syncthreads();
    uint32_t allna = 0, extra = 0;
    if (threadIdx.x < nrows) {
      const uint32_t slot = slot0 + threadIdx.x, row = row0 + threadIdx.x;
      const uint32_t rs = threadIdx.x == 0 ? start : __ldg(a.rec_start + slot);
      const uint32_t limit = (2u + row < R.n_rec) ? __ldg(a.rec_start + slot + 1) : D.len;
      RowWalk w(base, rs, limit, D.sep);
      uint32_t na = 0;
      const uint32_t inv_row = ~row;
      uint4* mrow = reinterpret_cast<uint4*>(a.cell_meta) + D.meta_base + (size_t)row * RFQ2_META_VECS(W);
      unsigned long long mlo = 0ull, mhi = 0ull;
      TypeSink k;
      k.s_arr = s_arr; k.W = W; k.inv_row = inv_row;
      for (uint32_t c = 0; c < W; ++c) {
        const CellSrc src = w.next();
        k.c = c; k.is_cold = false; k.is_na = false; k.extra = 0;
        classify_hot(src, k);
        {   // 16 bits per cell for the emit sweep, eight cells per 16-byte store
          const bool quoted = !src.missing && src.p != base + w.field_pos;
          const uint32_t me = cell_meta(src, quoted, k.is_cold, k.is_na);
          const uint32_t q = c & 7u;
          if (q < 4u) mlo |= (unsigned long long)me << (16u * q);
          else mhi |= (unsigned long long)me << (16u * (q - 4u));
          if (q == 7u || c + 1 == W) {
            mrow[c >> 3] = make_uint4((uint32_t)mlo, (uint32_t)(mlo >> 32), (uint32_t)mhi, (uint32_t)(mhi >> 32));
            mlo = 0ull; mhi = 0ull;
          }
        }
        if (k.is_cold) {
          const uint32_t i = atomicAdd(a.counters + 0, 1u);
          if (i < a.cold_cells_cap) a.cold_cells[i] = ColdCell{d, slot, c};
          else atomicOr(&a.run[d].flags, (uint32_t)RUN_GENERAL);
          continue;
        }
        // the lanes of a warp sit on the same column: one shared-memory add per warp for its NA cells
        const unsigned am = __activemask();
        const unsigned nm = __ballot_sync(am, k.is_na);
        if (k.is_na) {
          if ((threadIdx.x & 31u) == (uint32_t)(__ffs(nm) - 1)) atomicAdd(&s_arr[S_NA * W + c], (uint32_t)__popc(nm));
          if (!(s_arr[S_FLAGS * W + c] & SF_ANY_NA)) atomicOr(&s_arr[S_FLAGS * W + c], (uint32_t)SF_ANY_NA);
          if (c >= D.lead) ++na;
          continue;
        }
        if (c >= D.lead) extra += k.extra;
      }
      if (!w.eol) atomicOr(&a.run[d].flags, (uint32_t)RUN_FAIL);   // "Expected N fields in line K, saw M"
      a.row_na[slot] = na;
      a.row_extra[slot] = extra;
      allna = na == D.h ? 1u : 0u;
    }
    // item totals: rows whose output cells are all NA, extra bytes
    {
      const unsigned bm = __ballot_sync(0xffffffffu, allna != 0);
      uint32_t ex = extra;
#pragma unroll
      for (int o = 16; o; o >>= 1) ex += __shfl_xor_sync(0xffffffffu, ex, o);
      if ((threadIdx.x & 31u) == 0) { if (bm) atomicAdd(&s_red[0], (uint32_t)__popc(bm)); if (ex) atomicAdd(&s_red[1], ex); }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      a.item_allna[item] = s_red[0];
      a.item_extra[item] = s_red[1];
      if (s_red[0]) atomicAdd(a.chunk_allna + D.stat_base + chunk * W, s_red[0]);
    }
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) {
      StatV2* S = a.stat + D.stat_base + chunk * W + c;
#define RFQ2_SV(k) s_arr[(k) * W + c]
      if (RFQ2_SV(S_FLAGS)) atomicOr(&S->flags, RFQ2_SV(S_FLAGS));
      if (RFQ2_SV(S_BAD)) atomicMax(&S->inv_bad, RFQ2_SV(S_BAD));
      if (RFQ2_SV(S_UINV)) atomicMax(&S->inv_uinv, RFQ2_SV(S_UINV));
      if (RFQ2_SV(S_STR)) atomicMax(&S->str_len, RFQ2_SV(S_STR));
      if (RFQ2_SV(S_INT)) atomicMax(&S->int_len, RFQ2_SV(S_INT));
      if (RFQ2_SV(S_UINT)) atomicMax(&S->uint_len, RFQ2_SV(S_UINT));
      if (RFQ2_SV(S_BOOL)) atomicMax(&S->bool_len, RFQ2_SV(S_BOOL));
      if (RFQ2_SV(S_IP)) atomicMax(&S->ip_len, RFQ2_SV(S_IP));
      if (RFQ2_SV(S_TZ)) atomicMax(&S->inv_tz, RFQ2_SV(S_TZ));
      if (RFQ2_SV(S_OBJ)) atomicMax(&S->obj_len, RFQ2_SV(S_OBJ));
      if (RFQ2_SV(S_SCI)) atomicMax(&S->sci_len, RFQ2_SV(S_SCI));
      if (RFQ2_SV(S_FL)) atomicOr(&S->fl, RFQ2_SV(S_FL));
      if (RFQ2_SV(S_NA)) atomicAdd(&S->na_count, RFQ2_SV(S_NA));
#undef RFQ2_SV
    }
    __syncthreads();
  }
}

// ---- fix-up: the cells the hot sweep left, one thread each, general functions --------------------
__global__ void __launch_bounds__(128) csv2_type_cold_kernel(Csv2Args a) {
  const uint32_t n = min(a.counters[0], a.cold_cells_cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const ColdCell cc = a.cold_cells[i];
    const Doc2& D = a.docs[cc.doc];
    const DocRun& R = a.run[cc.doc];
    const uint32_t k_idx = cc.slot - D.slot_base;          // record index (1 = first data row)
    const uint32_t row = k_idx - 1u;
    const uint32_t rs = a.rec_start[cc.slot];
    const uint32_t limit = (k_idx + 1u < R.n_rec) ? a.rec_start[cc.slot + 1] : D.len;
    RowWalk w(a.arena + D.off, rs, limit, D.sep);
    CellSrc src = w.next();
    for (uint32_t c = 0; c < cc.col; ++c) src = w.next();
    const CellCand m = classify_measure_cell(src);
    const uint32_t chunk = row / D.bl;
    StatV2* S = a.stat + D.stat_base + chunk * D.W + cc.col;
    if (m.flags) atomicOr(&S->flags, m.flags);
    const uint32_t item = D.item_base + row / RFQ2_ROWS;
    if (m.na) {
      atomicAdd(&S->na_count, 1u);
      if (cc.col >= D.lead && atomicAdd(a.row_na + cc.slot, 1u) + 1u == D.h) {   // the row turned out all NA
        atomicAdd(a.item_allna + item, 1u);
        atomicAdd(a.chunk_allna + D.stat_base + chunk * D.W, 1u);
      }
      continue;
    }
    const uint32_t inv_row = ~row;
    if (m.int_status != INT_OK) {
      atomicMax(&S->inv_bad, inv_row);
      if (m.int_status == INT_OVERFLOW) atomicMax(&S->inv_ovf, inv_row);
      if (m.flags & SF_UINT_INVALID) atomicMax(&S->inv_uinv, inv_row);
      if (m.flags & SF_UINT_OVF) atomicMax(&S->inv_uovf, inv_row);
    }
    if (m.str_len) atomicMax(&S->str_len, m.str_len);
    if (m.int_len) atomicMax(&S->int_len, m.int_len);
    if (m.uint_len) atomicMax(&S->uint_len, m.uint_len);
    if (m.bool_len) atomicMax(&S->bool_len, m.bool_len);
    if (m.has_float) {
      if (m.ip_len) atomicMax(&S->ip_len, m.ip_len);
      if (m.tz < 6) atomicMax(&S->inv_tz, 6u - m.tz);
      if (m.inf_len) atomicMax(&S->inf_len, m.inf_len);
      if (m.obj_len) atomicMax(&S->obj_len, m.obj_len);
      if (m.sci_len) atomicMax(&S->sci_len, m.sci_len);
      if (m.fl) atomicOr(&S->fl, m.fl);
    }
    if (m.need_exact) S->need_exact = 1u;
    if (cc.col >= D.lead && m.extra) {
      atomicAdd(a.row_extra + cc.slot, m.extra);
      atomicAdd(a.item_extra + item, m.extra);
    }
  }
}

// ---- per document: chunk types, rows dropped, measures, column layout ---------------------------
__device__ __forceinline__ ChunkStat stat_of(const StatV2& s) {
  ChunkStat c;
  c.flags = s.flags;
  c.int_bad_min = s.inv_bad ? ~s.inv_bad : RFQ_NO_ROW;
  c.int_ovf_min = s.inv_ovf ? ~s.inv_ovf : RFQ_NO_ROW;
  c.uint_inv_min = s.inv_uinv ? ~s.inv_uinv : RFQ_NO_ROW;
  c.uint_ovf_min = s.inv_uovf ? ~s.inv_uovf : RFQ_NO_ROW;
  c.pad[0] = c.pad[1] = c.pad[2] = 0;
  return c;
}
__device__ __forceinline__ ChunkCand cand_of(const StatV2& s) {
  ChunkCand c;
  c.str_len = s.str_len; c.int_len = s.int_len; c.uint_len = s.uint_len; c.bool_len = s.bool_len;
  c.ip_len = s.ip_len; c.min_tz = 6u - s.inv_tz; c.inf_len = s.inf_len; c.obj_len = s.obj_len;
  c.sci_len = s.sci_len; c.fl = s.fl; c.na_count = s.na_count; c.need_exact = s.need_exact;
  return c;
}

__global__ void __launch_bounds__(256) csv2_plan_kernel(Csv2Args a) {
  const uint32_t d = blockIdx.x;
  const Doc2 D = a.docs[d];
  DocRun R = doc_view(D, a.run[d]);
  __shared__ uint32_t s_flags, s_dropped, s_carry;
  if (threadIdx.x == 0) { s_flags = 0; s_dropped = 0; s_carry = 0; }
  __syncthreads();
  const bool dead = (R.flags & (RUN_FAIL | RUN_UNSUP)) != 0;
  const uint32_t W = D.W;
  const uint32_t n_chunks = dead ? 0u : (R.n_data + D.bl - 1) / D.bl;
  const uint32_t n_ent = n_chunks * W;
  ChunkMeasure* meas = (ChunkMeasure*)a.meas;
  // 1. chunk types
  for (uint32_t e = threadIdx.x; e < n_ent; e += blockDim.x) {
    const uint32_t k = e / W, c = e - k * W;
    const StatV2 S = a.stat[D.stat_base + e];
    uint32_t fl = 0;
    if (S.flags & SF_BAD_UTF8) fl |= RUN_FAIL;               // UnicodeDecodeError in a cell
    const int t = decide_chunk_type(stat_of(S));
    a.ctype[D.stat_base + e] = (uint8_t)t;
    if (c >= D.lead) {
      if (t == CT_STRNF) a.chunk_nf[D.stat_base + k * W] = 1;
      const bool exact = !(t == CT_STR || t == CT_BOOL || t == CT_BOOLNA || t == CT_UINT || t == CT_INT ||
                           t == CT_INTNA || t == CT_FLOAT) || (t == CT_FLOAT && S.need_exact);
      if (exact) fl |= RUN_GENERAL;                           // measured by the three-sweep path instead
    }
    if (fl) atomicOr(&s_flags, fl);
  }
  __syncthreads();
  // 2. measures under the decided types, rows dropna removes
  for (uint32_t e = threadIdx.x; e < n_ent; e += blockDim.x) {
    const uint32_t k = e / W, c = e - k * W;
    const uint32_t e0 = D.stat_base + k * W;
    const uint32_t dropped = a.chunk_nf[e0] ? 0u : a.chunk_allna[e0];
    if (c == 0 && dropped) atomicAdd(&s_dropped, dropped);
    ChunkMeasure M;
    finish_measure(cand_of(a.stat[D.stat_base + e]), a.ctype[D.stat_base + e], dropped, &M);
    meas[D.stat_base + e] = M;
  }
  __syncthreads();
  const uint32_t n_keep = dead ? 0u : R.n_data - s_dropped;
  // 3. columns: class, width, decimals, offsets (block-wide running offset, 256 columns per round)
  bool unsup = false;
  for (uint32_t j0 = 0; j0 < D.h; j0 += blockDim.x) {
    const uint32_t j = j0 + threadIdx.x;
    uint32_t w1 = 0;
    ColOut2 o; o.cls = 0; o.dec = 0; o.sci = 0; o.width = 0; o.cp_off = 0;
    if (j < D.h && !dead) {
      const uint32_t e0 = D.stat_base + D.lead + j;
      bool nan_left = false;
      for (uint32_t k = 0; k < n_chunks; ++k) nan_left |= meas[e0 + (size_t)k * W].nan_left != 0;
      int cls = fold_column_class(a.ctype + e0, n_chunks, W, nan_left);
      if (cls == CC_UNSUP) { unsup = true; cls = CC_OBJ; }
      const ColumnLayout L = layout_column(cls, meas + e0, a.ctype + e0, n_chunks, W, a.name_cp[D.name_base + j]);
      o.cls = L.cls; o.dec = L.dec; o.sci = L.sci ? 1 : 0; o.width = L.width;
      w1 = L.width + 1;
    }
    uint32_t tot;
    const uint32_t ex = block_excl_scan<SumU32>(w1, &tot);
    if (j < D.h) { o.cp_off = s_carry + ex; a.col[D.col_base + j] = o; }
    __syncthreads();
    if (threadIdx.x == 0) s_carry += tot;
    __syncthreads();
  }
  if (unsup) atomicOr(&s_flags, (uint32_t)RUN_UNSUP);
  __syncthreads();
  if (threadIdx.x == 0) {
    R.flags |= s_flags;
    R.n_keep = n_keep;
    R.linew = s_carry ? s_carry - 1 : 0;
    R.hdr_bytes = (R.flags & (RUN_FAIL | RUN_UNSUP | RUN_GENERAL)) ? 0u
                  : D.banner_len + (n_keep ? R.linew + D.hdr_extra : D.empty_len);
    a.run[d] = R;
  }
}

// bytes of a work item's output range (the document's first item carries banner + header line): loads
// only, so that a caller with several items in flight overlaps their latencies
struct ItemBytes { unsigned long long b; uint32_t d, ex, big; };
__device__ __forceinline__ ItemBytes item_bytes_of(const Csv2Args& a, uint32_t item) {
  ItemBytes r{0ull, 0xFFFFFFFFu, 0u, 0u};
  if (item >= a.n_items) return r;
  r.d = ub_search(a.item_base, a.n_docs, item);
  const Doc2& D = a.docs[r.d];
  const DocRun& R = a.run[r.d];
  if (R.flags & (RUN_FAIL | RUN_UNSUP | RUN_GENERAL)) return r;
  const uint32_t row0 = (item - D.item_base) * RFQ2_ROWS;
  if (item == D.item_base) r.b += R.hdr_bytes;
  if (row0 < R.n_data && R.n_keep) {
    const uint32_t nrows = min(RFQ2_ROWS, R.n_data - row0);
    const uint32_t e0 = D.stat_base + (row0 / D.bl) * D.W;
    const uint32_t dropped = a.chunk_nf[e0] ? 0u : a.item_allna[item];
    r.ex = a.item_extra[item];
    const unsigned long long rb = (unsigned long long)(nrows - dropped) * (1ull + R.linew) + r.ex;
    r.b += rb;
    if (rb < 0xFFFFFFFFull) r.big = (uint32_t)rb;
  }
  return r;
}
// The two per-document side results (non-ASCII extra bytes, largest item), reduced over the warp before they
// touch global memory: neighbouring items belong to the same document nearly always.  All 32 lanes call it.
__device__ __forceinline__ void item_bytes_commit(const Csv2Args& a, const ItemBytes& r) {
  const uint32_t d0 = __shfl_sync(0xffffffffu, r.d, 0);
  if (__all_sync(0xffffffffu, r.d == d0 || (r.ex == 0 && r.big == 0))) {
    const uint32_t sum = __reduce_add_sync(0xffffffffu, r.ex), mx = __reduce_max_sync(0xffffffffu, r.big);
    if ((threadIdx.x & 31u) == 0 && d0 != 0xFFFFFFFFu) {
      if (sum) atomicAdd(&a.run[d0].extra, sum);
      if (mx > a.counters[4]) atomicMax(a.counters + 4, mx);
    }
  } else {
    if (r.ex) atomicAdd(&a.run[r.d].extra, r.ex);
    if (r.big > a.counters[4]) atomicMax(a.counters + 4, r.big);
  }
}

__global__ void __launch_bounds__(256) csv2_item_bytes_kernel(Csv2Args a) {
  const uint32_t item = blockIdx.x * blockDim.x + threadIdx.x;   // whole warps: the grid covers n_items + 1 rounded up
  const ItemBytes r = item_bytes_of(a, item);
  item_bytes_commit(a, r);
  if (item <= a.n_items) a.item_bytes[item] = r.b;
}

__global__ void __launch_bounds__(128) csv2_docout_kernel(Csv2Args a) {
  const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_docs) return;
  const unsigned long long o0 = a.item_off[a.item_base[d]], o1 = a.item_off[a.item_base[d + 1]];
  a.doc_out_off[d] = o0;
  a.doc_out_len[d] = o1 - o0;
  if (d == 0) *reinterpret_cast<unsigned long long*>(a.counters + 6) = a.item_off[a.n_items];   // total bytes of the attempt
}

static int sm_count() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}

void csv2_type(const Csv2Args& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  if (a.n_items == 0) return;
  const uint32_t smem = RFQ2_TYPE_STAGE + 32u + (uint32_t)S_COUNT * a.max_w * 4u;
  static std::atomic<unsigned long long> attr_done{0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_done.load() >> (dev & 63) & 1ull)) {   // per-device opt-in above 48 KB (wide tables)
    cudaFuncSetAttribute(csv2_type_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         RFQ2_TYPE_STAGE + 32 + S_COUNT * RFQ_MAX_W * 4);
    attr_done.fetch_or(1ull << (dev & 63));
  }
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, csv2_type_kernel, RFQ2_ROWS, smem);
  const uint32_t grid = std::min<uint32_t>(a.n_items, (uint32_t)(sm_count() * std::max(per_sm, 1)));
  prof_mark("csv2_type_kernel", s);
  RFQ_LAUNCH(csv2_type_kernel, grid, RFQ2_ROWS, smem, s, a, a.max_w);
  prof_mark("csv2_type_cold_kernel", s);
  RFQ_LAUNCH(csv2_type_cold_kernel, sm_count() * 4, 128, 0, s, a);
}

// item bytes -> offsets -> document offsets in ONE launch when the batch has few enough items for one CTA:
// five launches less on the critical path of a single big document.  The bytes of all items are
// computed first (independent loads, no barrier between rounds), then every thread scans its own
// consecutive items around ONE block scan.
__global__ void __launch_bounds__(1024) csv2_offsets_kernel(Csv2Args a) {
  __shared__ unsigned long long s_warp[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t per = (a.n_items + 1u + 1023u) / 1024u;
  const uint32_t lo = min(threadIdx.x * per, a.n_items + 1u), hi = min(lo + per, a.n_items + 1u);
  unsigned long long sum = 0;
  unsigned long long b8[8];
  const bool in_regs = per <= 8u;                 // up to 8191 items: a thread's items never leave its registers
  {
    // bytes of this thread's consecutive items, eight loads in flight; their side results are folded per
    // thread (consecutive items share a document nearly always) and committed once per warp
    ItemBytes acc{0ull, 0xFFFFFFFFu, 0u, 0u};
    for (uint32_t base = lo; base < hi; base += 8) {
      ItemBytes r[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) r[k] = item_bytes_of(a, base + k < hi ? base + k : 0xFFFFFFFFu);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (base + k >= hi) continue;
        sum += r[k].b;
        if (in_regs) b8[k] = r[k].b; else a.item_bytes[base + k] = r[k].b;
        if (r[k].d == 0xFFFFFFFFu || (r[k].ex == 0 && r[k].big == 0)) continue;
        if (acc.d == 0xFFFFFFFFu) acc.d = r[k].d;
        if (acc.d == r[k].d) { acc.ex += r[k].ex; acc.big = max(acc.big, r[k].big); }
        else {   // a second document inside one thread's items: straight to memory
          if (r[k].ex) atomicAdd(&a.run[r[k].d].extra, r[k].ex);
          if (r[k].big > a.counters[4]) atomicMax(a.counters + 4, r[k].big);
        }
      }
    }
    item_bytes_commit(a, acc);
  }
  const unsigned long long incl = warp_incl_scan<SumU64>(sum, lane);
  if (lane == 31) s_warp[w] = incl;
  __syncthreads();
  if (w == 0) s_warp[lane] = warp_incl_scan<SumU64>(s_warp[lane], lane);
  __syncthreads();
  unsigned long long run = (w ? s_warp[w - 1] : 0ull) + incl - sum;
  if (in_regs) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t i = lo + k;
      if (i >= hi) continue;
      a.item_off[i] = run;
      if (i == a.n_items) *reinterpret_cast<unsigned long long*>(a.counters + 6) = run;   // total bytes of the attempt
      run += b8[k];
    }
  } else {
    for (uint32_t i = lo; i < hi; ++i) {
      const unsigned long long b = a.item_bytes[i];
      a.item_off[i] = run;
      if (i == a.n_items) *reinterpret_cast<unsigned long long*>(a.counters + 6) = run;
      run += b;
    }
  }
  __syncthreads();
  for (uint32_t d = threadIdx.x; d < a.n_docs; d += blockDim.x) {
    const unsigned long long o0 = a.item_off[a.item_base[d]], o1 = a.item_off[a.item_base[d + 1]];
    a.doc_out_off[d] = o0;
    a.doc_out_len[d] = o1 - o0;
  }
}

void csv2_plan(const Csv2Args& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("csv2_plan_kernel", s);
  RFQ_LAUNCH(csv2_plan_kernel, a.n_docs, 256, 0, s, a);
  prof_mark("csv2_item_bytes+scan", s);
  if (a.n_items <= (1u << 16)) { RFQ_LAUNCH(csv2_offsets_kernel, 1, 1024, 0, s, a); return; }
  RFQ_LAUNCH(csv2_item_bytes_kernel, (a.n_items + 1 + 255) / 256, 256, 0, s, a);
  exclusive_scan<SumU64>(a.item_bytes, a.item_off, a.n_items + 1, a.scan_tmp64, nullptr, s);
  RFQ_LAUNCH(csv2_docout_kernel, (a.n_docs + 127) / 128, 128, 0, s, a);
}

}  // namespace rfq
