// Sweep 1 of 3 (classify + decide).  Cell typing + numeric parse, column decisions, and the fixed-width formatter / emitter
// (SURVEY.md §7.1b K3, K4, K5): everything `pd.read_csv` does after tokenising plus
// `dropna(how='all').fillna('')` and `to_string(index=False)` (app/file_parser.py:301-305).
//
// Work item = a block of `rb` consecutive data rows of one document that never crosses a
// low_memory chunk boundary; a CTA takes items grid-stride, one thread per cell; a warp works on
// consecutive rows of ONE column, so its lanes run the same parser on tokens of similar length
// (mixing columns in a warp left 6 of 32 lanes active: profiles/r01_emit_v1_*); the item's bytes
// are shared by its warps through L1.  Column statistics are reduced in
// shared memory per item and merged into global memory with one atomic per column per item.
// Two sweeps over the cells of a CSV document: typing + measuring in one (csv_classify_fused_kernel)
// -> decide, rows dropped, layout, row offsets -> emit; sheets and the documents the one-sweep
// candidates do not cover measure in a separate sweep under the decided type (csv_measure_kernel).
// Tokens are re-parsed by emit instead of being materialised.
#include <atomic>
#include "table_common.cuh"

namespace rfq {

// ---- sweep 1: classify --------------------------------------------------------------------------
__global__ void __launch_bounds__(256) csv_classify_kernel(TableArgs a) {
  __shared__ uint32_t s_flags[RFQ_MAX_W], s_bad[RFQ_MAX_W], s_ovf[RFQ_MAX_W], s_uinv[RFQ_MAX_W], s_uovf[RFQ_MAX_W];
  __shared__ uint32_t s_doc;
  ChunkStat* stat = (ChunkStat*)a.stat;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    ItemCtx ic = item_ctx(a, item, &s_doc);
    const TDoc D = a.docs[ic.d];
    uint32_t W = D.W;
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) {
      s_flags[c] = 0; s_bad[c] = RFQ_NO_ROW; s_ovf[c] = RFQ_NO_ROW; s_uinv[c] = RFQ_NO_ROW; s_uovf[c] = RFQ_NO_ROW;
    }
    __syncthreads();
    // a thread owns one row and walks its columns (a warp = 32 rows of the same column)
    uint32_t rbp = 32;
    while (rbp < ic.nrows) rbp <<= 1;
    const uint32_t ngrp = blockDim.x / rbp, rr = threadIdx.x % rbp, grp = threadIdx.x / rbp;
    if (rr < ic.nrows && grp < ngrp) {
      const uint32_t row = ic.row0 + rr, slot = D.r0 + 1 + row;
      const RowCur rc = row_open(a, slot);
      uint32_t na = 0;
      for (uint32_t c = grp; c < W; c += ngrp) {
        CellSrc src = row_cell(a, D, rc, c);
        int is;
        uint32_t f = classify_cell(src, &is);
        if ((s_flags[c] & f) != f) atomicOr(&s_flags[c], f);
        if (is != INT_OK) {
          if (row < s_bad[c]) atomicMin(&s_bad[c], row);
          if (is == INT_OVERFLOW && row < s_ovf[c]) atomicMin(&s_ovf[c], row);
          if ((f & SF_UINT_INVALID) && row < s_uinv[c]) atomicMin(&s_uinv[c], row);
          if ((f & SF_UINT_OVF) && row < s_uovf[c]) atomicMin(&s_uovf[c], row);
        }
        if ((f & SF_ANY_NA) && c >= D.lead) ++na;
      }
      if (ngrp == 1) a.row_na[slot] = na;
      else if (na) atomicAdd(a.row_na + slot, na);
    }
    __syncthreads();
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) {
      ChunkStat* S = stat + D.stat_base + ic.chunk * W + c;
      if (s_flags[c]) atomicOr(&S->flags, s_flags[c]);
      if (s_bad[c] != RFQ_NO_ROW) atomicMin(&S->int_bad_min, s_bad[c]);
      if (s_ovf[c] != RFQ_NO_ROW) atomicMin(&S->int_ovf_min, s_ovf[c]);
      if (s_uinv[c] != RFQ_NO_ROW) atomicMin(&S->uint_inv_min, s_uinv[c]);
      if (s_uovf[c] != RFQ_NO_ROW) atomicMin(&S->uint_ovf_min, s_uovf[c]);
    }
    __syncthreads();
  }
}

// ---- sweep 1 + 2 fused: flags AND measurement candidates in one visit of every cell -----------
// Shared memory: 16 reduction arrays of max_w words (dynamic), so narrow tables keep occupancy.
enum { FA_FLAGS = 0, FA_BAD, FA_OVF, FA_UINV, FA_UOVF, FA_STR, FA_INT, FA_UINT, FA_BOOL, FA_IP, FA_TZ,
       FA_INF, FA_OBJ, FA_SCI, FA_FL, FA_NA, FA_COUNT };
#define RFQ_SMAX(arr, v) do { uint32_t v_ = (v); if (v_ && v_ > (arr)[c]) atomicMax(&(arr)[c], v_); } while (0)

__global__ void __launch_bounds__(256) csv_classify_fused_kernel(TableArgs a) {
  extern __shared__ uint32_t s_arr[];
  __shared__ uint32_t s_doc;
  const uint32_t MW = a.max_w;
  uint32_t* s_flags = s_arr + FA_FLAGS * MW; uint32_t* s_bad = s_arr + FA_BAD * MW;
  uint32_t* s_ovf = s_arr + FA_OVF * MW; uint32_t* s_uinv = s_arr + FA_UINV * MW;
  uint32_t* s_uovf = s_arr + FA_UOVF * MW; uint32_t* s_str = s_arr + FA_STR * MW;
  uint32_t* s_int = s_arr + FA_INT * MW; uint32_t* s_uint = s_arr + FA_UINT * MW;
  uint32_t* s_bool = s_arr + FA_BOOL * MW; uint32_t* s_ip = s_arr + FA_IP * MW;
  uint32_t* s_tz = s_arr + FA_TZ * MW; uint32_t* s_inf = s_arr + FA_INF * MW;
  uint32_t* s_obj = s_arr + FA_OBJ * MW; uint32_t* s_sci = s_arr + FA_SCI * MW;
  uint32_t* s_fl = s_arr + FA_FL * MW; uint32_t* s_na = s_arr + FA_NA * MW;
  ChunkStat* stat = (ChunkStat*)a.stat;
  ChunkCand* cand = (ChunkCand*)a.cand;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    ItemCtx ic = item_ctx(a, item, &s_doc);
    const TDoc D = a.docs[ic.d];
    uint32_t W = D.W;
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) {
      s_flags[c] = 0; s_bad[c] = RFQ_NO_ROW; s_ovf[c] = RFQ_NO_ROW; s_uinv[c] = RFQ_NO_ROW; s_uovf[c] = RFQ_NO_ROW;
      s_str[c] = 0; s_int[c] = 0; s_uint[c] = 0; s_bool[c] = 0; s_ip[c] = 0; s_tz[c] = 6; s_inf[c] = 0;
      s_obj[c] = 0; s_sci[c] = 0; s_fl[c] = 0; s_na[c] = 0;
    }
    __syncthreads();
    uint32_t rbp = 32;
    while (rbp < ic.nrows) rbp <<= 1;
    const uint32_t ngrp = blockDim.x / rbp, rr = threadIdx.x % rbp, grp = threadIdx.x / rbp;
    if (rr < ic.nrows && grp < ngrp) {
      const uint32_t row = ic.row0 + rr, slot = D.r0 + 1 + row;
      const RowCur rc = row_open(a, slot);
      uint32_t na = 0, extra = 0;
      for (uint32_t c = grp; c < W; c += ngrp) {
        CellCand m = classify_measure_cell(row_cell_tag_write(a, D, rc, c));
        const uint32_t f = m.flags;
        if ((s_flags[c] & f) != f) atomicOr(&s_flags[c], f);
        if (m.int_status != INT_OK) {
          if (row < s_bad[c]) atomicMin(&s_bad[c], row);
          if (m.int_status == INT_OVERFLOW && row < s_ovf[c]) atomicMin(&s_ovf[c], row);
          if ((f & SF_UINT_INVALID) && row < s_uinv[c]) atomicMin(&s_uinv[c], row);
          if ((f & SF_UINT_OVF) && row < s_uovf[c]) atomicMin(&s_uovf[c], row);
        }
        // the lanes of a warp sit on the same column: one shared-memory add per warp for its NA cells
        const unsigned am = __activemask();
        const unsigned nm = __ballot_sync(am, m.na);
        if (m.na) {
          if ((threadIdx.x & 31u) == (uint32_t)(__ffs(nm) - 1)) atomicAdd(&s_na[c], (uint32_t)__popc(nm));
          if (c >= D.lead) ++na;
          continue;
        }
        RFQ_SMAX(s_str, m.str_len); RFQ_SMAX(s_int, m.int_len); RFQ_SMAX(s_uint, m.uint_len);
        RFQ_SMAX(s_bool, m.bool_len);
        if (m.has_float) {
          RFQ_SMAX(s_ip, m.ip_len); RFQ_SMAX(s_inf, m.inf_len); RFQ_SMAX(s_obj, m.obj_len);
          RFQ_SMAX(s_sci, m.sci_len);
          if (m.tz < s_tz[c]) atomicMin(&s_tz[c], m.tz);
          if ((s_fl[c] & m.fl) != m.fl) atomicOr(&s_fl[c], m.fl);
        }
        if (m.need_exact) cand[D.stat_base + ic.chunk * W + c].need_exact = 1;   // rare: straight to global
        if (c >= D.lead) extra += m.extra;
      }
      if (ngrp == 1) { a.row_na[slot] = na; a.row_extra[slot] = extra; }
      else { if (na) atomicAdd(a.row_na + slot, na); if (extra) atomicAdd(a.row_extra + slot, extra); }
    }
    __syncthreads();
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) {
      const uint32_t e = D.stat_base + ic.chunk * W + c;
      ChunkStat* S = stat + e;
      if (s_flags[c]) atomicOr(&S->flags, s_flags[c]);
      if (s_bad[c] != RFQ_NO_ROW) atomicMin(&S->int_bad_min, s_bad[c]);
      if (s_ovf[c] != RFQ_NO_ROW) atomicMin(&S->int_ovf_min, s_ovf[c]);
      if (s_uinv[c] != RFQ_NO_ROW) atomicMin(&S->uint_inv_min, s_uinv[c]);
      if (s_uovf[c] != RFQ_NO_ROW) atomicMin(&S->uint_ovf_min, s_uovf[c]);
      ChunkCand* C = cand + e;
      if (s_str[c]) atomicMax(&C->str_len, s_str[c]);
      if (s_int[c]) atomicMax(&C->int_len, s_int[c]);
      if (s_uint[c]) atomicMax(&C->uint_len, s_uint[c]);
      if (s_bool[c]) atomicMax(&C->bool_len, s_bool[c]);
      if (s_ip[c]) atomicMax(&C->ip_len, s_ip[c]);
      if (s_tz[c] < 6) atomicMin(&C->min_tz, s_tz[c]);
      if (s_inf[c]) atomicMax(&C->inf_len, s_inf[c]);
      if (s_obj[c]) atomicMax(&C->obj_len, s_obj[c]);
      if (s_sci[c]) atomicMax(&C->sci_len, s_sci[c]);
      if (s_fl[c]) atomicOr(&C->fl, s_fl[c]);
      if (s_na[c]) atomicAdd(&C->na_count, s_na[c]);
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) csv_cand_init_kernel(ChunkCand* cand, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ChunkCand c;
  c.str_len = 0; c.int_len = 0; c.uint_len = 0; c.bool_len = 0; c.ip_len = 0; c.min_tz = 6; c.inf_len = 0;
  c.obj_len = 0; c.sci_len = 0; c.fl = 0; c.na_count = 0; c.need_exact = 0;
  cand[i] = c;
}

__global__ void __launch_bounds__(256) csv_stat_init_kernel(ChunkStat* stat, ChunkMeasure* meas,
                                                            uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  stat[i] = ChunkStat{0u, RFQ_NO_ROW, RFQ_NO_ROW, RFQ_NO_ROW, RFQ_NO_ROW, {0u, 0u, 0u}};
  ChunkMeasure m;
  m.plain_len = 0; m.ip_len = 0; m.min_tz = 6; m.inf_len = 0; m.obj_len = 0; m.sci_len = 0;
  m.fl = 0; m.nan_left = 0;
  meas[i] = m;
}

void table_stat_init(const TableArgs& a, cudaStream_t s) {
  if (a.n_stat)
    RFQ_LAUNCH(csv_stat_init_kernel, (a.n_stat + 255) / 256, 256, 0, s, (ChunkStat*)a.stat,
               (ChunkMeasure*)a.meas, a.n_stat);
}

void csv_table_classify(const TableArgs& a, cudaStream_t s) {
  prof_mark("csv_stat_init_kernel", s);
  if (a.n_stat)
    RFQ_LAUNCH(csv_stat_init_kernel, (a.n_stat + 255) / 256, 256, 0, s, (ChunkStat*)a.stat,
               (ChunkMeasure*)a.meas, a.n_stat);
  if (a.cand) {   // fused sweep: flags + measurement candidates
    if (a.n_stat) RFQ_LAUNCH(csv_cand_init_kernel, (a.n_stat + 255) / 256, 256, 0, s, (ChunkCand*)a.cand, a.n_stat);
    const uint32_t smem = (uint32_t)FA_COUNT * a.max_w * 4u;
    static std::atomic<unsigned long long> attr_done{0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!(attr_done.load() >> (dev & 63) & 1ull)) {   // per-device opt-in above 48 KB (tables wider than 768 columns)
      cudaFuncSetAttribute(csv_classify_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           FA_COUNT * RFQ_MAX_W * 4);
      attr_done.fetch_or(1ull << (dev & 63));
    }
    prof_mark("csv_classify_fused_kernel", s);
    if (a.n_items) RFQ_LAUNCH(csv_classify_fused_kernel, table_grid(a.n_items), 256, smem, s, a);
    return;
  }
  prof_mark("csv_classify_kernel", s);
  if (a.n_items) RFQ_LAUNCH(csv_classify_kernel, table_grid(a.n_items), 256, 0, s, a);
}


}  // namespace rfq
