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
// Three sweeps over the cells: classify -> (decide) -> measure -> (layout, row offsets) -> emit.
// Tokens are re-parsed in each sweep instead of being materialised: the bytes stay in L2 and a
// typed copy would cost more HBM traffic than the re-parse costs issue slots.
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
  prof_mark("csv_classify_kernel", s);
  if (a.n_items) RFQ_LAUNCH(csv_classify_kernel, table_grid(a.n_items), 256, 0, s, a);
}


}  // namespace rfq
