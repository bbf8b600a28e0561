// Sweep 2 of 3 (measure + layout + row offsets).  Cell typing + numeric parse, column decisions, and the fixed-width formatter / emitter
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
#include "table_common.cuh"

namespace rfq {

// ---- decide chunk types: one thread per (document, chunk, column) entry ------------------------
// (chunk_nf is zeroed by the host before the run)
__global__ void __launch_bounds__(128) csv_decide_kernel(TableArgs a) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= a.n_stat) return;
  // entries are laid out document by document: largest d with docs[d].stat_base <= e
  uint32_t lo = 0, hi = a.n_docs;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (a.docs[mid].stat_base <= e) lo = mid; else hi = mid;
  }
  const TDoc D = a.docs[lo];
  if (!D.active || D.W == 0) return;
  uint32_t rel = e - D.stat_base;
  if (rel >= D.n_chunks * D.W) return;
  uint32_t k = rel / D.W, c = rel - k * D.W;
  ChunkStat S = ((const ChunkStat*)a.stat)[e];
  if (S.flags & SF_BAD_UTF8) atomicOr(a.doc_flags + lo, 1);
  int t = decide_chunk_type(S);
  a.ctype[e] = (uint8_t)t;
  if (t == CT_STRNF && c >= D.lead) a.chunk_nf[D.stat_base + k * D.W] = 1;
  if (a.cand && c >= D.lead) {
    // chunk types whose measurement the one-sweep candidates do not cover: the document goes
    // through the exact second sweep (doc_exact[n_docs] counts them: 0 = that sweep exits at once)
    bool exact = !(t == CT_STR || t == CT_BOOL || t == CT_BOOLNA || t == CT_UINT || t == CT_INT ||
                   t == CT_INTNA || t == CT_FLOAT) ||
                 (t == CT_FLOAT && ((const ChunkCand*)a.cand)[e].need_exact);
    if (exact && atomicExch(a.doc_exact + lo, 1u) == 0u) atomicAdd(a.doc_exact + a.n_docs, 1u);
  }
}

// ---- one-sweep mode: rows dropna removes, per chunk (the candidates cover all rows) ---------
__global__ void __launch_bounds__(256) csv_rowstat_kernel(TableArgs a) {
  uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  bool dropped = false;
  uint32_t key = 0;
  if (slot < a.n_slots) {
    uint32_t d = ub_search(a.slot_doc_r0, a.n_docs, slot);
    const TDoc D = a.docs[d];
    uint32_t L = slot - D.r0;
    if (D.active && L >= 1 && L <= D.n_data) {
      uint32_t chunk = (L - 1) / D.bl;
      key = D.stat_base + chunk * D.W;
      dropped = !row_kept(a, D, slot, chunk);
    }
  }
  // neighbouring rows share their chunk: one atomic per warp in the common case
  unsigned mask = __ballot_sync(0xffffffffu, dropped);
  if (mask == 0) return;
  uint32_t key0 = __shfl_sync(0xffffffffu, key, __ffs(mask) - 1);
  bool same = !dropped || key == key0;
  if (__all_sync(0xffffffffu, same)) {
    if ((threadIdx.x & 31) == (uint32_t)(__ffs(mask) - 1)) atomicAdd(a.chunk_dropped + key0, (uint32_t)__popc(mask));
  } else if (dropped) {
    atomicAdd(a.chunk_dropped + key, 1u);
  }
}

// ---- one-sweep mode: the ChunkMeasure of every entry from its candidates and decided type -----
__global__ void __launch_bounds__(128) csv_finish_kernel(TableArgs a) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= a.n_stat) return;
  uint32_t lo = 0, hi = a.n_docs;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (a.docs[mid].stat_base <= e) lo = mid; else hi = mid;
  }
  const TDoc D = a.docs[lo];
  if (!D.active || D.W == 0 || a.doc_exact[lo]) return;   // exact documents keep the initial values
  uint32_t rel = e - D.stat_base;
  if (rel >= D.n_chunks * D.W) return;
  uint32_t k = rel / D.W, c = rel - k * D.W;
  if (c < D.lead) return;
  ChunkMeasure M;
  finish_measure(((const ChunkCand*)a.cand)[e], a.ctype[e], a.chunk_dropped[D.stat_base + k * D.W], &M);
  ((ChunkMeasure*)a.meas)[e] = M;
}

// ---- sweep 2: measure ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 3) csv_measure_kernel(TableArgs a) {
  __shared__ uint32_t s_m[8][RFQ_MAX_W];  // plain, ip, tz(min), inf, obj, sci, fl(or), nan_left(or)
  __shared__ uint32_t s_doc, s_keep;
  ChunkMeasure* meas = (ChunkMeasure*)a.meas;
  const bool one_sweep = a.cand != nullptr;   // then only the documents marked exact are measured here
  if (one_sweep && a.doc_exact[a.n_docs] == 0) return;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    ItemCtx ic = item_ctx(a, item, &s_doc);
    if (one_sweep && !a.doc_exact[ic.d]) { __syncthreads(); continue; }
    const TDoc D = a.docs[ic.d];
    uint32_t W = D.W;
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) {
#pragma unroll
      for (int q = 0; q < 8; ++q) s_m[q][c] = q == 2 ? 6u : 0u;
    }
    if (threadIdx.x == 0) s_keep = 0;
    __syncthreads();
    if (!(a.doc_flags[ic.d] & 1)) {
      uint32_t rbp = 32;
      while (rbp < ic.nrows) rbp <<= 1;
      const uint32_t ngrp = blockDim.x / rbp, rr = threadIdx.x % rbp, grp = threadIdx.x / rbp;
      const uint32_t slot = D.r0 + 1 + ic.row0 + rr;
      if (rr < ic.nrows && grp < ngrp && row_kept(a, D, slot, ic.chunk)) {
        const RowCur rc = row_open(a, slot);
        if (grp == 0) atomicAdd(&s_keep, 1u);
        uint32_t extra = 0;
        for (uint32_t c = D.lead + grp; c < W; c += ngrp) {
          int ct = a.ctype[D.stat_base + ic.chunk * W + c];
          CellMeasure m = measure_cell(row_cell(a, D, rc, c), ct);
          if (m.is_nan) { if (!s_m[7][c]) s_m[7][c] = 1; continue; }
          if (m.plain_len > s_m[0][c]) atomicMax(&s_m[0][c], m.plain_len);
          if (m.ip_len > s_m[1][c]) atomicMax(&s_m[1][c], m.ip_len);
          if (m.tz < s_m[2][c]) atomicMin(&s_m[2][c], m.tz);
          if (m.inf_len > s_m[3][c]) atomicMax(&s_m[3][c], m.inf_len);
          if (m.obj_len > s_m[4][c]) atomicMax(&s_m[4][c], m.obj_len);
          if (m.sci_len > s_m[5][c]) atomicMax(&s_m[5][c], m.sci_len);
          if ((s_m[6][c] & m.fl) != m.fl) atomicOr(&s_m[6][c], m.fl);
          extra += m.extra_bytes;
        }
        if (extra && !one_sweep) { if (ngrp == 1) a.row_extra[slot] = extra; else atomicAdd(a.row_extra + slot, extra); }
      }
    }
    __syncthreads();
    for (uint32_t c = threadIdx.x + D.lead; c < W; c += blockDim.x) {
      ChunkMeasure* M = meas + D.stat_base + ic.chunk * W + c;
      if (s_m[0][c]) atomicMax(&M->plain_len, s_m[0][c]);
      if (s_m[1][c]) atomicMax(&M->ip_len, s_m[1][c]);
      if (s_m[2][c] < 6) atomicMin(&M->min_tz, s_m[2][c]);
      if (s_m[3][c]) atomicMax(&M->inf_len, s_m[3][c]);
      if (s_m[4][c]) atomicMax(&M->obj_len, s_m[4][c]);
      if (s_m[5][c]) atomicMax(&M->sci_len, s_m[5][c]);
      if (s_m[6][c]) atomicOr(&M->fl, s_m[6][c]);
      if (s_m[7][c]) atomicOr(&M->nan_left, 1u);
    }
    if (threadIdx.x == 0 && s_keep && !one_sweep) atomicAdd(a.n_keep + ic.d, s_keep);
    __syncthreads();
  }
}

// ---- layout: class, width, decimals of every output column; line width -----------------------
__global__ void __launch_bounds__(256) csv_layout_kernel(TableArgs a) {   // one warp per document
  const uint32_t d = blockIdx.x * 8 + (threadIdx.x >> 5);
  const uint32_t lane = threadIdx.x & 31u;
  if (d >= a.n_docs) return;
  const TDoc D = a.docs[d];
  if (!D.active) return;
  const ChunkMeasure* meas = (const ChunkMeasure*)a.meas;
  ColOut* col = (ColOut*)a.col;
  if (a.cand) {   // one-sweep mode: rows kept = data rows - rows dropna removed
    uint32_t dr = 0;
    for (uint32_t k = lane; k < D.n_chunks; k += 32) dr += a.chunk_dropped[D.stat_base + k * D.W];
#pragma unroll
    for (int o = 16; o; o >>= 1) dr += __shfl_xor_sync(0xffffffffu, dr, o);
    if (lane == 0) a.n_keep[d] = D.n_data - dr;
  }
  uint32_t carry = 0;   // code points of the columns to the left, separators included
  bool unsup = false;
  for (uint32_t j0 = 0; j0 < D.h; j0 += 32) {   // lanes take columns; a warp scan gives their offsets
    const uint32_t j = j0 + lane;
    uint32_t w1 = 0;
    ColOut o;
    o.cls = 0; o.dec = 0; o.sci = 0; o.width = 0; o.cp_off = 0;
    if (j < D.h) {
      uint32_t c = D.lead + j;
      uint32_t e0 = D.stat_base + c;
      bool nan_left = false;
      for (uint32_t k = 0; k < D.n_chunks; ++k) nan_left |= meas[e0 + (size_t)k * D.W].nan_left != 0;
      int cls = fold_column_class(a.ctype + e0, D.n_chunks, D.W, nan_left);
      if (cls == CC_UNSUP) { unsup = true; cls = CC_OBJ; }
      ColumnLayout L = layout_column(cls, meas + e0, a.ctype + e0, D.n_chunks, D.W,
                                     a.name_cp[D.col_base + j]);
      o.cls = L.cls; o.dec = L.dec; o.sci = L.sci ? 1 : 0; o.width = L.width;
      w1 = L.width + 1;
    }
    uint32_t incl = w1;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      uint32_t x = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= (uint32_t)s) incl += x;
    }
    if (j < D.h) { o.cp_off = carry + incl - w1; col[D.col_base + j] = o; }
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  const bool any_unsup = __any_sync(0xffffffffu, unsup);
  if (lane == 0) {
    a.doc_linew[d] = carry ? carry - 1 : 0;
    if (any_unsup) atomicOr(a.doc_flags + d, 2);
  }
}

// ---- bytes of every output line (indexed by record slot; slot r0 = banner + header line) ------
__global__ void __launch_bounds__(256) csv_rowbytes_kernel(TableArgs a) {
  uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot > a.n_slots) return;
  unsigned long long b = 0;
  if (slot < a.n_slots) {
    uint32_t d = ub_search(a.slot_doc_r0, a.n_docs, slot);
    const TDoc D = a.docs[d];
    if (D.active && a.doc_flags[d] == 0) {
      uint32_t L = slot - D.r0;
      uint32_t keep = a.n_keep[d];
      if (L == 0) b = D.banner_len + (keep ? a.doc_linew[d] + D.hdr_extra : D.empty_len);
      else if (L <= D.n_data && keep) {
        uint32_t chunk = (L - 1) / D.bl;
        if (row_kept(a, D, slot, chunk)) b = 1ull + a.doc_linew[d] + a.row_extra[slot];
      }
    }
  }
  a.row_bytes[slot] = b;
}

__global__ void __launch_bounds__(128) csv_docout_kernel(TableArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_docs) return;
  unsigned long long o0 = a.row_off[a.slot_doc_r0[d]], o1 = a.row_off[a.slot_doc_r0[d + 1]];
  a.doc_out_off[d] = o0;
  a.doc_out_len[d] = o1 - o0;
}

void csv_table_measure(const TableArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("csv_decide_kernel", s);
  if (!a.xcells && a.n_stat) RFQ_LAUNCH(csv_decide_kernel, (a.n_stat + 127) / 128, 128, 0, s, a);
  if (a.cand) {
    prof_mark("csv_rowstat+finish", s);
    RFQ_LAUNCH(csv_rowstat_kernel, (a.n_slots + 255) / 256, 256, 0, s, a);
    if (a.n_stat) RFQ_LAUNCH(csv_finish_kernel, (a.n_stat + 127) / 128, 128, 0, s, a);
  }
  prof_mark("csv_measure_kernel", s);
  if (a.n_items) RFQ_LAUNCH(csv_measure_kernel, table_grid(a.n_items), 256, 0, s, a);
  prof_mark("csv_layout+rowbytes+scan", s);
  RFQ_LAUNCH(csv_layout_kernel, (a.n_docs + 7) / 8, 256, 0, s, a);
  RFQ_LAUNCH(csv_rowbytes_kernel, (a.n_slots + 1 + 255) / 256, 256, 0, s, a);
  exclusive_scan<SumU64>(a.row_bytes, a.row_off, a.n_slots + 1, a.scan_tmp64, nullptr, s);
  RFQ_LAUNCH(csv_docout_kernel, (a.n_docs + 127) / 128, 128, 0, s, a);
  prof_mark("host+sync", s);
}


}  // namespace rfq
