// Cell typing + numeric parse, column decisions, and the fixed-width formatter / emitter
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
#include <cuda_runtime.h>
#include "cell_ops.cuh"
#include "launch.h"
#include "pipeline.h"
#include "scan.cuh"
#include "xlsx_fetch.cuh"

namespace rfq {

#define RFQ_MAX_W 1024

__device__ __forceinline__ uint32_t ub_search(const uint32_t* __restrict__ base, uint32_t n,
                                              uint32_t x) {
  // largest i in [0, n) with base[i] <= x
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (__ldg(base + mid) <= x) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ CellSrc fetch_cell(const TableArgs& a, const TDoc& D, uint32_t slot,
                                              uint32_t c) {
  if (a.xcells) return xfetch(a, D, slot - D.r0 + 1, c);  // sheet mode: slot r0+1+row <-> sheet row row+2
  CellSrc s;
  s.mode = 0; s.missing = false; s.clean = false; s.p = a.arena; s.n = 0; s.bits = 0;
  uint32_t fbeg = __ldg(a.rec_fend + slot - 1);
  uint32_t nf = __ldg(a.rec_fend + slot) - fbeg;
  if (c >= nf) { s.missing = true; return s; }
  uint32_t st = c == 0 ? __ldg(a.rec_start + slot) : __ldg(a.fend + fbeg + c - 1) + 1;
  uint32_t en = __ldg(a.fend + fbeg + c);
  s.p = a.arena + D.off + st;
  s.n = en - st;
  csv_simplify(s);
  return s;
}

// per-row state hoisted out of the column loop (a thread walks the cells of ONE row)
struct RowCur { uint32_t slot, fbeg, nf, start0; };
__device__ __forceinline__ RowCur row_open(const TableArgs& a, uint32_t slot) {
  RowCur r; r.slot = slot; r.fbeg = 0; r.nf = 0; r.start0 = 0;
  if (!a.xcells) {
    r.fbeg = __ldg(a.rec_fend + slot - 1);
    r.nf = __ldg(a.rec_fend + slot) - r.fbeg;
    r.start0 = __ldg(a.rec_start + slot);
  }
  return r;
}
__device__ __forceinline__ CellSrc row_cell(const TableArgs& a, const TDoc& D, const RowCur& r, uint32_t c) {
  if (a.xcells) return xfetch(a, D, r.slot - D.r0 + 1, c);
  CellSrc s;
  s.mode = 0; s.missing = false; s.clean = false; s.p = a.arena; s.n = 0; s.bits = 0;
  if (c >= r.nf) { s.missing = true; return s; }
  uint32_t st = c == 0 ? r.start0 : __ldg(a.fend + r.fbeg + c - 1) + 1;
  uint32_t en = __ldg(a.fend + r.fbeg + c);
  s.p = a.arena + D.off + st;
  s.n = en - st;
  csv_simplify(s);
  return s;
}

struct ItemCtx {
  uint32_t d, row0, nrows, chunk;
};
__device__ __forceinline__ ItemCtx item_ctx(const TableArgs& a, uint32_t item, uint32_t* s_doc) {
  if (threadIdx.x == 0) *s_doc = ub_search(a.item_doc_base, a.n_docs, item);
  __syncthreads();
  ItemCtx c;
  c.d = *s_doc;
  const TDoc& D = a.docs[c.d];
  c.row0 = (item - D.item_base) * D.rb;
  c.nrows = min(D.rb, D.n_data - c.row0);
  c.chunk = c.row0 / D.bl;
  return c;
}

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

// ---- decide chunk types (one thread per document; entries are few) ----------------------------
__global__ void __launch_bounds__(128) csv_decide_kernel(TableArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_docs) return;
  const TDoc D = a.docs[d];
  if (!D.active) return;
  const ChunkStat* stat = (const ChunkStat*)a.stat;
  int flags = 0;
  for (uint32_t k = 0; k < D.n_chunks; ++k) {
    bool nf = false;
    for (uint32_t c = 0; c < D.W; ++c) {
      uint32_t e = D.stat_base + k * D.W + c;
      ChunkStat S = stat[e];
      if (S.flags & SF_BAD_UTF8) flags |= 1;
      int t = decide_chunk_type(S);
      a.ctype[e] = (uint8_t)t;
      if (t == CT_STRNF && c >= D.lead) nf = true;
    }
    a.chunk_nf[D.stat_base + k * D.W] = nf ? 1 : 0;
  }
  if (flags) atomicOr(a.doc_flags + d, flags);
}

// ---- sweep 2: measure ---------------------------------------------------------------------------
__device__ __forceinline__ bool row_kept(const TableArgs& a, const TDoc& D, uint32_t slot,
                                         uint32_t chunk) {
  return !(__ldg(a.row_na + slot) == D.h && !a.chunk_nf[D.stat_base + chunk * D.W]);
}

__global__ void __launch_bounds__(256, 3) csv_measure_kernel(TableArgs a) {
  __shared__ uint32_t s_m[8][RFQ_MAX_W];  // plain, ip, tz(min), inf, obj, sci, fl(or), nan_left(or)
  __shared__ uint32_t s_doc, s_keep;
  ChunkMeasure* meas = (ChunkMeasure*)a.meas;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    ItemCtx ic = item_ctx(a, item, &s_doc);
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
        if (extra) { if (ngrp == 1) a.row_extra[slot] = extra; else atomicAdd(a.row_extra + slot, extra); }
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
    if (threadIdx.x == 0 && s_keep) atomicAdd(a.n_keep + ic.d, s_keep);
    __syncthreads();
  }
}

// ---- layout: class, width, decimals of every output column; line width -----------------------
__global__ void __launch_bounds__(128) csv_layout_kernel(TableArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_docs) return;
  const TDoc D = a.docs[d];
  if (!D.active) return;
  const ChunkMeasure* meas = (const ChunkMeasure*)a.meas;
  ColOut* col = (ColOut*)a.col;
  uint32_t off = 0;
  int flags = 0;
  for (uint32_t j = 0; j < D.h; ++j) {
    uint32_t c = D.lead + j;
    uint32_t e0 = D.stat_base + c;
    bool nan_left = false;
    for (uint32_t k = 0; k < D.n_chunks; ++k) nan_left |= meas[e0 + (size_t)k * D.W].nan_left != 0;
    int cls = fold_column_class(a.ctype + e0, D.n_chunks, D.W, nan_left);
    if (cls == CC_UNSUP) { flags |= 2; cls = CC_OBJ; }
    ColumnLayout L = layout_column(cls, meas + e0, a.ctype + e0, D.n_chunks, D.W,
                                   a.name_cp[D.col_base + j]);
    ColOut o;
    o.cls = L.cls; o.dec = L.dec; o.sci = L.sci ? 1 : 0; o.width = L.width; o.cp_off = off;
    col[D.col_base + j] = o;
    off += L.width + 1;
  }
  a.doc_linew[d] = off ? off - 1 : 0;
  if (flags) atomicOr(a.doc_flags + d, flags);
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

// ---- sweep 3: emit -------------------------------------------------------------------------------
// A work item's rows are one contiguous byte range of the output.  The CTA builds that range in
// shared memory (pre-filled with the padding blanks, so a cell only writes its own characters),
// then streams it out with aligned 16-byte stores.  Ranges that do not fit (very long cells)
// take the direct path: same formatting, byte stores to global memory.
#define RFQ_EMIT_SMEM (80u << 10)

// number -> text in a small local buffer; returns length, or -1 when the value needs the
// big-integer path (|v| >= 2^63 printed in fixed notation)
__device__ __forceinline__ int format_number(const CellSrc& s, int ct, int kind, int dec,
                                             const TextInfo& t, uint8_t* buf) {
  switch (kind) {
    case PR_BOOL: {
      const char* w = cell_bool(s, t) == 1 ? "True" : "False";
      int k = 0;
      while (w[k]) { buf[k] = (uint8_t)w[k]; ++k; }
      return k;
    }
    case PR_INT: {
      int64_t iv = 0; cell_int64(s, &iv);
      int n = int64_len(iv);
      uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
      write_u64_rev(mag, buf + n);
      if (iv < 0) buf[0] = '-';
      return n;
    }
    case PR_UINT: {
      UintTok u; cell_uint64(s, &u);
      int n = ndigits_u64(u.v);
      write_u64_rev(u.v, buf + n);
      return n;
    }
    case PR_FIXED: case PR_OBJFLOAT: {
      double v = cell_as_double(s, ct);
      Fixed6 f = fixed6(v);
      if (f.big) return -1;
      return fixed_write(f, v, kind == PR_FIXED ? dec : objfloat_dec(f), kind == PR_OBJFLOAT, buf);
    }
    case PR_SCI: return sci_write(sci7(cell_as_double(s, ct)), buf);
    case PR_PYINT: return -1;   // arbitrary length: measured and written in place
    case PR_DATE64: return (int)date64_write((int64_t)s.bits, dec, buf);
    case PR_DATETIME: return (int)pydatetime_write((int64_t)s.bits, buf);
    case PR_TIME: return (int)pytime_write((int64_t)s.bits, buf);
    default: return 0;
  }
}

// writes the cell right-justified into its field; `field` points at the first byte of the
// column's field inside the row; `fill` = also write the padding blanks (direct path)
// returns printed bytes - printed code points (non-zero only for non-ASCII strings)
__device__ __forceinline__ uint32_t emit_cell(const CellSrc& src, int ct, int kind, int dec,
                                              uint32_t width, uint8_t* field, bool fill) {
  TextInfo t = cell_text(src);
  bool na = cell_is_nan(src, t, ct);
  if (na) {
    if (fill) for (uint32_t q = 0; q < width; ++q) field[q] = ' ';
    if (kind == PR_DATE64) { field[width - 3] = 'N'; field[width - 2] = 'a'; field[width - 1] = 'T'; }
    return 0;
  }
  if (kind == PR_STR) {
    uint32_t ncp = t.ncp + t.nesc, pad = width - ncp;
    if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
    print_cell(src, ct, kind, dec, field + pad);
    return t.nbytes - t.ncp;
  }
  uint8_t buf[40];
  int n = format_number(src, ct, kind, dec, t, buf);
  if (n < 0) {  // rare: huge value in fixed notation
    uint32_t ncp, nb;
    printed_size(src, ct, kind, dec, &ncp, &nb);
    uint32_t pad = width - ncp;
    if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
    print_cell(src, ct, kind, dec, field + pad);
    return 0;
  }
  uint32_t pad = width - (uint32_t)n;
  if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
  for (int q = 0; q < n; ++q) field[pad + q] = buf[q];
  return 0;
}

__global__ void __launch_bounds__(256) csv_emit_kernel(TableArgs a, uint8_t* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t s_buf[];
  __shared__ uint32_t s_doc;
  const ColOut* col = (const ColOut*)a.col;
  for (uint32_t item = a.item_lo + blockIdx.x; item < a.item_hi; item += gridDim.x) {
    ItemCtx ic = item_ctx(a, item, &s_doc);
    const TDoc D = a.docs[ic.d];
    if (a.doc_flags[ic.d] == 0 && a.n_keep[ic.d] != 0) {
      uint32_t W = D.W;
      uint32_t slot0 = D.r0 + 1 + ic.row0;
      unsigned long long g0 = a.row_off[slot0], g1 = a.row_off[slot0 + ic.nrows];
      unsigned long long nbytes = g1 - g0;
      uint32_t mis = (uint32_t)((uintptr_t)(out + g0) & 15u);  // keep smem and global co-aligned
      bool staged = nbytes + mis + 16 <= RFQ_EMIT_SMEM;
      if (staged) {
        uint32_t nvec = (uint32_t)((mis + nbytes + 15) >> 4);
        uint4 sp = make_uint4(0x20202020u, 0x20202020u, 0x20202020u, 0x20202020u);
        for (uint32_t i = threadIdx.x; i < nvec; i += blockDim.x) reinterpret_cast<uint4*>(s_buf)[i] = sp;
        __syncthreads();
      }
      uint8_t* base = staged ? s_buf + mis : out + g0;   // base[x] <-> out[g0 + x]
      // A thread owns one row and walks its columns; a warp therefore works on 32 consecutive
      // rows of the SAME column at any moment.  Items with fewer rows than threads split the
      // columns over thread groups.
      uint32_t rbp = 32;
      while (rbp < ic.nrows) rbp <<= 1;
      const uint32_t ngrp = blockDim.x / rbp, rr = threadIdx.x % rbp, grp = threadIdx.x / rbp;
      const uint32_t slot = slot0 + rr;
      if (rr < ic.nrows && grp < ngrp && row_kept(a, D, slot, ic.chunk)) {
        const RowCur rc = row_open(a, slot);
        uint8_t* rowp = base + (a.row_off[slot] - g0);
        const uint32_t extra = __ldg(a.row_extra + slot);
        const uint8_t* ctrow = a.ctype + D.stat_base + ic.chunk * W;
        uint32_t xb = 0;  // extra bytes of the columns to the left (rows with non-ASCII text only)
        for (uint32_t c = D.lead + grp; c < W; c += ngrp) {
          uint32_t j = c - D.lead;
          ColOut co = col[D.col_base + j];
          int ct = ctrow[c];
          CellSrc src = row_cell(a, D, rc, c);
          if (ct == CT_XOBJ) ct = cell_own_ctype(src);
          int kind = print_kind(co.cls, ct, co.sci != 0);
          if (ngrp > 1 && j && extra) {   // columns are split over groups: recompute the shift
            xb = 0;
            for (uint32_t c2 = D.lead; c2 < c; ++c2) {
              ColOut c2o = col[D.col_base + c2 - D.lead];
              int ct2 = ctrow[c2];
              CellSrc s2 = row_cell(a, D, rc, c2);
              if (ct2 == CT_XOBJ) ct2 = cell_own_ctype(s2);
              int k2 = print_kind(c2o.cls, ct2, c2o.sci != 0);
              if (k2 != PR_STR) continue;
              uint32_t cp2, nb2;
              printed_size(s2, ct2, k2, c2o.dec, &cp2, &nb2);
              xb += nb2 - cp2;
            }
          }
          uint8_t* p = rowp;
          if (j == 0) { *p = '\n'; p += 1; }
          else { p += co.cp_off + xb; if (!staged) *p = ' '; p += 1; }
          uint32_t ex = emit_cell(src, ct, kind, co.dec, co.width, p, !staged);
          if (ngrp == 1) xb += ex;
        }
      }
      if (staged) {
        __syncthreads();
        // head bytes up to the first 16-byte boundary, aligned body, tail bytes
        uint8_t* g = out + g0;
        uint32_t head = mis ? min((unsigned long long)(16 - mis), nbytes) : 0u;
        if (threadIdx.x < head) g[threadIdx.x] = s_buf[mis + threadIdx.x];
        unsigned long long body = (nbytes - head) >> 4;
        const uint4* sv = reinterpret_cast<const uint4*>(s_buf + mis + head);
        uint4* gv = reinterpret_cast<uint4*>(g + head);
        for (unsigned long long i = threadIdx.x; i < body; i += blockDim.x) gv[i] = sv[i];
        uint32_t tail = (uint32_t)((nbytes - head) & 15u);
        if (threadIdx.x < tail) g[head + body * 16 + threadIdx.x] = s_buf[mis + head + body * 16 + threadIdx.x];
      }
    }
    __syncthreads();
  }
}

// banner + header line (or the "Empty DataFrame" text): one warp per document
__global__ void __launch_bounds__(256) csv_header_emit_kernel(TableArgs a, uint8_t* __restrict__ out) {
  uint32_t d = blockIdx.x * 8 + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (d >= a.n_docs) return;
  const TDoc D = a.docs[d];
  if (!D.active || a.doc_flags[d] != 0) return;
  const ColOut* col = (const ColOut*)a.col;
  uint8_t* p = out + a.doc_out_off[d];
  for (uint32_t i = lane; i < D.banner_len; i += 32) p[i] = a.names[D.banner_off + i];
  p += D.banner_len;
  if (a.n_keep[d] == 0) {
    for (uint32_t i = lane; i < D.empty_len; i += 32) p[i] = a.names[D.empty_off + i];
    return;
  }
  for (uint32_t j = lane; j < D.h; j += 32) {
    ColOut co = col[D.col_base + j];
    uint32_t xb = 0;
    for (uint32_t j2 = 0; j2 < j; ++j2) xb += a.name_len[D.col_base + j2] - a.name_cp[D.col_base + j2];
    bool numeric = co.cls == CC_INT || co.cls == CC_UINT || co.cls == CC_FLOAT || co.cls == CC_BOOL;
    uint32_t cp = a.name_cp[D.col_base + j] + (numeric ? 1u : 0u);
    uint8_t* q = p + co.cp_off + xb;
    if (j) q[-1] = ' ';
    uint32_t pad = co.width - cp;
    for (uint32_t i = 0; i < pad; ++i) q[i] = ' ';
    q += pad;
    if (numeric) *q++ = ' ';
    const uint8_t* nm = a.names + a.name_off[D.col_base + j];
    uint32_t nl = a.name_len[D.col_base + j];
    for (uint32_t i = 0; i < nl; ++i) q[i] = nm[i];
  }
}

// ---------------------------------------------------------------------------------------------
static int table_grid(uint32_t n_items) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  uint32_t g = (uint32_t)sms * 8u;  // 8 resident CTAs of 256 threads per SM
  return (int)(n_items < g ? n_items : g);
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

void csv_table_measure(const TableArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("csv_decide_kernel", s);
  if (!a.xcells) RFQ_LAUNCH(csv_decide_kernel, (a.n_docs + 127) / 128, 128, 0, s, a);
  prof_mark("csv_measure_kernel", s);
  if (a.n_items) RFQ_LAUNCH(csv_measure_kernel, table_grid(a.n_items), 256, 0, s, a);
  prof_mark("csv_layout+rowbytes+scan", s);
  RFQ_LAUNCH(csv_layout_kernel, (a.n_docs + 127) / 128, 128, 0, s, a);
  RFQ_LAUNCH(csv_rowbytes_kernel, (a.n_slots + 1 + 255) / 256, 256, 0, s, a);
  exclusive_scan<SumU64>(a.row_bytes, a.row_off, a.n_slots + 1, a.scan_tmp64, nullptr, s);
  RFQ_LAUNCH(csv_docout_kernel, (a.n_docs + 127) / 128, 128, 0, s, a);
  prof_mark("host+sync", s);
}

void csv_table_emit_header(const TableArgs& a, uint8_t* out, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("csv_header_emit_kernel", s);
  RFQ_LAUNCH(csv_header_emit_kernel, (a.n_docs + 7) / 8, 256, 0, s, a, out);
}

void csv_table_emit_items(const TableArgs& a0, uint8_t* out, uint32_t lo, uint32_t hi, cudaStream_t s) {
  if (hi <= lo) return;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(csv_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RFQ_EMIT_SMEM);
    attr_set = true;
  }
  TableArgs a = a0;
  a.item_lo = lo; a.item_hi = hi;
  prof_mark("csv_emit_kernel", s);
  RFQ_LAUNCH(csv_emit_kernel, table_grid(hi - lo), 256, RFQ_EMIT_SMEM, s, a, out);
}

void csv_table_emit(const TableArgs& a, uint8_t* out, cudaStream_t s) {
  csv_table_emit_header(a, out, s);
  csv_table_emit_items(a, out, 0, a.n_items, s);
}

}  // namespace rfq
