// Shared by the three table sweeps (csv_classify.cu, csv_measure.cu, csv_emit.cu): work items,
// per-row cursors and cell fetch for the CSV and sheet paths.
#pragma once
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

// The fused sweep already ran csv_simplify on every field; it leaves the verdict in one byte per
// field (index = position in `fend`) so that emit does not scan the bytes a second time:
//   bits 0-2: 0 = needs the quote state machine (CM_CSV as found), 1 = clean ASCII unquoted span,
//             2 = "..." around a clean ASCII span, 3 = "..." whose only quotes are "" pairs,
//             4 = plain unquoted span with non-ASCII bytes, 5 = "..." around such a span
//   bits 3-7: number of "" pairs (kind 3; 31 or more pairs are filed under kind 0)
__device__ __forceinline__ CellSrc row_cell_tag_write(const TableArgs& a, const TDoc& D, const RowCur& r, uint32_t c) {
  if (!a.cell_tag || a.xcells) return row_cell(a, D, r, c);
  CellSrc s;
  s.mode = CM_CSV; s.missing = false; s.clean = false; s.p = a.arena; s.n = 0; s.bits = 0;
  if (c >= r.nf) { s.missing = true; return s; }
  const uint32_t st = c == 0 ? r.start0 : __ldg(a.fend + r.fbeg + c - 1) + 1;
  const uint32_t n0 = __ldg(a.fend + r.fbeg + c) - st;
  s.p = a.arena + D.off + st;
  s.n = n0;
  csv_simplify(s);
  {
    uint32_t tag = 0;
    if (s.mode == CM_RAW) tag = (s.clean ? 1u : 4u) + (s.n == n0 ? 0u : 1u);
    else if (s.mode == CM_CSVQ) {
      uint32_t pairs = s.n - (uint32_t)s.bits;
      if (pairs < 31u) tag = 3u | (pairs << 3);
      else { s.mode = CM_CSV; s.p -= 1; s.n += 2; s.bits = 0; }   // keep both sweeps on the same path
    }
    a.cell_tag[r.fbeg + c] = (uint8_t)tag;
  }
  return s;
}
__device__ __forceinline__ CellSrc row_cell_tag_read(const TableArgs& a, const TDoc& D, const RowCur& r, uint32_t c) {
  if (!a.cell_tag || a.xcells) return row_cell(a, D, r, c);
  CellSrc s;
  s.mode = CM_CSV; s.missing = false; s.clean = false; s.p = a.arena; s.n = 0; s.bits = 0;
  if (c >= r.nf) { s.missing = true; return s; }
  uint32_t st = c == 0 ? r.start0 : __ldg(a.fend + r.fbeg + c - 1) + 1;
  uint32_t en = __ldg(a.fend + r.fbeg + c);
  const uint32_t tag = a.cell_tag[r.fbeg + c], kind = tag & 7u;
  s.p = a.arena + D.off + st;
  s.n = en - st;
  if (kind == 2u || kind == 3u || kind == 5u) { s.p += 1; s.n -= 2; }   // inside the quotes
  if (kind == 1u || kind == 2u) { s.mode = CM_RAW; s.clean = true; }
  else if (kind == 4u || kind == 5u) s.mode = CM_RAW;
  else if (kind == 3u) { s.mode = CM_CSVQ; s.bits = s.n - (tag >> 3); }
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

__device__ __forceinline__ bool row_kept(const TableArgs& a, const TDoc& D, uint32_t slot,
                                         uint32_t chunk) {
  return !(__ldg(a.row_na + slot) == D.h && !a.chunk_nf[D.stat_base + chunk * D.W]);
}


// 8 resident CTAs of 256 threads per SM
static inline int table_grid(uint32_t n_items) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  uint32_t g = (uint32_t)sms * 8u;
  return (int)(n_items < g ? n_items : g);
}

}  // namespace rfq
