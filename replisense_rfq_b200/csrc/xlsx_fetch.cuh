// Cell access of the sheet path for the shared table kernels (measure / emit): sheet row ->
// <row> element -> typed cell record -> CellSrc.
#pragma once
#include "pipeline.h"
#include "xlsx_cells.cuh"

namespace rfq {

__device__ __forceinline__ CellSrc xfetch(const TableArgs& a, const TDoc& D, uint32_t sheet_row, uint32_t c) {
  CellSrc s; s.p = a.arena; s.n = 0; s.mode = CM_NULL; s.missing = false; s.clean = false; s.bits = 0;
  uint32_t e = __ldg(a.x_r2elem + D.f0 + sheet_row);
  if (e == 0xFFFFFFFFu) return s;
  uint32_t c0 = __ldg(a.x_row_c0 + e), c1 = __ldg(a.x_row_c0 + e + 1);
  const XCell* xc = (const XCell*)a.xcells;
  uint32_t k = c0 + c;
  if (!(k < c1 && xc[k].col == c + 1)) {
    // a row with gaps: columns are increasing and column c + 1 cannot sit after position c, so the cell is a
    // few records before the dense guess (one step per gap before it); a very sparse row is bisected
    if (c1 == c0) return s;
    uint32_t hi = min(k, c1 - 1u);
    int steps = 0;
    while (hi > c0 && xc[hi].col > c + 1 && steps < 4) { --hi; ++steps; }
    if (xc[hi].col > c + 1) {
      uint32_t lo = c0, h2 = hi;
      while (lo < h2) { uint32_t mid = (lo + h2) >> 1; if (xc[mid].col < c + 1) lo = mid + 1; else h2 = mid; }
      hi = lo;
    }
    if (xc[hi].col != c + 1) return s;
    k = hi;
  }
  XCell x = xc[k];
  s.mode = (uint8_t)x.mode; s.bits = x.bits;
  if (x.mode == CM_RAW) {
    s.p = a.x_str_arena + (uint32_t)(x.bits >> 32); s.n = (uint32_t)x.bits & ~RFQ_SST_CLEAN;
    s.clean = ((uint32_t)x.bits & RFQ_SST_CLEAN) != 0;   // plain ASCII text takes the inline paths of measure / emit
  }
  else if (x.mode == CM_XML) { s.p = a.arena + D.off + (uint32_t)(x.bits >> 32); s.n = (uint32_t)x.bits; }
  return s;
}


}  // namespace rfq
