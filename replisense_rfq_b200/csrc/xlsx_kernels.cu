// sharedStrings decoder and <row>/<c>/<v> cell decoder for sheet XML:
// what openpyxl's read-only reader + pandas' _convert_cell do per element (xlsx_core.cuh /
// xlsx_cells.cuh cite the lines), reached from app/file_parser.py:225.
//
// XML text never contains a literal '<', so every '<' starts a tag and tags can be classified
// independently: a warp owns a 512-byte tile, each lane looks for '<' in its 16 bytes and
// classifies the tag from the bytes that follow.  Counts per tile -> exclusive scans -> a
// second sweep writes the positions of every <row>, <c> and <si>.  Then one thread per element:
// decode a shared string (entities, line ends, rich-text runs, phonetic runs skipped), read a
// row number, type a cell.  HBM-bound byte work: no tensor cores.
#include <cuda_runtime.h>
#include "launch.h"
#include "pipeline.h"
#include "scan.cuh"
#include "xlsx_cells.cuh"
#include "xlsx_fetch.cuh"

namespace rfq {

__device__ __forceinline__ uint32_t x_find(const uint32_t* __restrict__ base, uint32_t n, uint32_t x) {
  uint32_t lo = 0, hi = n;   // largest i in [0, n) with base[i] <= x
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (__ldg(base + mid) <= x) lo = mid; else hi = mid;
  }
  return lo;
}

// mask of the lane's 16 bytes that are '<'
__device__ __forceinline__ uint32_t lt_mask(const uint8_t* __restrict__ doc, uint32_t len, uint32_t base) {
  if (base >= len) return 0u;
  uint4 v = __ldg(reinterpret_cast<const uint4*>(doc + base));
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t m = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q)   // 0xFF per matching byte -> one bit per byte
    m |= (((__vcmpeq4(w[q], 0x3C3C3C3Cu) & 0x08040201u) * 0x01010101u) >> 24) << (4 * q);
  const uint32_t left = len - base;   // base < len here
  return left >= 16 ? m : m & ((1u << left) - 1u);
}

template <bool WRITE>
__global__ void __launch_bounds__(256) xml_tags_kernel(XmlArgs a) {
  uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (t >= a.n_tiles) return;
  uint32_t d = warp_find_doc(a.tile_base, a.n_docs, t, lane);
  uint32_t tile = t - a.tile_base[d];
  uint32_t len = a.doc_len[d];
  const uint8_t* doc = a.arena + a.doc_off[d];
  uint32_t base = tile * RFQ_TILE + lane * 16;
  uint32_t nr = 0, nc = 0, ns = 0, bang = 0;
  uint32_t kinds = 0;  // 2 bits per byte: 1 row, 2 cell, 3 si
  if (!WRITE) {
    const uint32_t m = lt_mask(doc, len, base);
    const bool is_sst = (a.kind[d] & 1) != 0;
    for (uint32_t mm = m; mm; mm &= mm - 1) {
      int k = __ffs(mm) - 1;
      int tg = xml_classify(doc + base + k, doc + len);
      if (tg == XT_BANG) bang = 1;
      if (!is_sst) {
        if (tg == XT_ROW) { ++nr; kinds |= 1u << (2 * k); }
        else if (tg == XT_C) { ++nc; kinds |= 2u << (2 * k); }
      } else if (tg == XT_SI) { ++ns; kinds |= 3u << (2 * k); }
    }
    a.lane_kinds[(size_t)t * 32 + lane] = kinds;      // coalesced; the index pass never looks at the bytes again
  } else {
    kinds = __ldg(a.lane_kinds + (size_t)t * 32 + lane);
    const uint32_t lo = kinds & 0x55555555u, hi = (kinds >> 1) & 0x55555555u;
    nr = (uint32_t)__popc(lo & ~hi); nc = (uint32_t)__popc(hi & ~lo); ns = (uint32_t)__popc(lo & hi);
  }
  if (!WRITE) {
    if (bang) atomicOr(&a.shape[d].flags, XU_BANG);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      nr += __shfl_xor_sync(0xffffffffu, nr, o);
      nc += __shfl_xor_sync(0xffffffffu, nc, o);
      ns += __shfl_xor_sync(0xffffffffu, ns, o);
    }
    if (lane == 0) { a.tile_cnt[0][t] = nr; a.tile_cnt[1][t] = nc; a.tile_cnt[2][t] = ns; }
  } else {
    uint32_t pr = nr, pc = nc, ps = ns;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t x = __shfl_up_sync(0xffffffffu, pr, o);
      uint32_t y = __shfl_up_sync(0xffffffffu, pc, o);
      uint32_t z = __shfl_up_sync(0xffffffffu, ps, o);
      if (lane >= o) { pr += x; pc += y; ps += z; }
    }
    uint32_t ir = a.tile_pre[0][t] + pr - nr, ic = a.tile_pre[1][t] + pc - nc, is = a.tile_pre[2][t] + ps - ns;
    for (uint32_t mm = (kinds | (kinds >> 1)) & 0x55555555u; mm; mm &= mm - 1) {
      int k = (__ffs(mm) - 1) >> 1;
      uint32_t kd = (kinds >> (2 * k)) & 3u;
      if (kd == 1) { a.rowpos[ir] = base + k; a.row_c0[ir] = ic; ++ir; }
      else if (kd == 2) {   // its document and the <row> element it follows are known right here
        a.cpos[ic] = base + k; a.cell_doc[ic] = d;
        a.cell_row[ic] = ir > a.tile_pre[0][a.tile_base[d]] ? ir - 1 : 0xFFFFFFFFu;
        ++ic;
      }
      else if (kd == 3) { a.sipos[is++] = base + k; }
    }
  }
}

__global__ void __launch_bounds__(256) xml_shape_kernel(XmlArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d > a.n_docs) return;
  uint32_t t0 = a.tile_base[d < a.n_docs ? d : a.n_docs];
  uint32_t rb = a.tile_pre[0][t0], cb = a.tile_pre[1][t0], sb = a.tile_pre[2][t0];
  a.row_base[d] = rb; a.cell_base[d] = cb; a.si_base[d] = sb;
  if (d == a.n_docs) return;
  uint32_t t1 = a.tile_base[d + 1];
  XDocShape& s = a.shape[d];
  s.n_rows = a.tile_pre[0][t1] - rb; s.n_cells = a.tile_pre[1][t1] - cb; s.n_si = a.tile_pre[2][t1] - sb;
  s.row_base = rb; s.cell_base = cb; s.si_base = sb;
}

// ---- shared strings ---------------------------------------------------------------------------
template <bool WRITE>
__global__ void __launch_bounds__(128) sst_kernel(XmlArgs a) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_si) return;
  uint32_t d = x_find(a.si_base, a.n_docs + 1, i);
  const uint8_t* doc = a.arena + a.doc_off[d];
  uint32_t flags = 0;
  if (!WRITE) {
    a.si_len[i] = sst_decode_si(doc + a.sipos[i], doc + a.doc_len[d], nullptr, &flags);
    if (flags) atomicOr(&a.shape[d].flags, flags);
  } else if (a.si_len[i]) {
    uint8_t* out = a.str_arena + a.si_off[i];
    const uint32_t n = a.si_len[i];
    sst_decode_si(doc + a.sipos[i], doc + a.doc_len[d], out, &flags);
    // plain ASCII without control bytes, decided once per shared string (bit 31 of its length): every cell
    // that refers to it is fetched three times, and each fetch used to scan the string for exactly this
    bool clean = true;
    for (uint32_t k = 0; k < n; ++k) if (out[k] < 0x0E || out[k] >= 0x80) { clean = false; break; }
    if (clean) a.si_len[i] = n | RFQ_SST_CLEAN;
  }
}

// ---- rows -----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) xrow_kernel(XmlArgs a) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_rows) return;
  uint32_t d = x_find(a.row_base, a.n_docs + 1, i);
  const uint8_t* doc = a.arena + a.doc_off[d];
  uint32_t flags = 0;
  a.rownum[i] = xlsx_row_number(doc + a.rowpos[i], doc + a.doc_len[d], &flags);
  if (flags) atomicOr(&a.shape[d].flags, flags);
}

// rows without r= continue the count of the row before them (openpyxl worksheet/_reader.py parse_row:
// row_counter = int(r) when r= is there, else row_counter += 1).  A part that never writes r= is numbered
// by position; in a part that writes it on SOME rows a row walks back to the nearest row that has one
// (a number other than 0 is final: either it was read from r= or another thread of this kernel derived it).
#define RFQ_XROW_WALK 4096u
__global__ void __launch_bounds__(128) xrow_fix_kernel(XmlArgs a) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_rows) return;
  uint32_t d = x_find(a.row_base, a.n_docs + 1, i);
  const uint32_t fl = a.shape[d].flags, rb = a.row_base[d];
  if (xu_rows_by_position(fl)) { a.rownum[i] = i - rb + 1; return; }
  if (!(fl & XU_ROW_NOREF)) return;
  volatile uint32_t* rn = a.rownum;
  if (rn[i] != 0) return;
  uint32_t j = i, steps = 0;
  while (j > rb && rn[j - 1] == 0 && steps < RFQ_XROW_WALK) { --j; ++steps; }
  if (steps >= RFQ_XROW_WALK) { atomicOr(&a.shape[d].flags, XU_ORDER); return; }   // a run this long: declined, not crawled
  const uint32_t before = j > rb ? rn[j - 1] : 0u;     // the counter starts at 0
  rn[i] = before + (i - j) + 1;
}

// ---- cells ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) xcell_kernel(XmlArgs a) {
  uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.n_cells) return;
  uint32_t d = a.cell_doc[k];
  const uint8_t* doc = a.arena + a.doc_off[d];
  SstView sv; sv.off = nullptr; sv.len = nullptr; sv.n = 0;
  int sd = a.sst_of[d];
  if (sd >= 0) {
    uint32_t sb = a.si_base[sd];
    sv.off = a.si_off + sb; sv.len = a.si_len + sb; sv.n = a.si_base[sd + 1] - sb;
  }
  uint32_t flags = 0, defer = 0;
  XCell x = xlsx_parse_cell(doc, a.doc_len[d], a.cpos[k], sv, a.date_ids + a.date_off[d],
                            a.date_off[d + 1] - a.date_off[d], (a.kind[d] & 2) != 0, &flags, &defer);
  // a number that needs the general parsers (exponent, 19+ digits, ...) is finished by xnum_kernel: there every
  // lane has one, here it would be two lanes of the warp running a long path while thirty wait
  if (defer) a.num_list[atomicAdd(a.num_count, 1u)] = make_uint2(k, defer - 1u);
  // row element of this cell, ordering checks
  uint32_t rb = a.row_base[d], nr = a.row_base[d + 1] - rb;
  uint32_t e = a.cell_row[k];
  if (e == 0xFFFFFFFFu) { flags |= XU_ORDER; e = rb; nr = 0; }   // a <c> before any <row> of its document
  uint32_t r = nr ? a.rownum[e] : 0;
  if (e > rb && a.rownum[e - 1] >= r) flags |= XU_ORDER;
  if ((flags & XU_C_NOREF) && nr) x.col = k - a.row_c0[e] + 1;   // col_counter += 1 from the row start (xcell_fix_kernel
                                                                 // corrects it when an earlier cell of the row has r=)
  ((XCell*)a.xcells)[k] = x;
  if (flags) atomicOr(&a.shape[d].flags, flags);
  if (!xcell_is_empty(x) && r >= 1 && r <= a.cap_rows[d]) {
    if (x.col > a.shape[d].W) atomicMax(&a.shape[d].W, x.col);
    if (r > a.shape[d].lastrow) atomicMax(&a.shape[d].lastrow, r);
  }
}

// cells without r= in a part that writes r= on SOME cells: col_counter += 1 from the nearest earlier cell of
// the row that has one (openpyxl worksheet/_reader.py parse_cell).  Columns taken from r= are final.
__device__ __forceinline__ bool xcell_has_ref(const uint8_t* doc, uint32_t len, uint32_t cpos) {
  const uint8_t* p = doc + cpos;
  const uint8_t* e = doc + len;
  const uint8_t* g = find_byte(p, e, '>');
  if (g >= e) return false;
  XAttrRTS at;
  xml_attrs_rts(p, g, &at);
  return (at.have & 1u) && at.vn[0] != 0;
}
__global__ void __launch_bounds__(128) xcell_fix_kernel(XmlArgs a) {
  uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.n_cells) return;
  const uint32_t d = a.cell_doc[k];
  const uint32_t fl = a.shape[d].flags;
  if (!((fl & XU_C_NOREF) && (fl & XU_C_HASREF)) || (fl & ~(uint32_t)RFQ_XU_INFO)) return;
  const uint32_t e = a.cell_row[k];
  if (e == 0xFFFFFFFFu) return;
  const uint8_t* doc = a.arena + a.doc_off[d];
  const uint32_t len = a.doc_len[d];
  if (xcell_has_ref(doc, len, a.cpos[k])) return;
  XCell* xc = (XCell*)a.xcells;
  const uint32_t c0 = a.row_c0[e];
  for (uint32_t j = k; j > c0; --j) {
    if (xcell_has_ref(doc, len, a.cpos[j - 1])) {
      const uint32_t col = xc[j - 1].col + (k - (j - 1));
      xc[k].col = col;
      const uint32_t r = a.rownum[e];
      if (!xcell_is_empty(xc[k]) && r >= 1 && r <= a.cap_rows[d] && col > a.shape[d].W) atomicMax(&a.shape[d].W, col);
      return;
    }
  }
}

// the numbers xcell_kernel put aside, densely: one per thread
__global__ void __launch_bounds__(128) xnum_kernel(XmlArgs a) {
  const uint32_t n = *a.num_count;
  XCell* xc = (XCell*)a.xcells;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 e = a.num_list[i];
    const uint32_t k = e.x, d = a.cell_doc[k];
    XCell x = xc[k];
    const uint8_t* lit = a.arena + a.doc_off[d] + (uint32_t)(x.bits >> 32);
    uint32_t flags = 0;
    x = xlsx_number_general(x, lit, (uint32_t)x.bits, (int)e.y, (a.kind[d] & 2) != 0, &flags);
    xc[k] = x;
    if (flags) atomicOr(&a.shape[d].flags, flags);
  }
}

// column order inside a row (needs the finished cell records)
__global__ void __launch_bounds__(128) xcell_order_kernel(XmlArgs a) {
  uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k + 1 >= a.n_cells) return;
  uint32_t d = a.cell_doc[k];
  if (a.cell_doc[k + 1] != d) return;
  uint32_t e0 = a.cell_row[k], e1 = a.cell_row[k + 1];
  if (e0 == 0xFFFFFFFFu) return;
  const XCell* xc = (const XCell*)a.xcells;
  if (e0 == e1 && xc[k + 1].col <= xc[k].col) atomicOr(&a.shape[d].flags, XU_ORDER);
}

__global__ void __launch_bounds__(128) xrow_order_kernel(XmlArgs a) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_rows) return;
  uint32_t d = x_find(a.row_base, a.n_docs + 1, i);
  if (i > a.row_base[d] && a.rownum[i - 1] >= a.rownum[i]) atomicOr(&a.shape[d].flags, XU_ORDER);
}

// sheet row number -> <row> element, for the rows that made it into the table
__global__ void __launch_bounds__(128) xscatter_kernel(XmlArgs a, const uint32_t* __restrict__ r2_base,
                                                       uint32_t* __restrict__ r2elem) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_rows) return;
  uint32_t d = x_find(a.row_base, a.n_docs + 1, i);
  uint32_t r = a.rownum[i];
  if (!xu_problem(a.shape[d].flags) && r >= 1 && r <= a.shape[d].lastrow) r2elem[r2_base[d] + r] = i;
}

// header cells (sheet row 1): decoded text, 256 bytes per column; meta = {mode, length}
__global__ void __launch_bounds__(128) xheader_kernel(TableArgs a, uint8_t* __restrict__ hdr_buf,
                                                      uint32_t* __restrict__ hdr_meta) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= a.n_cols) return;
  // column -> document
  uint32_t lo = 0, hi = a.n_docs;
  while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (a.docs[mid].col_base <= j) lo = mid; else hi = mid; }
  const TDoc D = a.docs[lo];
  uint32_t c = j - D.col_base;
  CellSrc s = xfetch(a, D, 1, c);
  uint32_t n = 0;
  uint8_t* o = hdr_buf + (size_t)j * 256;
  if (s.mode == CM_RAW) { for (; n < s.n && n < 256; ++n) o[n] = s.p[n]; if (s.n > 256) n = 0xFFFFFFFFu; }
  else if (s.mode == CM_XML) {
    XmlIter it(s.p, s.n); int ch;
    while ((ch = it.next()) >= 0) { if (n < 256) o[n] = (uint8_t)ch; ++n; }
    if (n > 256 || it.bad) n = 0xFFFFFFFFu;
  }
  else if (s.mode == CM_INT) {   // integer column label: its decimal digits
    int64_t iv = (int64_t)s.bits;
    n = (uint32_t)int64_len(iv);
    uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
    write_u64_rev(mag, o + n);
    if (iv < 0) o[0] = '-';
  }
  else if (s.mode == CM_BOOL) { const char* w = (s.bits & 1) ? "True" : "False"; while (w[n]) { o[n] = (uint8_t)w[n]; ++n; } }
  else if (s.mode == CM_FLOAT) {   // a float label of an object Index prints as f"{x: .6f}" with its zeros trimmed; the
    const double v = bits_as_double(s.bits);   // host also needs repr(x): the bits travel behind the text
    const Fixed6 f = fixed6(v);
    if (f.big || f.inf || dbl_is_nan(v)) n = 0xFFFFFFFFu;
    else n = (uint32_t)fixed_write(f, v, objfloat_dec(f), false, o);
    for (int q = 0; q < 8; ++q) o[240 + q] = (uint8_t)(s.bits >> (8 * q));
  }
  else if (s.mode == CM_DATE) n = pydatetime_write((int64_t)s.bits, o);   // str(label)
  else if (s.mode == CM_TIME) n = pytime_write((int64_t)s.bits, o);
  else if (s.mode == CM_TDELTA) n = pytd_write((int64_t)s.bits, o);
  hdr_meta[2 * j] = s.mode;
  hdr_meta[2 * j + 1] = n;
}

// ---- python-engine classification: XF_* flags per output column, NaN count per row ------------
__global__ void __launch_bounds__(256) xclassify_kernel(TableArgs a) {
  __shared__ uint32_t s_flags[1024];
  __shared__ uint32_t s_doc;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    if (threadIdx.x == 0) {
      uint32_t lo = 0, hi = a.n_docs;
      while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (__ldg(a.item_doc_base + mid) <= item) lo = mid; else hi = mid; }
      s_doc = lo;
    }
    __syncthreads();
    const TDoc D = a.docs[s_doc];
    uint32_t W = D.W;
    uint32_t row0 = (item - D.item_base) * D.rb;
    uint32_t nrows = min(D.rb, D.n_data - row0);
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x) s_flags[c] = 0;
    __syncthreads();
    uint32_t rbp = 32;
    while (rbp < nrows) rbp <<= 1;
    const uint32_t ngrp = blockDim.x / rbp, rr = threadIdx.x % rbp, grp = threadIdx.x / rbp;
    if (rr < nrows && grp < ngrp) {
      const uint32_t row = row0 + rr;
      uint32_t na = 0;
      for (uint32_t c = grp; c < W; c += ngrp) {
        uint32_t f = xclassify_cell(xfetch(a, D, row + 2, c), row == 0);
        if ((s_flags[c] & f) != f) atomicOr(&s_flags[c], f);
        if (f & XF_NULL) ++na;
      }
      if (na) atomicAdd(a.row_na + D.r0 + 1 + row, na);
    }
    __syncthreads();
    for (uint32_t c = threadIdx.x; c < W; c += blockDim.x)
      if (s_flags[c]) atomicOr(a.x_colflags + D.col_base + c, s_flags[c]);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128) xdecide_kernel(TableArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_docs) return;
  const TDoc D = a.docs[d];
  if (!D.active) return;
  int flags = 0;
  for (uint32_t c = 0; c < D.W; ++c) {
    uint32_t f = a.x_colflags[D.col_base + c];
    int t = xdecide_column(f);
    if (t == CT_UNSUP || (f & (XF_BAD_UTF8 | XF_ENTITY))) flags |= 2;
    a.ctype[D.stat_base + c] = (uint8_t)t;
  }
  a.chunk_nf[D.stat_base] = 0;
  if (flags) atomicOr(a.doc_flags + d, flags);
}

__global__ void set_u32_kernel(uint32_t* p, uint32_t v) { *p = v; }

// ---------------------------------------------------------------------------------------------
void xml_count(const XmlArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  for (int q = 0; q < 3; ++q) cudaMemsetAsync(a.tile_cnt[q] + a.n_tiles, 0, 4, s);
  prof_mark("xml_tags_kernel<count>", s);
  if (a.n_tiles) RFQ_LAUNCH(xml_tags_kernel<false>, (a.n_tiles + 7) / 8, 256, 0, s, a);
  prof_mark("scan(xml counts)", s);
  for (int q = 0; q < 3; ++q)
    exclusive_scan<SumU32>(a.tile_cnt[q], a.tile_pre[q], a.n_tiles + 1, a.scan_tmp, nullptr, s);
  RFQ_LAUNCH(xml_shape_kernel, (a.n_docs + 1 + 255) / 256, 256, 0, s, a);
  prof_mark("host+sync", s);
}

void xml_decode(const XmlArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("xml_tags_kernel<index>", s);
  if (a.n_tiles) RFQ_LAUNCH(xml_tags_kernel<true>, (a.n_tiles + 7) / 8, 256, 0, s, a);
  RFQ_LAUNCH(set_u32_kernel, 1, 1, 0, s, a.row_c0 + a.n_rows, a.n_cells);
  prof_mark("sst_kernel", s);
  if (a.n_si) {
    RFQ_LAUNCH(sst_kernel<false>, (a.n_si + 127) / 128, 128, 0, s, a);
    cudaMemsetAsync(a.si_len + a.n_si, 0, 4, s);
    exclusive_scan<SumU32>(a.si_len, a.si_off, a.n_si + 1, a.scan_tmp, nullptr, s);
    RFQ_LAUNCH(sst_kernel<true>, (a.n_si + 127) / 128, 128, 0, s, a);
  }
  prof_mark("xrow_kernel", s);
  if (a.n_rows) {
    RFQ_LAUNCH(xrow_kernel, (a.n_rows + 127) / 128, 128, 0, s, a);
    RFQ_LAUNCH(xrow_fix_kernel, (a.n_rows + 127) / 128, 128, 0, s, a);
    RFQ_LAUNCH(xrow_order_kernel, (a.n_rows + 127) / 128, 128, 0, s, a);
  }
  prof_mark("xcell_kernel", s);
  if (a.n_cells) {
    cudaMemsetAsync(a.num_count, 0, 4, s);
    RFQ_LAUNCH(xcell_kernel, (a.n_cells + 127) / 128, 128, 0, s, a);
    RFQ_LAUNCH(xnum_kernel, 148 * 8, 128, 0, s, a);
    RFQ_LAUNCH(xcell_fix_kernel, (a.n_cells + 127) / 128, 128, 0, s, a);
    RFQ_LAUNCH(xcell_order_kernel, (a.n_cells + 127) / 128, 128, 0, s, a);
  }
  prof_mark("host+sync", s);
}

void xlsx_scatter_rows(const XmlArgs& x, const uint32_t* r2_base, uint32_t* r2elem, cudaStream_t s) {
  if (x.n_rows) RFQ_LAUNCH(xscatter_kernel, (x.n_rows + 127) / 128, 128, 0, s, x, r2_base, r2elem);
}

void xlsx_header_cells(const TableArgs& a, uint8_t* hdr_buf, uint32_t* hdr_meta, cudaStream_t s) {
  if (a.n_cols) RFQ_LAUNCH(xheader_kernel, (a.n_cols + 127) / 128, 128, 0, s, a, hdr_buf, hdr_meta);
}

void xlsx_table_classify(const TableArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  uint32_t g = (uint32_t)sms * 8u;
  prof_mark("xclassify_kernel", s);
  if (a.n_items) RFQ_LAUNCH(xclassify_kernel, a.n_items < g ? a.n_items : g, 256, 0, s, a);
  RFQ_LAUNCH(xdecide_kernel, (a.n_docs + 127) / 128, 128, 0, s, a);
}

}  // namespace rfq
