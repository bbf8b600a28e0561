// Sweep 3 of 3 (emit).  Cell typing + numeric parse, column decisions, and the fixed-width formatter / emitter
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

// ---- sweep 3: emit -------------------------------------------------------------------------------
// A work item's rows are one contiguous byte range of the output.  The CTA builds that range in
// shared memory (pre-filled with the padding blanks, so a cell only writes its own characters),
// then streams it out with aligned 16-byte stores.  Ranges that do not fit (very long cells)
// take the direct path: same formatting, byte stores to global memory.
#define RFQ_EMIT_SMEM (80u << 10)   // 2 CTAs per SM (3 per SM with 72 KB and an 80-register cap measured slower: 0.70 vs 0.49 ms)

// The emit kernel keeps only the common cells inline: absent cells, typed sheet values, clean ASCII
// tokens (parsed with the plain byte iterator) and ""-pair text.  Everything else — quoted text
// that needs the state machine, XML entities, escapes, non-ASCII, huge fixed-notation values,
// Python-int columns, the shift recomputation of split rows — goes through the general
// formatters out of line.  Inlining all of it made cicc spend 8 minutes on this kernel and bought
// nothing at run time.
__device__ __noinline__ uint32_t print_cell_cold(CellSrc s, int ct, int kind, int dec, uint8_t* out) {
  return print_cell(s, ct, kind, dec, out);
}
struct SizeRes { uint32_t ncp, nbytes; };
__device__ __noinline__ SizeRes printed_size_cold(CellSrc s, int ct, int kind, int dec) {
  SizeRes r;
  printed_size(s, ct, kind, dec, &r.ncp, &r.nbytes);
  return r;
}
__device__ __forceinline__ void put_nat(int kind, uint32_t width, uint8_t* field) {
  if (pr_prints_nat(kind)) { field[width - 3] = 'N'; field[width - 2] = 'a'; field[width - 1] = 'T'; }
}
// any cell, any kind; returns printed bytes - printed code points
// (arguments by value: a reference would pin the caller's copy to local memory)
__device__ __noinline__ uint32_t emit_cell_cold(CellSrc src, int ct, int kind, int dec,
                                                uint32_t width, uint8_t* field, bool fill) {
  TextInfo t = cell_text(src);
  if (cell_is_nan(src, t, ct)) {
    if (fill) for (uint32_t q = 0; q < width; ++q) field[q] = ' ';
    put_nat(kind, width, field);
    return 0;
  }
  uint32_t ncp, nb;
  if (kind == PR_STR) { ncp = t.ncp + t.nesc; nb = t.nbytes + t.nesc; }   // the scan above already measured it
  else printed_size(src, ct, kind, dec, &ncp, &nb);
  uint32_t pad = width - ncp;
  if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
  print_cell(src, ct, kind, dec, field + pad);
  return nb - ncp;
}

// values of typed cells and clean tokens (RawIter only)
__device__ __forceinline__ int64_t fast_i64(const CellSrc& s) {
  if (s.mode == CM_INT) return (int64_t)s.bits;
  if (s.mode == CM_BOOL) return (int64_t)(s.bits & 1);
  int64_t v = 0;
  parse_int64(RawIter(s.p, s.n), &v);
  return v;
}
__device__ __forceinline__ double fast_f64(const CellSrc& s, int ct) {
  if (ct == CT_INT || ct == CT_INTNA) return __ll2double_rn(fast_i64(s));
  if (s.mode == CM_FLOAT) return bits_as_double(s.bits);
  if (s.mode == CM_INT) return __ll2double_rn((long long)s.bits);
  if (s.mode == CM_BOOL) return (s.bits & 1) ? 1.0 : 0.0;
  double v = 0.;
  if (parse_double(RawIter(s.p, s.n), &v)) return v;
  int k = inf_token(RawIter(s.p, s.n));
  return k > 0 ? __longlong_as_double(0x7FF0000000000000LL)
               : k < 0 ? __longlong_as_double(0xFFF0000000000000LL) : 0.;
}

// number -> text in a small local buffer; returns length, or -1 when the value needs the
// general path (|v| >= 2^63 printed in fixed notation, Python ints)
__device__ __forceinline__ int format_number(const CellSrc& s, int ct, int kind, int dec,
                                             uint32_t nbytes, uint64_t first8, uint8_t* buf) {
  switch (kind) {
    case PR_BOOL: {
      int bv = s.mode == CM_BOOL ? (int)(s.bits & 1) : bool_token(nbytes, first8);
      const char* w = bv == 1 ? "True" : "False";
      int k = 0;
      while (w[k]) { buf[k] = (uint8_t)w[k]; ++k; }
      return k;
    }
    case PR_INT: {
      int64_t iv = fast_i64(s);
      int n = int64_len(iv);
      uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
      write_u64_rev(mag, buf + n);
      if (iv < 0) buf[0] = '-';
      return n;
    }
    case PR_UINT: {
      UintTok u; u.v = 0;
      if (cell_is_text(s)) parse_uint64(RawIter(s.p, s.n), &u);
      int n = ndigits_u64(u.v);
      write_u64_rev(u.v, buf + n);
      return n;
    }
    case PR_FIXED: case PR_OBJFLOAT: {
      if (s.mode == CM_RAW && s.clean) {   // a short decimal literal prints its own digits (cell_format.cuh)
        Fixed6 d6;
        if (decimal_literal(s.p, s.n, &d6))
          return fixed_write(d6, 0., kind == PR_FIXED ? dec : objfloat_dec(d6), kind == PR_OBJFLOAT, buf);
      }
      double v = fast_f64(s, ct);
      Fixed6 f = fixed6(v);
      if (f.big) return -1;
      return fixed_write(f, v, kind == PR_FIXED ? dec : objfloat_dec(f), kind == PR_OBJFLOAT, buf);
    }
    case PR_SCI: return sci_write(sci7(fast_f64(s, ct)), buf);
    case PR_DATE64: return (int)date64_write((int64_t)s.bits, dec, buf);
    case PR_DATETIME: return (int)pydatetime_write((int64_t)s.bits, buf);
    case PR_TIME: return (int)pytime_write((int64_t)s.bits, buf);
    case PR_TD64: return (int)td64_write((int64_t)s.bits, dec, buf);
    case PR_PYTD: return (int)pytd_write((int64_t)s.bits, buf);
    default: return -1;   // PR_PYINT: arbitrary length, measured and written in place
  }
}

// writes the cell right-justified into its field; `field` points at the first byte of the
// column's field inside the row; `fill` = also write the padding blanks (direct path)
// returns printed bytes - printed code points (non-zero only for non-ASCII strings)
__device__ __forceinline__ uint32_t emit_cell(const CellSrc& src, int ct, int kind, int dec,
                                              uint32_t width, uint8_t* field, bool fill) {
  const bool absent = src.missing || src.mode == CM_NULL;
  const bool text = cell_is_text(src);
  // a plain span that is not clean ASCII (a shared string with accents, say) printed as text: count
  // its continuation bytes and look for \t \n \r eight bytes at a time; without those three it is
  // copied as it stands (UTF-8 validity was settled by the typing sweep)
  uint32_t cont = 0;
  bool raw_text = false;
  if (!absent && src.mode == CM_RAW && !src.clean && kind == PR_STR) {
    const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
    uint64_t esc = 0;
    for (uint32_t i = 0; i < src.n; i += 8) {
      uint64_t x = load8_masked(src.p + i, src.n - i);
      cont += (uint32_t)__popcll(x & ~(x << 1) & highs);              // 10xxxxxx
      uint64_t a = x ^ (ones * 0x09), b = x ^ (ones * 0x0A), c = x ^ (ones * 0x0D);
      uint64_t z = ((a - ones) & ~a) | ((b - ones) & ~b) | ((c - ones) & ~c);
      if (src.n - i < 8) z &= (1ull << (8 * (src.n - i))) - 1ull;      // the zero padding is not a tab
      esc |= z & highs;
    }
    raw_text = esc == 0;
  }
  if (!absent && text && !src.clean && !raw_text && !(src.mode == CM_CSVQ && kind == PR_STR))
    return emit_cell_cold(src, ct, kind, dec, width, field, fill);
  // what is left: absent, typed value, clean token, plain or ""-pair text printed as text
  uint32_t nbytes = 0;
  uint64_t first8 = 0;
  bool na = absent;
  if (!absent && text) {
    if (src.clean || raw_text) { nbytes = src.n; first8 = src.n ? load8_masked(src.p, src.n) : 0; }
    else {
      nbytes = (uint32_t)src.bits;
      if (nbytes <= 8) {   // only short tokens can be NA words
        PairIter it(src.p, src.n);
        for (int k = 0; k < 8; ++k) { int c = it.next(); if (c < 0) break; first8 |= (uint64_t)c << (8 * k); }
      } else first8 = 0x2020202020202020ull;
    }
    na = ct != CT_STRNF && is_na_token(nbytes, first8);
  } else if (!absent && src.mode == CM_FLOAT) {
    na = dbl_is_nan(bits_as_double(src.bits));
  }
  if (na) {
    if (fill) for (uint32_t q = 0; q < width; ++q) field[q] = ' ';
    put_nat(kind, width, field);
    return 0;
  }
  if (kind == PR_STR && text) {   // nothing to escape; code points = bytes - continuation bytes
    uint32_t pad = width - (nbytes - cont);
    if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
    uint8_t* o = field + pad;
    if (src.clean || raw_text) {
      uint32_t i = 0;
      for (; i + 8 <= src.n; i += 8) {          // whole windows: eight stores, no inner loop
        const uint64_t x = load8_masked(src.p + i, 8);
        const uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
        uint8_t* d = o + i;
        d[0] = (uint8_t)lo; d[1] = (uint8_t)(lo >> 8); d[2] = (uint8_t)(lo >> 16); d[3] = (uint8_t)(lo >> 24);
        d[4] = (uint8_t)hi; d[5] = (uint8_t)(hi >> 8); d[6] = (uint8_t)(hi >> 16); d[7] = (uint8_t)(hi >> 24);
      }
      if (i < src.n) {
        const uint64_t x = load8_masked(src.p + i, src.n - i);
        const uint32_t m = src.n - i;
        for (uint32_t q = 0; q < m; ++q) o[i + q] = (uint8_t)(x >> (8 * q));
      }
    } else {
      PairIter it(src.p, src.n);
      int c;
      while ((c = it.next()) >= 0) *o++ = (uint8_t)c;
    }
    return cont;
  }
  uint8_t buf[40];
  int n = (kind == PR_STR) ? -1 : format_number(src, ct, kind, dec, nbytes, first8, buf);
  if (n < 0) {  // rare: huge value in fixed notation, Python int, typed value printed as text
    SizeRes sz = printed_size_cold(src, ct, kind, dec);
    uint32_t pad = width - sz.ncp;
    if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
    print_cell_cold(src, ct, kind, dec, field + pad);
    return sz.nbytes - sz.ncp;
  }
  uint32_t pad = width - (uint32_t)n;
  if (fill) for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
  for (int q = 0; q < n; ++q) field[pad + q] = buf[q];
  return 0;
}

// bytes - code points of the text cells left of column c in this row (rows with non-ASCII text
// whose columns are split over thread groups)
__device__ __noinline__ uint32_t row_shift_before(TableArgs a, uint32_t d, RowCur rc,
                                                  const uint8_t* ctrow, const ColOut* col, uint32_t c) {
  const TDoc D = a.docs[d];
  uint32_t xb = 0;
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
  return xb;
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
          CellSrc src = row_cell_tag_read(a, D, rc, c);
          if (ct == CT_XOBJ) ct = cell_own_ctype(src);
          int kind = print_kind(co.cls, ct, co.sci != 0);
          if (ngrp > 1 && j && extra)   // columns are split over groups: recompute the shift
            xb = row_shift_before(a, ic.d, rc, ctrow, col, c);
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

void csv_table_emit_header(const TableArgs& a, uint8_t* out, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("csv_header_emit_kernel", s);
  RFQ_LAUNCH(csv_header_emit_kernel, (a.n_docs + 7) / 8, 256, 0, s, a, out);
}

void csv_table_emit_items(const TableArgs& a0, uint8_t* out, uint32_t lo, uint32_t hi, cudaStream_t s) {
  if (hi <= lo) return;
  // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute
  static std::atomic<unsigned long long> attr_done{0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_done.load() >> (dev & 63) & 1ull)) {
    cudaFuncSetAttribute(csv_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RFQ_EMIT_SMEM);
    attr_done.fetch_or(1ull << (dev & 63));
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
