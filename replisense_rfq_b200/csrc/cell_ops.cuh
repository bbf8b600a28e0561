// The three per-cell operations of the table pipeline — classify, measure, print — as
// host/device functions over a decoded-byte iterator factory, so the CSV kernels, the sheet-XML
// kernels and the CPU test harness run the very same code.
#pragma once
#include "cell_format.cuh"
#include "cell_parse.cuh"
#include "col_decide.cuh"
#include "datetime_fmt.cuh"
#include "xlsx_core.cuh"

namespace rfq {

// A cell source: where the token bytes live and how to decode them.
// Text modes carry a byte span; typed modes (sheet cells already typed by the XML: numbers,
// booleans, empty / error cells) carry their value in `bits`.
// CM_CSVQ: inside of a quoted CSV field that is clean ASCII and whose only quotes are "" pairs
// (span = between the outer quotes, bits = decoded length)
enum { CM_CSV = 0, CM_RAW = 1, CM_INT = 2, CM_FLOAT = 3, CM_BOOL = 4, CM_XML = 5, CM_NULL = 6, CM_CSVQ = 7,
       CM_DATE = 8, CM_TIME = 9,     // sheet cells: ms since 0001-01-01 / ms of the day (datetime_fmt.cuh)
       CM_TDELTA = 10,               // a timedelta-styled number: signed microseconds
       CM_NUM_PENDING = 11 };        // (between two kernels of the sheet path only) a number still to be parsed

// decoded bytes of a CM_CSVQ span: every "" is one "
struct PairIter {
  const uint8_t* p;
  const uint8_t* e;
  RFQ_HD PairIter(const uint8_t* b, uint32_t n) : p(b), e(b + n) {}
  RFQ_HD int next() {
    if (p >= e) return -1;
    int c = *p++;
    if (c == '"') ++p;
    return c;
  }
};
struct CellSrc {
  const uint8_t* p;
  uint32_t n;
  uint8_t mode;     // CM_*
  bool missing;     // short row: pandas pads with an empty token
  bool clean;       // text known to be plain ASCII without control bytes (set by csv_simplify)
  uint64_t bits;    // CM_INT: int64, CM_FLOAT: double bits, CM_BOOL: 0/1
};

// 8 bytes at p (little endian), bytes at index >= rem forced to zero. On the device this is two
// aligned 8-byte loads + a funnel shift (the arenas carry 64 bytes of slack, so reading up to 15
// bytes past p is safe); on the host a bounded byte loop.
RFQ_HD uint64_t load8_masked(const uint8_t* p, uint32_t rem) {
  uint64_t x;
#ifdef __CUDA_ARCH__
  const unsigned long long* q = (const unsigned long long*)((uintptr_t)p & ~(uintptr_t)7);
  uint32_t sh = (uint32_t)((uintptr_t)p & 7u) * 8u;
  x = *q;                                   // generic loads: p may point into a shared-memory stage
  if (sh) x = (x >> sh) | ((uint64_t)q[1] << (64u - sh));
#else
  x = 0;
  for (uint32_t i = 0; i < 8 && i < rem; ++i) x |= (uint64_t)p[i] << (8 * i);
#endif
  if (rem < 8) x &= (1ull << (8 * rem)) - 1ull;
  return x;
}

// true iff every one of the n bytes at p is >= 0x0E, none equals `forbid`, and (unless allow_high)
// all are ASCII
RFQ_HD bool span_is_clean(const uint8_t* p, uint32_t n, uint32_t forbid, bool allow_high = false) {
  const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
  for (uint32_t i = 0; i < n; i += 8) {
    uint32_t rem = n - i;
    uint64_t x = load8_masked(p + i, rem);
    if (rem < 8) x |= 0x2020202020202020ull & ~((1ull << (8 * rem)) - 1ull);  // padding = blanks
    uint64_t bad = allow_high ? 0ull : (x & highs);        // non-ASCII
    bad |= (x - ones * 0x0E) & ~x & highs;                 // a byte < 0x0E (NUL, \t, \n, \r, ...)
    uint64_t y = x ^ (ones * forbid);
    bad |= (y - ones) & ~y & highs;                        // a byte == forbid
    if (bad) return false;
  }
  return true;
}

// Turn a CSV field into a plain span when no quote processing can change it: unquoted clean
// ASCII, or "..." around clean ASCII without inner quotes.  Everything downstream (NA match,
// parsers, widths, copy-out) then runs without the quote state machine.
RFQ_HD void csv_simplify(CellSrc& s) {
  if (s.mode != CM_CSV || s.missing) return;
  if (s.n == 0) { s.mode = CM_RAW; s.clean = true; return; }
  if (s.p[0] != '"') {
    if (span_is_clean(s.p, s.n, 0xFF)) { s.mode = CM_RAW; s.clean = true; }
    // accents, currency signs ...: still a plain span when no quote and no control byte is inside
    else if (span_is_clean(s.p, s.n, '"', true)) s.mode = CM_RAW;
    return;
  }
  if (s.n < 2 || s.p[s.n - 1] != '"') return;
  if (span_is_clean(s.p + 1, s.n - 2, '"')) { s.p += 1; s.n -= 2; s.mode = CM_RAW; s.clean = true; return; }
  if (!span_is_clean(s.p + 1, s.n - 2, 0xFF)) {
    if (span_is_clean(s.p + 1, s.n - 2, '"', true)) { s.p += 1; s.n -= 2; s.mode = CM_RAW; }   // "..." around plain non-ASCII text
    return;
  }
  // clean ASCII with quotes inside: fine iff they all come as "" pairs
  uint32_t pairs = 0;
  const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
  for (uint32_t i = 1; i + 1 < s.n;) {
    uint32_t rem = s.n - 1 - i;
    uint64_t y = load8_masked(s.p + i, rem) ^ (ones * '"');
    uint64_t hit = (y - ones) & ~y & highs;                // exact for the first match
    if (rem < 8) hit &= (1ull << (8 * rem)) - 1ull;
    if (!hit) { i += 8; continue; }
#ifdef __CUDA_ARCH__
    i += (uint32_t)(__ffsll((long long)hit) - 1) >> 3;
#else
    i += (uint32_t)__builtin_ctzll(hit) >> 3;
#endif
    if (i + 2 < s.n && s.p[i + 1] == '"') { ++pairs; i += 2; } else return;
  }
  s.p += 1; s.n -= 2; s.mode = CM_CSVQ; s.bits = s.n - pairs;
}
RFQ_HD bool cell_is_text(const CellSrc& s) { return s.mode == CM_CSV || s.mode == CM_RAW || s.mode == CM_XML || s.mode == CM_CSVQ; }
RFQ_HD double bits_as_double(uint64_t b) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)b);
#else
  double d; __builtin_memcpy(&d, &b, 8); return d;
#endif
}

#define RFQ_WITH_ITER(src, F, ...)                                            \
  ((src).mode == CM_CSV ? F(CsvIter((src).p, (src).n), ##__VA_ARGS__)         \
   : (src).mode == CM_XML ? F(XmlIter((src).p, (src).n), ##__VA_ARGS__)       \
   : (src).mode == CM_CSVQ ? F(PairIter((src).p, (src).n), ##__VA_ARGS__)     \
                          : F(RawIter((src).p, (src).n), ##__VA_ARGS__))

// The parsers are inlined only for plain spans (CM_RAW: clean CSV tokens, shared strings); the
// three decoding iterators share ONE out-of-line copy of each parser per translation unit.
// Arguments and results travel by value so that the caller's CellSrc stays in registers.
#define RFQ_WITH_SLOW_ITER(mode, p, n, F, ...)                         \
  ((mode) == CM_CSV ? F(CsvIter((p), (n)), ##__VA_ARGS__)              \
   : (mode) == CM_XML ? F(XmlIter((p), (n)), ##__VA_ARGS__)            \
                      : F(PairIter((p), (n)), ##__VA_ARGS__))
struct I64Res { int64_t v; int st; };
struct U64Res { UintTok u; int st; };
struct F64Res { double v; int ok; };
RFQ_HD_COLD TextInfo scan_text_slow(const uint8_t* p, uint32_t n, int mode) {
  return RFQ_WITH_SLOW_ITER(mode, p, n, scan_text);
}
RFQ_HD_COLD I64Res parse_int64_slow(const uint8_t* p, uint32_t n, int mode) {
  I64Res r; r.v = 0;
  r.st = RFQ_WITH_SLOW_ITER(mode, p, n, parse_int64, &r.v);
  return r;
}
RFQ_HD_COLD U64Res parse_uint64_slow(const uint8_t* p, uint32_t n, int mode) {
  U64Res r;
  r.st = RFQ_WITH_SLOW_ITER(mode, p, n, parse_uint64, &r.u);
  return r;
}
RFQ_HD_COLD F64Res parse_double_slow(const uint8_t* p, uint32_t n, int mode) {   // 2 = an inf word
  F64Res r; r.v = 0.;
  r.ok = RFQ_WITH_SLOW_ITER(mode, p, n, parse_double, &r.v) ? 1 : 0;
  if (!r.ok) {
    int k = RFQ_WITH_SLOW_ITER(mode, p, n, inf_token);
    if (k) { r.ok = 2; r.v = (double)k; }
  }
  return r;
}
RFQ_HD_COLD PyInt parse_pyint_any(const uint8_t* p, uint32_t n, int mode) {
  return mode == CM_RAW ? parse_pyint(RawIter(p, n)) : RFQ_WITH_SLOW_ITER(mode, p, n, parse_pyint);
}
RFQ_HD_COLD uint32_t write_pyint_any(const uint8_t* p, uint32_t n, int mode, uint8_t* out) {
  return mode == CM_RAW ? write_pyint(RawIter(p, n), out) : RFQ_WITH_SLOW_ITER(mode, p, n, write_pyint, out);
}
RFQ_HD double dbl_inf(int sign) {
#ifdef __CUDA_ARCH__
  return sign > 0 ? __longlong_as_double(0x7FF0000000000000LL) : __longlong_as_double(0xFFF0000000000000LL);
#else
  return sign > 0 ? __builtin_huge_val() : -__builtin_huge_val();
#endif
}

// scan_text for a plain span: lengths and escape counts eight bytes at a time; only windows that
// hold non-ASCII bytes are walked byte by byte for the UTF-8 validation
RFQ_HD uint64_t zero_bytes(uint64_t y) {   // 0x80 in every byte of y that is zero (exact)
  const uint64_t lo7 = 0x7F7F7F7F7F7F7F7Full;
  return ~(((y & lo7) + lo7) | y | lo7);
}
RFQ_HD TextInfo scan_text_raw(const uint8_t* p, uint32_t n) {
  TextInfo t;
  t.nbytes = n; t.ncp = 0; t.nesc = 0; t.first8 = n ? load8_masked(p, n) : 0; t.utf8_ok = true; t.has_fchar = false;
  const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
  uint32_t cont = 0;
  int need = 0;
  uint32_t lo = 0x80, hi = 0xBF;
  for (uint32_t i = 0; i < n; i += 8) {
    const uint32_t rem = n - i, m = rem < 8 ? rem : 8;
    const uint64_t x = i == 0 ? t.first8 : load8_masked(p + i, rem);
    const uint64_t esc = zero_bytes(x ^ (ones * 0x09)) | zero_bytes(x ^ (ones * 0x0A)) | zero_bytes(x ^ (ones * 0x0D));
#ifdef __CUDA_ARCH__
    t.nesc += (uint32_t)__popcll(esc);
#else
    t.nesc += (uint32_t)__builtin_popcountll(esc);
#endif
    if (!(x & highs)) {
      if (need) { t.utf8_ok = false; need = 0; }
      continue;
    }
    const uint64_t cb = x & ~(x << 1) & highs;   // 10xxxxxx
#ifdef __CUDA_ARCH__
    cont += (uint32_t)__popcll(cb);
#else
    cont += (uint32_t)__builtin_popcountll(cb);
#endif
#pragma unroll 1
    for (uint32_t q = 0; q < m; ++q) {   // kept rolled: eight copies of this state machine would be 8 KB of code
      const uint32_t c = (uint32_t)(x >> (8 * q)) & 0xFFu;
      if (need) {
        if (c < lo || c > hi) { t.utf8_ok = false; need = 0; }
        else { --need; lo = 0x80; hi = 0xBF; }
        continue;
      }
      if (c < 0x80u) continue;
      lo = 0x80; hi = 0xBF;
      if (c >= 0xC2u && c <= 0xDFu) need = 1;
      else if (c >= 0xE0u && c <= 0xEFu) { need = 2; if (c == 0xE0u) lo = 0xA0; if (c == 0xEDu) hi = 0x9F; }
      else if (c >= 0xF0u && c <= 0xF4u) { need = 3; if (c == 0xF0u) lo = 0x90; if (c == 0xF4u) hi = 0x8F; }
      else t.utf8_ok = false;
    }
  }
  if (need) t.utf8_ok = false;
  t.ncp = n - cont;
  return t;
}

RFQ_HD TextInfo cell_text(const CellSrc& s) {
  if (s.missing || !cell_is_text(s)) {
    TextInfo t; t.nbytes = t.ncp = t.nesc = 0; t.first8 = 0; t.utf8_ok = true; t.has_fchar = false;
    if (!s.missing && s.mode != CM_NULL) { t.nbytes = 1; t.ncp = 1; t.first8 = '#'; }  // typed value: never an NA token
    return t;
  }
  if (s.clean) {
    TextInfo t; t.nbytes = s.n; t.ncp = s.n; t.nesc = 0; t.utf8_ok = true; t.has_fchar = false;
    t.first8 = s.n ? load8_masked(s.p, s.n) : 0;
    return t;
  }
  if (s.mode == CM_CSVQ) {
    TextInfo t; t.nbytes = (uint32_t)s.bits; t.ncp = t.nbytes; t.nesc = 0; t.utf8_ok = true; t.has_fchar = false;
    t.first8 = 0;
    uint64_t raw = load8_masked(s.p, s.n);
    const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
    uint64_t y = raw ^ (ones * '"');
    if (s.n >= 8 && !((y - ones) & ~y & highs)) { t.first8 = raw; return t; }   // no quote in the first 8
    PairIter it(s.p, s.n);
    for (int k = 0; k < 8; ++k) { int c = it.next(); if (c < 0) break; t.first8 |= (uint64_t)c << (8 * k); }
    return t;
  }
  if (s.mode == CM_RAW) return scan_text_raw(s.p, s.n);
  return scan_text_slow(s.p, s.n, s.mode);
}
RFQ_HD int cell_int64(const CellSrc& s, int64_t* v) {
  if (s.mode == CM_INT) { *v = (int64_t)s.bits; return INT_OK; }
  if (s.mode == CM_BOOL) { *v = (int64_t)(s.bits & 1); return INT_OK; }
  if (!cell_is_text(s)) return INT_INVALID;
  if (s.mode == CM_RAW) return parse_int64(RawIter(s.p, s.n), v);
  I64Res r = parse_int64_slow(s.p, s.n, s.mode);
  *v = r.v;
  return r.st;
}
RFQ_HD int cell_uint64(const CellSrc& s, UintTok* v) {
  if (!cell_is_text(s)) { v->v = 0; v->neg = false; v->gt_i64 = false; return INT_INVALID; }
  if (s.mode == CM_RAW) return parse_uint64(RawIter(s.p, s.n), v);
  U64Res r = parse_uint64_slow(s.p, s.n, s.mode);
  *v = r.u;
  return r.st;
}
RFQ_HD bool cell_double(const CellSrc& s, double* v) {
  if (s.mode == CM_FLOAT) { *v = bits_as_double(s.bits); return true; }
  if (s.mode == CM_INT) {
#ifdef __CUDA_ARCH__
    *v = __ll2double_rn((long long)s.bits);
#else
    *v = (double)(int64_t)s.bits;
#endif
    return true;
  }
  if (s.mode == CM_BOOL) { *v = (s.bits & 1) ? 1.0 : 0.0; return true; }
  if (!cell_is_text(s)) return false;
  if (s.mode != CM_RAW) {
    F64Res r = parse_double_slow(s.p, s.n, s.mode);
    if (!r.ok) return false;
    *v = r.ok == 2 ? dbl_inf(r.v > 0. ? 1 : -1) : r.v;
    return true;
  }
  if (parse_double(RawIter(s.p, s.n), v)) return true;
  {   // the inf words start with [+-]?[iI]
    uint64_t w = load8_masked(s.p, s.n);
    uint32_t b0 = (uint32_t)(w & 0xFFu), b1 = (uint32_t)((w >> 8) & 0xFFu);
    if ((((b0 == '+' || b0 == '-') ? b1 : b0) | 0x20u) != 'i') return false;
  }
  int k = inf_token(RawIter(s.p, s.n));
  if (k == 0) return false;
  *v = dbl_inf(k);
  return true;
}

// True/False of a boolean cell (typed, or one of pandas' bool words)
RFQ_HD int cell_bool(const CellSrc& s, const TextInfo& t) {
  if (s.mode == CM_BOOL) return (int)(s.bits & 1);
  return bool_token(t.nbytes, t.first8);
}
// does the cell print '' (NaN after fillna) under chunk type `ctype`?
RFQ_HD bool cell_is_nan(const CellSrc& s, const TextInfo& t, int ctype) {
  if (s.missing || s.mode == CM_NULL) return true;
  if (s.mode == CM_FLOAT) return dbl_is_nan(bits_as_double(s.bits));
  if (!cell_is_text(s)) return false;
  return ctype != CT_STRNF && is_na_token(t.nbytes, t.first8);
}
// chunk type a cell prints as inside a mixed object column of a sheet (CT_XOBJ)
RFQ_HD int cell_own_ctype(const CellSrc& s) {
  switch (s.mode) {
    case CM_INT: return CT_INT;
    case CM_FLOAT: return CT_FLOAT;
    case CM_BOOL: return CT_BOOL;
    case CM_DATE: return CT_DATE;
    case CM_TIME: return CT_TIME;
    case CM_TDELTA: return CT_TDELTA;
    default: return CT_STR;
  }
}

// ---- 1. classify: SF_* flags of this cell + int64 status -----------------------------------
RFQ_HD uint32_t classify_cell(const CellSrc& s, int* int_status) {
  TextInfo t = cell_text(s);
  *int_status = INT_OK;
  if (is_na_token(t.nbytes, t.first8)) return SF_ANY_NA;
  uint32_t f = SF_ANY_NONNA;
  if (!t.utf8_ok) f |= SF_BAD_UTF8;
  // Every numeric grammar here (str_to_int64 / str_to_uint64 / precise_xstrtod / the inf words /
  // int()) starts with white space, a sign, a digit, '.' or 'i': any other first byte is a plain
  // word, decided without running the parsers.
  const uint32_t c0 = (uint32_t)(t.first8 & 0xFFu);
  if (cell_is_text(s) && !(is_digit_ascii((int)c0) || c0 == '+' || c0 == '-' || c0 == '.' ||
                           is_space_ascii((int)c0) || (c0 | 0x20u) == 'i')) {
    *int_status = INT_INVALID;
    f |= SF_UINT_INVALID | SF_FLOAT_BAD | SF_PYINT_BAD;
    if (bool_token(t.nbytes, t.first8) < 0) f |= SF_BOOL_BAD;
    return f;
  }
  int64_t iv;
  int st = cell_int64(s, &iv);
  *int_status = st;
  if (st == INT_OK) {
    // still needed for the uint64 pass of a column that overflows elsewhere
    if (iv < 0) f |= SF_NEG;
    else if (iv == 0) {  // "-0" is negative for str_to_uint64
      UintTok u; cell_uint64(s, &u);
      if (u.neg) f |= SF_NEG;
    }
    f |= SF_BOOL_BAD;
    return f;
  }
  UintTok u;
  int us;
  if (st == INT_INVALID && cell_is_text(s) && (is_digit_ascii((int)c0) || c0 == '+' || c0 == '.')) {
    // no minus sign and int64 stopped on a bad character before any overflow: uint64 stops there too
    us = INT_INVALID; u.v = 0; u.neg = false; u.gt_i64 = false;
  } else {
    us = cell_uint64(s, &u);
  }
  if (us == INT_INVALID) f |= SF_UINT_INVALID;
  else if (us == INT_OVERFLOW) f |= SF_UINT_OVF;
  else { if (u.neg) f |= SF_NEG; if (u.gt_i64) f |= SF_GT_I64; }
  double dv;
  if (!cell_double(s, &dv)) f |= SF_FLOAT_BAD;
  if (bool_token(t.nbytes, t.first8) < 0) f |= SF_BOOL_BAD;
  // Python-int fallback (only matters for columns that overflow uint64): a plain non-negative
  // uint64 token is int()-parsable; a token that cannot even start an integer is not
  if (!(us == INT_OK && !u.neg)) {
    int pp = pyint_prefix(t.first8, t.nbytes);   // dates, decimals, words: decided in registers
    if (pp == 0 || (pp == 2 && !parse_pyint_any(s.p, s.n, s.mode).ok)) f |= SF_PYINT_BAD;
  }
  return f;
}

// ---- value of a numeric cell under its chunk type ---------------------------------------------
RFQ_HD double cell_as_double(const CellSrc& s, int ctype) {
  double v = 0.;
  if (ctype == CT_INT || ctype == CT_INTNA) {
    int64_t iv = 0; cell_int64(s, &iv);
#ifdef __CUDA_ARCH__
    v = __ll2double_rn(iv);
#else
    v = (double)iv;
#endif
  } else {
    cell_double(s, &v);
  }
  return v;
}

RFQ_HD int sci_len_of(double v) {
  // len of "%.6e": exact exponent digit count needs sci7 only next to the 1e100 / 1e-99 borders
  if (dbl_is_inf(v)) return v < 0 ? 4 : 3;
  bool neg = (dbl_bits(v) >> 63) != 0;
  double a = v < 0 ? -v : v;
  if (a == 0.) return neg + 12;
  if (a >= 1e-98 && a < 9e99) return neg + 12;
  if (a >= 1.1e100 || a < 9e-100) return neg + 13;
  return sci_len(sci7(v));
}

// ---- 2. measure ---------------------------------------------------------------------------------
struct CellMeasure {
  bool is_nan;          // prints '' after fillna
  uint32_t plain_len, ip_len, tz, inf_len, obj_len, sci_len, fl;
  uint32_t extra_bytes; // printed bytes - printed code points (non-ASCII strings only)
};

RFQ_HD CellMeasure measure_cell(const CellSrc& s, int ctype) {
  CellMeasure m;
  m.is_nan = false; m.plain_len = 0; m.ip_len = 0; m.tz = 6; m.inf_len = 0; m.obj_len = 0;
  m.sci_len = 0; m.fl = 0; m.extra_bytes = 0;
  TextInfo t = cell_text(s);
  const bool xobj = ctype == CT_XOBJ;
  if (xobj) ctype = cell_own_ctype(s);
  if (cell_is_nan(s, t, ctype)) { m.is_nan = true; return m; }
  if (ctype == CT_STR || ctype == CT_STRNF) {
    m.plain_len = t.ncp + t.nesc;
    m.extra_bytes = t.nbytes - t.ncp;
    return m;
  }
  if (ctype == CT_BOOL || ctype == CT_BOOLNA) {
    m.plain_len = cell_bool(s, t) == 1 ? 4 : 5;
    return m;
  }
  if (ctype == CT_TDELTA) {   // ip_len / obj_len of a float are free here: the two timedelta64 forms
    const int64_t us = (int64_t)s.bits;
    m.plain_len = pytd_len(us);
    if (!xobj) { m.ip_len = td64_len(us, 0); m.obj_len = td64_len(us, 1); }   // (a mixed object column prints str(td))
    m.fl = MF_FINITE | ((us % RFQ_US_PER_DAY) ? MF_DT_TIME : 0);
    return m;
  }
  if (ctype == CT_DATE || ctype == CT_TIME) {
    int64_t ms = (int64_t)s.bits;
    if (ctype == CT_TIME) { m.plain_len = pytime_len(ms); return m; }
    m.plain_len = pydatetime_len(ms);
    m.fl = MF_FINITE | ((ms % RFQ_MS_PER_DAY) ? MF_DT_TIME : 0) | ((ms % 1000) ? MF_DT_MS : 0);
    return m;
  }
  if (ctype == CT_UINT) {
    UintTok u; cell_uint64(s, &u);
    m.plain_len = (uint32_t)ndigits_u64(u.v);
    return m;
  }
  if (ctype == CT_PYINT) {
    PyInt pi = parse_pyint_any(s.p, s.n, s.mode);
    m.plain_len = pi.ndig + (pi.neg ? 1u : 0u);
    return m;
  }
  if (ctype == CT_INT || ctype == CT_INTNA) {
    int64_t iv = 0; cell_int64(s, &iv);
    m.plain_len = (uint32_t)int64_len(iv);
  }
  double v = cell_as_double(s, ctype);
  Fixed6 f = fixed6(v);
  double a = v < 0 ? -v : v;
  if (f.inf) {
    m.inf_len = 3 + f.neg; m.obj_len = 4; m.sci_len = m.inf_len; m.fl = MF_LARGE;
  } else {
    m.ip_len = f.neg + f.ip_digits; m.tz = f.tz; m.obj_len = objfloat_len(f);
    m.fl = MF_FINITE | (a > 1e6 ? MF_LARGE : 0) | ((a < 1e-6 && a > 0.) ? MF_SMALL : 0);
    m.sci_len = (uint32_t)sci_len_of(v);
  }
  if (xobj && ctype == CT_INT) m.obj_len = 0;  // an int inside a mixed object column prints as int
  return m;
}

// ---- 1 + 2 in one visit: the SF_* flags of a CSV cell AND what it would measure under every
// chunk type it can end up in.  The column's type is only known after all its cells were seen, so
// the candidates are reduced per (chunk, column) side by side and `finish_measure` picks the ones
// the decided type uses.  Identical, value for value, to classify_cell + measure_cell (the
// sequential model runs both ways in tests/test_hostcore.py); saves the second
// tokenise-and-parse sweep.
struct CellCand {
  uint32_t flags;       // SF_*
  int int_status;
  bool na;              // an NA token (NA filter on)
  bool has_float;       // float view valid: the cell is an int64 or parses as a double
  bool need_exact;      // its float view depends on which parser the column ends up using
  uint32_t str_len, extra;   // as text: printed code points, bytes - code points
  uint32_t int_len, uint_len, bool_len;   // as int64 / uint64 / bool (0 = not such a token)
  uint32_t ip_len, tz, inf_len, obj_len, sci_len, fl;   // as float64 (ChunkMeasure semantics)
};

RFQ_HD void cand_float_view(double v, CellCand* m) {
  Fixed6 f = fixed6(v);
  double a = v < 0 ? -v : v;
  m->has_float = true;
  if (f.inf) {
    m->inf_len = 3 + f.neg; m->obj_len = 4; m->sci_len = m->inf_len; m->fl = MF_LARGE;
  } else {
    m->ip_len = f.neg + f.ip_digits; m->tz = f.tz; m->obj_len = objfloat_len(f);
    m->fl = MF_FINITE | (a > 1e6 ? MF_LARGE : 0) | ((a < 1e-6 && a > 0.) ? MF_SMALL : 0);
    m->sci_len = (uint32_t)sci_len_of(v);
  }
}

RFQ_HD CellCand classify_measure_cell(const CellSrc& s) {
  CellCand m;
  m.flags = 0; m.int_status = INT_OK; m.na = false; m.has_float = false; m.need_exact = false;
  m.str_len = 0; m.extra = 0; m.int_len = 0; m.uint_len = 0; m.bool_len = 0;
  m.ip_len = 0; m.tz = 6; m.inf_len = 0; m.obj_len = 0; m.sci_len = 0; m.fl = 0;
  TextInfo t = cell_text(s);
  if (is_na_token(t.nbytes, t.first8)) { m.flags = SF_ANY_NA; m.na = true; return m; }
  uint32_t f = SF_ANY_NONNA;
  if (!t.utf8_ok) f |= SF_BAD_UTF8;
  m.str_len = t.ncp + t.nesc;
  m.extra = t.nbytes - t.ncp;
  const int bt = bool_token(t.nbytes, t.first8);
  if (bt >= 0) m.bool_len = bt == 1 ? 4u : 5u;
  const uint32_t c0 = (uint32_t)(t.first8 & 0xFFu);
  if (!(is_digit_ascii((int)c0) || c0 == '+' || c0 == '-' || c0 == '.' || is_space_ascii((int)c0) ||
        (c0 | 0x20u) == 'i')) {          // a plain word (see classify_cell)
    m.int_status = INT_INVALID;
    f |= SF_UINT_INVALID | SF_FLOAT_BAD | SF_PYINT_BAD;
    if (bt < 0) f |= SF_BOOL_BAD;
    m.flags = f;
    return m;
  }
  int64_t iv;
  int st = cell_int64(s, &iv);
  m.int_status = st;
  if (st == INT_OK) {
    if (iv < 0) f |= SF_NEG;
    else if (iv == 0) {   // "-0": int 0, but -0.0 for xstrtod
      UintTok u; cell_uint64(s, &u);
      if (u.neg) { f |= SF_NEG; m.need_exact = true; }
    }
    f |= SF_BOOL_BAD;
    m.flags = f;
    m.int_len = (uint32_t)int64_len(iv);
    if (iv >= 0) m.uint_len = m.int_len;     // a uint64 column prints its small members the same way
    uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
    if (mag > (1ull << 53)) {
      m.need_exact = true;   // xstrtod(token) may differ from (double)int64
#ifdef __CUDA_ARCH__
      cand_float_view(__ll2double_rn(iv), &m);
#else
      cand_float_view((double)iv, &m);
#endif
      return m;
    }
    // an exactly representable integer seen as float64: "%.6f" is its digits + ".000000", the
    // exponent of "%.6e" has two digits (what fixed6 / sci_len_of compute, without the 128-bit math)
    m.has_float = true;
    m.ip_len = m.int_len; m.tz = 6; m.obj_len = m.int_len - (iv < 0 ? 1u : 0u) + 3u;
    m.sci_len = 12u + (iv < 0 ? 1u : 0u);
    m.fl = MF_FINITE | (mag > 1000000ull ? MF_LARGE : 0);
    return m;
  }
  UintTok u;
  int us;
  if (st == INT_INVALID && (is_digit_ascii((int)c0) || c0 == '+' || c0 == '.')) {
    us = INT_INVALID; u.v = 0; u.neg = false; u.gt_i64 = false;
  } else {
    us = cell_uint64(s, &u);
  }
  if (us == INT_INVALID) f |= SF_UINT_INVALID;
  else if (us == INT_OVERFLOW) f |= SF_UINT_OVF;
  else {
    if (u.neg) f |= SF_NEG;
    if (u.gt_i64) f |= SF_GT_I64;
    if (!u.neg) m.uint_len = (uint32_t)ndigits_u64(u.v);
  }
  double dv;
  if (!cell_double(s, &dv)) f |= SF_FLOAT_BAD;
  else cand_float_view(dv, &m);
  if (bt < 0) f |= SF_BOOL_BAD;
  if (!(us == INT_OK && !u.neg)) {
    int pp = pyint_prefix(t.first8, t.nbytes);
    if (pp == 0 || (pp == 2 && !parse_pyint_any(s.p, s.n, s.mode).ok)) f |= SF_PYINT_BAD;
  }
  m.flags = f;
  return m;
}

// ---- 3. print -----------------------------------------------------------------------------------
// How a cell prints given the class of its column and the type of its chunk.
enum { PR_EMPTY = 0, PR_INT, PR_UINT, PR_FIXED, PR_SCI, PR_OBJFLOAT, PR_BOOL, PR_STR, PR_PYINT,
       PR_DATE64, PR_DATETIME, PR_TIME,     // datetime64 column (dec = resolution) / str(datetime) / str(time)
       PR_TD64, PR_PYTD };                  // timedelta64 column (dec = 0 whole days, 1 long form) / str(timedelta)
RFQ_HD bool pr_prints_nat(int kind) { return kind == PR_DATE64 || kind == PR_TD64; }

RFQ_HD int print_kind(int cls, int ctype, bool sci) {
  switch (cls) {
    case CC_INT: return PR_INT;
    case CC_UINT: return PR_UINT;
    case CC_FLOAT: return sci ? PR_SCI : PR_FIXED;
    case CC_BOOL: return PR_BOOL;
    case CC_STR: return PR_STR;
    case CC_OBJF: return PR_OBJFLOAT;
    case CC_DATE: return PR_DATE64;
    case CC_TDELTA: return PR_TD64;
    default:
      switch (ctype) {
        case CT_INT: return PR_INT;
        case CT_UINT: return PR_UINT;
        case CT_INTNA: case CT_FLOAT: return PR_OBJFLOAT;
        case CT_BOOL: case CT_BOOLNA: return PR_BOOL;
        case CT_PYINT: return PR_PYINT;
        case CT_DATE: return PR_DATETIME;
        case CT_TIME: return PR_TIME;
        case CT_TDELTA: return PR_PYTD;
        default: return PR_STR;
      }
  }
}

// Printed length in code points and bytes (for right-justification), without writing.
RFQ_HD void printed_size(const CellSrc& s, int ctype, int kind, int dec, uint32_t* ncp,
                         uint32_t* nbytes) {
  *ncp = 0; *nbytes = 0;
  TextInfo t = cell_text(s);
  if (cell_is_nan(s, t, ctype)) { if (pr_prints_nat(kind)) { *ncp = 3; *nbytes = 3; } return; }   // NaT
  uint32_t n = 0;
  switch (kind) {
    case PR_DATE64: n = date64_len(dec); break;
    case PR_DATETIME: n = pydatetime_len((int64_t)s.bits); break;
    case PR_TIME: n = pytime_len((int64_t)s.bits); break;
    case PR_TD64: n = td64_len((int64_t)s.bits, dec); break;
    case PR_PYTD: n = pytd_len((int64_t)s.bits); break;
    case PR_STR: *ncp = t.ncp + t.nesc; *nbytes = t.nbytes + t.nesc; return;
    case PR_BOOL: n = cell_bool(s, t) == 1 ? 4 : 5; break;
    case PR_INT: { int64_t iv = 0; cell_int64(s, &iv); n = (uint32_t)int64_len(iv); break; }
    case PR_UINT: { UintTok u; cell_uint64(s, &u); n = (uint32_t)ndigits_u64(u.v); break; }
    case PR_PYINT: { PyInt pi = parse_pyint_any(s.p, s.n, s.mode); n = pi.ndig + (pi.neg ? 1u : 0u); break; }
    case PR_FIXED: { Fixed6 f = fixed6(cell_as_double(s, ctype)); n = (uint32_t)fixed_len(f, dec); break; }
    case PR_OBJFLOAT: { Fixed6 f = fixed6(cell_as_double(s, ctype)); n = (uint32_t)objfloat_len(f); break; }
    case PR_SCI: n = (uint32_t)sci_len_of(cell_as_double(s, ctype)); break;
    default: break;
  }
  *ncp = n; *nbytes = n;
}

// Write the printed form at `out` (exactly the bytes printed_size reported). Returns bytes.
RFQ_HD uint32_t print_cell(const CellSrc& s, int ctype, int kind, int dec, uint8_t* out) {
  if (kind == PR_STR) {
    TextInfo t = cell_text(s);
    if (cell_is_nan(s, t, ctype)) return 0;
    uint32_t k = 0;
    int c;
    if (s.clean) {
      for (uint32_t i = 0; i < s.n; i += 8) {
        uint64_t x = load8_masked(s.p + i, s.n - i);
        uint32_t m = s.n - i < 8 ? s.n - i : 8;
        for (uint32_t q = 0; q < m; ++q) out[i + q] = (uint8_t)(x >> (8 * q));
      }
      return s.n;
    }
    if (s.mode == CM_CSVQ) {
      PairIter it(s.p, s.n);
      while ((c = it.next()) >= 0) out[k++] = (uint8_t)c;
      return k;
    }
    if (s.mode == CM_XML) {
      XmlIter it(s.p, s.n);
      while ((c = it.next()) >= 0) {
        if (c == '\t') { out[k++] = '\\'; out[k++] = 't'; }
        else if (c == '\n') { out[k++] = '\\'; out[k++] = 'n'; }
        else if (c == '\r') { out[k++] = '\\'; out[k++] = 'r'; }
        else out[k++] = (uint8_t)c;
      }
    } else if (s.mode == CM_CSV) {
      CsvIter it(s.p, s.n);
      while ((c = it.next()) >= 0) {
        if (c == '\t') { out[k++] = '\\'; out[k++] = 't'; }
        else if (c == '\n') { out[k++] = '\\'; out[k++] = 'n'; }
        else if (c == '\r') { out[k++] = '\\'; out[k++] = 'r'; }
        else out[k++] = (uint8_t)c;
      }
    } else {
      RawIter it(s.p, s.n);
      while ((c = it.next()) >= 0) {
        if (c == '\t') { out[k++] = '\\'; out[k++] = 't'; }
        else if (c == '\n') { out[k++] = '\\'; out[k++] = 'n'; }
        else if (c == '\r') { out[k++] = '\\'; out[k++] = 'r'; }
        else out[k++] = (uint8_t)c;
      }
    }
    return k;
  }
  TextInfo t = cell_text(s);
  if (cell_is_nan(s, t, ctype)) {
    if (!pr_prints_nat(kind)) return 0;
    out[0] = 'N'; out[1] = 'a'; out[2] = 'T';
    return 3;
  }
  switch (kind) {
    case PR_DATE64: return date64_write((int64_t)s.bits, dec, out);
    case PR_DATETIME: return pydatetime_write((int64_t)s.bits, out);
    case PR_TIME: return pytime_write((int64_t)s.bits, out);
    case PR_TD64: return td64_write((int64_t)s.bits, dec, out);
    case PR_PYTD: return pytd_write((int64_t)s.bits, out);
    case PR_BOOL: {
      const char* w = cell_bool(s, t) == 1 ? "True" : "False";
      uint32_t k = 0;
      while (w[k]) { out[k] = (uint8_t)w[k]; ++k; }
      return k;
    }
    case PR_INT: {
      int64_t iv = 0; cell_int64(s, &iv);
      uint32_t n = (uint32_t)int64_len(iv);
      uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
      write_u64_rev(mag, out + n);
      if (iv < 0) out[0] = '-';
      return n;
    }
    case PR_UINT: {
      UintTok u; cell_uint64(s, &u);
      uint32_t n = (uint32_t)ndigits_u64(u.v);
      write_u64_rev(u.v, out + n);
      return n;
    }
    case PR_PYINT: return write_pyint_any(s.p, s.n, s.mode, out);
    case PR_FIXED: {
      double v = cell_as_double(s, ctype);
      Fixed6 f = fixed6(v);
      return (uint32_t)fixed_write(f, v, dec, false, out);
    }
    case PR_OBJFLOAT: {
      double v = cell_as_double(s, ctype);
      Fixed6 f = fixed6(v);
      return (uint32_t)fixed_write(f, v, objfloat_dec(f), true, out);
    }
    case PR_SCI: {
      double v = cell_as_double(s, ctype);
      return (uint32_t)sci_write(sci7(v), out);
    }
    default: return 0;
  }
}

}  // namespace rfq
