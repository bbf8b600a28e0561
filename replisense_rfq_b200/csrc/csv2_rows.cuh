// Row walker of the CSV path: a thread that knows where a record starts finds its fields itself,
// eight bytes at a time, instead of reading a field index from memory.
//
// Inside one record pandas' tokenizer (csv_dfa.cuh; tokenizer.c `tokenize_bytes`, reached from
// app/file_parser.py:301) only moves between START_FIELD / IN_FIELD / IN_QUOTED_FIELD /
// QUOTE_IN_QUOTED_FIELD: a field that starts with '"' runs to the quote that is not followed by another
// quote, whatever follows that quote up to the next delimiter / LF / CR is literal; any other field runs
// to the next delimiter / LF / CR and its quotes are literal.  The record start itself (blank lines,
// CR LF pairs, the blanks that open a line) comes from the parallel DFA scan (csv2_scan.cu), so the two
// always agree on where a record ends.
//
// While it looks for the end of a field the walker also learns what downstream code needs to know
// about its bytes: control bytes (NUL cuts a token, TAB / LF / CR print as escapes), non-ASCII bytes,
// "" pairs.  The result is the same CellSrc shapes csv_simplify() makes out of an indexed field
// (CM_RAW clean / plain, CM_CSVQ, CM_CSV), so every per-cell function of cell_ops.cuh applies.
#pragma once
#include "cell_ops.cuh"

namespace rfq {

// 8 bytes at p + i, little endian, for a record that ends at `limit` (exclusive): bytes at or
// beyond the limit read as LF (the virtual record terminator).  Reads up to 15 bytes past p + i:
// every buffer this runs on (arena, shared-memory stage) carries 16 bytes of slack.
RFQ_HD uint64_t row_window(const uint8_t* p, uint32_t i, uint32_t limit) {
  const uint32_t rem = limit > i ? limit - i : 0u;
  uint64_t x = load8_masked(p + i, rem);
  if (rem < 8) x |= 0x0A0A0A0A0A0A0A0Aull & ~(rem ? ((1ull << (8 * rem)) - 1ull) : 0ull);
  return x;
}

RFQ_HD uint32_t first_flag_byte(uint64_t flags) {   // index of the lowest byte whose 0x80 bit is set
#ifdef __CUDA_ARCH__
  return (uint32_t)(__ffsll((long long)flags) - 1) >> 3;
#else
  return (uint32_t)__builtin_ctzll(flags) >> 3;
#endif
}

// bytes of x below index k kept, the others replaced by blanks (so they are neither control nor high)
RFQ_HD uint64_t keep_below(uint64_t x, uint32_t k) {
  const uint64_t m = k >= 8 ? ~0ull : ((1ull << (8 * k)) - 1ull);
  return (x & m) | (0x2020202020202020ull & ~m);
}

struct FieldScan {
  CellSrc src;
  uint32_t next;   // position of the byte after the field's terminator
  bool eol;        // the terminator ended the record (LF, CR or the limit)
};

// Field that starts at p[pos]; the record cannot extend beyond `limit`.  pos == limit is an empty
// last field.
RFQ_HD FieldScan scan_field(const uint8_t* p, uint32_t pos, uint32_t limit, uint32_t sep) {
  const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
  const uint64_t sep8 = ones * sep;
  FieldScan r;
  r.src.mode = CM_RAW; r.src.missing = false; r.src.clean = true; r.src.bits = 0;
  r.src.p = p + pos; r.src.n = 0;
  uint64_t acc = 0;        // 0x80 bits: some byte is non-ASCII; bit 0: some byte is a control byte
  uint32_t i = pos;
  uint32_t body = pos;     // where the unquoted scan starts (after a closed quote section: the junk)
  bool quoted = false, junk = false;
  uint32_t pairs = 0, close = 0;
  if (pos < limit && p[pos] == '"') {
    quoted = true;
    i = pos + 1;
    bool closed = false;
    while (!closed) {
      if (i >= limit) break;                               // unterminated: the record ends inside the quotes
      const uint64_t x = load8_masked(p + i, limit - i);   // bytes beyond the limit read as zero (never a quote)
      const uint64_t q = zero_bytes(x ^ (ones * '"'));
      const uint32_t rem = limit - i;
      if (!q) {
        const uint64_t xm = rem < 8 ? keep_below(x, rem) : x;
        acc |= (xm & highs) | (((xm - ones * 0x0E) & ~xm & highs) ? 1ull : 0ull);
        i += 8;
        continue;
      }
      const uint32_t k = first_flag_byte(q);
      const uint64_t xm = keep_below(x, k);
      acc |= (xm & highs) | (((xm - ones * 0x0E) & ~xm & highs) ? 1ull : 0ull);
      const uint32_t j = i + k;                            // a quote inside the quoted section
      const uint32_t c1 = j + 1 < limit ? p[j + 1] : '\n';
      if (c1 == '"') { ++pairs; i = j + 2; continue; }
      closed = true; close = j;
      if (c1 == sep || c1 == '\n' || c1 == '\r') {         // the usual end: quote, then the terminator
        const uint32_t inner = j - (pos + 1);
        r.next = j + 2;
        r.eol = c1 != sep;
        if (acc & 1ull) { r.src.mode = CM_CSV; r.src.clean = false; r.src.n = j + 1 - pos; return r; }
        r.src.p = p + pos + 1; r.src.n = inner;
        if (pairs == 0) { r.src.clean = !(acc & highs); return r; }
        if (!(acc & highs)) { r.src.mode = CM_CSVQ; r.src.clean = false; r.src.bits = inner - pairs; return r; }
        r.src.mode = CM_CSV; r.src.clean = false; r.src.p = p + pos; r.src.n = j + 1 - pos;
        return r;
      }
      junk = true; body = j + 1;                           // "x"y : literal bytes up to the terminator
    }
    if (!closed) {   // EOF inside string / record limit inside the quotes
      r.src.mode = CM_CSV; r.src.clean = false; r.src.n = limit - pos;
      r.next = limit + 1; r.eol = true;
      return r;
    }
    i = body;
  }
  // unquoted bytes up to the delimiter / LF / CR / limit
  // One test per window finds the delimiter and every byte below 0x0E (LF and CR end the field, the other
  // control bytes only mark it); the record's limit counts as a terminator at its own position, so the
  // window itself is read unmasked (every buffer carries 16 bytes of slack).
  for (;;) {
    const uint32_t rem = limit > i ? limit - i : 0u;
    const uint64_t x = load8_masked(p + i, 8);
    uint64_t t = zero_bytes(x ^ sep8) | (~((x | highs) - ones * 0x0Eu) & ~x & highs);
    if (rem < 8u) t |= 0x80ull << (8u * rem);
    while (t) {
      const uint32_t k = first_flag_byte(t);
      const uint32_t tb = (uint32_t)(x >> (8 * k)) & 0xFFu;
      const bool virt = k >= rem;
      if (virt || tb == sep || tb == '\n' || tb == '\r') {
        acc |= x & highs & ((1ull << (8u * k)) - 1ull);
        const uint32_t e = i + k;
        r.next = e + 1;
        r.eol = virt || tb != sep;
        r.src.n = e - pos;
        if (quoted || junk || (acc & 1ull)) { r.src.mode = CM_CSV; r.src.clean = false; }
        else r.src.clean = !(acc & highs);
        (void)close;
        return r;
      }
      acc |= 1ull;       // TAB, NUL, VT, ...: a control byte inside the field
      t &= t - 1ull;
    }
    acc |= x & highs;
    i += 8;
  }
}

// The same field as the index-based fetch (fetch_cell + csv_simplify) would describe it: used by the
// cold kernels, which re-find a cell from the start of its row.
struct RowWalk {
  const uint8_t* p;     // base pointer: p[pos] are document bytes
  uint32_t pos, limit, sep;
  bool eol;             // the record's terminator has been consumed
  RFQ_HD RowWalk(const uint8_t* base, uint32_t start, uint32_t lim, uint32_t s) : p(base), pos(start), limit(lim), sep(s), eol(false), field_pos(start) {}
  uint32_t field_pos;   // where the field returned last started
  RFQ_HD CellSrc next() {
    if (eol) {            // short row: pandas pads with empty tokens
      CellSrc s; s.mode = CM_CSV; s.missing = true; s.clean = false; s.p = p; s.n = 0; s.bits = 0;
      return s;
    }
    field_pos = pos;
    FieldScan f = scan_field(p, pos, limit, sep);
    pos = f.next; eol = f.eol;
    return f.src;
  }
};

// ---- the cells the hot kernels finish themselves ------------------------------------------------
// Everything else (numbers that are not short decimal literals or small integers, text that needs the
// quote state machine, non-ASCII numeric-looking tokens) goes to the fix-up kernels, which run the
// general functions of cell_ops.cuh.  Keeping those out of the hot kernels is what lets them compile
// to a small register count.
// Does precise_xstrtod accept this token (cell_parse.cuh parse_double, without computing the value)?
// 0 = no, 1 = yes, 2 = left to the general parser (an inf word, an exponent longer than 17 digits)
RFQ_HD int float_shape(const uint8_t* p, uint32_t n) {
  uint32_t i = 0;
  while (i < n && is_space_ascii(p[i])) ++i;
  if (i < n && (p[i] == '-' || p[i] == '+')) ++i;
  if (i < n && (p[i] | 0x20u) == 'i') return 2;
  uint32_t nd = 0;
  while (i < n && is_digit_ascii(p[i])) { ++i; ++nd; }
  if (i < n && p[i] == '.') {
    ++i;
    while (i < n && is_digit_ascii(p[i])) { ++i; ++nd; }
  }
  if (nd == 0) return 0;
  if (i < n && (p[i] | 0x20u) == 'e') {
    ++i;
    if (i < n && (p[i] == '-' || p[i] == '+')) ++i;
    uint32_t ne = 0;
    while (i < n && is_digit_ascii(p[i])) { ++i; ++ne; }
    if (ne == 0) return 0;
    if (ne > 17) return 2;
  }
  while (i < n && is_space_ascii(p[i])) ++i;
  return i < n ? 0 : 1;
}

// the inf words _try_double accepts after xstrtod fails: [+-]?(inf|infinity), any case, nothing around
RFQ_HD bool inf_word(const uint8_t* p, uint32_t n) {
  uint32_t i = (n && (p[0] == '+' || p[0] == '-')) ? 1u : 0u;
  const uint32_t l = n - i;
  if (l != 3 && l != 8) return false;
  const char* w = "infinity";
  for (uint32_t k = 0; k < l; ++k) if ((p[i + k] | 0x20u) != (uint32_t)w[k]) return false;
  return true;
}

struct HotCand {
  bool cold;            // not decided here
  CellCand m;
};

// The hot classifier reports what it found through a sink, so that the typing sweep can fold each
// outcome straight into its statistics (no struct of seventeen candidates travelling through
// registers) while the CPU cross-check collects the same calls into a CellCand:
//   cold()                                       left to the fix-up kernel
//   na()                                         missing cell or NA token
//   word(flags, str_len, extra, bool_len)        not a number (int64 status: invalid)
//   integer(flags, str_len, int_len, neg, large) an int64 of magnitude <= 2^53 (float view: digits + ".000000")
//   decimal(flags, str_len, ip_len, tz, obj_len, neg, large)   [+-]d{1,8}.d{1,6}
// A decision of the hot classifier, packed so that the rarely-run general part can live out of line:
//   bits 0-2 outcome, 3 neg, 4 large, 5-9 int_len / ip_len, 10-12 tz, 13-17 obj_len, 32-63 SF_* flags
enum { HD_COLD = 0, HD_WORD = 2, HD_INT = 3, HD_DEC = 4 };
RFQ_HD uint64_t hd_pack(uint32_t kind, uint32_t f, uint32_t len, uint32_t tz, uint32_t obj_len, bool neg, bool large) {
  return (uint64_t)kind | (neg ? 8ull : 0ull) | (large ? 16ull : 0ull) | ((uint64_t)len << 5) | ((uint64_t)tz << 10) |
         ((uint64_t)obj_len << 13) | ((uint64_t)f << 32);
}

// Numeric-looking clean ASCII tokens the eight-byte shortcuts of classify_hot do not settle: the parsers'
// own byte loops.  Out of line on the device: C2-like tables (quantities, prices, dates) never come
// here, and the hot sweep's code stays inside the instruction cache.
RFQ_HD_COLD uint64_t classify_hot_rest(const uint8_t* p, uint32_t n, uint64_t first8, uint32_t f) {
  const uint32_t c0 = (uint32_t)(first8 & 0xFFu);
  int64_t iv;
  const int st = parse_int64(RawIter(p, n), &iv);
  if (st == INT_OK) {
    const uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
    if (mag > (1ull << 53) || (iv == 0 && c0 != '0')) return HD_COLD;   // float view needs the exact parser; "-0", " 0", "+0"
    if (iv < 0) f |= SF_NEG;
    f |= SF_BOOL_BAD;
    return hd_pack(HD_INT, f, (uint32_t)int64_len(iv), 0, 0, iv < 0, mag > 1000000ull);
  }
  if (st == INT_OVERFLOW) return HD_COLD;
  Fixed6 d6;
  if (!decimal_literal(p, n, &d6)) {
    // not an integer: a date, a part number that starts with a digit, ... decided here when no
    // grammar takes it; real floats beyond the literal shape go to the fix-up kernel
    const uint32_t c1 = (c0 == '+' || c0 == '-') ? (uint32_t)((first8 >> 8) & 0xFFu) : c0;
    if ((c1 | 0x20u) == 'i') {                 // INR, item, -inf: only the inf words are numbers
      if (inf_word(p, n)) return HD_COLD;
    } else if (float_shape(p, n) != 0) return HD_COLD;
    uint32_t i = 0;
    while (i < n && is_space_ascii(p[i])) ++i;
    if (i < n && p[i] == '-') f |= SF_NEG; else f |= SF_UINT_INVALID;   // str_to_uint64: a leading '-' is "negative", not an error
    f |= SF_FLOAT_BAD | SF_BOOL_BAD;           // no bool word starts like a number
    const int pp = pyint_prefix(first8, n);
    if (pp == 2) return HD_COLD;
    if (pp == 0) f |= SF_PYINT_BAD;
    return hd_pack(HD_WORD, f, 0, 0, 0, false, false);
  }
  // [+-]d{1,8}.d{1,6}: not an int64, not a uint64 (invalid, or "negative" for a leading '-'), never
  // int()-parsable, not a bool; as float64 its "%.6f" digits are its own (cell_format.cuh)
  if (c0 == '-') f |= SF_NEG; else f |= SF_UINT_INVALID;
  f |= SF_PYINT_BAD | SF_BOOL_BAD;
  return hd_pack(HD_DEC, f, (d6.neg ? 1u : 0u) + (uint32_t)d6.ip_digits, (uint32_t)d6.tz, (uint32_t)objfloat_len(d6), d6.neg,
                 d6.ip > 1000000ull || (d6.ip == 1000000ull && d6.frac));
}

RFQ_HD_COLD bool is_na_word(uint32_t n, uint64_t first8) { return is_na_token(n, first8); }

template <class Sink>
RFQ_HD void classify_hot(const CellSrc& s, Sink& k) {
  if (s.missing) { k.na(); return; }
  if (s.mode == CM_CSV) { k.cold(); return; }
  const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
  uint32_t nbytes, ncp, f = SF_ANY_NONNA;
  uint64_t first8;
  if (s.clean) { nbytes = s.n; ncp = s.n; first8 = s.n ? load8_masked(s.p, s.n) : 0; }
  else if (s.mode == CM_CSVQ) {
    nbytes = (uint32_t)s.bits; ncp = nbytes;
    const uint64_t raw8 = load8_masked(s.p, s.n < 8u ? s.n : 8u);
    if (s.n >= 8u && !zero_bytes(raw8 ^ (ones * '"'))) first8 = raw8;   // no "" among the first eight bytes: they are the first eight characters
    else {
      first8 = 0;
      PairIter it(s.p, s.n);
      for (int q = 0; q < 8; ++q) { int c = it.next(); if (c < 0) break; first8 |= (uint64_t)c << (8 * q); }
    }
  } else {   // plain span with non-ASCII bytes
    TextInfo t = scan_text_raw(s.p, s.n);
    if (!t.utf8_ok) f |= SF_BAD_UTF8;
    if (t.nesc) { k.cold(); return; }   // cannot happen (no control bytes), kept for safety
    nbytes = t.nbytes; ncp = t.ncp; first8 = t.first8;
  }
  const uint32_t c0 = (uint32_t)(first8 & 0xFFu);
  if (nbytes == 0) { k.na(); return; }
  // every NA word starts with one of # - 1 < N n and has at most eight bytes; the comparison with the
  // nineteen words is out of line
  if (nbytes <= 8u && (c0 == 'N' || c0 == 'n' || c0 == '#' || c0 == '-' || c0 == '1' || c0 == '<') &&
      is_na_word(nbytes, first8)) { k.na(); return; }
  uint64_t dec;
  if (!(is_digit_ascii((int)c0) || c0 == '+' || c0 == '-' || c0 == '.' || is_space_ascii((int)c0) ||
        (c0 | 0x20u) == 'i')) {          // a plain word: no numeric grammar starts like this
    const int bt = bool_token(nbytes, first8);
    f |= SF_UINT_INVALID | SF_FLOAT_BAD | SF_PYINT_BAD;
    if (bt < 0) f |= SF_BOOL_BAD;
    k.word(f, ncp, nbytes - ncp, bt < 0 ? 0u : bt == 1 ? 4u : 5u);
    return;
  }
  if (!s.clean || s.mode != CM_RAW) { k.cold(); return; }   // numeric-looking but quoted-with-pairs / non-ASCII
  // ---- decided from the first eight bytes alone (quantities, prices, dates): the same answers as the
  // parsers of classify_hot_rest, without their byte loops
  {
    const uint64_t live = s.n >= 8u ? ~0ull : ((1ull << (8u * s.n)) - 1ull);
    const uint64_t y = first8 ^ (ones * 0x30u);                              // '0'..'9' -> 0..9 (clean ASCII: no byte carries)
    const uint64_t nondig = ((y + ones * 0x76u) | y) & highs & live;         // 0x80 where the byte is not a digit
    const bool d0 = (uint32_t)(c0 - '1') < 9u;                               // starts with 1..9
    const uint64_t dot = zero_bytes(first8 ^ (ones * '.')) & live;
    const uint64_t sg = (zero_bytes(first8 ^ (ones * '-')) | zero_bytes(first8 ^ (ones * '+'))) & live;
    if (s.n <= 8u && !nondig && (d0 || s.n == 1u)) {
      // d{1,8} without a leading zero: an int64 that prints as itself
      dec = hd_pack(HD_INT, f | SF_BOOL_BAD, s.n, 0, 0, false,
                    s.n == 8u || (s.n == 7u && first8 != RFQ_P8('1', '0', '0', '0', '0', '0', '0', 0)));
    } else if (s.n <= 8u && d0 && nondig && nondig == dot && !(dot & (dot - 1ull)) && first_flag_byte(dot) + 1u < s.n) {
      // [1-9]d*.d+ in at most eight bytes: one '.', at least one decimal (at most six)
      const uint32_t kd = first_flag_byte(dot);                              // digits before the point
      const uint64_t nz = ~zero_bytes(first8 ^ (ones * '0')) & highs & live; // bytes other than '0' (the point is one)
#ifdef __CUDA_ARCH__
      const uint32_t hb = (uint32_t)(63 - __clzll((long long)nz)) >> 3;
#else
      const uint32_t hb = (uint32_t)(63 - __builtin_clzll(nz)) >> 3;
#endif
      const uint32_t nf = s.n - 1u - kd;
      const uint32_t tz = hb == kd ? 6u : (6u - nf) + (s.n - 1u - hb);       // trailing zeros of the six-digit fraction
      dec = hd_pack(HD_DEC, f | SF_UINT_INVALID | SF_PYINT_BAD | SF_BOOL_BAD, kd, tz, 2u + kd + (tz >= 6u ? 1u : 6u - tz), false, false);
    } else if ((c0 | 0x20u) == 'i') {
      // INR, item, Infinity: only the inf words ([+-]?inf / infinity, any case) are numbers; the others fail
      // every grammar at their first byte
      const uint64_t lw = first8 | (ones * 0x20u);
      const bool infw = (s.n == 3u && (lw & 0xFFFFFFull) == RFQ_P8('i', 'n', 'f', 0, 0, 0, 0, 0)) ||
                        (s.n == 8u && lw == RFQ_P8('i', 'n', 'f', 'i', 'n', 'i', 't', 'y'));
      dec = infw ? (uint64_t)HD_COLD
                 : hd_pack(HD_WORD, f | SF_UINT_INVALID | SF_FLOAT_BAD | SF_BOOL_BAD | SF_PYINT_BAD, 0, 0, 0, false, false);
    } else if (is_digit_ascii((int)c0) && (sg & ((~((y + ones * 0x76u) | y) & highs & live) << 8))) {
      // 2025-09-12, 10-32UNF: a sign right after a digit ends every numeric grammar
      dec = hd_pack(HD_WORD, f | SF_UINT_INVALID | SF_FLOAT_BAD | SF_BOOL_BAD | SF_PYINT_BAD, 0, 0, 0, false, false);
    } else {
      dec = classify_hot_rest(s.p, s.n, first8, f);
    }
  }
  const uint32_t df = (uint32_t)(dec >> 32), len = (uint32_t)(dec >> 5) & 31u;
  const bool neg = (dec & 8ull) != 0, large = (dec & 16ull) != 0;
  switch ((uint32_t)dec & 7u) {
    case HD_WORD: k.word(df, ncp, 0u, 0u); return;
    case HD_INT: k.integer(df, ncp, len, neg, large); return;
    case HD_DEC: k.decimal(df, ncp, len, (uint32_t)(dec >> 10) & 7u, (uint32_t)(dec >> 13) & 31u, neg, large); return;
    default: k.cold(); return;
  }
}

// the calls collected into the candidates classify_measure_cell() reports (CPU cross-check, fix-ups)
struct CandSink {
  HotCand h;
  RFQ_HD CandSink() {
    h.cold = false;
    CellCand& m = h.m;
    m.flags = 0; m.int_status = INT_OK; m.na = false; m.has_float = false; m.need_exact = false;
    m.str_len = 0; m.extra = 0; m.int_len = 0; m.uint_len = 0; m.bool_len = 0;
    m.ip_len = 0; m.tz = 6; m.inf_len = 0; m.obj_len = 0; m.sci_len = 0; m.fl = 0;
  }
  RFQ_HD void cold() { h.cold = true; }
  RFQ_HD void na() { h.m.flags = SF_ANY_NA; h.m.na = true; }
  RFQ_HD void word(uint32_t f, uint32_t str_len, uint32_t extra, uint32_t bool_len) {
    h.m.flags = f; h.m.int_status = INT_INVALID; h.m.str_len = str_len; h.m.extra = extra; h.m.bool_len = bool_len;
  }
  RFQ_HD void integer(uint32_t f, uint32_t str_len, uint32_t int_len, bool neg, bool large) {
    CellCand& m = h.m;
    m.flags = f; m.str_len = str_len; m.int_len = int_len;
    if (!neg) m.uint_len = int_len;
    m.has_float = true;
    m.ip_len = int_len; m.tz = 6; m.obj_len = int_len - (neg ? 1u : 0u) + 3u;
    m.sci_len = 12u + (neg ? 1u : 0u);
    m.fl = MF_FINITE | (large ? MF_LARGE : 0);
  }
  RFQ_HD void decimal(uint32_t f, uint32_t str_len, uint32_t ip_len, uint32_t tz, uint32_t obj_len, bool neg, bool large) {
    CellCand& m = h.m;
    m.flags = f; m.int_status = INT_INVALID; m.str_len = str_len;
    m.has_float = true;
    m.ip_len = ip_len; m.tz = tz; m.obj_len = obj_len; m.sci_len = 12u + (neg ? 1u : 0u);
    m.fl = MF_FINITE | (large ? MF_LARGE : 0);
  }
};
RFQ_HD HotCand classify_measure_hot(const CellSrc& s) {
  CandSink k;
  classify_hot(s, k);
  return k.h;
}

// -?(0|[1-9]d{0,7}).d{dec}: the shape decimal_literal() accepts (at most 8 integer digits, dec <= 6
// decimals, not below 1e-5) written the way "%.{dec}f" writes it — no '+', no leading zeros, exactly
// `dec` decimals — so the printed field is the token itself.
RFQ_HD bool canonical_decimal(const uint8_t* p, uint32_t n, int dec) {
  if (n < (uint32_t)dec + 2u || n > 17u) return false;
  if (n <= 8u && (uint32_t)(p[0] - '1') < 9u) {   // [1-9]d*.d{dec} in eight bytes: the only non-digit is the point, in its place
    const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
    const uint64_t x = load8_masked(p, n);
    const uint64_t y = x ^ (ones * 0x30u);
    const uint64_t live = n >= 8u ? ~0ull : ((1ull << (8u * n)) - 1ull);
    const uint64_t nondig = ((y + ones * 0x76u) | y | x) & highs & live;
    return nondig == (0x80ull << (8u * (n - 1u - (uint32_t)dec))) && p[n - 1u - (uint32_t)dec] == '.';
  }
  uint32_t i = p[0] == '-' ? 1u : 0u;
  const uint32_t dot = n - 1u - (uint32_t)dec;
  if (dot <= i || dot - i > 8u || p[dot] != '.') return false;
  if (p[i] == '0' && dot - i != 1u) return false;                 // leading zero
  uint32_t nz = 0;
  for (uint32_t k = i; k < n; ++k) {
    if (k == dot) continue;
    const uint32_t dgt = (uint32_t)p[k] - '0';
    if (dgt > 9u) return false;
    if (k < dot || k + 1u < n) nz |= dgt;                          // any non-zero digit above the last decimal
  }
  if (p[i] == '0' && nz == 0) return false;                        // below 1e-5 (or zero): left to the general path
  return true;
}

// ---- the cells the hot emit sweep prints itself -------------------------------------------------
// n bytes from src (any address space, any alignment) to dst (byte stores)
RFQ_HD void copy_span8(uint8_t* d, uint64_t x) {
  const uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
  d[0] = (uint8_t)lo; d[1] = (uint8_t)(lo >> 8); d[2] = (uint8_t)(lo >> 16); d[3] = (uint8_t)(lo >> 24);
  d[4] = (uint8_t)hi; d[5] = (uint8_t)(hi >> 8); d[6] = (uint8_t)(hi >> 16); d[7] = (uint8_t)(hi >> 24);
}
RFQ_HD void copy_span(uint8_t* dst, const uint8_t* src, uint32_t n) {
  if (n >= 8) {
    uint32_t i = 0;
    for (; i + 8 <= n; i += 8) copy_span8(dst + i, load8_masked(src + i, 8));
    // the last 1-7 bytes as one more group of eight that ends at n (it rewrites bytes of the group before with
    // the same values): no loop whose length differs from lane to lane
    if (i < n) copy_span8(dst + (n - 8), load8_masked(src + (n - 8), 8));
    return;
  }
#ifdef __CUDA_ARCH__
  const uint64_t x = load8_masked(src, 8);      // only the n low bytes are stored
#else
  const uint64_t x = load8_masked(src, n);
#endif
  for (uint32_t q = 0; q < n; ++q) dst[q] = (uint8_t)(x >> (8 * q));
}

// What the typing sweep found out about a cell, 16 bits, so that the emit sweep neither looks for
// the end of the field again nor repeats the NA test:
//   bits 0-10 length of the span (2047 and more: filed under mode 3), bit 11 the span sits inside quotes
//   (starts one byte later, the field is two bytes longer), bits 12-13 mode: 0 plain clean ASCII span,
//   1 plain span with non-ASCII bytes, 2 scan the field again (""-pair text), 3 the row goes to the
//   fix-up kernel; bit 14 missing (short row), bit 15 an NA token.
enum { CMETA_QUOTED = 1u << 11, CMETA_MODE_SHIFT = 12, CMETA_MISSING = 1u << 14, CMETA_NA = 1u << 15 };
RFQ_HD uint32_t cell_meta(const CellSrc& s, bool quoted, bool cold, bool na) {
  if (s.missing) return CMETA_MISSING | CMETA_NA;
  if (cold || s.mode == CM_CSV || s.n > 2046u) return 3u << CMETA_MODE_SHIFT;
  if (s.mode == CM_CSVQ) return (2u << CMETA_MODE_SHIFT) | (na ? CMETA_NA : 0u);
  return s.n | (quoted ? CMETA_QUOTED : 0u) | ((s.clean ? 0u : 1u) << CMETA_MODE_SHIFT) | (na ? CMETA_NA : 0u);
}

// Numeric cells that do not print as their own token ("+5", "007", "1.5" in a two-decimal column, integers
// of a float column): parsed and formatted.  Out of line on the device, like classify_hot_rest.
RFQ_HD_COLD int emit_hot_number(const uint8_t* p, uint32_t n, int ct, int kind, int dec, uint32_t width, uint8_t* field) {
  if (kind == PR_INT) {
    int64_t iv = 0;
    if (parse_int64(RawIter(p, n), &iv) != INT_OK) return 1;
    const uint32_t len = (uint32_t)int64_len(iv);
    uint8_t* o = field + (width - len);
    write_u64_rev(iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv, o + len);
    if (iv < 0) o[0] = '-';
    return 0;
  }
  Fixed6 d6;
  if (ct == CT_INT || ct == CT_INTNA) {   // integer tokens of a float column: digits + ".000000"
    int64_t iv = 0;
    if (parse_int64(RawIter(p, n), &iv) != INT_OK) return 1;
    const uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
    if (mag > (1ull << 53)) return 1;
    d6.neg = iv < 0; d6.inf = false; d6.big = false; d6.ip = mag; d6.frac = 0; d6.ip_digits = ndigits_u64(mag); d6.tz = 6;
  } else if (!decimal_literal(p, n, &d6)) {   // a short decimal literal prints its own digits (cell_format.cuh)
    int64_t iv = 0;                           // "100" next to "99.5": xstrtod makes the same value of it
    if (parse_int64(RawIter(p, n), &iv) != INT_OK) return 1;
    const uint64_t mag = iv < 0 ? 0ull - (uint64_t)iv : (uint64_t)iv;
    if (mag > (1ull << 53) || (iv == 0 && p[0] != '0')) return 1;   // beyond exact doubles; "-0"
    d6.neg = iv < 0; d6.inf = false; d6.big = false; d6.ip = mag; d6.frac = 0; d6.ip_digits = ndigits_u64(mag); d6.tz = 6;
  }
  uint8_t buf[24];
  const int dd = kind == PR_FIXED ? dec : objfloat_dec(d6);
  const int len = fixed_write(d6, 0., dd, kind == PR_OBJFLOAT, buf);
  uint8_t* o = field + (width - (uint32_t)len);
  for (int q = 0; q < len; ++q) o[q] = buf[q];
  return 0;
}

// One cell printed right-justified into its (blank-filled) field.  Returns 0 = done, 1 = the row
// goes to the fix-up kernel.  *extra receives printed bytes - printed code points.
// na_known: -1 = find out here, 0 / 1 = what the typing sweep found (is_na_token of the token).
RFQ_HD int emit_hot(const CellSrc& s, int ct, int kind, int dec, uint32_t width,
                                        uint8_t* field, uint32_t* extra, int na_known = -1) {
  *extra = 0;
  if (s.missing) return 0;
  if (s.mode == CM_CSV) return 1;
  if (na_known == 1 && ct != CT_STRNF) return 0;
  const uint32_t nbytes = s.mode == CM_CSVQ ? (uint32_t)s.bits : s.n;
  uint64_t first8 = 0;
  if (na_known < 0 || kind == PR_BOOL || kind == PR_INT) {   // text whose NA test is done is copied without a look at its head
    if (s.mode == CM_CSVQ) {
      first8 = 0x2020202020202020ull;
      if (nbytes <= 8) {   // only short tokens can be NA words
        first8 = 0;
        PairIter it(s.p, s.n);
        for (int k = 0; k < 8; ++k) { int c = it.next(); if (c < 0) break; first8 |= (uint64_t)c << (8 * k); }
      }
    } else if (s.n) first8 = load8_masked(s.p, s.n);
  }
  if (na_known < 0 && ct != CT_STRNF && is_na_token(nbytes, first8)) return 0;   // NaN -> '' after fillna: the blanks are there already
  switch (kind) {
    case PR_STR: {
      if (s.mode == CM_CSVQ) {
        uint8_t* o = field + (width - nbytes);
        PairIter it(s.p, s.n);
        int c;
        while ((c = it.next()) >= 0) *o++ = (uint8_t)c;
        return 0;
      }
      uint32_t cont = 0;
      if (!s.clean) {   // plain span with non-ASCII bytes (valid UTF-8: the typing sweep checked): count continuation bytes
        const uint64_t highs = 0x8080808080808080ull;
        for (uint32_t i = 0; i < s.n; i += 8) {
          const uint64_t x = load8_masked(s.p + i, s.n - i);
#ifdef __CUDA_ARCH__
          cont += (uint32_t)__popcll(x & ~(x << 1) & highs);
#else
          cont += (uint32_t)__builtin_popcountll(x & ~(x << 1) & highs);
#endif
        }
      }
      copy_span(field + (width - (s.n - cont)), s.p, s.n);
      *extra = cont;
      return 0;
    }
    case PR_BOOL: {
      const int bv = bool_token(nbytes, first8);
      if (bv < 0 || !s.clean) return 1;
      const uint32_t w4 = bv == 1 ? 0x65757254u /* True */ : 0x736C6146u /* Fals */;
      uint8_t* o = field + (width - (bv == 1 ? 4u : 5u));
      o[0] = (uint8_t)w4; o[1] = (uint8_t)(w4 >> 8); o[2] = (uint8_t)(w4 >> 16); o[3] = (uint8_t)(w4 >> 24);
      if (bv == 0) o[4] = 'e';
      return 0;
    }
    case PR_INT: {
      if (!s.clean || s.mode != CM_RAW) return 1;
      const uint32_t b0 = (uint32_t)(first8 & 0xFFu), b1 = (uint32_t)((first8 >> 8) & 0xFFu);
      const uint32_t last = s.p[s.n - 1];
      const bool canon = (uint32_t)(last - '0') < 10u &&
                         ((uint32_t)(b0 - '1') < 9u || (b0 == '-' && (uint32_t)(b1 - '1') < 9u) || (s.n == 1 && b0 == '0'));
      if (canon) { copy_span(field + (width - s.n), s.p, s.n); return 0; }   // an int64 token in canonical form prints as itself
      return emit_hot_number(s.p, s.n, ct, kind, dec, width, field);
    }
    case PR_FIXED: case PR_OBJFLOAT: {
      if (!s.clean || s.mode != CM_RAW) return 1;
      if (kind == PR_FIXED && ct != CT_INT && ct != CT_INTNA && canonical_decimal(s.p, s.n, dec)) {
        copy_span(field + (width - s.n), s.p, s.n);   // "959.40" in a two-decimal column is its own "%.2f"
        return 0;
      }
      return emit_hot_number(s.p, s.n, ct, kind, dec, width, field);
    }
    default: return 1;   // uint64, "%.6e", Python ints
  }
}

}  // namespace rfq
