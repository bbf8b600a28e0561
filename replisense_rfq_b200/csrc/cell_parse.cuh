// Per-cell lexing: CSV field unescape iterator, NA-string match, and the numeric field parsers
// for quantity and price — bit-exact restatements of what pandas' C parser applies to a token
// (third-party, un-vendored: pandas/_libs/src/parser/tokenizer.c `str_to_int64`,
// `str_to_uint64`, `precise_xstrtod`; call sites pandas/_libs/parsers.pyx `_try_int64`,
// `_try_uint64`, `_try_double`, `_try_bool_flex`; reached from app/file_parser.py:301).
// Behaviour pinned by differential tests against pd.read_csv (tests/test_hostcore.py) because the
// C source is not in this image.
#pragma once
#include "rfq_common.cuh"
#include "pow10_table.cuh"

namespace rfq {

// ---- decoded-byte iterators -----------------------------------------------------------------
// CSV field: raw span between delimiters. A leading '"' opens a quoted section; '""' inside it
// is one '"'; after the closing quote further bytes are literal; a NUL ends the token (C strlen).
struct CsvIter {
  const uint8_t* p;
  const uint8_t* e;
  int st;  // 0 start, 1 in quotes, 2 quote seen inside quotes, 3 literal
  RFQ_HD CsvIter(const uint8_t* b, uint32_t n) : p(b), e(b + n), st(0) {}
  RFQ_HD int next() {
    while (p < e) {
      uint32_t c = *p++;
      if (st == 3) {
      } else if (st == 0) {
        if (c == '"') { st = 1; continue; }
        st = 3;
      } else if (st == 1) {
        if (c == '"') { st = 2; continue; }
      } else {  // st == 2
        st = (c == '"') ? 1 : 3;
      }
      if (c == 0) { p = e; return -1; }
      return (int)c;
    }
    return -1;
  }
};

// Plain bytes (already-decoded strings, numeric literals of sheet XML)
struct RawIter {
  const uint8_t* p;
  const uint8_t* e;
  RFQ_HD RawIter(const uint8_t* b, uint32_t n) : p(b), e(b + n) {}
  RFQ_HD int next() { return p < e ? (int)*p++ : -1; }
};

RFQ_HD bool is_space_ascii(int c) { return c == ' ' || (unsigned)(c - '\t') < 5u; }
RFQ_HD bool is_digit_ascii(int c) { return (unsigned)(c - '0') < 10u; }

// ---- text scan --------------------------------------------------------------------------------
struct TextInfo {
  uint32_t nbytes;    // decoded bytes
  uint32_t ncp;       // code points (valid only if utf8_ok)
  uint32_t nesc;      // \t \r \n occurrences (printed as two characters by to_string)
  uint64_t first8;    // first 8 decoded bytes, little endian
  bool utf8_ok;
  bool has_fchar;     // contains '.', 'e' or 'E'
};

template <class It>
RFQ_HD TextInfo scan_text(It it) {
  TextInfo t;
  t.nbytes = 0; t.ncp = 0; t.nesc = 0; t.first8 = 0; t.utf8_ok = true; t.has_fchar = false;
  int need = 0;          // continuation bytes still expected
  uint32_t lo = 0x80, hi = 0xBF;
  int c;
  while ((c = it.next()) >= 0) {
    if (t.nbytes < 8) t.first8 |= (uint64_t)c << (8 * t.nbytes);
    ++t.nbytes;
    if (need) {
      if ((uint32_t)c < lo || (uint32_t)c > hi) { t.utf8_ok = false; need = 0; }
      else { --need; lo = 0x80; hi = 0xBF; }
      continue;
    }
    if (c < 0x80) {
      ++t.ncp;
      if (c == '\t' || c == '\n' || c == '\r') ++t.nesc;
      else if (c == '.' || c == 'e' || c == 'E') t.has_fchar = true;
    } else {
      ++t.ncp;
      lo = 0x80; hi = 0xBF;
      if (c >= 0xC2 && c <= 0xDF) need = 1;
      else if (c >= 0xE0 && c <= 0xEF) { need = 2; if (c == 0xE0) lo = 0xA0; if (c == 0xED) hi = 0x9F; }
      else if (c >= 0xF0 && c <= 0xF4) { need = 3; if (c == 0xF0) lo = 0x90; if (c == 0xF4) hi = 0x8F; }
      else t.utf8_ok = false;
    }
  }
  if (need) t.utf8_ok = false;
  return t;
}

// ---- NA strings (pandas STR_NA_VALUES, 19 entries incl. '') ---------------------------------
RFQ_HD uint64_t pack8(const char* s) {
  uint64_t v = 0;
  for (int i = 0; i < 8 && s[i]; ++i) v |= (uint64_t)(uint8_t)s[i] << (8 * i);
  return v;
}
#define RFQ_P8(a, b, c, d, e, f, g, h)                                                        \
  ((uint64_t)(uint8_t)(a) | (uint64_t)(uint8_t)(b) << 8 | (uint64_t)(uint8_t)(c) << 16 |      \
   (uint64_t)(uint8_t)(d) << 24 | (uint64_t)(uint8_t)(e) << 32 | (uint64_t)(uint8_t)(f) << 40 | \
   (uint64_t)(uint8_t)(g) << 48 | (uint64_t)(uint8_t)(h) << 56)

RFQ_HD bool is_na_token(uint32_t n, uint64_t f) {
  if (n == 0) return true;
  if (n > 8) return false;
  {   // every NA word starts with one of # - 1 < N n: most cells leave here
    const uint32_t b = (uint32_t)(f & 0xFFu);
    if (!(b == 'N' || b == 'n' || b == '#' || b == '-' || b == '1' || b == '<')) return false;
  }
  switch (n) {
    case 0: return true;
    case 2: return f == RFQ_P8('N', 'A', 0, 0, 0, 0, 0, 0);
    case 3: return f == RFQ_P8('#', 'N', 'A', 0, 0, 0, 0, 0) || f == RFQ_P8('N', '/', 'A', 0, 0, 0, 0, 0) ||
                   f == RFQ_P8('N', 'a', 'N', 0, 0, 0, 0, 0) || f == RFQ_P8('n', '/', 'a', 0, 0, 0, 0, 0) ||
                   f == RFQ_P8('n', 'a', 'n', 0, 0, 0, 0, 0);
    case 4: return f == RFQ_P8('#', 'N', '/', 'A', 0, 0, 0, 0) || f == RFQ_P8('-', 'N', 'a', 'N', 0, 0, 0, 0) ||
                   f == RFQ_P8('-', 'n', 'a', 'n', 0, 0, 0, 0) || f == RFQ_P8('<', 'N', 'A', '>', 0, 0, 0, 0) ||
                   f == RFQ_P8('N', 'U', 'L', 'L', 0, 0, 0, 0) || f == RFQ_P8('N', 'o', 'n', 'e', 0, 0, 0, 0) ||
                   f == RFQ_P8('n', 'u', 'l', 'l', 0, 0, 0, 0);
    case 6: return f == RFQ_P8('1', '.', '#', 'I', 'N', 'D', 0, 0);
    case 7: return f == RFQ_P8('-', '1', '.', '#', 'I', 'N', 'D', 0) || f == RFQ_P8('1', '.', '#', 'Q', 'N', 'A', 'N', 0);
    case 8: return f == RFQ_P8('#', 'N', '/', 'A', ' ', 'N', '/', 'A') || f == RFQ_P8('-', '1', '.', '#', 'Q', 'N', 'A', 'N');
    default: return false;
  }
}

// bool tokens of _try_bool_flex: True/TRUE/true -> 1, False/FALSE/false -> 0, else -1
RFQ_HD int bool_token(uint32_t n, uint64_t f) {
  if (n == 4 && (f == RFQ_P8('T', 'r', 'u', 'e', 0, 0, 0, 0) || f == RFQ_P8('T', 'R', 'U', 'E', 0, 0, 0, 0) ||
                 f == RFQ_P8('t', 'r', 'u', 'e', 0, 0, 0, 0))) return 1;
  if (n == 5 && (f == RFQ_P8('F', 'a', 'l', 's', 'e', 0, 0, 0) || f == RFQ_P8('F', 'A', 'L', 'S', 'E', 0, 0, 0) ||
                 f == RFQ_P8('f', 'a', 'l', 's', 'e', 0, 0, 0))) return 0;
  return -1;
}

// case-insensitive inf tokens accepted by _try_double after xstrtod fails:
// inf +inf infinity +infinity -> +1 ; -inf -infinity -> -1 ; else 0 (no surrounding spaces)
template <class It>
RFQ_HD int inf_token(It it) {
  int c = it.next();
  int sign = 1;
  if (c == '+' || c == '-') { if (c == '-') sign = -1; c = it.next(); }
  const char* w = "infinity";
  int k = 0;
  while (c >= 0 && k < 8) {
    if ((c | 0x20) != w[k]) return 0;
    ++k;
    c = it.next();
  }
  if (c >= 0) return 0;
  return (k == 3 || k == 8) ? sign : 0;
}

// ---- integers ---------------------------------------------------------------------------------
enum { INT_OK = 0, INT_INVALID = 1, INT_OVERFLOW = 2 };

// str_to_int64: [space]* [+-] digits [space]* ; the first failure scanning left to right wins,
// except that an overflow followed by anything but digits is an invalid token (probed on pandas
// 3.0.2: "9223372036854775808 " and "111111111111111111111.5" send the column to the float pass).
template <class It>
RFQ_HD int parse_int64(It it, int64_t* out) {
  int c = it.next();
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  bool neg = (c == '-');
  if (neg || c == '+') c = it.next();
  if (!is_digit_ascii(c)) return INT_INVALID;
  uint64_t mag = 0;
  const uint64_t lim = neg ? 9223372036854775808ull : 9223372036854775807ull;
  int nd = 0;
  while (is_digit_ascii(c) && nd < 18) {      // 18 digits cannot overflow: no check per digit
    mag = mag * 10 + (uint32_t)(c - '0');
    ++nd;
    c = it.next();
  }
  while (is_digit_ascii(c)) {
    uint32_t d = (uint32_t)(c - '0');
    if (mag > (lim - d) / 10) {
      // reported as an overflow only when nothing but digits follows ("1e20 digits" + ".5", a
      // trailing blank, a letter: ERROR_INVALID_CHARS — the column goes on to the float pass)
      c = it.next();
      while (is_digit_ascii(c)) c = it.next();
      return c >= 0 ? INT_INVALID : INT_OVERFLOW;
    }
    mag = mag * 10 + d;
    c = it.next();
  }
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  if (c >= 0) return INT_INVALID;
  *out = neg ? (int64_t)(0ull - mag) : (int64_t)mag;
  return INT_OK;
}

// str_to_uint64: a leading '-' marks the token negative (value 0, no error); else [+] digits.
struct UintTok { uint64_t v; bool neg; bool gt_i64; };
template <class It>
RFQ_HD int parse_uint64(It it, UintTok* out) {
  out->v = 0; out->neg = false; out->gt_i64 = false;
  int c = it.next();
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  if (c == '-') { out->neg = true; return INT_OK; }
  if (c == '+') c = it.next();
  if (!is_digit_ascii(c)) return INT_INVALID;
  uint64_t v = 0;
  int nd = 0;
  while (is_digit_ascii(c) && nd < 19) {      // 19 digits cannot overflow
    v = v * 10 + (uint32_t)(c - '0');
    ++nd;
    c = it.next();
  }
  while (is_digit_ascii(c)) {
    uint32_t d = (uint32_t)(c - '0');
    if (v > (18446744073709551615ull - d) / 10) {
      c = it.next();                      // same rule as str_to_int64: digits only after the overflow
      while (is_digit_ascii(c)) c = it.next();
      return c >= 0 ? INT_INVALID : INT_OVERFLOW;
    }
    v = v * 10 + d;
    c = it.next();
  }
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  if (c >= 0) return INT_INVALID;
  out->v = v;
  out->gt_i64 = v > 9223372036854775807ull;
  return INT_OK;
}

// Python's int(str): [space]* [+-] digit (_? digit)* [space]*  (what pandas falls back to for
// integer columns beyond uint64).  ndig = digits of the canonical form (no leading zeros).
struct PyInt { bool ok, neg; uint32_t ndig; };
template <class It>
RFQ_HD PyInt parse_pyint(It it) {
  PyInt r; r.ok = false; r.neg = false; r.ndig = 0;
  int c = it.next();
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  if (c == '-') { r.neg = true; c = it.next(); }
  else if (c == '+') c = it.next();
  if (!is_digit_ascii(c)) return r;
  bool nonzero = false, prev_us = false;
  uint32_t nd = 0;
  while (c >= 0) {
    if (is_digit_ascii(c)) {
      if (c != '0') nonzero = true;
      if (nonzero) ++nd;
      prev_us = false;
    } else if (c == '_') {
      if (prev_us) return r;
      prev_us = true;
    } else break;
    c = it.next();
  }
  if (prev_us) return r;
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  if (c >= 0) return r;
  r.ok = true;
  if (!nonzero) { r.neg = false; nd = 1; }
  r.ndig = nd;
  return r;
}
// parse_pyint's state machine over the first min(n, 8) decoded bytes of a token (TextInfo.first8):
// 0 = rejected, 1 = accepted (n <= 8), 2 = the first 8 bytes are a valid prefix, run parse_pyint
RFQ_HD int pyint_prefix(uint64_t w, uint32_t n) {
#define RFQ_B8(k) ((int)((w >> (8 * (k))) & 0xFFu))
  uint32_t m = n < 8 ? n : 8, i = 0;
  while (i < m && is_space_ascii(RFQ_B8(i))) ++i;
  if (i < m && (RFQ_B8(i) == '-' || RFQ_B8(i) == '+')) ++i;
  if (i == m) return n > 8 ? 2 : 0;
  if (!is_digit_ascii(RFQ_B8(i))) return 0;
  bool prev_us = false;
  for (; i < m; ++i) {
    int c = RFQ_B8(i);
    if (is_digit_ascii(c)) prev_us = false;
    else if (c == '_') { if (prev_us) return 0; prev_us = true; }
    else break;
  }
  if (i == m) return n > 8 ? 2 : (prev_us ? 0 : 1);
  if (prev_us) return 0;
  for (; i < m; ++i) if (!is_space_ascii(RFQ_B8(i))) return 0;
#undef RFQ_B8
  return n > 8 ? 2 : 1;
}
// writes str(int(token)); returns bytes
template <class It>
RFQ_HD uint32_t write_pyint(It it, uint8_t* out) {
  uint32_t k = 0;
  int c = it.next();
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  bool neg = false;
  if (c == '-') { neg = true; c = it.next(); }
  else if (c == '+') c = it.next();
  if (neg) out[k++] = '-';
  bool nonzero = false;
  while (c >= 0 && (is_digit_ascii(c) || c == '_')) {
    if (c != '_') { if (c != '0') nonzero = true; if (nonzero) out[k++] = (uint8_t)c; }
    c = it.next();
  }
  if (!nonzero) { k = 0; out[k++] = '0'; }
  return k;
}

// ---- doubles: precise_xstrtod -----------------------------------------------------------------
template <class It>
RFQ_HD bool parse_double(It it, double* out) {
  const int max_digits = 17;
  int c = it.next();
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  bool negative = false;
  if (c == '-') { negative = true; c = it.next(); }
  else if (c == '+') c = it.next();
  double number = 0.;
  long long exponent = 0;
  int num_digits = 0, num_decimals = 0;
  while (is_digit_ascii(c)) {
    if (num_digits < max_digits) {
      number = RFQ_DADD(RFQ_DMUL(number, 10.), (double)(c - '0'));
      ++num_digits;
    } else {
      ++exponent;
    }
    c = it.next();
  }
  if (c == '.') {
    c = it.next();
    while (num_digits < max_digits && is_digit_ascii(c)) {
      number = RFQ_DADD(RFQ_DMUL(number, 10.), (double)(c - '0'));
      c = it.next();
      ++num_digits; ++num_decimals;
    }
    if (num_digits >= max_digits)
      while (is_digit_ascii(c)) c = it.next();
    exponent -= num_decimals;
  }
  if (num_digits == 0) return false;
  if (negative) number = -number;
  if (c == 'e' || c == 'E') {
    c = it.next();
    bool eneg = false;
    if (c == '-') { eneg = true; c = it.next(); }
    else if (c == '+') c = it.next();
    int nd = 0;
    long long n = 0;
    while (nd < max_digits && is_digit_ascii(c)) {
      n = n * 10 + (c - '0');
      ++nd;
      c = it.next();
    }
    if (nd == 0) return false;  // 'e' un-consumed -> trailing garbage -> not a number
    exponent += eneg ? -n : n;
  }
  if (exponent > 308) {
    if (number != 0.) number = negative ? -__builtin_huge_val() : __builtin_huge_val();
  } else if (exponent > 0) {
    number = RFQ_DMUL(number, RFQ_POW10(exponent));
  } else if (exponent < -308) {
    if (exponent < -616) number = 0.;
    else {
      number = RFQ_DDIV(number, RFQ_POW10(-308 - exponent));
      number = RFQ_DDIV(number, RFQ_POW10(308));
    }
  } else {
    number = RFQ_DDIV(number, RFQ_POW10(-exponent));
  }
  while (c >= 0 && is_space_ascii(c)) c = it.next();
  if (c >= 0) return false;
  *out = number;
  return true;
}

}  // namespace rfq
