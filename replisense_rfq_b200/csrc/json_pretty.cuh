// The .json path of FileParser (app/file_parser.py:315-331):
//
//     content = open(path, 'r', encoding='utf-8').read()          (strict UTF-8, universal newlines)
//     data = json.loads(content)
//     return "=== JSON Data ===\n" + json.dumps(data, indent=2, ensure_ascii=False)
//
// restated as ONE streaming pass over the bytes of a document (host/device): CPython's decoder grammar
// (Lib/json/decoder.py JSONObject / JSONArray, Lib/json/scanner.py, Modules/_json.c scanstring — NaN,
// Infinity and -Infinity are values, control characters inside strings are errors) and its error
// messages with the position json.JSONDecodeError reports; the encoder's layout for indent=2
// (Lib/json/encoder.py _make_iterencode: items on their own lines, ", " never used, "[]" / "{}" for empty
// containers, ": " after a key), `encode_basestring` escapes (\" \\ \n \r \t \b \f, \u00XX below 0x20,
// everything else raw), int and float repr.  The function is run twice per document: once to measure
// (out == nullptr), once to write.
//
// Declined (RFQ_DOC_UNSUPPORTED, never guessed):
//   * bytes that are not strict UTF-8 (the reference raises UnicodeDecodeError with CPython's wording),
//   * an object with two equal keys (a dict keeps the LAST value at the FIRST position: not a streaming
//     transformation),
//   * a lone surrogate escape (the result is a str that has no UTF-8 form),
//   * floats written with more than 15 significant digits or beyond 1e+-300 unless the shortest-repr
//     search below settles them, integers of more than 4300 digits (int() refuses them), nesting deeper
//     than RFQ_JSON_MAX_DEPTH.
#pragma once
#include "cell_format.cuh"
#include "utf8.cuh"
#include "xlsx_core.cuh"

namespace rfq {

#define RFQ_JSON_MAX_DEPTH 96
#ifdef __CUDA_ARCH__
#define RFQ_JSON_ANY(x) __any_sync(0xffffffffu, (x))
#else
#define RFQ_JSON_ANY(x) (x)
#endif
enum { JS_OK = 0, JS_INVALID = 1, JS_UNSUPPORTED = 2 };
// json.JSONDecodeError messages
enum { JE_NONE = 0, JE_VALUE, JE_PROPNAME, JE_COLON, JE_COMMA, JE_EXTRA, JE_UNTERMINATED, JE_CTRL, JE_ESCAPE,
       JE_UESCAPE, JE_BOM };

struct JsonResult {
  uint32_t status;      // JS_*
  uint32_t err;         // JE_*
  uint32_t pos;         // character index json.JSONDecodeError reports (code points, after newline translation)
  uint32_t line, col;   // its lineno / colno
  uint64_t out_bytes;   // bytes of the pretty-printed text (without the "=== JSON Data ===\n" banner)
};

struct JsonSink {
  uint8_t* out;
  uint64_t n;
  RFQ_HD void put(uint32_t b) { if (out) out[n] = (uint8_t)b; ++n; }
  RFQ_HD void puts(const char* s) { while (*s) put((uint8_t)*s++); }
  RFQ_HD void copy(const uint8_t* p, uint32_t len) { if (out) for (uint32_t i = 0; i < len; ++i) out[n + i] = p[i]; n += len; }
  RFQ_HD void newline_indent(uint32_t depth) { put('\n'); for (uint32_t i = 0; i < 2 * depth; ++i) put(' '); }
};

RFQ_HD bool json_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
RFQ_HD int json_hex(uint32_t c) {
  if (c >= '0' && c <= '9') return (int)(c - '0');
  c |= 0x20u;
  if (c >= 'a' && c <= 'f') return (int)(c - 'a' + 10);
  return -1;
}

// One JSON string starting at its opening quote p[i] == '"'.  Walks it like scanstring does; with a sink
// writes the re-encoded string (quotes included).  Returns the index after the closing quote, or sets
// *err / *err_at and returns 0.  *lone is set when a lone surrogate appears.
RFQ_HD uint32_t json_string(const uint8_t* p, uint32_t len, uint32_t i, JsonSink* sink, uint32_t* err, uint32_t* err_at,
                            bool* lone) {
  const uint32_t begin = i;
  ++i;
  if (sink) sink->put('"');
  for (;;) {
    if (i >= len) { *err = JE_UNTERMINATED; *err_at = begin; return 0; }
    uint32_t c = p[i];
    if (c == '"') { if (sink) sink->put('"'); return i + 1; }
    if (c < 0x20) { *err = JE_CTRL; *err_at = i; return 0; }
    if (c != '\\') { if (sink) sink->put(c); ++i; continue; }   // any other byte (UTF-8 was validated) is copied
    if (i + 1 >= len) { *err = JE_UNTERMINATED; *err_at = begin; return 0; }
    const uint32_t e = p[i + 1];
    uint32_t cp;
    if (e != 'u') {
      switch (e) {
        case '"': cp = '"'; break;
        case '\\': cp = '\\'; break;
        case '/': cp = '/'; break;
        case 'b': cp = 8; break;
        case 'f': cp = 12; break;
        case 'n': cp = 10; break;
        case 'r': cp = 13; break;
        case 't': cp = 9; break;
        default: *err = JE_ESCAPE; *err_at = i; return 0;
      }
      i += 2;
    } else {
      // \uXXXX
      if (i + 6 >= len) { *err = JE_UESCAPE; *err_at = i + 1; return 0; }   // (_json.c wants a character after the four digits)
      cp = 0;
      for (int k = 0; k < 4; ++k) {
        const int h = json_hex(p[i + 2 + k]);
        if (h < 0) { *err = JE_UESCAPE; *err_at = i + 1; return 0; }
        cp = cp * 16 + (uint32_t)h;
      }
      i += 6;
      if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 < len && p[i] == '\\' && p[i + 1] == 'u') {
        uint32_t c2 = 0;
        for (int k = 0; k < 4; ++k) {
          const int h = json_hex(p[i + 2 + k]);
          if (h < 0) { *err = JE_UESCAPE; *err_at = i + 1; return 0; }
          c2 = c2 * 16 + (uint32_t)h;
        }
        if (c2 >= 0xDC00 && c2 <= 0xDFFF) { cp = 0x10000 + (((cp - 0xD800) << 10) | (c2 - 0xDC00)); i += 6; }
      }
      if (cp >= 0xD800 && cp <= 0xDFFF) *lone = true;
    }
    if (!sink) continue;
    // encode_basestring
    if (cp == '"') { sink->put('\\'); sink->put('"'); }
    else if (cp == '\\') { sink->put('\\'); sink->put('\\'); }
    else if (cp == 10) { sink->put('\\'); sink->put('n'); }
    else if (cp == 13) { sink->put('\\'); sink->put('r'); }
    else if (cp == 9) { sink->put('\\'); sink->put('t'); }
    else if (cp == 8) { sink->put('\\'); sink->put('b'); }
    else if (cp == 12) { sink->put('\\'); sink->put('f'); }
    else if (cp < 0x20) {
      sink->put('\\'); sink->put('u'); sink->put('0'); sink->put('0');
      sink->put("0123456789abcdef"[cp >> 4]); sink->put("0123456789abcdef"[cp & 15]);
    }
    else if (cp < 0x80) sink->put(cp);
    else if (cp < 0x800) { sink->put(0xC0 | (cp >> 6)); sink->put(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { sink->put(0xE0 | (cp >> 12)); sink->put(0x80 | ((cp >> 6) & 0x3F)); sink->put(0x80 | (cp & 0x3F)); }
    else { sink->put(0xF0 | (cp >> 18)); sink->put(0x80 | ((cp >> 12) & 0x3F)); sink->put(0x80 | ((cp >> 6) & 0x3F)); sink->put(0x80 | (cp & 0x3F)); }
  }
}

// decoded UTF-8 bytes of an (already validated) JSON string body: for the duplicate-key test
struct JsonDecoded {
  const uint8_t* p; uint32_t i, end;   // [i, end) = the bytes between the quotes
  uint8_t q[4]; int nq, iq;
  RFQ_HD int next() {
    if (iq < nq) return q[iq++];
    if (i >= end) return -1;
    const uint32_t c = p[i];
    if (c != '\\') { ++i; return (int)c; }
    const uint32_t e = p[i + 1];
    if (e != 'u') {
      i += 2;
      switch (e) { case 'b': return 8; case 'f': return 12; case 'n': return 10; case 'r': return 13; case 't': return 9; default: return (int)e; }
    }
    uint32_t cp = 0;
    for (int k = 0; k < 4; ++k) cp = cp * 16 + (uint32_t)json_hex(p[i + 2 + k]);
    i += 6;
    if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 <= end && p[i] == '\\' && p[i + 1] == 'u') {
      uint32_t c2 = 0;
      for (int k = 0; k < 4; ++k) c2 = c2 * 16 + (uint32_t)json_hex(p[i + 2 + k]);
      if (c2 >= 0xDC00 && c2 <= 0xDFFF) { cp = 0x10000 + (((cp - 0xD800) << 10) | (c2 - 0xDC00)); i += 6; }
    }
    if (cp < 0x80) return (int)cp;
    if (cp < 0x800) { q[0] = (uint8_t)(0xC0 | (cp >> 6)); q[1] = (uint8_t)(0x80 | (cp & 0x3F)); nq = 2; }
    else if (cp < 0x10000) { q[0] = (uint8_t)(0xE0 | (cp >> 12)); q[1] = (uint8_t)(0x80 | ((cp >> 6) & 0x3F)); q[2] = (uint8_t)(0x80 | (cp & 0x3F)); nq = 3; }
    else { q[0] = (uint8_t)(0xF0 | (cp >> 18)); q[1] = (uint8_t)(0x80 | ((cp >> 12) & 0x3F)); q[2] = (uint8_t)(0x80 | ((cp >> 6) & 0x3F)); q[3] = (uint8_t)(0x80 | (cp & 0x3F)); nq = 4; }
    iq = 1;
    return q[0];
  }
};
// Two keys are the same dict key iff their decoded texts are equal.  Spans without a backslash are
// compared bytewise.
RFQ_HD bool json_keys_equal(const uint8_t* p, uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1, bool esc) {
  if (!esc) {
    if (a1 - a0 != b1 - b0) return false;
    for (uint32_t k = 0; k < a1 - a0; ++k) if (p[a0 + k] != p[b0 + k]) return false;
    return true;
  }
  JsonDecoded A{p, a0, a1, {0, 0, 0, 0}, 0, 0}, B{p, b0, b1, {0, 0, 0, 0}, 0, 0};
  for (;;) {
    const int x = A.next(), y = B.next();
    if (x != y) return false;
    if (x < 0) return true;
  }
}

// ---- numbers --------------------------------------------------------------------------------------
// d[.ddd]e+XX / fixed notation of repr(float) from the shortest digits and the exponent of the first one
RFQ_HD void json_put_float(JsonSink* sink, bool neg, const uint8_t* dig, int n, int e10) {
  if (neg) sink->put('-');
  if (e10 < -4 || e10 >= 16) {
    sink->put(dig[0]);
    if (n > 1) { sink->put('.'); for (int k = 1; k < n; ++k) sink->put(dig[k]); }
    sink->put('e');
    sink->put(e10 < 0 ? '-' : '+');
    const int ae = e10 < 0 ? -e10 : e10;
    if (ae >= 100) sink->put('0' + ae / 100);
    sink->put('0' + (ae / 10) % 10);
    sink->put('0' + ae % 10);
    return;
  }
  if (e10 < 0) {
    sink->put('0'); sink->put('.');
    for (int k = 0; k < -e10 - 1; ++k) sink->put('0');
    for (int k = 0; k < n; ++k) sink->put(dig[k]);
    return;
  }
  for (int k = 0; k <= e10; ++k) sink->put(k < n ? dig[k] : '0');
  sink->put('.');
  if (n <= e10 + 1) sink->put('0');
  else for (int k = e10 + 1; k < n; ++k) sink->put(dig[k]);
}

// floor(M * 2^E2 * 10^s) for a quotient below 2^62, in exact big-integer arithmetic; *rem: 0 the value is
// that integer exactly, 1 / 2 / 3 its fraction is below / equal to / above one half.
RFQ_HD_COLD uint64_t json_scaled(uint64_t M, int E2, int s, int* rem) {
  Big N, D;
  N.set_u64(M); D.set_u64(1);
  if (E2 >= 0) N.shl(E2); else D.shl(-E2);
  if (s >= 0) N.mul_pow10(s); else D.mul_pow10(-s);
  uint64_t q = 0;
  for (int bit = 61; bit >= 0; --bit)
    if (big_sub_shifted_if_ge(N, D, bit)) q |= 1ull << bit;
  if (N.n == 0) { *rem = 0; return q; }
  Big twice = N; twice.shl(1);
  const int c = big_cmp(twice, D);
  *rem = c < 0 ? 1 : c == 0 ? 2 : 3;
  return q;
}
// repr(v) digits for any finite positive double: the shortest decimal that reads back as v (David Gay's
// mode 0, what float_repr prints), the closest to v at that length.  The ends of v's rounding interval and v
// itself are scaled ONCE to 17-digit integers with exact big-integer divisions; every shorter length is then
// integer arithmetic on those three (floor / ceil of nested divisions are exact).
RFQ_HD_COLD bool json_shortest(double v, uint8_t* dig, int* n_out, int* e10_out) {
  uint64_t m; int e; bool neg;
  dbl_decompose(v, &m, &e, &neg);
  if (m == 0) return false;
  const int E2 = e - 2;
  const bool closed = (m & 1ull) == 0;
  const bool pow2 = m == (1ull << 52) && (dbl_bits(v) >> 52 & 0x7FF) > 1;
  const uint64_t mv = m * 4u, ml = mv - (pow2 ? 1u : 2u), mh = mv + 2u;
  int e10 = approx_log10(m, e);
  uint64_t Qv = 0; int rv = 0;
  const uint64_t P16 = 10000000000000000ull, P17 = 100000000000000000ull;
  bool ok = false;
  for (int tries = 0; tries < 4 && !ok; ++tries) {
    Qv = json_scaled(mv, E2, 16 - e10, &rv);
    if (Qv >= P17) ++e10; else if (Qv < P16) --e10; else ok = true;
  }
  if (!ok) return false;
  int rl = 0, rh = 0;
  const uint64_t Ql = json_scaled(ml, E2, 16 - e10, &rl), Qh = json_scaled(mh, E2, 16 - e10, &rh);
  uint64_t P = P16;                       // 10^(17 - k)
  for (int k = 1; k <= 17; ++k, P /= 10) {
    uint64_t lo, hi;
    if (rl == 0) { lo = (Ql + P - 1) / P; if (Ql % P == 0 && !closed) lo += 1; } else lo = Ql / P + 1;
    hi = Qh / P;
    if (rh == 0 && Qh % P == 0 && !closed) { if (hi == 0) continue; hi -= 1; }
    if (lo > hi) continue;
    // the k-digit decimal closest to v, ties to even, kept inside [lo, hi]
    uint64_t c = Qv / P;
    const uint64_t r = Qv % P;
    bool up;
    if (P == 1) up = rv == 3 || (rv == 2 && (c & 1ull));
    else {
      const uint64_t half = P / 2;
      up = r > half || (r == half && (rv != 0 || (c & 1ull)));
    }
    if (up) c += 1;
    if (c < lo) c = lo;
    if (c > hi) c = hi;
    int nd = k;
    uint64_t top = 1;
    for (int q = 0; q < k; ++q) top *= 10;
    if (c >= top) { c /= 10; e10 += 1; }           // rounded up into the next power of ten: "1" one place higher
    while (nd > 1 && c % 10 == 0) { c /= 10; --nd; }
    if (c >= top / 10 * 10 && nd == k) { /* unreachable: kept for clarity */ }
    // digits of c (it may have fewer than k after the strip)
    uint8_t tmp[20]; int n = 0;
    do { tmp[n++] = (uint8_t)('0' + c % 10); c /= 10; } while (c);
    for (int i = 0; i < n; ++i) dig[i] = tmp[n - 1 - i];
    *n_out = n; *e10_out = e10;
    return true;
  }
  return false;
}

// The same search in 128-bit integers for the magnitudes documents are made of (about 1e-5 .. 1e21): with
// v = m * 2^e, every k-digit decimal D * 10^p inside v's rounding interval satisfies
//     ml * A <= D * S <= mh * A,   A = 10^max(-p, 0),  S = 10^max(p, 0) << (2 - e),
// ml / mh = the interval's ends in quarter units of the last place (closed when m is even: ties go to the
// even mantissa).  Returns false outside the window (the big-integer search takes over).
RFQ_HD bool json_shortest_fast(double v, uint8_t* dig, int* n_out, int* e10_out) {
  uint64_t m; int e; bool neg;
  dbl_decompose(v, &m, &e, &neg);
  if (m == 0) return false;
  const int E = e - 2;
  if (E > 0 || E < -100) return false;
  const bool closed = (m & 1ull) == 0;
  const bool pow2 = m == (1ull << 52) && (dbl_bits(v) >> 52 & 0x7FF) > 1;   // the gap below is half the gap above
  const unsigned __int128 mv = (unsigned __int128)m * 4u, ml = mv - (pow2 ? 1u : 2u), mh = mv + 2u;
  const int k10 = approx_log10(m, e);                                        // floor(log10 v) within +-1
  auto attempt = [&](int k, int e10, uint64_t* D_out, int* st) -> bool {
    // *st: 0 fits / does not fit decided, +1 e10 too small, -1 e10 too large
    *st = 0;
    const int p = e10 - (k - 1);
    if (p < -21 || p > 17) { *st = 2; return false; }
    unsigned __int128 A = 1, B = 1;
    for (int q = 0; q < -p; ++q) A *= 10u;
    for (int q = 0; q < p; ++q) B *= 10u;
    if (-E >= 70 && p > 0) { *st = 2; return false; }
    if (-E > 120) { *st = 2; return false; }
    // S = B << -E must stay below 2^127, mh * A below 2^128
    int bb = 0; for (unsigned __int128 t = B; t; t >>= 1) ++bb;
    if (bb + (-E) > 126) { *st = 2; return false; }
    int ba = 0; for (unsigned __int128 t = A; t; t >>= 1) ++ba;
    if (ba + 56 > 127) { *st = 2; return false; }
    const unsigned __int128 S = B << (-E);
    const unsigned __int128 nl = ml * A, nh = mh * A, nv = mv * A;
    unsigned __int128 lo = nl / S, hi = nh / S;
    const bool lo_exact = lo * S == nl, hi_exact = hi * S == nh;
    if (!lo_exact) lo += 1; else if (!closed) lo += 1;
    if (hi_exact && !closed) { if (hi == 0) return false; hi -= 1; }
    // the decimal closest to v (ties to even), then into [lo, hi]
    unsigned __int128 c = nv / S;
    const unsigned __int128 r2 = (nv - c * S) * 2u;
    if (r2 > S || (r2 == S && (c & 1u))) c += 1;
    uint64_t lim_lo = 1, lim_hi = 10;
    for (int q = 1; q < k; ++q) { lim_lo *= 10; lim_hi *= 10; }
    if (c >= (unsigned __int128)lim_hi * 1u && lo >= lim_hi) { *st = 1; return false; }
    if (hi < lim_lo) { *st = -1; return false; }
    if (lo > hi) return false;
    if (c < lo) c = lo;
    if (c > hi) c = hi;
    if (c >= lim_hi) { *st = 1; return false; }      // the interval reaches the next power of ten: fewer digits there
    if (c < lim_lo) { *st = -1; return false; }
    *D_out = (uint64_t)c;
    return true;
  };
  auto fits = [&](int k, uint64_t* D_out, int* e10_out2, bool* outside) -> bool {
    int e10 = k10;
    for (int tries = 0; tries < 4; ++tries) {
      int st;
      if (attempt(k, e10, D_out, &st)) { *e10_out2 = e10; return true; }
      if (st == 2) { *outside = true; return false; }
      if (st == 0) return false;
      e10 += st;
    }
    return false;
  };
  bool outside = false;
  uint64_t D = 0; int e10 = 0;
  if (!fits(17, &D, &e10, &outside)) return false;     // (17 digits always read back; a miss means the window)
  int lo = 1, hi = 17;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    uint64_t D1; int e1;
    if (fits(mid, &D1, &e1, &outside)) { hi = mid; D = D1; e10 = e1; }
    else { if (outside) return false; lo = mid + 1; }
  }
  int nd = hi;
  while (nd > 1 && D % 10 == 0) { D /= 10; --nd; }
  for (int i = nd - 1; i >= 0; --i) { dig[i] = (uint8_t)('0' + D % 10); D /= 10; }
  *n_out = nd; *e10_out = e10;
  return true;
}

// A float literal p[0, n) (grammar already checked) printed as repr(float(literal)).  Returns false when the
// value cannot be settled (declined).
RFQ_HD bool json_float(const uint8_t* p, uint32_t n, JsonSink* sink) {
  uint32_t i = 0;
  const bool neg = p[0] == '-';
  if (neg) ++i;
  uint8_t dig[17];
  int nsig = 0, last_nz = 0;            // digits from the first non-zero one on; index (1-based) of the last non-zero
  long long first_pos = 0, pos = 0, int_len = 0;
  bool seen = false, in_frac = false;
  for (; i < n; ++i) {
    const uint32_t c = p[i];
    if (c == '.') { in_frac = true; continue; }
    if (c < '0' || c > '9') break;
    if (!in_frac) ++int_len;
    if (!seen && c != '0') { seen = true; first_pos = pos; }
    if (seen) {
      if (nsig < 17) dig[nsig] = (uint8_t)c;
      ++nsig;
      if (c != '0') last_nz = nsig;
    }
    ++pos;
  }
  long long ex = 0;
  if (i < n) {           // exponent
    ++i;
    bool eneg = false;
    if (p[i] == '+' || p[i] == '-') { eneg = p[i] == '-'; ++i; }
    for (; i < n; ++i) { ex = ex * 10 + (p[i] - '0'); if (ex > 1000000) ex = 1000000; }
    if (eneg) ex = -ex;
  }
  if (!seen) { if (neg) sink->put('-'); sink->put('0'); sink->put('.'); sink->put('0'); return true; }
  const long long e10 = int_len - 1 - first_pos + ex;
  if (e10 >= 309) { if (neg) sink->put('-'); sink->puts("Infinity"); return true; }
  if (e10 <= -325) { if (neg) sink->put('-'); sink->put('0'); sink->put('.'); sink->put('0'); return true; }
  if (last_nz <= 15 && e10 >= -300 && e10 <= 300) {
    // at most 15 significant digits: distinct such decimals are distinct doubles (DBL_DIG), so these ARE the
    // shortest digits that read back
    json_put_float(sink, neg, dig, last_nz, (int)e10);
    return true;
  }
  double v;
  if (!strtod_exact(p, n, &v)) return false;
  if (dbl_is_inf(v)) { if (neg) sink->put('-'); sink->puts("Infinity"); return true; }
  if (v == 0.0) { if (neg) sink->put('-'); sink->put('0'); sink->put('.'); sink->put('0'); return true; }
  uint8_t sd[20]; int sn, se;
  const double av = v < 0 ? -v : v;
  if (!json_shortest_fast(av, sd, &sn, &se) && !json_shortest(av, sd, &sn, &se)) return false;
  json_put_float(sink, neg, sd, sn, se);
  return true;
}

// ---- the document -----------------------------------------------------------------------------------
// character index / line / column of byte offset `at`, as json.JSONDecodeError computes them on the text
// Python read: code points, "\r\n" and a lone "\r" read as one "\n" (universal newlines)
RFQ_HD void json_locate(const uint8_t* p, uint32_t len, uint32_t at, JsonResult* r) {
  uint32_t pos = 0, line = 1, last_nl = 0xFFFFFFFFu;   // last_nl: character index of the last newline before pos
  for (uint32_t i = 0; i < at && i < len; ++i) {
    const uint32_t c = p[i];
    if ((c & 0xC0u) == 0x80u) continue;                 // continuation byte: same character
    if (c == '\r' && i + 1 < len && p[i + 1] == '\n') {
      if (i + 1 >= at) { /* the error sits on the LF of a CRLF: same character as the CR */ break; }
      continue;                                          // the LF that follows is the character
    }
    if (c == '\n' || c == '\r') { ++line; last_nl = pos; }
    ++pos;
  }
  r->pos = pos; r->line = line;
  r->col = last_nl == 0xFFFFFFFFu ? pos + 1 : pos - last_nl;
}

// keys of the objects that are open: where each object's keys start in `keys` (pairs begin, end)
struct JsonFrame { uint32_t key0; };

// Parses p[0, len) and pretty-prints it into out (or only measures when out == nullptr).
// keys: scratch of 2 * (len / 4 + 2) uint32 for the duplicate-key test (may be nullptr: then every document
// with an object of two keys or more is declined).
// (first formulation: nested loops per token.  Kept for the host tests, which run it beside the flat machine
// below; a GPU thread runs the flat one, whose lanes meet again after every byte.)
RFQ_HD void json_pretty_nested(const uint8_t* p, uint32_t len, uint8_t* out, uint32_t* keys, JsonResult* res) {
  res->status = JS_OK; res->err = JE_NONE; res->pos = 0; res->line = 1; res->col = 1; res->out_bytes = 0;
  JsonSink sink{out, 0};
  uint32_t err = 0, at = 0;
  // strict UTF-8 over the whole file comes first (open(..., encoding='utf-8').read())
  {
    struct Acc { const uint8_t* p; uint32_t n; RFQ_HD uint32_t operator()(int64_t k) const { return (k >= 0 && k < (int64_t)n) ? p[k] : 0u; } };
    const Acc B{p, len};
    for (uint32_t k = 0; k < len;) {
      const uint32_t c = p[k];
      if (c < 0x80) { ++k; continue; }
      const int l = u8_lead_len(c);
      if (l == 0 || !u8_seq_valid(B, (int64_t)k)) { res->status = JS_UNSUPPORTED; return; }
      k += (uint32_t)l;
    }
  }
  if (len >= 3 && p[0] == 0xEF && p[1] == 0xBB && p[2] == 0xBF) { err = JE_BOM; at = 0; goto fail; }
  {
    // bit d of is_obj: the container open at depth d is an object
    uint64_t is_obj[(RFQ_JSON_MAX_DEPTH + 63) / 64] = {0};
    uint32_t key0[RFQ_JSON_MAX_DEPTH];
    uint32_t depth = 0, n_keys = 0, i = 0;
    bool lone = false;
    enum { ST_VALUE, ST_AFTER, ST_KEY } st = ST_VALUE;
    for (;;) {
      while (i < len && json_ws(p[i])) ++i;
      if (st == ST_VALUE) {
        if (i >= len) { err = JE_VALUE; at = i; goto fail; }
        const uint32_t c = p[i];
        if (c == '"') {
          const uint32_t e = json_string(p, len, i, &sink, &err, &at, &lone);
          if (!e) goto fail;
          i = e; st = ST_AFTER;
        } else if (c == '{' || c == '[') {
          uint32_t j = i + 1;
          while (j < len && json_ws(p[j])) ++j;
          if (j < len && p[j] == (c == '{' ? '}' : ']')) { sink.put(c); sink.put(p[j]); i = j + 1; st = ST_AFTER; continue; }
          if (depth >= RFQ_JSON_MAX_DEPTH) { res->status = JS_UNSUPPORTED; return; }
          if (c == '{') is_obj[depth >> 6] |= 1ull << (depth & 63); else is_obj[depth >> 6] &= ~(1ull << (depth & 63));
          key0[depth] = n_keys;
          ++depth;
          sink.put(c); sink.newline_indent(depth);
          i = j;
          if (c == '{') {
            if (i >= len || p[i] != '"') { err = JE_PROPNAME; at = i; goto fail; }
            st = ST_KEY;
          }
        } else if (c == '-' || (c >= '0' && c <= '9')) {
          if (c == '-' && i + 9 <= len && p[i + 1] == 'I' && p[i + 2] == 'n' && p[i + 3] == 'f' && p[i + 4] == 'i' && p[i + 5] == 'n' &&
              p[i + 6] == 'i' && p[i + 7] == 't' && p[i + 8] == 'y') { sink.puts("-Infinity"); i += 9; st = ST_AFTER; continue; }
          uint32_t j = i + (c == '-' ? 1u : 0u);
          if (j >= len || p[j] < '0' || p[j] > '9') { err = JE_VALUE; at = i; goto fail; }
          if (p[j] == '0') ++j; else while (j < len && p[j] >= '0' && p[j] <= '9') ++j;
          const uint32_t int_end = j;
          bool is_float = false;
          if (j + 1 < len && p[j] == '.' && p[j + 1] >= '0' && p[j + 1] <= '9') {
            j += 2;
            while (j < len && p[j] >= '0' && p[j] <= '9') ++j;
            is_float = true;
          }
          if (j < len && (p[j] == 'e' || p[j] == 'E')) {
            uint32_t k = j + 1;
            if (k < len && (p[k] == '+' || p[k] == '-')) ++k;
            if (k < len && p[k] >= '0' && p[k] <= '9') {
              while (k < len && p[k] >= '0' && p[k] <= '9') ++k;
              j = k; is_float = true;
            }
          }
          if (!is_float) {
            const uint32_t nd = int_end - i - (c == '-' ? 1u : 0u);
            if (nd > 4300) { res->status = JS_UNSUPPORTED; return; }     // int() refuses: ValueError, another message
            if (c == '-' && nd == 1 && p[i + 1] == '0') sink.put('0');    // int("-0") is 0
            else sink.copy(p + i, int_end - i);
          } else if (!json_float(p + i, j - i, &sink)) { res->status = JS_UNSUPPORTED; return; }
          i = j; st = ST_AFTER;
        } else {
          const char* w = c == 'n' ? "null" : c == 't' ? "true" : c == 'f' ? "false" : c == 'N' ? "NaN" : c == 'I' ? "Infinity" : nullptr;
          uint32_t k = 0;
          if (w) while (w[k] && i + k < len && p[i + k] == (uint8_t)w[k]) ++k;
          if (!w || w[k]) { err = JE_VALUE; at = i; goto fail; }
          sink.puts(w); i += k; st = ST_AFTER;
        }
        continue;
      }
      if (st == ST_KEY) {
        // p[i] == '"' (checked by whoever set ST_KEY)
        const uint32_t k_begin = i + 1;
        const uint32_t e = json_string(p, len, i, &sink, &err, &at, &lone);
        if (!e) goto fail;
        const uint32_t k_end = e - 1;
        // duplicate keys of this object: dict semantics are not a streaming transformation -> declined
        {
          const uint32_t first = key0[depth - 1];
          bool esc = false;
          for (uint32_t q = k_begin; q < k_end; ++q) if (p[q] == '\\') esc = true;
          if (n_keys > first) {
            if (!keys) { res->status = JS_UNSUPPORTED; return; }
            for (uint32_t q = first; q < n_keys; ++q) {
              const uint32_t b0 = keys[2 * q] & 0x7FFFFFFFu, b1 = keys[2 * q + 1];
              if (json_keys_equal(p, b0, b1, k_begin, k_end, esc || (keys[2 * q] >> 31))) { res->status = JS_UNSUPPORTED; return; }
            }
          }
          if (keys) { keys[2 * n_keys] = k_begin | (esc ? 0x80000000u : 0u); keys[2 * n_keys + 1] = k_end; }
          ++n_keys;
        }
        i = e;
        while (i < len && json_ws(p[i])) ++i;
        if (i >= len || p[i] != ':') { err = JE_COLON; at = i; goto fail; }
        ++i;
        sink.put(':'); sink.put(' ');
        st = ST_VALUE;
        continue;
      }
      // ST_AFTER: a value just ended
      if (depth == 0) {
        if (i != len) { err = JE_EXTRA; at = i; goto fail; }
        break;
      }
      {
        const bool obj = (is_obj[(depth - 1) >> 6] >> ((depth - 1) & 63)) & 1ull;
        if (i < len && p[i] == ',') {
          ++i;
          while (i < len && json_ws(p[i])) ++i;
          if (obj && (i >= len || p[i] != '"')) { err = JE_PROPNAME; at = i; goto fail; }
          sink.put(','); sink.newline_indent(depth);
          st = obj ? ST_KEY : ST_VALUE;
        } else if (i < len && p[i] == (obj ? '}' : ']')) {
          ++i;
          --depth;
          n_keys = key0[depth];
          sink.newline_indent(depth); sink.put(obj ? '}' : ']');
        } else { err = JE_COMMA; at = i; goto fail; }
      }
    }
    if (lone) { res->status = JS_UNSUPPORTED; return; }
    res->out_bytes = sink.n;
    return;
  }
fail:
  res->status = JS_INVALID; res->err = err;
  json_locate(p, len, at, res);
}

// ---- the same walk as ONE loop over the bytes -----------------------------------------------------------
// Every iteration takes one byte (an escape sequence, a literal or the end of a number take a bounded few
// more) and switches on the state, so the 32 documents of a warp, each somewhere else in the grammar, run
// the iteration together and part only inside it.  Same results as json_pretty_nested, byte for byte and
// error for error (tests/test_json.py runs both).
RFQ_HD void json_pretty(const uint8_t* p, uint32_t len, uint8_t* out, uint32_t* keys, JsonResult* res) {
  res->status = JS_OK; res->err = JE_NONE; res->pos = 0; res->line = 1; res->col = 1; res->out_bytes = 0;
  JsonSink sink{out, 0};
  uint32_t err = 0, at = 0;
  bool unsup = false;
  {
    struct Acc { const uint8_t* p; uint32_t n; RFQ_HD uint32_t operator()(int64_t k) const { return (k >= 0 && k < (int64_t)n) ? p[k] : 0u; } };
    const Acc B{p, len};
    for (uint32_t k = 0; k < len;) {
      const uint32_t c = p[k];
      if (c < 0x80) { ++k; continue; }
      const int l = u8_lead_len(c);
      if (l == 0 || !u8_seq_valid(B, (int64_t)k)) { unsup = true; break; }
      k += (uint32_t)l;
    }
  }
  {
    enum { S_VALUE, S_AFTER, S_OBJ_OPEN, S_ARR_OPEN, S_KEY_EXPECT, S_COLON, S_STR, S_NUM_MINUS, S_NUM_ZERO, S_NUM_INT,
           S_NUM_FRAC, S_NUM_EXP, S_DONE, S_FAIL };
    const uint32_t END = 0x100u;                       // the byte after the last one
    uint64_t is_obj[(RFQ_JSON_MAX_DEPTH + 63) / 64] = {0};
    uint32_t key0[RFQ_JSON_MAX_DEPTH];
    uint32_t depth = 0, n_keys = 0, i = 0, st = S_VALUE;
    if (unsup) st = S_FAIL;
    else if (len >= 3 && p[0] == 0xEF && p[1] == 0xBB && p[2] == 0xBF) { err = JE_BOM; at = 0; st = S_FAIL; }
    uint32_t str_begin = 0, num_begin = 0;
    bool lone = false, is_key = false, key_esc = false, num_float = false;
    // On the GPU the loop is left by the whole warp at once: the vote brings the 32 documents back together
    // after every byte (without it the compiler threads each state straight into the next and the lanes never
    // meet again).  All 32 lanes of a warp must therefore call this function together.
    while (RFQ_JSON_ANY(st < S_DONE)) {
      if (st < S_DONE) {
      const uint32_t c = i < len ? p[i] : END;
      switch (st) {
        case S_VALUE: {
          if (json_ws(c)) { ++i; break; }
          if (c == '"') { str_begin = i; is_key = false; sink.put('"'); ++i; st = S_STR; break; }
          if (c == '{') { ++i; st = S_OBJ_OPEN; break; }
          if (c == '[') { ++i; st = S_ARR_OPEN; break; }
          if (c == '-') {
            if (i + 9 <= len && p[i + 1] == 'I' && p[i + 2] == 'n' && p[i + 3] == 'f' && p[i + 4] == 'i' && p[i + 5] == 'n' &&
                p[i + 6] == 'i' && p[i + 7] == 't' && p[i + 8] == 'y') { sink.puts("-Infinity"); i += 9; st = S_AFTER; break; }
            num_begin = i; num_float = false; ++i; st = S_NUM_MINUS; break;
          }
          if (c >= '0' && c <= '9') { num_begin = i; num_float = false; ++i; st = c == '0' ? S_NUM_ZERO : S_NUM_INT; break; }
          {
            const char* w = c == 'n' ? "null" : c == 't' ? "true" : c == 'f' ? "false" : c == 'N' ? "NaN" : c == 'I' ? "Infinity" : nullptr;
            uint32_t k = 0;
            if (w) while (w[k] && i + k < len && p[i + k] == (uint8_t)w[k]) ++k;
            if (!w || w[k]) { err = JE_VALUE; at = i; st = S_FAIL; goto next; }
            sink.puts(w); i += k; st = S_AFTER;
          }
          break;
        }
        case S_OBJ_OPEN: case S_ARR_OPEN: {
          if (json_ws(c)) { ++i; break; }
          const bool obj = st == S_OBJ_OPEN;
          if (c == (obj ? (uint32_t)'}' : (uint32_t)']')) { sink.put(obj ? '{' : '['); sink.put(c); ++i; st = S_AFTER; break; }
          if (depth >= RFQ_JSON_MAX_DEPTH) { unsup = true; st = S_FAIL; goto next; }
          if (obj) is_obj[depth >> 6] |= 1ull << (depth & 63); else is_obj[depth >> 6] &= ~(1ull << (depth & 63));
          key0[depth] = n_keys;
          ++depth;
          sink.put(obj ? '{' : '['); sink.newline_indent(depth);
          if (!obj) { st = S_VALUE; break; }                       // the same byte again, as a value
          if (c != '"') { err = JE_PROPNAME; at = i; st = S_FAIL; goto next; }
          str_begin = i; is_key = true; key_esc = false; sink.put('"'); ++i; st = S_STR;
          break;
        }
        case S_KEY_EXPECT: {
          if (json_ws(c)) { ++i; break; }
          if (c != '"') { err = JE_PROPNAME; at = i; st = S_FAIL; goto next; }
          str_begin = i; is_key = true; key_esc = false; sink.put('"'); ++i; st = S_STR;
          break;
        }
        case S_COLON: {
          if (json_ws(c)) { ++i; break; }
          if (c != ':') { err = JE_COLON; at = i; st = S_FAIL; goto next; }
          ++i; sink.put(':'); sink.put(' '); st = S_VALUE;
          break;
        }
        case S_STR: {
          if (c == END) { err = JE_UNTERMINATED; at = str_begin; st = S_FAIL; goto next; }
          if (c == '"') {
            sink.put('"'); ++i;
            if (!is_key) { st = S_AFTER; break; }
            // a key ends: equal keys inside one object are declined (dict semantics are not a streaming transformation)
            const uint32_t k_begin = str_begin + 1, k_end = i - 1, first = key0[depth - 1];
            if (n_keys > first) {
              if (!keys) { unsup = true; st = S_FAIL; goto next; }
              for (uint32_t q = first; q < n_keys; ++q) {
                const uint32_t b0 = keys[2 * q] & 0x7FFFFFFFu, b1 = keys[2 * q + 1];
                if (json_keys_equal(p, b0, b1, k_begin, k_end, key_esc || (keys[2 * q] >> 31))) { unsup = true; st = S_FAIL; goto next; }
              }
            }
            if (keys) { keys[2 * n_keys] = k_begin | (key_esc ? 0x80000000u : 0u); keys[2 * n_keys + 1] = k_end; }
            ++n_keys;
            st = S_COLON;
            break;
          }
          if (c < 0x20) { err = JE_CTRL; at = i; st = S_FAIL; goto next; }
          if (c != '\\') { sink.put(c); ++i; break; }
          // an escape: two bytes, six for \uXXXX, twelve for a surrogate pair
          key_esc = true;
          if (i + 1 >= len) { err = JE_UNTERMINATED; at = str_begin; st = S_FAIL; goto next; }
          const uint32_t e = p[i + 1];
          uint32_t cp;
          if (e != 'u') {
            switch (e) {
              case '"': cp = '"'; break;
              case '\\': cp = '\\'; break;
              case '/': cp = '/'; break;
              case 'b': cp = 8; break;
              case 'f': cp = 12; break;
              case 'n': cp = 10; break;
              case 'r': cp = 13; break;
              case 't': cp = 9; break;
              default: err = JE_ESCAPE; at = i; st = S_FAIL; goto next;
            }
            i += 2;
          } else {
            if (i + 6 >= len) { err = JE_UESCAPE; at = i + 1; st = S_FAIL; goto next; }   // (_json.c wants a character after the four digits)
            cp = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int h = json_hex(p[i + 2 + k]);
              if (h < 0) { err = JE_UESCAPE; at = i + 1; st = S_FAIL; goto next; }
              cp = cp * 16 + (uint32_t)h;
            }
            i += 6;
            if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 < len && p[i] == '\\' && p[i + 1] == 'u') {
              uint32_t c2 = 0;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const int h = json_hex(p[i + 2 + k]);
                if (h < 0) { err = JE_UESCAPE; at = i + 1; st = S_FAIL; goto next; }
                c2 = c2 * 16 + (uint32_t)h;
              }
              if (c2 >= 0xDC00 && c2 <= 0xDFFF) { cp = 0x10000 + (((cp - 0xD800) << 10) | (c2 - 0xDC00)); i += 6; }
            }
            if (cp >= 0xD800 && cp <= 0xDFFF) lone = true;
          }
          // encode_basestring
          if (cp == '"' || cp == '\\') { sink.put('\\'); sink.put(cp); }
          else if (cp == 10) { sink.put('\\'); sink.put('n'); }
          else if (cp == 13) { sink.put('\\'); sink.put('r'); }
          else if (cp == 9) { sink.put('\\'); sink.put('t'); }
          else if (cp == 8) { sink.put('\\'); sink.put('b'); }
          else if (cp == 12) { sink.put('\\'); sink.put('f'); }
          else if (cp < 0x20) {
            sink.put('\\'); sink.put('u'); sink.put('0'); sink.put('0');
            sink.put("0123456789abcdef"[cp >> 4]); sink.put("0123456789abcdef"[cp & 15]);
          }
          else if (cp < 0x80) sink.put(cp);
          else if (cp < 0x800) { sink.put(0xC0 | (cp >> 6)); sink.put(0x80 | (cp & 0x3F)); }
          else if (cp < 0x10000) { sink.put(0xE0 | (cp >> 12)); sink.put(0x80 | ((cp >> 6) & 0x3F)); sink.put(0x80 | (cp & 0x3F)); }
          else { sink.put(0xF0 | (cp >> 18)); sink.put(0x80 | ((cp >> 12) & 0x3F)); sink.put(0x80 | ((cp >> 6) & 0x3F)); sink.put(0x80 | (cp & 0x3F)); }
          break;
        }
        case S_NUM_MINUS: {
          if (c >= '0' && c <= '9') { ++i; st = c == '0' ? S_NUM_ZERO : S_NUM_INT; break; }
          err = JE_VALUE; at = num_begin; st = S_FAIL; goto next;
        }
        case S_NUM_ZERO: case S_NUM_INT: case S_NUM_FRAC: case S_NUM_EXP: {
          if (c >= '0' && c <= '9' && st != S_NUM_ZERO) { ++i; break; }
          if (c == '.' && (st == S_NUM_ZERO || st == S_NUM_INT) && i + 1 < len && p[i + 1] >= '0' && p[i + 1] <= '9') {
            i += 2; num_float = true; st = S_NUM_FRAC; break;
          }
          if ((c == 'e' || c == 'E') && st != S_NUM_EXP) {
            uint32_t k = i + 1;
            if (k < len && (p[k] == '+' || p[k] == '-')) ++k;
            if (k < len && p[k] >= '0' && p[k] <= '9') { i = k + 1; num_float = true; st = S_NUM_EXP; break; }
          }
          // the number ends before this byte
          if (!num_float) {
            const bool neg = p[num_begin] == '-';
            const uint32_t nd = i - num_begin - (neg ? 1u : 0u);
            if (nd > 4300) { unsup = true; st = S_FAIL; goto next; }     // int() refuses: ValueError, another message
            if (neg && nd == 1 && p[num_begin + 1] == '0') sink.put('0');  // int("-0") is 0
            else sink.copy(p + num_begin, i - num_begin);
          } else if (!json_float(p + num_begin, i - num_begin, &sink)) { unsup = true; st = S_FAIL; goto next; }
          st = S_AFTER;
          break;                                                          // the same byte again, after a value
        }
        case S_AFTER: {
          if (json_ws(c)) { ++i; break; }
          if (depth == 0) {
            if (c != END) { err = JE_EXTRA; at = i; st = S_FAIL; goto next; }
            st = S_DONE;
            break;
          }
          const bool obj = (is_obj[(depth - 1) >> 6] >> ((depth - 1) & 63)) & 1ull;
          if (c == ',') { ++i; sink.put(','); sink.newline_indent(depth); st = obj ? S_KEY_EXPECT : S_VALUE; break; }
          if (c == (obj ? (uint32_t)'}' : (uint32_t)']')) {
            ++i; --depth; n_keys = key0[depth];
            sink.newline_indent(depth); sink.put(c);
            break;
          }
          err = JE_COMMA; at = i; st = S_FAIL; goto next;
        }
        default: st = S_DONE; break;
      }
      next:;
      }
    }
    if (unsup || (st == S_DONE && lone)) { res->status = JS_UNSUPPORTED; return; }
    if (st == S_DONE) { res->out_bytes = sink.n; return; }
  }
  res->status = JS_INVALID; res->err = err;
  json_locate(p, len, at, res);
}

}  // namespace rfq
