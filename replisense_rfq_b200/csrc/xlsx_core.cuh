// Sheet-XML / sharedStrings decoding primitives (host/device), restating what openpyxl 3.1.5's
// read-only reader + pandas' `_convert_cell` do per element (third-party, un-vendored:
// openpyxl/worksheet/_reader.py WorkSheetParser.parse_row/parse_cell + _cast_number,
// openpyxl/reader/strings.py read_string_table, openpyxl/cell/text.py Text.content;
// pandas/io/excel/_openpyxl.py:596-612; reached from app/file_parser.py:225).
#pragma once
#include "cell_format.cuh"
#include "cell_parse.cuh"

namespace rfq {

// ---- decoded text of an XML character-data span: entities + line-end normalisation -----------
struct XmlIter {
  const uint8_t* p;
  const uint8_t* e;
  uint32_t pend;   // pending UTF-8 bytes of a decoded code point (little endian), npend of them
  int npend;
  bool bad;        // unknown / malformed entity (expat would raise)
  RFQ_HD XmlIter(const uint8_t* b, uint32_t n) : p(b), e(b + n), pend(0), npend(0), bad(false) {}
  RFQ_HD void push_cp(uint32_t cp) {
    if (cp < 0x80) { pend = cp; npend = 1; }
    else if (cp < 0x800) { pend = (0xC0 | (cp >> 6)) | ((0x80 | (cp & 0x3F)) << 8); npend = 2; }
    else if (cp < 0x10000) {
      pend = (0xE0 | (cp >> 12)) | ((0x80 | ((cp >> 6) & 0x3F)) << 8) | ((0x80 | (cp & 0x3F)) << 16);
      npend = 3;
    } else {
      pend = (0xF0 | (cp >> 18)) | ((0x80 | ((cp >> 12) & 0x3F)) << 8) |
             ((0x80 | ((cp >> 6) & 0x3F)) << 16) | ((0x80u | (cp & 0x3F)) << 24);
      npend = 4;
    }
  }
  RFQ_HD int next() {
    if (npend) { int c = (int)(pend & 0xFF); pend >>= 8; --npend; return c; }
    if (p >= e) return -1;
    uint32_t c = *p++;
    if (c == '\r') { if (p < e && *p == '\n') ++p; return '\n'; }
    if (c != '&') return (int)c;
    // entity: up to ';'
    const uint8_t* q = p;
    while (q < e && *q != ';' && q - p < 12) ++q;
    if (q >= e || *q != ';') { bad = true; return '&'; }
    uint32_t n = (uint32_t)(q - p);
    uint32_t cp = 0;
    if (n >= 2 && p[0] == '#') {
      if (p[1] == 'x') { for (uint32_t i = 2; i < n; ++i) { uint32_t d = p[i]; d = d <= '9' ? d - '0' : (d | 0x20) - 'a' + 10; if (d > 15) bad = true; cp = cp * 16 + d; } if (n == 2) bad = true; }
      else { for (uint32_t i = 1; i < n; ++i) { uint32_t d = p[i] - '0'; if (d > 9) bad = true; cp = cp * 10 + d; } }
    } else if (n == 3 && p[0] == 'a' && p[1] == 'm' && p[2] == 'p') cp = '&';
    else if (n == 2 && p[0] == 'l' && p[1] == 't') cp = '<';
    else if (n == 2 && p[0] == 'g' && p[1] == 't') cp = '>';
    else if (n == 4 && p[0] == 'q' && p[1] == 'u' && p[2] == 'o' && p[3] == 't') cp = '"';
    else if (n == 4 && p[0] == 'a' && p[1] == 'p' && p[2] == 'o' && p[3] == 's') cp = '\'';
    else bad = true;
    p = q + 1;
    push_cp(cp);
    int r = (int)(pend & 0xFF); pend >>= 8; --npend;
    return r;
  }
};

// ---- tags ---------------------------------------------------------------------------------------
enum { XT_NONE = 0, XT_ROW, XT_C, XT_V, XT_IS, XT_T, XT_SI, XT_RPH, XT_RPH_END, XT_SI_END, XT_C_END,
       XT_ROW_END, XT_F, XT_BANG, XT_R, XT_OTHER };

RFQ_HD bool xml_name_end(uint32_t c) { return c == ' ' || c == '>' || c == '/' || c == '\t' || c == '\n' || c == '\r'; }

// classify the tag that starts at p[0] == '<' (p may be read up to 7 bytes ahead, bounded by e)
RFQ_HD int xml_classify(const uint8_t* p, const uint8_t* e) {
  uint32_t b[8];
#ifdef __CUDA_ARCH__
  {   // the 8 bytes in two aligned loads (the arena carries slack past the last document)
    const unsigned long long* q = (const unsigned long long*)((uintptr_t)p & ~(uintptr_t)7);
    const uint32_t sh = (uint32_t)((uintptr_t)p & 7u) * 8u;
    unsigned long long x = __ldg(q);
    if (sh) x = (x >> sh) | (__ldg(q + 1) << (64u - sh));
    const size_t rem = (size_t)(e - p);
    if (rem < 8) x &= (1ull << (8 * rem)) - 1ull;
#pragma unroll
    for (int i = 0; i < 8; ++i) b[i] = (uint32_t)(x >> (8 * i)) & 0xFFu;
  }
#else
  for (int i = 0; i < 8; ++i) b[i] = (p + i < e) ? p[i] : 0u;
#endif
  if (b[1] == '!') return XT_BANG;
  if (b[1] == '/') {
    if (b[2] == 'c' && b[3] == '>') return XT_C_END;
    if (b[2] == 'r' && b[3] == 'o' && b[4] == 'w' && b[5] == '>') return XT_ROW_END;
    if (b[2] == 's' && b[3] == 'i' && b[4] == '>') return XT_SI_END;
    if (b[2] == 'r' && b[3] == 'P' && b[4] == 'h' && b[5] == '>') return XT_RPH_END;
    return XT_OTHER;
  }
  if (b[1] == 'c' && xml_name_end(b[2])) return XT_C;
  if (b[1] == 'v' && xml_name_end(b[2])) return XT_V;
  if (b[1] == 't' && xml_name_end(b[2])) return XT_T;
  if (b[1] == 'f' && xml_name_end(b[2])) return XT_F;
  if (b[1] == 'r' && xml_name_end(b[2])) return XT_R;
  if (b[1] == 'r' && b[2] == 'o' && b[3] == 'w' && xml_name_end(b[4])) return XT_ROW;
  if (b[1] == 'i' && b[2] == 's' && xml_name_end(b[3])) return XT_IS;
  if (b[1] == 's' && b[2] == 'i' && xml_name_end(b[3])) return XT_SI;
  if (b[1] == 'r' && b[2] == 'P' && b[3] == 'h' && xml_name_end(b[4])) return XT_RPH;
  return XT_OTHER;
}

// first position in [p, e) holding byte c, or e: eight bytes per step on the device (two aligned
// loads; the arena carries slack past its last document)
RFQ_HD const uint8_t* find_byte(const uint8_t* p, const uint8_t* e, uint8_t c) {
#ifdef __CUDA_ARCH__
  const unsigned long long ones = 0x0101010101010101ull, lo7 = 0x7F7F7F7F7F7F7F7Full;
  while (p < e) {
    const unsigned long long* q = (const unsigned long long*)((uintptr_t)p & ~(uintptr_t)7);
    const uint32_t sh = (uint32_t)((uintptr_t)p & 7u) * 8u;
    unsigned long long x = __ldg(q);
    if (sh) x = (x >> sh) | (__ldg(q + 1) << (64u - sh));
    const unsigned long long y = x ^ (ones * c);
    unsigned long long hit = ~(((y & lo7) + lo7) | y | lo7);   // 0x80 in every byte equal to c
    const size_t rem = (size_t)(e - p);
    if (rem < 8) hit &= (1ull << (8 * rem)) - 1ull;
    if (hit) return p + ((__ffsll((long long)hit) - 1) >> 3);
    p += 8;
  }
  return e;
#else
  while (p < e && *p != c) ++p;
  return p;
#endif
}

// find attribute ` name="value"` inside the tag [p, e) (e = position of the closing '>').
// name is 1 or 2 chars (r, t, s, ht ...). Returns value span; false if absent.
RFQ_HD bool xml_attr(const uint8_t* p, const uint8_t* e, char n0, const uint8_t** vb, uint32_t* vn) {
  for (const uint8_t* q = p; q + 3 < e; ++q) {
    if ((q[0] == ' ' || q[0] == '\t' || q[0] == '\n' || q[0] == '\r') && q[1] == (uint8_t)n0 && q[2] == '=' &&
        (q[3] == '"' || q[3] == '\'')) {
      uint8_t quote = q[3];
      const uint8_t* s = q + 4;
      const uint8_t* t = s;
      while (t < e && *t != quote) ++t;
      *vb = s; *vn = (uint32_t)(t - s);
      return true;
    }
    if (q[0] == '"' || q[0] == '\'') {  // skip a quoted value so its content is not mistaken for a name
      uint8_t quote = q[0];
      ++q;
      while (q < e && *q != quote) ++q;
    }
  }
  return false;
}

// the r=, t= and s= attributes of a <c ...> tag in ONE pass over its bytes (first occurrence of each,
// quoted values skipped): what three xml_attr calls return
struct XAttrRTS { const uint8_t* vb[3]; uint32_t vn[3]; uint32_t have; };   // index 0 r, 1 t, 2 s
RFQ_HD void xml_attrs_rts(const uint8_t* p, const uint8_t* e, XAttrRTS* o) {
  o->have = 0;
  o->vb[0] = o->vb[1] = o->vb[2] = p; o->vn[0] = o->vn[1] = o->vn[2] = 0;
  const uint8_t* q = p;
  while (q + 3 < e) {
    const uint8_t c = *q;
    if (c == '"' || c == '\'') {            // a value whose name was not single-letter: skip it
      ++q;
      while (q < e && *q != c) ++q;
      ++q;
      continue;
    }
    if ((c == ' ' || c == '\t' || c == '\n' || c == '\r') && q[2] == '=' && (q[3] == '"' || q[3] == '\'')) {
      const uint8_t n0 = q[1], quote = q[3];
      const int k = n0 == 'r' ? 0 : n0 == 't' ? 1 : n0 == 's' ? 2 : -1;
      const uint8_t* s = q + 4;
      const uint8_t* t = s;
      while (t < e && *t != quote) ++t;
      if (k >= 0 && !((o->have >> k) & 1u)) { o->vb[k] = s; o->vn[k] = (uint32_t)(t - s); o->have |= 1u << k; }
      q = t + 1;
      continue;
    }
    ++q;
  }
}

// "AB12" -> column 28 (1-based), row 12; false if malformed
RFQ_HD bool a1_parse(const uint8_t* s, uint32_t n, uint32_t* col, uint32_t* row) {
  uint32_t i = 0, c = 0, r = 0;
  if (i < n && s[i] == '$') ++i;
  uint32_t nl = 0;
  while (i < n && ((s[i] | 0x20) >= 'a' && (s[i] | 0x20) <= 'z')) { c = c * 26 + ((s[i] | 0x20) - 'a' + 1); ++i; ++nl; }
  if (nl == 0 || nl > 3) return false;
  if (i < n && s[i] == '$') ++i;
  uint32_t nd = 0;
  while (i < n && s[i] >= '0' && s[i] <= '9') { r = r * 10 + (s[i] - '0'); ++i; ++nd; if (nd > 9) return false; }
  if (nd == 0 || i != n) return false;
  *col = c; *row = r;
  return true;
}

// ---- correctly rounded decimal -> double (what Python's float(str) gives) -------------------
// digits/exponent from a plain numeric literal [sign] digits [. digits] [(e|E) [sign] digits]
RFQ_HD_COLD bool strtod_exact(const uint8_t* s, uint32_t n, double* out) {
  uint32_t i = 0;
  bool neg = false;
  if (i < n && (s[i] == '-' || s[i] == '+')) { neg = s[i] == '-'; ++i; }
  const uint32_t dig0 = i;
  // pass 1: up to 19 significant digits in a u64, total significant digits, decimal exponent
  uint64_t w = 0; int nd = 0, nsig = 0; bool any = false, truncated = false;
  long long e10 = 0;
  for (; i < n && s[i] >= '0' && s[i] <= '9'; ++i) {
    any = true;
    if (nsig || s[i] != '0') ++nsig;
    if (nd < 19) { w = w * 10 + (s[i] - '0'); if (w) ++nd; }
    else { ++e10; if (s[i] != '0') truncated = true; }
  }
  if (i < n && s[i] == '.') {
    ++i;
    for (; i < n && s[i] >= '0' && s[i] <= '9'; ++i) {
      any = true;
      if (nsig || s[i] != '0') ++nsig;
      if (nd < 19) { w = w * 10 + (s[i] - '0'); if (w) ++nd; --e10; }
      else if (s[i] != '0') truncated = true;
    }
  }
  const uint32_t dig1 = i;
  if (!any) return false;
  if (i < n && (s[i] == 'e' || s[i] == 'E')) {
    ++i;
    bool eneg = false;
    if (i < n && (s[i] == '-' || s[i] == '+')) { eneg = s[i] == '-'; ++i; }
    long long ev = 0; int ned = 0;
    for (; i < n && s[i] >= '0' && s[i] <= '9'; ++i) { if (ev < 100000) ev = ev * 10 + (s[i] - '0'); ++ned; }
    if (!ned) return false;
    e10 += eneg ? -ev : ev;
  }
  if (i != n) return false;
  double z;
  if (w == 0) { *out = neg ? -0.0 : 0.0; return true; }
  if (w < (1ull << 53) && !truncated && e10 >= -22 && e10 <= 22) {
    z = (double)w;
    z = e10 >= 0 ? RFQ_DMUL(z, RFQ_POW10(e10)) : RFQ_DDIV(z, RFQ_POW10(-e10));
    *out = neg ? -z : z;
    return true;
  }
  if (!truncated && e10 < 0 && e10 >= -19) {
    // up to 19 digits over a power of ten that fits 64 bits (the 16-17 digit literals writers emit):
    // correctly rounded quotient with one 128-bit division instead of the big-integer path
    uint64_t D = 1;
    for (long long q = 0; q < -e10; ++q) D *= 10ull;
    int bw = 0, bd = 0;
    for (uint64_t t = w; t; t >>= 1) ++bw;
    for (uint64_t t = D; t; t >>= 1) ++bd;
    int sft = 56 + bd - bw;                    // quotient gets 56..57 significant bits
    if (sft < 0) sft = 0;
    unsigned __int128 N = (unsigned __int128)w << sft;
    uint64_t q = (uint64_t)(N / D);
    bool sticky = (N % D) != 0;
    int nb = 0;
    for (uint64_t t = q; t; t >>= 1) ++nb;
    int drop = nb - 53;                        // >= 3 whenever bits were shifted in
    uint64_t mant = q;
    if (drop > 0) {
      uint64_t rem = q & ((1ull << drop) - 1ull), half = 1ull << (drop - 1);
      mant = q >> drop;
      if (rem > half || (rem == half && (sticky || (mant & 1ull)))) ++mant;
    } else {
      drop = 0;                                // short exact quotient (sticky implies nb >= 56, so not here)
    }
    z = ldexp((double)mant, drop - sft);       // mant <= 2^53: exact; exponent far inside the normal range
    *out = neg ? -z : z;
    return true;
  }
  if (e10 > 310) { *out = neg ? -__builtin_huge_val() : __builtin_huge_val(); return true; }
  if (e10 < -345) { *out = neg ? -0.0 : 0.0; return true; }
  // exact digits for the slow path: all significant digits (capped at 40 + sticky) as a big integer
  Big DIG; DIG.set_u64(0);
  long long eb = 0;      // value = DIG * 10^eb
  bool sticky = false;
  {
    int taken = 0; bool seen_point = false, started = false;
    for (uint32_t q = dig0; q < dig1; ++q) {
      if (s[q] == '.') { seen_point = true; continue; }
      uint32_t d = s[q] - '0';
      if (!started && d == 0) { if (seen_point) --eb; continue; }
      started = true;
      if (taken < 40) { DIG.mul_small(10); DIG.add_small(d); ++taken; if (seen_point) --eb; }
      else { if (!seen_point) ++eb; if (d) sticky = true; }
    }
  }
  // e10 currently = (exponent field) + adjustments for the 19-digit view; recompute for DIG:
  // exponent field = e10 - (adjustment of pass 1). Derive it by re-reading the field.
  {
    long long field = 0;
    uint32_t q = dig1;
    if (q < n && (s[q] == 'e' || s[q] == 'E')) {
      ++q; bool eneg = false;
      if (q < n && (s[q] == '-' || s[q] == '+')) { eneg = s[q] == '-'; ++q; }
      long long ev = 0;
      for (; q < n; ++q) if (ev < 100000) ev = ev * 10 + (s[q] - '0');
      field = eneg ? -ev : ev;
    }
    eb += field;
  }
  {
    double a = (double)w;
    long long ee = e10;
    while (ee > 0) { int k = ee > 300 ? 300 : (int)ee; a *= RFQ_POW10(k); ee -= k; }
    while (ee < 0) { int k = ee < -300 ? 300 : (int)-ee; a /= RFQ_POW10(k); ee += k; }
    z = a;
  }
  if (dbl_is_inf(z)) z = 1.7976931348623157e308;
  if (z == 0.0) z = 4.9406564584124654e-324;
  for (int iter = 0; iter < 8; ++iter) {
    uint64_t m; int k; bool ng;
    dbl_decompose(z, &m, &k, &ng);
    if (m == 0) break;
    // Everything at scale 2^(k-2): z = 4m, upper midpoint 4m+2, lower midpoint 4m-2 (4m-1 when z is
    // the smallest value of its binade, where the gap below is half as wide).
    bool boundary = (m == (1ull << 52)) && k > -1074;
    int k2 = k - 2;
    Big D = DIG;
    if (eb > 0) D.mul_pow10((int)eb);
    if (k2 < 0) D.shl(-k2);
    Big hi, lo;
    hi.set_u64(4 * m + 2);
    lo.set_u64(boundary ? 4 * m - 1 : 4 * m - 2);
    if (eb < 0) { hi.mul_pow10((int)-eb); lo.mul_pow10((int)-eb); }
    if (k2 > 0) { hi.shl(k2); lo.shl(k2); }
    int chi = big_cmp(D, hi);
    bool up = chi > 0 || (chi == 0 && (sticky || (m & 1)));
    int clo = big_cmp(D, lo);
    bool down = !up && (clo < 0 || (clo == 0 && !sticky && (m & 1)));
    if (!up && !down) break;
    uint64_t bts = dbl_bits(z);
    bts = up ? bts + 1 : bts - 1;
#ifdef __CUDA_ARCH__
    z = __longlong_as_double((long long)bts);
#else
    __builtin_memcpy(&z, &bts, 8);
#endif
    if (dbl_is_inf(z)) break;
  }
  *out = neg ? -z : z;
  return true;
}

}  // namespace rfq
