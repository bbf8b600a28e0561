// Number → text exactly as `DataFrame.to_string(index=False)` prints it (app/file_parser.py:305,
// :243; pandas/io/formats/format.py `_IntArrayFormatter` f"{x:d}", `FloatArrayFormatter`
// "{value:.6f}" / "{value:.6e}" + `_trim_zeros_float`, `_GenericArrayFormatter` f"{x: .6f}" +
// `_trim_zeros_single_float`).  Python's float formatting is correctly rounded (half-even on the
// exact binary value), so everything here is exact integer arithmetic: 128-bit for the fixed
// form in the range that can actually be printed fixed, a small big-integer for the rest.
#pragma once
#include "rfq_common.cuh"

namespace rfq {

RFQ_HD int ndigits_u64(uint64_t v) {
  int n = 1;
  if (v >= 10000000000000000ull) { v /= 10000000000000000ull; n += 16; }
  uint32_t x;
  if (v >= 100000000ull) { x = (uint32_t)(v / 100000000ull); n += 8; } else x = (uint32_t)v;
  if (x >= 10000u) { x /= 10000u; n += 4; }
  if (x >= 100u) { x /= 100u; n += 2; }
  return n + (x >= 10u);
}

// writes decimal digits of v ending at end (exclusive); returns pointer to first digit
RFQ_HD uint8_t* write_u64_rev(uint64_t v, uint8_t* end) {
  while (v > 0xFFFFFFFFull) { *--end = (uint8_t)('0' + v % 10); v /= 10; }
  uint32_t x = (uint32_t)v;   // 32-bit divisions from here on
  do { *--end = (uint8_t)('0' + x % 10u); x /= 10u; } while (x);
  return end;
}

RFQ_HD int int64_len(int64_t v) {
  uint64_t mag = v < 0 ? 0ull - (uint64_t)v : (uint64_t)v;
  return ndigits_u64(mag) + (v < 0);
}

// ---- tiny big-integer (little endian 32-bit words) ---------------------------------------------
struct Big {
  static const int W = 40;
  uint32_t w[W];
  int n;  // used words (no leading zeros; n == 0 means zero)
  RFQ_HD void set_u64(uint64_t v) {
    n = 0;
    for (int i = 0; i < W; ++i) w[i] = 0;
    if (v) { w[0] = (uint32_t)v; w[1] = (uint32_t)(v >> 32); n = w[1] ? 2 : 1; }
  }
  RFQ_HD void mul_small(uint32_t m) {
    uint64_t carry = 0;
    for (int i = 0; i < n; ++i) {
      uint64_t t = (uint64_t)w[i] * m + carry;
      w[i] = (uint32_t)t; carry = t >> 32;
    }
    if (carry && n < W) w[n++] = (uint32_t)carry;
  }
  RFQ_HD void add_small(uint32_t a) {
    uint64_t carry = a;
    for (int i = 0; i < n && carry; ++i) { uint64_t t = (uint64_t)w[i] + carry; w[i] = (uint32_t)t; carry = t >> 32; }
    if (carry && n < W) w[n++] = (uint32_t)carry;
  }
  RFQ_HD void mul_pow10(int p) {
    while (p >= 9) { mul_small(1000000000u); p -= 9; }
    uint32_t m = 1;
    while (p-- > 0) m *= 10u;
    if (m > 1) mul_small(m);
  }
  RFQ_HD void shl(int bits) {
    if (n == 0 || bits == 0) return;
    int ws = bits >> 5, bs = bits & 31;
    int nn = n + ws + 1;
    if (nn > W) nn = W;
    for (int i = nn - 1; i >= 0; --i) {
      int s = i - ws;
      uint32_t lo = (s >= 0 && s < n) ? w[s] : 0u;
      uint32_t lo2 = (s - 1 >= 0 && s - 1 < n) ? w[s - 1] : 0u;
      w[i] = bs ? (lo << bs) | (lo2 >> (32 - bs)) : lo;
    }
    n = nn;
    while (n > 0 && w[n - 1] == 0) --n;
  }
  RFQ_HD uint32_t divmod_small(uint32_t d) {  // this /= d, returns remainder
    uint64_t rem = 0;
    for (int i = n - 1; i >= 0; --i) {
      uint64_t cur = (rem << 32) | w[i];
      w[i] = (uint32_t)(cur / d); rem = cur % d;
    }
    while (n > 0 && w[n - 1] == 0) --n;
    return (uint32_t)rem;
  }
};
RFQ_HD int big_cmp(const Big& a, const Big& b) {
  if (a.n != b.n) return a.n < b.n ? -1 : 1;
  for (int i = a.n - 1; i >= 0; --i)
    if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
  return 0;
}
// compare a with (b << s) without materialising the shift, then a -= (b << s) if a >= it
RFQ_HD bool big_sub_shifted_if_ge(Big& a, const Big& b, int s) {
  Big t = b;
  t.shl(s);
  if (big_cmp(a, t) < 0) return false;
  int64_t borrow = 0;
  for (int i = 0; i < a.n; ++i) {
    int64_t d = (int64_t)a.w[i] - (i < t.n ? (int64_t)t.w[i] : 0) - borrow;
    borrow = d < 0;
    a.w[i] = (uint32_t)(d + (borrow << 32));
  }
  while (a.n > 0 && a.w[a.n - 1] == 0) --a.n;
  return true;
}

// ---- double decomposition ---------------------------------------------------------------------
RFQ_HD uint64_t dbl_bits(double v) {
#ifdef __CUDA_ARCH__
  return (uint64_t)__double_as_longlong(v);
#else
  uint64_t b; __builtin_memcpy(&b, &v, 8); return b;
#endif
}
// |v| = m * 2^e, m < 2^53 (v finite)
RFQ_HD void dbl_decompose(double v, uint64_t* m, int* e, bool* neg) {
  uint64_t b = dbl_bits(v);
  *neg = (b >> 63) != 0;
  int be = (int)((b >> 52) & 0x7FF);
  uint64_t frac = b & 0xFFFFFFFFFFFFFull;
  if (be == 0) { *m = frac; *e = -1074; }
  else { *m = frac | (1ull << 52); *e = be - 1075; }
}
RFQ_HD bool dbl_is_inf(double v) { return (dbl_bits(v) & 0x7FFFFFFFFFFFFFFFull) == 0x7FF0000000000000ull; }
RFQ_HD bool dbl_is_nan(double v) { return (dbl_bits(v) & 0x7FFFFFFFFFFFFFFFull) > 0x7FF0000000000000ull; }

// ---- "%.6f" -------------------------------------------------------------------------------------
struct Fixed6 {
  bool neg, inf, big;   // big: |v| >= 2^63 (integer part does not fit u64; exact digits on demand)
  uint64_t ip;          // integer part (when !big)
  uint32_t frac;        // 6 decimals, 0..999999
  int ip_digits;        // number of digits of the integer part (exact, also when big)
  int tz;               // trailing decimal zeros of frac (0..6)
};

RFQ_HD_COLD int big_int_digits(uint64_t m, int e, uint8_t* out /* may be null */) {
  // decimal digits of m << e (e >= 0); returns count, writes digits to out[0..count) if out
  Big b; b.set_u64(m); b.shl(e);
  // count digits: peel 9 at a time into a local buffer (max 309 digits + slack)
  uint8_t tmp[330];
  int pos = 330;
  while (b.n > 0) {
    uint32_t r = b.divmod_small(1000000000u);
    if (b.n > 0) { for (int i = 0; i < 9; ++i) { tmp[--pos] = (uint8_t)('0' + r % 10); r /= 10; } }
    else { do { tmp[--pos] = (uint8_t)('0' + r % 10); r /= 10; } while (r); }
  }
  int cnt = 330 - pos;
  if (out) for (int i = 0; i < cnt; ++i) out[i] = tmp[pos + i];
  return cnt;
}

RFQ_HD Fixed6 fixed6(double v) {
  Fixed6 f;
  f.inf = dbl_is_inf(v); f.big = false; f.ip = 0; f.frac = 0; f.ip_digits = 1; f.tz = 6;
  uint64_t m; int e;
  dbl_decompose(v, &m, &e, &f.neg);
  if (f.inf) return f;
  if (e >= 0) {
    if (e > 10 || (m << e) >> e != m) {  // >= 2^63
      f.big = true;
      f.ip_digits = big_int_digits(m, e, nullptr);
      return f;
    }
    f.ip = m << e;
  } else {
    int s = -e;
    if (s >= 64) f.ip = 0; else f.ip = m >> s;
    uint64_t fr = s >= 64 ? m : (m & ((1ull << s) - 1ull));
    if (s < 75 && fr) {
      unsigned __int128 P = (unsigned __int128)fr * 1000000u;
      unsigned __int128 q = P >> s;
      unsigned __int128 rem = P - (q << s);
      unsigned __int128 half = (unsigned __int128)1 << (s - 1);
      uint64_t qq = (uint64_t)q;
      if (rem > half || (rem == half && (qq & 1))) ++qq;
      if (qq >= 1000000u) { qq -= 1000000u; ++f.ip; }
      f.frac = (uint32_t)qq;
    }
  }
  f.ip_digits = ndigits_u64(f.ip);
  uint32_t fr = f.frac; int tz = 0;
  if (fr == 0) tz = 6; else while (fr % 10 == 0) { fr /= 10; ++tz; }
  f.tz = tz;
  return f;
}

// Short decimal literal  [+-] d{1,8} . d{1,6}  (a price): its "%.6f" digits are the literal's own
// digits padded with zeros.  The double the parser makes of it is within a few ulp (< 5e-8 below
// 1e8) of the literal, and the nearest rounding boundary of the sixth decimal is 5e-7 away, so the
// correctly rounded "%.6f" of that double cannot print anything else: no 128-bit arithmetic and no
// double needed.  Literals below 1e-5 are left to the general path (their |v| < 1e-6 test is made
// on the parsed double).
RFQ_HD bool decimal_literal(const uint8_t* p, uint32_t n, Fixed6* f) {
  if (n < 3 || n > 16) return false;
  uint32_t i = 0;
  f->neg = false;
  if (p[0] == '-' || p[0] == '+') { f->neg = p[0] == '-'; i = 1; }
  uint64_t ip = 0; uint32_t nd = 0;
  for (; i < n && (uint32_t)(p[i] - '0') < 10u; ++i) { ip = ip * 10u + (uint32_t)(p[i] - '0'); ++nd; }
  if (nd == 0 || nd > 8 || i >= n || p[i] != '.') return false;
  ++i;
  uint32_t fr = 0, nf = 0;
  for (; i < n && (uint32_t)(p[i] - '0') < 10u; ++i) { fr = fr * 10u + (uint32_t)(p[i] - '0'); ++nf; }
  if (nf == 0 || nf > 6 || i != n) return false;
  for (uint32_t q = nf; q < 6; ++q) fr *= 10u;
  if (ip == 0 && fr < 10u) return false;
  f->inf = false; f->big = false; f->ip = ip; f->frac = fr;
  f->ip_digits = ndigits_u64(ip);
  int tz = 0;
  if (fr == 0) tz = 6; else { uint32_t t = fr; while (t % 10u == 0) { t /= 10u; ++tz; } }
  f->tz = tz;
  return true;
}

// length of "{:.6f}" cut to `dec` decimals (dec in 1..6); inf -> "inf"/"-inf"
RFQ_HD int fixed_len(const Fixed6& f, int dec) {
  if (f.inf) return 3 + f.neg;
  return f.neg + f.ip_digits + 1 + dec;
}
// writes it; `v` is needed only for the big case. returns length
RFQ_HD int fixed_write(const Fixed6& f, double v, int dec, bool space_for_plus, uint8_t* out) {
  int n = 0;
  if (f.neg) out[n++] = '-'; else if (space_for_plus) out[n++] = ' ';
  if (f.inf) { out[n++] = 'i'; out[n++] = 'n'; out[n++] = 'f'; return n; }
  if (f.big) {
    uint64_t m; int e; bool ng;
    dbl_decompose(v, &m, &e, &ng);
    n += big_int_digits(m, e, out + n);
  } else {
    uint8_t* p = write_u64_rev(f.ip, out + n + f.ip_digits);
    (void)p;
    n += f.ip_digits;
  }
  out[n++] = '.';
  uint32_t fr = f.frac;
  uint32_t div = 100000u;
  for (int i = 0; i < dec; ++i) { out[n++] = (uint8_t)('0' + (fr / div) % 10); div /= 10; }
  return n;
}
// object-column float: f"{x: .6f}" then rstrip('0') (+ '0' if it ends with '.')
RFQ_HD int objfloat_dec(const Fixed6& f) { int d = 6 - f.tz; return d < 1 ? 1 : d; }
RFQ_HD int objfloat_len(const Fixed6& f) {
  if (f.inf) return 4;
  return 1 + f.ip_digits + 1 + objfloat_dec(f);
}

// ---- "%.6e" -------------------------------------------------------------------------------------
struct Sci7 { bool neg, inf; uint32_t digits; int exp10; };  // digits: 7-digit integer d.dddddd

RFQ_HD int approx_log10(uint64_t m, int e) {
  // floor(log10(m * 2^e)) within +-1: log10(2) ~ 1233/4096
  int bl = 64;
  uint64_t t = m;
  while (bl > 0 && !(t >> 63)) { t <<= 1; --bl; }  // bl = bit length of m
  long long lg2 = (long long)(bl - 1) + e;            // floor(log2 v)
  long long k = (lg2 * 1233) >> 12;
  return (int)k;
}

RFQ_HD_COLD Sci7 sci7(double v) {
  Sci7 s; s.inf = dbl_is_inf(v); s.digits = 0; s.exp10 = 0;
  uint64_t m; int e;
  dbl_decompose(v, &m, &e, &s.neg);
  if (s.inf || m == 0) return s;
  int k = approx_log10(m, e);
  for (int iter = 0; iter < 6; ++iter) {
    Big N, D;
    N.set_u64(m); D.set_u64(1);
    if (e >= 0) N.shl(e); else D.shl(-e);
    int p = k - 6;
    if (p >= 0) D.mul_pow10(p); else N.mul_pow10(-p);
    uint32_t q = 0;
    for (int bit = 27; bit >= 0; --bit)
      if (big_sub_shifted_if_ge(N, D, bit)) q |= 1u << bit;
    // if N was >= D << 28 the quotient is saturated: k is too small
    if (big_cmp(N, D) >= 0 || q >= 10000000u) { ++k; continue; }
    if (q < 1000000u) { --k; continue; }
    // round half even on remainder N vs D
    Big twice = N; twice.shl(1);
    int c = big_cmp(twice, D);
    if (c > 0 || (c == 0 && (q & 1u))) ++q;
    if (q == 10000000u) { q = 1000000u; ++k; }
    s.digits = q; s.exp10 = k;
    return s;
  }
  return s;
}
RFQ_HD int sci_len(const Sci7& s) {
  if (s.inf) return 3 + s.neg;
  int ae = s.exp10 < 0 ? -s.exp10 : s.exp10;
  return s.neg + 8 + 2 + (ae >= 100 ? 3 : 2);
}
RFQ_HD int sci_write(const Sci7& s, uint8_t* out) {
  int n = 0;
  if (s.neg) out[n++] = '-';
  if (s.inf) { out[n++] = 'i'; out[n++] = 'n'; out[n++] = 'f'; return n; }
  uint32_t d = s.digits;
  uint32_t div = 1000000u;
  out[n++] = (uint8_t)('0' + d / div);
  out[n++] = '.';
  for (int i = 0; i < 6; ++i) { div /= 10; out[n++] = (uint8_t)('0' + (d / div) % 10); }
  out[n++] = 'e';
  int ex = s.exp10;
  out[n++] = ex < 0 ? '-' : '+';
  if (ex < 0) ex = -ex;
  if (ex >= 100) { out[n++] = (uint8_t)('0' + ex / 100); ex %= 100; }
  out[n++] = (uint8_t)('0' + ex / 10);
  out[n++] = (uint8_t)('0' + ex % 10);
  return n;
}

}  // namespace rfq
