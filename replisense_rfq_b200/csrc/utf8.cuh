// UTF-8 "decode with errors='ignore'" keep/drop decision, universal-newline rule and Unicode
// whitespace classes, as host/device functions over a byte accessor.
//
// Reference semantics: app/file_parser.py:146-148 (`open(..., 'r', encoding='utf-8',
// errors='ignore').read().strip()`): CPython's UTF-8 decoder drops every maximal invalid
// prefix and resumes at the offending byte, so a byte survives iff it is ASCII or belongs to a
// complete, well-formed sequence (no overlongs, no surrogates, <= U+10FFFF).
#pragma once
#include "rfq_common.cuh"

// length implied by a lead byte; 0 = not a lead (ASCII, continuation or invalid)
RFQ_HD int u8_lead_len(uint32_t b) {
  if (b >= 0xC2 && b <= 0xDF) return 2;
  if (b >= 0xE0 && b <= 0xEF) return 3;
  if (b >= 0xF0 && b <= 0xF4) return 4;
  return 0;
}
RFQ_HD bool u8_is_cont(uint32_t b) { return (b & 0xC0) == 0x80; }

// Acc: uint32_t operator()(int64_t i) -> byte at i, 0 outside the document.
template <class Acc>
RFQ_HD bool u8_seq_valid(const Acc& B, int64_t p) {
  uint32_t lead = B(p);
  int n = u8_lead_len(lead);
  if (n == 0) return false;
  uint32_t b1 = B(p + 1);
  uint32_t lo = 0x80, hi = 0xBF;
  if (lead == 0xE0) lo = 0xA0;
  else if (lead == 0xED) hi = 0x9F;
  else if (lead == 0xF0) lo = 0x90;
  else if (lead == 0xF4) hi = 0x8F;
  if (b1 < lo || b1 > hi) return false;
  if (n >= 3 && !u8_is_cont(B(p + 2))) return false;
  if (n == 4 && !u8_is_cont(B(p + 3))) return false;
  return true;
}

// Does the byte at p survive decode(errors='ignore')?
template <class Acc>
RFQ_HD bool u8_keep(const Acc& B, int64_t p) {
  uint32_t b = B(p);
  if (b < 0x80) return true;
  if (!u8_is_cont(b)) return u8_seq_valid(B, p);  // lead (or invalid byte -> false)
  for (int back = 1; back <= 3; ++back) {
    uint32_t c = B(p - back);
    if (u8_is_cont(c)) continue;
    return u8_lead_len(c) > back && u8_seq_valid(B, p - back);
  }
  return false;
}

// str.isspace() for the character starting at p (p must be a kept character start).
// ASCII \t\n\v\f\r space \x1c-\x1f; U+0085 U+00A0 U+1680 U+2000-200A U+2028 U+2029 U+202F
// U+205F U+3000.  NOT: NUL, U+200B, U+FEFF.
template <class Acc>
RFQ_HD bool u8_is_space_at(const Acc& B, int64_t p) {
  uint32_t b = B(p);
  if (b < 0x80) return (b >= 0x09 && b <= 0x0D) || (b >= 0x1C && b <= 0x20);
  if (b == 0xC2) { uint32_t c = B(p + 1); return c == 0x85 || c == 0xA0; }
  if (b == 0xE1) return B(p + 1) == 0x9A && B(p + 2) == 0x80;
  if (b == 0xE2) {
    uint32_t c = B(p + 1), d = B(p + 2);
    if (c == 0x80) return (d >= 0x80 && d <= 0x8A) || d == 0xA8 || d == 0xA9 || d == 0xAF;
    if (c == 0x81) return d == 0x9F;
    return false;
  }
  if (b == 0xE3) return B(p + 1) == 0x80 && B(p + 2) == 0x80;
  return false;
}

// Universal newlines act on the DECODED text: a '\n' is swallowed iff the previous surviving
// byte is '\r' (dropped invalid bytes in between do not separate the pair).
template <class Acc>
RFQ_HD bool u8_lf_swallowed(const Acc& B, int64_t p) {
  int64_t q = p - 1;
  while (q >= 0) {
    uint32_t c = B(q);
    if (c < 0x80) return c == '\r';
    if (u8_keep(B, q)) return false;
    --q;
  }
  return false;
}

// Final per-byte decision for the text path: emitted or not, and the emitted value.
template <class Acc>
RFQ_HD bool txt_emit(const Acc& B, int64_t p, uint32_t* out) {
  uint32_t b = B(p);
  if (b == '\n') { *out = '\n'; return !u8_lf_swallowed(B, p); }
  if (b == '\r') { *out = '\n'; return true; }
  *out = b;
  return u8_keep(B, p);
}
