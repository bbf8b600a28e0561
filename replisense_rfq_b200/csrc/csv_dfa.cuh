// The delimiter / quote-state machine of pandas' C tokenizer restated as an 8-state DFA over
// byte classes, in a form that can be scanned in parallel (transition functions compose).
//
// Reference semantics: pandas/_libs/src/parser/tokenizer.c `tokenize_bytes` with read_csv's
// defaults (quotechar '"', doublequote, QUOTE_MINIMAL, no escapechar, skip_blank_lines,
// terminators \n, \r, \r\n), reached from app/file_parser.py:301.  Source not in this image;
// behaviour pinned by the differential tests.
//
//   R  start of record          F  start of field          U  in unquoted field
//   Q  in quoted field          QQ quote seen inside quotes
//   C  after the CR that ended a record (swallows one LF)
//   N  after the CR of a skipped blank line (swallows one LF *or one delimiter*: pandas quirk)
//   W  line that is whitespace-only so far (skipped if it stays so; otherwise the blanks are
//      the beginning of the first field)
// The end of the document is a virtual LF; a final state of Q means "EOF inside string".
#pragma once
#include "rfq_common.cuh"

namespace rfq {

enum { ST_R = 0, ST_F = 1, ST_U = 2, ST_Q = 3, ST_QQ = 4, ST_C = 5, ST_N = 6, ST_W = 7 };
enum { CL_O = 0, CL_D = 1, CL_QT = 2, CL_LF = 3, CL_CR = 4, CL_WS = 5, CL_PAD = 6, N_CLASSES = 7 };

RFQ_HD int csv_class(uint32_t b, uint32_t sep) {
  if (b == sep) return CL_D;
  if (b == '"') return CL_QT;
  if (b == '\n') return CL_LF;
  if (b == '\r') return CL_CR;
  if (b == ' ' || b == '\t') return CL_WS;
  return CL_O;
}

// next state from "effective start of record"
RFQ_HD int csv_from_R(int cl) {
  switch (cl) {
    case CL_LF: return ST_R;
    case CL_CR: return ST_N;
    case CL_WS: return ST_W;
    case CL_D: return ST_F;
    case CL_QT: return ST_Q;
    default: return ST_U;
  }
}

RFQ_HD int csv_next(int st, int cl) {
  if (cl == CL_PAD) return st;
  switch (st) {
    case ST_R: return csv_from_R(cl);
    case ST_F:
      if (cl == CL_LF) return ST_R;
      if (cl == CL_CR) return ST_C;
      if (cl == CL_QT) return ST_Q;
      if (cl == CL_D) return ST_F;
      return ST_U;
    case ST_U:
      if (cl == CL_LF) return ST_R;
      if (cl == CL_CR) return ST_C;
      if (cl == CL_D) return ST_F;
      return ST_U;
    case ST_Q: return cl == CL_QT ? ST_QQ : ST_Q;
    case ST_QQ:
      if (cl == CL_QT) return ST_Q;
      if (cl == CL_D) return ST_F;
      if (cl == CL_LF) return ST_R;
      if (cl == CL_CR) return ST_C;
      return ST_U;
    case ST_C: return cl == CL_LF ? ST_R : csv_from_R(cl);
    case ST_N: return (cl == CL_LF || cl == CL_D) ? ST_R : csv_from_R(cl);
    default:  // ST_W
      if (cl == CL_LF) return ST_R;
      if (cl == CL_CR) return ST_N;
      if (cl == CL_WS) return ST_W;
      if (cl == CL_D) return ST_F;
      return ST_U;
  }
}

// events at a byte, given the state before it
enum { EV_FIELD_END = 1, EV_LINE_END = 2, EV_REC_START = 4, EV_REC_START_WS = 8 };
RFQ_HD int csv_events(int st, int cl) {
  if (cl == CL_PAD) return 0;
  int ev = 0;
  bool infield = (st == ST_F || st == ST_U || st == ST_QQ);
  if (infield) {
    if (cl == CL_D) ev |= EV_FIELD_END;
    if (cl == CL_LF || cl == CL_CR) ev |= EV_FIELD_END | EV_LINE_END;
    return ev;
  }
  if (st == ST_Q) return 0;
  if (st == ST_W) {
    if (cl == CL_D) ev |= EV_FIELD_END | EV_REC_START_WS;
    else if (cl == CL_QT || cl == CL_O) ev |= EV_REC_START_WS;
    return ev;
  }
  // R, C, N: effective start of record unless the byte is swallowed
  if (st == ST_C && cl == CL_LF) return 0;
  if (st == ST_N && (cl == CL_LF || cl == CL_D)) return 0;
  if (cl == CL_D) ev |= EV_FIELD_END | EV_REC_START;
  else if (cl == CL_QT || cl == CL_O) ev |= EV_REC_START;
  return ev;
}

// ---- transition functions as 8 nibbles (function value for state s in bits 4s..4s+3) ---------
RFQ_HD uint32_t csv_class_fn(int cl) {
  uint32_t f = 0;
  for (int s = 0; s < 8; ++s) f |= (uint32_t)csv_next(s, cl) << (4 * s);
  return f;
}
// events of a class for every state, 4 bits per state
RFQ_HD uint32_t csv_class_ev(int cl) {
  uint32_t f = 0;
  for (int s = 0; s < 8; ++s) f |= (uint32_t)csv_events(s, cl) << (4 * s);
  return f;
}
#define RFQ_FN_IDENTITY 0x76543210u

// (a then b)(s) = b(a(s))
RFQ_HD uint32_t fn_compose(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  // b as a byte table (two PRMTs interleave its even and odd nibbles), looked up with a's nibbles
  // (PRMT again: a selector nibble below 8 picks a byte), bytes squeezed back into nibbles
  const uint32_t ev = b & 0x0F0F0F0Fu, od = (b >> 4) & 0x0F0F0F0Fu;
  const uint32_t blo = __byte_perm(ev, od, 0x5140u);   // b(0), b(1), b(2), b(3)
  const uint32_t bhi = __byte_perm(ev, od, 0x7362u);   // b(4) .. b(7)
  const uint32_t r0 = __byte_perm(blo, bhi, a);         // b(a(0..3))
  const uint32_t r1 = __byte_perm(blo, bhi, a >> 16);   // b(a(4..7))
  const uint32_t p0 = r0 | (r0 >> 4), p1 = r1 | (r1 >> 4);   // byte 0 = n0 | n1 << 4, byte 2 = n2 | n3 << 4
  return __byte_perm(p0, p1, 0x6420u);
#else
  uint32_t r = 0;
  for (int s = 0; s < 8; ++s) {
    uint32_t as = (a >> (4 * s)) & 0xFu;
    r |= ((b >> (4 * as)) & 0xFu) << (4 * s);
  }
  return r;
#endif
}
// Which of a lane's 16 bytes (classes as nibbles, byte k in bits 4k..4k+3) a replay has to visit to
// see every record event and every state change a LATER visited byte can tell apart.  With bit 4k
// standing for byte k: A = quote / LF / CR, N = anything but an ordinary byte, D = delimiter.  Visited:
// A; a non-ordinary byte whose predecessor is non-ordinary (a delimiter or blank right after an ordinary
// byte leaves the machine where an ordinary byte leaves it, as far as any later visited byte can
// tell: U / F / QQ all go to the same place on LF, CR, blank and delimiter, and a quote right after that
// delimiter makes the delimiter visited); a delimiter right before a quote or at the end of the lane;
// the byte after any of these; byte 0.  Returns bit k = byte k is visited.
RFQ_HD uint32_t visit_mask16(unsigned long long n) {
  const unsigned long long M1 = 0x1111111111111111ull;
  const unsigned long long b0 = n & M1, b1 = (n >> 1) & M1, b2 = (n >> 2) & M1;
  const unsigned long long N = b0 | b1 | b2;
  const unsigned long long QTm = b1 & ~b0 & ~b2;
  const unsigned long long A = (b1 & ~b2) | (b2 & ~b0 & ~b1);
  const unsigned long long Dm = b0 & ~b1 & ~b2;
  const unsigned long long V1 = A | (N & ((N << 4) | 1ull)) | (Dm & ((QTm >> 4) | (1ull << 60)));
  unsigned long long tt = V1 | (V1 << 4) | 1ull;
  tt = (tt | (tt >> 3)) & 0x0303030303030303ull;
  tt = (tt | (tt >> 6)) & 0x000F000F000F000Full;
  tt = (tt | (tt >> 12)) & 0x000000FF000000FFull;
  return (uint32_t)(tt | (tt >> 24)) & 0xFFFFu;
}

RFQ_HD int fn_apply(uint32_t f, int s) { return (int)((f >> (4 * s)) & 0xFu); }
RFQ_HD uint32_t fn_const(int s) { return 0x11111111u * (uint32_t)s; }

}  // namespace rfq
