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
  // expand b to a byte table, look up with a's nibbles via PRMT
  uint32_t blo = (b & 0xFu) | ((b & 0xF0u) << 4) | ((b & 0xF00u) << 8) | ((b & 0xF000u) << 12);
  uint32_t bh = b >> 16;
  uint32_t bhi = (bh & 0xFu) | ((bh & 0xF0u) << 4) | ((bh & 0xF00u) << 8) | ((bh & 0xF000u) << 12);
  uint32_t r0 = __byte_perm(blo, bhi, a & 0xFFFFu);   // bytes b(a(0..3))
  uint32_t r1 = __byte_perm(blo, bhi, a >> 16);       // bytes b(a(4..7))
  uint32_t p0 = (r0 | (r0 >> 4));                     // byte0 = n0|n1<<4, byte2 = n2|n3<<4
  uint32_t p1 = (r1 | (r1 >> 4));
  uint32_t lo = (p0 & 0xFFu) | ((p0 >> 8) & 0xFF00u);
  uint32_t hi = (p1 & 0xFFu) | ((p1 >> 8) & 0xFF00u);
  return lo | (hi << 16);
#else
  uint32_t r = 0;
  for (int s = 0; s < 8; ++s) {
    uint32_t as = (a >> (4 * s)) & 0xFu;
    r |= ((b >> (4 * as)) & 0xFu) << (4 * s);
  }
  return r;
#endif
}
RFQ_HD int fn_apply(uint32_t f, int s) { return (int)((f >> (4 * s)) & 0xFu); }
RFQ_HD uint32_t fn_const(int s) { return 0x11111111u * (uint32_t)s; }

}  // namespace rfq
