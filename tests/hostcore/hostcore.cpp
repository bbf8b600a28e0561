// TEST HARNESS ONLY: compiles the host/device core headers of the CUDA library with g++ so the
// per-cell semantics (numeric parsers, float formatting, UTF-8 rules) can be checked against
// pandas / CPython on the CPU.  Never shipped, never on the product path.
#include <stdint.h>
#include <string.h>
#include "../../replisense_rfq_b200/csrc/cell_parse.cuh"
#include "../../replisense_rfq_b200/csrc/cell_format.cuh"
#include "../../replisense_rfq_b200/csrc/utf8.cuh"

using namespace rfq;

struct HcCell {
  int32_t is_na, int_status, uint_status, uint_neg, uint_gt, float_ok, bool_tok, inf_tok;
  int64_t ival; uint64_t uval; double fval;
  uint32_t nbytes, ncp, nesc; int32_t utf8_ok, has_fchar;
};

extern "C" {

void hc_parse_cell(const uint8_t* p, uint32_t n, int csv, HcCell* o) {
  memset(o, 0, sizeof(*o));
  TextInfo t = csv ? scan_text(CsvIter(p, n)) : scan_text(RawIter(p, n));
  o->nbytes = t.nbytes; o->ncp = t.ncp; o->nesc = t.nesc; o->utf8_ok = t.utf8_ok; o->has_fchar = t.has_fchar;
  o->is_na = is_na_token(t.nbytes, t.first8);
  o->bool_tok = bool_token(t.nbytes, t.first8);
  UintTok u;
  if (csv) {
    o->int_status = parse_int64(CsvIter(p, n), &o->ival);
    o->uint_status = parse_uint64(CsvIter(p, n), &u);
    o->float_ok = parse_double(CsvIter(p, n), &o->fval);
    o->inf_tok = inf_token(CsvIter(p, n));
  } else {
    o->int_status = parse_int64(RawIter(p, n), &o->ival);
    o->uint_status = parse_uint64(RawIter(p, n), &u);
    o->float_ok = parse_double(RawIter(p, n), &o->fval);
    o->inf_tok = inf_token(RawIter(p, n));
  }
  o->uval = u.v; o->uint_neg = u.neg; o->uint_gt = u.gt_i64;
}

int hc_csv_decode(const uint8_t* p, uint32_t n, uint8_t* out) {
  CsvIter it(p, n); int c, k = 0;
  while ((c = it.next()) >= 0) out[k++] = (uint8_t)c;
  return k;
}

// fixed "%.6f" pieces
void hc_fixed_info(double v, int32_t* tz, int32_t* ipd, int32_t* big) {
  Fixed6 f = fixed6(v); *tz = f.tz; *ipd = f.ip_digits; *big = f.big;
}
int hc_fixed(double v, int dec, int space, uint8_t* out) {
  Fixed6 f = fixed6(v);
  return fixed_write(f, v, dec, space != 0, out);
}
int hc_objfloat(double v, uint8_t* out) {
  Fixed6 f = fixed6(v);
  int n = fixed_write(f, v, objfloat_dec(f), true, out);
  return n == objfloat_len(f) ? n : -n;
}
int hc_sci(double v, uint8_t* out) {
  Sci7 s = sci7(v);
  int n = sci_write(s, out);
  return n == sci_len(s) ? n : -n;
}

// text path, sequential: returns emitted length (before strip) and strip range
struct HAcc { const uint8_t* p; int64_t len; uint32_t operator()(int64_t i) const { return (i >= 0 && i < len) ? p[i] : 0u; } };
int64_t hc_txt(const uint8_t* p, int64_t len, uint8_t* out) {
  HAcc B{p, len};
  int64_t k = 0, first = -1, last_end = 0;
  for (int64_t i = 0; i < len; ++i) {
    uint32_t b;
    if (!txt_emit(B, i, &b)) continue;
    bool start = p[i] < 0x80 || !u8_is_cont(p[i]);
    if (start && !u8_is_space_at(B, i)) {
      if (first < 0) first = k;
      int cl = p[i] < 0x80 ? 1 : u8_lead_len(p[i]);
      last_end = k + cl;
    }
    out[k++] = (uint8_t)b;
  }
  if (first < 0) return 0;
  memmove(out, out + first, last_end - first);
  return last_end - first;
}

}  // extern "C"
