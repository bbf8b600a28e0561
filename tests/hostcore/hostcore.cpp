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

// ================================================================================================
// Sequential model of the whole CSV table pipeline built from the SAME host/device functions the
// kernels call (DFA, classify/measure/print, column decisions, header naming).  It exists so the
// semantics can be fuzzed against pandas on the CPU; the kernels add only the parallel plumbing.
#include <string>
#include <vector>
#include "../../replisense_rfq_b200/csrc/cell_ops.cuh"
#include "../../replisense_rfq_b200/csrc/header_names.h"

#include "../../replisense_rfq_b200/csrc/csv2_rows.cuh"
namespace {
static bool g_fused = false;
static uint64_t g_fused_used = 0;
// row-walker mode: cells are found by scan_field from the record start (what the v2 kernels do)
// instead of through the field index; hot-path results are compared with the general functions
static bool g_walk = false;
static uint64_t g_walk_mismatch = 0, g_hot_mismatch = 0, g_hot_cells = 0, g_cold_cells = 0;
static uint64_t g_emit_hot = 0, g_emit_cold = 0, g_emit_mismatch = 0;

static bool same_src(const CellSrc& a, const CellSrc& b) {
  if (a.missing != b.missing) return false;
  if (a.missing) return true;
  // the walker may call a span CM_RAW where csv_simplify keeps CM_CSV (same decoded bytes): compare decoded text
  auto dec = [](const CellSrc& s) {
    std::string o; int c;
    if (s.mode == CM_CSV) { CsvIter it(s.p, s.n); while ((c = it.next()) >= 0) o.push_back((char)c); }
    else if (s.mode == CM_CSVQ) { PairIter it(s.p, s.n); while ((c = it.next()) >= 0) o.push_back((char)c); }
    else o.assign((const char*)s.p, s.n);
    return o;
  };
  return dec(a) == dec(b);
}
static bool same_cand(const CellCand& a, const CellCand& b) {
  return a.flags == b.flags && a.int_status == b.int_status && a.na == b.na && a.has_float == b.has_float &&
         a.need_exact == b.need_exact && a.str_len == b.str_len && a.extra == b.extra && a.int_len == b.int_len &&
         a.uint_len == b.uint_len && a.bool_len == b.bool_len && a.ip_len == b.ip_len && a.tz == b.tz &&
         a.inf_len == b.inf_len && a.obj_len == b.obj_len && a.sci_len == b.sci_len && a.fl == b.fl;
}

struct Rec { uint64_t start; uint32_t f0, f1; };

// status: 0 ok, 1 attempt fails (exception or <= 1 column), 2 unsupported on the GPU path
int csv_attempt(const uint8_t* doc, uint64_t len, uint32_t sep, int64_t nrows, std::string& out) {
  std::vector<uint64_t> fend;
  std::vector<Rec> recs;
  uint64_t i = 0;
  if (len >= 3 && doc[0] == 0xEF && doc[1] == 0xBB && doc[2] == 0xBF) i = 3;
  const uint64_t doc_begin = i;
  int st = ST_R;
  uint64_t ws_start = 0, cur_start = 0;
  uint32_t cur_f0 = 0;
  bool stopped = false;
  for (; i <= len; ++i) {
    int cl = (i == len) ? CL_LF : csv_class(doc[i], sep);
    int ev = csv_events(st, cl);
    int ns = csv_next(st, cl);
    if (st != ST_W && ns == ST_W) ws_start = i;
    if (ev & EV_REC_START) { cur_start = i; cur_f0 = (uint32_t)fend.size(); }
    if (ev & EV_REC_START_WS) {
      cur_start = ws_start; cur_f0 = (uint32_t)fend.size();
      // pandas walks back to the previous LF here: after a bare CR the outcome is an accident (unsupported)
      if (ws_start > doc_begin && doc[ws_start - 1] != '\n') return 2;
    }
    if (ev & EV_FIELD_END) fend.push_back(i);
    if (ev & EV_LINE_END) {
      recs.push_back({cur_start, cur_f0, (uint32_t)fend.size()});
      if ((int64_t)recs.size() == nrows + 1) { stopped = true; st = ns; break; }
    }
    st = ns;
  }
  if (!stopped && st == ST_Q) return 1;  // EOF inside string
  if (recs.empty()) return 1;            // EmptyDataError
  HeaderInfo hdr = parse_header(doc, len, sep);
  if (!hdr.found || !hdr.utf8_ok) return 1;
  uint32_t h = (uint32_t)hdr.names.size();
  uint32_t W = h;
  if (recs.size() >= 2) { uint32_t f1 = recs[1].f1 - recs[1].f0; if (f1 > W) W = f1; }
  for (size_t r = 2; r < recs.size(); ++r)
    if (recs[r].f1 - recs[r].f0 > W) return 1;  // "Expected N fields, saw M"
  if (h <= 1) return 1;
  uint32_t lead = W - h;
  uint32_t ndata = (uint32_t)recs.size() - 1;
  uint32_t bl = chunk_rows_for_width(W);
  uint32_t nch = (ndata + bl - 1) / bl;
  auto cell_idx = [&](uint32_t row, uint32_t c) {
    const Rec& R = recs[row + 1];
    CellSrc s; s.mode = 0; s.missing = false; s.clean = false; s.bits = 0; s.p = doc; s.n = 0;
    uint32_t nf = R.f1 - R.f0;
    if (c >= nf) { s.missing = true; return s; }
    uint64_t a = c == 0 ? R.start : fend[R.f0 + c - 1] + 1;
    uint64_t b = fend[R.f0 + c];
    s.p = doc + a; s.n = (uint32_t)(b - a);
    csv_simplify(s);
    return s;
  };
  // padded copy: the walker's window loads may read a few bytes past the document
  std::vector<uint8_t> padded;
  if (g_walk) { padded.assign(doc, doc + len); padded.resize(len + 32, 0); }
  auto cell_walk = [&](uint32_t row, uint32_t c) {
    const Rec& R = recs[row + 1];
    uint32_t limit = (row + 2 < recs.size()) ? (uint32_t)recs[row + 2].start : (uint32_t)len;
    RowWalk w(padded.data(), (uint32_t)R.start, limit, sep);
    CellSrc s = w.next();
    for (uint32_t k = 0; k < c; ++k) s = w.next();
    if (c + 1 == W) {   // after the last column the record must be over (rows longer than W were rejected above)
      if (!w.eol) ++g_walk_mismatch;
    }
    // map back onto the caller's buffer so pointers stay valid after `padded` is gone (same offsets)
    if (!s.missing) s.p = doc + (s.p - padded.data());
    return s;
  };
  auto cell = [&](uint32_t row, uint32_t c) {
    if (!g_walk) return cell_idx(row, c);
    CellSrc a = cell_walk(row, c), b = cell_idx(row, c);
    if (!same_src(a, b)) ++g_walk_mismatch;
    return a;
  };
  std::vector<ChunkStat> stat((size_t)nch * W, ChunkStat{0, RFQ_NO_ROW, RFQ_NO_ROW, RFQ_NO_ROW, RFQ_NO_ROW, {0, 0, 0}});
  std::vector<uint32_t> row_na(ndata, 0);
  std::vector<ChunkCand> cand((size_t)nch * W, ChunkCand{0, 0, 0, 0, 0, 6, 0, 0, 0, 0, 0, 0});
  for (uint32_t r = 0; r < ndata; ++r)
    for (uint32_t c = 0; c < W; ++c) {
      int is;
      uint32_t f;
      if (g_fused) {   // one visit: flags + measurement candidates (the fused GPU sweep)
        CellSrc cs = cell(r, c);
        CellCand m = classify_measure_cell(cs);
        if (g_walk) {  // the hot kernels' shortcut must agree with the general function wherever it decides
          HotCand hc = classify_measure_hot(cs);
          if (hc.cold) ++g_cold_cells;
          else { ++g_hot_cells; if (!same_cand(hc.m, m)) ++g_hot_mismatch; }
        }
        f = m.flags; is = m.int_status;
        ChunkCand& C = cand[(size_t)(r / bl) * W + c];
        if (m.na) C.na_count++;
        else {
          C.str_len = std::max(C.str_len, m.str_len); C.int_len = std::max(C.int_len, m.int_len);
          C.uint_len = std::max(C.uint_len, m.uint_len); C.bool_len = std::max(C.bool_len, m.bool_len);
          if (m.has_float) {
            C.ip_len = std::max(C.ip_len, m.ip_len); C.min_tz = std::min(C.min_tz, m.tz);
            C.inf_len = std::max(C.inf_len, m.inf_len); C.obj_len = std::max(C.obj_len, m.obj_len);
            C.sci_len = std::max(C.sci_len, m.sci_len); C.fl |= m.fl;
          }
          if (m.need_exact) C.need_exact = 1;
        }
      } else {
        f = classify_cell(cell(r, c), &is);
      }
      ChunkStat& S = stat[(size_t)(r / bl) * W + c];
      S.flags |= f;
      if (is != INT_OK) {
        if (r < S.int_bad_min) S.int_bad_min = r;
        if (is == INT_OVERFLOW && r < S.int_ovf_min) S.int_ovf_min = r;
        if ((f & SF_UINT_INVALID) && r < S.uint_inv_min) S.uint_inv_min = r;
        if ((f & SF_UINT_OVF) && r < S.uint_ovf_min) S.uint_ovf_min = r;
      }
      if ((f & SF_ANY_NA) && c >= lead) row_na[r]++;
    }
  std::vector<uint8_t> ctype((size_t)nch * W);
  std::vector<bool> chunk_has_nf(nch, false);
  for (size_t k = 0; k < stat.size(); ++k) {
    if (stat[k].flags & SF_BAD_UTF8) return 1;  // UnicodeDecodeError
    ctype[k] = (uint8_t)decide_chunk_type(stat[k]);
    if (ctype[k] == CT_STRNF && (k % W) >= lead) chunk_has_nf[k / W] = true;
  }
  std::vector<bool> keep(ndata);
  uint32_t nkeep = 0;
  for (uint32_t r = 0; r < ndata; ++r) {
    keep[r] = !(row_na[r] == h && !chunk_has_nf[r / bl]);
    nkeep += keep[r];
  }
  std::vector<ChunkMeasure> meas((size_t)nch * W, ChunkMeasure{0, 0, 6, 0, 0, 0, 0, 0});
  bool exact = !g_fused;
  if (g_fused) {
    std::vector<uint32_t> dropped(nch, 0);
    for (uint32_t r = 0; r < ndata; ++r) if (!keep[r]) dropped[r / bl]++;
    for (size_t k = 0; k < meas.size(); ++k)
      if ((k % W) >= lead && !finish_measure(cand[k], ctype[k], dropped[k / W], &meas[k])) exact = true;
    if (exact) std::fill(meas.begin(), meas.end(), ChunkMeasure{0, 0, 6, 0, 0, 0, 0, 0});
    else g_fused_used++;
  }
  for (uint32_t r = 0; r < ndata && exact; ++r) {
    if (!keep[r]) continue;
    for (uint32_t c = lead; c < W; ++c) {
      size_t k = (size_t)(r / bl) * W + c;
      CellMeasure m = measure_cell(cell(r, c), ctype[k]);
      ChunkMeasure& M = meas[k];
      if (m.is_nan) { M.nan_left = 1; continue; }
      if (m.plain_len > M.plain_len) M.plain_len = m.plain_len;
      if (m.ip_len > M.ip_len) M.ip_len = m.ip_len;
      if (m.tz < M.min_tz) M.min_tz = m.tz;
      if (m.inf_len > M.inf_len) M.inf_len = m.inf_len;
      if (m.obj_len > M.obj_len) M.obj_len = m.obj_len;
      if (m.sci_len > M.sci_len) M.sci_len = m.sci_len;
      M.fl |= m.fl;
    }
  }
  const char* encs = "utf-8";
  out = std::string("=== CSV Data (using ") + encs + ", separator '" + (char)sep + "') ===\n";
  if (nkeep == 0) { out += empty_frame_text(hdr.names); return 0; }
  std::vector<ColumnLayout> lay(h);
  for (uint32_t j = 0; j < h; ++j) {
    uint32_t c = lead + j;
    bool nan_left = false;
    for (uint32_t k = 0; k < nch; ++k) nan_left |= meas[(size_t)k * W + c].nan_left != 0;
    int cls = fold_column_class(&ctype[c], nch, W, nan_left);
    if (cls == CC_UNSUP) return 2;
    lay[j] = layout_column(cls, &meas[c], &ctype[c], nch, W, hdr.ncp[j]);
  }
  // header line
  for (uint32_t j = 0; j < h; ++j) {
    bool numeric = lay[j].cls == CC_INT || lay[j].cls == CC_UINT || lay[j].cls == CC_FLOAT || lay[j].cls == CC_BOOL;
    uint32_t cp = hdr.ncp[j] + (numeric ? 1 : 0);
    if (j) out += ' ';
    out.append(lay[j].width - cp, ' ');
    if (numeric) out += ' ';
    out += hdr.disp[j];
  }
  std::vector<uint8_t> buf;
  for (uint32_t r = 0; r < ndata; ++r) {
    if (!keep[r]) continue;
    out += '\n';
    for (uint32_t j = 0; j < h; ++j) {
      uint32_t c = lead + j;
      int ct = ctype[(size_t)(r / bl) * W + c];
      int kind = print_kind(lay[j].cls, ct, lay[j].sci);
      CellSrc s = cell(r, c);
      uint32_t ncp, nb;
      printed_size(s, ct, kind, lay[j].dec, &ncp, &nb);
      if (j) out += ' ';
      out.append(lay[j].width - ncp, ' ');
      buf.resize(nb + 400);
      uint32_t wrote = print_cell(s, ct, kind, lay[j].dec, buf.data());
      if (wrote != nb) { out += "<<SIZE MISMATCH>>"; }
      out.append((const char*)buf.data(), wrote);
      if (g_walk) {   // the hot emit sweep's shortcut must print the same field wherever it prints at all
        const uint32_t width = lay[j].width;
        std::vector<uint8_t> field(width + nb + 64, ' ');
        std::vector<uint8_t> pad(s.missing ? 0 : s.n + 32, 0);   // window loads read a few bytes past the token
        CellSrc s2 = s;
        if (!s.missing) { memcpy(pad.data(), s.p, s.n); s2.p = pad.data(); }
        uint32_t ex = 0;
        if (emit_hot(s2, ct, kind, lay[j].dec, width, field.data(), &ex) == 0) {
          ++g_emit_hot;
          std::string want(width - ncp, ' ');
          want.append((const char*)buf.data(), wrote);
          if (ex != nb - ncp || std::string((const char*)field.data(), width + ex) != want) ++g_emit_mismatch;
        } else ++g_emit_cold;
      }
    }
  }
  return 0;
}

}  // namespace

extern "C" {

// Full reference trial loop restricted to what can differ structurally: utf-8 then latin-1, each
// with ',', ';', TAB (cp1252 can never win where latin-1 failed: same bytes, same structure).
// returns status (0 ok / 1 unparseable / 2 unsupported); *out_len bytes written to out (cap).
// 1 when the windowed scan of a plain span agrees with the byte-by-byte scan_text
int hc_scan_text_cmp(const uint8_t* p, uint32_t n) {
  TextInfo a = scan_text(RawIter(p, n)), b = scan_text_raw(p, n);
  if (a.nbytes != b.nbytes || a.first8 != b.first8 || a.utf8_ok != b.utf8_ok) return 0;
  if (a.utf8_ok && (a.ncp != b.ncp || a.nesc != b.nesc)) return 0;   // counts only mean something for valid text
  return 1;
}

// 1 when the hot classifier leaves this row's field `col` to the fix-up kernel (debugging aid)
int hc_hot_is_cold(const uint8_t* row, uint32_t n, uint32_t sep, uint32_t col) {
  std::vector<uint8_t> b(row, row + n); b.resize(n + 32, 0);
  RowWalk w(b.data(), 0, n, sep);
  CellSrc s = w.next();
  for (uint32_t k = 0; k < col; ++k) s = w.next();
  return classify_measure_hot(s).cold ? 1 : 0;
}
// Reduced replay of the row scan (csv_dfa.cuh visit_mask16) against the full DFA walk, lane by lane
// (16 bytes, exact start state): record starts / ends must come out the same.  Returns mismatching lanes.
uint64_t hc_visit_check(const uint8_t* doc, uint64_t len, uint32_t sep) {
  uint64_t bad = 0;
  int st = ST_R;
  for (uint64_t base = 0; base <= len; base += 16) {
    unsigned long long cls = 0;
    for (int k = 0; k < 16; ++k) {
      uint64_t pos = base + k;
      int c = pos < len ? csv_class(doc[pos], sep) : (pos == len ? CL_LF : CL_PAD);
      cls |= (unsigned long long)c << (4 * k);
    }
    uint32_t sm = 0, wm = 0, lm = 0, sm2 = 0, wm2 = 0, lm2 = 0;
    int s1 = st, s2 = st;
    for (int k = 0; k < 16; ++k) {
      int c = (int)((cls >> (4 * k)) & 0xF);
      int e = csv_events(s1, c);
      if (e & EV_REC_START) sm |= 1u << k;
      if (e & EV_REC_START_WS) wm |= 1u << k;
      if (e & EV_LINE_END) lm |= 1u << k;
      s1 = csv_next(s1, c);
    }
    for (uint32_t m = visit_mask16(cls); m; m &= m - 1) {
      int k = __builtin_ctz(m);
      int c = (int)((cls >> (4 * k)) & 0xF);
      int e = csv_events(s2, c);
      if (e & EV_REC_START) sm2 |= 1u << k;
      if (e & EV_REC_START_WS) wm2 |= 1u << k;
      if (e & EV_LINE_END) lm2 |= 1u << k;
      s2 = csv_next(s2, c);
    }
    if (sm != sm2 || wm != wm2 || lm != lm2) ++bad;
    st = s1;
  }
  return bad;
}
void hc_set_walk(int on) { g_walk = on != 0; }
void hc_walk_stats(uint64_t* o) { o[0] = g_walk_mismatch; o[1] = g_hot_mismatch; o[2] = g_hot_cells; o[3] = g_cold_cells; o[4] = g_emit_mismatch; o[5] = g_emit_hot; o[6] = g_emit_cold; }
void hc_set_fused(int on) { g_fused = on != 0; }
uint64_t hc_fused_used() { return g_fused_used; }

int hc_csv_parse(const uint8_t* doc, uint64_t len, int64_t nrows, uint8_t* out, uint64_t cap,
                 uint64_t* out_len) {
  static const uint32_t seps[3] = {',', ';', '\t'};
  std::string text;
  *out_len = 0;
  for (int enc = 0; enc < 2; ++enc) {
    std::vector<uint8_t> tr;
    const uint8_t* d = doc; uint64_t n = len;
    if (enc == 1) {  // latin-1 -> UTF-8 transcode of the stream
      tr.reserve(len * 2);
      for (uint64_t i = 0; i < len; ++i) {
        uint8_t b = doc[i];
        if (b < 0x80) tr.push_back(b);
        else { tr.push_back((uint8_t)(0xC0 | (b >> 6))); tr.push_back((uint8_t)(0x80 | (b & 0x3F))); }
      }
      d = tr.data(); n = tr.size();
    }
    for (int k = 0; k < 3; ++k) {
      int st = csv_attempt(d, n, seps[k], nrows, text);
      if (st == 2) return 2;
      if (st == 0) {
        if (enc == 1) text.replace(text.find("utf-8"), 5, "latin-1");
        if (text.size() > cap) return -1;
        memcpy(out, text.data(), text.size());
        *out_len = text.size();
        return 0;
      }
    }
  }
  return 1;
}

}  // extern "C"

// ---- sheet XML primitives ---------------------------------------------------------------------
#include "../../replisense_rfq_b200/csrc/xlsx_core.cuh"
extern "C" {
int hc_strtod(const uint8_t* s, uint32_t n, double* out) { return rfq::strtod_exact(s, n, out) ? 1 : 0; }
long long hc_timedelta_us(double v) { int64_t us = 0; return rfq::excel_to_timedelta(v, &us) ? (long long)us : (long long)0x7FFFFFFFFFFFFFFFll; }
int hc_timedelta_str(long long us, char* out) { return (int)rfq::pytd_write((int64_t)us, (uint8_t*)out); }
int hc_float_repr(double v, char* out) { std::string t = rfq::py_float_repr(v); memcpy(out, t.data(), t.size()); return (int)t.size(); }
int hc_xml_decode(const uint8_t* p, uint32_t n, uint8_t* out) {
  rfq::XmlIter it(p, n); int c, k = 0;
  while ((c = it.next()) >= 0) out[k++] = (uint8_t)c;
  return it.bad ? -1 : k;
}
}

// ================================================================================================
// Sequential model of the sheet pipeline (same per-element functions as the kernels).
#include "../../replisense_rfq_b200/csrc/xlsx_cells.cuh"
namespace {
// python-engine header names: str() of the first-row values, blanks -> Unnamed: k, de-duplicated
std::string py_repr_double(double v) {
  char buf[64];
  for (int prec = 1; prec <= 17; ++prec) {
    snprintf(buf, sizeof buf, "%.*e", prec - 1, v);
    if (strtod(buf, nullptr) == v) break;
  }
  // buf = d.ddddde[+-]XX ; re-render with Python's repr rules
  std::string m(buf);
  size_t epos = m.find('e');
  int ex = atoi(m.c_str() + epos + 1);
  std::string digits;
  bool neg = m[0] == '-';
  for (size_t i = neg ? 1 : 0; i < epos; ++i) if (m[i] != '.') digits.push_back(m[i]);
  std::string out = neg ? "-" : "";
  if (ex < -4 || ex >= 16) {
    out += digits.substr(0, 1);
    if (digits.size() > 1) { out += "."; out += digits.substr(1); }
    char eb[16]; snprintf(eb, sizeof eb, "e%c%02d", ex < 0 ? '-' : '+', ex < 0 ? -ex : ex);
    out += eb;
  } else if (ex < 0) {
    out += "0."; out.append(-ex - 1, '0'); out += digits;
  } else {
    if ((int)digits.size() <= ex + 1) { out += digits; out.append(ex + 1 - digits.size(), '0'); out += ".0"; }
    else { out += digits.substr(0, ex + 1); out += "."; out += digits.substr(ex + 1); }
  }
  return out;
}
}  // namespace

extern "C" {
// status: 0 ok, 2 unsupported, 3 empty sheet (df.empty -> skipped by the reference)
int hc_sheet_parse(const uint8_t* doc, uint32_t len, const uint8_t* sst_doc, uint32_t sst_len,
                   const uint32_t* date_ids, uint32_t n_date, int date1904, int64_t nrows, uint8_t* out, uint64_t cap,
                   uint64_t* out_len) {
  *out_len = 0;
  uint32_t flags = 0;
  // shared strings
  std::vector<uint32_t> s_off, s_len; std::vector<uint8_t> s_arena;
  if (sst_doc) {
    const uint8_t* e = sst_doc + sst_len;
    for (uint32_t i = 0; i + 3 < sst_len; ++i) {
      if (sst_doc[i] != '<') continue;
      int tg = xml_classify(sst_doc + i, e);
      if (tg == XT_BANG) flags |= XU_BANG;
      if (tg != XT_SI) continue;
      uint32_t n = sst_decode_si(sst_doc + i, e, nullptr, &flags);
      s_off.push_back((uint32_t)s_arena.size()); s_len.push_back(n);
      s_arena.resize(s_arena.size() + n);
      if (n) sst_decode_si(sst_doc + i, e, s_arena.data() + s_off.back(), &flags);
    }
  }
  SstView sv{s_off.data(), s_len.data(), (uint32_t)s_off.size()};
  // rows / cells in document order
  struct RowE { uint32_t r; uint32_t c0, c1; };
  std::vector<RowE> rows; std::vector<XCell> cells;
  const uint8_t* e = doc + len;
  for (uint32_t i = 0; i + 2 < len; ++i) {
    if (doc[i] != '<') continue;
    int tg = xml_classify(doc + i, e);
    if (tg == XT_BANG) flags |= XU_BANG;
    if (tg == XT_ROW) {
      if (!rows.empty()) rows.back().c1 = (uint32_t)cells.size();
      rows.push_back({xlsx_row_number(doc + i, e, &flags), (uint32_t)cells.size(), (uint32_t)cells.size()});
    } else if (tg == XT_C) {
      cells.push_back(xlsx_parse_cell(doc, len, i, sv, date_ids, n_date, date1904 != 0, &flags));
    }
  }
  if (!rows.empty()) rows.back().c1 = (uint32_t)cells.size();
  // elements without r= continue the count of the element before them (row_counter / col_counter)
  if (flags & XU_ROW_NOREF) {
    uint32_t counter = 0;
    for (auto& R : rows) { if (R.r == 0) R.r = counter + 1; counter = R.r; }
  }
  if (flags & XU_C_NOREF)
    for (auto& R : rows) {
      uint32_t counter = 0;
      for (uint32_t k = R.c0; k < R.c1; ++k) { if (cells[k].col == 0) cells[k].col = counter + 1; counter = cells[k].col; }
    }
  uint32_t cap_rows = (uint32_t)std::min<int64_t>(nrows, 0x7FFFFFF0) + 1;  // file_rows_needed
  uint32_t W = 0, lastrow = 0, prev_r = 0;
  for (auto& R : rows) {
    if (R.r <= prev_r) flags |= XU_ORDER;
    prev_r = R.r;
    uint32_t pc = 0;
    for (uint32_t k = R.c0; k < R.c1; ++k) {
      if (cells[k].col <= pc) flags |= XU_ORDER;
      pc = cells[k].col;
      if (R.r <= cap_rows && !xcell_is_empty(cells[k])) { W = std::max(W, cells[k].col); lastrow = std::max(lastrow, R.r); }
    }
  }
  if (xu_problem(flags)) return 2;
  if (lastrow <= 1) return 3;  // no data rows -> empty frame -> sheet skipped
  uint32_t ndata = lastrow - 1;
  std::vector<int> r2e(lastrow + 1, -1);
  for (size_t k = 0; k < rows.size(); ++k) if (rows[k].r <= lastrow) r2e[rows[k].r] = (int)k;
  auto cell = [&](uint32_t r, uint32_t c) {  // r 1-based sheet row, c 0-based column
    CellSrc s; s.p = nullptr; s.n = 0; s.mode = CM_NULL; s.missing = false; s.clean = false; s.bits = 0;
    int k = r2e[r];
    if (k < 0) return s;
    for (uint32_t q = rows[k].c0; q < rows[k].c1; ++q) if (cells[q].col == c + 1) {
      const XCell& x = cells[q];
      s.mode = (uint8_t)x.mode; s.bits = x.bits;
      if (x.mode == CM_RAW) { s.p = s_arena.data() + (x.bits >> 32); s.n = (uint32_t)x.bits & ~RFQ_SST_CLEAN; }
      if (x.mode == CM_XML) { s.p = doc + (x.bits >> 32); s.n = (uint32_t)x.bits; }
      break;
    }
    return s;
  };
  // header
  std::vector<std::string> raw(W); std::vector<size_t> unnamed;
  bool all_int = true;
  std::vector<int> lkind(W, 0);
  std::vector<std::string> ftext(W);
  std::vector<uint8_t> no_blank(W, 0);
  std::vector<double> lval(W, 0.0);
  for (uint32_t c = 0; c < W; ++c) {
    CellSrc s = cell(1, c);
    std::string nm;
    lkind[c] = s.mode == CM_INT ? 1 : s.mode == CM_BOOL ? 2 : s.mode == CM_DATE ? 3 : s.mode == CM_TIME ? 4
               : s.mode == CM_FLOAT ? 5 : s.mode == CM_TDELTA ? 6 : 0;
    if (s.mode == CM_FLOAT) {   // float label: repr() as a name, f"{x: .6f}" trimmed when printed
      const double v = bits_as_double(s.bits);
      const Fixed6 f = fixed6(v);
      if (f.big || f.inf || dbl_is_nan(v)) return 2;
      uint8_t b[64];
      const int n = fixed_write(f, v, objfloat_dec(f), false, b);
      ftext[c].assign((const char*)b, n);
      nm = py_float_repr(v);
      lval[c] = v;
    }
    if (s.mode == CM_INT) lval[c] = (double)(long long)s.bits;
    if (s.mode == CM_BOOL) lval[c] = (s.bits & 1) ? 1.0 : 0.0;
    if (s.mode == CM_BOOL) nm = (s.bits & 1) ? "True" : "False";
    if (s.mode == CM_DATE || s.mode == CM_TIME || s.mode == CM_TDELTA) {
      uint8_t b[48];
      uint32_t n = s.mode == CM_DATE ? pydatetime_write((int64_t)s.bits, b)
                   : s.mode == CM_TIME ? pytime_write((int64_t)s.bits, b) : pytd_write((int64_t)s.bits, b);
      nm.assign((const char*)b, n);
    }
    if (s.mode != CM_INT) all_int = false;
    if (s.mode == CM_INT) nm = std::to_string((long long)s.bits);
    else if (s.mode == CM_RAW) nm.assign((const char*)s.p, s.n);
    else if (s.mode == CM_XML) { XmlIter it(s.p, s.n); int ch; while ((ch = it.next()) >= 0) nm.push_back((char)ch); }
    if (s.mode == CM_NULL) { nm = "Unnamed: " + std::to_string(c); unnamed.push_back(c); }
    raw[c] = nm;
  }
  if (!sheet_labels_supported(lkind, raw, &lval)) return 2;
  const std::vector<std::string> before = raw;
  {
    std::vector<size_t> order; std::vector<bool> un(W, false);
    for (size_t k : unnamed) un[k] = true;
    for (size_t k = 0; k < W; ++k) if (!un[k]) order.push_back(k);
    for (size_t k : unnamed) order.push_back(k);
    std::unordered_map<std::string, int> counts;
    for (size_t i2 : order) {
      std::string col = raw[i2], old = col; int cur = 0;
      auto f = counts.find(col); if (f != counts.end()) cur = f->second;
      while (cur > 0) {
        counts[old] = cur + 1;
        col = old + "." + std::to_string(cur);
        bool in = false; for (auto& x : raw) if (x == col) in = true;
        if (in) cur += 1; else { auto g = counts.find(col); cur = g == counts.end() ? 0 : g->second; }
      }
      raw[i2] = col; counts[col] = cur + 1;
    }
  }
  bool int_index = false;
  if (all_int) {  // Int64 Index: labels left-justified to one width (only if dedupe changed nothing)
    bool distinct = true;
    for (uint32_t a = 0; a < W; ++a) for (uint32_t b = a + 1; b < W; ++b) if (raw[a] == raw[b]) distinct = false;
    bool digits_only = true;
    for (auto& x : raw) for (char ch : x) if (!((ch >= '0' && ch <= '9') || ch == '-')) digits_only = false;
    if (distinct && digits_only) {
      int_index = true;
      size_t mx = 0;
      for (auto& x : raw) mx = std::max(mx, x.size());
      for (auto& x : raw) x.append(mx - x.size(), ' ');
    }
  }
  std::vector<uint32_t> ncp(W);
  bool names_valid = true;
  for (uint32_t c = 0; c < W; ++c) if (!utf8_valid_count(raw[c], &ncp[c])) names_valid = false;
  std::vector<std::string> disp = raw;
  if (names_valid && !int_index) {
    std::vector<std::string> shown = raw;
    for (uint32_t c = 0; c < W; ++c)
      if (lkind[c] == 5 && raw[c] == before[c]) { shown[c] = ftext[c]; no_blank[c] = ftext[c][0] == '-'; }
    display_labels(shown, disp, ncp, &no_blank);
  }
  // inference
  std::vector<uint32_t> xf(W, 0); std::vector<uint32_t> row_na(ndata, 0);
  for (uint32_t r = 0; r < ndata; ++r)
    for (uint32_t c = 0; c < W; ++c) {
      uint32_t f = xclassify_cell(cell(r + 2, c), r == 0);
      xf[c] |= f;
      if (f & XF_NULL) row_na[r]++;
    }
  std::vector<uint8_t> ctype(W);
  for (uint32_t c = 0; c < W; ++c) { ctype[c] = (uint8_t)xdecide_column(xf[c]); if (ctype[c] == CT_UNSUP || (xf[c] & (XF_BAD_UTF8 | XF_ENTITY))) return 2; }
  std::vector<bool> keep(ndata); uint32_t nkeep = 0;
  for (uint32_t r = 0; r < ndata; ++r) { keep[r] = row_na[r] != W; nkeep += keep[r]; }
  std::string text;
  if (nkeep == 0) text = empty_frame_text(raw);
  else {
    std::vector<ChunkMeasure> meas(W, ChunkMeasure{0, 0, 6, 0, 0, 0, 0, 0});
    for (uint32_t r = 0; r < ndata; ++r) {
      if (!keep[r]) continue;
      for (uint32_t c = 0; c < W; ++c) {
        CellMeasure m = measure_cell(cell(r + 2, c), ctype[c]);
        ChunkMeasure& M = meas[c];
        if (m.is_nan) { M.nan_left = 1; continue; }
        M.plain_len = std::max(M.plain_len, m.plain_len); M.ip_len = std::max(M.ip_len, m.ip_len);
        M.min_tz = std::min(M.min_tz, m.tz); M.inf_len = std::max(M.inf_len, m.inf_len);
        M.obj_len = std::max(M.obj_len, m.obj_len); M.sci_len = std::max(M.sci_len, m.sci_len); M.fl |= m.fl;
      }
    }
    std::vector<ColumnLayout> lay(W);
    for (uint32_t c = 0; c < W; ++c) {
      int cls = fold_column_class(&ctype[c], 1, W, meas[c].nan_left != 0);
      lay[c] = layout_column(cls, &meas[c], &ctype[c], 1, W, ncp[c]);
    }
    for (uint32_t c = 0; c < W; ++c) {
      bool numeric = lay[c].cls == CC_INT || lay[c].cls == CC_UINT || lay[c].cls == CC_FLOAT || lay[c].cls == CC_BOOL;
      uint32_t cp = ncp[c] + (numeric ? 1 : 0);
      if (c) text += ' ';
      text.append(lay[c].width - cp, ' ');
      if (numeric) text += ' ';
      text += disp[c];
    }
    std::vector<uint8_t> buf;
    for (uint32_t r = 0; r < ndata; ++r) {
      if (!keep[r]) continue;
      text += '\n';
      for (uint32_t c = 0; c < W; ++c) {
        CellSrc s = cell(r + 2, c);
        int ct = ctype[c];
        if (ct == CT_XOBJ) ct = cell_own_ctype(s);
        int kind = print_kind(lay[c].cls, ct, lay[c].sci);
        uint32_t cp, nb;
        printed_size(s, ct, kind, lay[c].dec, &cp, &nb);
        if (c) text += ' ';
        text.append(lay[c].width - cp, ' ');
        buf.resize(nb + 400);
        uint32_t wrote = print_cell(s, ct, kind, lay[c].dec, buf.data());
        if (wrote != nb) text += "<<SIZE MISMATCH>>";
        text.append((const char*)buf.data(), wrote);
      }
    }
  }
  if (text.size() > cap) return -1;
  memcpy(out, text.data(), text.size());
  *out_len = text.size();
  return 0;
}
}

// ---- the .json path ----------------------------------------------------------------------------
#include "../../replisense_rfq_b200/csrc/json_pretty.cuh"
extern "C" {
// returns status (0 ok, 1 invalid, 2 unsupported); info = {err, pos, line, col}; text written to out (cap bytes)
static int g_json_nested = 0;
void hc_json_nested(int on) { g_json_nested = on; }
int hc_json_pretty(const uint8_t* doc, uint32_t len, uint8_t* out, uint64_t cap, uint64_t* out_len, uint32_t* info) {
  std::vector<uint32_t> keys(2 * ((size_t)len / 4 + 2));
  rfq::JsonResult r;
  auto run = [&](uint8_t* o, rfq::JsonResult* rr) {
    if (g_json_nested) rfq::json_pretty_nested(doc, len, o, keys.data(), rr); else rfq::json_pretty(doc, len, o, keys.data(), rr);
  };
  run(nullptr, &r);
  info[0] = r.err; info[1] = r.pos; info[2] = r.line; info[3] = r.col;
  *out_len = r.out_bytes;
  if (r.status != rfq::JS_OK) return (int)r.status;
  if (r.out_bytes > cap) return -1;
  rfq::JsonResult r2;
  run(out, &r2);
  if (r2.status != rfq::JS_OK || r2.out_bytes != r.out_bytes) return -2;
  return 0;
}
}

