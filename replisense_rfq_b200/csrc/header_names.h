// Host-side (C++) handling of the header record of a CSV document: tokenise the first record,
// unescape, name blank columns, de-duplicate — what pandas' TextReader._get_header does
// (pandas/_libs/parsers.pyx; reached from app/file_parser.py:301) — and render the
// "Empty DataFrame" text of pandas/io/formats/string.py:56-62.  O(columns) per document; the
// data rows never pass through here.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <charconv>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "cell_parse.cuh"
#include "csv_dfa.cuh"

namespace rfq {

struct HeaderInfo {
  bool found = false;                 // a header record exists
  bool utf8_ok = true;
  std::vector<std::string> names;     // final (mangled) column names ("Empty DataFrame" prints these)
  std::vector<std::string> disp;      // labels as to_string prints them (display_labels)
  std::vector<uint32_t> ncp;          // code points of each display label
  uint32_t first_nf = 0;              // fields of the second record (0 = there is none): with the header it fixes the table width
};

inline bool utf8_valid_count(const std::string& s, uint32_t* ncp) {
  TextInfo t = scan_text(RawIter((const uint8_t*)s.data(), (uint32_t)s.size()));
  *ncp = t.ncp;
  return t.utf8_ok;
}

// Python's str.isspace() for one code point (the set str.lstrip() removes)
inline bool py_isspace_cp(uint32_t cp) {
  return (cp >= 0x09 && cp <= 0x0D) || (cp >= 0x1C && cp <= 0x20) || cp == 0x85 || cp == 0xA0 || cp == 0x1680 ||
         (cp >= 0x2000 && cp <= 0x200A) || cp == 0x2028 || cp == 0x2029 || cp == 0x202F || cp == 0x205F ||
         cp == 0x3000;
}

// Column labels as `to_string` prints them: pandas formats the columns Index through
// `Index._format_flat` -> `format_array(values, None, justify='all')` (every label becomes
// " " + pprint_thing(label, escape_chars=('\t', '\r', '\n'))) followed by `trim_front`, which cuts
// the smallest common run of leading whitespace (str.lstrip() semantics) from ALL labels
// (pandas/core/indexes/base.py `_format_with_header`, pandas/io/formats/format.py
// `_GenericArrayFormatter._format_strings`, pandas/io/formats/printing.py `trim_front`).  So TAB / CR /
// LF inside a label print as the two characters \t \r \n, and a header row whose names all start
// with blanks loses the common part.  The "Empty DataFrame" text prints the raw names instead.
// Labels must be valid UTF-8 (the caller checks); returns the code points of each display label.
// no_blank (optional): labels that format_array prints WITHOUT the blank (a negative float label of an
// object Index, "-4.5": the sign takes the place of the blank) — trim_front then finds nothing common and
// every other label keeps its blank.
inline void display_labels(const std::vector<std::string>& names, std::vector<std::string>& disp,
                           std::vector<uint32_t>& ncp, const std::vector<uint8_t>* no_blank = nullptr) {
  const size_t n = names.size();
  disp.assign(n, std::string());
  ncp.assign(n, 0);
  std::vector<uint32_t> lead_cp(n, 0), lead_bytes(n, 0);
  uint32_t cut = 0xFFFFFFFFu;
  for (size_t k = 0; k < n; ++k) {
    std::string& d = disp[k];
    for (char ch : names[k]) {
      if (ch == '\t') d += "\\t"; else if (ch == '\r') d += "\\r"; else if (ch == '\n') d += "\\n"; else d.push_back(ch);
    }
    // leading whitespace of the escaped label, in code points (+1 for the blank format_array adds)
    size_t i = 0; uint32_t cps = 0, tot = 0;
    bool in_lead = true;
    while (i < d.size()) {
      uint8_t b = (uint8_t)d[i];
      uint32_t cp = b; size_t l = 1;
      if (b >= 0xF0 && i + 3 < d.size()) { cp = ((b & 7u) << 18) | (((uint8_t)d[i + 1] & 0x3Fu) << 12) | (((uint8_t)d[i + 2] & 0x3Fu) << 6) | ((uint8_t)d[i + 3] & 0x3Fu); l = 4; }
      else if (b >= 0xE0 && i + 2 < d.size()) { cp = ((b & 0xFu) << 12) | (((uint8_t)d[i + 1] & 0x3Fu) << 6) | ((uint8_t)d[i + 2] & 0x3Fu); l = 3; }
      else if (b >= 0xC0 && i + 1 < d.size()) { cp = ((b & 0x1Fu) << 6) | ((uint8_t)d[i + 1] & 0x3Fu); l = 2; }
      if (in_lead && py_isspace_cp(cp)) { ++cps; lead_bytes[k] = (uint32_t)(i + l); } else in_lead = false;
      ++tot; i += l;
    }
    lead_cp[k] = cps; ncp[k] = tot;
    const uint32_t own = (no_blank && (*no_blank)[k]) ? cps : cps + 1;
    if (own < cut) cut = own;
  }
  if (n && cut == 0) {                // some label starts at once: the others keep the blank format_array added
    for (size_t k = 0; k < n; ++k)
      if (!(no_blank && (*no_blank)[k])) { disp[k].insert(disp[k].begin(), ' '); ncp[k] += 1; }
    return;
  }
  if (n == 0 || cut <= 1) return;     // only the blank format_array added is common
  cut -= 1;                           // code points to drop from every label
  for (size_t k = 0; k < n; ++k) {
    std::string& d = disp[k];
    size_t i = 0;
    const uint32_t drop = (no_blank && (*no_blank)[k]) ? cut + 1 : cut;
    for (uint32_t q = 0; q < drop && i < d.size(); ++q) {
      uint8_t b = (uint8_t)d[i];
      i += b >= 0xF0 ? 4 : b >= 0xE0 ? 3 : b >= 0xC0 ? 2 : 1;
    }
    d.erase(0, i);
    ncp[k] -= drop;
  }
}

// pandas' duplicate-name mangling (`a`, `a.1`, ...): named columns first, then the unnamed ones
// (pandas/_libs/parsers.pyx _get_header; pandas/io/parsers/python_parser.py _infer_columns)
inline void dedupe_names(std::vector<std::string>& raw, const std::vector<size_t>& unnamed) {
  std::vector<size_t> order;
  {
    std::vector<bool> is_un(raw.size(), false);
    for (size_t k : unnamed) is_un[k] = true;
    for (size_t k = 0; k < raw.size(); ++k) if (!is_un[k]) order.push_back(k);
    for (size_t k : unnamed) order.push_back(k);
  }
  std::unordered_map<std::string, int> counts;
  auto in_header = [&](const std::string& s) {
    for (auto& x : raw) if (x == s) return true;
    return false;
  };
  for (size_t i2 : order) {
    std::string col = raw[i2], old_col = col;
    int cur = 0;
    auto f = counts.find(col);
    if (f != counts.end()) cur = f->second;
    while (cur > 0) {
      counts[old_col] = cur + 1;
      col = old_col + "." + std::to_string(cur);
      if (in_header(col)) cur += 1;
      else { auto g = counts.find(col); cur = g == counts.end() ? 0 : g->second; }
    }
    raw[i2] = col;
    counts[col] = cur + 1;
  }
}

// The first two records of a document without building anything: where the header record lies, how
// many fields it and the record after it have.  O(bytes of two records), no allocation — what decides
// whether an (encoding, separator) attempt is worth sending to the device at all.
struct HeaderScan {
  bool found = false;          // a header record exists
  uint64_t begin = 0, end = 0; // bytes of the header record: [begin, end), end = its terminator (or len)
  uint32_t h = 0;              // its fields
  uint32_t first_nf = 0;       // fields of the second record (0 = none)
  uint64_t hash = 0;           // FNV-1a of the header record's bytes (which documents share a header)
};
// the DFA as tables: byte -> class for one separator, (state, class) -> next state / events
struct DfaTables {
  uint8_t cls_[256];
  uint8_t next_[8][8], ev_[8][8];
  explicit DfaTables(uint32_t sep) {
    for (int b = 0; b < 256; ++b) cls_[b] = (uint8_t)csv_class((uint32_t)b, sep);
    for (int st = 0; st < 8; ++st)
      for (int c = 0; c < 8; ++c) {
        next_[st][c] = c < N_CLASSES ? (uint8_t)csv_next(st, c) : (uint8_t)st;
        ev_[st][c] = c < N_CLASSES ? (uint8_t)csv_events(st, c) : 0;
      }
  }
  int cls(uint32_t b) const { return cls_[b]; }
  int next(int st, int c) const { return next_[st][c]; }
  int ev(int st, int c) const { return ev_[st][c]; }
};
// the same automaton straight from its definition (csv_dfa.cuh): what a GPU thread uses
struct DfaDirect {
  uint32_t sep;
  RFQ_HD int cls(uint32_t b) const { return csv_class(b, sep); }
  RFQ_HD int next(int st, int c) const { return csv_next(st, c); }
  RFQ_HD int ev(int st, int c) const { return csv_events(st, c); }
};
template <class Dfa>
RFQ_HD HeaderScan scan_header(const uint8_t* doc, uint64_t len, const Dfa& T, bool latin1 = false) {
  HeaderScan r;
  r.found = false; r.begin = 0; r.end = 0; r.h = 0; r.first_nf = 0; r.hash = 0;
  uint64_t i = 0;
  if (!latin1 && len >= 3 && doc[0] == 0xEF && doc[1] == 0xBB && doc[2] == 0xBF) i = 3;
  int st = ST_R;
  uint64_t ws_start = 0;
  bool open = false, done = false;
  for (; i <= len && !done; ++i) {
    const int cl = (i == len) ? CL_LF : T.cls(doc[i]);
    const int ev = T.ev(st, cl);
    const int ns = T.next(st, cl);
    if (st != ST_W && ns == ST_W) ws_start = i;
    if (ev) {
      if (ev & EV_REC_START) { open = true; r.begin = i; }
      if (ev & EV_REC_START_WS) { open = true; r.begin = ws_start; }
      if (ev & EV_FIELD_END) ++r.h;
      if (ev & EV_LINE_END) { done = true; r.end = i; }
    }
    st = ns;
  }
  if (!done || !open) return r;
  r.found = true;
  bool done2 = false;
  uint32_t nf2 = 0;
  for (; i <= len && !done2; ++i) {
    const int cl = (i == len) ? CL_LF : T.cls(doc[i]);
    const int ev = T.ev(st, cl);
    if (ev & EV_FIELD_END) ++nf2;
    if (ev & EV_LINE_END) done2 = true;
    st = T.next(st, cl);
  }
  r.first_nf = done2 ? nf2 : 0;
  uint64_t h = 1469598103934665603ull;
  for (uint64_t q = r.begin; q < r.end; ++q) { h ^= doc[q]; h *= 1099511628211ull; }
  r.hash = h;
  return r;
}
inline HeaderScan scan_header(const uint8_t* doc, uint64_t len, uint32_t sep, bool latin1 = false) {
  return scan_header(doc, len, DfaTables(sep), latin1);
}

// fn(i) for i in [0, n) on a few host threads (the per-document host work of big batches)
template <class F>
inline void parallel_for(uint32_t n, F&& fn) {
  unsigned hw = std::thread::hardware_concurrency();
  const uint32_t nt = n < 8192 ? 1u : std::min<uint32_t>(hw ? hw : 4u, 16u);
  if (nt <= 1) { for (uint32_t i = 0; i < n; ++i) fn(i); return; }
  std::vector<std::thread> th;
  const uint32_t per = (n + nt - 1) / nt;
  for (uint32_t t = 0; t < nt; ++t)
    th.emplace_back([=, &fn] { for (uint32_t i = t * per; i < std::min<uint32_t>(n, (t + 1) * per); ++i) fn(i); });
  for (auto& x : th) x.join();
}

// Tokenise the first record of `doc` with the same DFA the kernels run.
// latin1: the document bytes are latin-1 (the stream pandas tokenises is their UTF-8 transcode:
// same ASCII structure, no BOM possible, high bytes become two-byte sequences in the names).
inline HeaderInfo parse_header(const uint8_t* doc, uint64_t len, uint32_t sep, bool latin1 = false) {
  HeaderInfo h;
  uint64_t i = 0;
  if (!latin1 && len >= 3 && doc[0] == 0xEF && doc[1] == 0xBB && doc[2] == 0xBF) i = 3;
  int st = ST_R;
  std::vector<std::pair<uint64_t, uint64_t>> spans;
  uint64_t fstart = 0, ws_start = 0;
  bool open = false, done = false;
  for (; i <= len && !done; ++i) {
    int cl = (i == len) ? CL_LF : csv_class(doc[i], sep);
    int ev = csv_events(st, cl);
    if (st != ST_W && csv_next(st, cl) == ST_W) ws_start = i;
    if (ev & EV_REC_START) { open = true; fstart = i; }
    if (ev & EV_REC_START_WS) { open = true; fstart = ws_start; }
    if (ev & EV_FIELD_END) { spans.push_back({fstart, i}); fstart = i + 1; }
    if (ev & EV_LINE_END) done = true;
    st = csv_next(st, cl);
  }
  if (!done || !open) return h;
  {   // the second record's field count (pandas sizes the table from the header and the first data row)
    bool done2 = false;
    uint32_t nf2 = 0;
    for (; i <= len && !done2; ++i) {
      int cl = (i == len) ? CL_LF : csv_class(doc[i], sep);
      int ev = csv_events(st, cl);
      if (ev & EV_FIELD_END) ++nf2;
      if (ev & EV_LINE_END) done2 = true;
      st = csv_next(st, cl);
    }
    h.first_nf = done2 ? nf2 : 0;
  }
  h.found = true;
  std::vector<std::string> raw;
  std::vector<size_t> unnamed;
  for (size_t k = 0; k < spans.size(); ++k) {
    CsvIter it(doc + spans[k].first, (uint32_t)(spans[k].second - spans[k].first));
    std::string s;
    int c;
    while ((c = it.next()) >= 0) {
      if (latin1 && c >= 0x80) { s.push_back((char)(0xC0 | (c >> 6))); s.push_back((char)(0x80 | (c & 0x3F))); }
      else s.push_back((char)c);
    }
    if (s.empty()) { s = "Unnamed: " + std::to_string(k); unnamed.push_back(k); }
    raw.push_back(s);
  }
  dedupe_names(raw, unnamed);
  h.names = raw;
  h.ncp.resize(raw.size());
  for (size_t k = 0; k < raw.size(); ++k)
    if (!utf8_valid_count(raw[k], &h.ncp[k])) h.utf8_ok = false;
  if (h.utf8_ok) display_labels(h.names, h.disp, h.ncp);
  else h.disp = h.names;
  return h;
}

// Header row of a sheet: labels keep their Python type (pandas/io/parsers/python_parser.py
// _infer_columns).  kind per label: 0 str (or blank -> "Unnamed: k"), 1 int, 2 bool, 3 datetime,
// 4 time, 5 float (text = repr(label)), 6 timedelta (alone: a TimedeltaIndex), anything else = a type whose Index formatting is not restated
// (NaN).  `value` = the number behind an int / bool / float label.
// Supported: all labels int (an Int64 Index), or an object Index — labels of two type families or more,
// any str, any time: every label prints as str(label), a float as f"{x: .6f}" with its zeros trimmed
// (format.py _GenericArrayFormatter).  Declined: Index dtypes pandas formats as a whole (all bool, all
// datetime, int + float = a Float64 Index), labels that are ONE dict key for the de-duplication under
// different types (True == 1 == 1.0), and equal text under different types (2024 vs "2024").
inline bool sheet_labels_supported(const std::vector<int>& kind, const std::vector<std::string>& text,
                                   const std::vector<double>* value = nullptr) {
  // type families: an Index of ONE family other than str / time has its own dtype and formatter
  // (bool, int64, float64 = int + float, datetime64); two families or more, any str, any time -> object
  uint32_t fam = 0;
  bool has_float = false;
  for (size_t i = 0; i < kind.size(); ++i) {
    const int k = kind[i];
    if (k == 0) fam |= 1u; else if (k == 1) fam |= 2u; else if (k == 5) { fam |= 2u; has_float = true; }
    else if (k == 2) fam |= 4u; else if (k == 3) fam |= 8u; else if (k == 4) fam |= 16u; else if (k == 6) fam |= 32u;
    else return false;
  }
  const bool all_int = fam == 2u && !has_float;
  const bool object_index = (fam & 1u) || (fam & 16u) || (fam & (fam - 1u)) != 0;
  if (!all_int && !object_index) return false;
  for (size_t i = 0; i < kind.size(); ++i)
    for (size_t j = i + 1; j < kind.size(); ++j) {
      if (kind[i] == kind[j]) continue;
      if (text[i] == text[j]) return false;
      // True == 1 == 1.0 is ONE dict key for the de-duplication
      const bool ni = kind[i] == 1 || kind[i] == 2 || kind[i] == 5, nj = kind[j] == 1 || kind[j] == 2 || kind[j] == 5;
      if (ni && nj && (!value || (*value)[i] == (*value)[j])) return false;
    }
  return true;
}

// repr(float) of CPython (float_repr_style 'short': the shortest digits that round-trip, Objects/floatobject.c
// float_repr -> PyOS_double_to_string(x, 'r', 0, Py_DTSF_ADD_DOT_0)): fixed notation for 1e-4 <= |x| < 1e16,
// else d[.ddd]e+XX with at least two exponent digits.  std::to_chars gives the same shortest digits.
inline std::string py_float_repr(double v) {
  if (v != v) return "nan";
  if (v == 1.0 / 0.0) return "inf";
  if (v == -1.0 / 0.0) return "-inf";
  char buf[64];
  auto r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
  std::string t(buf, r.ptr);                       // [-]d[.ddd]e[+-]XX
  std::string out;
  size_t i = 0;
  if (t[i] == '-') { out = "-"; ++i; }
  std::string digits;
  size_t e = t.find('e');
  for (size_t k = i; k < e; ++k) if (t[k] != '.') digits.push_back(t[k]);
  const int exp10 = atoi(t.c_str() + e + 1);
  if (exp10 < -4 || exp10 >= 16) {
    out.push_back(digits[0]);
    if (digits.size() > 1) { out.push_back('.'); out.append(digits, 1, std::string::npos); }
    char eb[16];
    snprintf(eb, sizeof(eb), "e%c%02d", exp10 < 0 ? '-' : '+', exp10 < 0 ? -exp10 : exp10);
    return out + eb;
  }
  if (exp10 < 0) {
    out += "0.";
    out.append((size_t)(-exp10 - 1), '0');
    return out + digits;
  }
  if (digits.size() <= (size_t)exp10 + 1) {
    out += digits;
    out.append((size_t)exp10 + 1 - digits.size(), '0');
    return out + ".0";
  }
  out.append(digits, 0, (size_t)exp10 + 1);
  out.push_back('.');
  out.append(digits, (size_t)exp10 + 1, std::string::npos);
  return out;
}

// "Empty DataFrame\nColumns: [a, b]\nIndex: []" (display.max_seq_items = 100)
inline std::string empty_frame_text(const std::vector<std::string>& names) {
  std::string s = "Empty DataFrame\nColumns: [";
  size_t n = names.size(), lim = n < 100 ? n : 100;
  for (size_t k = 0; k < lim; ++k) { if (k) s += ", "; s += names[k]; }
  if (n > lim) s += ", ...";
  s += "]\nIndex: []";
  return s;
}

}  // namespace rfq
