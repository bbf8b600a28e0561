// Per-element decoders of the sheet path (host/device): one <si> of sharedStrings, one <row>,
// one <c>, and the python-engine column inference pandas applies to the cell objects
// (pandas/io/parsers/base_parser.py:449-563 `_infer_types` -> lib.maybe_convert_numeric /
// ops.maybe_convert_bool; pandas/io/excel/_base.py:913-937).
#pragma once
#include "cell_ops.cuh"
#include "xlsx_core.cuh"

namespace rfq {

// document-level problems (never guessed: the sheet is reported RFQ_DOC_UNSUPPORTED)
enum {
  XU_BANG = 1u << 0,        // comment / CDATA / doctype inside the part
  XU_NO_REF = 1u << 1,      // malformed r= on a <row> or <c>
  XU_ORDER = 1u << 2,       // rows / columns not strictly increasing
  XU_DATE = 1u << 3,        // timedelta-styled number, negative date serial, t="d" text outside the accepted ISO forms
  XU_RICH_INLINE = 1u << 4, // inline string with rich-text runs
  XU_BAD_NUMBER = 1u << 5,  // numeric literal we cannot type exactly (overflow, nan, malformed)
  XU_BAD_SST = 1u << 6,     // shared-string index out of range / no table
  XU_ENTITY = 1u << 7,      // unknown entity
  XU_TYPE = 1u << 8,        // unknown t=
  XU_X005F = 1u << 9,       // 'x005F_' escape inside a shared string
  // not problems by themselves: which of its rows / cells carry r=.  openpyxl counts rows and
  // columns from the previous element when r= is missing (worksheet/_reader.py parse_row /
  // parse_cell); a part whose rows (cells) ALL lack r= is decoded by position, in a mix an element
  // without r= continues from the nearest one before it that has it
  XU_ROW_NOREF = 1u << 10, XU_ROW_HASREF = 1u << 11, XU_C_NOREF = 1u << 12, XU_C_HASREF = 1u << 13
};
#define RFQ_XU_INFO (XU_ROW_NOREF | XU_ROW_HASREF | XU_C_NOREF | XU_C_HASREF)
RFQ_HD bool xu_rows_by_position(uint32_t f) { return (f & XU_ROW_NOREF) && !(f & XU_ROW_HASREF); }
RFQ_HD bool xu_cells_by_position(uint32_t f) { return (f & XU_C_NOREF) && !(f & XU_C_HASREF); }
RFQ_HD bool xu_problem(uint32_t f) { return (f & ~(uint32_t)RFQ_XU_INFO) != 0; }

// bit 31 of a shared string's length (sst table, and the length half of a CM_RAW cell record): the string is
// plain ASCII without control bytes (span_is_clean), decided once by the kernel that decoded it
#define RFQ_SST_CLEAN 0x80000000u

struct XCell {
  uint32_t col;     // 1-based column
  uint32_t mode;    // CM_*
  uint64_t bits;    // value, or (offset << 32 | length) for text: offset into the string arena
                    // (CM_RAW) or into the document arena (CM_XML)
};

// ---- <si>: text of its <t> children and of the <t> of its <r> runs; phonetic runs excluded ----
// p points at "<si"; e = end of the document. Writes decoded bytes to out when non-null.
// Returns decoded length.
RFQ_HD uint32_t sst_decode_si(const uint8_t* p, const uint8_t* e, uint8_t* out, uint32_t* flags) {
  uint32_t n = 0;
  const uint8_t* q = p + 3;
  // self-closing <si/> ?
  while (q < e && *q != '>') ++q;
  if (q >= e) return 0;
  if (q[-1] == '/') return 0;
  ++q;
  int in_rph = 0;
  uint32_t x5 = 0;  // progress of the literal "x005F_" in the decoded stream
  while (q < e) {
    if (*q != '<') { ++q; continue; }   // text outside <t> is ignored
    int tg = xml_classify(q, e);
    if (tg == XT_SI_END) break;
    if (tg == XT_BANG) { *flags |= XU_BANG; }
    if (tg == XT_RPH) { const uint8_t* g = q; while (g < e && *g != '>') ++g; if (!(g < e && g[-1] == '/')) in_rph = 1; q = g; continue; }
    if (tg == XT_RPH_END) { in_rph = 0; ++q; continue; }
    if (tg == XT_T && !in_rph) {
      const uint8_t* g = q;
      g = find_byte(g, e, '>');
      if (g >= e) break;
      if (g[-1] == '/') { q = g; continue; }  // <t/>
      const uint8_t* tb = g + 1;
      const uint8_t* te = tb;
      while (te < e && *te != '<') ++te;
      XmlIter it(tb, (uint32_t)(te - tb));
      int c;
      while ((c = it.next()) >= 0) {
        if (c == "x005F_"[x5]) { if (++x5 == 6) { *flags |= XU_X005F; x5 = 0; } }
        else x5 = (c == 'x') ? 1 : 0;
        if (out) out[n] = (uint8_t)c;
        ++n;
      }
      if (it.bad) *flags |= XU_ENTITY;
      q = te;
      continue;
    }
    ++q;
  }
  return n;
}

// ---- <row ...>: its r= number ------------------------------------------------------------------
RFQ_HD uint32_t xlsx_row_number(const uint8_t* p, const uint8_t* e, uint32_t* flags) {
  const uint8_t* g = p;
  g = find_byte(g, e, '>');
  const uint8_t* vb; uint32_t vn;
  if (!xml_attr(p, g, 'r', &vb, &vn) || vn == 0) { *flags |= XU_ROW_NOREF; return 0; }
  *flags |= XU_ROW_HASREF;
  if (vn > 9) { *flags |= XU_NO_REF; return 0; }
  uint32_t r = 0;
  for (uint32_t i = 0; i < vn; ++i) { uint32_t d = vb[i] - '0'; if (d > 9) { *flags |= XU_NO_REF; return 0; } r = r * 10 + d; }
  if (r == 0) *flags |= XU_NO_REF;   // 0 is "no number yet" here
  return r;
}

struct SstView {             // shared-string table of the workbook this sheet belongs to
  const uint32_t* off;       // [n] offset of string i in the string arena
  const uint32_t* len;       // [n]
  uint32_t n;
};

// a date-styled number: from_excel; an out-of-range date is openpyxl's '#VALUE!' error cell -> NaN
RFQ_HD XCell xlsx_date_cell(XCell x, double v, bool date1904, uint32_t* flags) {
  if (v < 0.) { *flags |= XU_DATE; return x; }   // before the epoch: not produced by spreadsheets
  ExcelDate ed = excel_to_datetime(v, date1904);
  if (ed.kind == XD_ERROR) { x.mode = CM_FLOAT; x.bits = 0x7FF8000000000000ull; return x; }
  x.mode = ed.kind == XD_TIME ? CM_TIME : CM_DATE;
  x.bits = (uint64_t)ed.ms;
  return x;
}

// a timedelta-styled number: from_excel(value, timedelta=True); beyond the range the conversion is exact
// for (datetime_fmt.cuh excel_to_timedelta) the sheet is declined
RFQ_HD XCell xlsx_timedelta_cell(XCell x, double v, bool exact, uint32_t* flags) {
  int64_t us;
  // timedelta(days=v) beyond 999 999 999 days: OverflowError -> openpyxl's '#VALUE!' error cell -> NaN
  if (v >= 1.0e9 || v <= -1.0e9) { x.mode = CM_FLOAT; x.bits = 0x7FF8000000000000ull; return x; }
  if (!exact || !excel_to_timedelta(v, &us)) { *flags |= XU_DATE; return x; }
  x.mode = CM_TDELTA; x.bits = (uint64_t)us;
  return x;
}

// A numeric <v> that is not a plain decimal literal: int when it has no '.', 'e' or 'E' (openpyxl _cast_number),
// else float(text) exactly; date / timedelta styles applied.  x.col is kept.
RFQ_HD XCell xlsx_number_general(XCell x, const uint8_t* val_b, uint32_t val_n, int dstyle, bool date1904, uint32_t* flags) {
  x.mode = CM_NULL; x.bits = 0;
  bool is_float = false;
  for (uint32_t i = 0; i < val_n; ++i) if (val_b[i] == '.' || val_b[i] == 'e' || val_b[i] == 'E') is_float = true;
  if (!is_float) {
    int64_t iv;
    if (parse_int64(RawIter(val_b, val_n), &iv) != INT_OK) { *flags |= XU_BAD_NUMBER; return x; }
    if (dstyle == 2) return xlsx_timedelta_cell(x, (double)iv, iv > -(1ll << 53) && iv < (1ll << 53), flags);
    if (dstyle) return xlsx_date_cell(x, (double)iv, date1904, flags);
    x.mode = CM_INT; x.bits = (uint64_t)iv;
    return x;
  }
  double v;
  if (!strtod_exact(val_b, val_n, &v) || dbl_is_inf(v) || dbl_is_nan(v)) { *flags |= XU_BAD_NUMBER; return x; }
  if (dstyle == 2) return xlsx_timedelta_cell(x, v, true, flags);
  if (dstyle) return xlsx_date_cell(x, v, date1904, flags);
  // pandas _convert_cell: integral floats become ints
  if (v > -9.2e18 && v < 9.2e18) {
    int64_t iv = (int64_t)v;
    if ((double)iv == v) { x.mode = CM_INT; x.bits = (uint64_t)iv; return x; }
  } else { *flags |= XU_BAD_NUMBER; return x; }  // would be a Python big int
  x.mode = CM_FLOAT; x.bits = dbl_bits(v);
  return x;
}

// ---- <c ...>...</c> -----------------------------------------------------------------------------
// doc/len: the sheet part; cpos: offset of "<c". date_ids: sorted style ids that are dates.
RFQ_HD XCell xlsx_parse_cell(const uint8_t* doc, uint32_t len, uint32_t cpos, const SstView& sst,
                             const uint32_t* date_ids, uint32_t n_date, bool date1904,
                             uint32_t* flags, uint32_t* defer = nullptr) {
  XCell x; x.col = 0; x.mode = CM_NULL; x.bits = 0;
  const uint8_t* p = doc + cpos;
  const uint8_t* e = doc + len;
  const uint8_t* g = p;
  g = find_byte(g, e, '>');
  if (g >= e) return x;
  const uint8_t* vb; uint32_t vn;
  uint32_t row = 0;
  XAttrRTS at;
  xml_attrs_rts(p, g, &at);
  if (!(at.have & 1u) || at.vn[0] == 0) *flags |= XU_C_NOREF;   // col stays 0: set by position
  else if (!a1_parse(at.vb[0], at.vn[0], &x.col, &row)) { *flags |= XU_NO_REF; return x; }
  else *flags |= XU_C_HASREF;
  // type
  int ty = 'n';
  if (at.have & 2u) {
    vb = at.vb[1]; vn = at.vn[1];
    if (vn == 1) ty = vb[0];                                   // n s b e d
    else if (vn == 3 && vb[0] == 's' && vb[1] == 't' && vb[2] == 'r') ty = 'S';
    else if (vn == 9 && vb[0] == 'i') ty = 'I';               // inlineStr
    else { *flags |= XU_TYPE; return x; }
  }
  uint32_t style = 0;
  if (at.have & 4u) for (uint32_t i = 0; i < at.vn[2]; ++i) style = style * 10 + (at.vb[2][i] - '0');
  if (g[-1] == '/') return x;  // <c .../> : no value
  // children until </c>
  const uint8_t* q = g + 1;
  const uint8_t* val_b = nullptr; uint32_t val_n = 0; bool have_v = false;
  const uint8_t* is_b = nullptr; uint32_t is_n = 0; int n_t = 0; bool in_is = false;
  while (q < e) {
    if (*q != '<') { ++q; continue; }
    int tg = xml_classify(q, e);
    if (tg == XT_C_END || tg == XT_C || tg == XT_ROW_END) break;
    if (tg == XT_BANG) *flags |= XU_BANG;
    const uint8_t* h = q;
    h = find_byte(h, e, '>');
    if (h >= e) break;
    if (tg == XT_V && !have_v) {
      have_v = true;
      if (h[-1] != '/') { val_b = h + 1; const uint8_t* t2 = find_byte(val_b, e, '<'); val_n = (uint32_t)(t2 - val_b); }
    } else if (tg == XT_IS) { in_is = h[-1] != '/'; }
    else if (tg == XT_R && in_is) { *flags |= XU_RICH_INLINE; }
    else if (tg == XT_T && in_is) {
      if (h[-1] != '/' && n_t == 0) { is_b = h + 1; const uint8_t* t2 = find_byte(is_b, e, '<'); is_n = (uint32_t)(t2 - is_b); }
      ++n_t;
    }
    q = h + 1;
  }
  if (ty == 'I') {
    // value = Text.content of <is>; openpyxl sets it only when <is> exists; "" stays ""
    if (n_t > 1) *flags |= XU_RICH_INLINE;
    if (is_b && is_n) { x.mode = CM_XML; x.bits = ((uint64_t)(uint32_t)(is_b - doc) << 32) | is_n; }
    return x;
  }
  if (!val_b || val_n == 0) return x;  // findtext(...) or None -> None -> ""
  switch (ty) {
    case 'n': {
      // date_ids: cellXfs indices with a date number format; bit 31 marks a timedelta format
      int dstyle = 0;
      for (uint32_t i = 0; i < n_date; ++i)
        if ((date_ids[i] & 0x7FFFFFFFu) == style) dstyle = (date_ids[i] >> 31) ? 2 : 1;
      // The spellings writers emit — [-]digits[.digits], at most 18 digits in all — are read in ONE pass:
      // an integer as it stands, a fraction as (double)digits / 10^k (both exact doubles, one correctly
      // rounded division: what float(text) gives).  Everything else goes through the general parsers below.
      {
        uint32_t i = 0, nd = 0, nfrac = 0;
        const bool neg = val_n && val_b[0] == '-';
        if (neg) i = 1;
        uint64_t w = 0;
        bool frac = false, plain = i < val_n;
        for (; i < val_n; ++i) {
          const uint32_t ch = val_b[i];
          if (ch == '.' && !frac && nd) { frac = true; continue; }
          const uint32_t dg = ch - '0';
          if (dg > 9u || nd == 18u) { plain = false; break; }
          w = w * 10u + dg; ++nd;
          if (frac) ++nfrac;
        }
        if (plain && nd && !(frac && nfrac == 0)) {
          if (!frac) {
            const int64_t iv = neg ? -(int64_t)w : (int64_t)w;
            if (dstyle == 2) return xlsx_timedelta_cell(x, (double)iv, iv > -(1ll << 53) && iv < (1ll << 53), flags);
            if (dstyle) return xlsx_date_cell(x, (double)iv, date1904, flags);
            x.mode = CM_INT; x.bits = (uint64_t)iv;
            return x;
          }
          if (w < (1ull << 53)) {
            double v = RFQ_DDIV((double)w, RFQ_POW10((int)nfrac));
            if (neg) v = -v;
            if (dstyle == 2) return xlsx_timedelta_cell(x, v, true, flags);
            if (dstyle) return xlsx_date_cell(x, v, date1904, flags);
            const int64_t iv = (int64_t)v;                       // pandas _convert_cell: integral floats become ints
            if ((double)iv == v) { x.mode = CM_INT; x.bits = (uint64_t)iv; return x; }
            x.mode = CM_FLOAT; x.bits = dbl_bits(v);
            return x;
          }
        }
      }
      // everything else (exponents, 19+ digits, signs, blanks): the general parsers.  A caller that gave a
      // `defer` slot gets the literal back instead (mode CM_NUM_PENDING, bits = offset << 32 | length, *defer =
      // 1 + date style) and runs xlsx_number_general over all such cells together, with full warps.
      if (defer) {
        x.mode = CM_NUM_PENDING; x.bits = ((uint64_t)(uint32_t)(val_b - doc) << 32) | val_n;
        *defer = 1u + (uint32_t)dstyle;
        return x;
      }
      return xlsx_number_general(x, val_b, val_n, dstyle, date1904, flags);
    }
    case 's': {
      int64_t idx;
      if (parse_int64(RawIter(val_b, val_n), &idx) != INT_OK || idx < 0 || (uint64_t)idx >= sst.n) { *flags |= XU_BAD_SST; return x; }
      uint32_t l = sst.len[idx];   // bit 31 = RFQ_SST_CLEAN, kept in the cell record
      if ((l & ~RFQ_SST_CLEAN) == 0) return x;  // "" prints and trims like an empty cell
      x.mode = CM_RAW; x.bits = ((uint64_t)sst.off[idx] << 32) | l;
      return x;
    }
    case 'S': x.mode = CM_XML; x.bits = ((uint64_t)(uint32_t)(val_b - doc) << 32) | val_n; return x;
    case 'b': {
      int64_t iv;
      if (parse_int64(RawIter(val_b, val_n), &iv) != INT_OK) { *flags |= XU_BAD_NUMBER; return x; }
      x.mode = CM_BOOL; x.bits = iv != 0;
      return x;
    }
    case 'e': x.mode = CM_FLOAT; x.bits = 0x7FF8000000000000ull; return x;  // np.nan
    case 'd': {   // ISO 8601 text -> datetime.fromisoformat
      int64_t ms;
      if (!parse_iso_datetime(val_b, val_n, &ms)) { *flags |= XU_DATE; return x; }
      x.mode = CM_DATE; x.bits = (uint64_t)ms;
      return x;
    }
    default: *flags |= XU_TYPE; return x;
  }
}

// is the converted value "" (trimmed from the right of a row / from the bottom of the sheet)?
RFQ_HD bool xcell_is_empty(const XCell& x) { return x.mode == CM_NULL; }

// ---- python-engine inference ----------------------------------------------------------------
enum {
  XF_NULL = 1u << 0,          // "", NaN, NA string
  XF_FLOAT = 1u << 1,         // float object or float-looking string
  XF_INT = 1u << 2,           // int object or int-looking string
  XF_BOOL = 1u << 3,          // bool object
  XF_STR_NONNUM = 1u << 4,    // string that floatify rejects
  XF_TYPED_NUM = 1u << 5,     // int / float object (not bool-like for maybe_convert_bool)
  XF_STR_NOT_BOOLWORD = 1u << 6,
  XF_NON_STR = 1u << 7,       // any non-null non-string object
  XF_FIRST_IS_INT = 1u << 8,  // first data row holds a Python int (or bool)
  XF_BAD_UTF8 = 1u << 9,
  XF_ENTITY = 1u << 10,
  XF_BIG = 1u << 11,          // int-looking string beyond int64
  XF_DATE = 1u << 12,         // datetime.datetime object
  XF_TIME = 1u << 13,         // datetime.time object
  XF_TDELTA = 1u << 14        // datetime.timedelta object
};

RFQ_HD uint32_t xclassify_cell(const CellSrc& s, bool first_row) {
  if (s.missing || s.mode == CM_NULL) return XF_NULL;
  if (s.mode == CM_INT) return XF_INT | XF_TYPED_NUM | XF_NON_STR | (first_row ? XF_FIRST_IS_INT : 0);
  if (s.mode == CM_BOOL) return XF_BOOL | XF_NON_STR | (first_row ? XF_FIRST_IS_INT : 0);
  if (s.mode == CM_DATE) return XF_DATE | XF_NON_STR;
  if (s.mode == CM_TIME) return XF_TIME | XF_NON_STR;
  if (s.mode == CM_TDELTA) return XF_TDELTA | XF_NON_STR;
  if (s.mode == CM_FLOAT) {
    if (dbl_is_nan(bits_as_double(s.bits))) return XF_NULL;
    return XF_FLOAT | XF_TYPED_NUM | XF_NON_STR;
  }
  TextInfo t = cell_text(s);
  uint32_t f = 0;
  if (!t.utf8_ok) f |= XF_BAD_UTF8;
  if (is_na_token(t.nbytes, t.first8)) return f | XF_NULL;
  if (bool_token(t.nbytes, t.first8) < 0) f |= XF_STR_NOT_BOOLWORD;
  int64_t iv;
  int is = cell_int64(s, &iv);
  if (is == INT_OK) return f | XF_INT;
  double dv;
  if (cell_double(s, &dv)) return f | (is == INT_OVERFLOW ? XF_BIG : 0) | (dbl_is_nan(dv) ? XF_NULL : XF_FLOAT);
  return f | XF_STR_NONNUM;
}

RFQ_HD int xdecide_column(uint32_t f) {
  if (f & XF_BIG) return CT_UNSUP;
  // a datetime / time object makes maybe_convert_numeric raise TypeError and is no bool either;
  // the DataFrame constructor then turns an all-datetime (+NaN) object column into datetime64
  // (maybe_infer_to_datetimelike); the same for timedeltas -> timedelta64
  const bool dt = (f & (XF_DATE | XF_TIME | XF_TDELTA)) != 0;
  if (dt) {
    if ((f & XF_DATE) && !(f & (XF_TIME | XF_TDELTA | XF_FLOAT | XF_INT | XF_BOOL | XF_STR_NONNUM))) return CT_DATE;
    if ((f & XF_TDELTA) && !(f & (XF_DATE | XF_TIME | XF_FLOAT | XF_INT | XF_BOOL | XF_STR_NONNUM))) return CT_TDELTA;
    return CT_XOBJ;
  }
  if (!(f & XF_STR_NONNUM)) {            // lib.maybe_convert_numeric succeeds
    if (f & (XF_NULL | XF_FLOAT)) return CT_FLOAT;
    if (f & XF_INT) return CT_INT;
    if (f & XF_BOOL) return CT_BOOL;
    return CT_FLOAT;
  }
  // object column; ops.maybe_convert_bool unless the first value is a Python int
  if (!(f & XF_FIRST_IS_INT) && !(f & (XF_TYPED_NUM | XF_STR_NOT_BOOLWORD)))
    return (f & XF_NULL) ? CT_BOOLNA : CT_BOOL;
  if (!(f & XF_NON_STR)) return CT_STR;  // all strings (or NaN): str dtype
  return CT_XOBJ;
}

}  // namespace rfq
