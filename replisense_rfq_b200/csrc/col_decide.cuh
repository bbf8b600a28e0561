// Column typing and layout decisions, shared by the kernels and the CPU test harness.
//
// Restates (a) pandas' C-parser dtype inference per `low_memory` chunk — try int64, on overflow
// uint64, then float64, then bool, else string; NA strings → NaN with int→float / bool→object
// upcasts (pandas/_libs/parsers.pyx `_convert_tokens`/`_convert_with_dtype`, chunking in
// `_read_low_memory`, buffer_lines heuristic in `TextReader.__cinit__`), (b) chunk
// concatenation (`_concatenate_chunks`, pandas/io/parsers/c_parser_wrapper.py:336-372),
// (c) `dropna(how='all').fillna('')` (app/file_parser.py:303) and (d) the column-wide rules of
// `to_string`: width, trailing-zero trim, fixed↔scientific switch
// (pandas/io/formats/format.py:1348-1477, 1816-1849).
#pragma once
#include "rfq_common.cuh"

namespace rfq {

// type of one column within one chunk
enum {
  CT_INT = 0,     // int64, no NA
  CT_INTNA = 1,   // int64 tokens + NA -> float64 (values converted from the integers)
  CT_UINT = 2,    // uint64
  CT_FLOAT = 3,   // float64 (tokens parsed by precise_xstrtod; NA -> NaN); also the all-NA chunk
  CT_BOOL = 4,    // bool, no NA
  CT_BOOLNA = 5,  // object: True / False / NaN
  CT_STR = 6,     // str, NA -> NaN
  CT_STRNF = 7,   // str with the NA filter off (uint64 conflict): nothing is NaN
  CT_UNSUP = 8,   // Python-int object columns (beyond uint64) — not produced on the GPU path
  CT_XOBJ = 9,    // sheet column of mixed Python objects: every cell prints as its own type
  CT_DATE = 11,   // sheet column whose non-blank cells are all datetimes: datetime64, blanks NaT
  CT_TIME = 12,   // (cell level only) a datetime.time inside a mixed object column
  CT_TDELTA = 13, // sheet column whose non-blank cells are all timedeltas: timedelta64, blanks NaT
  CT_PYINT = 10   // object column of Python ints (a value beyond uint64 / a negative next to an overflow):
                  // every token is int()-parsable and prints as str(int(token)); NA -> NaN
};

// printing class of a whole column
// CC_OBJF: float64 column with holes after fillna('') -> object column whose cells are all floats
enum { CC_INT = 0, CC_UINT = 1, CC_FLOAT = 2, CC_BOOL = 3, CC_STR = 4, CC_OBJ = 5, CC_OBJF = 6, CC_UNSUP = 7, CC_DATE = 8, CC_TDELTA = 9 };

// per (chunk, column) classification flags, OR-reduced over the cells
enum {
  SF_ANY_NA = 1u << 0,
  SF_ANY_NONNA = 1u << 1,
  SF_FLOAT_BAD = 1u << 2,    // a non-NA token that is neither a float nor an inf token
  SF_BOOL_BAD = 1u << 3,     // a non-NA token that is not a bool token
  SF_UINT_INVALID = 1u << 4, // a non-NA token str_to_uint64 rejects as invalid
  SF_UINT_OVF = 1u << 5,     // > UINT64_MAX
  SF_NEG = 1u << 6,          // a non-NA token whose first non-space char is '-'
  SF_GT_I64 = 1u << 7,       // a valid unsigned value > INT64_MAX
  SF_BAD_UTF8 = 1u << 8,     // a non-NA token that is not valid UTF-8
  SF_PYINT_BAD = 1u << 9     // a non-NA token Python's int() rejects
};
#define RFQ_NO_ROW 0xFFFFFFFFu

struct ChunkStat {
  uint32_t flags;
  uint32_t int_bad_min;   // first row whose non-NA token is not an int64 (invalid OR overflow)
  uint32_t int_ovf_min;   // first row whose non-NA token overflows int64
  uint32_t uint_inv_min;  // first row whose non-NA token str_to_uint64 rejects as invalid
  uint32_t uint_ovf_min;  // first row whose non-NA token exceeds UINT64_MAX
  uint32_t pad[3];
};

RFQ_HD int decide_chunk_type(const ChunkStat& s) {
  uint32_t f = s.flags;
  if (!(f & SF_ANY_NONNA)) return CT_FLOAT;  // every token NA: int64 pass succeeds, upcast to NaN
  if (s.int_bad_min == RFQ_NO_ROW) return (f & SF_ANY_NA) ? CT_INTNA : CT_INT;
  if (s.int_ovf_min == s.int_bad_min) {
    // the int64 pass stopped on an overflow -> uint64 pass, which stops at ITS first failure:
    // an invalid token sends the column on to the float pass, a value beyond uint64 raises
    // OverflowError -> Python-int pass (str(int(token)) objects, or plain strings if any token is
    // not int()-parsable)
    bool fall_to_float = false;
    if (s.uint_inv_min != RFQ_NO_ROW && s.uint_inv_min < s.uint_ovf_min) fall_to_float = true;
    else if (s.uint_ovf_min != RFQ_NO_ROW) return (f & SF_PYINT_BAD) ? CT_STR : CT_PYINT;
    else {
      bool conflict = (f & SF_GT_I64) && (f & (SF_NEG | SF_ANY_NA));
      if (conflict) return CT_STRNF;                       // ValueError: strings, NA filter off
      if (f & SF_NEG) return (f & SF_PYINT_BAD) ? CT_STR : CT_PYINT;   // OverflowError
      return CT_UINT;
    }
    (void)fall_to_float;
  }
  if (!(f & SF_FLOAT_BAD)) return CT_FLOAT;
  if (!(f & SF_BOOL_BAD)) return (f & SF_ANY_NA) ? CT_BOOLNA : CT_BOOL;
  return CT_STR;
}

// buffer_lines of the C parser for a table of `width` columns
RFQ_HD uint32_t chunk_rows_for_width(uint32_t width) {
  uint32_t heuristic = width ? (1u << 20) / width : 0u;
  uint32_t bl = 1;
  while ((uint64_t)bl * 2 < heuristic) bl *= 2;
  return bl;
}

// fold the chunk types of one column into its printing class.
// `any_nan_left`: a NaN survives dropna in this column (-> fillna('') makes it object).
RFQ_HD int fold_column_class(const uint8_t* ctypes, uint32_t n_chunks, uint32_t stride,
                             bool any_nan_left) {
  bool all_int = true, all_uint = true, all_num = true, all_bool = true, all_str = true;
  bool unsup = false, any_uint = false, all_date = true, all_td = true;
  for (uint32_t k = 0; k < n_chunks; ++k) {
    int t = ctypes[(size_t)k * stride];
    if (t == CT_UNSUP) unsup = true;
    if (t == CT_UINT) any_uint = true;
    if (t != CT_INT) all_int = false;
    if (t != CT_UINT) all_uint = false;
    if (!(t == CT_INT || t == CT_INTNA || t == CT_FLOAT)) all_num = false;
    if (t != CT_BOOL) all_bool = false;
    if (t != CT_STR) all_str = false;
    if (t != CT_DATE) all_date = false;
    if (t != CT_TDELTA) all_td = false;
  }
  if (unsup) return CC_UNSUP;
  if (n_chunks == 0) return CC_OBJ;  // no data rows: dtype object, prints nothing
  if (all_int) return CC_INT;
  if (all_uint) return CC_UINT;
  if (any_uint) return CC_UNSUP;  // uint64 mixed with other chunk dtypes: numpy promotion corner
  if (all_num) return any_nan_left ? CC_OBJF : CC_FLOAT;
  if (all_bool) return CC_BOOL;
  if (all_str) return CC_STR;
  if (all_date) return CC_DATE;
  if (all_td) return CC_TDELTA;
  return CC_OBJ;
}

// per (chunk, column) measurement, max/min/or-reduced over the non-NaN cells
struct ChunkMeasure {
  uint32_t plain_len;     // max len printed as own type: int digits, True/False, escaped string
  uint32_t ip_len;        // max (sign + integer digits) of finite values seen as float64
  uint32_t min_tz;        // min trailing zeros of the 6 decimals over finite values (6 if none)
  uint32_t inf_len;       // max len of inf tokens (3/4), 0 if none
  uint32_t obj_len;       // max len printed as object-column float
  uint32_t sci_len;       // max len in "%.6e"
  uint32_t fl;            // bit0 any |v| > 1e6 (incl. inf), bit1 any 0 < |v| < 1e-6, bit2 any finite
  uint32_t nan_left;      // a NaN cell in a row that survives dropna
};
enum { MF_LARGE = 1, MF_SMALL = 2, MF_FINITE = 4, MF_DT_TIME = 8, MF_DT_MS = 16 };  // MF_DT_*: datetime cells

// ---- one-sweep mode (cell_ops.cuh classify_measure_cell) ---------------------------------------
// candidates of one (chunk, column), reduced over its cells
struct ChunkCand {
  uint32_t str_len, int_len, uint_len, bool_len;
  uint32_t ip_len, min_tz, inf_len, obj_len, sci_len, fl;
  uint32_t na_count;     // NA cells (all rows)
  uint32_t need_exact;   // some cell's float view needs the exact sweep
};

// the ChunkMeasure the second sweep would have produced for a chunk of type `ctype`;
// returns false when that type is not covered by the candidates (exact sweep needed)
RFQ_HD bool finish_measure(const ChunkCand& c, int ctype, uint32_t dropped_rows, ChunkMeasure* out) {
  ChunkMeasure m;
  m.plain_len = 0; m.ip_len = 0; m.min_tz = 6; m.inf_len = 0; m.obj_len = 0; m.sci_len = 0; m.fl = 0;
  m.nan_left = c.na_count > dropped_rows ? 1u : 0u;   // a dropped row is NA in every column
  bool ok = true;
  switch (ctype) {
    case CT_STR: m.plain_len = c.str_len; break;
    case CT_BOOL: case CT_BOOLNA: m.plain_len = c.bool_len; break;
    case CT_UINT: m.plain_len = c.uint_len; break;
    case CT_INT: case CT_INTNA:
      m.plain_len = c.int_len;
      m.ip_len = c.ip_len; m.min_tz = c.min_tz; m.inf_len = c.inf_len; m.obj_len = c.obj_len;
      m.sci_len = c.sci_len; m.fl = c.fl;
      break;
    case CT_FLOAT:
      if (c.need_exact) ok = false;
      m.ip_len = c.ip_len; m.min_tz = c.min_tz; m.inf_len = c.inf_len; m.obj_len = c.obj_len;
      m.sci_len = c.sci_len; m.fl = c.fl;
      break;
    default: ok = false; break;       // CT_STRNF (NA words print), CT_PYINT, CT_UNSUP
  }
  *out = m;
  return ok;
}

struct ColumnLayout {
  int cls;
  int dec;        // decimals for CC_FLOAT fixed
  bool sci;       // CC_FLOAT printed "%.6e"
  uint32_t width; // code points, header included
};

// combine the chunk measures of one column (stride = table width) given its class
RFQ_HD ColumnLayout layout_column(int cls, const ChunkMeasure* m, const uint8_t* ctypes,
                                  uint32_t n_chunks, uint32_t stride, uint32_t header_cp) {
  ColumnLayout L;
  L.cls = cls; L.dec = 6; L.sci = false;
  uint32_t w = 0;
  if (cls == CC_FLOAT) {
    uint32_t ip = 0, tz = 6, inf = 0, sci = 0, fl = 0;
    for (uint32_t k = 0; k < n_chunks; ++k) {
      const ChunkMeasure& c = m[(size_t)k * stride];
      if (c.ip_len > ip) ip = c.ip_len;
      if (c.min_tz < tz) tz = c.min_tz;
      if (c.inf_len > inf) inf = c.inf_len;
      if (c.sci_len > sci) sci = c.sci_len;
      fl |= c.fl;
    }
    int dec = 6 - (int)tz;
    if (dec < 1) dec = 1;
    if (!(fl & MF_FINITE)) dec = 6;
    uint32_t fixed_max = (fl & MF_FINITE) ? ip + 1 + (uint32_t)dec : 0;
    if (inf > fixed_max) fixed_max = inf;
    bool too_long = fixed_max > 12;
    L.sci = (fl & MF_SMALL) || (too_long && (fl & MF_LARGE));
    L.dec = dec;
    w = L.sci ? (sci > inf ? sci : inf) : fixed_max;
    header_cp += 1;  // numeric dtypes get " " + name
  } else if (cls == CC_INT || cls == CC_UINT || cls == CC_BOOL) {
    for (uint32_t k = 0; k < n_chunks; ++k)
      if (m[(size_t)k * stride].plain_len > w) w = m[(size_t)k * stride].plain_len;
    header_cp += 1;
  } else if (cls == CC_DATE) {
    // one format for the column, from the finest resolution present (NaT never sets the width)
    uint32_t fl = 0, nat = 0;
    for (uint32_t k = 0; k < n_chunks; ++k) { fl |= m[(size_t)k * stride].fl; nat |= m[(size_t)k * stride].nan_left; }
    L.dec = (fl & MF_DT_MS) ? 2 : (fl & MF_DT_TIME) ? 1 : 0;
    w = (fl & MF_FINITE) ? (L.dec == 0 ? 10u : L.dec == 1 ? 19u : 23u) : (nat ? 3u : 0u);
  } else if (cls == CC_TDELTA) {
    // "{days} days" when every value is whole days, else the long form per value (NaT never sets the width)
    uint32_t fl = 0, nat = 0;
    for (uint32_t k = 0; k < n_chunks; ++k) { fl |= m[(size_t)k * stride].fl; nat |= m[(size_t)k * stride].nan_left; }
    L.dec = (fl & MF_DT_TIME) ? 1 : 0;
    for (uint32_t k = 0; k < n_chunks; ++k) {
      const uint32_t l = L.dec ? m[(size_t)k * stride].obj_len : m[(size_t)k * stride].ip_len;
      if (l > w) w = l;
    }
    if (!(fl & MF_FINITE)) w = nat ? 3u : 0u;
    else if (nat && w < 3u) w = 3u;
  } else if (cls == CC_STR) {
    for (uint32_t k = 0; k < n_chunks; ++k)
      if (m[(size_t)k * stride].plain_len > w) w = m[(size_t)k * stride].plain_len;
  } else {  // CC_OBJ: every chunk prints as its own type; float cells with the object rule
    for (uint32_t k = 0; k < n_chunks; ++k) {
      const ChunkMeasure& c = m[(size_t)k * stride];
      int t = ctypes[(size_t)k * stride];
      bool as_float = (cls == CC_OBJF) || t == CT_INTNA || t == CT_FLOAT;
      uint32_t l = as_float ? c.obj_len : c.plain_len;
      if (t == CT_XOBJ && cls != CC_OBJF && c.obj_len > l) l = c.obj_len;
      if (l > w) w = l;
    }
  }
  L.width = w > header_cp ? w : header_cp;
  return L;
}

}  // namespace rfq
