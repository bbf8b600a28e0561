// Date-styled sheet numbers (host/device): Excel serial -> datetime / time as openpyxl's
// from_excel does it (openpyxl/utils/datetime.py, reached through pd.read_excel at
// app/file_parser.py:225), and the three ways pandas prints the result:
//   * a datetime64[us] column (every non-blank cell of the column is a datetime): one format for
//     the whole column, chosen from the finest resolution present — `%Y-%m-%d` when every value
//     is midnight, else `%Y-%m-%d %H:%M:%S`, plus `.mmm` when any value has milliseconds; blanks
//     stay `NaT` (pandas/_libs/tslib.pyx format_array_from_datetime, pandas 3.0.2 fillna('')
//     leaves datetime64 NaT in place);
//   * a datetime.datetime inside a mixed object column: str(datetime);
//   * a datetime.time (serials in [0, 1)): str(time).
// Instants are carried as milliseconds since 0001-01-01T00:00 (from_excel rounds to the
// millisecond), times as milliseconds of the day.
#pragma once
#include <math.h>
#include "rfq_common.cuh"

namespace rfq {

enum { XD_DATETIME = 0, XD_TIME = 1, XD_ERROR = 2 };
struct ExcelDate { int kind; int64_t ms; };

#define RFQ_MS_PER_DAY 86400000ll
// datetime.max = 9999-12-31 23:59:59.999999 -> last whole millisecond
#define RFQ_DT_MAX_MS (3652059ll * RFQ_MS_PER_DAY - 1)

RFQ_HD int64_t rfq_rint_ll(double x) {
#ifdef __CUDA_ARCH__
  return __double2ll_rn(x);
#else
  return (int64_t)__builtin_nearbyint(x);   // round-half-even in the default rounding mode
#endif
}

// from_excel(value, epoch) for a non-negative serial (an int serial arrives as the same double:
// divmod(int, 1) and divmod(float, 1) agree below 2^53, and anything larger is out of range).
RFQ_HD ExcelDate excel_to_datetime(double value, bool date1904) {
  ExcelDate r; r.kind = XD_ERROR; r.ms = 0;
  double day = floor(value);
  double frac = value - day;                                  // exact: fmod(value, 1)
  int64_t ms = rfq_rint_ll(RFQ_DMUL(RFQ_DMUL(frac, 86400.0), 1000.0));
  if (value < 1.0 && ms < RFQ_MS_PER_DAY) { r.kind = XD_TIME; r.ms = ms; return r; }
  if (day > 4.0e6) return r;                                  // OverflowError -> '#VALUE!' -> NaN
  int64_t d = (int64_t)day;
  if (value > 0.0 && value < 60.0 && !date1904) d += 1;       // the 1900 leap-year bug
  // ordinal (0001-01-01 = day 0) of the epochs: 1899-12-30 -> 693593, 1904-01-01 -> 695055
  int64_t total = ((date1904 ? 695055ll : 693593ll) + d) * RFQ_MS_PER_DAY + ms;
  if (total > RFQ_DT_MAX_MS) return r;
  r.kind = XD_DATETIME; r.ms = total;
  return r;
}

// t="d" cells: ISO 8601 text as datetime.fromisoformat reads it.  Accepted here: YYYY-MM-DD, optionally
// followed by 'T' or a blank and HH:MM[:SS[.f]] with 1-6 fraction digits that amount to whole
// milliseconds, years 1000-9999, no zone.  Returns false for every other spelling (declined by the
// caller, never guessed).  *ms = milliseconds since 0001-01-01T00:00.
RFQ_HD bool parse_iso_datetime(const uint8_t* s, uint32_t n, int64_t* ms) {
  auto dig = [&](uint32_t i) -> int { return (i < n && s[i] >= '0' && s[i] <= '9') ? s[i] - '0' : -1; };
  auto num = [&](uint32_t i, int k) -> int { int v = 0; for (int q = 0; q < k; ++q) { int d = dig(i + q); if (d < 0) return -1; v = v * 10 + d; } return v; };
  if (n < 10 || s[4] != '-' || s[7] != '-') return false;
  int y = num(0, 4), mo = num(5, 2), da = num(8, 2);
  if (y < 1000 || mo < 1 || mo > 12 || da < 1) return false;
  const int mdays[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
  bool leap = (y % 4 == 0 && y % 100 != 0) || y % 400 == 0;
  if (da > mdays[mo - 1] + ((mo == 2 && leap) ? 1 : 0)) return false;
  int h = 0, mi = 0, se = 0, frac_ms = 0;
  if (n > 10) {
    if (!(s[10] == 'T' || s[10] == ' ') || n < 16 || s[13] != ':') return false;
    h = num(11, 2); mi = num(14, 2);
    if (h < 0 || h > 23 || mi < 0 || mi > 59) return false;
    uint32_t i = 16;
    if (i < n) {
      if (s[i] != ':' || n < 19) return false;
      se = num(17, 2);
      if (se < 0 || se > 59) return false;
      i = 19;
      if (i < n) {
        if (s[i] != '.') return false;
        uint32_t nd = n - (i + 1);
        if (nd < 1 || nd > 6) return false;
        int f = 0;
        for (uint32_t q = 0; q < 6; ++q) { int d = q < nd ? dig(i + 1 + q) : 0; if (d < 0) return false; f = f * 10 + d; }
        if (f % 1000) return false;            // microseconds: a finer column resolution than we carry
        frac_ms = f / 1000;
      }
    }
  }
  // days since 0001-01-01 (proleptic Gregorian)
  int64_t yy = y - (mo <= 2 ? 1 : 0);
  int64_t era = yy / 400, yoe = yy - era * 400;
  int64_t doy = (153 * (mo + (mo > 2 ? -3 : 9)) + 2) / 5 + da - 1;
  int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  int64_t days = era * 146097 + doe - 306;
  *ms = days * RFQ_MS_PER_DAY + ((int64_t)h * 3600 + mi * 60 + se) * 1000 + frac_ms;
  return true;
}

struct Civil { int y, mo, d, h, mi, s, ms; };
RFQ_HD Civil civil_from_ms(int64_t total) {
  Civil c;
  int64_t days = total / RFQ_MS_PER_DAY;
  int64_t rem = total - days * RFQ_MS_PER_DAY;
  c.ms = (int)(rem % 1000); rem /= 1000;
  c.s = (int)(rem % 60); rem /= 60;
  c.mi = (int)(rem % 60); c.h = (int)(rem / 60);
  // days since 0001-01-01 -> days since 0000-03-01 (proleptic Gregorian, 400-year eras)
  int64_t z = days + 306;
  int64_t era = z / 146097;
  int64_t doe = z - era * 146097;
  int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  int64_t mp = (5 * doy + 2) / 153;
  c.d = (int)(doy - (153 * mp + 2) / 5 + 1);
  c.mo = (int)(mp < 10 ? mp + 3 : mp - 9);
  c.y = (int)(yoe + era * 400 + (c.mo <= 2 ? 1 : 0));
  return c;
}

RFQ_HD void put2(uint8_t* o, int v) { o[0] = (uint8_t)('0' + v / 10); o[1] = (uint8_t)('0' + v % 10); }
RFQ_HD uint32_t put_hms(uint8_t* o, const Civil& c) {
  put2(o, c.h); o[2] = ':'; put2(o + 3, c.mi); o[5] = ':'; put2(o + 6, c.s);
  return 8;
}
RFQ_HD uint32_t put_ymd(uint8_t* o, const Civil& c) {   // years 1..9999, zero padded to 4
  put2(o, c.y / 100); put2(o + 2, c.y % 100); o[4] = '-'; put2(o + 5, c.mo); o[7] = '-'; put2(o + 8, c.d);
  return 10;
}

// resolution of a datetime64 column: 0 = dates only, 1 = seconds, 2 = milliseconds
RFQ_HD uint32_t date64_len(int res) { return res == 0 ? 10u : res == 1 ? 19u : 23u; }
RFQ_HD uint32_t date64_write(int64_t total, int res, uint8_t* o) {
  Civil c = civil_from_ms(total);
  uint32_t k = put_ymd(o, c);
  if (res == 0) return k;
  o[k++] = ' ';
  k += put_hms(o + k, c);
  if (res == 2) { o[k++] = '.'; o[k++] = (uint8_t)('0' + c.ms / 100); put2(o + k, c.ms % 100); k += 2; }
  return k;
}
// str(datetime.datetime): microseconds (always a multiple of 1000 here) only when non-zero
RFQ_HD uint32_t pydatetime_len(int64_t total) { return total % 1000 ? 26u : 19u; }
RFQ_HD uint32_t pydatetime_write(int64_t total, uint8_t* o) {
  Civil c = civil_from_ms(total);
  uint32_t k = put_ymd(o, c);
  o[k++] = ' ';
  k += put_hms(o + k, c);
  if (c.ms) { o[k++] = '.'; o[k++] = (uint8_t)('0' + c.ms / 100); put2(o + k, c.ms % 100); k += 2; o[k++] = '0'; o[k++] = '0'; o[k++] = '0'; }
  return k;
}
// str(datetime.time)
RFQ_HD uint32_t pytime_len(int64_t ms) { return ms % 1000 ? 15u : 8u; }
RFQ_HD uint32_t pytime_write(int64_t ms, uint8_t* o) {
  Civil c = civil_from_ms(ms);
  uint32_t k = put_hms(o, c);
  if (c.ms) { o[k++] = '.'; o[k++] = (uint8_t)('0' + c.ms / 100); put2(o + k, c.ms % 100); k += 2; o[k++] = '0'; o[k++] = '0'; o[k++] = '0'; }
  return k;
}

}  // namespace rfq
