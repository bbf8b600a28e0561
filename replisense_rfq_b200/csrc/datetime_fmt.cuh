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

// ---- timedelta-styled numbers ([h]:mm:ss and friends) ----------------------------------------------
// from_excel(value, timedelta=True) of openpyxl 3.1 (utils/datetime.py):
//     td = datetime.timedelta(days=value)
//     if td.microseconds: td = timedelta(seconds=td.total_seconds() // 1, microseconds=round(td.microseconds, -3))
// over CPython's timedelta arithmetic (Modules/_datetimemodule.c accum(): the whole days are exact, the
// fraction is ONE double product frac * 86400e6 rounded half-even to a whole microsecond).  Accepted:
// |value| < 150 000 days: total_seconds() is the correctly rounded double of an exact quotient whose
// fraction is at most 0.999999, and below 2^34 seconds that rounding cannot reach the next whole second,
// so its floor is the exact integer floor taken here; anything longer is declined by the caller.
#define RFQ_US_PER_DAY 86400000000ll
RFQ_HD bool excel_to_timedelta(double value, int64_t* us_out) {
  if (!(value > -150000.0 && value < 150000.0)) return false;
  double ip;
  const double fr = modf(value, &ip);
  int64_t total = (int64_t)ip * RFQ_US_PER_DAY + rfq_rint_ll(RFQ_DMUL(fr, 86400000000.0));
  int64_t us = total % 1000000ll;
  if (us < 0) us += 1000000ll;
  if (us) {
    const int64_t w = (total - us) / 1000000ll;                    // total_seconds() // 1 (floor: us >= 0)
    const int64_t r = us % 1000ll;
    int64_t base = us - r;                                         // round(us, -3): half to even
    if (r > 500 || (r == 500 && ((base / 1000ll) & 1ll))) base += 1000ll;
    total = w * 1000000ll + base;
  }
  *us_out = total;
  return true;
}
// (days, seconds of the day, microseconds) with 0 <= seconds < 86400, 0 <= us < 1e6 (floor division)
struct TdParts { int64_t days; int h, mi, s; int us; };
RFQ_HD TdParts td_parts(int64_t total) {
  TdParts p;
  int64_t d = total / RFQ_US_PER_DAY, rem = total % RFQ_US_PER_DAY;
  if (rem < 0) { rem += RFQ_US_PER_DAY; d -= 1; }
  p.days = d;
  p.us = (int)(rem % 1000000ll);
  const int sec = (int)(rem / 1000000ll);
  p.h = sec / 3600; p.mi = (sec / 60) % 60; p.s = sec % 60;
  return p;
}
RFQ_HD uint32_t put_i64(uint8_t* o, int64_t v) {   // decimal digits with a '-' for negatives; returns the length
  uint8_t tmp[24];
  uint32_t n = 0;
  uint64_t m = v < 0 ? 0ull - (uint64_t)v : (uint64_t)v;
  do { tmp[n++] = (uint8_t)('0' + m % 10); m /= 10; } while (m);
  uint32_t k = 0;
  if (v < 0) o[k++] = '-';
  while (n) o[k++] = tmp[--n];
  return k;
}
RFQ_HD uint32_t put_us6(uint8_t* o, int us) { for (int q = 5; q >= 0; --q) { o[q] = (uint8_t)('0' + us % 10); us /= 10; } return 6; }
// A timedelta64 column as to_string prints it (pandas/io/formats/format.py get_format_timedelta64 ->
// Timedelta._repr_base): res 0 = every value of the column is whole days -> "{days} days"; res 1 =
// "{days} days HH:MM:SS" (" +" before the hours of a negative one), ".ffffff" where THIS value has a
// fraction of a second.
RFQ_HD uint32_t td64_write(int64_t total, int res, uint8_t* o) {
  const TdParts p = td_parts(total);
  uint32_t k = put_i64(o, p.days);
  o[k++] = ' '; o[k++] = 'd'; o[k++] = 'a'; o[k++] = 'y'; o[k++] = 's';
  if (res == 0) return k;
  o[k++] = ' ';
  if (p.days < 0) o[k++] = '+';
  put2(o + k, p.h); o[k + 2] = ':'; put2(o + k + 3, p.mi); o[k + 5] = ':'; put2(o + k + 6, p.s); k += 8;
  if (p.us) { o[k++] = '.'; k += put_us6(o + k, p.us); }
  return k;
}
RFQ_HD uint32_t td64_len(int64_t total, int res) { uint8_t b[48]; return td64_write(total, res, b); }
// str(datetime.timedelta): "[D day[s], ]H:MM:SS[.ffffff]" — what an object column or a label prints
RFQ_HD uint32_t pytd_write(int64_t total, uint8_t* o) {
  const TdParts p = td_parts(total);
  uint32_t k = 0;
  if (p.days) {
    k = put_i64(o, p.days);
    o[k++] = ' '; o[k++] = 'd'; o[k++] = 'a'; o[k++] = 'y';
    if (p.days != 1 && p.days != -1) o[k++] = 's';
    o[k++] = ','; o[k++] = ' ';
  }
  k += put_i64(o + k, p.h);
  o[k++] = ':'; put2(o + k, p.mi); k += 2; o[k++] = ':'; put2(o + k, p.s); k += 2;
  if (p.us) { o[k++] = '.'; k += put_us6(o + k, p.us); }
  return k;
}
RFQ_HD uint32_t pytd_len(int64_t total) { uint8_t b[48]; return pytd_write(total, b); }

}  // namespace rfq
