// rfc3339.hpp — host-side time.Parse(time.RFC3339, s) / Format(time.RFC3339).
//
// The culler stores and receives instants as strings and parses them with Go's
// time.Parse(time.RFC3339, ...) (CULL:212,229,332,339,353,358); the device
// only ever sees (sec, nsec).  This parser follows the rules of Go's layout
// parser for "2006-01-02T15:04:05Z07:00" (Go 1.25 stdlib `time`, not under the
// reference tree, so stated from the published behaviour):
//   * 4-digit year, 2-digit month/day/minute/second, hour of ONE or two digits;
//   * literal '-', 'T', ':' exactly (lower-case 't'/'z' are rejected);
//   * month 1-12, day valid for month/leap year, hour 0-23, minute/second 0-59;
//   * optional fraction after the seconds, introduced by '.' or ',', at least
//     one digit, digits beyond the ninth ignored;
//   * zone 'Z', or sign hh ':' mm with hh <= 24 and mm <= 60;
//   * nothing may follow.
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>

namespace nbh {

struct Instant {
  int64_t sec = 0;     // Unix seconds (UTC)
  int32_t nsec = 0;    // 0..999999999
  int32_t offset = 0;  // zone offset east of UTC, seconds (kept for Format)
};

inline bool is_leap(int y) { return (y % 4 == 0) && (y % 100 != 0 || y % 400 == 0); }
inline int days_in(int m, int y) {
  static const int d[] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
  return (m == 2 && is_leap(y)) ? 29 : d[m - 1];
}
// days since 1970-01-01 of a proleptic Gregorian date
inline int64_t days_from_civil(int64_t y, int m, int d) {
  y -= m <= 2;
  const int64_t era = (y >= 0 ? y : y - 399) / 400;
  const int64_t yoe = y - era * 400;
  const int64_t doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + doe - 719468;
}
inline void civil_from_days(int64_t z, int* y, int* m, int* d) {
  z += 719468;
  const int64_t era = (z >= 0 ? z : z - 146096) / 146097;
  const int64_t doe = z - era * 146097;
  const int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const int64_t yy = yoe + era * 400;
  const int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const int64_t mp = (5 * doy + 2) / 153;
  *d = (int)(doy - (153 * mp + 2) / 5 + 1);
  *m = (int)(mp < 10 ? mp + 3 : mp - 9);
  *y = (int)(yy + (*m <= 2));
}

inline bool parse_rfc3339(const char* s, size_t n, Instant* out) {
  size_t p = 0;
  auto digit = [&](size_t k) { return k < n && s[k] >= '0' && s[k] <= '9'; };
  auto num2 = [&](int* v) {  // exactly two digits
    if (!digit(p) || !digit(p + 1)) return false;
    *v = (s[p] - '0') * 10 + (s[p + 1] - '0');
    p += 2;
    return true;
  };
  auto lit = [&](char c) { if (p < n && s[p] == c) { ++p; return true; } return false; };

  if (n < 4 || !digit(0) || !digit(1) || !digit(2) || !digit(3)) return false;
  const int year = (s[0] - '0') * 1000 + (s[1] - '0') * 100 + (s[2] - '0') * 10 + (s[3] - '0');
  p = 4;
  int month, day, hour, minute, second;
  if (!lit('-') || !num2(&month) || month < 1 || month > 12) return false;
  if (!lit('-') || !num2(&day)) return false;
  if (!lit('T')) return false;
  if (!digit(p)) return false;               // hour: one or two digits
  hour = s[p++] - '0';
  if (digit(p)) hour = hour * 10 + (s[p++] - '0');
  if (hour > 23) return false;
  if (!lit(':') || !num2(&minute) || minute > 59) return false;
  if (!lit(':') || !num2(&second) || second > 59) return false;
  int64_t nsec = 0;
  if (p + 1 < n && (s[p] == '.' || s[p] == ',') && digit(p + 1)) {
    ++p;
    int k = 0;
    while (digit(p)) { if (k < 9) { nsec = nsec * 10 + (s[p] - '0'); ++k; } ++p; }
    for (; k < 9; ++k) nsec *= 10;
  }
  int offset = 0;
  if (p < n && s[p] == 'Z') {
    ++p;
  } else {
    if (n - p < 6 || s[p + 3] != ':') return false;
    const char sign = s[p];
    ++p;
    int hh, mm;
    if (!num2(&hh)) return false;
    ++p;  // ':'
    if (!num2(&mm)) return false;
    if (hh > 24 || mm > 60) return false;
    offset = (hh * 60 + mm) * 60;
    if (sign == '-') offset = -offset;
    else if (sign != '+') return false;
  }
  if (p != n) return false;                  // extra text
  if (day < 1 || day > days_in(month, year)) return false;
  out->sec = days_from_civil(year, month, day) * 86400 + hour * 3600 + minute * 60 + second - offset;
  out->nsec = (int32_t)nsec;
  out->offset = offset;
  return true;
}
inline bool parse_rfc3339(const std::string& s, Instant* out) { return parse_rfc3339(s.data(), s.size(), out); }

// t.Format(time.RFC3339): whole seconds, 'Z' when the offset is zero.
inline std::string format_rfc3339(int64_t sec, int32_t offset) {
  const int64_t local = sec + offset;
  int64_t days = local / 86400, rem = local % 86400;
  if (rem < 0) { rem += 86400; --days; }
  int y, m, d;
  civil_from_days(days, &y, &m, &d);
  char buf[48];
  int k = snprintf(buf, sizeof buf, "%04d-%02d-%02dT%02d:%02d:%02d", y, m, d, (int)(rem / 3600),
                   (int)(rem % 3600 / 60), (int)(rem % 60));
  if (offset == 0) {
    buf[k++] = 'Z'; buf[k] = 0;
  } else {
    const int a = offset < 0 ? -offset : offset;
    snprintf(buf + k, sizeof buf - k, "%c%02d:%02d", offset < 0 ? '-' : '+', a / 3600, a % 3600 / 60);
  }
  return buf;
}

}  // namespace nbh
