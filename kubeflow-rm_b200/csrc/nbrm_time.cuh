// nbrm_time.cuh — time.Parse(time.RFC3339, s) and t.Format(time.RFC3339) in
// integer arithmetic (Go stdlib behaviour the culling path relies on,
// components/notebook-controller/controllers/culling_controller.go:178,195,
// 298-314,339-344,472,502; pinned by culling_controller_test.go:187,197,207).
#pragma once
#include "nbrm_json.cuh"

namespace nbrm {

struct Time {
  long long secs;   // Unix seconds (UTC)
  uint32_t nanos;   // fractional part
  int off;          // zone offset in seconds, kept for Format
  bool ok;
};

NB_HD long long days_from_civil(long long y, unsigned m, unsigned d) {
  y -= m <= 2;
  const long long era = (y >= 0 ? y : y - 399) / 400;
  const unsigned yoe = (unsigned)(y - era * 400);
  const unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + (long long)doe - 719468;
}

NB_HD void civil_from_days(long long z, int& y, unsigned& m, unsigned& d) {
  z += 719468;
  const long long era = (z >= 0 ? z : z - 146096) / 146097;
  const unsigned doe = (unsigned)(z - era * 146097);
  const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const long long yy = (long long)yoe + era * 400;
  const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const unsigned mp = (5 * doy + 2) / 153;
  d = doy - (153 * mp + 2) / 5 + 1;
  m = mp < 10 ? mp + 3 : mp - 9;
  y = (int)(yy + (m <= 2));
}

NB_HD int digits(const Doc& d, uint32_t at, uint32_t end, int w) {  // -1 when not w digits
  if (at + (uint32_t)w > end) return -1;
  int v = 0;
  for (int k = 0; k < w; ++k) {
    char c = d.at(at + k);
    if (c < '0' || c > '9') return -1;
    v = v * 10 + (c - '0');
  }
  return v;
}

// s = the bytes between the quotes of a JSON string.  A string holding any
// escape is not a timestamp (RFC 3339 has no character that needs one).
NB_HD Time parse_rfc3339(const Doc& d, Span s) {
  Time t;
  t.secs = 0;
  t.nanos = 0;
  t.off = 0;
  t.ok = false;
  const uint32_t b = s.off, e = s.end();
  if (s.len < 20) return t;
  const int y = digits(d, b, e, 4), mo = digits(d, b + 5, e, 2), dd = digits(d, b + 8, e, 2);
  const int h = digits(d, b + 11, e, 2), mi = digits(d, b + 14, e, 2), sec = digits(d, b + 17, e, 2);
  if (y < 0 || mo < 0 || dd < 0 || h < 0 || mi < 0 || sec < 0) return t;
  if (d.at(b + 4) != '-' || d.at(b + 7) != '-' || d.at(b + 10) != 'T' || d.at(b + 13) != ':' ||
      d.at(b + 16) != ':')
    return t;
  if (mo < 1 || mo > 12 || h > 23 || mi > 59 || sec > 59) return t;
  const bool leap = (y % 4 == 0) && (y % 100 != 0 || y % 400 == 0);
  const int mdays[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
  const int dim = mdays[mo - 1] + ((mo == 2 && leap) ? 1 : 0);
  if (dd < 1 || dd > dim) return t;
  uint32_t i = b + 19;
  uint32_t ns = 0;
  if (i < e && (d.at(i) == '.' || d.at(i) == ',')) {
    uint32_t j = i + 1, nd = 0;
    while (j < e && d.at(j) >= '0' && d.at(j) <= '9') {
      if (nd < 9) {
        ns = ns * 10 + (uint32_t)(d.at(j) - '0');
        ++nd;
      }
      ++j;
    }
    if (j == i + 1) return t;
    while (nd < 9) {
      ns *= 10;
      ++nd;
    }
    i = j;
  }
  if (i >= e) return t;
  int off = 0;
  if (d.at(i) == 'Z') {
    if (i + 1 != e) return t;
  } else if (d.at(i) == '+' || d.at(i) == '-') {
    if (i + 6 != e || d.at(i + 3) != ':') return t;
    const int oh = digits(d, i + 1, e, 2), om = digits(d, i + 4, e, 2);
    if (oh < 0 || om < 0 || oh > 23 || om > 59) return t;
    off = (oh * 3600 + om * 60) * (d.at(i) == '+' ? 1 : -1);
  } else {
    return t;
  }
  t.secs = days_from_civil(y, (unsigned)mo, (unsigned)dd) * 86400 + h * 3600 + mi * 60 + sec - off;
  t.nanos = ns;
  t.off = off;
  t.ok = true;
  return t;
}

NB_HD bool time_after(const Time& a, const Time& b) {
  return a.secs > b.secs || (a.secs == b.secs && a.nanos > b.nanos);
}

NB_HD void put2(char* o, unsigned v) {
  o[0] = (char)('0' + v / 10);
  o[1] = (char)('0' + v % 10);
}

// t.Format(time.RFC3339): whole seconds; "Z" for UTC else +hh:mm.  Returns the
// number of bytes written (20 or 25).
NB_HD uint32_t format_rfc3339(long long secs, int off, char* o) {
  const long long local = secs + off;
  long long days = local / 86400;
  long long rem = local - days * 86400;
  if (rem < 0) {
    rem += 86400;
    --days;
  }
  int y;
  unsigned m, dd;
  civil_from_days(days, y, m, dd);
  o[0] = (char)('0' + (y / 1000) % 10);
  o[1] = (char)('0' + (y / 100) % 10);
  o[2] = (char)('0' + (y / 10) % 10);
  o[3] = (char)('0' + y % 10);
  o[4] = '-';
  put2(o + 5, m);
  o[7] = '-';
  put2(o + 8, dd);
  o[10] = 'T';
  put2(o + 11, (unsigned)(rem / 3600));
  o[13] = ':';
  put2(o + 14, (unsigned)((rem % 3600) / 60));
  o[16] = ':';
  put2(o + 17, (unsigned)(rem % 60));
  if (off == 0) {
    o[19] = 'Z';
    return 20;
  }
  const int a = off < 0 ? -off : off;
  o[19] = off < 0 ? '-' : '+';
  put2(o + 20, (unsigned)(a / 3600));
  o[22] = ':';
  put2(o + 23, (unsigned)((a % 3600) / 60));
  return 25;
}

}  // namespace nbrm
