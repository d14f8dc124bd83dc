// Time model of the query path: ISO-8601 instants/intervals, half-open ms
// intervals, Druid granularities and Joda "timeFormat" patterns.
//   intervals:     /root/reference/src/main/scala/org/sparklinedata/druid/QueryIntervals.scala:40-130
//   granularities: /root/reference/src/main/scala/org/sparklinedata/druid/DruidQueryGranularity.scala:42-107
//   formats:       /root/reference/src/main/scala/org/sparklinedata/druid/DateTimeExtractor.scala:358-372
// Time zones: UTC and fixed offsets only (the reference's tests pin UTC,
// AbstractTest.scala:83-88); a named zone is rejected as unsupported.
#pragma once
#include <algorithm>
#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace sdo {

struct Unsupported : std::runtime_error {
  explicit Unsupported(const std::string& m) : std::runtime_error(m) {}
};
struct Invalid : std::runtime_error {
  explicit Invalid(const std::string& m) : std::runtime_error(m) {}
};

const int64_t MS_DAY = 86400000ll;

inline int64_t days_from_civil(int64_t y, int m, int d) {
  y -= m <= 2;
  const int64_t era = (y >= 0 ? y : y - 399) / 400;
  const unsigned yoe = (unsigned)(y - era * 400);
  const unsigned doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + (int64_t)doe - 719468;
}
inline void civil_from_days(int64_t z, int64_t* y, int* m, int* d) {
  z += 719468;
  const int64_t era = (z >= 0 ? z : z - 146096) / 146097;
  const unsigned doe = (unsigned)(z - era * 146097);
  const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const int64_t yy = (int64_t)yoe + era * 400;
  const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const unsigned mp = (5 * doy + 2) / 153;
  *d = (int)(doy - (153 * mp + 2) / 5 + 1);
  *m = (int)(mp < 10 ? mp + 3 : mp - 9);
  *y = yy + (*m <= 2);
}
inline int64_t floordiv(int64_t a, int64_t b) { return a / b - ((a % b != 0) && ((a < 0) != (b < 0))); }
inline int64_t floormod(int64_t a, int64_t b) { return a - floordiv(a, b) * b; }

struct Civil {
  int64_t year;
  int month, day, hour, minute, second, millis;
  int dow;  // 1 = Monday .. 7 = Sunday (ISO)
  int doy;  // 1-based
};
inline Civil civil_from_ms(int64_t ms, int64_t tz_offset_ms) {
  ms += tz_offset_ms;
  int64_t days = floordiv(ms, MS_DAY);
  int64_t rem = ms - days * MS_DAY;
  Civil c;
  civil_from_days(days, &c.year, &c.month, &c.day);
  c.hour = (int)(rem / 3600000);
  c.minute = (int)(rem / 60000 % 60);
  c.second = (int)(rem / 1000 % 60);
  c.millis = (int)(rem % 1000);
  c.dow = (int)floormod(days + 3, 7) + 1;  // 1970-01-01 was a Thursday
  c.doy = (int)(days - days_from_civil(c.year, 1, 1)) + 1;
  return c;
}

// Parse a fixed-offset zone id: "UTC", "Z", "GMT", "+05:30", "-0800", "UTC+1".
inline int64_t tz_offset_ms(const std::string& tz_in) {
  std::string tz = tz_in;
  if (tz.empty() || tz == "UTC" || tz == "Z" || tz == "GMT" || tz == "Etc/UTC" || tz == "UCT" || tz == "Etc/GMT") return 0;
  if (!tz.compare(0, 3, "UTC") || !tz.compare(0, 3, "GMT")) tz = tz.substr(3);
  if (tz.size() >= 2 && (tz[0] == '+' || tz[0] == '-')) {
    int sign = tz[0] == '-' ? -1 : 1;
    int h = 0, m = 0;
    std::string r = tz.substr(1);
    if (sscanf(r.c_str(), "%d:%d", &h, &m) >= 1) {
      if (r.find(':') == std::string::npos && r.size() == 4) { m = h % 100; h /= 100; }
      return sign * (h * 3600000ll + m * 60000ll);
    }
  }
  throw Unsupported("time zone '" + tz_in + "' is not a fixed offset");
}

// ---- time zones -----------------------------------------------------------------------------------------------------
// A zone is a fixed offset ("UTC", "+05:30") or a named region ("America/Los_Angeles": what the planner sends when
// spark.sparklinedata.tz.id / the JVM default is one, SQD/DruidPlanner.scala:73-76).  Named zones come from the system's
// TZif files (/usr/share/zoneinfo, or $TZDIR): the 64-bit transition table of the version-2+ block; instants past the last
// listed transition keep its offset (the installed files list transitions to 2037).
struct ZoneRules {
  std::vector<int64_t> at_ms;      // transition instants (UTC ms), ascending
  std::vector<int64_t> offset_ms;  // offset_ms[i] applies from at_ms[i-1] (or forever before) up to at_ms[i]; size = at_ms.size() + 1
};
struct Zone {
  int64_t fixed_ms = 0;
  std::shared_ptr<const ZoneRules> rules;  // null = fixed offset
  int64_t offset_at(int64_t utc_ms) const {
    if (!rules) return fixed_ms;
    const size_t k = (size_t)(std::upper_bound(rules->at_ms.begin(), rules->at_ms.end(), utc_ms) - rules->at_ms.begin());
    return rules->offset_ms[k];
  }
  // Local wall-clock ms -> UTC ms.  A time that occurs twice (clocks set back) means its first occurrence, one that does
  // not exist (clocks set forward) is read with the offset in force before the change -- what java.time / Joda's lenient
  // conversion and Python's fold = 0 do.
  int64_t local_to_utc(int64_t local_ms) const {
    if (!rules) return local_ms - fixed_ms;
    const int64_t before = offset_at(local_ms - 36 * 3600000ll), after = offset_at(local_ms + 36 * 3600000ll);
    const int64_t u1 = local_ms - before, u2 = local_ms - after;
    const bool ok1 = offset_at(u1) == before, ok2 = offset_at(u2) == after;
    if (ok1 && ok2) return std::min(u1, u2);
    if (ok1) return u1;
    if (ok2) return u2;
    return u1;
  }
};

inline std::shared_ptr<const ZoneRules> load_tzif(const std::string& id) {
  for (char c : id)
    if (!(isalnum((unsigned char)c) || c == '_' || c == '-' || c == '+' || c == '/')) throw Invalid("bad time zone id '" + id + "'");
  if (id.find("..") != std::string::npos || id.empty() || id[0] == '/') throw Invalid("bad time zone id '" + id + "'");
  const char* dir = getenv("TZDIR");
  const std::string path = std::string(dir && *dir ? dir : "/usr/share/zoneinfo") + "/" + id;
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) throw Unsupported("time zone '" + id + "' is not known (no " + path + ")");
  std::vector<unsigned char> b;
  unsigned char buf[4096];
  size_t n;
  while ((n = fread(buf, 1, sizeof buf, f)) > 0) b.insert(b.end(), buf, buf + n);
  fclose(f);
  auto be32 = [&](size_t o) { return (int64_t)(int32_t)((uint32_t)b[o] << 24 | (uint32_t)b[o + 1] << 16 | (uint32_t)b[o + 2] << 8 | (uint32_t)b[o + 3]); };
  auto be64 = [&](size_t o) {
    uint64_t v = 0;
    for (int k = 0; k < 8; k++) v = v << 8 | b[o + k];
    return (int64_t)v;
  };
  auto bad = [&]() -> Invalid { return Invalid("time zone file '" + path + "' is not a version 2+ TZif file"); };
  if (b.size() < 44 || memcmp(b.data(), "TZif", 4) != 0 || b[4] < '2') throw bad();
  // skip the version-1 block: header (44) + its data sized with 4-byte transition times
  size_t o = 0;
  const int64_t isut1 = be32(20), isstd1 = be32(24), leap1 = be32(28), time1 = be32(32), type1 = be32(36), char1 = be32(40);
  o = 44 + (size_t)(time1 * 4 + time1 + type1 * 6 + char1 + leap1 * 8 + isstd1 + isut1);
  if (b.size() < o + 44 || memcmp(b.data() + o, "TZif", 4) != 0) throw bad();
  const int64_t timecnt = be32(o + 32), typecnt = be32(o + 36);
  o += 44;
  if (typecnt < 1 || b.size() < o + (size_t)(timecnt * 9 + typecnt * 6)) throw bad();
  auto rules = std::make_shared<ZoneRules>();
  std::vector<int64_t> type_off((size_t)typecnt);
  const size_t o_idx = o + (size_t)timecnt * 8, o_types = o_idx + (size_t)timecnt;
  for (int64_t t = 0; t < typecnt; t++) type_off[(size_t)t] = be32(o_types + (size_t)t * 6) * 1000;
  // before the first transition: the first standard-time type (RFC 8536 3.2), i.e. the first type without the DST flag
  size_t first = 0;
  for (int64_t t = 0; t < typecnt; t++)
    if (b[o_types + (size_t)t * 6 + 4] == 0) { first = (size_t)t; break; }
  rules->offset_ms.push_back(type_off[first]);
  for (int64_t k = 0; k < timecnt; k++) {
    const unsigned idx = b[o_idx + (size_t)k];
    if ((int64_t)idx >= typecnt) throw bad();
    rules->at_ms.push_back(be64(o + (size_t)k * 8) * 1000);
    rules->offset_ms.push_back(type_off[idx]);
  }
  return rules;
}

// any zone id: a fixed offset, else a region of the system's zone database
inline Zone load_zone(const std::string& id) {
  Zone z;
  try {
    z.fixed_ms = tz_offset_ms(id);
    return z;
  } catch (const Unsupported&) {
  }
  z.rules = load_tzif(id);
  return z;
}

// ISO-8601 instant -> epoch ms.  Accepts yyyy-MM-dd['T'HH:mm[:ss[.SSS]]][Z|±HH:mm|±HHmm];
// a missing zone means `default_tz_ms`.
inline int64_t parse_iso_instant(const std::string& s, int64_t default_tz_ms = 0) {
  int y = 0, mo = 1, d = 1, h = 0, mi = 0, se = 0, ms = 0;
  const char* p = s.c_str();
  int n = 0;
  int sign = 1;
  if (*p == '-') { sign = -1; p++; }
  if (sscanf(p, "%d%n", &y, &n) != 1) throw Invalid("bad ISO instant '" + s + "'");
  y *= sign;
  p += n;
  if (*p == '-') {
    if (sscanf(p, "-%2d%n", &mo, &n) != 1) throw Invalid("bad ISO instant '" + s + "'");
    p += n;
    if (*p == '-') {
      if (sscanf(p, "-%2d%n", &d, &n) != 1) throw Invalid("bad ISO instant '" + s + "'");
      p += n;
    }
  }
  int64_t tz = default_tz_ms;
  if (*p == 'T' || *p == ' ') {
    p++;
    if (sscanf(p, "%2d%n", &h, &n) == 1) {
      p += n;
      if (*p == ':' && sscanf(p, ":%2d%n", &mi, &n) == 1) {
        p += n;
        if (*p == ':' && sscanf(p, ":%2d%n", &se, &n) == 1) {
          p += n;
          if (*p == '.' || *p == ',') {
            p++;
            int digits = 0, v = 0;
            while (*p >= '0' && *p <= '9') {
              if (digits < 3) { v = v * 10 + (*p - '0'); digits++; }
              p++;
            }
            while (digits++ < 3) v *= 10;
            ms = v;
          }
        }
      }
    }
  }
  if (*p == 'Z' || *p == 'z') { tz = 0; p++; }
  else if (*p == '+' || *p == '-') { tz = tz_offset_ms(p); p += strlen(p); }
  if (*p) throw Invalid("bad ISO instant '" + s + "'");
  if (mo < 1 || mo > 12 || d < 1 || d > 31 || h > 24 || mi > 59 || se > 60) throw Invalid("bad ISO instant '" + s + "'");
  return days_from_civil(y, mo, d) * MS_DAY + h * 3600000ll + mi * 60000ll + se * 1000ll + ms - tz;
}

struct Interval { int64_t start, end; };  // half-open [start, end) epoch ms
inline Interval parse_iso_interval(const std::string& s) {
  size_t k = s.find('/');
  if (k == std::string::npos) throw Invalid("interval '" + s + "' is not start/end");
  Interval iv{parse_iso_instant(s.substr(0, k)), parse_iso_instant(s.substr(k + 1))};
  if (iv.end < iv.start) throw Invalid("interval '" + s + "' ends before it starts");
  return iv;
}
inline std::string format_iso_instant(int64_t ms) {
  Civil c = civil_from_ms(ms, 0);
  char b[48];
  snprintf(b, sizeof b, "%04lld-%02d-%02dT%02d:%02d:%02d.%03dZ", (long long)c.year, c.month, c.day, c.hour, c.minute, c.second, c.millis);
  return b;
}

// ---- granularity ---------------------------------------------------------------------------
struct Granularity {
  enum Kind { All, None, Duration, Period } kind = All;
  int64_t duration_ms = 0;  // Duration
  int64_t origin_ms = 0;    // Duration / Period origin
  int months = 0;           // Period: calendar part (years*12 + months)
  int64_t period_ms = 0;    // Period: fixed part (weeks/days/hours/minutes/seconds)
  Zone zone;                // Period: buckets are aligned in this zone's local time
  bool has_origin = false;
};

// ISO-8601 period "PnYnMnWnDTnHnMnS" -> (months, fixed ms)
inline void parse_period(const std::string& s, int* months, int64_t* fixed_ms) {
  *months = 0;
  *fixed_ms = 0;
  if (s.size() < 2 || s[0] != 'P') throw Invalid("bad period '" + s + "'");
  bool time = false;
  size_t i = 1;
  while (i < s.size()) {
    if (s[i] == 'T') { time = true; i++; continue; }
    size_t j = i;
    while (j < s.size() && s[j] >= '0' && s[j] <= '9') j++;
    if (j == i || j >= s.size()) throw Invalid("bad period '" + s + "'");
    int64_t v = strtoll(s.substr(i, j - i).c_str(), nullptr, 10);
    char u = s[j];
    if (!time) {
      if (u == 'Y') *months += (int)v * 12;
      else if (u == 'M') *months += (int)v;
      else if (u == 'W') *fixed_ms += v * 7 * MS_DAY;
      else if (u == 'D') *fixed_ms += v * MS_DAY;
      else throw Invalid("bad period '" + s + "'");
    } else {
      if (u == 'H') *fixed_ms += v * 3600000ll;
      else if (u == 'M') *fixed_ms += v * 60000ll;
      else if (u == 'S') *fixed_ms += v * 1000ll;
      else throw Invalid("bad period '" + s + "'");
    }
    i = j + 1;
  }
  if (*months && *fixed_ms) throw Unsupported("period '" + s + "' mixes calendar and fixed units");
  if (!*months && !*fixed_ms) throw Invalid("empty period '" + s + "'");
}

// Druid "simple" granularity names
inline Granularity simple_granularity(const std::string& name_in) {
  std::string name;
  for (char c : name_in) name += (char)tolower((unsigned char)c);
  Granularity g;
  if (name == "all") { g.kind = Granularity::All; return g; }
  if (name == "none") { g.kind = Granularity::None; return g; }
  g.kind = Granularity::Period;
  if (name == "second") g.period_ms = 1000;
  else if (name == "minute") g.period_ms = 60000;
  else if (name == "fifteen_minute") g.period_ms = 900000;
  else if (name == "thirty_minute") g.period_ms = 1800000;
  else if (name == "hour") g.period_ms = 3600000;
  else if (name == "day") g.period_ms = MS_DAY;
  else if (name == "week") g.period_ms = 7 * MS_DAY;
  else if (name == "month") g.months = 1;
  else if (name == "quarter") g.months = 3;
  else if (name == "year") g.months = 12;
  else throw Invalid("unknown granularity '" + name_in + "'");
  return g;
}

// Start (epoch ms) of the bucket that contains t.
inline int64_t bucket_start(const Granularity& g, int64_t t) {
  switch (g.kind) {
    case Granularity::All: return INT64_MIN;
    case Granularity::None: return t;
    case Granularity::Duration: return t - floormod(t - g.origin_ms, g.duration_ms);
    case Granularity::Period: {
      const int64_t off = g.zone.offset_at(t);
      if (g.months) {
        Civil c = civil_from_ms(t, off);
        int64_t midx = c.year * 12 + (c.month - 1);
        int64_t o = 0;
        if (g.has_origin) {
          Civil oc = civil_from_ms(g.origin_ms, g.zone.offset_at(g.origin_ms));
          o = oc.year * 12 + (oc.month - 1);
          // an origin off a month boundary is not supported (Druid allows it; no caller uses it)
          if (oc.day != 1 || oc.hour || oc.minute || oc.second || oc.millis) throw Unsupported("period origin must be a month start");
        }
        int64_t b = midx - floormod(midx - o, g.months);
        return g.zone.local_to_utc(days_from_civil(floordiv(b, 12), (int)floormod(b, 12) + 1, 1) * MS_DAY);
      }
      // Fixed periods: aligned in local time.  Weeks start on Monday (ISO, as Joda).
      if (g.has_origin) return t - floormod(t - g.origin_ms, g.period_ms);
      if (g.period_ms < MS_DAY) return t - floormod(t + off, g.period_ms);  // sub-day buckets: offsets are multiples of them in practice
      const int64_t local = t + off;
      const int64_t local_origin = g.period_ms % (7 * MS_DAY) == 0 ? -3 * MS_DAY : 0;
      return g.zone.local_to_utc(local - floormod(local - local_origin, g.period_ms));
    }
  }
  return t;
}

// ---- Joda-style timeFormat ----------------------------------------------------------------
inline int iso_week(const Civil& c, int64_t* weekyear) {
  // ISO-8601 week number (Joda weekOfWeekyear)
  int wk = (c.doy - c.dow + 10) / 7;
  int64_t wy = c.year;
  auto leap = [](int64_t y) { return (y % 4 == 0 && y % 100 != 0) || y % 400 == 0; };
  auto weeks_in = [&](int64_t y) {
    int64_t jan1 = days_from_civil(y, 1, 1);
    int dow = (int)floormod(jan1 + 3, 7) + 1;
    return (dow == 4 || (dow == 3 && leap(y))) ? 53 : 52;
  };
  if (wk < 1) { wy = c.year - 1; wk = weeks_in(wy); }
  else if (wk > weeks_in(c.year)) { wk = 1; wy = c.year + 1; }
  if (weekyear) *weekyear = wy;
  return wk;
}

inline std::string joda_format(const std::string& fmt, int64_t ms, int64_t tz_ms) {
  static const char* MON[] = {"January", "February", "March", "April", "May", "June", "July", "August", "September", "October", "November", "December"};
  static const char* DOW[] = {"Monday", "Tuesday", "Wednesday", "Thursday", "Friday", "Saturday", "Sunday"};
  Civil c = civil_from_ms(ms, tz_ms);
  std::string o;
  char b[32];
  size_t i = 0;
  auto num = [&](int64_t v, int width) {
    snprintf(b, sizeof b, "%0*lld", width, (long long)v);
    o += b;
  };
  while (i < fmt.size()) {
    char ch = fmt[i];
    if (ch == '\'') {
      size_t j = i + 1;
      if (j < fmt.size() && fmt[j] == '\'') { o += '\''; i += 2; continue; }
      while (j < fmt.size() && fmt[j] != '\'') o += fmt[j++];
      i = j + 1;
      continue;
    }
    if (!isalpha((unsigned char)ch)) { o += ch; i++; continue; }
    size_t j = i;
    while (j < fmt.size() && fmt[j] == ch) j++;
    int n = (int)(j - i);
    switch (ch) {
      case 'y': case 'Y':
        if (n == 2) num(floormod(c.year, 100), 2);
        else num(c.year, n);
        break;
      case 'x': {
        int64_t wy;
        iso_week(c, &wy);
        if (n == 2) num(floormod(wy, 100), 2);
        else num(wy, n);
        break;
      }
      case 'M':
        if (n >= 4) o += MON[c.month - 1];
        else if (n == 3) o += std::string(MON[c.month - 1]).substr(0, 3);
        else num(c.month, n);
        break;
      case 'd': num(c.day, n); break;
      case 'D': num(c.doy, n); break;
      case 'w': num(iso_week(c, nullptr), n); break;
      case 'e': num(c.dow, n); break;
      case 'E':
        if (n >= 4) o += DOW[c.dow - 1];
        else o += std::string(DOW[c.dow - 1]).substr(0, 3);
        break;
      case 'H': num(c.hour, n); break;
      case 'k': num(c.hour == 0 ? 24 : c.hour, n); break;
      case 'h': num(c.hour % 12 == 0 ? 12 : c.hour % 12, n); break;
      case 'K': num(c.hour % 12, n); break;
      case 'a': o += c.hour < 12 ? "AM" : "PM"; break;
      case 'm': num(c.minute, n); break;
      case 's': num(c.second, n); break;
      case 'S': {
        int v = c.millis;
        if (n == 1) v /= 100;
        else if (n == 2) v /= 10;
        num(v, n < 3 ? n : 3);
        for (int k = 3; k < n; k++) o += '0';
        break;
      }
      case 'Z': {
        int64_t off = tz_ms;
        char sg = off < 0 ? '-' : '+';
        if (off < 0) off = -off;
        snprintf(b, sizeof b, n >= 2 ? "%c%02d:%02d" : "%c%02d%02d", sg, (int)(off / 3600000), (int)(off / 60000 % 60));
        o += b;
        break;
      }
      default:
        throw Unsupported(std::string("timeFormat pattern letter '") + ch + "' is not supported");
    }
    i = j;
  }
  return o;
}

}  // namespace sdo
