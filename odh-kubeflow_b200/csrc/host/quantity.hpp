// quantity.hpp — resource.Quantity as reflect.DeepEqual sees it.
//
// CopyStatefulSetFields compares PodSpecs with reflect.DeepEqual (COMMON:128),
// which looks at every field of resource.Quantity, exported or not:
//     i int64Amount{value int64, scale int32}; d infDecAmount{*inf.Dec};
//     s string (cached canonical text); Format
// so "1" != "1000m" and "1Gi" != "1024Mi" even though they are numerically
// equal.  The struct is produced by ParseQuantity during JSON decoding, which
// is upstream of the hot path and lives in k8s.io/apimachinery v0.33.7
// (pkg/api/resource/quantity.go) — a go.mod dependency (NBC/go.mod:14) that is
// NOT under the reference tree.  The restatement below follows the published
// algorithm of that file.  PARITY UNPINNED: the reference's tests only pin
// .String() round trips (ODH notebook_webhook_test.go:384-387), not the struct.
// The device-side parity boundary is therefore the struct tuple, which tests
// and the synthetic fleets can also supply directly (make_quantity()).
#pragma once
#include <cstdint>
#include <string>

#include "gval.hpp"

namespace nbh {

// Build the GVal of a Quantity from its fields.  `dec_unscaled` non-empty
// means d.Dec is a non-nil *inf.Dec with that unscaled decimal text and scale.
inline GVal make_quantity(int64_t value, int32_t scale, const std::string& format,
                          const std::string& s, const std::string& dec_unscaled = std::string(),
                          int32_t dec_scale = 0) {
  GVal q = GVal::strct();
  q.tag = kTagQuantity;
  GVal i = GVal::strct();
  i.set("value", GVal::integer(value));
  i.set("scale", GVal::integer(scale));
  q.set("i", std::move(i));
  if (!dec_unscaled.empty()) {
    GVal dec = GVal::strct();
    dec.set("unscaled", GVal::str(dec_unscaled));
    dec.set("scale", GVal::integer(dec_scale));
    GVal d = GVal::strct();
    d.set("Dec", GVal::ptr(std::move(dec)));
    q.set("d", std::move(d));
  }
  q.set("s", GVal::str(s));
  q.set("Format", GVal::str(format));
  return q;
}

namespace qdetail {

inline bool is_digit(char c) { return c >= '0' && c <= '9'; }

// parseQuantityString: sign, digits, optional '.', digits, suffix.
inline bool scan(const std::string& str, bool* positive, std::string* value, std::string* num,
                 std::string* denom, std::string* suffix) {
  *positive = true;
  size_t pos = 0, end = str.size();
  if (pos < end) {
    if (str[0] == '-') { *positive = false; ++pos; }
    else if (str[0] == '+') { ++pos; }
  }
  while (pos < end && str[pos] == '0') ++pos;           // strip leading zeros
  if (pos >= end) { *num = "0"; *value = *num; return true; }
  size_t i = pos;
  while (i < end && is_digit(str[i])) ++i;
  *num = str.substr(pos, i - pos);
  if (i >= end) { *value = str; return true; }
  pos = i;
  if (num->empty()) *num = "0";
  if (pos < end && str[pos] == '.') {
    ++pos;
    i = pos;
    while (i < end && is_digit(str[i])) ++i;
    *denom = str.substr(pos, i - pos);
    if (i >= end) { *value = str; return true; }
    pos = i;
  }
  *value = str.substr(0, pos);
  const size_t suffix_start = pos;
  i = pos;
  while (true) {
    if (i >= end) { *suffix = str.substr(suffix_start); return true; }
    if (std::string("eEinumkKMGTP").find(str[i]) == std::string::npos) { pos = i; break; }
    ++i;
  }
  if (pos < end && (str[pos] == '-' || str[pos] == '+')) ++pos;
  i = pos;
  while (true) {
    if (i >= end) { *suffix = str.substr(suffix_start); return true; }
    if (!is_digit(str[i])) break;
    ++i;
  }
  return false;  // ErrFormatWrong
}

inline bool interpret(const std::string& suf, int* base, int* exponent, std::string* format) {
  static const struct { const char* s; int e; } dec[] = {{"n", -9}, {"u", -6}, {"m", -3}, {"", 0},  {"k", 3},
                                                         {"M", 6},  {"G", 9},  {"T", 12}, {"P", 15}, {"E", 18}};
  static const struct { const char* s; int e; } bin[] = {{"Ki", 10}, {"Mi", 20}, {"Gi", 30},
                                                         {"Ti", 40}, {"Pi", 50}, {"Ei", 60}};
  for (const auto& d : dec) if (suf == d.s) { *base = 10; *exponent = d.e; *format = "DecimalSI"; return true; }
  for (const auto& b : bin) if (suf == b.s) { *base = 2; *exponent = b.e; *format = "BinarySI"; return true; }
  if (suf.size() > 1 && (suf[0] == 'E' || suf[0] == 'e')) {
    size_t k = 1;
    bool neg = false;
    if (suf[k] == '+' || suf[k] == '-') { neg = suf[k] == '-'; ++k; }
    if (k >= suf.size()) return false;
    long long v = 0;
    for (; k < suf.size(); ++k) {
      if (!is_digit(suf[k])) return false;
      v = v * 10 + (suf[k] - '0');
      if (v > 1000000) return false;
    }
    *base = 10; *exponent = (int)(neg ? -v : v); *format = "DecimalExponent";
    return true;
  }
  return false;
}

typedef unsigned __int128 u128;
inline std::string u128_str(u128 v) {
  if (v == 0) return "0";
  std::string s;
  while (v) { s.insert(s.begin(), (char)('0' + (int)(v % 10))); v /= 10; }
  return s;
}

}  // namespace qdetail

// The number a Quantity holds, normalised so that numerically equal quantities give the same
// (digits, exponent): value = digits x 10^exponent, digits without trailing zeros, "0" for zero.
// This is what equality.Semantic compares (Quantity.Cmp == 0).
inline void quantity_number(const GVal& q, std::string* digits, int64_t* exponent) {
  std::string d;
  int64_t e = 0;
  const GVal* dd = q.find("d");
  const GVal* dec = dd ? dd->find("Dec") : nullptr;
  if (dec && dec->kind == GVal::Ptr && !dec->items.empty()) {
    d = dec->items[0].get_str("unscaled");
    const GVal* sc = dec->items[0].find("scale");
    e = -(sc ? sc->i : 0);
  } else {
    const GVal* i = q.find("i");
    const GVal* v = i ? i->find("value") : nullptr;
    const GVal* sc = i ? i->find("scale") : nullptr;
    d = std::to_string(v ? v->i : 0);
    e = sc ? sc->i : 0;
  }
  bool neg = !d.empty() && d[0] == '-';
  if (neg) d.erase(0, 1);
  while (d.size() > 1 && d[0] == '0') d.erase(0, 1);
  while (d.size() > 1 && d.back() == '0') { d.pop_back(); ++e; }
  if (d.empty() || d == "0") { d = "0"; e = 0; neg = false; }
  *digits = (neg ? "-" : "") + d;
  *exponent = e;
}

// Quantity.Cmp(b): -1, 0, 1 by numeric value.
inline int quantity_cmp(const GVal& a, const GVal& b) {
  std::string da, db;
  int64_t ea, eb;
  quantity_number(a, &da, &ea);
  quantity_number(b, &db, &eb);
  const bool na = !da.empty() && da[0] == '-', nb = !db.empty() && db[0] == '-';
  if (na) da.erase(0, 1);
  if (nb) db.erase(0, 1);
  const bool za = da == "0", zb = db == "0";
  if (za || zb || na != nb) {
    const int sa = za ? 0 : (na ? -1 : 1), sb = zb ? 0 : (nb ? -1 : 1);
    return sa < sb ? -1 : sa > sb ? 1 : 0;
  }
  // same sign, both non-zero: compare magnitudes digits x 10^exponent
  int mag;
  const int64_t top_a = (int64_t)da.size() + ea, top_b = (int64_t)db.size() + eb;   // position of the leading digit
  if (top_a != top_b) mag = top_a < top_b ? -1 : 1;
  else {
    const size_t n = da.size() > db.size() ? da.size() : db.size();
    da.resize(n, '0'); db.resize(n, '0');
    mag = da < db ? -1 : da > db ? 1 : 0;
  }
  return na ? -mag : mag;
}
inline int quantity_sign(const GVal& q) {
  std::string d; int64_t e;
  quantity_number(q, &d, &e);
  return d == "0" ? 0 : (d[0] == '-' ? -1 : 1);
}

// ParseQuantity(str) -> GVal.  Returns false where the Go function returns an error.
inline bool parse_quantity(const std::string& str, GVal* out) {
  using namespace qdetail;
  if (str.empty()) return false;
  if (str == "0") { *out = make_quantity(0, 0, "", "0"); return true; }
  bool positive;
  std::string value, num, denom, suf;
  if (!scan(str, &positive, &value, &num, &denom, &suf)) return false;
  int base, exponent;
  std::string format;
  if (!interpret(suf, &base, &exponent, &format)) return false;

  int precision = 0, scale = 0;
  int64_t mantissa = 1;
  if (format != "BinarySI") {
    scale = exponent;
    precision = 18 - (int)(num.size() + denom.size());
  } else {
    if (exponent >= 0 && denom.empty()) {
      mantissa = (int64_t)1 << exponent;
      precision = 15 - (int)num.size() - (int)((float)exponent * 3 / 10) - 1;
    } else {
      precision = -1;
    }
  }
  if (precision >= 0) {
    scale -= (int)denom.size();
    if (scale >= -9) {
      const std::string shifted = num + denom;
      // strconv.ParseInt(shifted, 10, 64): at most 18 digits here, cannot overflow
      int64_t v = 0;
      for (char c : shifted) v = v * 10 + (c - '0');
      __int128 wide = (__int128)v * mantissa;
      if (wide <= INT64_MAX) {
        int64_t result = (int64_t)wide;
        if (!positive) result = -result;
        bool keep = false;
        if (format == "BinarySI") keep = (exponent % 10 == 0) && ((v & 7) != 0);
        else
          keep = (scale % 3 == 0) && !(shifted.size() >= 3 && shifted.compare(shifted.size() - 3, 3, "000") == 0) &&
                 shifted[0] != '0';
        *out = make_quantity(result, scale, format, keep ? str : std::string());
        return true;
      }
    }
  }
  // Slow path: inf.Dec.  amount = value (digits) * 10^exponent or * 2^exponent,
  // rounded UP to 9 decimal places; held exactly while it fits 128 bits.
  u128 unscaled = 0;
  int dscale = (int)denom.size();  // inf.Dec scale: digits after the point
  bool exact = true;
  for (char c : num + denom) {
    if (unscaled > (~(u128)0) / 10 - 10) { exact = false; break; }
    unscaled = unscaled * 10 + (unsigned)(c - '0');
  }
  if (exact) {
    if (base == 10) {
      dscale -= exponent;                                  // amount.SetScale(scale - exponent)
    } else {
      for (int k = 0; k < exponent && exact; ++k) {
        if (unscaled >> 126) exact = false;
        unscaled <<= 1;
      }
    }
  }
  if (exact && unscaled != 0 && dscale != 9) {
    // amount.Round(amount, 9, RoundUp): always lands on scale 9
    while (dscale < 9 && exact) {
      if (unscaled > (~(u128)0) / 10) exact = false; else { unscaled *= 10; ++dscale; }
    }
    bool inexact_digit = false;
    while (dscale > 9) { inexact_digit |= (unscaled % 10) != 0; unscaled /= 10; --dscale; }
    if (inexact_digit) unscaled += 1;                      // away from zero (sign handled apart)
  }
  if (!exact) {
    // beyond 128 bits: keep the text; equal texts still compare equal
    *out = make_quantity(0, 0, format, "", "raw:" + str, 0);
    return true;
  }
  if (format == "BinarySI") {
    // cap at maxAllowed = 2^63-1 (scale 0); values in (0,1) switch to DecimalSI
    u128 pow = 1;
    for (int k = 0; k < dscale; ++k) pow *= 10;
    if (dscale >= 0 && unscaled / pow > (u128)INT64_MAX) { unscaled = (u128)INT64_MAX; dscale = 0; }
    if (dscale > 0 && unscaled != 0 && unscaled < pow) format = "DecimalSI";
  }
  std::string digits = u128_str(unscaled);
  if (!positive && unscaled != 0) digits = "-" + digits;
  *out = make_quantity(0, 0, format, "", digits.empty() ? "0" : digits, dscale);
  return true;
}

}  // namespace nbh
