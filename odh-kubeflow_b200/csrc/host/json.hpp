// json.hpp — JSON text -> GVal, with the dynamic types the Kubernetes
// "unstructured" decoder produces (what CopyVirtualService sees, COMMON:199):
//   object -> Map (map[string]interface{}), array -> List ([]interface{}),
//   string -> Str, true/false -> Bool, null -> Null (nil interface),
//   number -> Int (int64) when it is an integer literal that fits, else Float.
// So 80 and 80.0 decode to different dynamic types, as in Go.
#pragma once
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>

#include "gval.hpp"

namespace nbh {

class JsonParser {
 public:
  JsonParser(const char* p, size_t n) : p_(p), end_(p + n) {}
  bool parse(GVal* out, std::string* err) {
    skip();
    if (!value(out, 0)) { if (err) *err = err_; return false; }
    skip();
    if (p_ != end_) { if (err) *err = "trailing characters after JSON value"; return false; }
    return true;
  }

 private:
  const char* p_;
  const char* end_;
  std::string err_;

  bool fail(const char* m) { if (err_.empty()) err_ = m; return false; }
  void skip() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_; }

  bool value(GVal* out, int depth) {
    if (depth > 200) return fail("nesting too deep");
    if (p_ >= end_) return fail("unexpected end");
    switch (*p_) {
      case '{': return object(out, depth);
      case '[': return array(out, depth);
      case '"': { out->kind = GVal::Str; return string(&out->s); }
      case 't': return lit("true", out, GVal::boolean(true));
      case 'f': return lit("false", out, GVal::boolean(false));
      case 'n': return lit("null", out, GVal::null());
      default: return number(out);
    }
  }
  bool lit(const char* w, GVal* out, GVal v) {
    const size_t n = strlen(w);
    if ((size_t)(end_ - p_) < n || memcmp(p_, w, n) != 0) return fail("bad literal");
    p_ += n; *out = std::move(v); return true;
  }
  bool number(GVal* out) {
    const char* s = p_;
    bool integral = true;
    if (p_ < end_ && *p_ == '-') ++p_;
    if (p_ >= end_ || *p_ < '0' || *p_ > '9') return fail("bad number");
    while (p_ < end_ && *p_ >= '0' && *p_ <= '9') ++p_;
    if (p_ < end_ && *p_ == '.') { integral = false; ++p_; while (p_ < end_ && *p_ >= '0' && *p_ <= '9') ++p_; }
    if (p_ < end_ && (*p_ == 'e' || *p_ == 'E')) {
      integral = false; ++p_;
      if (p_ < end_ && (*p_ == '+' || *p_ == '-')) ++p_;
      while (p_ < end_ && *p_ >= '0' && *p_ <= '9') ++p_;
    }
    const std::string txt(s, p_ - s);
    if (integral) {
      errno = 0;
      char* e = nullptr;
      const long long v = strtoll(txt.c_str(), &e, 10);
      if (errno == 0 && e && *e == 0) { *out = GVal::integer(v); return true; }
    }
    *out = GVal::real(strtod(txt.c_str(), nullptr));
    return true;
  }
  static void utf8(std::string* s, unsigned cp) {
    if (cp < 0x80) s->push_back((char)cp);
    else if (cp < 0x800) { s->push_back((char)(0xC0 | (cp >> 6))); s->push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
      s->push_back((char)(0xE0 | (cp >> 12))); s->push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
      s->push_back((char)(0x80 | (cp & 0x3F)));
    } else {
      s->push_back((char)(0xF0 | (cp >> 18))); s->push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
      s->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); s->push_back((char)(0x80 | (cp & 0x3F)));
    }
  }
  bool hex4(unsigned* v) {
    if (end_ - p_ < 4) return fail("bad \\u escape");
    unsigned r = 0;
    for (int k = 0; k < 4; ++k) {
      const char c = p_[k];
      r <<= 4;
      if (c >= '0' && c <= '9') r |= c - '0';
      else if (c >= 'a' && c <= 'f') r |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') r |= c - 'A' + 10;
      else return fail("bad \\u escape");
    }
    p_ += 4; *v = r; return true;
  }
  bool string(std::string* out) {
    ++p_;  // opening quote
    out->clear();
    while (p_ < end_) {
      const char c = *p_++;
      if (c == '"') return true;
      if (c != '\\') { out->push_back(c); continue; }
      if (p_ >= end_) break;
      const char e = *p_++;
      switch (e) {
        case '"': out->push_back('"'); break;
        case '\\': out->push_back('\\'); break;
        case '/': out->push_back('/'); break;
        case 'b': out->push_back('\b'); break;
        case 'f': out->push_back('\f'); break;
        case 'n': out->push_back('\n'); break;
        case 'r': out->push_back('\r'); break;
        case 't': out->push_back('\t'); break;
        case 'u': {
          unsigned cp;
          if (!hex4(&cp)) return false;
          if (cp >= 0xD800 && cp < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
            const char* save = p_;
            p_ += 2;
            unsigned lo;
            if (!hex4(&lo)) return false;
            if (lo >= 0xDC00 && lo < 0xE000) cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            else { p_ = save; cp = 0xFFFD; }
          } else if (cp >= 0xD800 && cp < 0xE000) {
            cp = 0xFFFD;
          }
          utf8(out, cp);
          break;
        }
        default: return fail("bad escape");
      }
    }
    return fail("unterminated string");
  }
  bool array(GVal* out, int depth) {
    ++p_;
    *out = GVal::list();
    skip();
    if (p_ < end_ && *p_ == ']') { ++p_; return true; }
    while (true) {
      GVal v;
      skip();
      if (!value(&v, depth + 1)) return false;
      out->items.push_back(std::move(v));
      skip();
      if (p_ >= end_) return fail("unterminated array");
      if (*p_ == ',') { ++p_; continue; }
      if (*p_ == ']') { ++p_; return true; }
      return fail("expected , or ]");
    }
  }
  bool object(GVal* out, int depth) {
    ++p_;
    *out = GVal::map();
    skip();
    if (p_ < end_ && *p_ == '}') { ++p_; return true; }
    while (true) {
      skip();
      if (p_ >= end_ || *p_ != '"') return fail("expected object key");
      std::string k;
      if (!string(&k)) return false;
      skip();
      if (p_ >= end_ || *p_ != ':') return fail("expected :");
      ++p_;
      skip();
      GVal v;
      if (!value(&v, depth + 1)) return false;
      out->set(k, std::move(v));  // duplicate key: last wins, as in Go
      skip();
      if (p_ >= end_) return fail("unterminated object");
      if (*p_ == ',') { ++p_; continue; }
      if (*p_ == '}') { ++p_; return true; }
      return fail("expected , or }");
    }
  }
};

inline bool parse_json(const std::string& text, GVal* out, std::string* err) {
  JsonParser p(text.data(), text.size());
  return p.parse(out, err);
}

}  // namespace nbh
