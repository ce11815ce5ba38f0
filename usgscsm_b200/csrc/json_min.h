// json_min.h -- a small JSON reader for model-state strings (host only).
//
// The reference reads states with nlohmann::json and then types every value by the rules
// of variantMapFromJson (src/Utilities.cpp:2747-2775).  This reader keeps exactly the
// information those rules need: for each number whether its literal was an integer, and
// for arrays of numbers the flat double values plus the "integer-ness" of each.
// Numbers are converted with strtod (correctly rounded, like nlohmann's parser).
#ifndef USGSCSM_B200_JSON_MIN_H_
#define USGSCSM_B200_JSON_MIN_H_

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ub200 {
namespace json {

enum class Type { Null, Bool, Int, Float, String, Array, Object };

struct Value {
  Type type = Type::Null;
  bool b = false;
  double num = 0.0;  // Int and Float (Int values are exactly representable here)
  std::string str;
  // Arrays: all-number arrays are stored flat (nums + firstIsInt), other arrays in items.
  bool numericArray = false;
  bool firstIsInt = false;
  std::vector<double> nums;
  std::vector<unsigned char> numIsInt;  // per element of nums: the literal was an integer
  std::vector<Value> items;
  std::map<std::string, Value> members;

  bool isNumber() const { return type == Type::Int || type == Type::Float; }
  size_t arraySize() const { return numericArray ? nums.size() : items.size(); }
};

class Parser {
 public:
  Parser(const char *begin, const char *end) : p_(begin), end_(end) {}

  Value parseDocument() {
    Value v = parseValue(0);
    skipWs();
    if (p_ != end_) fail("trailing characters after JSON document");
    return v;
  }

 private:
  const char *p_;
  const char *end_;

  [[noreturn]] void fail(const std::string &what) const {
    throw std::runtime_error("JSON parse error: " + what);
  }
  void skipWs() {
    while (p_ != end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) ++p_;
  }
  bool consume(const char *lit) {
    size_t n = std::strlen(lit);
    if ((size_t)(end_ - p_) >= n && std::memcmp(p_, lit, n) == 0) {
      p_ += n;
      return true;
    }
    return false;
  }

  Value parseValue(int depth) {
    if (depth > 64) fail("nesting too deep");
    skipWs();
    if (p_ == end_) fail("unexpected end of input");
    char c = *p_;
    if (c == '{') return parseObject(depth);
    if (c == '[') return parseArray(depth);
    if (c == '"') {
      Value v;
      v.type = Type::String;
      v.str = parseString();
      return v;
    }
    if (c == 't' || c == 'f') {
      Value v;
      v.type = Type::Bool;
      if (consume("true"))
        v.b = true;
      else if (consume("false"))
        v.b = false;
      else
        fail("bad literal");
      return v;
    }
    if (c == 'n') {
      if (!consume("null")) fail("bad literal");
      return Value();
    }
    // nlohmann does not accept NaN/Infinity literals; neither do we
    return parseNumber();
  }

  Value parseNumber() {
    const char *s = p_;
    bool isInt = true;
    if (p_ != end_ && *p_ == '-') ++p_;
    if (p_ == end_ || !(*p_ >= '0' && *p_ <= '9')) fail("bad number");
    while (p_ != end_ && *p_ >= '0' && *p_ <= '9') ++p_;
    if (p_ != end_ && *p_ == '.') {
      isInt = false;
      ++p_;
      if (p_ == end_ || !(*p_ >= '0' && *p_ <= '9')) fail("bad number");
      while (p_ != end_ && *p_ >= '0' && *p_ <= '9') ++p_;
    }
    if (p_ != end_ && (*p_ == 'e' || *p_ == 'E')) {
      isInt = false;
      ++p_;
      if (p_ != end_ && (*p_ == '+' || *p_ == '-')) ++p_;
      if (p_ == end_ || !(*p_ >= '0' && *p_ <= '9')) fail("bad number");
      while (p_ != end_ && *p_ >= '0' && *p_ <= '9') ++p_;
    }
    std::string tok(s, p_);
    Value v;
    v.num = std::strtod(tok.c_str(), nullptr);
    // nlohmann keeps an integer literal as an integer only while it fits 64 bits
    if (isInt && std::fabs(v.num) >= 9.2e18) isInt = false;
    v.type = isInt ? Type::Int : Type::Float;
    return v;
  }

  static void appendUtf8(std::string &out, unsigned cp) {
    if (cp < 0x80) {
      out += (char)cp;
    } else if (cp < 0x800) {
      out += (char)(0xC0 | (cp >> 6));
      out += (char)(0x80 | (cp & 0x3F));
    } else if (cp < 0x10000) {
      out += (char)(0xE0 | (cp >> 12));
      out += (char)(0x80 | ((cp >> 6) & 0x3F));
      out += (char)(0x80 | (cp & 0x3F));
    } else {
      out += (char)(0xF0 | (cp >> 18));
      out += (char)(0x80 | ((cp >> 12) & 0x3F));
      out += (char)(0x80 | ((cp >> 6) & 0x3F));
      out += (char)(0x80 | (cp & 0x3F));
    }
  }
  unsigned parseHex4() {
    if (end_ - p_ < 4) fail("bad \\u escape");
    unsigned v = 0;
    for (int i = 0; i < 4; i++) {
      char c = *p_++;
      v <<= 4;
      if (c >= '0' && c <= '9')
        v |= (unsigned)(c - '0');
      else if (c >= 'a' && c <= 'f')
        v |= (unsigned)(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F')
        v |= (unsigned)(c - 'A' + 10);
      else
        fail("bad \\u escape");
    }
    return v;
  }
  std::string parseString() {
    ++p_;  // opening quote
    std::string out;
    for (;;) {
      if (p_ == end_) fail("unterminated string");
      char c = *p_++;
      if (c == '"') break;
      if (c != '\\') {
        out += c;
        continue;
      }
      if (p_ == end_) fail("unterminated escape");
      char e = *p_++;
      switch (e) {
        case '"': out += '"'; break;
        case '\\': out += '\\'; break;
        case '/': out += '/'; break;
        case 'b': out += '\b'; break;
        case 'f': out += '\f'; break;
        case 'n': out += '\n'; break;
        case 'r': out += '\r'; break;
        case 't': out += '\t'; break;
        case 'u': {
          unsigned cp = parseHex4();
          if (cp >= 0xD800 && cp < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
            p_ += 2;
            unsigned lo = parseHex4();
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          appendUtf8(out, cp);
        } break;
        default: fail("bad escape");
      }
    }
    return out;
  }

  Value parseArray(int depth) {
    ++p_;  // [
    Value v;
    v.type = Type::Array;
    v.numericArray = true;
    skipWs();
    if (p_ != end_ && *p_ == ']') {
      ++p_;
      return v;
    }
    for (;;) {
      Value e = parseValue(depth + 1);
      if (v.numericArray && e.isNumber()) {
        if (v.nums.empty()) v.firstIsInt = (e.type == Type::Int);
        v.nums.push_back(e.num);
        v.numIsInt.push_back(e.type == Type::Int);
      } else {
        if (v.numericArray) {  // demote: re-box what we have
          v.numericArray = false;
          for (size_t k = 0; k < v.nums.size(); k++) {
            Value n;
            n.type = v.numIsInt[k] ? Type::Int : Type::Float;
            n.num = v.nums[k];
            v.items.push_back(n);
          }
          v.nums.clear();
          v.numIsInt.clear();
        }
        v.items.push_back(std::move(e));
      }
      skipWs();
      if (p_ == end_) fail("unterminated array");
      if (*p_ == ',') {
        ++p_;
        continue;
      }
      if (*p_ == ']') {
        ++p_;
        break;
      }
      fail("expected , or ] in array");
    }
    return v;
  }

  Value parseObject(int depth) {
    ++p_;  // {
    Value v;
    v.type = Type::Object;
    skipWs();
    if (p_ != end_ && *p_ == '}') {
      ++p_;
      return v;
    }
    for (;;) {
      skipWs();
      if (p_ == end_ || *p_ != '"') fail("expected string key");
      std::string key = parseString();
      skipWs();
      if (p_ == end_ || *p_ != ':') fail("expected : after key");
      ++p_;
      v.members[key] = parseValue(depth + 1);  // later duplicates win, as in nlohmann
      skipWs();
      if (p_ == end_) fail("unterminated object");
      if (*p_ == ',') {
        ++p_;
        continue;
      }
      if (*p_ == '}') {
        ++p_;
        break;
      }
      fail("expected , or } in object");
    }
    return v;
  }
};

inline Value parse(const std::string &text) {
  Parser p(text.data(), text.data() + text.size());
  return p.parseDocument();
}

// Compact JSON text of a value tree: keys in sorted order, doubles with 17 significant digits
// (they parse back to the same bits), non-finite numbers as null like nlohmann::json::dump.
inline void dumpNumber(std::string &out, double v, bool isInt) {
  char buf[40];
  if (!(v == v) || v - v != 0.0) {
    out += "null";
    return;
  }
  if (isInt && v > -9.2e18 && v < 9.2e18) {
    std::snprintf(buf, sizeof buf, "%.0f", v);
    out += buf;
    return;
  }
  std::snprintf(buf, sizeof buf, "%.17g", v);
  out += buf;
  if (!std::strpbrk(buf, ".eEn")) out += ".0";
}

inline void dumpString(std::string &out, const std::string &s) {
  out += '"';
  for (unsigned char c : s) {
    switch (c) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\n': out += "\\n"; break;
      case '\r': out += "\\r"; break;
      case '\t': out += "\\t"; break;
      default:
        if (c < 0x20) {
          char buf[8];
          std::snprintf(buf, sizeof buf, "\\u%04x", c);
          out += buf;
        } else {
          out += (char)c;
        }
    }
  }
  out += '"';
}

inline void dumpValue(std::string &out, const Value &v) {
  switch (v.type) {
    case Type::Null: out += "null"; break;
    case Type::Bool: out += v.b ? "true" : "false"; break;
    case Type::Int: dumpNumber(out, v.num, true); break;
    case Type::Float: dumpNumber(out, v.num, false); break;
    case Type::String: dumpString(out, v.str); break;
    case Type::Array: {
      out += '[';
      const size_t n = v.arraySize();
      for (size_t k = 0; k < n; k++) {
        if (k) out += ',';
        if (v.numericArray)
          dumpNumber(out, v.nums[k], k < v.numIsInt.size() && v.numIsInt[k]);
        else
          dumpValue(out, v.items[k]);
      }
      out += ']';
    } break;
    case Type::Object: {
      out += '{';
      bool first = true;
      for (const auto &kv : v.members) {
        if (!first) out += ',';
        first = false;
        dumpString(out, kv.first);
        out += ':';
        dumpValue(out, kv.second);
      }
      out += '}';
    } break;
  }
}

inline std::string dump(const Value &v) {
  std::string out;
  dumpValue(out, v);
  return out;
}

inline Value numberArray(const std::vector<double> &nums) {
  Value v;
  v.type = Type::Array;
  v.numericArray = true;
  v.nums = nums;
  v.numIsInt.assign(nums.size(), 0);
  return v;
}

inline Value number(double d) {
  Value v;
  v.type = Type::Float;
  v.num = d;
  return v;
}

}  // namespace json
}  // namespace ub200
#endif
