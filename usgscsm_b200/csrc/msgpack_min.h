// msgpack_min.h -- reader for model states in msgpack binary form (host only).
//
// The reference writes a state as nlohmann::json::to_msgpack(state) (bin/usgscsm_cam_test.cc
// --output-binary-model-state; README.md:60-62 "much faster to load than JSON") and reads it
// back with json::from_msgpack (bin/usgscsm_cam_test.cc:157-170); a file is taken for msgpack
// when its first byte is a map marker (isMsgpack, src/Utilities.cpp:2511-2530).  This decoder
// produces the same json::Value tree as json_min.h does for the equivalent text, so the state
// loaders are unaware of the encoding.  Format: github.com/msgpack/msgpack/blob/master/spec.md.
#ifndef USGSCSM_B200_MSGPACK_MIN_H_
#define USGSCSM_B200_MSGPACK_MIN_H_

#include <cstdint>

#include "json_min.h"

namespace ub200 {
namespace msgpack {

// isMsgpack (src/Utilities.cpp:2511-2530) on a buffer.
inline bool looksLikeMap(const char *data, size_t len) {
  if (len == 0) return false;
  unsigned char b = (unsigned char)data[0];
  return (b >= 0x80 && b <= 0x8F) || b == 0xDE || b == 0xDF;
}

class Reader {
 public:
  Reader(const char *begin, const char *end)
      : p_((const unsigned char *)begin), end_((const unsigned char *)end) {}

  json::Value readDocument() {
    json::Value v = readValue(0);
    if (p_ != end_) fail("trailing bytes after the msgpack document");
    return v;
  }

 private:
  const unsigned char *p_;
  const unsigned char *end_;

  [[noreturn]] void fail(const std::string &what) const {
    throw std::runtime_error("msgpack parse error: " + what);
  }
  void need(size_t n) const {
    if ((size_t)(end_ - p_) < n) fail("unexpected end of input");
  }
  uint64_t be(int bytes) {
    need((size_t)bytes);
    uint64_t v = 0;
    for (int i = 0; i < bytes; i++) v = (v << 8) | *p_++;
    return v;
  }
  static json::Value integer(double d) {
    json::Value v;
    v.type = json::Type::Int;
    v.num = d;
    return v;
  }
  static json::Value real(double d) {
    json::Value v;
    v.type = json::Type::Float;
    v.num = d;
    return v;
  }
  json::Value str(size_t n) {
    need(n);
    json::Value v;
    v.type = json::Type::String;
    v.str.assign((const char *)p_, n);
    p_ += n;
    return v;
  }
  json::Value array(size_t n, int depth) {
    json::Value v;
    v.type = json::Type::Array;
    v.numericArray = true;
    for (size_t i = 0; i < n; i++) {
      json::Value e = readValue(depth + 1);
      if (v.numericArray && e.isNumber()) {
        if (v.nums.empty()) v.firstIsInt = (e.type == json::Type::Int);
        v.nums.push_back(e.num);
        v.numIsInt.push_back(e.type == json::Type::Int);
      } else {
        if (v.numericArray) {  // demote, as json_min.h does
          v.numericArray = false;
          for (size_t k = 0; k < v.nums.size(); k++)
            v.items.push_back(v.numIsInt[k] ? integer(v.nums[k]) : real(v.nums[k]));
          v.nums.clear();
          v.numIsInt.clear();
        }
        v.items.push_back(std::move(e));
      }
    }
    return v;
  }
  json::Value map(size_t n, int depth) {
    json::Value v;
    v.type = json::Type::Object;
    for (size_t i = 0; i < n; i++) {
      json::Value k = readValue(depth + 1);
      if (k.type != json::Type::String) fail("map key is not a string");
      v.members[k.str] = readValue(depth + 1);
    }
    return v;
  }

  json::Value readValue(int depth) {
    if (depth > 64) fail("nesting too deep");
    need(1);
    const unsigned char c = *p_++;
    if (c <= 0x7F) return integer((double)c);                       // positive fixint
    if (c >= 0xE0) return integer((double)(int8_t)c);               // negative fixint
    if (c >= 0x80 && c <= 0x8F) return map(c & 0x0F, depth);        // fixmap
    if (c >= 0x90 && c <= 0x9F) return array(c & 0x0F, depth);      // fixarray
    if (c >= 0xA0 && c <= 0xBF) return str(c & 0x1F);               // fixstr
    switch (c) {
      case 0xC0: return json::Value();  // nil
      case 0xC2:
      case 0xC3: {
        json::Value v;
        v.type = json::Type::Bool;
        v.b = (c == 0xC3);
        return v;
      }
      case 0xCA: {  // float32 (nlohmann writes it when the double is exactly a float)
        uint32_t u = (uint32_t)be(4);
        float f;
        std::memcpy(&f, &u, 4);
        return real((double)f);
      }
      case 0xCB: {
        uint64_t u = be(8);
        double d;
        std::memcpy(&d, &u, 8);
        return real(d);
      }
      case 0xCC: return integer((double)be(1));
      case 0xCD: return integer((double)be(2));
      case 0xCE: return integer((double)be(4));
      case 0xCF: return integer((double)be(8));
      case 0xD0: return integer((double)(int8_t)be(1));
      case 0xD1: return integer((double)(int16_t)be(2));
      case 0xD2: return integer((double)(int32_t)be(4));
      case 0xD3: return integer((double)(int64_t)be(8));
      case 0xD9: return str((size_t)be(1));
      case 0xDA: return str((size_t)be(2));
      case 0xDB: return str((size_t)be(4));
      case 0xDC: return array((size_t)be(2), depth);
      case 0xDD: return array((size_t)be(4), depth);
      case 0xDE: return map((size_t)be(2), depth);
      case 0xDF: return map((size_t)be(4), depth);
      default: fail("unsupported type byte (bin / ext are not used by model states)");
    }
  }
};

inline json::Value parse(const char *data, size_t len) { return Reader(data, data + len).readDocument(); }

// Writer: every value in the smallest msgpack form that holds it (the rule nlohmann::json's
// to_msgpack follows, so the bytes are the ones the reference writes): integers by range,
// a double as float32 when that is exact, strings / arrays / maps by length; map keys in
// sorted order (nlohmann's default object type is a std::map too).
class Writer {
 public:
  std::string out;

  void write(const json::Value &v) {
    switch (v.type) {
      case json::Type::Null: put(0xC0); break;
      case json::Type::Bool: put(v.b ? 0xC3 : 0xC2); break;
      case json::Type::Int: integer(v.num); break;
      case json::Type::Float: real(v.num); break;
      case json::Type::String: str(v.str); break;
      case json::Type::Array: {
        const size_t n = v.arraySize();
        if (n <= 15) put(0x90 | (unsigned)n);
        else if (n <= 0xFFFF) { put(0xDC); be(n, 2); }
        else { put(0xDD); be(n, 4); }
        if (v.numericArray) {
          for (size_t k = 0; k < n; k++) {
            if (v.numIsInt[k]) integer(v.nums[k]);
            else real(v.nums[k]);
          }
        } else {
          for (const json::Value &e : v.items) write(e);
        }
      } break;
      case json::Type::Object: {
        const size_t n = v.members.size();
        if (n <= 15) put(0x80 | (unsigned)n);
        else if (n <= 0xFFFF) { put(0xDE); be(n, 2); }
        else { put(0xDF); be(n, 4); }
        for (const auto &kv : v.members) {
          str(kv.first);
          write(kv.second);
        }
      } break;
    }
  }

 private:
  void put(unsigned b) { out += (char)(unsigned char)b; }
  void be(uint64_t v, int bytes) {
    for (int i = bytes - 1; i >= 0; i--) put((unsigned)((v >> (8 * i)) & 0xFF));
  }
  void integer(double d) {
    if (d >= 0) {
      uint64_t u = (uint64_t)d;
      if (u < 128) put((unsigned)u);
      else if (u <= 0xFF) { put(0xCC); be(u, 1); }
      else if (u <= 0xFFFF) { put(0xCD); be(u, 2); }
      else if (u <= 0xFFFFFFFFull) { put(0xCE); be(u, 4); }
      else { put(0xCF); be(u, 8); }
    } else {
      int64_t i = (int64_t)d;
      if (i >= -32) put((unsigned)(uint8_t)(int8_t)i);
      else if (i >= -128) { put(0xD0); be((uint64_t)(uint8_t)(int8_t)i, 1); }
      else if (i >= -32768) { put(0xD1); be((uint64_t)(uint16_t)(int16_t)i, 2); }
      else if (i >= -2147483648LL) { put(0xD2); be((uint64_t)(uint32_t)(int32_t)i, 4); }
      else { put(0xD3); be((uint64_t)i, 8); }
    }
  }
  void real(double d) {
    if (d >= -3.4028234663852886e38 && d <= 3.4028234663852886e38 && (double)(float)d == d) {
      float f = (float)d;
      uint32_t u;
      std::memcpy(&u, &f, 4);
      put(0xCA);
      be(u, 4);
    } else {
      uint64_t u;
      std::memcpy(&u, &d, 8);
      put(0xCB);
      be(u, 8);
    }
  }
  void str(const std::string &s) {
    const size_t n = s.size();
    if (n <= 31) put(0xA0 | (unsigned)n);
    else if (n <= 0xFF) { put(0xD9); be(n, 1); }
    else if (n <= 0xFFFF) { put(0xDA); be(n, 2); }
    else { put(0xDB); be(n, 4); }
    out += s;
  }
};

inline std::string dump(const json::Value &v) {
  Writer w;
  w.write(v);
  return w.out;
}

}  // namespace msgpack
}  // namespace ub200
#endif
