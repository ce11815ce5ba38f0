// state_reader.h -- typed access to a parsed model state with the reference's VariantMap
// conversion rules (host only).
//
//  * variantMapFromJson (src/Utilities.cpp:2747-2775) types a JSON array by its FIRST
//    element: [0, 1.5] becomes vector<int> {0, 1} -- the truncation is reproduced here
//    (SURVEY.md §9 item 17);
//  * VariantMap::get<int>/get<double> convert between int and double leniently
//    (src/VariantMap.cpp:103-137), get<vector<double>> accepts vector<int> and vice versa;
//  * a missing key throws (VariantMapKeyError in the reference).
#ifndef USGSCSM_B200_STATE_READER_H_
#define USGSCSM_B200_STATE_READER_H_

#include <algorithm>
#include <cctype>
#include <stdexcept>
#include <string>
#include <vector>

#include "json_min.h"
#include "msgpack_min.h"

namespace ub200 {

// stateAsJson (src/Utilities.cpp:2489-2500): non-printable bytes -> '\n', then the text
// between the first '{' and the last '}'.
inline std::string stateJsonText(const char *state, size_t len) {
  std::string s(state, len);
  for (char &c : s) {
    if (!::isprint((unsigned char)c)) c = '\n';
  }
  size_t first = s.find_first_of('{');
  size_t last = s.find_last_of('}');
  if (first == std::string::npos) first = 0;
  if (last == std::string::npos) throw std::runtime_error("model state holds no JSON object");
  return s.substr(first, last + 1 - first);
}

// A model state in either encoding -> value tree: msgpack when the first byte is a map marker
// (isMsgpack, src/Utilities.cpp:2511-2530), JSON text (possibly wrapped in a .sup) otherwise.
inline json::Value parseState(const char *state, size_t len) {
  if (msgpack::looksLikeMap(state, len)) return msgpack::parse(state, len);
  return json::parse(stateJsonText(state, len));
}

class StateReader {
 public:
  explicit StateReader(const json::Value &root) : root_(root) {
    if (root.type != json::Type::Object)
      throw std::runtime_error("model state is not a JSON object");
  }

  bool contains(const std::string &key) const { return root_.members.count(key) != 0; }

  std::string getString(const std::string &key) const {
    const json::Value &v = at(key);
    if (v.type != json::Type::String) typeError(key, "string");
    return v.str;
  }
  std::string getString(const std::string &key, const std::string &def) const {
    if (!contains(key)) return def;
    const json::Value &v = at(key);
    return v.type == json::Type::String ? v.str : def;
  }
  bool getBool(const std::string &key) const {
    const json::Value &v = at(key);
    if (v.type != json::Type::Bool) typeError(key, "bool");
    return v.b;
  }
  int getInt(const std::string &key) const {
    const json::Value &v = at(key);
    if (!v.isNumber()) typeError(key, "numeric");
    return (int)v.num;  // int stays int; double is truncated (static_cast<int>)
  }
  double getDouble(const std::string &key) const {
    const json::Value &v = at(key);
    if (!v.isNumber()) typeError(key, "numeric");
    return v.type == json::Type::Int ? (double)(int)v.num : v.num;
  }
  std::vector<double> getVecDouble(const std::string &key) const {
    const json::Value &v = at(key);
    if (v.type != json::Type::Array) typeError(key, "numeric vector");
    if (v.arraySize() == 0) return {};
    if (!v.numericArray) typeError(key, "numeric vector");
    std::vector<double> out(v.nums);
    if (v.firstIsInt) {
      // stored as vector<int> by the reference: every element truncated
      for (double &d : out) d = (double)(int)d;
    }
    return out;
  }
  std::vector<double> getVecDouble(const std::string &key, const std::vector<double> &def) const {
    if (!contains(key)) return def;
    return getVecDouble(key);
  }

 private:
  const json::Value &root_;

  const json::Value &at(const std::string &key) const {
    auto it = root_.members.find(key);
    if (it == root_.members.end())
      throw std::runtime_error("Key '" + key + "' not found in model state");
    return it->second;
  }
  [[noreturn]] static void typeError(const std::string &key, const char *want) {
    throw std::runtime_error("Bad type for key '" + key + "': expected " + want);
  }
};

}  // namespace ub200
#endif
