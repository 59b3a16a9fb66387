// A small JSON reader for the fit configuration (the reference vendors jsoncpp,
// contrib/jsoncpp-src-0.6.0-rc2; nothing of it is used here).  It accepts what jsoncpp accepts
// in the shipped configs: // and /* */ comments, and a missing comma between members
// (config/example.json:42-43).  Object members iterate in sorted key order, as jsoncpp's do --
// src/config.cpp:93-151 depends on that order.
#ifndef SXMC_B200_JSON_MIN_H
#define SXMC_B200_JSON_MIN_H

#include <cctype>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace jsonmin {

struct Value {
  enum Type { NUL, BOOL, NUMBER, STRING, ARRAY, OBJECT };
  Type type;
  bool b;
  double num;
  std::string str;
  std::vector<Value> arr;
  std::map<std::string, Value> obj;

  Value() : type(NUL), b(false), num(0) {}
  bool isMember(const std::string& k) const { return type == OBJECT && obj.count(k) > 0; }
  const Value& operator[](const std::string& k) const {
    static const Value none;
    if (type != OBJECT) return none;
    std::map<std::string, Value>::const_iterator it = obj.find(k);
    return it == obj.end() ? none : it->second;
  }
  const Value& operator[](size_t i) const {
    static const Value none;
    return (type == ARRAY && i < arr.size()) ? arr[i] : none;
  }
  size_t size() const { return type == ARRAY ? arr.size() : (type == OBJECT ? obj.size() : 0); }
  bool isNull() const { return type == NUL; }
  double asDouble(double dflt = 0) const { return type == NUMBER ? num : (type == BOOL ? (b ? 1 : 0) : dflt); }
  float asFloat(float dflt = 0) const { return (float)asDouble(dflt); }
  long asInt(long dflt = 0) const { return type == NUMBER ? (long)num : dflt; }
  bool asBool(bool dflt = false) const { return type == BOOL ? b : (type == NUMBER ? num != 0 : dflt); }
  std::string asString(const std::string& dflt = "") const { return type == STRING ? str : dflt; }
};

class Parser {
 public:
  explicit Parser(const std::string& text) : s_(text), i_(0) {}
  Value parse() {
    Value v = value();
    skip();
    if (i_ != s_.size()) fail("trailing characters");
    return v;
  }

 private:
  const std::string& s_;
  size_t i_;

  void fail(const std::string& what) const {
    size_t line = 1;
    for (size_t k = 0; k < i_ && k < s_.size(); ++k) line += s_[k] == '\n';
    throw std::runtime_error("JSON parse error at line " + std::to_string(line) + ": " + what);
  }
  void skip() {          // whitespace and comments
    for (;;) {
      while (i_ < s_.size() && std::isspace((unsigned char)s_[i_])) ++i_;
      if (i_ + 1 < s_.size() && s_[i_] == '/' && s_[i_ + 1] == '/') {
        while (i_ < s_.size() && s_[i_] != '\n') ++i_;
      } else if (i_ + 1 < s_.size() && s_[i_] == '/' && s_[i_ + 1] == '*') {
        i_ += 2;
        while (i_ + 1 < s_.size() && !(s_[i_] == '*' && s_[i_ + 1] == '/')) ++i_;
        i_ = i_ + 2 <= s_.size() ? i_ + 2 : s_.size();
      } else {
        return;
      }
    }
  }
  Value value() {
    skip();
    if (i_ >= s_.size()) fail("unexpected end");
    const char c = s_[i_];
    Value v;
    if (c == '{') {
      v.type = Value::OBJECT;
      ++i_;
      skip();
      if (i_ < s_.size() && s_[i_] == '}') { ++i_; return v; }
      for (;;) {
        skip();
        if (i_ >= s_.size() || s_[i_] != '"') fail("Missing '}' or object member name");
        const std::string key = string();
        skip();
        if (i_ >= s_.size() || s_[i_] != ':') fail("Missing ':' after object member name");
        ++i_;
        v.obj[key] = value();
        skip();
        // jsoncpp 0.6 (the reference's reader) insists on the separator: config/example.json as
        // shipped, with its missing comma, does not load in the reference either
        if (i_ < s_.size() && s_[i_] == ',') { ++i_; continue; }
        if (i_ < s_.size() && s_[i_] == '}') { ++i_; break; }
        fail("Missing ',' or '}' in object declaration");
      }
    } else if (c == '[') {
      v.type = Value::ARRAY;
      ++i_;
      skip();
      if (i_ < s_.size() && s_[i_] == ']') { ++i_; return v; }
      for (;;) {
        v.arr.push_back(value());
        skip();
        if (i_ < s_.size() && s_[i_] == ',') { ++i_; continue; }
        if (i_ < s_.size() && s_[i_] == ']') { ++i_; break; }
        fail("Missing ',' or ']' in array declaration");
      }
    } else if (c == '"') {
      v.type = Value::STRING;
      v.str = string();
    } else if (s_.compare(i_, 4, "true") == 0) {
      v.type = Value::BOOL; v.b = true; i_ += 4;
    } else if (s_.compare(i_, 5, "false") == 0) {
      v.type = Value::BOOL; v.b = false; i_ += 5;
    } else if (s_.compare(i_, 4, "null") == 0) {
      i_ += 4;
    } else {
      char* end = NULL;
      v.num = std::strtod(s_.c_str() + i_, &end);
      if (end == s_.c_str() + i_) fail(std::string("unexpected character '") + c + "'");
      v.type = Value::NUMBER;
      i_ = (size_t)(end - s_.c_str());
    }
    return v;
  }
  std::string string() {
    std::string out;
    ++i_;                       // opening quote
    while (i_ < s_.size() && s_[i_] != '"') {
      char c = s_[i_++];
      if (c == '\\' && i_ < s_.size()) {
        const char e = s_[i_++];
        switch (e) {
          case 'n': c = '\n'; break;
          case 't': c = '\t'; break;
          case 'r': c = '\r'; break;
          case 'b': c = '\b'; break;
          case 'f': c = '\f'; break;
          case 'u': {             // \uXXXX: ASCII range only (ROOT-LaTeX titles use '#', not unicode)
            const unsigned code = (unsigned)std::strtoul(s_.substr(i_, 4).c_str(), NULL, 16);
            i_ += 4;
            c = code < 128 ? (char)code : '?';
            break;
          }
          default: c = e;
        }
      }
      out.push_back(c);
    }
    if (i_ >= s_.size()) fail("unterminated string");
    ++i_;                       // closing quote
    return out;
  }
};

inline Value parse(const std::string& text) { return Parser(text).parse(); }

}  // namespace jsonmin

#endif
