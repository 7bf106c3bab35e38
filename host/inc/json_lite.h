// json_lite.h -- a small recursive-descent JSON reader for the kernel configuration files.
// The reference parses them with boost::property_tree (src/kernel_config.cpp:10-12); Boost is not a
// dependency of this build.  Objects keep their member order; every scalar is kept as text so that
// get<std::string>() on a number or boolean behaves like property_tree's string view of it.
#pragma once

#include <cstdint>
#include <fstream>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace json_lite {

struct Value {
  enum Kind { Null, Scalar, Array, Object } kind = Null;
  std::string text;                                      // Scalar
  std::vector<std::pair<std::string, Value>> members;   // Object (key, value) / Array ("", value)

  const Value *find(const std::string &key) const {
    for (auto &m : members)
      if (m.first == key) return &m.second;
    return nullptr;
  }
  const Value &at(const std::string &key) const {
    const Value *v = find(key);
    if (!v) throw std::runtime_error("No such node (" + key + ")");
    return *v;
  }
  std::string str(const std::string &key) const { return at(key).text; }
  bool has(const std::string &key) const { return find(key) != nullptr; }
};

class Parser {
public:
  explicit Parser(const std::string &s) : _s(s) {}

  Value parse() {
    Value v = value();
    ws();
    if (_p != _s.size()) fail("trailing characters");
    return v;
  }

private:
  [[noreturn]] void fail(const std::string &what) const {
    std::size_t line = 1;
    for (std::size_t i = 0; i < _p && i < _s.size(); ++i) line += _s[i] == '\n';
    throw std::runtime_error("JSON parse error at line " + std::to_string(line) + ": " + what);
  }
  void ws() {
    while (_p < _s.size() && (_s[_p] == ' ' || _s[_p] == '\t' || _s[_p] == '\n' || _s[_p] == '\r')) ++_p;
  }
  char peek() {
    ws();
    if (_p >= _s.size()) fail("unexpected end of input");
    return _s[_p];
  }
  void expect(char c) {
    if (peek() != c) fail(std::string("expected '") + c + "'");
    ++_p;
  }
  static void utf8(std::string &out, uint32_t cp) {
    if (cp < 0x80) out += (char)cp;
    else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
  }
  uint32_t hex4() {
    if (_p + 4 > _s.size()) fail("short \\u escape");
    uint32_t v = 0;
    for (int i = 0; i < 4; ++i) {
      const char c = _s[_p++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= (uint32_t)(c - '0');
      else if (c >= 'a' && c <= 'f') v |= (uint32_t)(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F') v |= (uint32_t)(c - 'A' + 10);
      else fail("bad \\u escape");
    }
    return v;
  }
  std::string string() {
    expect('"');
    std::string out;
    while (true) {
      if (_p >= _s.size()) fail("unterminated string");
      const char c = _s[_p++];
      if (c == '"') break;
      if (c != '\\') {
        out += c;
        continue;
      }
      if (_p >= _s.size()) fail("unterminated escape");
      const char e = _s[_p++];
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
          uint32_t cp = hex4();
          if (cp >= 0xD800 && cp <= 0xDBFF && _p + 1 < _s.size() && _s[_p] == '\\' && _s[_p + 1] == 'u') {
            _p += 2;
            const uint32_t lo = hex4();
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          utf8(out, cp);
          break;
        }
        default: fail("unknown escape");
      }
    }
    return out;
  }
  Value value() {
    const char c = peek();
    Value v;
    if (c == '{') {
      ++_p;
      v.kind = Value::Object;
      if (peek() == '}') { ++_p; return v; }
      while (true) {
        std::string key = string();
        expect(':');
        v.members.emplace_back(std::move(key), value());
        const char d = peek();
        ++_p;
        if (d == '}') break;
        if (d != ',') fail("expected ',' or '}'");
      }
    } else if (c == '[') {
      ++_p;
      v.kind = Value::Array;
      if (peek() == ']') { ++_p; return v; }
      while (true) {
        v.members.emplace_back(std::string(), value());
        const char d = peek();
        ++_p;
        if (d == ']') break;
        if (d != ',') fail("expected ',' or ']'");
      }
    } else if (c == '"') {
      v.kind = Value::Scalar;
      v.text = string();
    } else {
      // number / true / false / null, kept as text
      const std::size_t b = _p;
      while (_p < _s.size() && _s[_p] != ',' && _s[_p] != '}' && _s[_p] != ']' && _s[_p] != ' ' && _s[_p] != '\n' &&
             _s[_p] != '\t' && _s[_p] != '\r')
        ++_p;
      if (_p == b) fail("unexpected character");
      v.text = _s.substr(b, _p - b);
      v.kind = v.text == "null" ? Value::Null : Value::Scalar;
    }
    return v;
  }

  const std::string &_s;
  std::size_t _p = 0;
};

inline Value read_file(const std::string &filename) {
  std::ifstream in(filename, std::ios::binary);
  if (!in) throw std::runtime_error(filename + ": cannot open file");
  std::stringstream ss;
  ss << in.rdbuf();
  const std::string text = ss.str();
  return Parser(text).parse();
}

}  // namespace json_lite
