// arithexpr_evaluator.h -- evaluates the buffer-size expressions of the kernel JSONs, e.g.
// "(4*v_MHeight_2*v_MWidthC_1)" (reference: src/arithexpr_evaluator.cpp:10-29, which hands the string
// to the exprtk library).  Every shipped JSON uses only integer literals, '*', parentheses and the
// three size variables; this evaluator also accepts + - / and unary minus.
// Arithmetic is done in double like exprtk; the result is 64-bit here (the reference returns int,
// which caps a buffer at 2 GB).  An unparsable expression ("?", used for the RSA inputs and never
// evaluated by the reference) yields 0.
#pragma once

#include <cctype>
#include <cstdint>
#include <cstdlib>
#include <string>

#include "csds_timer.h"

class Evaluator {
public:
  static int64_t evaluate(const std::string &expr, int64_t v_MWidth_1, int64_t v_MHeight_2, int64_t v_VLength_3) {
    Evaluator e(expr, (double)v_MWidth_1, (double)v_MHeight_2, (double)v_VLength_3);
    const double v = e.sum();
    e.skip();
    if (!e._ok || e._p != e._s.size()) return 0;
    return (int64_t)v;
  }

private:
  Evaluator(const std::string &s, double w, double h, double l) : _s(s), _w(w), _h(h), _l(l) {}

  void skip() {
    while (_p < _s.size() && std::isspace((unsigned char)_s[_p])) ++_p;
  }
  bool eat(char c) {
    skip();
    if (_p < _s.size() && _s[_p] == c) {
      ++_p;
      return true;
    }
    return false;
  }
  double sum() {
    double v = product();
    while (_ok) {
      if (eat('+')) v += product();
      else if (eat('-')) v -= product();
      else break;
    }
    return v;
  }
  double product() {
    double v = atom();
    while (_ok) {
      if (eat('*')) v *= atom();
      else if (eat('/')) v /= atom();
      else break;
    }
    return v;
  }
  double atom() {
    skip();
    if (eat('(')) {
      const double v = sum();
      if (!eat(')')) _ok = false;
      return v;
    }
    if (eat('-')) return -atom();
    if (_p < _s.size() && (std::isdigit((unsigned char)_s[_p]) || _s[_p] == '.')) {
      char *end = nullptr;
      const double v = std::strtod(_s.c_str() + _p, &end);
      _p = (std::size_t)(end - _s.c_str());
      return v;
    }
    if (_p < _s.size() && (std::isalpha((unsigned char)_s[_p]) || _s[_p] == '_')) {
      const std::size_t b = _p;
      while (_p < _s.size() && (std::isalnum((unsigned char)_s[_p]) || _s[_p] == '_')) ++_p;
      const std::string name = _s.substr(b, _p - b);
      if (name == "v_MWidthC_1") return _w;
      if (name == "v_MHeight_2") return _h;
      if (name == "v_VLength_3") return _l;
    }
    _ok = false;
    return 0.0;
  }

  const std::string &_s;
  double _w, _h, _l;
  std::size_t _p = 0;
  bool _ok = true;
};
