// options.h -- command-line options with the reference's behaviour (inc/options.h:76-256):
//   * "-x value" / "--long value"; bool options are switches that take no value
//   * an unknown flag, or a flag with a missing value, prints "Error: ..." plus the usage text and exit(0)
//   * require() on an option that was not given prints `Required argument "<long>" not set.` and exit(-1)
//   * -h / --help prints the usage text and exit(0)
#pragma once

#include <algorithm>
#include <cstdlib>
#include <iomanip>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "common.h"

class OptionBase {
public:
  OptionBase(char s, const std::string &l, const std::string &d, bool has_default)
      : _shortopt(s), _longopt(l), _desc(d), _has_default(has_default) {}
  virtual ~OptionBase() {}

  char getShort() const { return _shortopt; }
  const std::string &getLong() const { return _longopt; }
  const std::string &getDesc() const { return _desc; }
  bool has_default() const { return _has_default; }
  bool provided() const { return _value_provided; }

  virtual void print(std::ostream &out) const = 0;
  // consumes this option's value (if it takes one); `current` indexes the flag itself on entry
  virtual bool parseArgs(int argc, int &current, char **argv) = 0;

protected:
  char _shortopt;
  std::string _longopt, _desc;
  bool _has_default;
  bool _value_provided = false;
};

template <typename T> class Option : public OptionBase {
public:
  Option(char s, const std::string &l, const std::string &d) : OptionBase(s, l, d, false), _value() {}
  Option(char s, const std::string &l, const std::string &d, const T &v) : OptionBase(s, l, d, true), _value(v) {}

  T get() const { return _value; }

  T require() const {
    if (!_value_provided) {
      std::cout << "Required argument \"" << _longopt << "\" not set." << ENDL;
      std::exit(-1);
    }
    return _value;
  }

  operator T() const { return _value; }

  void setValue(const T &v) {
    _value = v;
    _value_provided = true;
  }

  bool parseArgs(int argc, int &current, char **argv) override {
    ++current;
    if (current >= argc) return false;
    std::istringstream ss(argv[current]);
    ss >> _value;
    _value_provided = true;
    return true;
  }

  void print(std::ostream &out) const override { out << _longopt << ": " << _value; }

private:
  T _value;
};

template <> inline bool Option<bool>::parseArgs(int, int &, char **) {
  _value = !_value;
  _value_provided = true;
  return true;
}

class OptParser {
public:
  explicit OptParser(const std::string &description) : _desc(description) {
    _help = addOption<bool>({'h', "help", "Print help and exit.", false});
  }

  template <typename T> std::shared_ptr<Option<T>> addOption(Option<T> &&opt) {
    auto p = std::make_shared<Option<T>>(opt);
    _args.push_back(p);
    return p;
  }

  void parse(int argc, char **argv) {
    int c = 1;
    while (c < argc) {
      OptionBase *hit = nullptr;
      const std::string tok = argv[c];
      if (tok.size() >= 2 && tok[0] == '-') {
        for (auto &o : _args) {
          const bool match = tok[1] == '-' ? (o->getLong() == tok.substr(2)) : (o->getShort() && o->getShort() == tok[1]);
          if (match) {
            hit = o.get();
            break;
          }
        }
      }
      if (!hit) {
        std::cout << "Error: Invalid argument '" << argv[c] << "'." << ENDL;
        _help->setValue(true);
        break;
      }
      if (!hit->parseArgs(argc, c, argv)) {
        std::cout << "Error: invalid argument for option " << hit->getLong() << ENDL;
        _help->setValue(true);
        break;
      }
      ++c;
    }
    if (_help->get()) {
      usage(argv[0]);
      std::exit(0);
    }
  }

  void usage(const char *prog) const {
    std::size_t width = 0;
    for (auto &o : _args) width = std::max(width, o->getLong().length());
    std::cout << "Usage:\n  " << prog << " [OPTIONS]...\nDescription:\n  " << _desc << "\nOptions:\n";
    for (auto &o : _args) {
      if (o->getShort()) std::cout << "  -" << o->getShort();
      else std::cout << "    ";
      std::cout << "  --" << std::left << std::setw((int)width) << o->getLong() << "  " << o->getDesc() << std::endl;
    }
  }

  void print(std::ostream &out = std::cout) const {
    for (auto &o : _args) {
      o->print(out);
      out << ENDL << ENDL;
    }
  }

private:
  std::string _desc;
  std::vector<std::shared_ptr<OptionBase>> _args;
  std::shared_ptr<Option<bool>> _help;
};
