// csds_timer.h -- scope timers and externally measured timings in the reference's text format
// (reference: inc/csds_timer.h:10-14, src/csds_timer.cpp:30-58).
//
//   start_timer(name, ctx)         RAII; on stderr:  PFTimerStart("name", "ctx")
//                                                    PROFILING_DATUM("name", "ctx", <ms>, "C++")
//                                                    PFTimerEnd("name", "ctx")
//   report_timing(name, ctx, ns)   device-side time (cudaEvent pair inside libshb) on stdout:
//                                                    PROFILING_DATUM("name", "ctx", <ms>, "C++")
// scripts/experiments/analyse.sh greps these lines, so the spelling is kept byte for byte.
#pragma once

#include <chrono>
#include <iostream>
#include <string>

#include "common.h"

class CSDSTimer {
public:
  using clock_type = std::chrono::system_clock;

  CSDSTimer(const std::string &name, const std::string &context, std::ostream &stream)
      : _name(name), _context(context), _out(&stream), _start(clock_type::now()) {
    *_out << "PFTimerStart(\"" << _name << "\", \"" << _context << "\")" << ENDL;
  }
  CSDSTimer(const CSDSTimer &) = delete;
  CSDSTimer &operator=(const CSDSTimer &) = delete;

  ~CSDSTimer() {
    const auto ns = std::chrono::duration_cast<std::chrono::nanoseconds>(clock_type::now() - _start);
    datum(*_out, _name, _context, ns);
    *_out << "PFTimerEnd(\"" << _name << "\", \"" << _context << "\")" << ENDL;
  }

  static void reportTiming(const std::string &name, const std::string &context, std::chrono::nanoseconds ns) {
    datum(std::cout, name, context, ns);
  }

private:
  static void datum(std::ostream &os, const std::string &name, const std::string &context,
                    std::chrono::nanoseconds ns) {
    os << "PROFILING_DATUM(\"" << name << "\", \"" << context << "\", " << ((double)ns.count()) / 1000000.0
       << ", \"C++\")" << ENDL;
  }
  std::string _name, _context;
  std::ostream *_out;
  std::chrono::time_point<clock_type> _start;
};

#define SHB_TIMER_CAT2(a, b) a##b
#define SHB_TIMER_CAT(a, b) SHB_TIMER_CAT2(a, b)
#define start_timer(name, context) CSDSTimer SHB_TIMER_CAT(_csds_timer_, __LINE__)(#name, #context, std::cerr)
#define report_timing(name, context, time) CSDSTimer::reportTiming(#name, #context, std::chrono::nanoseconds(time))
