// logger.h -- LOG_ERROR / LOG_WARNING / LOG_INFO / LOG_DEBUG / LOG_DEBUG_INFO to std::clog with the
// reference's line header (src/util/Logger.cpp:226-251, inc/util/Logger.h:63-91):
//   [<file, right-aligned, width 16, fill '='>:<line, left, width 4> <SSS.mmm>s <severity, width 5>] message
// The reference builds with -DDEBUG and never NDEBUG, so its debug lines are live; here they are
// live unless SHB_QUIET_LOG is set in the environment (or NDEBUG is defined), because a debug dump
// per buffer is unusable on 2^24-row operators.
#pragma once

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>

namespace shb_log {
enum Severity { Debug, DebugInfo, Info, Warning, Error };

inline const char *severityName(Severity s) {
  switch (s) {
    case Debug: return "DEBUG";
    case DebugInfo: return "DINFO";
    case Info: return "INFO";
    case Warning: return "WARN";
    case Error: return "ERROR";
  }
  return "?";
}

inline std::chrono::steady_clock::time_point startTime() {
  static const std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  return t0;
}

inline bool debugEnabled() {
#ifdef NDEBUG
  return false;
#else
  static const bool on = std::getenv("SHB_QUIET_LOG") == nullptr;
  return on;
#endif
}

inline std::string header(Severity sev, const char *file, int line) {
  const char *base = std::strrchr(file, '/');
  base = base ? base + 1 : file;
  const auto ms =
      std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - startTime()).count();
  std::ostringstream s;
  s << "[" << std::setw(16) << std::setfill('=') << base << ":" << std::setw(4) << std::setfill(' ') << std::left << line
    << " ";
  s << std::right << std::setfill('0') << std::setw(3) << (ms / 1000) % 1000 << "." << std::setw(3) << ms % 1000 << "s";
  s << std::setfill(' ') << " " << std::setw(5) << std::left << severityName(sev) << "] ";
  return s.str();
}

inline void append(std::ostringstream &) {}
template <typename T, typename... Rest> void append(std::ostringstream &os, const T &v, const Rest &...rest) {
  os << v;
  append(os, rest...);
}

template <typename... Args> void log(Severity sev, const char *file, int line, const Args &...args) {
  if ((sev == Debug || sev == DebugInfo) && !debugEnabled()) return;
  std::ostringstream os;
  os << header(sev, file, line);
  append(os, args...);
  os << "\n";
  std::clog << os.str();
}
}  // namespace shb_log

#define LOG_ERROR(...) shb_log::log(shb_log::Error, __FILE__, __LINE__, __VA_ARGS__)
#define LOG_WARNING(...) shb_log::log(shb_log::Warning, __FILE__, __LINE__, __VA_ARGS__)
#define LOG_INFO(...) shb_log::log(shb_log::Info, __FILE__, __LINE__, __VA_ARGS__)
#define LOG_DEBUG(...) shb_log::log(shb_log::Debug, __FILE__, __LINE__, __VA_ARGS__)
#define LOG_DEBUG_INFO(...) shb_log::log(shb_log::DebugInfo, __FILE__, __LINE__, __VA_ARGS__)
