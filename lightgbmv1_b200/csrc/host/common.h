// Small host-side utilities shared by the C-API mirror: logging with the reference's error convention
// (Log::Fatal throws, the C API turns it into -1 + LGBM_GetLastError; include/LightGBM/utils/log.h:110-131,
// src/c_api.cpp:35-49), string helpers and the numeric constants of include/LightGBM/meta.h.
#pragma once
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace lgb {

typedef int32_t data_size_t;
typedef float score_t;
typedef float label_t;

// include/LightGBM/meta.h:50-54 (score_t constants: float literals)
const score_t kMinScore = -std::numeric_limits<score_t>::infinity();
const score_t kEpsilon = 1e-15f;
const double kZeroThreshold = 1e-35f;
const double kSparseThreshold = 0.7;  // include/LightGBM/bin.h:39

enum MissingType { kMissingNone = 0, kMissingZero = 1, kMissingNaN = 2 };

struct Log {
  static int& level() { static thread_local int lvl = 1; return lvl; }  // verbosity
  typedef void (*Callback)(const char*);
  static Callback& callback() { static thread_local Callback cb = nullptr; return cb; }  // LGBM_RegisterLogCallback (log.h:77-80)
  static void write(int lvl, const char* tag, const char* fmt, va_list ap) {
    if (lvl > level()) return;
    if (callback() != nullptr) {   // the reference hands the pieces of a line to the callback (log.h:134-150)
      char buf[1024];
      std::snprintf(buf, sizeof(buf), "[LightGBM-B200] [%s] ", tag);
      callback()(buf);
      std::vsnprintf(buf, sizeof(buf), fmt, ap);
      callback()(buf);
      callback()("\n");
      return;
    }
    std::fprintf(stdout, "[LightGBM-B200] [%s] ", tag);
    std::vfprintf(stdout, fmt, ap);
    std::fprintf(stdout, "\n");
    std::fflush(stdout);
  }
  static void Info(const char* fmt, ...) { va_list ap; va_start(ap, fmt); write(1, "Info", fmt, ap); va_end(ap); }
  static void Warning(const char* fmt, ...) { va_list ap; va_start(ap, fmt); write(0, "Warning", fmt, ap); va_end(ap); }
  static void Debug(const char* fmt, ...) { va_list ap; va_start(ap, fmt); write(2, "Debug", fmt, ap); va_end(ap); }
  [[noreturn]] static void Fatal(const char* fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); std::vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    throw std::runtime_error(std::string(buf));
  }
};

inline int RoundInt(double x) { return static_cast<int>(x + 0.5f); }                 // utils/common.h:985-987
inline double NextUp(double a) { return std::nextafter(a, INFINITY); }                // GetDoubleUpperBound
inline bool EqualOrdered(double a, double b) { return b <= std::nextafter(a, INFINITY); }  // CheckDoubleEqualOrdered

inline std::string Trim(const std::string& s) {
  size_t a = s.find_first_not_of(" \f\n\r\t\v");
  if (a == std::string::npos) return "";
  size_t b = s.find_last_not_of(" \f\n\r\t\v");
  return s.substr(a, b - a + 1);
}

inline std::vector<std::string> SplitAny(const std::string& s, const char* delims) {
  std::vector<std::string> out;
  size_t i = 0;
  while (i < s.size()) {
    size_t j = s.find_first_of(delims, i);
    if (j == std::string::npos) j = s.size();
    if (j > i) out.push_back(s.substr(i, j - i));
    i = j + 1;
  }
  return out;
}

// Common::ArrayToString / ArrayToStringFast (utils/common.h:411-502): "%.17g" for model doubles, "%g" otherwise
template <typename T>
inline std::string JoinInts(const std::vector<T>& a, size_t n) {
  std::string s;
  for (size_t i = 0; i < n && i < a.size(); ++i) { if (i) s += ' '; s += std::to_string(a[i]); }
  return s;
}
template <typename T>
inline std::string JoinG(const std::vector<T>& a, size_t n) {
  std::string s; char buf[32];
  for (size_t i = 0; i < n && i < a.size(); ++i) {
    std::snprintf(buf, sizeof(buf), "%g", static_cast<double>(a[i]));
    if (i) s += ' ';
    s += buf;
  }
  return s;
}
inline std::string Join17(const std::vector<double>& a, size_t n) {
  std::string s; char buf[40];
  for (size_t i = 0; i < n && i < a.size(); ++i) {
    std::snprintf(buf, sizeof(buf), "%.17g", a[i]);
    if (i) s += ' ';
    s += buf;
  }
  return s;
}

}  // namespace lgb
