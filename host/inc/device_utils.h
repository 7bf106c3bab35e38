// device_utils.h -- error handling and device queries over libshb's C ABI.
// Replaces the reference's inc/opencl_utils.h:
//   checkCLError(err)  (:15-23)   -> checkShbError(status): log + exit(1), same convention
//   getErrorString     (:25-171)  -> getErrorString(status): "<NAME> (code: n)" + libshb's detail text
//   deviceGetMaxAllocSize(platform, device) (:216-226) -> same name; `platform` is accepted and ignored
//   printCharVector<T> (:228-245), assertBuffersNotEqual (:247-258) -> unchanged behaviour, including the
//   "Buffers are identical - kernel has probably failed..." line scripts/experiments/all_kernel_test.sh:50 greps
#pragma once

#include <cstdlib>
#include <sstream>
#include <string>
#include <vector>

#include "logger.h"
#include "shb.h"

inline std::string getErrorString(int status) {
  std::ostringstream ostr;
  ostr << shb_status_string(status) << " (code: " << status << ")";
  const char *detail = shb_last_error();
  if (status != SHB_OK && detail && *detail) ostr << ": " << detail;
  return ostr.str();
}

#define checkShbError(expr)                                                                                    \
  do {                                                                                                         \
    const int _shb_status = (expr);                                                                            \
    if (_shb_status != SHB_OK) {                                                                               \
      LOG_ERROR("Device call failed with error ", getErrorString(_shb_status), " in ", __FILE__, " at line ", \
                __LINE__);                                                                                     \
      std::exit(1);                                                                                            \
    }                                                                                                          \
  } while (0)

inline unsigned long deviceGetMaxAllocSize(unsigned int platform, unsigned int device) {
  (void)platform;
  uint64_t bytes = 0;
  int count = 0;
  checkShbError(shb_device_count(&count));
  if (count == 0) {
    LOG_ERROR("No CUDA devices found!");
    std::exit(1);
  }
  checkShbError(shb_device_max_alloc_bytes((int)device, &bytes));
  return (unsigned long)bytes;
}

template <typename T> void printCharVector(const std::string &name, const std::vector<char> &v) {
  if (!shb_log::debugEnabled()) return;
  const T *data = reinterpret_cast<const T *>(v.data());
  const std::size_t length = v.size() / sizeof(T);
  if (length > 4096) {  // a 2^24-element dump helps nobody
    LOG_DEBUG_INFO("Buffer ", name, "[", length, " elements, not printed]");
    return;
  }
  std::ostringstream ostr;
  ostr << "[";
  for (std::size_t i = 0; i < length; ++i) {
    if (i) ostr << ",";
    ostr << data[i];
  }
  ostr << "]";
  LOG_DEBUG_INFO("Buffer ", name, ostr.str());
}

inline void assertBuffersNotEqual(const std::vector<char> &v1, const std::vector<char> &v2) {
  bool different_found = v1.size() != v2.size();
  for (std::size_t i = 0; !different_found && i < v1.size(); ++i) different_found = v1[i] != v2[i];
  if (!different_found) LOG_ERROR("Buffers are identical - kernel has probably failed...");
}
