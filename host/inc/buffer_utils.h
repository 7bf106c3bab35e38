// buffer_utils.h -- typed vector <-> raw byte buffer (reference: inc/buffer_utils.h:77-88 `enchar`).
#pragma once

#include <cstring>
#include <vector>

template <typename T> std::vector<char> enchar(const std::vector<T> &in) {
  std::vector<char> out(in.size() * sizeof(T));
  if (!in.empty()) std::memcpy(out.data(), in.data(), out.size());
  return out;
}

template <typename T> std::vector<T> dechar(const std::vector<char> &in) {
  std::vector<T> out(in.size() / sizeof(T));
  if (!out.empty()) std::memcpy(out.data(), in.data(), out.size() * sizeof(T));
  return out;
}
