// vector_generator.h -- generators for the dense x / y vectors (reference: inc/vector_generator.h:9-61).
// generate(len) evaluates get(0..len-1); the gold loop calls get(ix) directly with arbitrary indices.
#pragma once

#include <vector>

#include "csds_timer.h"

template <typename T> class VectorGenerator {
public:
  virtual ~VectorGenerator() {}
  virtual T get(int ix) = 0;

  std::vector<T> generate(int length) {
    start_timer(generate, VectorGenerator);
    std::vector<T> v((std::size_t)(length > 0 ? length : 0));
    for (int i = 0; i < length; ++i) v[(std::size_t)i] = get(i);
    return v;
  }
};

// the "x" of y = A x: as long as the (padded) matrix height
template <typename T> class XVectorGenerator : public VectorGenerator<T> {};
// the "y" of alpha*A x + beta*y: as long as the padded matrix height
template <typename T> class YVectorGenerator : public VectorGenerator<T> {};

template <typename T> class ConstXVectorGenerator : public XVectorGenerator<T> {
public:
  explicit ConstXVectorGenerator(T constv) : value(constv) {}
  T get(int) override { return value; }

private:
  T value;
};

template <typename T> class ConstYVectorGenerator : public YVectorGenerator<T> {
public:
  explicit ConstYVectorGenerator(T constv) : value(constv) {}
  T get(int) override { return value; }

private:
  T value;
};

// value `first` at index 0 and `rest` elsewhere: the initial distances of sssp (0, FLT_MAX...;
// app/sssp.cpp:179-209) and the initial frontier of bfs (1, 0...; app/bfs.cpp:177-207)
template <typename T, typename Base> class SourceVectorGenerator : public Base {
public:
  SourceVectorGenerator(T first, T rest) : _first(first), _rest(rest) {}
  T get(int ix) override { return ix == 0 ? _first : _rest; }

private:
  T _first, _rest;
};

// x[i] = i: the initial component labels of scc (app/scc.cpp:179-195)
template <typename T, typename Base> class IotaVectorGenerator : public Base {
public:
  T get(int ix) override { return (T)ix; }
};
