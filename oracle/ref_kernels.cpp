// ref_kernels.cpp -- feeds the reference's own encoder output to the reference's own kernel text.
//
// TEST INFRASTRUCTURE (oracle/): compiled by oracle/Makefile (target `ref`) into
// oracle/_ref/ref_kernels together with, taken where they lie under /root/reference and never copied:
//   src/sparse_matrix.cpp src/mmio.cpp src/csds_timer.cpp src/util/Logger.cpp   (loader, cl_encode)
//   example/<alg>/kernel{5,6}.json "source"                                    (via ref_kernels_dev.cpp)
//   the generator classes of app/{sssp,bfs,scc}.cpp                            (_ref/gen/ref_generators.inc)
// This file itself contains no reference code.  What it does, per algorithm and per kernel variant:
//   1. SparseMatrix<T>(file)  [+ pagerank_normalise / scc_normalise]     src/sparse_matrix.cpp:5-119,409-456
//   2. cl_encode(max_alloc, zero, chunk?, split?, ragged?, chunkSize, splitSize) with the flags that
//      inc/kernel_utils.h:47-55 derives from the kernel JSON's properties (kernel5: none set;
//      kernel6: arrayType == "ragged")                                   src/sparse_matrix.cpp:122-399
//   3. size scalars as inc/kernel_utils.h:57-65; x, y = generators evaluated on VLength entries (:75-76)
//   4. one launch (`step1`) with zero-filled temp/out buffers            inc/harness.h:197-264
//   5. the executeRun recurrence (app/sssp.cpp:112-153 and its three twins): out <- KERNEL(x = in, y),
//      host termination test on (in, out), swap, re-bind x, out AND y to the swapped buffers; y is the
//      original y buffer only for the first launch.  Termination predicates restated from
//      app/sssp.cpp:166-171, app/pr.cpp:166-171 (fabs(in - out) < delta, early exit) and
//      app/bfs.cpp:163-168, app/scc.cpp:165-170 (in == out).  The loop is capped (the reference has no
//      cap and its bfs/pr loops need not terminate, TODO.md:3-9); the cap is reported.
// Initial states (literal constants of the apps' main(), cited, not copied):
//   spmv  x = 1, y = 0, alpha 1, beta 0, zero 0                                    app/spmv.cpp:117-131
//   sssp  x = y = InitialDistancesGenerator(FLT_MAX), alpha = beta = 0, zero FLT_MAX app/sssp.cpp:214-232
//   bfs   x = y = InitialDistancesGenerator(0) (1 at index 0), alpha 1, beta 0, zero 0  app/bfs.cpp:213-227
//   pr    x = 1/N, y = 1, alpha 1, beta (1-0.85)/N, zero 0, pagerank_normalise(0.85, 0) app/pr.cpp:182-201
//   scc   x[i] = i, y = INT_MIN, alpha INT_MAX, beta = zero = INT_MIN, scc_normalise   app/scc.cpp:201-222
//
// usage: ref_kernels <mtx> <alg: spmv|sssp|bfs|pr|scc> <variant 5|6> <cap> <delta> <gsz> <xmode> <out-file>
//   xmode (spmv only): 0 -> the app's x = 1; 1 -> x[i] = ((7 i + 3) mod 11) - 5
//   writes: "H <cl_height> W <MWidthC> ITERS <launches> TERMINATED <0|1>", then "STEP1 n" + n hex words,
//   then "FINAL n" + n hex words (bit patterns; float and int alike).
// The reference's encoder prints the whole matrix to stdout (src/sparse_matrix.cpp:387-395): callers
// redirect stdout/stderr and read <out-file>.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "sparse_matrix.h"
#include "vector_generator.h"

#include "_ref/gen/ref_generators.inc"

int refk_run_f32(int alg, int variant, int gsz, const int *idx, const void *vals, const float *x, const float *y,
                 float alpha, float beta, float *out, float *temp, int mheight, int mwidth, int vlength);
int refk_run_i32(int alg, int variant, int gsz, const int *idx, const void *vals, const int *x, const int *y,
                 int alpha, int beta, int *out, int *temp, int mheight, int mwidth, int vlength);

template <typename T> class XMode1 : public XVectorGenerator<T> {
public:
  virtual T get(int ix) { return static_cast<T>(((7 * ix + 3) % 11) - 5); }
};

static int launch(int alg, int variant, int gsz, const CL_matrix &cm, const float *x, const float *y, float a, float b,
                  float *out, float *temp, int mh, int mw, int vl) {
  return refk_run_f32(alg, variant, gsz, (const int *)cm.indices.data(), cm.values.data(), x, y, a, b, out, temp, mh,
                      mw, vl);
}
static int launch(int alg, int variant, int gsz, const CL_matrix &cm, const int *x, const int *y, int a, int b, int *out,
                  int *temp, int mh, int mw, int vl) {
  return refk_run_i32(alg, variant, gsz, (const int *)cm.indices.data(), cm.values.data(), x, y, a, b, out, temp, mh,
                      mw, vl);
}

static bool terminate(const std::vector<float> &in, const std::vector<float> &out, double delta) {
  bool equal = true;
  for (size_t i = 0; equal && i < in.size() && i < out.size(); i++) equal = std::fabs(in[i] - out[i]) < delta;
  return equal;
}
static bool terminate(const std::vector<int> &in, const std::vector<int> &out, double) {
  bool equal = true;
  for (size_t i = 0; equal && i < in.size() && i < out.size(); i++) equal = in[i] == out[i];
  return equal;
}

template <typename T> static void dump(FILE *f, const char *tag, const std::vector<T> &v) {
  std::fprintf(f, "%s %zu\n", tag, v.size());
  for (size_t i = 0; i < v.size(); ++i) {
    unsigned int w;
    std::memcpy(&w, &v[i], 4);
    std::fprintf(f, "%08x\n", w);
  }
}

template <typename T>
static int run(SparseMatrix<T> &m, int alg, int variant, XVectorGenerator<T> &xg, YVectorGenerator<T> &yg, T alpha,
               T beta, T zero, long cap, double delta, int gsz, const char *outfile) {
  const bool rsa = variant == 6;
  const int chunk = -1, split = -1; // kernel5/6 JSON: no chunkSize, no splitSize (src/kernel_config.cpp:28-38)
  CL_matrix cm = m.cl_encode(0xFFFFFFFFu, zero, chunk != -1, split != -1, rsa, chunk, split);
  const int mw = rsa ? m.width() : cm.cl_width / std::abs(split);
  const int mh = cm.cl_height, vl = cm.cl_height;
  std::vector<T> x = xg.generate(vl), y = yg.generate(vl);
  const size_t ntemp = rsa ? (size_t)mh : (size_t)mh * (size_t)mw; // tempGlobals size strings of the two JSONs
  std::vector<T> out(mh, T(0)), temp(ntemp, T(0));
  if (launch(alg, variant, gsz, cm, x.data(), y.data(), alpha, beta, out.data(), temp.data(), mh, mw, vl)) return 3;
  std::vector<T> step1 = out;

  // executeRun: buffers A (= x) and B (= output) ping-pong; y is its own buffer for the first launch only
  std::vector<T> bufA = x, bufB(mh, T(0));
  std::vector<T> *in = &bufA, *o = &bufB;
  const std::vector<T> *yy = &y;
  long iters = 0;
  bool done = false;
  do {
    std::fill(temp.begin(), temp.end(), T(0)); // resetTempBuffers(): zero BYTES, inc/harness.h:254-264
    launch(alg, variant, gsz, cm, in->data(), yy->data(), alpha, beta, o->data(), temp.data(), mh, mw, vl);
    done = terminate(*in, *o, delta);
    std::swap(in, o);
    yy = in;
    iters++;
  } while (!done && iters < cap);

  FILE *f = std::fopen(outfile, "w");
  if (!f) return 4;
  std::fprintf(f, "H %d W %d ITERS %ld TERMINATED %d\n", mh, mw, iters, done ? 1 : 0);
  dump(f, "STEP1", step1);
  dump(f, "FINAL", *in);
  std::fclose(f);
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 9) {
    std::fprintf(stderr, "usage: see header\n");
    return 64;
  }
  const std::string file = argv[1], alg = argv[2];
  const int variant = std::atoi(argv[3]);
  const long cap = std::atol(argv[4]);
  const double delta = std::atof(argv[5]);
  const int gsz = std::atoi(argv[6]);
  const int xmode = std::atoi(argv[7]);
  const char *outfile = argv[8];
  if (alg == "spmv") {
    SparseMatrix<float> m(file);
    ConstXVectorGenerator<float> x1(1.0f);
    XMode1<float> x2;
    ConstYVectorGenerator<float> y(0);
    XVectorGenerator<float> &x = xmode ? static_cast<XVectorGenerator<float> &>(x2) : x1;
    return run<float>(m, 0, variant, x, y, 1.0f, 0.0f, 0.0f, 1, delta, gsz, outfile);
  }
  if (alg == "sssp") {
    SparseMatrix<float> m(file);
    g_sssp::InitialDistancesGeneratorX<float> x(std::numeric_limits<float>::max());
    g_sssp::InitialDistancesGeneratorY<float> y(std::numeric_limits<float>::max());
    return run<float>(m, 1, variant, x, y, 0.0f, 0.0f, std::numeric_limits<float>::max(), cap, delta, gsz, outfile);
  }
  if (alg == "bfs") {
    SparseMatrix<int> m(file);
    g_bfs::InitialDistancesGeneratorX<int> x(0);
    g_bfs::InitialDistancesGeneratorY<int> y(0);
    return run<int>(m, 2, variant, x, y, 1, 0, 0, cap, delta, gsz, outfile);
  }
  if (alg == "pr") {
    SparseMatrix<float> m(file);
    const float d = 0.85f;
    ConstXVectorGenerator<float> x(1.0f / (float)m.height());
    ConstYVectorGenerator<float> y(1.0f);
    const float beta = (1.0f - d) / (float)m.height();
    m.pagerank_normalise(d, 0.0f);
    return run<float>(m, 3, variant, x, y, 1.0f, beta, 0.0f, cap, delta, gsz, outfile);
  }
  if (alg == "scc") {
    SparseMatrix<int> m(file);
    g_scc::InitialComponentsGeneratorX<int> x;
    ConstYVectorGenerator<int> y(std::numeric_limits<int>::min());
    m.scc_normalise();
    return run<int>(m, 4, variant, x, y, std::numeric_limits<int>::max(), std::numeric_limits<int>::min(),
                    std::numeric_limits<int>::min(), cap, delta, gsz, outfile);
  }
  return 64;
}
