// kernel_utils.h -- turns a loaded matrix + a kernel configuration into the arguments of one harness
// (reference: inc/kernel_utils.h:18-168, `ArgContainer` and `executorEncodeMatrix`).
//
// Same steps as the reference: derive the sizes v_MWidth_1 / v_MHeight_2 / v_VLength_3 from the
// kernel's properties (:57-65), generate x and y for the PADDED height (:75-76), evaluate the size
// expressions of the output / temporary buffers (:109-140), order the size arguments by "paramVars"
// (:145-160).  What changes is the matrix layout: the reference encodes padded ELLPACK or RSA byte
// blobs (SparseMatrix::cl_encode); here arguments 0 "idx" and 1 "vals" are the CSR arrays libshb
// consumes.  Padded rows (chunked kernels, Q7: a full extra chunk even when the height divides) are
// kept as EMPTY CSR rows so that x, y, out have the reference's lengths and the iterative apps'
// termination test sees the same padded elements.
#pragma once

#include <cstdint>
#include <cstdlib>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "arithexpr_evaluator.h"
#include "buffer_utils.h"
#include "common.h"
#include "csds_timer.h"
#include "kernel_config.h"
#include "logger.h"
#include "shb.h"
#include "sparse_matrix.h"
#include "vector_generator.h"

typedef std::vector<char> raw_arg;

template <typename T> class ArgContainer {
public:
  // arguments 0/1 of the kernel ABI: the operator (CSR instead of the ELL / RSA blobs)
  CsrArrays<T> matrix;
  bool pattern = false;  // every stored value is the semiring one: upload no value array
  // arguments 2/3: x and y as raw bytes, padded length
  raw_arg x_vect, y_vect;
  // arguments 4/5
  T alpha, beta;
  // sizes of the buffers the kernel JSON declares (bytes); the CUDA kernels need no temporaries,
  // the numbers are kept for reporting and for the output length
  std::vector<uint64_t> temp_globals;
  uint64_t output = 0;
  std::vector<uint64_t> temp_locals;
  std::vector<unsigned int> size_args;
  // what replaces the JIT of "source": semiring + schedule picked from the JSON
  int semiring = SHB_PLUS_TIMES_F32;
  shb_launch_cfg launch = {};
  // when set, the operator is generated on the device (BASELINE configs 2-5) and `matrix` is empty
  std::string synthetic;
  // --gpu_encode: the tuples go to the device as they are and the CSR is built there (shb_csr_from_coo); `matrix`
  // then carries the sizes only.  The SparseMatrix outlives the harness (it lives in main).
  const SparseMatrix<T> *coo = nullptr;
  bool coo_truncate = true;
};

// padded height of the reference's cl_encode (src/sparse_matrix.cpp:147-150)
inline int padded_height(int height, int chunkSize) {
  return chunkSize > 0 ? height + (chunkSize - (height % chunkSize)) : height;
}

template <typename T>
ArgContainer<T> executorEncodeMatrix(uint64_t device_max_alloc_bytes, KernelConfig<T> &kernel, SparseMatrix<T> &matrix,
                                     T zero, XVectorGenerator<T> &xgen, YVectorGenerator<T> &ygen,
                                     T alpha = static_cast<T>(1), T beta = static_cast<T>(1),
                                     bool truncate_values = true, bool gpu_encode = false) {
  start_timer(executorEncodeMatrix, kernel_utils);
  (void)zero;  // the ELL padding value; CSR stores no padding entries
  const KernelProperties kprops = kernel.getProperties();
  const bool ragged = kprops.arrayType == "ragged";
  const int cl_height = padded_height(matrix.height(), kprops.chunkSize);

  ArgContainer<T> arg_cnt;
  if (gpu_encode) {
    arg_cnt.coo = &matrix;
    arg_cnt.coo_truncate = truncate_values;
    arg_cnt.matrix.n_rows = cl_height;
    arg_cnt.matrix.n_cols = std::max<int64_t>(matrix.width(), cl_height);
  } else {
    arg_cnt.matrix = matrix.csr_encode(truncate_values, cl_height);
  }

  // the reference bails out when the index buffer exceeds the device's max allocation
  // (src/sparse_matrix.cpp:231-233, caught in main)
  const uint64_t ixs_arr_size = (uint64_t)matrix.storedEntries() * sizeof(int32_t) + (uint64_t)(cl_height + 1) * 4;
  if (ixs_arr_size > device_max_alloc_bytes) throw (unsigned long)ixs_arr_size;

  // sizes the Lift kernels take as arguments (inc/kernel_utils.h:57-65).  cl_width is the ELL row
  // width the reference would have padded to.
  int cl_width = (int)matrix.maxRowLength();
  if (!ragged && kprops.splitSize > 0) cl_width = cl_width + (kprops.splitSize - (cl_width % kprops.splitSize));
  const int v_MWidth_1 = ragged ? matrix.width() : cl_width / std::abs(kprops.splitSize);
  const int v_MHeight_2 = cl_height;
  const int v_VLength_3 = cl_height;

  std::cerr << "Encoding matrix with sizes:"
            << "\n\tv_MWidth_1 = " << v_MWidth_1 << "\n\tv_MHeight_2 = " << v_MHeight_2
            << "\n\tv_VLength_3 = " << v_VLength_3 << "\n";
  std::cerr << "Filling with these sizes: \n\tx = " << matrix.height() << " \n\ty = " << cl_height << ENDL;
  std::vector<T> xvector = xgen.generate(v_VLength_3);
  std::vector<T> yvector = ygen.generate(v_VLength_3);
  LOG_DEBUG("Input vector arg: ", (size_t)xvector.size() * sizeof(T));
  LOG_DEBUG("Input vector arg: ", (size_t)yvector.size() * sizeof(T));
  arg_cnt.x_vect = enchar<T>(xvector);
  arg_cnt.y_vect = enchar<T>(yvector);
  arg_cnt.alpha = alpha;
  arg_cnt.beta = beta;

  {
    start_timer(outputBuffer, executorEncodeMatrix);
    const ArgDescr *o = kernel.getOutputArg();
    int64_t memsize = Evaluator::evaluate(o->size, v_MWidth_1, v_MHeight_2, v_VLength_3);
    // the CUDA kernels write one element per (padded) row whatever the JSON says
    if (memsize < (int64_t)cl_height * (int64_t)sizeof(T)) memsize = (int64_t)cl_height * (int64_t)sizeof(T);
    arg_cnt.output = (uint64_t)memsize;
    LOG_DEBUG("Global output arg - arg: ", o->variable, ", address space: ", o->addressSpace, ", size:", o->size,
              ", realsize: ", memsize);
  }
  {
    start_timer(tempGlobal, executorEncodeMatrix);
    for (auto &arg : kernel.getTempGlobals()) {
      const int64_t memsize = Evaluator::evaluate(arg.size, v_MWidth_1, v_MHeight_2, v_VLength_3);
      arg_cnt.temp_globals.push_back((uint64_t)memsize);
      LOG_DEBUG("Global temp arg - arg: ", arg.variable, ", address space: ", arg.addressSpace, ", size:", arg.size,
                ", realsize: ", memsize, " (not allocated: the CUDA kernels reduce in registers)");
    }
  }
  {
    start_timer(tempLocal, executorEncodeMatrix);
    for (auto &arg : kernel.getTempLocals()) {
      const int64_t memsize = Evaluator::evaluate(arg.size, v_MWidth_1, v_MHeight_2, v_VLength_3);
      arg_cnt.temp_locals.push_back((uint64_t)memsize);
      LOG_DEBUG("Local temp arg - arg: ", arg.variable, ", address space: ", arg.addressSpace, ", size:", arg.size,
                ", realsize: ", memsize);
    }
  }
  {
    start_timer(sizeArgs, executorEncodeMatrix);
    std::map<std::string, int> sizeMap{{"MWidthC", v_MWidth_1}, {"MHeight", v_MHeight_2}, {"VLength", v_VLength_3}};
    for (auto &sizeArg : kernel.getParamVars()) {
      const int size = sizeMap[sizeArg];
      LOG_DEBUG("Size argument - name: ", sizeArg, " value: ", size);
      arg_cnt.size_args.push_back((unsigned int)size);
    }
  }
  arg_cnt.semiring = kernel.getSemiring();
  arg_cnt.launch = kernel.getLaunchCfg();
  return arg_cnt;
}

// Arguments for an operator that is generated on the device (BASELINE.json configs 2-5; the
// reference can only ingest Matrix Market text, which cannot hold them).  No padding: height = n.
template <typename T>
ArgContainer<T> executorSyntheticArgs(KernelConfig<T> &kernel, const std::string &spec, int n, XVectorGenerator<T> &xgen,
                                      YVectorGenerator<T> &ygen, T alpha, T beta) {
  start_timer(executorSyntheticArgs, kernel_utils);
  ArgContainer<T> arg_cnt;
  arg_cnt.synthetic = spec;
  arg_cnt.x_vect = enchar<T>(xgen.generate(n));
  arg_cnt.y_vect = enchar<T>(ygen.generate(n));
  arg_cnt.alpha = alpha;
  arg_cnt.beta = beta;
  arg_cnt.output = (uint64_t)n * sizeof(T);
  arg_cnt.semiring = kernel.getSemiring();
  arg_cnt.launch = kernel.getLaunchCfg();
  return arg_cnt;
}
