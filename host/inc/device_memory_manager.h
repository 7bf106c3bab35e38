// device_memory_manager.h -- the harness's bag of device handles and host mirrors.
// Replaces the reference's inc/cl_memory_manager.h:6-29 (CLMemoryManager<T>): cl_mem handles become
// opaque libshb handles; the two matrix blobs become one CSR operator handle.
#pragma once

#include <vector>

#include "kernel_utils.h"
#include "shb.h"

template <typename SemiringType> class DeviceMemoryManager {
public:
  explicit DeviceMemoryManager(const ArgContainer<SemiringType> &args)
      : _input_host_buffer(args.x_vect.begin(), args.x_vect.end()), _output_host_buffer(args.output, 0),
        _temp_out_buffer(args.output, 0) {}

  shb_csr *_matrix = nullptr;  // arguments 0 + 1 (idx, vals)
  shb_buf *_x_vect = nullptr;  // argument 2
  shb_buf *_y_vect = nullptr;  // argument 3
  shb_buf *_output = nullptr;  // argument 6
  std::vector<shb_buf *> _temp_global;  // always empty: no CUDA schedule needs global temporaries

  unsigned int _input_idx = 2;
  unsigned int _output_idx = 6;

  std::vector<char> _input_host_buffer;
  std::vector<char> _output_host_buffer;
  std::vector<char> _temp_out_buffer;
};
