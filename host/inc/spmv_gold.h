// spmv_gold.h -- the harness's in-run CPU check for the plus-times SpMV
// (reference: inc/spmv_gold.h:9-28, called from app/spmv.cpp:144, compared by Harness::check_result).
//
// It exists only to fill the "correct / badvalues" column of the result rows, exactly as in the
// reference; no device result is ever replaced by it.  The loop is the reference's arithmetic:
// for each stored (col, v) of a row, in file order (values through int, as the reference's ragged rows hold them),
//     acc += (alpha * (x(col) * v)) + (beta * y((int)v))
// including its quirk that the beta term is added per non-zero and indexed by the VALUE (SURVEY.md
// Q5); the spmv app runs with beta = 0, where it coincides with the device formula.
#pragma once

#include <vector>

#include "csds_timer.h"
#include "sparse_matrix.h"
#include "vector_generator.h"

template <typename T> class Gold {
public:
  static std::vector<T> spmv(SparseMatrix<T> &A, XVectorGenerator<T> &x, YVectorGenerator<T> &y, T alpha, T beta,
                             T zero) {
    start_timer(spmv, gold);
    // The reference walks its ragged rows (calculate_ellpack: tuples appended row by row in file order, values
    // through int).  The stable counting sort of csr_encode(true) stores exactly that sequence per row, in three flat
    // arrays instead of one heap vector per row (1 s and 24 bytes per entry less on a 16.7 M-entry file).
    const CsrArrays<T> rows = A.csr_encode(true);
    const std::size_t height = (std::size_t)std::max(A.height(), 0);
    std::vector<T> result(height, 0);
    for (std::size_t i = 0; i < height; ++i) {
      T acc = zero;
      for (int64_t k = rows.row_ptr[i]; k < rows.row_ptr[i + 1]; ++k) {
        const T v = rows.vals[(std::size_t)k];
        acc += (alpha * (x.get(rows.col_idx[(std::size_t)k]) * v)) + (beta * y.get((int)v));
      }
      result[i] = acc;
    }
    return result;
  }

private:
  Gold() {}
};
