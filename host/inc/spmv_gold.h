// spmv_gold.h -- the harness's in-run CPU check for the plus-times SpMV
// (reference: inc/spmv_gold.h:9-28, called from app/spmv.cpp:144, compared by Harness::check_result).
//
// It exists only to fill the "correct / badvalues" column of the result rows, exactly as in the
// reference; no device result is ever replaced by it.  The loop is the reference's arithmetic:
// for each stored (col, v) of a row, in file order,
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
    auto &rows = A.ellpack_encode();
    std::vector<T> result(rows.size(), 0);
    for (std::size_t i = 0; i < rows.size(); ++i) {
      T acc = zero;
      for (std::size_t j = 0; j < rows[i].size(); ++j) {
        const auto &elem = rows[i][j];
        acc += (alpha * (x.get(elem.first) * elem.second)) + (beta * y.get((int)elem.second));
      }
      result[i] = acc;
    }
    return result;
  }

private:
  Gold() {}
};
