// just_parser_harness -- times the Matrix Market parse and the argument encoding `trials` times without
// touching the device (reference: app/just_parser.cpp:40-96).  Two additions used by the CPU tests:
//   --elem float|int      element type of the operator (default float)
//   --normalise none|pr|scc
//   --gpu_encode          encode COO -> CSR on the device (shb_csr_from_coo) instead of the host counting sort;
//                         --dump_csr then writes what the DEVICE built (downloaded), so the two encoders can be compared
//   --dump_csr PREFIX     write PREFIX.rowptr (int64), PREFIX.col (int32), PREFIX.val (4-byte elements),
//                         PREFIX.x / PREFIX.y (the generated vectors) and, for float, PREFIX.gold (the in-run gold
//                         values for x = 1, y = 0, alpha = 1, beta = 0) of the LAST trial
#include <limits>

#include "app_common.h"
#include "device_utils.h"
#include "kernel_utils.h"
#include "spmv_gold.h"

template <typename T> static void write_file(const std::string &path, const std::vector<T> &v) { dump_vector(path, v); }

template <typename T>
static int parse_trials(unsigned trials, const std::string &matrix_filename, const std::string &kernel_filename,
                        const std::string &normalise, const std::string &dump, bool truncate, bool gpu_encode) {
  const uint64_t one_gb = 1ull << 30;
  shb_ctx *ctx = nullptr;
  if (gpu_encode) checkShbError(shb_ctx_create(0, &ctx));
  for (unsigned int i = 0; i < trials; i++) {
    SparseMatrix<T> matrix(matrix_filename);
    KernelConfig<T> kernel(kernel_filename);
    if (matrix.height() != matrix.width()) {
      std::cout << "Matrix is not square. Failing computation." << ENDL;
      std::cerr << "Matrix is not square. Failing computation." << ENDL;
      std::exit(2);
    } else {
      std::cout << " Matrix is square - width = " << matrix.width() << " and height = " << matrix.height() << "\n";
    }
    if (normalise == "pr") matrix.pagerank_normalise(0.85f, (T)0);
    if (normalise == "scc") matrix.scc_normalise();
    ConstXVectorGenerator<T> onegen((T)1);
    ConstYVectorGenerator<T> zerogen((T)0);
    try {
      auto args = executorEncodeMatrix<T>(one_gb, kernel, matrix, (T)0, onegen, zerogen, (T)1, (T)0, truncate, gpu_encode);
      if (gpu_encode) {
        start_timer(gpu_encode, just_parser);
        shb_csr *A = nullptr;
        uint64_t ns = 0;
        checkShbError(shb_csr_from_coo(ctx, args.matrix.n_rows, args.matrix.n_cols, (int64_t)matrix.storedEntries(),
                                       matrix.tupleFirst().data(), matrix.tupleSecond().data(), matrix.tupleValues().data(),
                                       std::is_floating_point<T>::value ? 0 : 1, truncate ? 1 : 0, &A, &ns));
        report_timing(shb_csr_from_coo, just_parser, ns);
        if (!dump.empty() && i + 1 == trials) {  // what the device built, for the comparison with the host encoder
          args.matrix.row_ptr.resize((std::size_t)args.matrix.n_rows + 1);
          args.matrix.col_idx.resize(matrix.storedEntries());
          args.matrix.vals.resize(matrix.storedEntries());
          checkShbError(shb_csr_download(ctx, A, 0, args.matrix.n_rows, args.matrix.row_ptr.data(), args.matrix.col_idx.data(),
                                         args.matrix.vals.data()));
        }
        shb_csr_destroy(A);
      }
      if (!dump.empty() && i + 1 == trials) {
        write_file(dump + ".rowptr", args.matrix.row_ptr);
        write_file(dump + ".col", args.matrix.col_idx);
        write_file(dump + ".val", args.matrix.vals);
        write_file(dump + ".x", args.x_vect);
        write_file(dump + ".y", args.y_vect);
        write_file(dump + ".gold", Gold<T>::spmv(matrix, onegen, zerogen, (T)1, (T)0, (T)0));
        std::cout << "semiring " << args.semiring << " variant " << args.launch.variant << " lanes " << args.launch.lanes
                  << " flags " << args.launch.flags << " output_bytes " << args.output << " rows " << args.matrix.n_rows
                  << " nnz " << args.matrix.nnz() << ENDL;
      }
    } catch (unsigned long alloc) {
      LOG_ERROR("Tried to alloc ", alloc, " bytes of memory on the gpu, when maximum is ", one_gb);
    }
  }
  return 0;
}

int main(int argc, char **argv) {
  start_timer(main, global);
  OptParser op("Benchmark program to test the speed of the sparse matrix parser and specialisation routines without "
               "performing any actual GPU work.");
  auto opt_trials = op.addOption<unsigned>({'i', "trials", "Execute each kernel 'trials' times (default 10).", 10});
  auto opt_matrix_file = op.addOption<std::string>({'m', "matrix", "Input matrix"});
  auto opt_kernel_file = op.addOption<std::string>({'k', "kernel", "Input kernel"});
  auto opt_experiment_id =
      op.addOption<std::string>({'e', "experiment", "An experiment ID for data reporting", "null_experiment"});
  auto opt_elem = op.addOption<std::string>({0, "elem", "Element type: float (default) or int.", "float"});
  auto opt_norm = op.addOption<std::string>({0, "normalise", "none (default), pr or scc.", "none"});
  auto opt_dump = op.addOption<std::string>({0, "dump_csr", "Prefix of the files the encoded operator is written to.", ""});
  auto opt_keep = op.addOption<bool>({0, "keep_values", "Do not pass values through int (canonical operator).", false});
  auto opt_gpu = op.addOption<bool>({0, "gpu_encode", "Encode COO -> CSR on the device (needs a GPU).", false});
  op.parse(argc, argv);

  const std::string matrix_filename = opt_matrix_file->require();
  const std::string kernel_filename = opt_kernel_file->require();
  std::cerr << "matrix_filename " << matrix_filename << ENDL;
  std::cerr << "kernel_filename " << kernel_filename << ENDL;
  (void)opt_experiment_id;
  if (opt_elem->get() == "int")
    return parse_trials<int>(opt_trials->get(), matrix_filename, kernel_filename, opt_norm->get(), opt_dump->get(),
                             !opt_keep->get(), opt_gpu->get());
  return parse_trials<float>(opt_trials->get(), matrix_filename, kernel_filename, opt_norm->get(), opt_dump->get(),
                             !opt_keep->get(), opt_gpu->get());
}
