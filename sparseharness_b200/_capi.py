"""ctypes declarations for lib/libshb.so -- exactly the entry points of include/shb.h.

There is no fallback: if the CUDA library has not been built (python __graft_entry__.py build, or
make -C sparseharness_b200/csrc) importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# SHB_LIB=checked loads the build with device-side index checks (make -C sparseharness_b200/csrc CHECKED=1)
LIB_PATH = os.path.join(HERE, "lib", "libshb_checked.so" if os.environ.get("SHB_LIB") == "checked" else "libshb.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()' "
        "or make -C sparseharness_b200/csrc). sparseharness_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH)

c_void_pp = C.POINTER(C.c_void_p)
u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int)


class LaunchCfg(C.Structure):
    _fields_ = [("variant", C.c_int32), ("lanes", C.c_int32), ("rows_per_group", C.c_int32),
                ("block_threads", C.c_int32), ("grid_blocks", C.c_int32), ("flags", C.c_uint32)]


class CsrInfo(C.Structure):
    _fields_ = [("n_rows", C.c_int64), ("n_cols", C.c_int64), ("nnz", C.c_int64), ("max_row_len", C.c_int64),
                ("min_row_len", C.c_int64), ("empty_rows", C.c_int64), ("mean_row_len", C.c_double),
                ("has_values", C.c_int32), ("elem_bytes", C.c_int32), ("auto_variant", C.c_int32),
                ("auto_lanes", C.c_int32), ("stream_tiles", C.c_int64), ("straddling_rows", C.c_int64),
                ("device_bytes", C.c_uint64), ("algorithmic_bytes_spmv", C.c_uint64)]


class GraphStats(C.Structure):
    _fields_ = [("iterations", C.c_uint32), ("converged", C.c_uint32), ("edges_traversed", C.c_int64),
                ("reached", C.c_int64), ("total_ns", C.c_uint64), ("launches", C.c_uint64)]


# name -> (restype, argtypes); every symbol include/shb.h declares
SIGNATURES = {
    "shb_abi_version": (C.c_int, []),
    "shb_status_string": (C.c_char_p, [C.c_int]),
    "shb_last_error": (C.c_char_p, []),
    "shb_device_count": (C.c_int, [i32p]),
    "shb_device_max_alloc_bytes": (C.c_int, [C.c_int, u64p]),
    "shb_device_props": (C.c_int, [C.c_int, C.c_char_p, C.c_size_t, i32p, i32p, i32p, u64p, u64p]),
    "shb_ctx_create": (C.c_int, [C.c_int, c_void_pp]),
    "shb_ctx_destroy": (C.c_int, [C.c_void_p]),
    "shb_ctx_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "shb_ctx_synchronize": (C.c_int, [C.c_void_p]),
    "shb_ctx_device_name": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "shb_ctx_device": (C.c_int, [C.c_void_p, i32p]),
    "shb_buf_create": (C.c_int, [C.c_void_p, C.c_uint64, c_void_pp]),
    "shb_buf_wrap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, c_void_pp]),
    "shb_buf_destroy": (C.c_int, [C.c_void_p]),
    "shb_buf_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, u64p]),
    "shb_buf_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, u64p]),
    "shb_buf_fill_zero": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, u64p]),
    "shb_buf_fill_u32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32]),
    "shb_buf_copy": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "shb_buf_size": (C.c_int, [C.c_void_p, u64p]),
    "shb_buf_device_ptr": (C.c_int, [C.c_void_p, c_void_pp]),
    "shb_host_alloc_pinned": (C.c_int, [C.c_uint64, c_void_pp]),
    "shb_host_free_pinned": (C.c_int, [C.c_void_p]),
    "shb_csr_create": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_int, c_void_pp, u64p]),
    "shb_csr_from_coo": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                   c_void_pp, u64p]),
    "shb_csr_destroy": (C.c_int, [C.c_void_p]),
    "shb_csr_get_info": (C.c_int, [C.c_void_p, C.POINTER(CsrInfo)]),
    "shb_csr_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "shb_csr_transpose": (C.c_int, [C.c_void_p, C.c_void_p, c_void_pp]),
    "shb_csr_build_blocked": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "shb_csr_blocked_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64), u64p]),
    "shb_csr_build_panels": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, i32p]),
    "shb_csr_balance_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "shb_csr_slice_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, c_void_pp]),
    "shb_spmv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(LaunchCfg), C.c_void_p, C.c_void_p, C.c_void_p,
                           C.c_void_p, C.c_void_p, u64p]),
    "shb_spmv_stream_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, u64p]),
    "shb_spmv_stream_static_max_nnz": (C.c_int, [C.POINTER(C.c_int64)]),
    "shb_spmv_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(LaunchCfg), C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, u64p, u64p]),
    "shb_spmv_host_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(LaunchCfg), C.c_int, c_void_pp, c_void_pp,
                                      C.c_void_p, C.c_void_p, c_void_pp, u64p]),
    "shb_iterate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(LaunchCfg), C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_uint32, C.c_uint32, u64p, u32p,
                              i32p, u64p]),
    "shb_bfs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32, C.POINTER(GraphStats)]),
    "shb_sssp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32, C.POINTER(GraphStats)]),
    "shb_pagerank": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(LaunchCfg), C.c_void_p, C.c_void_p, C.c_float,
                               C.c_float, C.c_double, C.c_uint32, C.c_uint32, i32p, C.POINTER(GraphStats)]),
    "shb_power_iteration": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(LaunchCfg), C.c_void_p, C.c_void_p, C.c_double, C.c_uint32,
                                      C.POINTER(C.c_float), i32p, C.POINTER(GraphStats)]),
    "shb_synth_uniform": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_uint64, C.c_int, C.c_int64,
                                    C.c_int64, c_void_pp]),
    "shb_synth_vector": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint64, C.c_int]),
    "shb_synth_rmat": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_int, C.c_int, C.c_float, c_void_pp]),
    "shb_synth_grid2d": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_uint64, c_void_pp]),
    "shb_ipc_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "shb_ipc_open": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, c_void_pp]),
    "shb_ipc_close": (C.c_int, [C.c_void_p]),
    "shb_csr_row_lengths": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "shb_bfs_block_state_bytes": (C.c_int, [u64p]),
    "shb_bfs_block_counter_bytes": (C.c_int, [u64p, u32p]),
    "shb_bfs_block_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_int]),
    "shb_bfs_block_publish": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, c_void_pp, C.c_int, C.c_int,
                                        C.c_uint64]),
    "shb_bfs_block_level": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_uint32, C.c_void_p,
                                      C.c_void_p, c_void_pp, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, c_void_pp, C.c_uint64]),
    "shb_bfs_block_label_frontier": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_uint32, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "shb_bfs_block_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "shb_sssp_block_work_bytes": (C.c_int, [C.c_int64, u64p]),
    "shb_sssp_block_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.c_void_p]),
    "shb_sssp_block_seed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_uint64]),
    "shb_sssp_block_solve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, c_void_pp, C.c_int,
                                       C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_float,
                                       C.c_uint32]),
    "shb_sssp_block_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "shb_spmv_scatter": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(LaunchCfg), C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, c_void_pp, C.c_int, C.c_int64, C.c_double, C.c_int,
                                   C.c_void_p, C.c_void_p, u64p]),
    "shb_group_create": (C.c_int, [C.c_int, i32p, c_void_pp]),
    "shb_group_destroy": (C.c_int, [C.c_void_p]),
    "shb_group_size": (C.c_int, [C.c_void_p, i32p]),
    "shb_group_ctx": (C.c_int, [C.c_void_p, C.c_int, c_void_pp]),
    "shb_group_partition_host": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                           c_void_pp]),
    "shb_group_adopt": (C.c_int, [C.c_void_p, c_void_pp, c_void_pp, C.c_void_p, C.c_int64, c_void_pp]),
    "shb_gpart_destroy": (C.c_int, [C.c_void_p]),
    "shb_gpart_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_void_p]),
    "shb_gpart_block": (C.c_int, [C.c_void_p, C.c_int, c_void_pp, c_void_pp]),
    "shb_group_iterate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(LaunchCfg), C.c_void_p, C.c_uint32, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p,
                                    C.c_void_p, u32p, u64p]),
    "shb_group_pagerank": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(LaunchCfg), C.c_float, C.c_float, C.c_double, C.c_uint32,
                                     C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(GraphStats)]),
    "shb_group_bfs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(GraphStats)]),
    "shb_group_sssp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_float, C.c_uint32, C.c_uint32,
                                 C.POINTER(GraphStats)]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _f = getattr(lib, _name)  # AttributeError here = the library does not export a declared symbol
    _f.restype = _res
    _f.argtypes = _args
