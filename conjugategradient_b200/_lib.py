"""ctypes binding of the C-ABI library (include/cgb200.h -> conjugategradient_b200/libcgb200.so).

The library is the product; this module only declares its entry points for the Python-side
harness (tests, bench).  There is no fallback: if the shared library is missing, importing the
solver fails with the build instruction.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CGB_LIB_PATH") or os.path.join(_HERE, "libcgb200.so")  # override: A/B builds of the same library

c_i64 = C.c_int64
c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)
c_vp = C.c_void_p

CGB_OK, CGB_ERR_INVALID, CGB_ERR_CUDA, CGB_ERR_NCCL, CGB_ERR_STATE, CGB_ERR_IO, CGB_ERR_UNSUPPORTED = range(7)
FMT_AUTO, FMT_CSR, FMT_SELL, FMT_DICT = 0, 1, 2, 3
COO_MIRROR, COO_SUM_DUPLICATES, COO_COMPACT_ROWS, COO_ONE_BASED = 1, 2, 4, 8
COO_REFERENCE_DIALECT = COO_MIRROR | COO_COMPACT_ROWS | COO_ONE_BASED
LOOP_AUTO, LOOP_GRAPH, LOOP_PERSISTENT, LOOP_STEPWISE, LOOP_CLUSTER = 0, 1, 2, 3, 4
NCCL_ID_BYTES = 128


class CgOpts(C.Structure):
    _fields_ = [
        ("loop_mode", C.c_int), ("check_every", C.c_int), ("fuse_p_update", C.c_int),
        ("time_kernels", C.c_int), ("l2_persist_p", C.c_int), ("record_history", C.c_int),
        ("residual_replace_every", C.c_int), ("jacobi", C.c_int), ("reserved", C.c_int * 4),
    ]


class CgStats(C.Structure):
    _fields_ = [
        ("iterations", c_i64), ("rr", C.c_double), ("rr0", C.c_double), ("converged", C.c_int),
        ("nan_stop", C.c_int), ("solve_ms", C.c_double), ("spmv_ms", C.c_double),
        ("update_ms", C.c_double), ("pupdate_ms", C.c_double), ("timed_samples", c_i64),
        ("kernel_launches", c_i64), ("loop_mode", C.c_int), ("format", C.c_int),
        ("reserved", C.c_double * 4),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


# name -> (restype, argtypes); the single source the symbol-export test checks against the header
SIGNATURES = {
    "cgb_last_error": (C.c_char_p, []),
    "cgb_version": (C.c_int, []),
    "cgb_device_count": (C.c_int, []),
    "cgb_ctx_create": (C.c_int, [C.c_int, C.POINTER(c_vp)]),
    "cgb_ctx_destroy": (C.c_int, [c_vp]),
    "cgb_ctx_sync": (C.c_int, [c_vp]),
    "cgb_ctx_stream": (c_vp, [c_vp]),
    "cgb_ctx_device": (C.c_int, [c_vp]),
    "cgb_ctx_sm_count": (C.c_int, [c_vp]),
    "cgb_ctx_set_progress": (C.c_int, [c_vp, c_vp, c_vp]),
    "cgb_malloc": (C.c_int, [c_vp, C.c_size_t, C.POINTER(c_vp)]),
    "cgb_free": (C.c_int, [c_vp, c_vp]),
    "cgb_trim": (C.c_int, []),
    "cgb_memcpy_h2d": (C.c_int, [c_vp, c_vp, c_vp, C.c_size_t, C.c_int]),
    "cgb_memcpy_d2h": (C.c_int, [c_vp, c_vp, c_vp, C.c_size_t, C.c_int]),
    "cgb_memcpy_d2d": (C.c_int, [c_vp, c_vp, c_vp, C.c_size_t]),
    "cgb_memcpy": (C.c_int, [c_vp, c_vp, c_vp, C.c_size_t, C.c_int]),
    "cgb_memset": (C.c_int, [c_vp, c_vp, C.c_int, C.c_size_t]),
    "cgb_fill_f64": (C.c_int, [c_vp, c_vp, C.c_double, C.c_size_t]),
    "cgb_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(c_vp)]),
    "cgb_host_free": (C.c_int, [c_vp]),
    "cgb_event_record": (C.c_int, [c_vp, C.POINTER(c_vp)]),
    "cgb_event_sync": (C.c_int, [c_vp]),
    "cgb_event_elapsed_ms": (C.c_int, [c_vp, c_vp, c_dp]),
    "cgb_event_destroy": (C.c_int, [c_vp]),
    "cgb_csr_create": (C.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, C.POINTER(c_vp)]),
    "cgb_csr_from_device": (C.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, C.c_int, C.POINTER(c_vp)]),
    "cgb_csr_destroy": (C.c_int, [c_vp]),
    "cgb_csr_n": (c_i64, [c_vp]),
    "cgb_csr_nnz": (c_i64, [c_vp]),
    "cgb_csr_vals": (c_vp, [c_vp]),
    "cgb_csr_cols": (c_vp, [c_vp]),
    "cgb_csr_rowptr": (c_vp, [c_vp]),
    "cgb_csr_build_sell": (C.c_int, [c_vp, C.c_int]),
    "cgb_csr_sell_padded": (c_i64, [c_vp]),
    "cgb_csr_dict_bytes": (c_i64, [c_vp]),
    "cgb_csr_dict_patterns": (C.c_int, [c_vp]),
    "cgb_csr_set_format": (C.c_int, [c_vp, C.c_int]),
    "cgb_csr_get_format": (C.c_int, [c_vp]),
    "cgb_csr_set_panels": (C.c_int, [c_vp, C.c_int]),
    "cgb_csr_from_triplets": (C.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, C.c_int, C.POINTER(c_vp)]),
    "cgb_csr_from_triplets_device": (C.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, C.c_int, C.POINTER(c_vp)]),
    "cgb_csr_from_mm_file": (C.c_int, [c_vp, C.c_char_p, C.POINTER(c_vp)]),
    "cgb_csr_get_panels": (C.c_int, [c_vp]),
    "cgb_gen_stencil": (C.c_int, [c_vp, C.c_int, c_i64, c_i64, c_i64, c_i64, C.POINTER(c_vp)]),
    "cgb_gen_random_spd": (C.c_int, [c_vp, c_i64, C.c_uint64, C.c_int, C.c_int, C.c_double, c_i64, c_i64, C.POINTER(c_vp)]),
    "cgb_spmv": (C.c_int, [c_vp, c_vp, c_vp, c_vp]),
    "cgb_dot_acc": (C.c_int, [c_vp, c_i64, c_vp, c_vp, c_vp]),
    "cgb_norm2_acc": (C.c_int, [c_vp, c_i64, c_vp, c_vp]),
    "cgb_axpby": (C.c_int, [c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "cgb_xpby": (C.c_int, [c_vp, c_i64, c_vp, c_vp, c_vp, C.c_int, c_vp]),
    "cgb_scalar_set": (C.c_int, [c_vp, c_vp, C.c_double]),
    "cgb_scalar_get": (C.c_int, [c_vp, c_vp, c_dp]),
    "cgb_cg_default_opts": (None, [C.POINTER(CgOpts)]),
    "cgb_cg_solve": (C.c_int, [c_vp, c_vp, c_vp, c_vp, C.c_double, c_i64, C.POINTER(CgOpts), C.POINTER(CgStats)]),
    "cgb_cg_solve_host": (C.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, C.c_double, c_i64,
                                    C.POINTER(CgOpts), C.POINTER(CgStats)]),
    "cgb_cg_solve_static": (C.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, C.c_double, c_i64, C.POINTER(CgStats)]),
    "cgb_cg_history": (C.c_int, [c_vp, c_vp, c_vp, c_i64, C.POINTER(c_i64)]),
    "cgb_residual": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_dp, c_dp]),
    "cgb_memory_footprint": (C.c_size_t, [c_i64, c_i64]),
    "cgb_mm_read": (C.c_int, [C.c_char_p, C.POINTER(c_vp)]),
    "cgb_mm_n": (c_i64, [c_vp]),
    "cgb_mm_nnz": (c_i64, [c_vp]),
    "cgb_mm_vals": (c_vp, [c_vp]),
    "cgb_mm_cols": (c_vp, [c_vp]),
    "cgb_mm_rowptr": (c_vp, [c_vp]),
    "cgb_mm_free": (C.c_int, [c_vp]),
    "cgb_nccl_unique_id": (C.c_int, [c_vp]),
    "cgb_dist_create": (C.c_int, [c_vp, C.c_int, C.c_int, c_vp, c_vp, c_vp, C.POINTER(c_vp)]),
    "cgb_dist_set_matrix": (C.c_int, [c_vp, c_vp, c_vp]),
    "cgb_dist_destroy": (C.c_int, [c_vp]),
    "cgb_dist_cg_solve": (C.c_int, [c_vp, c_vp, c_vp, C.c_double, c_i64, C.POINTER(CgOpts), C.POINTER(CgStats)]),
    "cgb_dist_residual": (C.c_int, [c_vp, c_vp, c_vp, c_dp, c_dp]),
    "cgb_dist_halo_count": (c_i64, [c_vp]),
    "cgb_dist_exchange_mode": (C.c_int, [c_vp]),
    "cgb_plan_halo": (C.c_int, [C.c_int, C.c_int, c_vp, c_i64, c_vp, c_vp, c_i64, C.POINTER(c_i64), c_vp]),
}

_lib = None


class CgbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def load():
    """Load libcgb200.so (once) and attach the prototypes.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C conjugategradient_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != CGB_OK:
        raise CgbError(rc, load().cgb_last_error().decode(errors="replace"))
