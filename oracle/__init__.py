"""TEST INFRASTRUCTURE ONLY — ctypes loaders for the parity oracle.

``port``  : oracle/_build/libcg_oracle.so, the plain-C restatement (oracle/cg_oracle.c).
``ref(flavour)`` : oracle/_ref/libref_{serial,omp}.so, the reference's own unmodified sources compiled
            against oracle/sycl_standin (only exists where /root/reference was available at
            build time; the built files travel to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / reference legs may import this
package. The product (conjugategradient_b200/, include/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(ref: bool = True) -> None:
    """Compile the oracle (and the reference bridge when /root/reference exists)."""
    targets = ["port"] + (["ref"] if ref else [])
    subprocess.run(["make", "-C", _HERE, "-s"] + targets, check=True)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class Port:
    """Plain-C restatement (oracle/cg_oracle.c)."""

    def __init__(self, path=None):
        path = path or os.path.join(_HERE, "_build", "libcg_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = self.lib = C.CDLL(path)
        L.oracle_spmv.argtypes = [C.c_long, _dp, _ip, _ip, _dp, _dp]
        L.oracle_dot_acc.argtypes = [C.c_long, _dp, _dp, C.POINTER(C.c_double)]
        L.oracle_norm_acc.argtypes = [C.c_long, _dp, C.POINTER(C.c_double)]
        L.oracle_saxpby.argtypes = [C.c_long, _dp, _dp, C.c_double, C.c_double, _dp]
        L.oracle_sapbx.argtypes = [C.c_long, _dp, _dp, C.c_double, _dp]
        L.oracle_sambx.argtypes = [C.c_long, _dp, _dp, C.c_double, _dp]
        L.oracle_accuracy.argtypes = [C.c_long, _dp, _ip, _ip, _dp, _dp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.oracle_accuracy.restype = C.c_double
        L.oracle_memory_footprint.argtypes = [C.c_long, C.c_long]
        L.oracle_memory_footprint.restype = C.c_long
        L.oracle_cg_solve.argtypes = [C.c_long, _dp, _ip, _ip, _dp, _dp, C.c_double, C.c_long,
                                      C.POINTER(C.c_double), C.c_void_p]
        L.oracle_cg_solve.restype = C.c_long
        L.oracle_mm_open.argtypes = [C.c_char_p, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.oracle_mm_open.restype = C.c_void_p
        L.oracle_mm_copy.argtypes = [C.c_void_p, _dp, _ip, _ip]
        L.oracle_mm_close.argtypes = [C.c_void_p]

    def spmv(self, vals, cols, rows, v):
        n = len(rows) - 1
        out = np.empty(n)
        self.lib.oracle_spmv(n, _f64(vals), _i32(cols), _i32(rows), _f64(v), out)
        return out

    def dot_acc(self, x, y, initial=0.0):
        r = C.c_double(initial)
        self.lib.oracle_dot_acc(len(x), _f64(x), _f64(y), C.byref(r))
        return r.value

    def norm_acc(self, x, initial=0.0):
        r = C.c_double(initial)
        self.lib.oracle_norm_acc(len(x), _f64(x), C.byref(r))
        return r.value

    def saxpby(self, x, y, a, b):
        out = np.empty(len(x))
        self.lib.oracle_saxpby(len(x), _f64(x), _f64(y), a, b, out)
        return out

    def sapbx(self, x, y, b):
        out = np.empty(len(x))
        self.lib.oracle_sapbx(len(x), _f64(x), _f64(y), b, out)
        return out

    def sambx(self, x, y, b):
        out = np.empty(len(x))
        self.lib.oracle_sambx(len(x), _f64(x), _f64(y), b, out)
        return out

    def accuracy(self, vals, cols, rows, x, b):
        n = len(rows) - 1
        r2, x2 = C.c_double(0), C.c_double(0)
        acc = self.lib.oracle_accuracy(n, _f64(vals), _i32(cols), _i32(rows), _f64(x), _f64(b), C.byref(r2), C.byref(x2))
        return acc, r2.value, x2.value

    def memory_footprint(self, n, nnz):
        return self.lib.oracle_memory_footprint(n, nnz)

    def cg_solve(self, vals, cols, rows, b, tol, x0=None, max_iter=-1, history=False):
        """Returns dict(x, iterations, rxr, hist)."""
        n = len(rows) - 1
        x = np.zeros(n) if x0 is None else _f64(x0).copy()
        cap = (n + 1) if max_iter < 0 else max_iter
        hist = np.zeros(cap + 1) if history else None
        rxr = C.c_double(0)
        it = self.lib.oracle_cg_solve(n, _f64(vals), _i32(cols), _i32(rows), _f64(b), x, float(tol), int(max_iter),
                                      C.byref(rxr), hist.ctypes.data if history else None)
        return dict(x=x, iterations=int(it), rxr=rxr.value, hist=None if hist is None else hist[: it + 1])

    def mm_read(self, path):
        n, nnz = C.c_long(0), C.c_long(0)
        h = self.lib.oracle_mm_open(os.fsencode(path), C.byref(n), C.byref(nnz))
        if not h:
            raise FileNotFoundError(path)
        vals, cols, rows = np.empty(nnz.value), np.empty(nnz.value, np.int32), np.empty(n.value + 1, np.int32)
        self.lib.oracle_mm_copy(h, vals, cols, rows)
        self.lib.oracle_mm_close(h)
        return vals, cols, rows


class Ref:
    """The reference's own sources (oracle/ref_bridge.cpp), serial or OpenMP flavour."""

    def __init__(self, flavour="serial"):
        path = os.path.join(_HERE, "_ref", f"libref_{flavour}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.flavour = flavour
        L = self.lib = C.CDLL(path)
        L.ref_cg_solve.argtypes = [C.c_long, C.c_long, _dp, _ip, _ip, _dp, C.c_void_p, C.c_double, _dp,
                                   C.POINTER(C.c_long), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                   C.POINTER(C.c_long)]
        L.ref_set_trip_limit.argtypes = [C.c_long]
        L.ref_cg_create.argtypes = [C.c_long, C.c_long, _dp, _ip, _ip]
        L.ref_cg_create.restype = C.c_void_p
        L.ref_cg_run.argtypes = [C.c_void_p, _dp, C.c_double, C.c_long, C.c_void_p, C.POINTER(C.c_long),
                                 C.POINTER(C.c_double)]
        L.ref_cg_destroy.argtypes = [C.c_void_p]
        L.ref_cg_solve_missing.argtypes = [C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.ref_reduce.argtypes = [C.c_int, C.c_long, _dp, C.c_void_p, C.c_double]
        L.ref_reduce.restype = C.c_double
        L.ref_axpy.argtypes = [C.c_int, C.c_int, C.c_long, _dp, _dp, C.c_double, C.c_double, _dp]
        L.ref_spmv.argtypes = [C.c_long, C.c_long, _dp, _ip, _ip, _dp, _dp]
        L.ref_mm_open.argtypes = [C.c_char_p, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.ref_mm_open.restype = C.c_void_p
        L.ref_mm_copy.argtypes = [C.c_void_p, _dp, _ip, _ip]
        L.ref_mm_close.argtypes = [C.c_void_p]

    @property
    def threads(self):
        return int(self.lib.ref_num_threads())

    def set_threads(self, t):
        self.lib.ref_set_threads(int(t))

    def cg_solve(self, vals, cols, rows, b, tol, x0=None, trip_limit=0):
        """trip_limit > 0 (bench timing samples only): leave the reference loop after that many trips."""
        n = len(rows) - 1
        self.lib.ref_set_trip_limit(int(trip_limit))
        x = np.empty(n)
        it, foot = C.c_long(0), C.c_long(0)
        acc, ms = C.c_double(0), C.c_double(0)
        x0a = None if x0 is None else _f64(x0)
        rc = self.lib.ref_cg_solve(n, len(vals), _f64(vals), _i32(cols), _i32(rows), _f64(b),
                                   None if x0a is None else x0a.ctypes.data, float(tol), x,
                                   C.byref(it), C.byref(acc), C.byref(ms), C.byref(foot))
        if rc != 0:
            raise RuntimeError(f"ref_cg_solve rc={rc}")
        return dict(x=x, iterations=int(it.value), accuracy=acc.value, solve_ms=ms.value, footprint=int(foot.value))

    def cg_create(self, vals, cols, rows):
        """Persistent reference solver (matrix set once) for timing loops; free with cg_destroy."""
        return self.lib.ref_cg_create(len(rows) - 1, len(vals), _f64(vals), _i32(cols), _i32(rows))

    def cg_run(self, handle, b, tol, trip_limit=0, want_x=False):
        b = _f64(b)
        x = np.empty(b.size) if want_x else None
        it, ms = C.c_long(0), C.c_double(0)
        rc = self.lib.ref_cg_run(handle, b, float(tol), int(trip_limit), None if x is None else x.ctypes.data,
                                 C.byref(it), C.byref(ms))
        if rc != 0:
            raise RuntimeError(f"ref_cg_run rc={rc}")
        return dict(x=x, iterations=int(it.value), solve_ms=ms.value)

    def cg_destroy(self, handle):
        self.lib.ref_cg_destroy(handle)

    def solve_missing(self, have_matrix, have_rhs):
        buf = C.create_string_buffer(256)
        threw = self.lib.ref_cg_solve_missing(int(have_matrix), int(have_rhs), buf, 256)
        return bool(threw), buf.value.decode()

    def reduce(self, which, x, y=None, initial=0.0):
        which = {"dot_product_trivial": 0, "dot_product_optimised": 1, "dot_product": 2, "norm": 3}[which]
        ya = None if y is None else _f64(y)
        return self.lib.ref_reduce(which, len(x), _f64(x), None if ya is None else ya.ctypes.data, float(initial))

    def axpy(self, which, x, y, a=0.0, b=0.0, alias=0):
        which = {"saxpby": 0, "sapbx": 1, "sambx": 2}[which]
        out = np.empty(len(x))
        self.lib.ref_axpy(which, alias, len(x), _f64(x), _f64(y), float(a), float(b), out)
        return out

    def spmv(self, vals, cols, rows, v):
        n = len(rows) - 1
        out = np.empty(n)
        self.lib.ref_spmv(n, len(vals), _f64(vals), _i32(cols), _i32(rows), _f64(v), out)
        return out

    def mm_read(self, path):
        n, nnz = C.c_long(0), C.c_long(0)
        h = self.lib.ref_mm_open(os.fsencode(path), C.byref(n), C.byref(nnz))
        vals, cols, rows = np.empty(nnz.value), np.empty(nnz.value, np.int32), np.empty(n.value + 1, np.int32)
        self.lib.ref_mm_copy(h, vals, cols, rows)
        self.lib.ref_mm_close(h)
        return vals, cols, rows


def have_ref(flavour="serial") -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", f"libref_{flavour}.so"))
