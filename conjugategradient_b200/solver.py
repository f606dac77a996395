"""Python harness over the C-ABI, mirroring the reference's C++ API names so tests read like the
reference's driver (reference test/Tester.cpp:36-53, src/CG.hpp:53-601, src/VectorOperations.hpp).

Nothing here computes: every method forwards to libcgb200.so (CUDA, sm_100a).  NumPy arrays are
host buffers only.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """Device + stream (reference: acpp::sycl::queue, src/CG.hpp:61,70-77)."""

    def __init__(self, device: int = 0):
        self.lib = L.load()
        h = C.c_void_p()
        L.check(self.lib.cgb_ctx_create(device, C.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            self.lib.cgb_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        L.check(self.lib.cgb_ctx_sync(self.h))

    @property
    def stream(self) -> int:
        return int(self.lib.cgb_ctx_stream(self.h) or 0)

    @property
    def sm_count(self) -> int:
        return self.lib.cgb_ctx_sm_count(self.h)

    # -- raw buffers -------------------------------------------------------------------------
    def malloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        L.check(self.lib.cgb_malloc(self.h, nbytes, C.byref(p)))
        return p.value

    def free(self, ptr):
        if ptr and self.h:  # after close() there is no stream to order the free on (the driver reclaims at exit)
            L.check(self.lib.cgb_free(self.h, ptr))


class Vector:
    """Device vector of doubles (reference: CGSolver::Vector, src/LinearAlgebraTypes.hpp:143-203)."""

    def __init__(self, ctx: Context, data=None, n: int | None = None):
        self.ctx = ctx
        if data is not None:
            data = _f64(data)
            n = data.size
        self.n = int(n)
        self.ptr = ctx.malloc(8 * max(self.n, 1))
        if data is not None:
            self.upload(data)
        else:
            L.check(ctx.lib.cgb_memset(ctx.h, self.ptr, 0, 8 * self.n))

    def upload(self, data):
        data = _f64(data)
        assert data.size == self.n
        L.check(self.ctx.lib.cgb_memcpy_h2d(self.ctx.h, self.ptr, _p(data), 8 * self.n, 1))

    def download(self) -> np.ndarray:
        out = np.empty(self.n)
        L.check(self.ctx.lib.cgb_memcpy_d2h(self.ctx.h, _p(out), self.ptr, 8 * self.n, 1))
        return out

    def free(self):
        if self.ptr:
            self.ctx.free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Scalar(Vector):
    """One device scalar (reference: CGSolver::Scalar, src/LinearAlgebraTypes.hpp:210-250)."""

    def __init__(self, ctx: Context, value: float = 0.0):
        super().__init__(ctx, data=np.array([value]))

    @property
    def value(self) -> float:
        v = C.c_double()
        L.check(self.ctx.lib.cgb_scalar_get(self.ctx.h, self.ptr, C.byref(v)))
        return v.value

    def set(self, value: float):
        L.check(self.ctx.lib.cgb_scalar_set(self.ctx.h, self.ptr, float(value)))


class Matrix:
    """Device CSR matrix (reference: CGSolver::Matrix, src/LinearAlgebraTypes.hpp:57-132)."""

    def __init__(self, ctx: Context, data=None, cols=None, rows=None, handle=None):
        self.ctx = ctx
        if handle is None:
            data, cols, rows = _f64(data), _i32(cols), _i32(rows)
            h = C.c_void_p()
            L.check(ctx.lib.cgb_csr_create(ctx.h, rows.size - 1, data.size, _p(data), _p(cols), _p(rows), C.byref(h)))
            handle = h
        self.h = handle

    @classmethod
    def stencil(cls, ctx: Context, kind: str, m: int, mz: int = 0, row_begin: int = 0, row_end: int = -1):
        k = {"2d5": 0, "3d7": 1, "3d27": 2}[kind]
        h = C.c_void_p()
        L.check(ctx.lib.cgb_gen_stencil(ctx.h, k, m, mz, row_begin, row_end, C.byref(h)))
        return cls(ctx, handle=h)

    @classmethod
    def from_triplets(cls, ctx: Context, rows, cols, vals, n: int = -1, flags: int = 0):
        """COO -> CSR on the GPU (cgb_csr_from_triplets; replaces the reference reader's host assembly,
        test/mm_reader.cpp:45-107, and its HiFlow3 bridge, src/CG.hpp:104-151)."""
        rows, cols, vals = _i32(rows), _i32(cols), _f64(vals)
        h = C.c_void_p()
        L.check(ctx.lib.cgb_csr_from_triplets(ctx.h, n, vals.size, _p(rows), _p(cols), _p(vals), flags, C.byref(h)))
        return cls(ctx, handle=h)

    @classmethod
    def from_mm_file(cls, ctx: Context, path: str):
        """Matrix Market file -> device CSR (parallel parse on the host, assembly on the GPU)."""
        h = C.c_void_p()
        L.check(ctx.lib.cgb_csr_from_mm_file(ctx.h, path.encode(), C.byref(h)))
        return cls(ctx, handle=h)

    @classmethod
    def random_spd(cls, ctx: Context, n: int, seed: int = 12345, d_min: int = 2, d_max: int = 4096, gamma: float = 2.2,
                   row_begin: int = 0, row_end: int = -1):
        """Random SPD matrix with power-law row lengths generated on the device (config C5)."""
        h = C.c_void_p()
        L.check(ctx.lib.cgb_gen_random_spd(ctx.h, n, seed, d_min, d_max, gamma, row_begin, row_end, C.byref(h)))
        return cls(ctx, handle=h)

    def N(self) -> int:
        return int(self.ctx.lib.cgb_csr_n(self.h))

    def NNZ(self) -> int:
        return int(self.ctx.lib.cgb_csr_nnz(self.h))

    def to_host(self):
        n, nnz = self.N(), self.NNZ()
        vals, cols, rows = np.empty(nnz), np.empty(nnz, np.int32), np.empty(n + 1, np.int32)
        lib, c = self.ctx.lib, self.ctx.h
        L.check(lib.cgb_memcpy_d2h(c, _p(vals), lib.cgb_csr_vals(self.h), 8 * nnz, 1))
        L.check(lib.cgb_memcpy_d2h(c, _p(cols), lib.cgb_csr_cols(self.h), 4 * nnz, 1))
        L.check(lib.cgb_memcpy_d2h(c, _p(rows), lib.cgb_csr_rowptr(self.h), 4 * (n + 1), 1))
        return vals, cols, rows

    def set_format(self, fmt: int, sigma: int | None = None):
        """sigma: SELL sorting window (None: keep the copy already built, or the library default)."""
        if fmt == L.FMT_SELL and sigma is not None:
            L.check(self.ctx.lib.cgb_csr_build_sell(self.h, sigma))
        L.check(self.ctx.lib.cgb_csr_set_format(self.h, fmt))

    def set_panels(self, panels: int):
        """Column panels of the SELL engine: 0 automatic, 1 none, P exactly P (include/cgb200.h)."""
        L.check(self.ctx.lib.cgb_csr_set_panels(self.h, panels))

    def panels(self) -> int:
        return int(self.ctx.lib.cgb_csr_get_panels(self.h))

    def dict_bytes(self) -> int:
        """Size of the row-pattern dictionary copy (CGB_FMT_DICT), 0 when not built."""
        return int(self.ctx.lib.cgb_csr_dict_bytes(self.h))

    def dict_patterns(self) -> int:
        return int(self.ctx.lib.cgb_csr_dict_patterns(self.h))

    def sell_padded(self) -> int:
        return int(self.ctx.lib.cgb_csr_sell_padded(self.h))

    def free(self):
        if self.h and self.ctx.h:  # a handle that outlives its context (closed first) must not touch it
            self.ctx.lib.cgb_csr_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class VectorOperations:
    """Per-op kernels (reference: CGSolver::VectorOperations, src/VectorOperations.hpp:39-488)."""

    def __init__(self, ctx: Context):
        self.ctx = ctx
        self.vector_size = 0

    def setVectorSize(self, n: int):
        self.vector_size = int(n)

    def spmv(self, A: Matrix, vec: Vector, result: Vector):
        assert self.vector_size != 0 and A.N() == self.vector_size  # reference :444
        L.check(self.ctx.lib.cgb_spmv(self.ctx.h, A.h, vec.ptr, result.ptr))

    def dot_product_trivial(self, left: Vector, right: Vector, result: Scalar):
        L.check(self.ctx.lib.cgb_dot_acc(self.ctx.h, self.vector_size, left.ptr, right.ptr, result.ptr))

    dot_product_optimised = dot_product_trivial
    dot_product = dot_product_trivial

    def norm(self, vec: Vector, result: Scalar):
        L.check(self.ctx.lib.cgb_norm2_acc(self.ctx.h, self.vector_size, vec.ptr, result.ptr))

    def saxpby(self, X: Vector, Y: Vector, a: Scalar, b: Scalar, result: Vector):
        L.check(self.ctx.lib.cgb_axpby(self.ctx.h, self.vector_size, X.ptr, Y.ptr, a.ptr, b.ptr, result.ptr))

    def sapbx(self, X: Vector, Y: Vector, b: Scalar, result: Vector):
        L.check(self.ctx.lib.cgb_xpby(self.ctx.h, self.vector_size, X.ptr, Y.ptr, b.ptr, 1, result.ptr))

    def sambx(self, X: Vector, Y: Vector, b: Scalar, result: Vector):
        L.check(self.ctx.lib.cgb_xpby(self.ctx.h, self.vector_size, X.ptr, Y.ptr, b.ptr, -1, result.ptr))


def make_opts(**kw) -> L.CgOpts:
    o = L.CgOpts()
    L.load().cgb_cg_default_opts(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError(f"unknown CG option {k}")
        setattr(o, k, int(v))
    return o


class CG:
    """Reference-shaped solver object (reference: CGSolver::CG, src/CG.hpp:53-601)."""

    def __init__(self, ctx: Context | None = None, **opts):
        self.ctx = ctx or Context(0)
        self.A: Matrix | None = None
        self.b: Vector | None = None
        self.x: Vector | None = None
        self.opts = opts
        self.stats: dict | None = None

    @classmethod
    def createCG(cls, **opts):
        return cls(Context(0), **opts)

    def setMatrix(self, data, columns=None, rows=None):
        self.A = data if isinstance(data, Matrix) else Matrix(self.ctx, data, columns, rows)

    def getDimension(self) -> int:
        return self.A.N() if self.A is not None else 0

    def setTarget(self, data):
        self.b = data if isinstance(data, Vector) else Vector(self.ctx, data)

    def setInital(self, data):  # sic — the reference spells it this way (src/CG.hpp:215)
        self.x = data if isinstance(data, Vector) else Vector(self.ctx, data)

    setInitial = setInital

    def calculateExpectedStepCount(self, accuracy):  # no-op in the reference too (src/CG.hpp:235)
        return None

    def solve(self, improvement: float = 0.0, max_iter: int = -1):
        if self.b is None:
            raise RuntimeError("No right hand side to solve for")  # src/CG.hpp:266-268
        if self.A is None:
            raise RuntimeError("No Matrix given")  # src/CG.hpp:270-272
        if self.x is None:
            self.x = Vector(self.ctx, n=self.A.N())  # zeros (src/CG.hpp:291-297)
        o = make_opts(**self.opts)
        st = L.CgStats()
        L.check(self.ctx.lib.cgb_cg_solve(self.ctx.h, self.A.h, self.b.ptr, self.x.ptr, float(improvement),
                                          int(max_iter), C.byref(o), C.byref(st)))
        self.stats = st.as_dict()
        return self.stats

    def history(self) -> np.ndarray:
        cnt = C.c_int64()
        L.check(self.ctx.lib.cgb_cg_history(self.ctx.h, self.A.h, None, 0, C.byref(cnt)))
        out = np.empty(cnt.value)
        if cnt.value:
            L.check(self.ctx.lib.cgb_cg_history(self.ctx.h, self.A.h, _p(out), cnt.value, C.byref(cnt)))
        return out

    def accuracy(self) -> float:
        rr, xx = self.residual()
        return abs(rr / xx) if xx != 0 else abs(np.float64(rr) / np.float64(xx))

    def residual(self):
        rr, xx = C.c_double(), C.c_double()
        L.check(self.ctx.lib.cgb_residual(self.ctx.h, self.A.h, self.x.ptr, self.b.ptr, C.byref(rr), C.byref(xx)))
        return rr.value, xx.value

    def extract(self) -> np.ndarray:
        return self.x.download()

    def extractTo(self, out: np.ndarray):
        out[...] = self.x.download()

    def memoryFootprint(self) -> int:
        return int(self.ctx.lib.cgb_memory_footprint(self.A.N(), self.A.NNZ()))

    # additive accessors (the reference returns none of these)
    def iterations(self) -> int:
        return int(self.stats["iterations"]) if self.stats else 0

    def residualNorm(self) -> float:
        return float(np.sqrt(self.stats["rr"])) if self.stats else float("nan")


def solve_host(ctx: Context, vals, cols, rows, b, tol, x0=None, max_iter=-1, **opts):
    """Host-buffer solve through cgb_cg_solve_host (the end-to-end path)."""
    vals, cols, rows, b = _f64(vals), _i32(cols), _i32(rows), _f64(b)
    n = rows.size - 1
    x = np.empty(n)
    x0a = None if x0 is None else _f64(x0)
    o = make_opts(**opts)
    st = L.CgStats()
    L.check(ctx.lib.cgb_cg_solve_host(ctx.h, n, vals.size, _p(vals), _p(cols), _p(rows), _p(b),
                                      None if x0a is None else _p(x0a), _p(x), float(tol), int(max_iter),
                                      C.byref(o), C.byref(st)))
    return x, st.as_dict()


def mm_read(path: str):
    """Matrix Market -> CSR through the library's dialect-compatible reader (test/mm_reader.cpp:154)."""
    import os
    lib = L.load()
    h = C.c_void_p()
    L.check(lib.cgb_mm_read(os.fsencode(path), C.byref(h)))
    try:
        n, nnz = lib.cgb_mm_n(h), lib.cgb_mm_nnz(h)
        vals = np.ctypeslib.as_array(C.cast(lib.cgb_mm_vals(h), L.c_dp), shape=(nnz,)).copy()
        cols = np.ctypeslib.as_array(C.cast(lib.cgb_mm_cols(h), L.c_ip), shape=(nnz,)).copy()
        rows = np.ctypeslib.as_array(C.cast(lib.cgb_mm_rowptr(h), L.c_ip), shape=(n + 1,)).copy()
    finally:
        lib.cgb_mm_free(h)
    return vals, cols, rows


# ---- row-partitioned multi-GPU (one process per GPU; new) ---------------------------
def partition_rows(n: int, world: int, align: int = 1, rowptr=None, edge_extra: int = 0) -> np.ndarray:
    """world+1 global row offsets of contiguous row blocks.  Balanced by rows (multiples of `align`,
    e.g. one grid plane for a stencil) or, when `rowptr` is given, by non-zeros.

    edge_extra (rows; world >= 3 only): the first and the last rank get this many more rows than the others.  In a
    slab decomposition of a stencil they have one neighbour instead of two (one halo plane to pack, send and apply),
    finish each trip early and wait in the all-reduce; the extra rows even the trip time out."""
    if rowptr is None and edge_extra > 0 and world >= 3:
        units, extra = n // align, edge_extra // align
        inner = max((units - 2 * extra) // world, 0)
        sizes = [inner + extra] + [inner] * (world - 2) + [inner + extra]
        left = units - sum(sizes)           # hand the remainder out from the middle
        for k in range(left):
            sizes[1 + (k % (world - 2))] += 1
        starts = align * np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        starts[-1] = n
        return starts
    if rowptr is not None:
        rp = np.asarray(rowptr, dtype=np.int64)
        targets = rp[-1] * np.arange(1, world, dtype=np.float64) / world
        cuts = np.searchsorted(rp, targets, side="left")
        starts = np.concatenate([[0], cuts, [n]]).astype(np.int64)
        return np.maximum.accumulate(starts)
    units = n // align
    starts = np.array([align * ((units * r) // world) for r in range(world + 1)], dtype=np.int64)
    starts[-1] = n
    return starts


def local_block(vals, cols, rows, lo: int, hi: int):
    """Rows [lo, hi) of a host CSR matrix with GLOBAL column indices (what cgb_dist_create takes)."""
    rows = np.asarray(rows)
    a, b = int(rows[lo]), int(rows[hi])
    return _f64(vals[a:b]), _i32(cols[a:b]), _i32(rows[lo:hi + 1] - rows[lo])


def broadcast_nccl_id(lib, rank: int, group=None) -> bytes:
    """ncclUniqueId created on rank 0 and broadcast through torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist
    buf = (C.c_ubyte * L.NCCL_ID_BYTES)()
    if rank == 0:
        L.check(lib.cgb_nccl_unique_id(buf))
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
    if dist.get_backend(group) == "nccl":
        t = t.cuda()
    dist.broadcast(t, 0, group=group)
    return bytes(t.cpu().tolist())


class DistCG:
    """CG over a row-partitioned matrix: this rank owns global rows [row_starts[rank], row_starts[rank+1]).
    Halo entries of p travel over NVLink, the two scalars per iteration are all-reduced (cgb_dist_*)."""

    def __init__(self, ctx: Context, rank: int, world: int, nccl_id: bytes, row_starts, A_local: Matrix, **opts):
        self.ctx, self.rank, self.world = ctx, rank, world
        self.row_starts = np.ascontiguousarray(row_starts, dtype=np.int64)
        self.A = A_local
        self.opts = opts
        self.stats = None
        h = C.c_void_p()
        L.check(ctx.lib.cgb_dist_create(ctx.h, rank, world, nccl_id, _p(self.row_starts), A_local.h, C.byref(h)))
        self.h = h
        self.n_local = int(self.row_starts[rank + 1] - self.row_starts[rank])

    def set_matrix(self, row_starts, A_local: Matrix):
        self.row_starts = np.ascontiguousarray(row_starts, dtype=np.int64)
        self.A = A_local
        L.check(self.ctx.lib.cgb_dist_set_matrix(self.h, _p(self.row_starts), A_local.h))

    def halo_count(self) -> int:
        return int(self.ctx.lib.cgb_dist_halo_count(self.h))

    def solve(self, b: Vector, x: Vector, tol: float, max_iter: int = -1):
        o = make_opts(**self.opts)
        st = L.CgStats()
        L.check(self.ctx.lib.cgb_dist_cg_solve(self.h, b.ptr, x.ptr, float(tol), int(max_iter), C.byref(o), C.byref(st)))
        self.stats = st.as_dict()
        return self.stats

    def residual(self, x: Vector, b: Vector):
        rr, xx = C.c_double(), C.c_double()
        L.check(self.ctx.lib.cgb_dist_residual(self.h, x.ptr, b.ptr, C.byref(rr), C.byref(xx)))
        return rr.value, xx.value

    def close(self):
        if self.h:
            self.ctx.lib.cgb_dist_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
