"""cgb_cg_solve_static: the CG solve for hosts that must not allocate device memory.

Some users of libcgb200.so cannot let it call cudaMalloc: applications that give (almost) all of the GPU's memory
to their own pool when they start, so a later allocation from a plugin fails; time-step loops that cannot take the
device-wide synchronisation cudaMalloc / cudaFree bring; deployments that forbid driver allocations once
initialisation is over.  For them the library carries a fixed work area (a __device__ array, CGB_STATIC_MAX_ROWS
rows) and this entry point, which takes only the caller's own device arrays.

The test calls it through the C-ABI and holds its x, trip count and final r.r against the existing entry point
(cgb_cg_solve, one-kernel loop) on the same device arrays: two calls in a row on different systems, so that what
the first call left in the shared work area must not leak into the second, then the first system again, which must
reproduce its own earlier result bit for bit.  The harness (not the entry point under test) allocates the inputs
and outputs, as a caller with its own pool would.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the
parity tests.  It passes both on its own in a fresh process (the first call then performs the per-device set-up)
and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err
from conjugategradient_b200 import _lib as L
from conjugategradient_b200 import generators as G
from conjugategradient_b200 import solver as S

pytestmark = pytest.mark.gpu

X_RTOL = 1e-10          # the solution tolerance of tests/test_solve_gpu.py (BASELINE.json north_star)
STATIC_MAX_ROWS = 131072  # CGB_STATIC_MAX_ROWS in include/cgb200.h


def iters_close(a, b):
    return abs(a - b) <= max(1, int(round(0.02 * b)))


class System:
    """One linear system resident on the device: CSR handle (for the existing entry point) + b."""

    def __init__(self, ctx, gen, rel):
        v, c, r = gen()
        self.ctx, self.n, self.nnz = ctx, r.size - 1, v.size
        self.A = S.Matrix(ctx, v, c, r)
        bh = G.default_rhs(self.n)
        self.b = S.Vector(ctx, bh)
        self.tol = rel * float(np.linalg.norm(bh))
        lib = ctx.lib
        self.vals, self.cols, self.rowptr = lib.cgb_csr_vals(self.A.h), lib.cgb_csr_cols(self.A.h), lib.cgb_csr_rowptr(self.A.h)

    def solve_existing(self):
        lib, ctx = self.ctx.lib, self.ctx
        x = S.Vector(ctx, n=self.n)
        opts, st = L.CgOpts(), L.CgStats()
        lib.cgb_cg_default_opts(C.byref(opts))
        opts.loop_mode = L.LOOP_PERSISTENT
        L.check(lib.cgb_cg_solve(ctx.h, self.A.h, self.b.ptr, x.ptr, self.tol, -1, C.byref(opts), C.byref(st)))
        assert st.loop_mode == L.LOOP_PERSISTENT
        return x.download(), st.as_dict()

    def solve_static(self, stream):
        lib, ctx = self.ctx.lib, self.ctx
        x = S.Vector(ctx, n=self.n)  # zero initial guess, uploaded before the call
        ctx.sync()
        st = L.CgStats()
        L.check(lib.cgb_cg_solve_static(stream, self.n, self.nnz, self.vals, self.cols, self.rowptr, self.b.ptr, x.ptr,
                                        self.tol, -1, C.byref(st)))
        return x.download(), st.as_dict()


def same_answer(got, want):
    (xg, sg), (xw, sw) = got, want
    assert sg["converged"] == 1 and sg["nan_stop"] == 0 and sw["converged"] == 1
    assert iters_close(sg["iterations"], sw["iterations"]), (sg["iterations"], sw["iterations"])
    assert rel_err(xg, xw) <= X_RTOL
    assert abs(sg["rr0"] - sw["rr0"]) <= 1e-12 * sw["rr0"]
    assert sg["kernel_launches"] == 2 and sg["loop_mode"] == L.LOOP_PERSISTENT


def test_static_solve_matches_the_allocating_solve_twice_in_a_row(ctx):
    big = System(ctx, lambda: G.poisson3d_7pt(40), 1e-8)     # 64,000 rows
    small = System(ctx, lambda: G.poisson2d_5pt(128), 1e-8)  # 16,384 rows (BASELINE config C1)
    want_big, want_small = big.solve_existing(), small.solve_existing()
    first = big.solve_static(ctx.stream)     # call 1 (in a fresh process: also the per-device set-up)
    second = small.solve_static(ctx.stream)  # call 2: a shorter system over the first call's leftovers
    same_answer(first, want_big)
    same_answer(second, want_small)
    third = big.solve_static(None)           # default stream; must reproduce call 1 exactly
    assert third[1]["iterations"] == first[1]["iterations"] and third[1]["rr"] == first[1]["rr"]
    assert np.array_equal(third[0], first[0])


def test_static_solve_refuses_what_does_not_fit(ctx):
    lib = ctx.lib
    d = S.Vector(ctx, n=4)
    st = L.CgStats()
    n = STATIC_MAX_ROWS + 1
    rc = lib.cgb_cg_solve_static(None, n, 0, d.ptr, d.ptr, d.ptr, d.ptr, d.ptr, 1e-8, -1, C.byref(st))
    assert rc == L.CGB_ERR_UNSUPPORTED
    assert "CGB_STATIC_MAX_ROWS" in lib.cgb_last_error().decode()
    # a missing matrix / right-hand side is reported the way cgb_cg_solve reports it (src/CG.hpp:266-272)
    assert lib.cgb_cg_solve_static(None, 4, 0, d.ptr, d.ptr, None, d.ptr, d.ptr, 1e-8, -1, None) == L.CGB_ERR_STATE
    assert lib.cgb_cg_solve_static(None, 4, 0, d.ptr, d.ptr, d.ptr, None, d.ptr, 1e-8, -1, None) == L.CGB_ERR_STATE
