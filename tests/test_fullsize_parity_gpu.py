"""Parity at the BASELINE sizes, against the reference itself.

The reference pins nothing of its own (no golden vectors, no asserting tests: CMakeLists.txt:46, test/Tester.cpp:54), so
the pin is the reference's sources run here (oracle/_ref, built from /root/reference by oracle/Makefile; the OpenMP
flavour so the full-size systems finish in seconds) on exactly the matrices the GPU solves:

  C2  3D 7-pt 128^3     full solve: x <= 1e-10, trips +-2 %, r.r history (C restatement, bit-pinned to the reference)
  C3  3D 27-pt 256^3    full solve: x <= 1e-10, trips +-2 %, r.r history of the first trips
  C4  3D 7-pt 512^3     the first K trips (the reference loop left after K trips, oracle/ref_bridge.cpp ref_cg_run):
  C5  random SPD 10^7   x after K trips <= 1e-10, true residual norms agree

The engines checked are the ones the bench uses at these sizes: TMA-staged CSR tiles on 444 persistent CTAs, SELL-32
with two column panels, CUDA-graph batches of 32 trips (K = 48 crosses a batch boundary and overshoots the second).
Protocol as test/Tester.cpp:27-46: b = 1..n, x0 = 0."""
import numpy as np
import pytest

from conftest import rel_err
from conjugategradient_b200 import _lib as L
from conjugategradient_b200 import generators as G
from conjugategradient_b200 import solver as S

pytestmark = pytest.mark.gpu
X_RTOL = 1e-10      # north_star: solution within 1e-10 relative
K = 48


@pytest.fixture(scope="module")
def ref_omp():
    import oracle
    if not oracle.have_ref("omp"):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return oracle.Ref("omp")


def iters_close(a, b):
    return abs(a - b) <= max(1, int(round(0.02 * b)))


def gpu_solve(ctx, A, b, tol, max_iter=-1, **opts):
    cg = S.CG(ctx, **opts)
    cg.setMatrix(A)
    cg.setTarget(b)
    st = cg.solve(tol, max_iter=max_iter)
    return cg, st


def test_C2_full_solve_vs_reference(ctx, ref_omp, port):
    m = 128
    A = S.Matrix.stencil(ctx, "3d7", m)
    v, c, r = A.to_host()
    hv, hc, hr = G.poisson3d_7pt(m)
    assert np.array_equal(v, hv) and np.array_equal(c, hc) and np.array_equal(r, hr)    # device generator == host generator
    n = r.size - 1
    b = G.default_rhs(n)
    tol = 1e-8 * np.linalg.norm(b)
    cg, st = gpu_solve(ctx, A, b, tol, record_history=1)
    assert st["loop_mode"] == L.LOOP_GRAPH and st["format"] == L.FMT_CSR
    x = cg.extract()
    o = ref_omp.cg_solve(v, c, r, b, tol)                       # the reference's CG::solve, 442 trips
    assert iters_close(st["iterations"], o["iterations"]) and st["converged"] == 1
    assert rel_err(x, o["x"]) <= X_RTOL
    p = port.cg_solve(v, c, r, b, tol, history=True)            # serial restatement (bit-pinned to the reference): history
    assert p["iterations"] == o["iterations"] or iters_close(p["iterations"], o["iterations"])
    assert rel_err(x, p["x"]) <= X_RTOL
    h = cg.history()
    k = min(h.size, p["hist"].size, 200)
    assert np.allclose(h[:k], p["hist"][:k], rtol=1e-8, atol=0)
    # final relative residuals agree to 1e-10 (true residual, computed by the oracle for both)
    nb = np.linalg.norm(b)
    res_g = np.linalg.norm(b - port.spmv(v, c, r, x)) / nb
    res_o = np.linalg.norm(b - port.spmv(v, c, r, o["x"])) / nb
    assert abs(res_g - res_o) <= 1e-10


def test_C3_full_solve_vs_reference(ctx, ref_omp, port):
    m = 256
    A = S.Matrix.stencil(ctx, "3d27", m)
    v, c, r = A.to_host()
    n = r.size - 1
    assert n == m ** 3 and v.size == G.stencil_nnz("3d27", m)
    b = G.default_rhs(n)
    nb = np.linalg.norm(b)
    tol = 1e-8 * nb
    cg, st = gpu_solve(ctx, A, b, tol, record_history=1)
    x = cg.extract()
    o = ref_omp.cg_solve(v, c, r, b, tol)                       # ~30 s on 16 host threads
    assert iters_close(st["iterations"], o["iterations"]) and st["converged"] == 1
    assert rel_err(x, o["x"]) <= X_RTOL
    kk = 12
    p = port.cg_solve(v, c, r, b, 0.0, max_iter=kk, history=True)
    assert np.allclose(cg.history()[:kk + 1], p["hist"][:kk + 1], rtol=1e-8, atol=0)
    res_g = np.linalg.norm(b - port.spmv(v, c, r, x)) / nb
    res_o = np.linalg.norm(b - port.spmv(v, c, r, o["x"])) / nb
    assert abs(res_g - res_o) <= 1e-10


def first_k_vs_reference(ctx, ref, A, k):
    """x after k trips: GPU (max_iter = k) against the reference's solve() left after k trips, on A as the device holds it."""
    v, c, r = A.to_host()
    n = r.size - 1
    b = G.default_rhs(n)
    nb = float(np.sqrt(n * (n + 1) * (2 * n + 1) / 6.0))
    h = ref.cg_create(v, c, r)
    del v, c, r
    o = ref.cg_run(h, b, 0.0, trip_limit=k, want_x=True)
    ref.cg_destroy(h)
    assert o["iterations"] == k
    cg, st = gpu_solve(ctx, A, b, 0.0, max_iter=k)
    assert st["iterations"] == k and st["converged"] == 0 and st["nan_stop"] == 0
    x = cg.extract()
    err = rel_err(x, o["x"])
    # true residual norms of the two iterates, both formed by the GPU's residual kernel (cgb_residual)
    rr_g = cg.residual()[0]
    cg.x.upload(o["x"])
    rr_o = cg.residual()[0]
    return st, err, np.sqrt(rr_g) / nb, np.sqrt(rr_o) / nb


def test_C4_first_trips_vs_reference(ctx, ref_omp):
    """The configuration the metric is quoted on: 134,217,728 rows, TMA tile engine, 444 persistent CTAs."""
    A = S.Matrix.stencil(ctx, "3d7", 512)
    assert A.N() == 512 ** 3 and A.NNZ() == 937951232
    st, err, res_g, res_o = first_k_vs_reference(ctx, ref_omp, A, K)
    assert st["loop_mode"] == L.LOOP_GRAPH and st["format"] == L.FMT_CSR
    assert err <= X_RTOL, err
    assert abs(res_g - res_o) <= 1e-10 * max(res_o, 1.0)


def test_C5_first_trips_vs_reference(ctx, ref_omp):
    """Random SPD with power-law rows, n = 10^7: SELL-32-sigma engine with column panels (CGB_FMT_AUTO's choice)."""
    A = S.Matrix.random_spd(ctx, 10_000_000)
    st, err, res_g, res_o = first_k_vs_reference(ctx, ref_omp, A, K)
    assert st["format"] == L.FMT_SELL and A.panels() >= 2
    assert err <= X_RTOL, err
    assert abs(res_g - res_o) <= 1e-10 * max(res_o, 1.0)
