"""SELL-32-sigma engine and the on-device random-SPD generator (BASELINE config C5 path).
SpMV / accuracy / whole solves through the SELL copy must agree with the oracle exactly like the
CSR engines do; the row sum order of SELL is the reference's own (sequential, left to right)."""
import numpy as np
import pytest

from conftest import rel_err
from conjugategradient_b200 import _lib as L
from conjugategradient_b200 import generators as G
from conjugategradient_b200 import solver as S

pytestmark = pytest.mark.gpu

MATS = {
    "p2d_33": lambda: G.poisson2d_5pt(33),
    "p3d27_9": lambda: G.stencil3d_27pt(9),
    "rspd_20000": lambda: G.random_spd_powerlaw(20000, seed=9),
    "rspd_777": lambda: G.random_spd_powerlaw(777, seed=2),     # n not a multiple of 32
    "diag_1": lambda: (np.array([2.0]), np.array([0], np.int32), np.array([0, 1], np.int32)),
}


@pytest.mark.parametrize("sigma", [1, 32, 4096])
@pytest.mark.parametrize("name", sorted(MATS))
def test_sell_spmv_and_accuracy(ctx, port, name, sigma):
    v, c, r = MATS[name]()
    n = r.size - 1
    rng = np.random.default_rng(11)
    z, b = rng.standard_normal(n), rng.standard_normal(n)
    A = S.Matrix(ctx, v, c, r)
    A.set_format(L.FMT_SELL, sigma)
    assert A.sell_padded() % 32 == 0
    ops = S.VectorOperations(ctx)
    ops.setVectorSize(n)
    Z, O = S.Vector(ctx, z), S.Vector(ctx, n=n)
    ops.spmv(A, Z, O)
    exp = port.spmv(v, c, r, z)
    absrow = port.spmv(np.abs(v), c, r, np.abs(z))
    # same summation order as the oracle, but fused multiply-add: within an ulp or two of the row's absolute sum
    assert np.all(np.abs(O.download() - exp) <= 4e-16 * absrow * 4)
    cg = S.CG(ctx)
    cg.setMatrix(A)
    cg.setTarget(b)
    cg.setInital(z)
    acc, rr, xx = port.accuracy(v, c, r, z, b)
    grr, gxx = cg.residual()
    assert abs(grr - rr) <= 1e-13 * rr and abs(gxx - xx) <= 1e-13 * xx


def test_sell_padding_shrinks_with_sigma(ctx):
    v, c, r = G.random_spd_powerlaw(50000, seed=5)
    A = S.Matrix(ctx, v, c, r)
    pads = []
    for sigma in (1, 256, 8192):
        A.set_format(L.FMT_CSR)
        L.check(ctx.lib.cgb_csr_build_sell(A.h, sigma))
        pads.append(A.sell_padded())
    assert pads[0] > pads[1] >= pads[2] > 0
    assert pads[2] <= 1.25 * v.size          # sorting inside windows removes most of the padding (long rows live outside)


@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("name,rel", [("C1_p2d_128", 1e-8), ("p3d7_30", 1e-8), ("rspd_50000", 1e-10)])
def test_sell_solve_vs_oracle(ctx, port, name, rel, fused):
    gen = {"C1_p2d_128": lambda: G.poisson2d_5pt(128), "p3d7_30": lambda: G.poisson3d_7pt(30),
           "rspd_50000": lambda: G.random_spd_powerlaw(50000, seed=21)}[name]
    v, c, r = gen()
    n = r.size - 1
    b = G.default_rhs(n)
    tol = rel * np.linalg.norm(b)
    o = port.cg_solve(v, c, r, b, tol)
    A = S.Matrix(ctx, v, c, r)
    A.set_format(L.FMT_SELL, 1024)
    cg = S.CG(ctx, fuse_p_update=fused)
    cg.setMatrix(A)
    cg.setTarget(b)
    st = cg.solve(tol)
    assert st["format"] == L.FMT_SELL and st["converged"] == 1
    slack = 0.05 if name.startswith("rspd") else 0.02
    assert abs(st["iterations"] - o["iterations"]) <= max(1, round(slack * o["iterations"]))
    assert rel_err(cg.extract(), o["x"]) <= 1e-10
    # switching back to CSR on the same handle gives the same answer (graphs are rebuilt)
    A.set_format(L.FMT_CSR)
    cg.x = None
    st2 = cg.solve(tol)
    assert st2["format"] == L.FMT_CSR and rel_err(cg.extract(), o["x"]) <= 1e-10


def test_random_spd_generator_properties(ctx, port):
    n = 30000
    A = S.Matrix.random_spd(ctx, n, seed=7)
    v, c, r = A.to_host()
    assert r.size == n + 1 and r[0] == 0 and r[-1] == v.size
    lens = np.diff(r)
    assert lens.min() >= 1 and lens.max() > 8 * lens.mean()          # power-law tail
    rows = np.repeat(np.arange(n), lens)
    assert np.all(np.diff(c.astype(np.int64) + rows * n) > 0)          # columns strictly ascending inside rows, no duplicates
    import scipy.sparse as sp
    M = sp.csr_matrix((v, c, r), shape=(n, n))
    assert abs(M - M.T).max() == 0.0                                   # exactly symmetric
    d = M.diagonal()
    off = np.asarray(abs(M).sum(axis=1)).ravel() - np.abs(d)
    assert np.all(d > 0) and np.allclose(d, 1.0 + off, rtol=1e-14)    # strictly diagonally dominant => SPD
    assert np.all(v[c != rows] < 0) and np.all(v[c != rows] >= -1)
    # deterministic, and a row block equals the slice of the whole matrix (global columns)
    B = S.Matrix.random_spd(ctx, n, seed=7)
    assert all(np.array_equal(a, b) for a, b in zip(B.to_host(), (v, c, r)))
    blk = S.Matrix.random_spd(ctx, n, seed=7, row_begin=1000, row_end=5000)
    bv, bc, br = blk.to_host()
    assert np.array_equal(br, r[1000:5001] - r[1000]) and np.array_equal(bc, c[r[1000]:r[5000]])
    assert np.array_equal(bv, v[r[1000]:r[5000]])
    C2 = S.Matrix.random_spd(ctx, n, seed=8)
    assert C2.NNZ() != A.NNZ() or not np.array_equal(C2.to_host()[1], c)
    # and CG on it converges like the oracle
    b = G.default_rhs(n)
    tol = 1e-10 * np.linalg.norm(b)
    o = port.cg_solve(v, c, r, b, tol)
    cg = S.CG(ctx)
    cg.setMatrix(A)
    cg.setTarget(b)
    st = cg.solve(tol)
    assert abs(st["iterations"] - o["iterations"]) <= max(1, round(0.05 * o["iterations"]))
    assert rel_err(cg.extract(), o["x"]) <= 1e-10


# ---- column panels: A = [A_0 | ... | A_{P-1}], one pass per panel accumulating into the output ------------
@pytest.mark.parametrize("panels", [2, 3, 7])
@pytest.mark.parametrize("name", ["p2d_33", "rspd_20000", "rspd_777", "diag_1"])
def test_sell_panels_spmv_and_accuracy(ctx, port, name, panels):
    v, c, r = MATS[name]()
    n = r.size - 1
    rng = np.random.default_rng(5)
    z, b = rng.standard_normal(n), rng.standard_normal(n)
    A = S.Matrix(ctx, v, c, r)
    A.set_panels(panels)
    A.set_format(L.FMT_SELL, 1024)
    assert A.panels() == panels
    ops = S.VectorOperations(ctx)
    ops.setVectorSize(n)
    Z, O = S.Vector(ctx, z), S.Vector(ctx, n=n)
    ops.spmv(A, Z, O)
    exp = port.spmv(v, c, r, z)
    absrow = port.spmv(np.abs(v), c, r, np.abs(z))
    # the row sum is (panel 0) + (panel 1) + ...: a different association, still a few ulp of the absolute sum
    assert np.all(np.abs(O.download() - exp) <= 4e-16 * absrow * (4 + panels))
    first = O.download().copy()
    ops.spmv(A, Z, O)
    assert np.array_equal(first, O.download())                        # bitwise reproducible
    cg = S.CG(ctx)
    cg.setMatrix(A)
    cg.setTarget(b)
    cg.setInital(z)
    acc, rr, xx = port.accuracy(v, c, r, z, b)
    grr, gxx = cg.residual()
    assert abs(grr - rr) <= 1e-13 * rr and abs(gxx - xx) <= 1e-13 * xx
    # back to one piece on the same handle
    A.set_panels(1)
    ops.spmv(A, Z, O)
    assert A.panels() == 0
    assert np.all(np.abs(O.download() - exp) <= 4e-16 * absrow * 4)


@pytest.mark.parametrize("panels", [2, 4])
@pytest.mark.parametrize("name,rel", [("p3d7_30", 1e-8), ("rspd_50000", 1e-10)])
def test_sell_panels_solve_vs_oracle(ctx, port, name, rel, panels):
    gen = {"p3d7_30": lambda: G.poisson3d_7pt(30), "rspd_50000": lambda: G.random_spd_powerlaw(50000, seed=21)}[name]
    v, c, r = gen()
    n = r.size - 1
    b = G.default_rhs(n)
    tol = rel * np.linalg.norm(b)
    o = port.cg_solve(v, c, r, b, tol)
    A = S.Matrix(ctx, v, c, r)
    A.set_panels(panels)
    A.set_format(L.FMT_SELL)
    cg = S.CG(ctx, record_history=1)
    cg.setMatrix(A)
    cg.setTarget(b)
    st = cg.solve(tol)
    assert A.panels() == panels and st["format"] == L.FMT_SELL and st["converged"] == 1
    assert st["kernel_launches"] >= st["iterations"] * (3 + panels - 1)
    slack = 0.05 if name.startswith("rspd") else 0.02
    assert abs(st["iterations"] - o["iterations"]) <= max(1, round(slack * o["iterations"]))
    x1 = cg.extract()
    assert rel_err(x1, o["x"]) <= 1e-10
    rr, _ = cg.residual()
    assert np.sqrt(rr) <= 10 * tol
    cg.x = None
    cg.solve(tol)
    assert np.array_equal(cg.extract(), x1)                           # deterministic run to run
