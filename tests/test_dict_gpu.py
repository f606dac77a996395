"""Row-pattern dictionary engine (CGB_FMT_DICT): an explicit, lossless storage option for
matrices that repeat a few row patterns (constant-coefficient stencils): one 16-bit pattern id per row.  Same
arithmetic as the CSR engines (entries in CSR order, one FMA chain per row), so SpMV is bit-identical to the
one-lane-per-row CSR kernel and whole solves agree with the oracle exactly like the default path does."""
import numpy as np
import pytest

from conftest import rel_err
from conjugategradient_b200 import _lib as L
from conjugategradient_b200 import generators as G
from conjugategradient_b200 import solver as S

pytestmark = pytest.mark.gpu

MATS = {
    "p2d_33": lambda: G.poisson2d_5pt(33),          # n = 1089: not a multiple of 32
    "p3d7_20": lambda: G.poisson3d_7pt(20),
    "p3d27_11": lambda: G.stencil3d_27pt(11),       # 27 entries per row: 4 groups of 8
    "diag_1": lambda: (np.array([2.0]), np.array([0], np.int32), np.array([0, 1], np.int32)),
}


def banded(n, half, seed=0, period=3, p_keep=1.0):
    """Band matrix whose coefficients repeat with `period` rows (few row patterns when p_keep == 1); every 13th row
    is empty.  p_keep < 1 drops entries at random: as many patterns as rows."""
    rng = np.random.default_rng(seed)
    palette = np.array([4.0, -1.0, -0.5, 0.25, 1e-3, -7.0])
    rows, cols, vals = [], [], []
    for i in range(n):
        if i % 13 == 5:
            continue                                  # an empty row
        for d in range(-half, half + 1):
            j = i + d
            if 0 <= j < n and (d == 0 or rng.random() < p_keep):
                rows.append(i); cols.append(j); vals.append(palette[(abs(d) * 7 + i % period) % palette.size] if d else 9.0)
    rows, cols, vals = np.array(rows), np.array(cols, np.int32), np.array(vals)
    r = np.zeros(n + 1, np.int32)
    np.add.at(r, rows + 1, 1)
    return vals, cols, np.cumsum(r).astype(np.int32)


PATTERNS = {"p2d_33": 9, "p3d7_20": 27, "p3d27_11": 27, "diag_1": 1}


@pytest.mark.parametrize("name", sorted(MATS) + ["banded_12", "banded_3"])
def test_dict_spmv_accuracy_and_bitwise_vs_csr(ctx, port, name):
    v, c, r = MATS[name]() if name in MATS else banded(2500, int(name.split("_")[1]), seed=3)
    n = r.size - 1
    rng = np.random.default_rng(2)
    z, b = rng.standard_normal(n), rng.standard_normal(n)
    A = S.Matrix(ctx, v, c, r)
    ops = S.VectorOperations(ctx)
    ops.setVectorSize(n)
    Z, O = S.Vector(ctx, z), S.Vector(ctx, n=n)
    A.set_format(L.FMT_DICT)
    assert 2 * n < A.dict_bytes() <= 2 * n + 40 * 1024
    assert A.dict_patterns() == PATTERNS.get(name, A.dict_patterns()) and 1 <= A.dict_patterns() <= 1024
    ops.spmv(A, Z, O)
    got = O.download()
    exp = port.spmv(v, c, r, z)
    absrow = port.spmv(np.abs(v), c, r, np.abs(z))
    assert np.all(np.abs(got - exp) <= 4e-16 * absrow * 4)
    # the same FMA chain as numpy-free sequential evaluation: compare with an explicit left-to-right fma emulation
    # through the library's own one-lane CSR engine (row lengths <= 7 use one lane per row)
    if int(np.diff(r).max()) <= 7:
        A.set_format(L.FMT_CSR)
        ops.spmv(A, Z, O)
        assert np.array_equal(O.download(), got)
        A.set_format(L.FMT_DICT)
    cg = S.CG(ctx)
    cg.setMatrix(A)
    cg.setTarget(b)
    cg.setInital(z)
    acc, rr, xx = port.accuracy(v, c, r, z, b)
    grr, gxx = cg.residual()
    assert abs(grr - rr) <= 1e-13 * rr and abs(gxx - xx) <= 1e-13 * xx


@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("name,rel", [("C1_p2d_128", 1e-8), ("p3d7_40", 1e-8), ("p3d27_16", 1e-10)])
def test_dict_solve_vs_oracle(ctx, port, name, rel, fused):
    gen = {"C1_p2d_128": lambda: G.poisson2d_5pt(128), "p3d7_40": lambda: G.poisson3d_7pt(40),
           "p3d27_16": lambda: G.stencil3d_27pt(16)}[name]
    v, c, r = gen()
    n = r.size - 1
    b = G.default_rhs(n)
    tol = rel * np.linalg.norm(b)
    o = port.cg_solve(v, c, r, b, tol)
    A = S.Matrix(ctx, v, c, r)
    A.set_format(L.FMT_DICT)
    cg = S.CG(ctx, fuse_p_update=fused)
    cg.setMatrix(A)
    cg.setTarget(b)
    st = cg.solve(tol)
    assert st["format"] == L.FMT_DICT and st["converged"] == 1
    assert abs(st["iterations"] - o["iterations"]) <= max(1, round(0.02 * o["iterations"]))
    x1 = cg.extract()
    assert rel_err(x1, o["x"]) <= 1e-10
    cg.x = None
    cg.solve(tol)
    assert np.array_equal(cg.extract(), x1)                           # deterministic run to run
    A.set_format(L.FMT_CSR)                                           # and back to CSR on the same handle
    cg.x = None
    st2 = cg.solve(tol)
    assert st2["format"] == L.FMT_CSR and rel_err(cg.extract(), o["x"]) <= 1e-10


def test_dict_refuses_what_does_not_fit(ctx):
    v, c, r = G.random_spd_powerlaw(3000, seed=2)                      # thousands of distinct values
    A = S.Matrix(ctx, v, c, r)
    with pytest.raises(L.CgbError) as e:
        A.set_format(L.FMT_DICT)
    assert e.value.code == L.CGB_ERR_UNSUPPORTED
    assert ctx.lib.cgb_csr_get_format(A.h) == L.FMT_AUTO and A.dict_bytes() == 0
    v, c, r = banded(400, 40, p_keep=0.7)                              # hundreds of patterns, 20,000 pattern entries
    B = S.Matrix(ctx, v, c, r)
    with pytest.raises(L.CgbError) as e:
        B.set_format(L.FMT_DICT)
    assert e.value.code == L.CGB_ERR_UNSUPPORTED
    # both still solve through the default engines
    z = np.ones(400)
    ops = S.VectorOperations(ctx)
    ops.setVectorSize(400)
    O = S.Vector(ctx, n=400)
    ops.spmv(B, S.Vector(ctx, z), O)
    import scipy.sparse as sp
    assert np.allclose(O.download(), sp.csr_matrix((v, c, r), shape=(400, 400)) @ z, rtol=1e-13)
