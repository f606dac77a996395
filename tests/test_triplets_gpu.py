"""Triplets (COO) -> device CSR and Matrix Market file -> device CSR: the GPU
assembly must produce exactly the arrays the reference reader's host assembly produces (test/mm_reader.cpp:45-107,
checked against the golden fixture, the oracle port and the library's own host reader), and CG on the result must
behave like CG on the host-built matrix."""
import os

import numpy as np
import pytest

from conftest import rel_err
from conjugategradient_b200 import _lib as L
from conjugategradient_b200 import generators as G
from conjugategradient_b200 import solver as S

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def same(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a, b))


def lower_triplets(v, c, r):
    rows = np.repeat(np.arange(r.size - 1), np.diff(r))
    keep = c <= rows
    return rows[keep].astype(np.int32), c[keep].astype(np.int32), v[keep]


def test_mm_file_to_device_matches_golden_reader_fixture(ctx, tmp_path, port):
    g = np.load(os.path.join(GOLD, "reader.npz"))
    p = tmp_path / "q.mtx"
    p.write_bytes(g["text"].tobytes())
    A = S.Matrix.from_mm_file(ctx, str(p))
    assert same(A.to_host(), (g["vals"], g["cols"], g["rows"]))
    assert same(A.to_host(), port.mm_read(str(p)))


@pytest.mark.parametrize("gen", [lambda: G.poisson2d_5pt(40), lambda: G.stencil3d_27pt(7),
                                 lambda: G.random_spd_powerlaw(30000, seed=3)])
def test_mm_file_to_device_equals_host_reader_and_solves(ctx, tmp_path, port, gen):
    v, c, r = gen()
    p = str(tmp_path / "a.mtx")
    G.write_mtx_reference_dialect(p, v, c, r)
    A = S.Matrix.from_mm_file(ctx, p)
    assert same(A.to_host(), S.mm_read(p)) and same(A.to_host(), (v, c, r))
    n = r.size - 1
    b = G.default_rhs(n)
    tol = 1e-10 * np.linalg.norm(b)
    o = port.cg_solve(v, c, r, b, tol)
    cg = S.CG(ctx)
    cg.setMatrix(A)
    cg.setTarget(b)
    st = cg.solve(tol)
    assert abs(st["iterations"] - o["iterations"]) <= max(1, round(0.05 * o["iterations"]))
    assert rel_err(cg.extract(), o["x"]) <= 1e-10


def test_mm_file_quirks_on_device(ctx, tmp_path):
    cases = {
        "nocomment": "%%MatrixMarket matrix coordinate real symmetric\n3 3 4\n1 1 2\n2 1 -1\n2 2 2\n3 3 2\n",
        "unsorted": "%%MatrixMarket matrix coordinate real symmetric\n%\n3 3 4\n3 3 2.5e0\n 2 1   -1e-1\n1 1 2\n2 2 +2\n",
    }
    for name, text in cases.items():
        p = tmp_path / (name + ".mtx")
        p.write_text(text)
        assert same(S.Matrix.from_mm_file(ctx, str(p)).to_host(), S.mm_read(str(p))), name
    # a file with an empty row: the reference reader compacts the rows and keeps the column ids (test/mm_reader.cpp:92-104),
    # so its spmv would read outside the vector; the host reader still returns the reference's arrays (tests/test_reader.py),
    # the device path, whose result goes straight into a solve, refuses the file
    p = tmp_path / "emptyrow.mtx"
    p.write_text("%%MatrixMarket matrix coordinate real symmetric\n%\n3 3 2\n1 1 2\n3 3 2\n")
    with pytest.raises(L.CgbError) as e:
        S.Matrix.from_mm_file(ctx, str(p))
    assert e.value.code == L.CGB_ERR_INVALID
    with pytest.raises(L.CgbError) as e:
        S.Matrix.from_mm_file(ctx, "/nonexistent/file.mtx")
    assert e.value.code == L.CGB_ERR_IO


def test_triplets_flags(ctx):
    v, c, r = G.random_spd_powerlaw(5000, seed=8)
    n = r.size - 1
    rows = np.repeat(np.arange(n), np.diff(r)).astype(np.int32)
    rng = np.random.default_rng(1)
    # full matrix, shuffled, zero-based, explicit n
    perm = rng.permutation(v.size)
    A = S.Matrix.from_triplets(ctx, rows[perm], c[perm], v[perm], n=n)
    assert same(A.to_host(), (v, c, r))
    # one triangle, one-based, mirrored
    lr, lc, lv = lower_triplets(v, c, r)
    perm = rng.permutation(lv.size)
    B = S.Matrix.from_triplets(ctx, lr[perm] + 1, lc[perm] + 1, lv[perm], n=n, flags=L.COO_MIRROR | L.COO_ONE_BASED)
    assert same(B.to_host(), (v, c, r))
    # empty rows are kept without COMPACT_ROWS and vanish with it
    keep = (rows % 7 != 3) & (c % 7 != 3)
    C1 = S.Matrix.from_triplets(ctx, rows[keep], c[keep], v[keep], n=n)
    hv, hc, hr = C1.to_host()
    assert hr.size == n + 1 and np.all(np.diff(hr)[np.arange(n) % 7 == 3] == 0) and hr[-1] == keep.sum()
    C2 = S.Matrix.from_triplets(ctx, rows[keep], c[keep], v[keep], flags=L.COO_COMPACT_ROWS)
    cv, cc, cr = C2.to_host()
    assert cr.size == n - np.count_nonzero(np.arange(n) % 7 == 3) + 1 and same((cv, cc), (hv, hc))
    # ... which leaves a rectangular matrix (column ids are not renumbered, test/mm_reader.cpp:92-104): the solver refuses it
    cg = S.CG(ctx)
    cg.setMatrix(C2)
    cg.setTarget(np.ones(C2.N()))
    with pytest.raises(L.CgbError) as e:
        cg.solve(1e-8)
    assert e.value.code == L.CGB_ERR_INVALID and "not square" in str(e.value)
    # n = -1: dimension from the largest index
    D = S.Matrix.from_triplets(ctx, rows, c, v)
    assert D.N() == n


def test_triplets_duplicates_are_summed_in_input_order(ctx):
    """Finite-element style assembly: element contributions to the same (i, j) add up."""
    rng = np.random.default_rng(3)
    n, m = 2000, 60000
    r = rng.integers(0, n, m).astype(np.int32)
    c = rng.integers(0, n, m).astype(np.int32)
    v = rng.standard_normal(m)
    A = S.Matrix.from_triplets(ctx, r, c, v, n=n, flags=L.COO_SUM_DUPLICATES)
    hv, hc, hr = A.to_host()
    import scipy.sparse as sp
    ref = sp.coo_matrix((v, (r, c)), shape=(n, n)).tocsr()
    ref.sort_indices()
    assert np.array_equal(hr, ref.indptr) and np.array_equal(hc, ref.indices)
    assert np.allclose(hv, ref.data, rtol=0, atol=1e-13)
    # input order inside a run: reproducible bit for bit
    B = S.Matrix.from_triplets(ctx, r, c, v, n=n, flags=L.COO_SUM_DUPLICATES)
    assert np.array_equal(B.to_host()[0], hv)
    # without the flag duplicates stay separate entries
    K = S.Matrix.from_triplets(ctx, r, c, v, n=n)
    assert K.NNZ() == m and A.NNZ() == ref.nnz < m


def test_triplets_errors(ctx):
    r = np.array([0, 1, 5], np.int32)
    v = np.ones(3)
    with pytest.raises(L.CgbError):
        S.Matrix.from_triplets(ctx, r, r, v, n=3)                      # index outside the matrix
    with pytest.raises(L.CgbError):
        S.Matrix.from_triplets(ctx, r, r, v, flags=L.COO_ONE_BASED)    # index below 1
    with pytest.raises(L.CgbError):
        S.Matrix.from_triplets(ctx, r, r, v, n=6, flags=L.COO_COMPACT_ROWS)
