"""Pins the C restatement (oracle/cg_oracle.c) against (a) the committed golden fixtures that were
generated from the reference's own sources (tests/golden/make_golden.py) and (b), when
oracle/_ref is built, the compiled reference itself on fresh inputs.  CPU only."""
import glob
import os

import numpy as np
import pytest

from conftest import rel_err
from conjugategradient_b200 import generators as G

GOLD = os.path.join(os.path.dirname(__file__), "golden")
SOLVES = sorted(p for p in glob.glob(os.path.join(GOLD, "*.npz")) if os.path.basename(p) not in ("ops.npz", "reader.npz"))


@pytest.mark.parametrize("path", SOLVES, ids=[os.path.basename(p)[:-4] for p in SOLVES])
def test_port_reproduces_golden_solves(port, path):
    g = np.load(path)
    x0 = g["x0"] if g["x0"].size else None
    out = port.cg_solve(g["vals"], g["cols"], g["rows"], g["b"], float(g["tol"]), x0=x0)
    # serial left-to-right arithmetic on both sides: identical trip count, bit-identical x
    assert out["iterations"] == int(g["iterations"])
    assert np.array_equal(out["x"], g["x"])
    acc, _, _ = port.accuracy(g["vals"], g["cols"], g["rows"], out["x"], g["b"])
    assert acc == float(g["accuracy"])
    n, nnz = g["rows"].size - 1, g["vals"].size
    assert port.memory_footprint(n, nnz) == int(g["footprint"])


def test_port_reproduces_golden_ops(port):
    g = np.load(os.path.join(GOLD, "ops.npz"))
    x, y = g["x"], g["y"]
    assert port.dot_acc(x, y, 2.5) == float(g["dot_trivial"])
    assert port.norm_acc(x, 1.25) == float(g["norm"])
    assert np.array_equal(port.saxpby(x, y, 0.75, -1.5), g["saxpby"])
    assert np.array_equal(port.sapbx(x, y, 0.3), g["sapbx"])
    assert np.array_equal(port.sambx(x, y, 0.3), g["sambx"])
    # in-place forms of the reference give the same numbers as out-of-place (element-wise ops)
    assert np.array_equal(g["sapbx_alias_x"], g["sapbx"]) and np.array_equal(g["sapbx_alias_y"], g["sapbx"])
    assert np.array_equal(g["sambx_alias_x"], g["sambx"])
    assert np.array_equal(port.spmv(g["sp_vals"], g["sp_cols"], g["sp_rows"], g["sp_vec"]), g["sp_out"])
    # the reference's tree reductions differ from the serial sum only by reassociation
    for k in ("dot_opt", "dot_dep"):
        assert abs(float(g[k]) - float(g["dot_trivial"])) <= 1e-12 * np.abs(x * y).sum()


def test_survey_probe_numbers(port):
    """SURVEY.md §8c probe results: C1 (2D 5-pt 128^2, b = 1..n, x0 = 0)."""
    v, c, r = G.poisson2d_5pt(128)
    assert v.size == 81408 and r.size - 1 == 16384
    b = G.default_rhs(16384)
    nb = np.linalg.norm(b)
    assert abs(nb - 1.210847e6) < 1.0
    for rel, iters in ((1e-8, 366), (1e-10, 408), (1e-12, 445)):
        assert port.cg_solve(v, c, r, b, rel * nb)["iterations"] == iters
    assert port.memory_footprint(16384, 81408) == 1957888


def test_port_matches_compiled_reference_on_fresh_inputs(port, ref_serial):
    rng = np.random.default_rng(11)
    for v, c, r in (G.poisson2d_5pt(24), G.poisson3d_7pt(9), G.stencil3d_27pt(7), G.random_spd_powerlaw(1500, seed=5)):
        n = r.size - 1
        b = rng.standard_normal(n)
        tol = 1e-9 * np.linalg.norm(b)
        a, o = ref_serial.cg_solve(v, c, r, b, tol), port.cg_solve(v, c, r, b, tol)
        assert a["iterations"] == o["iterations"] and np.array_equal(a["x"], o["x"])
        z = rng.standard_normal(n)
        assert np.array_equal(ref_serial.spmv(v, c, r, z), port.spmv(v, c, r, z))


def test_stop_rule_semantics(port):
    """Absolute tolerance, one-iteration lag, N+1 cap, NaN stop (reference src/CG.hpp:396-404, :436)."""
    v, c, r = G.poisson2d_5pt(8)
    b = G.default_rhs(64)
    out = port.cg_solve(v, c, r, b, 1e-3, history=True)
    h = np.sqrt(out["hist"])
    k = int(np.argmax(h <= 1e-3))  # first k with ||r_k|| <= tol
    assert out["iterations"] == k + 1  # the check lags one iteration
    assert port.cg_solve(v, c, r, b, 1e300)["iterations"] == 1
    capped = port.cg_solve(v, c, r, b, 0.0)  # default improvement = 0: runs to NaN or the N+1 cap
    assert capped["iterations"] <= 65
    assert np.isnan(capped["x"]).all() or capped["iterations"] == 65


def test_error_messages_of_reference(ref_serial):
    assert ref_serial.solve_missing(True, False) == (True, "No right hand side to solve for")
    assert ref_serial.solve_missing(False, True) == (True, "No Matrix given")
