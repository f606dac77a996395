"""Matrix Market reader of the library vs the reference reader (golden fixture + compiled
reference when present) and vs the oracle port.  CPU only (host code of libcgb200.so)."""
import os

import numpy as np
import pytest

from conjugategradient_b200 import _lib as L
from conjugategradient_b200 import generators as G
from conjugategradient_b200 import solver as S

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def same(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a, b))


def test_reader_golden_fixture(tmp_path, port):
    g = np.load(os.path.join(GOLD, "reader.npz"))
    p = tmp_path / "q.mtx"
    p.write_bytes(g["text"].tobytes())
    exp = (g["vals"], g["cols"], g["rows"])
    assert same(S.mm_read(str(p)), exp)      # general banner is still mirrored; 2nd comment skipped
    assert same(port.mm_read(str(p)), exp)


@pytest.mark.parametrize("gen", [lambda: G.poisson2d_5pt(17), lambda: G.poisson3d_7pt(6), lambda: G.stencil3d_27pt(5),
                                 lambda: G.random_spd_powerlaw(400, seed=3)])
def test_reader_roundtrip(tmp_path, port, gen):
    v, c, r = gen()
    p = str(tmp_path / "a.mtx")
    G.write_mtx_reference_dialect(p, v, c, r)
    got = S.mm_read(p)
    assert same(got, (v, c, r))
    assert same(port.mm_read(p), (v, c, r))


def test_reader_quirks_match_compiled_reference(tmp_path, ref_serial):
    cases = {
        # no comment line: the size line is eaten by the unconditional discard, first entry lost
        "nocomment": "%%MatrixMarket matrix coordinate real symmetric\n3 3 4\n1 1 2\n2 1 -1\n2 2 2\n3 3 2\n",
        # unsorted input, scientific notation, extra blanks
        "unsorted": "%%MatrixMarket matrix coordinate real symmetric\n%\n3 3 4\n3 3 2.5e0\n 2 1   -1e-1\n1 1 2\n2 2 +2\n",
        # an empty row (row 2 absent) vanishes from the offsets
        "emptyrow": "%%MatrixMarket matrix coordinate real symmetric\n%\n3 3 2\n1 1 2\n3 3 2\n",
    }
    for name, text in cases.items():
        p = tmp_path / (name + ".mtx")
        p.write_text(text)
        assert same(S.mm_read(str(p)), ref_serial.mm_read(str(p))), name


def test_reader_missing_file_is_an_error_not_an_abort():
    with pytest.raises(L.CgbError) as e:
        S.mm_read("/nonexistent/file.mtx")
    assert e.value.code == L.CGB_ERR_IO


def test_reader_parallel_parse_equals_serial(tmp_path, port, monkeypatch):
    """A body large enough to be cut into several line-aligned chunks parses to the same arrays as one thread
    and as the oracle's reader (which follows the reference's token loop, test/mm_reader.cpp:62-66)."""
    v, c, r = G.random_spd_powerlaw(40000, seed=4)
    p = str(tmp_path / "big.mtx")
    G.write_mtx_reference_dialect(p, v, c, r)
    assert os.path.getsize(p) > 4 * 256 * 1024
    monkeypatch.setenv("CGB_MM_THREADS", "7")
    many = S.mm_read(p)
    monkeypatch.setenv("CGB_MM_THREADS", "1")
    one = S.mm_read(p)
    assert same(many, one) and same(many, (v, c, r)) and same(port.mm_read(p), many)


def test_reader_odd_layouts_fall_back_to_the_token_stream(tmp_path, ref_serial, monkeypatch):
    """Triples that straddle lines, and a body that stops parsing half way: the line-parallel fast path must
    step aside and reproduce the reference's own reading loop."""
    monkeypatch.setenv("CGB_MM_THREADS", "4")
    head = "%%MatrixMarket matrix coordinate real symmetric\n%\n4 4 5\n"
    cases = {
        "straddle": head + "1 1\n2.0 2 1\n-1 2 2 2\n3 3 2 4 4\n2\n",
        "stops": head + "1 1 2\n2 1 -1\n2 2 2\nthis is not a number\n3 3 2\n4 4 2\n",
        "blank_lines": head + "1 1 2\n\n2 1 -1\n   \n2 2 2\n3 3 2\n4 4 2\n\n",
    }
    for name, text in cases.items():
        p = tmp_path / (name + ".mtx")
        p.write_text(text)
        assert same(S.mm_read(str(p)), ref_serial.mm_read(str(p))), name


def test_reader_vs_reference_speed(tmp_path, ref_serial, capsys):
    """Same file through the reference's own read_file (compiled where it lies, oracle/_ref) and through the library:
    identical arrays, and the library's reader must not be the slower one (GPU box, 147 MB C2-sized file: 4.56 s
    against 0.25 s)."""
    import time
    v, c, r = G.poisson3d_7pt(40)
    p = str(tmp_path / "p.mtx")
    G.write_mtx_reference_dialect(p, v, c, r)
    t0 = time.perf_counter()
    ref = ref_serial.mm_read(p)
    t_ref = time.perf_counter() - t0
    S.mm_read(p)                                   # page cache / first-touch out of the picture
    t0 = time.perf_counter()
    ours = S.mm_read(p)
    t_ours = time.perf_counter() - t0
    assert same(ours, ref) and same(ours, (v, c, r))
    with capsys.disabled():
        print(f"\n[reader] n={r.size - 1} nnz={v.size}: reference read_file {t_ref:.3f} s, cgb_mm_read {t_ours:.3f} s")
    assert t_ours < t_ref


@pytest.mark.gpu
def test_empty_row_is_refused(tmp_path):
    """The reference dialect compacts empty rows away without renumbering columns (test/mm_reader.cpp:92-104): its
    spmv then reads outside the vector.  The device path refuses such a file instead of faulting."""
    from conjugategradient_b200 import _lib as L
    from conjugategradient_b200 import solver as S
    p = tmp_path / "hole.mtx"
    p.write_text("%%MatrixMarket matrix coordinate real symmetric\n% row 3 is empty\n4 4 4\n1 1 2.0\n2 2 2.0\n2 1 -1.0\n4 4 2.0\n")
    ctx = S.Context(0)
    with pytest.raises(L.CgbError) as e:
        S.Matrix.from_mm_file(ctx, str(p))
    assert e.value.code == L.CGB_ERR_INVALID and "empty row" in str(e.value)
