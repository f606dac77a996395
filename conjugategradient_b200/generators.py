"""Deterministic synthetic SPD matrices for the CG hot path (host side, NumPy).

These are the inputs SURVEY.md §8(d) names: 2D 5-point, 3D 7-point and 3D 27-point
Dirichlet stencils in natural ordering, and a strictly diagonally dominant random SPD
matrix with power-law row lengths.  Every generator returns CSR in the exact layout the
reference's ``CG::setMatrix(data, columns, rows)`` expects (reference src/CG.hpp:87-93):
fp64 values, int32 column indices ascending inside a row, int32 row offsets of length n+1.

The device-side stencil generators in ``csrc/generate.cu`` produce bit-identical arrays
(tested in tests/test_generators.py); use those for sizes whose host CSR would not fit.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "poisson2d_5pt", "poisson3d_7pt", "stencil3d_27pt", "random_spd_powerlaw",
    "write_mtx_reference_dialect", "csr_to_lower_coo", "default_rhs", "stencil_nnz",
]


def _assemble(n, rows_list, cols_list, vals_list):
    rows = np.concatenate(rows_list)
    cols = np.concatenate(cols_list)
    vals = np.concatenate(vals_list)
    order = np.lexsort((cols, rows))
    rows, cols, vals = rows[order], cols[order], vals[order]
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.add.at(rowptr, rows + 1, 1)
    rowptr = np.cumsum(rowptr)
    assert rowptr[-1] < 2 ** 31, "reference API uses 32-bit row offsets (LinearAlgebraTypes.hpp:101-102)"
    return vals.astype(np.float64), cols.astype(np.int32), rowptr.astype(np.int32)


def _stencil(dims, offsets, diag, offdiag):
    """Generic Dirichlet-truncated stencil on a dims[0] x dims[1] x ... grid, x fastest."""
    dims = tuple(int(d) for d in dims)
    n = int(np.prod(dims))
    idx = np.arange(n, dtype=np.int64)
    coords = []
    rem = idx
    for d in dims:
        coords.append(rem % d)
        rem = rem // d
    strides = [1]
    for d in dims[:-1]:
        strides.append(strides[-1] * d)
    rows_l, cols_l, vals_l = [idx], [idx], [np.full(n, float(diag))]
    for off in offsets:
        ok = np.ones(n, dtype=bool)
        shift = 0
        for c, o, d, s in zip(coords, off, dims, strides):
            ok &= (c + o >= 0) & (c + o < d)
            shift += o * s
        r = idx[ok]
        rows_l.append(r)
        cols_l.append(r + shift)
        vals_l.append(np.full(r.size, float(offdiag)))
    return _assemble(n, rows_l, cols_l, vals_l)


def poisson2d_5pt(m: int):
    """2D 5-point Laplacian on an m x m grid: diag 4, neighbours -1, row = i + m*j (config C1)."""
    offs = [(-1, 0), (1, 0), (0, -1), (0, 1)]
    return _stencil((m, m), offs, 4.0, -1.0)


def poisson3d_7pt(m: int, mz: int | None = None):
    """3D 7-point Laplacian on m x m x mz: diag 6, neighbours -1, row = i + m*j + m^2*k (C2, C4)."""
    mz = m if mz is None else mz
    offs = [(-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1)]
    return _stencil((m, m, mz), offs, 6.0, -1.0)


def stencil3d_27pt(m: int, mz: int | None = None):
    """3D 27-point stencil: diag 26, all 26 neighbours -1, Dirichlet truncation (config C3)."""
    mz = m if mz is None else mz
    offs = [(a, b, c) for c in (-1, 0, 1) for b in (-1, 0, 1) for a in (-1, 0, 1) if (a, b, c) != (0, 0, 0)]
    return _stencil((m, m, mz), offs, 26.0, -1.0)


def stencil_nnz(kind: str, m: int) -> int:
    if kind == "2d5":
        return 5 * m * m - 4 * m
    if kind == "3d7":
        return 7 * m ** 3 - 6 * m * m
    if kind == "3d27":
        return (3 * m - 2) ** 3
    raise ValueError(kind)


def random_spd_powerlaw(n: int, seed: int = 12345, gamma: float = 2.2, d_min: int = 2, d_max: int = 4096):
    """Random strictly diagonally dominant SPD matrix with power-law row lengths (config C5).

    Strictly-lower degree of row i: min(d_max, floor(d_min * u^(-1/(gamma-1)))), u~U(0,1], capped
    at i; columns uniform in [0, i) without repetition; values U(-1, 0); mirrored; diagonal
    a_ii = 1 + sum_j |a_ij|.
    """
    rng = np.random.default_rng(seed)
    u = 1.0 - rng.random(n)  # (0, 1]
    deg = np.minimum(d_max, np.floor(d_min * u ** (-1.0 / (gamma - 1.0)))).astype(np.int64)
    deg = np.minimum(deg, np.arange(n, dtype=np.int64))
    total = int(deg.sum())
    rows = np.repeat(np.arange(n, dtype=np.int64), deg)
    # draw columns uniform in [0, row); dedupe (row, col) pairs afterwards
    cols = (rng.random(total) * rows).astype(np.int64)
    key = rows * n + cols
    key = np.unique(key)
    rows, cols = key // n, key % n
    vals = -rng.random(rows.size)
    vals[vals == 0.0] = -0.5
    absum = np.zeros(n)
    np.add.at(absum, rows, np.abs(vals))
    np.add.at(absum, cols, np.abs(vals))
    diag_idx = np.arange(n, dtype=np.int64)
    return _assemble(n, [rows, cols, diag_idx], [cols, rows, diag_idx], [vals, vals, 1.0 + absum])


def default_rhs(n: int) -> np.ndarray:
    """b[i] = i + 1, the right-hand side the reference driver uses (test/Tester.cpp:27-30)."""
    return np.arange(1, n + 1, dtype=np.float64)


def csr_to_lower_coo(vals, cols, rowptr):
    n = rowptr.size - 1
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(rowptr.astype(np.int64)))
    keep = cols <= rows
    return rows[keep], cols[keep].astype(np.int64), vals[keep]


def write_mtx_reference_dialect(path, vals, cols, rowptr, comment="generated by conjugategradient_b200"):
    """Write a symmetric matrix as Matrix Market in the dialect the reference reader needs.

    Rules (reference test/mm_reader.cpp): 5-word banner (:118); the line after the banner is
    discarded unconditionally (:163-164) so exactly one comment line must follow; one triangle
    only because every off-diagonal is mirrored regardless of the banner (:68-74); 1-based
    indices (:62-66); no duplicates, no empty rows.
    """
    n = rowptr.size - 1
    r, c, v = csr_to_lower_coo(vals, cols, rowptr)
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real symmetric\n")
        f.write(f"% {comment}\n")
        f.write(f"{n} {n} {r.size}\n")
        lines = [f"{int(a) + 1} {int(b) + 1} {float(x):.17g}\n" for a, b, x in zip(r, c, v)]
        f.writelines(lines)
    return r.size
