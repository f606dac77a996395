"""Generates tests/golden/*.npz from the REFERENCE'S OWN SOURCES compiled in this container
(oracle/_ref/libref_serial.so = /root/reference/src/*.hpp + test/mm_reader.cpp against the serial
SYCL stand-in).  Run from the repo root:  python tests/golden/make_golden.py
The fixtures are small on purpose; they pin the C restatement (oracle/cg_oracle.c) and, on the GPU
box where /root/reference does not exist, the CUDA path.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from conjugategradient_b200 import generators as G  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def case(name, vals, cols, rows, tol_rel, ref, x0=None):
    n = rows.size - 1
    b = G.default_rhs(n)
    tol = tol_rel * np.linalg.norm(b) if tol_rel < 1e-20 or tol_rel > 0 else 0.0
    if tol_rel == -24:
        tol = 1e-24  # what test/Tester.cpp:45 passes
    out = ref.cg_solve(vals, cols, rows, b, tol, x0=x0)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), vals=vals, cols=cols, rows=rows, b=b, tol=tol,
                        x0=np.zeros(0) if x0 is None else x0, x=out["x"], iterations=out["iterations"],
                        accuracy=out["accuracy"], footprint=out["footprint"])
    print(f"{name}: n={n} nnz={vals.size} tol={tol:.6e} iterations={out['iterations']} accuracy={out['accuracy']:.3e}")


def main():
    ref = oracle.Ref("serial")
    case("p2d_8_rel1e-8", *G.poisson2d_5pt(8), 1e-8, ref)
    case("p2d_8_tester", *G.poisson2d_5pt(8), -24, ref)
    case("p2d_32_rel1e-8", *G.poisson2d_5pt(32), 1e-8, ref)
    case("p2d_32_rel1e-12", *G.poisson2d_5pt(32), 1e-12, ref)
    case("p3d7_12_rel1e-8", *G.poisson3d_7pt(12), 1e-8, ref)
    case("p3d27_10_rel1e-8", *G.stencil3d_27pt(10), 1e-8, ref)
    v, c, r = G.random_spd_powerlaw(3000, seed=7)
    case("rspd_3000_rel1e-10", v, c, r, 1e-10, ref)
    rng = np.random.default_rng(3)
    case("p2d_16_warm", *G.poisson2d_5pt(16), 1e-9, ref, x0=rng.standard_normal(256))
    # per-op vectors: reductions with a non-zero initial value (accumulate-into), axpy family, spmv
    x, y = rng.standard_normal(1000), rng.standard_normal(1000)
    v, c, r = G.poisson2d_5pt(20)
    z = rng.standard_normal(400)
    np.savez_compressed(
        os.path.join(HERE, "ops.npz"), x=x, y=y,
        dot_trivial=ref.reduce("dot_product_trivial", x, y, 2.5), dot_opt=ref.reduce("dot_product_optimised", x, y, 2.5),
        dot_dep=ref.reduce("dot_product", x, y, 2.5), norm=ref.reduce("norm", x, None, 1.25),
        saxpby=ref.axpy("saxpby", x, y, 0.75, -1.5), sapbx=ref.axpy("sapbx", x, y, b=0.3), sambx=ref.axpy("sambx", x, y, b=0.3),
        sapbx_alias_x=ref.axpy("sapbx", x, y, b=0.3, alias=1), sapbx_alias_y=ref.axpy("sapbx", x, y, b=0.3, alias=2),
        sambx_alias_x=ref.axpy("sambx", x, y, b=0.3, alias=1),
        sp_vals=v, sp_cols=c, sp_rows=r, sp_vec=z, sp_out=ref.spmv(v, c, r, z))
    # reader fixture: the dialect quirks (general banner still mirrored, line 2 discarded)
    d = tempfile.mkdtemp()
    p = os.path.join(d, "q.mtx")
    with open(p, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n% c1\n% c2\n4 4 6\n1 1 4\n2 1 -1\n2 2 4\n4 2 -0.5\n3 3 2e0\n4 4 3.25\n")
    mv, mc, mr = ref.mm_read(p)
    np.savez_compressed(os.path.join(HERE, "reader.npz"), text=np.frombuffer(open(p, "rb").read(), dtype=np.uint8),
                        vals=mv, cols=mc, rows=mr)
    print("reader:", mv, mc, mr)


if __name__ == "__main__":
    main()
