#!/usr/bin/env python
"""World-size-1 run of the row-partitioned solver (cgb_dist_*): the peer windows, flags and the persistent
cooperative loop all live on one GPU, so compute-sanitizer can watch them (it cannot follow cudaIpc mappings).
  python tools/dist_single_probe.py [m]          3D 7-point m^3 (default 48); prints the loop mode and the error vs cgb_cg_solve
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 48
    ctx = S.Context(0)
    lib = ctx.lib
    v, c, r = G.poisson3d_7pt(m)
    n = r.size - 1
    b = G.default_rhs(n)
    tol = 1e-8 * np.linalg.norm(b)
    single = S.CG(ctx, loop_mode=L.LOOP_GRAPH)
    single.setMatrix(v, c, r)
    single.setTarget(b)
    st0 = single.solve(tol)
    x0 = single.extract()
    idbuf = (C.c_ubyte * L.NCCL_ID_BYTES)()
    L.check(lib.cgb_nccl_unique_id(idbuf))
    starts = np.array([0, n], dtype=np.int64)
    A = S.Matrix(ctx, v, c, r)
    d = S.DistCG(ctx, 0, 1, bytes(idbuf), starts, A)
    bv, xv = S.Vector(ctx, b), S.Vector(ctx, n=n)
    for rep in range(2):
        xv.upload(np.zeros(n))
        st = d.solve(bv, xv, tol)
        x = xv.download()
        err = np.linalg.norm(x - x0) / np.linalg.norm(x0)
        print(f"dist world=1 m={m}: loop_mode={st['loop_mode']} iterations={st['iterations']} (single {st0['iterations']}) "
              f"converged={st['converged']} rel diff vs single-GPU solve={err:.2e} launches={st['kernel_launches']}", flush=True)
        assert st["iterations"] == st0["iterations"] and err < 1e-12
    d.close()
    print("PROBE_OK")


if __name__ == "__main__":
    main()
