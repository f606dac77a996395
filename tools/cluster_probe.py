"""Cluster-resident loop against the cooperative kernel on a 2D 5-point Poisson matrix: python tools/cluster_probe.py [m]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
ctx = S.Context(0)
v, c, r = G.poisson2d_5pt(m)
n = r.size - 1
b = G.default_rhs(n)
tol = 1e-8 * np.linalg.norm(b)
for mode in (L.LOOP_CLUSTER, L.LOOP_PERSISTENT):
    cg = S.CG(ctx, loop_mode=mode)
    cg.setMatrix(v, c, r)
    cg.setTarget(b)
    cg.solve(tol)
    best = 1e9
    for _ in range(5):
        cg.x = None
        st = cg.solve(tol)
        best = min(best, st["solve_ms"])
    ctx.sync()
    print(f"mode={mode} it={st['iterations']} solve_ms={best:.4f} us/it={1e3 * best / st['iterations']:.3f}", flush=True)
