"""Wall time of cgb_cg_solve_host (new handle per call) on a 2D 5-point Poisson matrix: python tools/host_solve_probe.py [m]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
ctx = S.Context(0)
v, c, r = G.poisson2d_5pt(m)
b = G.default_rhs(r.size - 1)
tol = 1e-8 * np.linalg.norm(b)
for k in range(6):
    t0 = time.perf_counter()
    x, st = S.solve_host(ctx, v, c, r, b, tol)
    dt = time.perf_counter() - t0
    print(f"call {k}: wall {1e3 * dt:.3f} ms, solve_ms {st['solve_ms']:.3f}, loop_mode {st['loop_mode']}, it {st['iterations']}", flush=True)
