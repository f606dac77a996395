#!/usr/bin/env python
"""Both arms of the L2 access-policy-window A/B in ONE process, for a single ncu pass over the SpMV launches
(VERDICT r1 item 8a): a few trips with the library's default (evict-first hints on the matrix stream, no window), then
the same trips with the persisting window over p (CGB_L2_WINDOW=1), then with a carve-out sized to the window.
The launches appear in the ncu log in that order, `trips` + 1 per arm (K0's row pass is the same kernel family).
  ncu --metrics ... -k regex:cg_spmv_kernel python tools/l2window_ncu.py <3d7|3d27> <m> [trips = 4]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    kind, m = sys.argv[1], int(sys.argv[2])
    trips = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    ctx = S.Context(0)
    A = S.Matrix.stencil(ctx, kind, m)
    n = A.N()
    b = S.Vector(ctx, np.arange(1, n + 1, dtype=np.float64))
    arms = [("default", {}), ("window", {"CGB_L2_WINDOW": "1"}),
            ("window_carve", {"CGB_L2_WINDOW": "1", "CGB_L2_CARVE_MB": str(max(1, (8 * n + (1 << 20) - 1) >> 20))})]
    for label, env in arms:
        for k in [k for k in os.environ if k.startswith("CGB_L2")]:
            del os.environ[k]
        os.environ.update(env)
        cg = S.CG(ctx, loop_mode=L.LOOP_STEPWISE)
        cg.setMatrix(A)
        cg.setTarget(b)
        st = cg.solve(0.0, max_iter=trips)
        print(f"ARM {label} env={env} trips={st['iterations']} rr={st['rr']:.6e}", flush=True)


if __name__ == "__main__":
    main()
