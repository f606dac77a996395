#!/usr/bin/env python
"""One arm of the L2 access-policy-window A/B: solves a stencil system for a fixed number of
trips with whatever CGB_L2_* environment the caller set and prints it/s and the per-kernel CUDA-event times.
  CGB_L2_WINDOW=0|1  CGB_L2_CARVE_MB=<MB>  CGB_L2_HITRATIO=<0..1>  CGB_L2_WINDOW_MODE=k1   (read by cgb_cg_solve)
  python tools/l2window_probe.py <3d7|3d27> <m> [trips = 256] [label]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    kind, m = sys.argv[1], int(sys.argv[2])
    trips = int(sys.argv[3]) if len(sys.argv) > 3 else 256
    label = sys.argv[4] if len(sys.argv) > 4 else ""
    ctx = S.Context(0)
    A = S.Matrix.stencil(ctx, kind, m)
    n = A.N()
    b = S.Vector(ctx, np.arange(1, n + 1, dtype=np.float64))
    out = {}
    for timed in (0, 1):  # untimed (pipelined batches) for it/s, then CUDA-event brackets per kernel
        cg = S.CG(ctx, loop_mode=L.LOOP_GRAPH, time_kernels=timed)
        cg.setMatrix(A)
        cg.setTarget(b)
        best = None
        for _ in range(3):
            cg.x = None
            st = cg.solve(0.0, max_iter=trips)
            if best is None or st["solve_ms"] < best["solve_ms"]:
                best = st
        if timed:
            out.update(spmv_ms=best["spmv_ms"], update_ms=best["update_ms"], pupdate_ms=best["pupdate_ms"])
        else:
            out.update(its_per_s=best["iterations"] / (best["solve_ms"] * 1e-3), rr=best["rr"])
    out.update(label=label, kind=kind, m=m, trips=trips,
               env={k: v for k, v in os.environ.items() if k.startswith("CGB_L2")})
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
