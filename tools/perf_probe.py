"""Quick per-kernel timing probe (not the bench): python tools/perf_probe.py [m ...]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    ms = [int(a) for a in sys.argv[1:] if a.isdigit()] or [128, 256]
    kinds = [a for a in sys.argv[1:] if a in ("3d7", "3d27", "2d5")] or ["3d7"]
    iters = int(os.environ.get("PROBE_ITERS", "200"))
    ctx = S.Context(0)
    for kind in kinds:
        for m in ms:
            A = S.Matrix.stencil(ctx, kind, m)
            n, nnz = A.N(), A.NNZ()
            b = G.default_rhs(n)
            fl = [int(c) for c in os.environ.get("PROBE_FUSED", "01")]
            tl = [int(c) for c in os.environ.get("PROBE_TIMED", "01")]
            for fused in fl:
                for timed in tl:
                    cg = S.CG(ctx, fuse_p_update=fused, time_kernels=timed, check_every=25, loop_mode=int(os.environ.get("PROBE_LOOP", "1")))
                    cg.setMatrix(A)
                    cg.setTarget(b)
                    cg.solve(0.0, max_iter=50)  # warm-up (graph build)
                    cg.x = None
                    st = cg.solve(0.0, max_iter=iters)
                    it = st["iterations"]
                    B_iter = 12 * nnz + 4 * (n + 1) + 88 * n
                    sec = st["solve_ms"] / 1e3
                    line = (f"{kind} m={m} n={n} nnz={nnz} fused={fused} timed={timed} it={it} {it / sec:9.1f} it/s "
                            f"{it * B_iter / sec / 1e9:8.1f} GB/s(B_iter)")
                    if timed:
                        k1 = (12 * nnz + 4 * (n + 1) + (48 if fused else 16) * n) / (st["spmv_ms"] * 1e6) if st["spmv_ms"] else 0
                        k2 = 24 * n / (st["update_ms"] * 1e6) if st["update_ms"] else 0
                        k3 = 40 * n / (st["pupdate_ms"] * 1e6) if st["pupdate_ms"] else 0
                        line += (f" | spmv {st['spmv_ms']:.4f} ms {k1:7.1f} GB/s | update {st['update_ms']:.4f} ms {k2:7.1f} GB/s"
                                 f" | pupd {st['pupdate_ms']:.4f} ms {k3:7.1f} GB/s | samples {st['timed_samples']}")
                    print(line, flush=True)
            A.free()


if __name__ == "__main__":
    main()
