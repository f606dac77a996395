"""C5 probe: random SPD n rows on the device, CSR-TMA vs SELL engines.  python tools/c5_probe.py [n]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
    ctx = S.Context(0)
    t0 = time.time()
    kind = os.environ.get("C5_KIND", "rspd")              # rspd | 3d27 | 3d7 | 2d5 (the argument is the grid size m then)
    A = S.Matrix.random_spd(ctx, n) if kind == "rspd" else S.Matrix.stencil(ctx, kind, n, n, 0, -1)
    n = A.N()
    ctx.sync()
    nnz = A.NNZ()
    print(f"generated n={n} nnz={nnz} mean={nnz / n:.2f} in {time.time() - t0:.2f}s", flush=True)
    b = np.arange(1, n + 1, dtype=np.float64)
    tol = 1e-8 * float(np.sqrt(n * (n + 1) * (2 * n + 1) / 6.0))
    B_iter = 12 * nnz + 4 * (n + 1) + 88 * n
    # format:sigma[:panels]   (panels: 0 automatic, 1 none, P column panels)
    cfgs = os.environ.get("C5_CONFIGS", "csr:0,sell:32768:1,sell:32768:2,sell:32768:3,sell:32768:4,sell:32768:6,sell:32768:8")
    for cfg in cfgs.split(","):
        f = cfg.split(":")
        fmt, sigma, panels = {"sell": L.FMT_SELL, "dict": L.FMT_DICT}.get(f[0], L.FMT_CSR), int(f[1]), int(f[2]) if len(f) > 2 else 1
        A.set_format(L.FMT_CSR)
        if fmt == L.FMT_SELL:
            t0 = time.time()
            A.set_panels(panels)
            L.check(ctx.lib.cgb_csr_build_sell(A.h, sigma))
            A.set_format(L.FMT_SELL)
            ctx.sync()
            tb = time.time() - t0
        if fmt == L.FMT_DICT:
            A.set_format(L.FMT_DICT)
        for fused in [int(v) for v in os.environ.get("C5_FUSED", "0").split(",")]:
            cg = S.CG(ctx, fuse_p_update=fused, time_kernels=1)
            cg.setMatrix(A)
            cg.setTarget(b)
            cg.solve(tol)
            cg.x = None
            st = cg.solve(tol)
            sec = st["solve_ms"] / 1e3
            it = st["iterations"]
            rr, xx = cg.residual()
            extra = f" sell_padded={A.sell_padded() / nnz:.3f}x build={tb:.2f}s sigma={sigma} panels={A.panels()}" if fmt == L.FMT_SELL else ""
            print(f"fmt={ {L.FMT_SELL: 'SELL', L.FMT_DICT: 'DICT'}.get(fmt, 'CSR ') } fused={fused} it={it} {it / sec:8.1f} it/s {it * B_iter / sec / 1e9:7.1f} GB/s(B_iter) "
                  f"spmv {st['spmv_ms']:.4f} ms true_rel_res={np.sqrt(rr) / (tol / 1e-8):.2e}{extra}", flush=True)


if __name__ == "__main__":
    main()
