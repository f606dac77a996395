#!/usr/bin/env python
"""Generates tests/golden/bench_parity.json: checksums of the REFERENCE's x after the first K_PARITY trips of
every bench.py workload (oracle/_ref: the reference's own CG::solve, left after k trips through the SYCL stand-in's
trip limit).  bench.py holds the GPU's x after the same k trips against these numbers at every GPU count, so the
driver's SCALE record carries a parity statement without a CPU leg on N > 1.

  python tools/make_bench_parity.py [workload ...]        (stencils: host generators, no GPU needed)
  python tools/make_bench_parity.py --device random_spd_10m   (C5: the matrix comes from the device generator)
  python tools/make_bench_parity.py --full [workload ...]    adds entry["full"]: the reference's CG::solve run to
      convergence at bench.py's tolerance (rel * |b|): its trip count and the same three checksums of its final x
      (C4: about 20 minutes and 20 GB of host memory on 8 threads)

TEST INFRASTRUCTURE: uses oracle/; never imported by the product."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import oracle  # noqa: E402


def main():
    argv = sys.argv[1:]
    device = "--device" in argv
    full = "--full" in argv
    names = [a for a in argv if not a.startswith("--")] or [w for w in bench.WORKLOADS if w != "random_spd_10m"]
    path = bench.GOLDEN_PARITY
    out = json.load(open(path)) if os.path.exists(path) else {}
    flavour = "omp" if oracle.have_ref("omp") else "serial"
    ref = oracle.Ref(flavour)
    for wl in names:
        kind, m, rel = bench.WORKLOADS[wl]
        if device:
            from conjugategradient_b200 import solver as S
            ctx = S.Context(0)
            A = S.Matrix.random_spd(ctx, m) if kind == "rspd" else S.Matrix.stencil(ctx, kind, m, 0 if kind == "2d5" else m)
            vals, cols, rows = A.to_host()
            A.free()
        else:
            assert kind != "rspd", "C5's matrix exists only on the device generator: run with --device on a GPU box"
            vals, cols, rows = bench.gen_host_csr(kind, m)
        n = rows.size - 1
        b = np.arange(1, n + 1, dtype=np.float64)
        h = ref.cg_create(vals, cols, rows)
        w = bench.counter_w(0, n)
        if full:  # the whole solve, as bench.py runs it on the GPU (src/CG.hpp:396-404, :436)
            tol = rel * float(np.linalg.norm(b))
            r = ref.cg_run(h, b, tol, want_x=True)
            ref.cg_destroy(h)
            x = r["x"]
            res = b - oracle.Port().spmv(vals, cols, rows, x)
            out.setdefault(wl, {})["full"] = {
                "iterations": int(r["iterations"]), "tol": tol, "x_norm2": float(np.linalg.norm(x)), "x_dot_w": float(x @ w),
                "true_residual_norm": float(np.linalg.norm(res)), "solve_s": r["solve_ms"] / 1e3,
                "source": f"oracle/_ref libref_{flavour}.so (reference CG::solve to convergence, {ref.threads} threads)"}
            print(wl, "full", out[wl]["full"], flush=True)
            json.dump(out, open(path, "w"), indent=1, sort_keys=True)
            del vals, cols, rows, x, w, res
            continue
        r = ref.cg_run(h, b, 0.0, trip_limit=bench.K_PARITY, want_x=True)
        ref.cg_destroy(h)
        x = r["x"]
        res = b - oracle.Port().spmv(vals, cols, rows, x)
        prev_full = out.get(wl, {}).get("full")
        out[wl] = {"k": int(r["iterations"]), "x_norm2": float(np.linalg.norm(x)), "x_dot_w": float(x @ w),
                   "true_residual_norm": float(np.linalg.norm(res)), "n": int(n), "nnz": int(vals.size),
                   "source": f"oracle/_ref libref_{flavour}.so (reference CG::solve, {ref.threads} threads), "
                             f"matrix from the {'device' if device else 'host'} generator"}
        if prev_full:
            out[wl]["full"] = prev_full
        print(wl, out[wl], flush=True)
        del vals, cols, rows, x, w, res
    json.dump(out, open(path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
