"""Worker of tests/test_dist_gpu.py: one rank of a row-partitioned CG solve (run under torchrun).
Rank 0 gathers x and compares with the oracle; prints 'DIST_OK <case> ...' per passing case."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = S.Context(local)
    nccl_id = S.broadcast_nccl_id(ctx.lib, rank)
    cases = {
        "p3d7_24": (lambda: G.poisson3d_7pt(24), 1e-8, "rows"),
        "p2d_64": (lambda: G.poisson2d_5pt(64), 1e-8, "rows"),
        "p3d27_12": (lambda: G.stencil3d_27pt(12), 1e-8, "nnz"),
        "rspd_30000": (lambda: G.random_spd_powerlaw(30000, seed=4), 1e-11, "nnz"),  # x to 1e-10 needs a tighter stop here
    }
    dcg = None
    for name, (gen, rel, how) in cases.items():
        v, c, r = gen()
        n = r.size - 1
        b = G.default_rhs(n)
        tol = rel * np.linalg.norm(b)
        starts = S.partition_rows(n, world, rowptr=r if how == "nnz" else None)
        if name.startswith("rspd") and world > 1:   # odd first rows: the all-gather's 16-byte stores start one entry in
            starts = np.asarray(starts).copy()
            starts[1:-1] |= 1
        lo, hi = int(starts[rank]), int(starts[rank + 1])
        lv, lc, lr = S.local_block(v, c, r, lo, hi)
        A = S.Matrix(ctx, lv, lc, lr)
        want = os.environ.get("CGB_TEST_FORMAT", "")
        if want == "dict":                       # explicit storage option, inherited by the rank's local block
            try:
                A.set_format(L.FMT_DICT)
            except RuntimeError:
                want = ""                        # random SPD: as many patterns as rows, stays CSR / SELL
        if dcg is None:
            dcg = S.DistCG(ctx, rank, world, nccl_id, starts, A)
        else:
            dcg.set_matrix(starts, A)   # same communicator, new matrix
        bv, xv = S.Vector(ctx, b[lo:hi]), S.Vector(ctx, n=max(hi - lo, 1))
        if os.environ.get("CGB_TEST_HOST_JITTER"):
            import time
            time.sleep(float(np.random.default_rng(rank * 7919 + len(name)).uniform(0.0, 0.3)))
        st = dcg.solve(bv, xv, tol)
        expect = os.environ.get("CGB_TEST_EXPECT_LOOP", "")
        stencil = not name.startswith("rspd")
        if expect == "persistent" and name in ("p3d7_24", "p2d_64"):   # (a 27-point block one plane thick is mostly halo: it keeps the graph loop)
            assert st["loop_mode"] == L.LOOP_PERSISTENT, st
        if name.startswith("rspd") and os.environ.get("CGB_DIST_TRANSPORT", "") != "nccl":
            want_mode = 0 if os.environ.get("CGB_DIST_ALLGATHER", "") == "0" else 1   # dense coupling: all-gather unless switched off
            assert ctx.lib.cgb_dist_exchange_mode(dcg.h) == want_mode, (name, want_mode)
        if expect == "graph":
            assert st["loop_mode"] == L.LOOP_GRAPH, st
        if os.environ.get("CGB_DIST_JITTER_NS") and stencil:   # bitwise reproducible whatever the timing
            x_first = xv.download().copy()
            xv.upload(np.zeros(max(hi - lo, 1)))
            st2 = dcg.solve(bv, xv, tol)
            assert st2["iterations"] == st["iterations"] and st2["rr"] == st["rr"] and np.array_equal(xv.download(), x_first)
        rr, xx = dcg.residual(xv, bv)
        xl = xv.download()[: hi - lo]
        parts = [None] * world
        dist.all_gather_object(parts, xl)
        if rank == 0:
            import oracle
            x = np.concatenate(parts)
            o = oracle.Port().cg_solve(v, c, r, b, tol)
            err = np.linalg.norm(x - o["x"]) / np.linalg.norm(o["x"])
            slack = 0.05 if name.startswith("rspd") else 0.02
            ok_it = abs(st["iterations"] - o["iterations"]) <= max(1, round(slack * o["iterations"]))
            true_rr = float(np.sum((b - oracle.Port().spmv(v, c, r, x)) ** 2))
            ok = ok_it and err <= 1e-10 and st["converged"] == 1 and abs(rr - true_rr) <= (1e-13 / rel) * true_rr + 1e-30
            ok = ok and (want != "dict" or st["format"] == L.FMT_DICT)
            print(f"{'DIST_OK' if ok else 'DIST_FAIL'} {name} world={world} it={st['iterations']} oracle_it={o['iterations']} "
                  f"err={err:.2e} halo={dcg.halo_count()} rr={rr:.3e} true_rr={true_rr:.3e}", flush=True)
        dist.barrier()
    dcg.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
