"""Host-side logic of the row-partitioned multi-GPU path, on CPU with world_size-2/3 gloo process groups:
row partitioning, local blocks with global columns, the halo plan of the library (cgb_plan_halo, host code
of libcgb200.so), send/receive symmetry between ranks, and a NumPy emulation of the distributed CG that
uses exactly that plan (interior/halo split, halo exchange, rank-ordered scalar all-reduce) against the
single-process oracle.  The CUDA kernels of the same path are covered by tests/test_dist_gpu.py."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def plan_halo(rank, world, starts, cols):
    lib = L.load()
    starts = np.ascontiguousarray(starts, dtype=np.int64)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    cnt = C.c_int64()
    per = np.zeros(world, dtype=np.int64)
    L.check(lib.cgb_plan_halo(rank, world, starts.ctypes.data, cols.size, cols.ctypes.data, None, 0, C.byref(cnt),
                              per.ctypes.data))
    recv = np.empty(cnt.value, dtype=np.int32)
    L.check(lib.cgb_plan_halo(rank, world, starts.ctypes.data, cols.size, cols.ctypes.data, recv.ctypes.data, recv.size,
                              C.byref(cnt), per.ctypes.data))
    return recv, per


def test_partition_rows():
    assert list(S.partition_rows(10, 3)) == [0, 3, 6, 10]
    assert list(S.partition_rows(512 ** 3, 8, align=512 * 512)) == [k * 64 * 512 * 512 for k in range(9)]
    st = S.partition_rows(512 ** 3, 8, align=512, edge_extra=int(1.25 * 512 * 512))    # edge ranks: one halo plane, more rows
    sz = np.diff(st)
    assert st[0] == 0 and st[-1] == 512 ** 3 and (sz % 512 == 0).all() and sz[0] == sz[-1]
    assert sz[0] - sz[1:-1].max() in (int(1.25 * 512) * 512 - 512, int(1.25 * 512) * 512) and sz[1:-1].max() - sz[1:-1].min() <= 512
    assert list(S.partition_rows(10, 2, edge_extra=4)) == [0, 5, 10]                     # two ranks: both are edges
    st = S.partition_rows(27, 4, align=9)                  # 3 planes over 4 ranks: one rank gets nothing
    assert st[0] == 0 and st[-1] == 27 and np.all(np.diff(st) >= 0) and np.all(st % 9 == 0)
    v, c, r = G.random_spd_powerlaw(5000, seed=1)
    st = S.partition_rows(5000, 4, rowptr=r)               # balanced by non-zeros
    nnz = np.diff(r[st])
    assert st[0] == 0 and st[-1] == 5000 and nnz.max() <= 1.1 * nnz.mean() + r.max() / 5000 * 64


def test_plan_halo_single_process():
    v, c, r = G.poisson3d_7pt(6)                           # 216 rows, planes of 36
    st = S.partition_rows(216, 3, align=36)
    lv, lc, lr = S.local_block(v, c, r, int(st[1]), int(st[2]))
    recv, per = plan_halo(1, 3, st, lc)
    # the middle slab needs exactly the plane below and the plane above
    assert list(per) == [36, 0, 36]
    assert np.array_equal(recv, np.concatenate([np.arange(36, 72), np.arange(144, 180)]))
    recv0, per0 = plan_halo(0, 3, st, S.local_block(v, c, r, 0, int(st[1]))[1])
    assert list(per0) == [0, 36, 0] and np.array_equal(recv0, np.arange(72, 108))


def _worker(rank, world, port, case, out_q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        gen, how = {"p3d7": (lambda: G.poisson3d_7pt(10), "rows"), "rspd": (lambda: G.random_spd_powerlaw(1500, seed=3), "nnz")}[case]
        v, c, r = gen()
        n = r.size - 1
        b = G.default_rhs(n)
        tol = 1e-9 * np.linalg.norm(b)
        starts = S.partition_rows(n, world, rowptr=r if how == "nnz" else None)
        lo, hi = int(starts[rank]), int(starts[rank + 1])
        lv, lc, lr = S.local_block(v, c, r, lo, hi)
        recv, per = plan_halo(rank, world, starts, lc)
        # every rank learns who needs what from it: the send lists are the other ranks' receive lists
        allrecv = [None] * world
        dist.all_gather_object(allrecv, recv)
        send = [np.array([g for g in allrecv[q] if lo <= g < hi], dtype=np.int64) for q in range(world)]
        counts = [None] * world
        dist.all_gather_object(counts, per)
        for q in range(world):                              # symmetry: what q expects from me is what I send to q
            assert send[q].size == counts[q][rank]
        assert send[rank].size == 0 and per[rank] == 0
        # interior / halo split exactly as dist.cu does it
        local = (lc >= lo) & (lc < hi)
        halo_slot = np.searchsorted(recv, lc[~local])
        rows_of = np.repeat(np.arange(hi - lo), np.diff(lr))

        def spmv(p_loc):
            halo = np.empty(recv.size)
            # halo exchange: gather what each rank asked for (object collectives keep the test simple)
            outgoing = [p_loc[send[q] - lo] for q in range(world)]
            gathered = [None] * world
            dist.all_gather_object(gathered, outgoing)
            off = 0
            for q in range(world):
                part = gathered[q][rank]
                halo[off:off + part.size] = part
                off += part.size
            y = np.zeros(hi - lo)
            np.add.at(y, rows_of[local], lv[local] * p_loc[lc[local] - lo])
            np.add.at(y, rows_of[~local], lv[~local] * halo[halo_slot])
            return y

        def allsum(x):
            parts = [None] * world
            dist.all_gather_object(parts, float(x))
            s = 0.0
            for q in range(world):                          # rank order, like the in-kernel all-reduce
                s += parts[q]
            return s

        x = np.zeros(hi - lo)
        rr_ = b[lo:hi] - spmv(x)
        p = np.zeros(hi - lo)
        rr = allsum(rr_ @ rr_)
        beta, it, done = 0.0, 0, False
        while not done and it < n + 1:                      # the reference's loop and stop rule (src/CG.hpp:359-436)
            p = rr_ + beta * p
            Ap = spmv(p)
            alpha = rr / allsum(p @ Ap)
            x += alpha * p
            rr_ -= alpha * Ap
            rr_new = allsum(rr_ @ rr_)
            if np.isnan(rr) or np.sqrt(rr) <= tol:
                done = True
            beta, rr = rr_new / rr, rr_new
            it += 1
        xs = [None] * world
        dist.all_gather_object(xs, x)
        if rank == 0:
            import oracle
            o = oracle.Port().cg_solve(v, c, r, b, tol)
            xg = np.concatenate(xs)
            err = np.linalg.norm(xg - o["x"]) / np.linalg.norm(o["x"])
            out_q.put((it, o["iterations"], err))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", ["p3d7", "rspd"])
def test_distributed_plan_emulation_gloo(world, case):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + world * 10 + (0 if case == "p3d7" else 1)
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    it, oit, err = q.get(timeout=10)
    assert abs(it - oit) <= max(1, round(0.03 * oit)) and err <= 1e-9
