#!/usr/bin/env python
"""bench.py — CG iterations/s on the reference's headline workload (BASELINE.json):
3D 7-point Poisson 512^3, fp64 CSR, b = 1..n, x0 = 0, absolute tolerance 1e-8 * ||b||.

  python bench.py --gpus 1 --steps K --warmup W          one B200
  torchrun ... bench.py --gpus N ...                      row-partitioned over N B200 (strong scaling)
  python bench.py --impl reference ...                    the reference's own CPU solver (oracle/_ref)

A "step" is one CG solve of the system.  `value` = iterations / second with matrix and vectors resident in
HBM, run with the library's default options; `e2e` = the same through the host-buffer C-ABI call
(cgb_cg_solve_host: upload CSR + b from pinned host memory, solve, download x) inside the timed region.
The line also carries
  roofline      the SpMV + p.Ap kernel against the measured HBM peak (live CUDA-event time per launch),
  cpu_baseline  the reference's own CG::solve on the host cores (N = 1, rank 0; bounded sample of >= 32 trips),
  parity        checksums of x (||x||, x.w for a fixed counter-based w) after the full solve and after the first
                K_PARITY trips, the latter held against the reference's own x after the same trips (committed
                fixture tests/golden/bench_parity.json at every N; live element-wise comparison at N = 1),
  configs       one short run each of the other BASELINE configurations (C1, C2, C3, C5).
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cg_iterations_per_sec"
UNIT = "iterations/s"
WORKLOADS = {
    # name: (stencil kind, m, rel tol)
    "poisson3d_7pt_512": ("3d7", 512, 1e-8),   # BASELINE config C4 (north star)
    "poisson3d_7pt_128": ("3d7", 128, 1e-8),   # C2
    "stencil3d_27pt_256": ("3d27", 256, 1e-8), # C3
    "poisson2d_5pt_128": ("2d5", 128, 1e-8),   # C1
    "random_spd_10m": ("rspd", 10_000_000, 1e-8),  # C5: power-law row lengths (SELL-32-sigma engine)
}
BASELINE_NAME = {"poisson2d_5pt_128": "C1", "poisson3d_7pt_128": "C2", "stencil3d_27pt_256": "C3",
                 "poisson3d_7pt_512": "C4", "random_spd_10m": "C5"}
FULL_X_TOL = 1e-7      # checksums of the converged x against the reference's when the two stop a trip apart (both stop at rel 1e-8)
K_PARITY = 32          # trips of the first-k parity leg ("k ~ 32")
MIN_CPU_TRIPS = 32     # the reference's set-up (8 vector allocations + zero fills per solve) must not dominate its sample
PARITY_TOL = 1e-10     # north_star: solution within 1e-10 relative
GOLDEN_PARITY = os.path.join(ROOT, "tests", "golden", "bench_parity.json")


def b_iter(n, nnz):
    """Algorithmic bytes of one CG iteration."""
    return 12 * nnz + 4 * (n + 1) + 88 * n


def stencil_dims(kind, m):
    from conjugategradient_b200 import generators as G
    if kind == "rspd":
        return m, None  # nnz is generator-defined: taken from the generated matrix
    n = m * m if kind == "2d5" else m ** 3
    return n, G.stencil_nnz(kind, m)


def counter_w(lo, hi):
    """w[i] in [-1, 1): splitmix64 of the GLOBAL index i — the fixed vector of the x.w checksum, the same on
    any partition (tools/make_bench_parity.py uses this very function)."""
    out = np.empty(hi - lo)
    step = 1 << 24
    for a in range(lo, hi, step):
        b = min(hi, a + step)
        z = (np.arange(a, b, dtype=np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
        z ^= z >> np.uint64(30)
        z *= np.uint64(0xBF58476D1CE4E5B9)
        z ^= z >> np.uint64(27)
        z *= np.uint64(0x94D049BB133111EB)
        z ^= z >> np.uint64(31)
        out[a - lo:b - lo] = (z >> np.uint64(11)).astype(np.float64) * (2.0 ** -52) - 1.0
    return out


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:  # gone before anything else is timed: a query in flight holds driver locks that launches wait for
            self.proc.wait(timeout=3)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            self.proc.wait()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


L2_BYTES = 126e6  # B200 L2; decides which regime a one-kernel loop is in


def one_kernel_roofline(cluster, world, iterations, bytes_per_iteration, ms_per_launch, us_per_iteration, peak, peak_src):
    """Roofline block of a solve that ONE kernel (per rank) runs from start to end: its launch = all iterations.

    `achieved` and `frac` are PER GPU (every rank streams 1/world of the bytes, against one GPU's peak); the aggregate
    is kept beside them.  The regime follows from what one rank moves per iteration: below the L2 size the loop is
    latency-bound and the figure of merit is microseconds per iteration, above it the kernel is HBM-bound."""
    by = iterations * bytes_per_iteration
    agg = by / (ms_per_launch * 1e6)
    ach = agg / world
    per_rank_iter = bytes_per_iteration / world
    if cluster:
        kname = "cg_cluster_kernel (whole CG loop in one 16-CTA cluster, matrix and vectors in distributed shared memory)"
        regime = "latency-bound (working set in shared memory; no HBM traffic inside the loop)"
    else:
        kname = ("cg_persistent_kernel (whole CG loop, cooperative launch)" if world == 1 else
                 "cg_dist_persistent_kernel (whole row-partitioned CG loop, one cooperative kernel per rank: pack, TMA SpMV, "
                 "halo rows, peer all-reduces)")
        regime = ("latency-bound (working set in L2)" if per_rank_iter <= L2_BYTES else
                  f"hbm-bound ({per_rank_iter / 1e9:.2f} GB per rank and iteration, L2 holds {L2_BYTES / 1e6:.0f} MB)")
    return {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "per_gpu": True, "aggregate_GBps": agg, "n_gpus": world, "peak_source": peak_src, "traffic": None,
            "bytes_per_launch": by / world, "ms_per_launch": ms_per_launch, "share_of_iteration": 1.0,
            "us_per_iteration": us_per_iteration, "regime": regime}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs: copy, read + write bytes)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(workload):
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        t = json.load(open(p)).get(workload)
        return (t["bytes_per_launch"], t.get("source")) if t else (None, None)
    return None, None


_GATHER_PEAK = {}


def l2_roofline(nnz, k1_ms, rig):
    """The SELL engine on a random matrix is bound by the L2 -> SM path, not by HBM: every gathered double pulls a
    32-byte sector out of the L2-resident column panel of p.  Peak = tools/gather_bench.cu run live (random 8-byte
    gathers from a 40 MB vector, nothing else); achieved = (32 B per gather + the 12 B per entry of the matrix stream,
    which crosses the same path) / the kernel's time."""
    exe = os.path.join(ROOT, "tools", "_build", "gather_bench")
    if "v" not in _GATHER_PEAK:
        _GATHER_PEAK["v"] = None
        if os.path.exists(exe) and rig.world == 1:
            try:
                env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(rig.local_rank))
                out = subprocess.run([exe, "40", str(1 << 27), "3"], capture_output=True, text=True, timeout=120, env=env)
                _GATHER_PEAK["v"] = json.loads(out.stdout.strip().splitlines()[-1])
            except Exception:  # noqa: BLE001
                _GATHER_PEAK["v"] = None
    g = _GATHER_PEAK["v"]
    if not g:
        return None
    by = 32 * nnz + 12 * nnz
    ach = by / (k1_ms * 1e6)
    return {"bound": "l2_to_sm", "bytes_per_launch": by, "achieved": ach, "unit": "GB/s",
            "peak": g["l2_to_sm_sector_GBps"], "frac": ach / g["l2_to_sm_sector_GBps"],
            "gathers_per_s_kernel": nnz / (k1_ms * 1e-3), "gathers_per_s_peak": g["gathers_per_s"],
            "peak_source": "tools/gather_bench.cu, live: random 8-byte gathers out of a 40 MB vector, 32 B sector per gather"}


_STREAM_PEAKS = {}


def stream_peaks(rig):
    """HBM rates of this box on the access mixes the CG kernels have (tools/stream_bench.cu, run live once): read only,
    13 reads : 1 write (the SpMV kernel's mix) and a 1 : 1 copy.  The driver's MEASURED_PEAKS figure is a copy."""
    exe = os.path.join(ROOT, "tools", "_build", "stream_bench")
    if "v" not in _STREAM_PEAKS:
        _STREAM_PEAKS["v"] = None
        if os.path.exists(exe) and rig.world == 1:
            try:
                env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(rig.local_rank))
                out = subprocess.run([exe, "4", "3"], capture_output=True, text=True, timeout=120, env=env)
                _STREAM_PEAKS["v"] = json.loads(out.stdout.strip().splitlines()[-1])
            except Exception:  # noqa: BLE001
                _STREAM_PEAKS["v"] = None
    return _STREAM_PEAKS["v"]


def golden_parity(workload):
    if os.path.exists(GOLDEN_PARITY):
        return json.load(open(GOLDEN_PARITY)).get(workload)
    return None


def config_dict(workload, kind, m, n, nnz, rel, n_gpus):
    """The workload, in the SAME words in both arms (what ran on top of it is in the line's `run` block)."""
    return {"workload": workload, "baseline_config": BASELINE_NAME.get(workload),
            "matrix": f"{kind} stencil m={m}" if kind != "rspd" else
            f"random SPD, power-law row lengths (gamma 2.2, d_min 2, d_max 4096), n={m}", "n": n, "nnz": nnz,
            "format": "CSR fp64/int32", "rhs": "b[i]=i+1, x0=0",
            "tolerance": f"absolute {rel:g}*||b||2 on the recurrence residual (reference stop rule)",
            "step": "one CG solve of this system (GPU arm: the full solve; reference CPU arm: a bounded sample, the "
                    "first trips of the same solve, see cpu_baseline.sample)",
            "l2_policy": "inputs larger than L2 (CSR %.1f GB, vectors %.2f GB each)" % ((12 * nnz + 4 * n) / 1e9, 8 * n / 1e9)
            if 12 * nnz + 96 * n > 126e6 else "working set fits L2 (latency regime); no flush: every solve rewrites all its vectors",
            "B_iter_bytes": b_iter(n, nnz), "parallelism": f"row-partitioned x{n_gpus}" if n_gpus > 1 else "single GPU"}


# ------------------------------------------------------------------------------------------------
def gen_host_csr(kind, m):
    """Host CSR of the stencil, built directly (no lexsort) so 512^3 takes seconds, not minutes."""
    if kind == "rspd":  # NumPy twin of the device generator's distribution (same parameters, different RNG stream)
        from conjugategradient_b200 import generators as G
        return G.random_spd_powerlaw(m)
    n = m * m if kind == "2d5" else m ** 3
    if kind == "3d27" or m < 64:
        from conjugategradient_b200 import generators as G
        return {"3d7": G.poisson3d_7pt, "3d27": G.stencil3d_27pt, "2d5": G.poisson2d_5pt}[kind](m)
    idx = np.arange(n, dtype=np.int64)
    i, j = idx % m, (idx // m) % m
    k = idx // (m * m)
    offs = [(-m * m, k > 0), (-m, j > 0), (-1, i > 0), (0, None), (1, i < m - 1), (m, j < m - 1),
            (m * m, k < (m - 1))] if kind == "3d7" else [(-m, j > 0), (-1, i > 0), (0, None), (1, i < m - 1), (m, j < m - 1)]
    diag = 6.0 if kind == "3d7" else 4.0
    cnt = np.zeros(n, dtype=np.int32)
    for _, ok in offs:
        cnt += 1 if ok is None else ok.astype(np.int32)
    rows = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(cnt, out=rows[1:])
    nnz = int(rows[-1])
    cols = np.empty(nnz, dtype=np.int32)
    vals = np.empty(nnz, dtype=np.float64)
    pos = rows[:-1].copy()
    for off, ok in offs:
        sel = slice(None) if ok is None else ok
        p = pos[sel]
        cols[p] = (idx[sel] + off).astype(np.int32)
        vals[p] = diag if off == 0 else -1.0
        if ok is None:
            pos += 1
        else:
            pos[ok] += 1
    return vals, cols, rows.astype(np.int32)


def time_cpu(vals, cols, rows, b, target_seconds, steps, warmup, want_x_after=0):
    """Times the reference CPU solver on a bounded sample; returns iterations/s.

    Every timed step runs the first `trips` iterations of the solve, trips >= MIN_CPU_TRIPS whatever `steps` is, timed
    as test/Tester.cpp:43-46 times it (solve() alone, its own allocations included).  Warm-up steps run 2 trips (they
    only have to touch the matrix).  want_x_after = k > 0: one extra, untimed run of k trips returns the reference's x."""
    import oracle
    n = rows.size - 1
    if oracle.have_ref("omp"):
        ref = oracle.Ref("omp")
        h = ref.cg_create(vals, cols, rows)
        ncpu = os.cpu_count() or 1
        # the fastest thread count among {all, 1/2, 1/4} of the host cores (oversubscribed hosts differ)
        best_t, best_ms = None, None
        for t in sorted({ncpu, max(1, ncpu // 2), max(1, ncpu // 4)}, reverse=True):
            ref.set_threads(t)
            r2 = ref.cg_run(h, b, 0.0, trip_limit=2)["solve_ms"]
            r6 = ref.cg_run(h, b, 0.0, trip_limit=6)["solve_ms"]
            ms = max((r6 - r2) / 4.0, 1e-4)  # per trip, set-up cancelled
            if best_ms is None or ms < best_ms:
                best_t, best_ms = t, ms
        ref.set_threads(best_t)
        full = n + 1
        trips = int(min(full, max(MIN_CPU_TRIPS, (target_seconds * 1e3 / max(1, steps)) / best_ms)))
        trips = min(trips, 2000)
        for _ in range(warmup):
            ref.cg_run(h, b, 0.0, trip_limit=2)
        its, ms = 0, 0.0
        for _ in range(steps):
            r = ref.cg_run(h, b, 0.0, trip_limit=trips)
            its += r["iterations"]
            ms += r["solve_ms"]
        x_ref = None
        if want_x_after > 0:
            x_ref = ref.cg_run(h, b, 0.0, trip_limit=want_x_after, want_x=True)["x"]
        ref.cg_destroy(h)
        cores = best_t
        what = "reference CG::solve (unmodified sources, OpenMP SYCL stand-in)"
    else:
        port = oracle.Port()
        trips, cores = MIN_CPU_TRIPS, 1
        t0 = time.time()
        its = port.cg_solve(vals, cols, rows, b, 0.0, max_iter=trips)["iterations"]
        ms = (time.time() - t0) * 1e3
        steps = 1
        x_ref = port.cg_solve(vals, cols, rows, b, 0.0, max_iter=want_x_after)["x"] if want_x_after > 0 else None
        what = "C restatement oracle/cg_oracle.c (serial)"
    return {"value": its / (ms / 1e3), "ms_per_step": ms / max(steps, 1), "cores": cores, "trips": trips, "x_ref": x_ref,
            "sample": f"first {trips} CG iterations per step of the same system (n={n}), solve() timed as "
                      f"test/Tester.cpp:43-46 (its per-solve allocations included), {what}, {cores} threads"}


def run_reference(args, kind, m, rel):
    """The reference's own CPU CG (unmodified sources + OpenMP SYCL stand-in, oracle/_ref) on the box's
    host cores.  Each step is a bounded sample: the first `trips` (>= 32) iterations of the same system."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    from conjugategradient_b200 import generators as G
    n, nnz = stencil_dims(kind, m)
    t0 = time.time()
    vals, cols, rows = gen_host_csr(kind, m)
    nnz = int(vals.size)
    b = G.default_rhs(n)
    kind_ref = "reference" if oracle.have_ref("omp") else "port"
    out = time_cpu(vals, cols, rows, b, target_seconds=args.cpu_seconds, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": out["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": out["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args.workload, kind, m, n, nnz, rel, args.gpus),
        "cpu_baseline": {"value": out["value"], "unit": UNIT, "cores": out["cores"], "kind": kind_ref,
                         "sample": out["sample"]},
        "e2e": {"value": out["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "run": {"where": "host cores of the GPU box", "trips_per_step": out["trips"], "setup_s": round(time.time() - t0, 1)},
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
class Rig:
    """What every workload of one bench process shares: context, library, ranks, collectives for the timing."""

    def __init__(self, args):
        from conjugategradient_b200 import solver as S
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if args.gpus > 1 and self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={self.world})")
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(self.local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist
        self.ctx = S.Context(self.local_rank)
        self.lib = self.ctx.lib
        self.dh = None  # cgb_dist handle, created by the first workload, re-used (set_matrix) by the others

    def barrier(self):
        self.ctx.sync()
        if self.dist is not None:
            import torch
            self.dist.barrier()
            torch.cuda.synchronize()

    def _reduce(self, v, op):
        if self.dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(self, v):
        return self._reduce(v, self.dist.ReduceOp.MAX) if self.dist is not None else v

    def sum_over_ranks(self, v):
        return self._reduce(v, self.dist.ReduceOp.SUM) if self.dist is not None else v

    def timed(self, fn, steps):
        """K steps bracketed by barrier+sync, CUDA events on the library's stream, max over ranks."""
        from conjugategradient_b200 import _lib as L
        lib, ctx = self.lib, self.ctx
        self.barrier()
        e0, e1 = C.c_void_p(), C.c_void_p()
        L.check(lib.cgb_event_record(ctx.h, C.byref(e0)))
        outs = [fn() for _ in range(steps)]
        L.check(lib.cgb_event_record(ctx.h, C.byref(e1)))
        ms = C.c_double()
        L.check(lib.cgb_event_elapsed_ms(e0, e1, C.byref(ms)))
        self.barrier()
        lib.cgb_event_destroy(e0)
        lib.cgb_event_destroy(e1)
        return self.max_over_ranks(ms.value), outs

    def close(self):
        if self.dh is not None:
            self.lib.cgb_dist_destroy(self.dh)
            self.dh = None
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def run_workload(rig, args, workload, steps, warmup, e2e_steps, cpu_seconds, headline):
    """One BASELINE configuration on this rig; returns the measurement blocks (rank 0 uses them)."""
    from conjugategradient_b200 import _lib as L
    from conjugategradient_b200 import generators as G
    from conjugategradient_b200 import solver as S

    kind, m, rel = WORKLOADS[workload]
    if args.m and headline:
        m = args.m
    ctx, lib, world, rank = rig.ctx, rig.lib, rig.world, rig.rank
    n, nnz = stencil_dims(kind, m)
    if kind == "rspd":
        row_starts = S.partition_rows(n, world)
    else:
        plane = m if kind == "2d5" else m * m
        # contiguous slabs of grid planes per rank; CGB_BENCH_EDGE_PLANES (fraction of a plane allowed, cut along grid
        # lines): extra rows for the two edge ranks, which have one halo plane instead of two
        edge = float(os.environ.get("CGB_BENCH_EDGE_PLANES", "0"))
        if edge > 0 and kind != "2d5" and world >= 3:
            row_starts = S.partition_rows(n, world, align=m, edge_extra=int(edge * plane))
        else:
            row_starts = S.partition_rows(n, world, align=plane)
    lo, hi = int(row_starts[rank]), int(row_starts[rank + 1])
    nloc = hi - lo
    if kind == "rspd":
        A = S.Matrix.random_spd(ctx, n, row_begin=lo, row_end=hi)
    else:
        A = S.Matrix.stencil(ctx, kind, m, 0 if kind == "2d5" else m, lo, hi)
    b_host = np.arange(lo + 1, hi + 1, dtype=np.float64)
    nb = float(np.sqrt(n * (n + 1) * (2 * n + 1) / 6.0))  # ||1..n||_2 in closed form
    tol = rel * nb
    bvec = S.Vector(ctx, b_host)
    xvec = S.Vector(ctx, n=nloc)
    small = n <= 131072  # latency regime: the library's AUTO loop is one cluster-resident / cooperative kernel per solve
    opts = S.make_opts(check_every=args.check_every)  # the library's defaults (pipelined graph batches)
    stats = L.CgStats()

    if world > 1:
        if os.environ.get("CGB_BENCH_FORMAT", "") == "dict":  # debug: the opt-in storage format on every rank's block
            A.set_format(L.FMT_DICT)
        if rig.dh is None:
            idbytes = S.broadcast_nccl_id(lib, rank)
            dh = C.c_void_p()
            L.check(lib.cgb_dist_create(ctx.h, rank, world, idbytes, row_starts.ctypes.data_as(C.c_void_p), A.h, C.byref(dh)))
            rig.dh = dh
        else:
            L.check(lib.cgb_dist_set_matrix(rig.dh, row_starts.ctypes.data_as(C.c_void_p), A.h))
    dh = rig.dh if world > 1 else None

    def solve_resident(o=opts, max_iter=-1):
        L.check(lib.cgb_memset(ctx.h, xvec.ptr, 0, 8 * nloc))
        if dh is None:
            L.check(lib.cgb_cg_solve(ctx.h, A.h, bvec.ptr, xvec.ptr, tol, max_iter, C.byref(o), C.byref(stats)))
        else:
            L.check(lib.cgb_dist_cg_solve(dh, bvec.ptr, xvec.ptr, tol, max_iter, C.byref(o), C.byref(stats)))
        return stats.as_dict()

    def residual():
        rr, xx = C.c_double(), C.c_double()
        if dh is None:
            L.check(lib.cgb_residual(ctx.h, A.h, xvec.ptr, bvec.ptr, C.byref(rr), C.byref(xx)))
        else:
            L.check(lib.cgb_dist_residual(dh, xvec.ptr, bvec.ptr, C.byref(rr), C.byref(xx)))
        return rr.value, xx.value

    if nnz is None:
        nnz = int(rig.sum_over_ranks(float(A.NNZ())))

    # ---- device-resident solves, default options --------------------------------------------------
    for _ in range(warmup):
        solve_resident()
    sampler = ClockSampler(rig.local_rank) if headline and not os.environ.get("CGB_BENCH_NO_CLOCKS") else None
    if sampler and rank == 0:
        sampler.start()
    ms_total, outs = rig.timed(solve_resident, steps)
    clocks = sampler.stop() if (sampler and rank == 0) else None
    iters = sum(o["iterations"] for o in outs)
    value = iters / (ms_total / 1e3)
    if os.environ.get("CGB_BENCH_TRACE") and rank == 0:  # device time inside each solve against the whole timed region
        print(f"[bench trace] {workload}: timed region {ms_total:.2f} ms, solves " +
              " ".join(f"{o['solve_ms']:.2f}" for o in outs), file=sys.stderr, flush=True)
    launches = rig.sum_over_ranks(float(sum(o["kernel_launches"] for o in outs)))
    last = outs[-1]
    rr_full, xx_full = residual()  # true residual of the last solve (global)
    true_rel = float(np.sqrt(rr_full)) / nb

    # ---- parity: checksums of x, full solve and first K_PARITY trips --------------------------------
    wvec = S.Vector(ctx, counter_w(lo, hi))
    w_norm = float(np.sqrt(rig.sum_over_ranks(float(np.sum(np.square(wvec.download()))))))
    acc = S.Scalar(ctx, 0.0)

    def x_dot_w():
        acc.set(0.0)
        L.check(lib.cgb_dot_acc(ctx.h, nloc, xvec.ptr, wvec.ptr, acc.ptr))
        return rig.sum_over_ranks(acc.value)
    parity = {"w": "w[i] = splitmix64(i) mapped to [-1, 1), i the global row", "x_norm2": float(np.sqrt(xx_full)),
              "x_dot_w": x_dot_w(), "iterations": last["iterations"], "true_rel_residual": true_rel}
    st_k = solve_resident(max_iter=K_PARITY)
    rr_k, xx_k = residual()
    parity["first_k"] = {"k": int(st_k["iterations"]), "x_norm2": float(np.sqrt(xx_k)), "x_dot_w": x_dot_w(),
                         "true_residual_norm": float(np.sqrt(rr_k)), "recurrence_rr": st_k["rr"]}
    gold = golden_parity(workload) if not (args.m and headline) else None
    checks = []
    if gold and gold.get("k") == parity["first_k"]["k"]:
        fk = parity["first_k"]
        d_norm = abs(fk["x_norm2"] - gold["x_norm2"]) / gold["x_norm2"]
        d_dot = abs(fk["x_dot_w"] - gold["x_dot_w"]) / (gold["x_norm2"] * w_norm)
        d_res = abs(fk["true_residual_norm"] - gold["true_residual_norm"]) / gold["true_residual_norm"]
        parity["vs_reference_fixture"] = {"fixture": "tests/golden/bench_parity.json (the reference's own x after k trips; tools/make_bench_parity.py)",
                                          "x_norm2_rel_diff": d_norm, "x_dot_w_diff_over_norms": d_dot,
                                          "true_residual_rel_diff": d_res}
        checks += [d_norm <= PARITY_TOL, d_dot <= PARITY_TOL, d_res <= 1e-8]
    if gold and gold.get("full"):
        # the reference's own CG::solve run to convergence at this tolerance (tools/make_bench_parity.py --full):
        # north_star's "iteration count within +-2 %", and the checksums of the two final x.  Both solves stop on
        # |r| <= 1e-8 |b|, possibly a trip apart, so the final x agree to about the tolerance, not to rounding.
        gf = gold["full"]
        it_ok = abs(parity["iterations"] - gf["iterations"]) <= max(1, int(round(0.02 * gf["iterations"])))
        x_tol = PARITY_TOL if parity["iterations"] == gf["iterations"] else FULL_X_TOL  # same trip count: same iterate to rounding
        f_norm = abs(parity["x_norm2"] - gf["x_norm2"]) / gf["x_norm2"]
        f_dot = abs(parity["x_dot_w"] - gf["x_dot_w"]) / (gf["x_norm2"] * w_norm)
        parity["full_solve"] = {"iterations": parity["iterations"], "iterations_ref": gf["iterations"], "iterations_ok": it_ok,
                                "x_norm2_rel_diff": f_norm, "x_dot_w_diff_over_norms": f_dot, "x_tol": x_tol,
                                "fixture": "tests/golden/bench_parity.json[full] (reference CG::solve to convergence)"}
        checks += [it_ok, f_norm <= x_tol, f_dot <= x_tol]
        parity["full_solve"]["ok"] = bool(it_ok and f_norm <= x_tol and f_dot <= x_tol)
    x_k_host = xvec.download() if (world == 1 and not args.no_cpu) else None  # for the live comparison below
    wvec.free()
    acc.free()

    # ---- roofline of the dominant kernel (SpMV + p.Ap): a separate short solve with per-kernel CUDA-event brackets ----
    peak, peak_src = measured_peak()
    roof = None
    if last["loop_mode"] in (L.LOOP_PERSISTENT, L.LOOP_CLUSTER):
        roof = one_kernel_roofline(last["loop_mode"] == L.LOOP_CLUSTER, world, last["iterations"], b_iter(n, nnz),
                                   ms_total / steps, 1e3 * ms_total / max(iters, 1), peak, peak_src)
    elif world == 1:
        t_opts = S.make_opts(time_kernels=1, check_every=args.check_every)
        cap = -1 if last["iterations"] <= 16 * args.check_every else 16 * args.check_every
        t_out = solve_resident(t_opts, max_iter=cap)
        if t_out["timed_samples"] > 0:
            k1_bytes = 12 * nnz + 4 * (n + 1) + 16 * n  # vals + cols + rowptr + p read + Ap write
            k1_ms = t_out["spmv_ms"]
            ach = k1_bytes / (k1_ms * 1e6)
            eng = "SELL-32-sigma copy, warp per slice" if last["format"] == L.FMT_SELL else "CSR, TMA-staged tiles"
            it_gbps = iters * b_iter(n, nnz) / (ms_total * 1e6)
            roof = {"bound": "hbm", "kernel": f"cg_spmv_kernel (SpMV + p.Ap; {eng})", "achieved": ach, "peak": peak,
                    "unit": "GB/s", "frac": ach / peak, "peak_source": peak_src, "traffic": ncu_traffic(workload)[0],
                    "traffic_source": ncu_traffic(workload)[1],
                    "bytes_per_launch": k1_bytes, "ms_per_launch": k1_ms, "timed_launches": t_out["timed_samples"],
                    "timing": "CUDA events around the kernel inside a separate solve (opts.time_kernels; the headline solves run untimed, pipelined)",
                    "share_of_iteration": k1_ms / (ms_total / max(iters, 1)),
                    "update_ms": t_out["update_ms"], "pupdate_ms": t_out["pupdate_ms"],
                    "frac_of_8TBps_spec": ach / 8000.0,
                    "iteration_GBps_B_iter": it_gbps, "iteration_frac_of_peak": it_gbps / peak,
                    "iteration_frac_of_8TBps": it_gbps / 8000.0,
                    "l2": l2_roofline(nnz, k1_ms, rig) if last["format"] == L.FMT_SELL else None,
                    "note": "frac can exceed 1: the peak is a COPY (half reads, half writes); this kernel's traffic is 93 % reads, "
                            "which HBM3e serves faster than a 50/50 mix. The spec-based fraction is frac_of_8TBps_spec; "
                            "stream_peaks holds this box's own streaming rates measured in this run (tools/stream_bench.cu); "
                            "frac_of_read_only_stream is against the read-only one, an upper bound for any mix."}
            if headline:
                sp = stream_peaks(rig)
                roof["stream_peaks"] = sp
                if sp and sp.get("read_only_GBps"):  # the tightest ceiling the probe gives: nothing but reads (LDG.128 or TMA bulk)
                    roof["frac_of_read_only_stream"] = ach / max(sp["read_only_GBps"], sp.get("read_only_tma_GBps") or 0.0)

    # ---- opt-in storage format (not the headline): the same solve through the row-pattern dictionary copy ----
    # (CGB_FMT_DICT, include/cgb200.h: 2 bytes per ROW instead of 12 per non-zero for constant-coefficient stencils;
    # same arithmetic, bit-identical row sums).  Reported beside the CSR number, never instead of it.
    extra_fmt = None
    if headline and world == 1 and kind != "rspd" and not args.no_extra and not small:
        try:
            A.set_format(L.FMT_DICT)
            solve_resident()
            d_ms, d_outs = rig.timed(solve_resident, 1)
            rr2, _ = residual()
            d_it = d_outs[0]["iterations"]
            extra_fmt = {"format": "row-pattern dictionary (CGB_FMT_DICT, explicit option; the headline value streams the CSR arrays)",
                         "value": d_it / (d_ms / 1e3), "unit": UNIT, "iterations_per_solve": d_it,
                         "true_rel_residual": float(np.sqrt(rr2)) / nb, "patterns": A.dict_patterns(),
                         "matrix_bytes": A.dict_bytes(), "csr_bytes": 12 * nnz + 4 * (n + 1),
                         "bytes_per_iteration_moved": A.dict_bytes() + 80 * n,
                         "GBps_moved": d_it * (A.dict_bytes() + 80 * n) / (d_ms * 1e6)}
        except L.CgbError as e:
            extra_fmt = {"format": "row-pattern dictionary", "unavailable": str(e)}
        A.set_format(L.FMT_AUTO)

    # ---- end to end: host buffers in, host buffer out -----------------------------------------------
    e2e = None
    pinned_handles = []
    pv = pc = pr = None
    if not args.no_e2e:
        hv, hc, hr = A.to_host()  # this rank's block (global columns), then pinned
        nnz_loc = hv.size

        def pinned(a):
            p = C.c_void_p()
            L.check(lib.cgb_host_alloc(a.nbytes, C.byref(p)))
            pinned_handles.append(p)
            out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(a.nbytes,)).view(a.dtype)
            out[...] = a
            return out
        pv, pc, pr, pb, px = pinned(hv), pinned(hc), pinned(hr), pinned(b_host), pinned(np.zeros(nloc))
        del hv, hc, hr
        e_opts = S.make_opts(check_every=args.check_every)
        e_stats = L.CgStats()

        def solve_e2e():
            if dh is None:
                L.check(lib.cgb_cg_solve_host(ctx.h, nloc, nnz_loc, pv.ctypes.data, pc.ctypes.data, pr.ctypes.data,
                                              pb.ctypes.data, None, px.ctypes.data, tol, -1, C.byref(e_opts), C.byref(e_stats)))
            else:
                trace = rank == 0 and os.environ.get("CGB_DIST_TIMING")  # debug: wall time of each call (adds syncs)
                t_prev = [time.perf_counter()]

                def mark(what):
                    if trace:
                        ctx.sync()
                        now = time.perf_counter()
                        print(f"[bench e2e timing] {what:<28s} {1e3 * (now - t_prev[0]):8.2f} ms", file=sys.stderr, flush=True)
                        t_prev[0] = now
                a2 = C.c_void_p()
                L.check(lib.cgb_csr_create(ctx.h, nloc, nnz_loc, pv.ctypes.data, pc.ctypes.data, pr.ctypes.data, C.byref(a2)))
                mark("cgb_csr_create (H2D)")
                L.check(lib.cgb_dist_set_matrix(dh, row_starts.ctypes.data_as(C.c_void_p), a2))
                mark("cgb_dist_set_matrix")
                lib.cgb_csr_destroy(a2)
                mark("cgb_csr_destroy")
                L.check(lib.cgb_memcpy_h2d(ctx.h, bvec.ptr, pb.ctypes.data, 8 * nloc, 0))
                L.check(lib.cgb_memset(ctx.h, xvec.ptr, 0, 8 * nloc))
                L.check(lib.cgb_dist_cg_solve(dh, bvec.ptr, xvec.ptr, tol, -1, C.byref(e_opts), C.byref(e_stats)))
                mark("b H2D + cgb_dist_cg_solve")
                L.check(lib.cgb_memcpy_d2h(ctx.h, px.ctypes.data, xvec.ptr, 8 * nloc, 1))
                mark("x D2H")
            return e_stats.as_dict()

        A.free()  # make room: the e2e path uploads its own copy every step (N > 1: the split copies live in the dist handle)
        e_steps = max(1, min(steps, e2e_steps))
        for _ in range(min(warmup, 1)):
            solve_e2e()
        e_ms, e_outs = rig.timed(solve_e2e, e_steps)
        e_iters = sum(o["iterations"] for o in e_outs)
        h2d = rig.sum_over_ranks(float(12 * nnz_loc + 4 * (nloc + 1) + 8 * nloc))
        d2h = rig.sum_over_ranks(float(8 * nloc))
        e2e = {"value": e_iters / (e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "steps": e_steps, "ms_per_step": e_ms / e_steps,
               "api": "cgb_cg_solve_host (pinned host CSR + b in, x out)" if dh is None else
                      "cgb_csr_create + cgb_dist_set_matrix + cgb_dist_cg_solve + D2H per rank"}

    # ---- CPU baseline: the reference's own solver on the host cores (N = 1, rank 0) ----------------
    cpu = None
    if world == 1 and not args.no_cpu:
        # the very matrix the GPU solved (downloaded for the e2e leg), else the host generator's
        vals, cols, rows = (pv, pc, pr) if e2e is not None else (A.to_host() if kind == "rspd" else gen_host_csr(kind, m))
        b_full = G.default_rhs(n)
        import oracle
        out = time_cpu(vals, cols, rows, b_full, target_seconds=cpu_seconds, steps=1, warmup=0, want_x_after=K_PARITY)
        cpu = {"value": out["value"], "unit": UNIT, "cores": out["cores"],
               "kind": "reference" if oracle.have_ref("omp") else "port", "sample": out["sample"]}
        if out["x_ref"] is not None and x_k_host is not None:
            xr = out["x_ref"]
            err = float(np.linalg.norm(x_k_host - xr) / np.linalg.norm(xr))
            parity["vs_reference_live"] = {"what": f"x after the first {K_PARITY} trips, GPU (max_iter = k) against the reference's "
                                                   "CG::solve left after k trips, element-wise, same run", "x_rel_err": err}
            checks.append(err <= PARITY_TOL)
        del vals, cols, rows
    for hnd in pinned_handles:
        lib.cgb_host_free(hnd)
    pv = pc = pr = None
    parity["tolerance"] = PARITY_TOL
    parity["ok"] = bool(checks) and all(checks)
    if not checks:
        parity["ok"] = None
        parity["note"] = "no reference fixture for this workload and no CPU leg in this run"

    bvec.free()
    xvec.free()
    A.free()
    lib.cgb_trim()  # the library caches freed device blocks for re-use: hand them back before the next workload
    run = {"where": f"{world}xB200", "iterations_per_solve": last["iterations"], "true_rel_residual": true_rel,
           "loop": ("one thread-block cluster per solve (cluster barriers, distributed shared memory)" if last["loop_mode"] == L.LOOP_CLUSTER
                    else "one persistent cooperative kernel per solve (grid-wide barriers)" if last["loop_mode"] == L.LOOP_PERSISTENT
                    else "CUDA graph batches of %d iterations, device-side stop flag, polled one batch behind" % args.check_every),
           "engine": {L.FMT_SELL: "SELL-32-sigma", L.FMT_DICT: "row-pattern dictionary (explicit option)"}.get(last["format"], "CSR TMA-staged"),
           "transport": (os.environ.get("CGB_DIST_TRANSPORT", "peer") if world > 1 else None),
           "exchange": (None if world == 1 else "all-gather of p over NVLink (dense coupling)" if lib.cgb_dist_exchange_mode(rig.dh) == 1
                        else "halo entries only")}
    return {"kind": kind, "m": m, "rel": rel, "n": n, "nnz": nnz, "value": value, "ms_total": ms_total, "steps": steps,
            "warmup": warmup, "iters": iters, "launches": int(launches), "clocks": clocks, "roofline": roof, "e2e": e2e,
            "cpu_baseline": cpu, "parity": parity, "run": run, "opt_in_format": extra_fmt,
            "achieved_GBps_B_iter": iters * b_iter(n, nnz) / (ms_total * 1e6)}


def run_ours(args):
    rig = Rig(args)
    world, rank = rig.world, rig.rank
    head = run_workload(rig, args, args.workload, args.steps, args.warmup, args.e2e_steps, args.cpu_seconds, headline=True)
    # ---- the other BASELINE configurations: one short run each (north_star: "throughput on synthetic stencil and
    # random-SPD matrices at 1 GPU and at 2/4/8 GPUs") --------------------------------------------------------
    others = {}
    if args.configs != "none":
        names = ["poisson2d_5pt_128", "poisson3d_7pt_128", "stencil3d_27pt_256", "random_spd_10m"] if world == 1 else \
                ["poisson3d_7pt_128", "random_spd_10m"]
        if args.configs != "all":
            names = [w for w in args.configs.split(",") if w in WORKLOADS]
        for wl in names:
            if wl == args.workload:
                continue
            t0 = time.time()
            try:
                r = run_workload(rig, args, wl, 3, 3, 2, min(args.cpu_seconds, 6.0), headline=False)
            except Exception as e:  # noqa: BLE001 — the headline has been measured: report it, say what failed, stop here
                others[BASELINE_NAME[wl]] = {"error": f"{type(e).__name__}: {e}"[:400]}
                break
            entry = {"config": config_dict(wl, r["kind"], r["m"], r["n"], r["nnz"], r["rel"], world), "value": r["value"],
                     "unit": UNIT, "steps": r["steps"], "warmup": r["warmup"], "ms_per_step": r["ms_total"] / r["steps"],
                     "iterations_per_solve": r["run"]["iterations_per_solve"], "true_rel_residual": r["run"]["true_rel_residual"],
                     "achieved_GBps_B_iter": r["achieved_GBps_B_iter"], "roofline": r["roofline"], "e2e": r["e2e"],
                     "cpu_baseline": r["cpu_baseline"], "parity": r["parity"], "run": r["run"], "wall_s": round(time.time() - t0, 1)}
            others[BASELINE_NAME[wl]] = entry
    # ---- the drop-in C++ headers end to end from pageable std::vector storage (C2) ----------
    e2e_cpp = None
    exe = os.path.join(ROOT, "examples", "e2e_cpp")
    if world == 1 and not args.no_e2e and args.configs != "none" and os.path.exists(exe):
        try:
            out = subprocess.run([exe, "128", "3", "1"], capture_output=True, text=True, timeout=300)
            e2e_cpp = json.loads(out.stdout.strip().splitlines()[-1])
            e2e_cpp["api"] = ("include/CG.hpp: createCG + setMatrix(std::vector) + setTarget + solve + extract per step, pageable host "
                              "memory (examples/e2e_cpp.cpp; reference protocol test/Tester.cpp:36-49), wall clock")
            e2e_cpp["config"] = "C2 (3D 7-pt 128^3)"
            pinned = (others.get("C2") or {}).get("e2e") or {}
            if pinned.get("value"):
                e2e_cpp["vs_pinned_c_abi_e2e"] = e2e_cpp["value"] / pinned["value"]
        except Exception as e:  # noqa: BLE001
            e2e_cpp = {"unavailable": f"{type(e).__name__}: {e}"}
    if rank == 0:
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_total"] / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": config_dict(args.workload, head["kind"], head["m"], head["n"], head["nnz"], head["rel"], world),
            "run": head["run"], "roofline": head["roofline"], "cpu_baseline": head["cpu_baseline"], "e2e": head["e2e"],
            "parity": head["parity"], "gpu_launches": head["launches"], "clocks": head["clocks"],
            "achieved_GBps_B_iter": head["achieved_GBps_B_iter"], "opt_in_format": head["opt_in_format"],
            "configs": others, "e2e_cpp": e2e_cpp,
        }
        emit(line)
    rig.close()


_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line, on the process's real stdout (see main)."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    # Libraries write to fd 1 behind Python's back (NCCL prints "NCCL version ..." there on multi-GPU runs): keep
    # the contract of exactly one JSON line on stdout by pointing fd 1 at stderr for the duration of the run.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("CGB_BENCH_WORKLOAD", "poisson3d_7pt_512"), choices=sorted(WORKLOADS))
    ap.add_argument("--m", type=int, default=int(os.environ.get("CGB_BENCH_M", "0")), help="override grid size (debug)")
    ap.add_argument("--check-every", type=int, default=32)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=20.0)
    ap.add_argument("--configs", default=os.environ.get("CGB_BENCH_CONFIGS", "all"),
                    help="'all' (default), 'none', or a comma list of workloads for the `configs` block")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the opt-in storage-format leg")
    args = ap.parse_args()
    if args.impl == "reference":
        kind, m, rel = WORKLOADS[args.workload]
        run_reference(args, kind, m if not args.m else args.m, rel)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
