"""bench.py prints exactly ONE JSON line on stdout carrying the keys the driver reads (both arms)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"}


def run_bench(*args, timeout=600):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                         timeout=timeout, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-3000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, out.stdout[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line():
    """The reference arm: the reference's own CPU solver (oracle/_ref, else the C restatement) on the host cores."""
    d = run_bench("--impl", "reference", "--workload", "poisson2d_5pt_128", "--steps", "1", "--warmup", "1")
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "cg_iterations_per_sec" and d["unit"] == "iterations/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 1 and d["dtype"] == "f64"
    assert d["config"]["workload"] == "poisson2d_5pt_128" and d["config"]["n"] == 16384
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    assert d["run"]["trips_per_step"] >= 32          # the reference's per-solve set-up must not dominate the sample
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


@pytest.mark.gpu
def test_our_arm_line():
    """Our arm on the small configuration: roofline, cpu_baseline, e2e, launch count and clocks are all there."""
    d = run_bench("--workload", "poisson2d_5pt_128", "--steps", "3", "--warmup", "3", "--cpu-seconds", "2", "--configs", "none")
    assert BASE_KEYS | {"roofline", "gpu_launches", "clocks"} <= set(d) and d.get("impl", "ours") == "ours"
    assert d["value"] > 0 and d["n_gpus"] == 1 and d["gpu_launches"] > 0
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and r["peak"] > 0 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    cb = d["cpu_baseline"]
    assert cb["value"] > 0 and cb["cores"] >= 1 and cb["kind"] in ("reference", "port") and cb["sample"]
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] == 8 * d["config"]["n"]
    assert d["run"]["iterations_per_solve"] == 366 and d["run"]["true_rel_residual"] <= 2e-8
    assert 0 < r["frac"] <= 1.2
    # parity block: checksums of x after the full solve and after the first k trips, held against the reference
    p = d["parity"]
    assert p["ok"] is True and p["first_k"]["k"] == 32 and p["vs_reference_live"]["x_rel_err"] <= 1e-10
    assert p["vs_reference_fixture"]["x_norm2_rel_diff"] <= 1e-10
    assert d["configs"] == {}            # --configs none below


def test_both_arms_describe_the_same_config():
    """`config` is the workload in the same words in both arms (the driver compares them)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    c = b.config_dict("poisson3d_7pt_512", "3d7", 512, 512 ** 3, 937951232, 1e-8, 1)
    assert c["B_iter_bytes"] == 23603445764 and c["baseline_config"] == "C4"
    assert not ({"where", "iterations_per_solve", "true_rel_residual", "loop", "engine", "transport"} & set(c))   # per-run facts live in `run`
    w = b.counter_w(5, 1000)
    assert w.min() >= -1.0 and w.max() < 1.0 and abs(w.mean()) < 0.1 and (b.counter_w(0, 1000)[5:] == w).all()


def test_one_kernel_roofline_is_per_gpu():
    """At N > 1 the roofline block of the one-kernel loop is per GPU (round 2 divided the aggregate bytes by ONE GPU's
    peak: frac 7.93 at N = 8) and its regime follows the per-rank working set.  Figures: SCALE_r02 (C4, 8 GPUs)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    bi = b.b_iter(512 ** 3, 937951232)
    for world, its in ((1, 290.0), (2, 593.9), (4, 1151.4), (8, 2199.5)):
        r = b.one_kernel_roofline(False, world, 1798, bi, 1798 / its * 1e3, 1e6 / its, 6545.3, "MEASURED_PEAKS.json")
        assert 0.9 < r["frac"] <= 1.2 and r["per_gpu"] and r["n_gpus"] == world
        assert abs(r["aggregate_GBps"] - world * r["achieved"]) < 1e-6 * r["aggregate_GBps"]
        assert r["regime"].startswith("hbm-bound")
    c1 = b.one_kernel_roofline(False, 1, 366, b.b_iter(16384, 81408), 366 / 147.0, 6.8, 6545.3, "x")
    assert c1["regime"].startswith("latency-bound") and c1["frac"] < 0.2
