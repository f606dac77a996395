"""Row-partitioned multi-GPU CG vs the single-process oracle (needs >= 2 B200; skipped otherwise)."""
import os
import subprocess
import sys

import pytest

from conjugategradient_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("variant", ["default", "jitter", "graph", "halo_engine", "no_allgather", "nccl", "dict"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_dist_solve_matches_oracle(world, variant):
    if L.load().cgb_device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ)
    if variant == "default":
        env["CGB_TEST_EXPECT_LOOP"] = "persistent"   # stencil blocks run the whole loop in one cooperative kernel per rank
    if variant == "jitter":
        # flag-protocol stress: every CTA of every rank sleeps a pseudo-random time (< 20 us) before each phase, and the
        # ranks enter every solve after a random host delay; results must not change by a bit
        env["CGB_DIST_JITTER_NS"] = "20000"
        env["CGB_TEST_HOST_JITTER"] = "1"
        env["CGB_TEST_EXPECT_LOOP"] = "persistent"
    if variant == "graph":
        env["CGB_DIST_PERSISTENT"] = "0"          # the launch-per-kernel loop (CUDA-graph batches) stays correct
        env["CGB_TEST_EXPECT_LOOP"] = "graph"
    if variant == "no_allgather":
        env["CGB_DIST_ALLGATHER"] = "0"           # the random matrix through the halo split (default for it: all-gather of p)
    if variant == "halo_engine":
        env["CGB_HALO_ENGINE_MIN"] = "1"      # A_halo through the regular row engines even for these small cases
    if variant == "nccl":
        env["CGB_DIST_TRANSPORT"] = "nccl"    # the fallback transport stays correct
    if variant == "dict":
        env["CGB_TEST_FORMAT"] = "dict"       # row-pattern dictionary on every rank's local block
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + world), os.path.join(ROOT, "tests", "dist_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert out.stdout.count("DIST_OK") == 4 and "DIST_FAIL" not in out.stdout, out.stdout
