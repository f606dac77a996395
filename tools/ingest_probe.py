"""Matrix Market ingest probe: the same file through
  (a) cgb_mm_read with one host thread and with all of them (host CSR),
  (b) cgb_csr_from_mm_file (parallel parse + CSR assembled on the GPU), when a GPU is present.
The reference's own read_file on the same kind of file is timed by tests/test_reader.py::test_reader_vs_reference_speed
(the oracle is test infrastructure and is not imported from here).
python tools/ingest_probe.py [m]      3D 7-point Poisson m^3 (default 128 = BASELINE config C2)"""
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import generators as G  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    v, c, r = G.poisson3d_7pt(m)
    path = os.path.join(tempfile.gettempdir(), f"p3d7_{m}.mtx")
    t0 = time.time()
    G.write_mtx_reference_dialect(path, v, c, r)
    print(f"wrote {path}: n={r.size - 1} nnz={v.size} {os.path.getsize(path) / 1e6:.1f} MB in {time.time() - t0:.1f}s", flush=True)

    def best(fn, reps=3):
        ts = []
        for _ in range(reps):
            t0 = time.time()
            out = fn()
            ts.append(time.time() - t0)
        return min(ts), out

    for th in (1, os.cpu_count() or 1):
        os.environ["CGB_MM_THREADS"] = str(th)
        t, out = best(lambda: S.mm_read(path))
        assert all(np.array_equal(a, b) for a, b in zip(out, (v, c, r)))
        print(f"cgb_mm_read, {th} thread(s), host CSR: {t:.2f} s", flush=True)
    if L.load().cgb_device_count() > 0:
        ctx = S.Context(0)
        S.Matrix.from_mm_file(ctx, path).free()  # warm-up (context, cub temp storage)
        ts = []
        for _ in range(3):
            t0 = time.time()
            A = S.Matrix.from_mm_file(ctx, path)
            ctx.sync()
            ts.append(time.time() - t0)
            ok = all(np.array_equal(a, b) for a, b in zip(A.to_host(), (v, c, r)))
            A.free()
            assert ok
        t = min(ts)
        print(f"cgb_csr_from_mm_file, {os.cpu_count()} threads + GPU assembly, device CSR: {t:.2f} s", flush=True)
        rows = np.repeat(np.arange(r.size - 1), np.diff(r)).astype(np.int32)
        t, B = best(lambda: (S.Matrix.from_triplets(ctx, rows, c, v, n=r.size - 1), ctx.sync())[0])
        print(f"cgb_csr_from_triplets ({v.size} triplets from host memory): {t:.3f} s", flush=True)
    os.remove(path)


if __name__ == "__main__":
    main()
