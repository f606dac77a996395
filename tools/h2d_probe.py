#!/usr/bin/env python
"""Host <-> device copy rates through the C-ABI (cgb_memcpy_h2d / _d2h) at 1 GiB: pinned host memory, pageable host
memory through the library's staged path (pinned bounce buffers + host worker threads, core.cu), and pageable with
staging switched off (CGB_STAGE=0: the driver's own path).  VERDICT r1 item 6 asks for >= 35 GB/s pageable H2D.
  python tools/h2d_probe.py [GiB = 1]"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def rate(fn, nbytes, reps=5):
    fn()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return nbytes / best / 1e9


def main():
    gib = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    nbytes = int(gib * (1 << 30)) & ~7
    ctx = S.Context(0)
    lib = ctx.lib
    d = ctx.malloc(nbytes)
    pageable = np.random.default_rng(0).random(nbytes // 8)
    back = np.zeros(nbytes // 8)
    hp = C.c_void_p()
    L.check(lib.cgb_host_alloc(nbytes, C.byref(hp)))
    pinned = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_double)), shape=(nbytes // 8,))
    pinned[...] = pageable
    out = {}
    out["h2d_pinned_GBps"] = rate(lambda: L.check(lib.cgb_memcpy_h2d(ctx.h, d, pinned.ctypes.data, nbytes, 1)), nbytes)
    out["d2h_pinned_GBps"] = rate(lambda: L.check(lib.cgb_memcpy_d2h(ctx.h, pinned.ctypes.data, d, nbytes, 1)), nbytes)
    out["h2d_pageable_staged_GBps"] = rate(lambda: L.check(lib.cgb_memcpy_h2d(ctx.h, d, pageable.ctypes.data, nbytes, 1)), nbytes)
    out["d2h_pageable_staged_GBps"] = rate(lambda: L.check(lib.cgb_memcpy_d2h(ctx.h, back.ctypes.data, d, nbytes, 1)), nbytes)
    assert np.array_equal(back, pageable)          # the staged round trip is exact
    os.environ["CGB_STAGE"] = "0"
    out["h2d_pageable_driver_GBps"] = rate(lambda: L.check(lib.cgb_memcpy_h2d(ctx.h, d, pageable.ctypes.data, nbytes, 1)), nbytes, reps=3)
    out["d2h_pageable_driver_GBps"] = rate(lambda: L.check(lib.cgb_memcpy_d2h(ctx.h, back.ctypes.data, d, nbytes, 1)), nbytes, reps=3)
    del os.environ["CGB_STAGE"]
    out["bytes"] = nbytes
    out["stage_threads"] = os.environ.get("CGB_STAGE_THREADS", "default")
    import json
    print(json.dumps(out))


if __name__ == "__main__":
    main()
