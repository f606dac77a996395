"""Pinned H2D / D2H bandwidth through the C-ABI copies (what the e2e leg pays per solve)."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import _lib as L  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402

ctx = S.Context(0)
lib = ctx.lib
for gb in (1, 4):
    nbytes = gb << 30
    h = C.c_void_p()
    L.check(lib.cgb_host_alloc(nbytes, C.byref(h)))
    C.memset(h, 1, nbytes)
    d = ctx.malloc(nbytes)
    for name, fn in (("h2d", lambda: lib.cgb_memcpy_h2d(ctx.h, d, h, nbytes, 1)), ("d2h", lambda: lib.cgb_memcpy_d2h(ctx.h, h, d, nbytes, 1))):
        fn()
        t0 = time.time()
        for _ in range(3):
            L.check(fn())
        dt = (time.time() - t0) / 3
        print(f"{name} {gb} GiB pinned: {nbytes / dt / 1e9:.1f} GB/s", flush=True)
    t0 = time.time()
    p2 = ctx.malloc(nbytes)
    ctx.free(p2)
    print(f"malloc+free {gb} GiB: {(time.time() - t0) * 1e3:.1f} ms")
    ctx.free(d)
    lib.cgb_host_free(h)
a = np.ones(1 << 27)  # pageable 1 GiB
d = ctx.malloc(a.nbytes)
t0 = time.time()
L.check(lib.cgb_memcpy_h2d(ctx.h, d, a.ctypes.data, a.nbytes, 1))
print(f"h2d 1 GiB pageable: {a.nbytes / (time.time() - t0) / 1e9:.1f} GB/s")
