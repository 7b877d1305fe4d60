"""The case profiled with ncu: three Lloyd iterations on one shard-sized table of blobs (default 12.5M x 256, k = 1000 --
what one of 8 GPUs holds of the headline config), through the C ABI.  Launch order per iteration on the default path:
coarse one-product pass, gather, three-product fast pass, gather, full-bookkeeping pass, re-check, bucketing, accumulate."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, ".")
import oap_mllib_b200 as m
from oap_mllib_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 256
k = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
_lib.check(m.lib().kmb200_set_device(0))
for opt in os.environ.get("KMB_OPTIONS", "").split(","):     # e.g. KMB_OPTIONS=fused=2,cascade=0
    if opt:
        name, val = opt.split("=")
        _lib.check(m.lib().kmb200_set_option(name.encode(), float(val)))
centres = np.random.default_rng(42).uniform(-10, 10, (k, d)).astype(np.float32)
t = m.Table(n, d)
t.fill_blobs(centres, seed=42)
init = np.ascontiguousarray(np.vstack([t.read_rows(int(i), 1) for i in (np.arange(k, dtype=np.int64) * n) // k]).astype(np.float64))
out = np.empty((k, d))
cost, it = C.c_double(), C.c_int32()
_lib.check(m.lib().kmb200_fit(t.handle, None, init.ctypes.data_as(C.c_void_p), k, d, -1.0, iters,
                              out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
p = _lib.profile()
print("iterations", it.value, "cost", cost.value, "variant", p["assign_variant"], "assign ms/iter", p["assign_ms"] / iters, "main", p["assign_main_ms"] / iters)
t.release()
