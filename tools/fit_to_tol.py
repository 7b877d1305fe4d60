"""BASELINE config 3 run the way a user runs it: Lloyd to tol = 1e-4 (maxIter 300) on the synthetic 100M x 256
blobs, k = 1000, one B200; prints iterations, cost and time per iteration (device timings of the library)."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oap_mllib_b200 as m  # noqa: E402
from oap_mllib_b200 import _lib  # noqa: E402

rows, d, k = int(os.environ.get("ROWS", 100_000_000)), 256, 1000
centres = np.random.default_rng(42).uniform(-10, 10, (k, d)).astype(np.float32)
t = m.Table(rows, d)
t.fill_blobs(centres, seed=42)
idx = (np.arange(k, dtype=np.int64) * rows) // k
init = np.ascontiguousarray(np.vstack([t.read_rows(int(i), 1) for i in idx]).astype(np.float64))
out = np.empty((k, d))
cost, it = C.c_double(), C.c_int32()
t0 = time.perf_counter()
_lib.check(m.lib().kmb200_fit(t.handle, None, init.ctypes.data_as(C.c_void_p), k, d, 1e-4, int(os.environ.get("MAXIT", 300)),
                              out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
wall = time.perf_counter() - t0
p = _lib.profile()
print(json.dumps({"workload": f"{rows}x{d} k={k} tol=1e-4 maxIter=" + os.environ.get("MAXIT", "300"), "iterations": it.value, "cost": cost.value,
                  "wall_s": wall, "device_ms_per_iteration": p["total_ms"] / max(it.value, 1),
                  "rechecked_rows_total": p["rechecked_rows"], "full_rechecks_total": p["full_rechecks"],
                  "empty_cluster_events": p["empty_cluster_events"],
                  "phases_ms_total": {kk: p[kk] for kk in ("assign_ms", "recheck_ms", "bucket_ms", "accumulate_ms", "finalize_ms")}}))
t.release()
