"""Diagnostics: random (n, d, k) shapes through kmb200_assign, labels / sizes / cost against the FP64 oracle,
single-pass and two-level bookkeeping.  usage: python tools/random_sweep.py [cases] [seed]"""
import sys

import numpy as np

sys.path.insert(0, ".")
import oap_mllib_b200 as m
from oracle import kmeans_oracle as ko

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 20
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for case in range(cases):
    d = int(rng.choice([8, 24, 32, 33, 64, 72, 96, 100, 128, 129, 160, 192, 200, 224, 256]))
    k = int(rng.choice([17, 31, 33, 64, 65, 100, 128, 129, 255, 256, 257, 384, 500, 513, 1000, 1100, 2049]))
    n = int(rng.integers(max(k, 300), 40000))
    X, _, _ = ko.gaussian_blobs(n, d, max(2, k // 2), seed=case)   # fewer blobs than centres: many near-ties
    C0 = ko.row_init_centers(X, k)
    want, best, second = ko.assign(X.astype(np.float64), C0)
    tie = ko.near_tie_mask(best, second)
    want_cost = float(best.sum())
    for two_level in (0, 2):
        m._lib.check(m.lib().kmb200_set_option(b"two_level", float(two_level)))
        t = m.OneDAL.vectorsToMergedNumericTables(X)
        try:
            labels = np.empty(n, np.int32)
            sizes = np.empty(k, np.int64)
            import ctypes as C
            cost = C.c_double()
            Cc = np.ascontiguousarray(C0, np.float64)
            m._lib.check(m.lib().kmb200_assign(t.handle, None, Cc.ctypes.data, k, d, labels.ctypes.data, sizes.ctypes.data, C.byref(cost)))
        finally:
            t.release()
        nbad = int(((labels != want) & ~tie).sum())          # exact-distance near-ties (<= 1e-6 relative) are excused
        ok = nbad == 0 and np.array_equal(sizes, np.bincount(labels, minlength=k)) and abs(cost.value - want_cost) <= 1e-6 * want_cost
        bad += not ok
        print(f"case {case}: n={n} d={d} k={k} two_level={two_level}: {nbad} label mismatches, cost {cost.value:.6e}/{want_cost:.6e} {'OK' if ok else 'FAIL'}", flush=True)
m._lib.check(m.lib().kmb200_set_option(b"two_level", 1.0))
print("failures:", bad)
sys.exit(1 if bad else 0)
