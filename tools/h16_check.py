"""Diagnostics: labels of kmb200_assign (tensor-core path) against the FP64 oracle for one shape, per kernel option."""
import sys
import numpy as np
sys.path.insert(0, ".")
import oap_mllib_b200 as m
from oracle import kmeans_oracle as ko

n, d, k = (int(a) for a in sys.argv[1:4])
X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + k)
C0 = ko.row_init_centers(X, k)
want, _, _ = ko.assign(X.astype(np.float64), C0)
for h16 in (1, 0, 1, 1):
    m._lib.check(m.lib().kmb200_set_option(b"h16", float(h16)))
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        labels = np.empty(n, np.int32)
        sizes = np.empty(k, np.int64)
        Cc = np.ascontiguousarray(C0, np.float64)
        m._lib.check(m.lib().kmb200_assign(t.handle, None, Cc.ctypes.data, k, d, labels.ctypes.data, sizes.ctypes.data, None))
    finally:
        t.release()
    bad = np.flatnonzero(labels != want)
    print(f"h16={h16}: {bad.size} label mismatches", bad[:10], labels[bad[:10]], want[bad[:10]])
