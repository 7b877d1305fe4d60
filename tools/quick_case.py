"""Small end-to-end case (quick check after a kernel change): columnar ingest, fits on the FP16-split kernel with the
two-level scheme forced on / off, a fused small-k fit.  Exits non-zero on a parity failure."""
import sys

import numpy as np

sys.path.insert(0, ".")
import oap_mllib_b200 as m
from oracle import kmeans_oracle as ko

ok = True
for n, d, k, tl in [(3000, 64, 40, 2), (2500, 256, 130, 2), (2000, 96, 33, 0), (4000, 32, 8, 1)]:
    m._lib.check(m.lib().kmb200_set_option(b"two_level", float(tl)))
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n)
    init = ko.row_init_centers(X, k)
    t = m.OneDAL.vectorsToMergedNumericTables(tuple(np.ascontiguousarray(X[:, c]) for c in range(d)))
    model = m.KMeansDALImpl(k, 3, 1e-4, "euclidean", init, 1, 1).train(t)
    t.release()
    want_c, want_cost, want_it = ko.lloyd(X.astype(np.float64), init, 1e-4, 3)
    good = model.numIter == want_it and abs(model.trainingCost - want_cost) <= 1e-5 * want_cost
    print(n, d, k, tl, "OK" if good else "FAIL", flush=True)
    ok &= good
m._lib.check(m.lib().kmb200_set_option(b"two_level", 1.0))
sys.exit(0 if ok else 1)
