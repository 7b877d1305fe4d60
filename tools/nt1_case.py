"""Quick check (GPU): fits on shapes whose row tiles hold a single centroid tile (k <= 128) and a few others, against the
FP64 oracle -- the one-tile case exercises the MMA issuer that owns no tile of a row tile."""
import sys, numpy as np
sys.path.insert(0, ".")
import oap_mllib_b200 as m
from oracle import kmeans_oracle as ko
ok = True
for n, d, k, tl in [(40000, 256, 100, 0), (40000, 256, 100, 2), (30000, 200, 33, 2), (50000, 64, 100, 2), (9000, 96, 70, 0)]:
    m._lib.check(m.lib().kmb200_set_option(b"two_level", float(tl)))
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + d)
    init = ko.row_init_centers(X, k)
    model = m.KMeansDALImpl(k, 4, 1e-4, "euclidean", init, 1, 1).train(X)
    wc, wcost, wit = ko.lloyd(X.astype(np.float64), init, 1e-4, 4)
    good = model.numIter == wit and abs(model.trainingCost - wcost) <= 1e-5 * wcost and np.abs(model.clusterCenters - wc).max() <= 1e-5 * np.abs(wc).max()
    print(n, d, k, tl, "OK" if good else "FAIL", flush=True)
    ok &= good
sys.exit(0 if ok else 1)
