"""Writes tests/golden/oracle_fixtures.npz: seeded inputs are regenerated from (n, d, k, seed) by
oracle.kmeans_oracle.gaussian_blobs; the EXPECTED centres / cost / iterations come from the FP64 oracle
(oracle/kmeans_oracle.py) at the time this script was run.  The fixtures (a) let the GPU parity tests
compare against committed numbers, not only against a live oracle run, and (b) pin the oracle itself:
tests/test_oracle.py fails if a later edit of the oracle changes these results."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import kmeans_oracle as ko  # noqa: E402

CASES = [(6000, 24, 12, 101, 15), (5000, 96, 40, 202, 10), (3000, 256, 120, 303, 8), (8000, 32, 16, 404, 12)]


def main():
    out = {"cases": np.array(CASES, dtype=np.int64)}
    for i, (n, d, k, seed, iters) in enumerate(CASES):
        X, _, _ = ko.gaussian_blobs(n, d, k, seed=seed)
        init = ko.row_init_centers(X, k)
        C, cost, it, lab = ko.lloyd(X.astype(np.float64), init, 1e-4, iters, return_labels=True)
        out[f"centers_{i}"] = C
        out[f"cost_{i}"] = np.float64(cost)
        out[f"iters_{i}"] = np.int64(it)
        out[f"sizes_{i}"] = np.bincount(lab, minlength=k).astype(np.int64)
        out[f"xsum_{i}"] = X.astype(np.float64).sum(0)       # guards the input generator
    np.savez_compressed(os.path.join(HERE, "oracle_fixtures.npz"), **out)
    print("wrote", len(CASES), "cases")


if __name__ == "__main__":
    main()
