"""CPU tests: the oracle against every known answer the reference's own tests hold for the path
(tests/golden/reference_kats.json), and the two independent restatements against each other."""
import numpy as np
import pytest

from helpers import c_oracle_lloyd, load_kats
from oracle import kmeans_oracle as ko

KATS = load_kats()


def test_kat_sample_data_all_inits():
    kat = KATS["kat_sample_data"]
    X = np.array(kat["data"])
    expected_iters = {(0, 3): 2, (0, 1): 3, (4, 5): 3}
    for init, iters in expected_iters.items():
        C, cost, it, lab = ko.lloyd(X, X[list(init)], 1e-4, 20, return_labels=True)
        C = C[np.argsort(C[:, 0])]
        np.testing.assert_allclose(C, kat["expect"]["centers"], rtol=0, atol=1e-12)
        assert abs(cost - kat["expect"]["cost"]) < 1e-9
        assert it == iters
        assert sorted(np.bincount(lab).tolist()) == [3, 3]


def test_kat_fit_summary_one_iteration():
    # IntelKMeansSuite.scala:110-141: k=5 on the 50x3 fixture -> numIter == 1, cost < 0.1, sizes sum to N
    kat = KATS["kat_fit_summary"]
    X = np.array(kat["data"])
    init = np.unique(X, axis=0)                       # what k-means|| yields on 5 distinct points
    C, cost, it, lab = ko.lloyd(X, init, 1e-4, 20, return_labels=True)
    assert it == kat["expect"]["num_iter"]
    assert cost < kat["expect"]["cost_lt"] and cost == 0.0
    assert sorted(set(lab.tolist())) == kat["expect"]["labels_set"]
    assert np.bincount(lab).sum() == kat["expect"]["sizes_sum"]
    assert C.shape[0] == kat["expect"]["n_centers"]


def test_kat_four_centers_pair_means():
    # IntelKMeansSuite.scala:413-480: DAL result == vanilla result == the four pair means
    kat = KATS["kat_four_centers"]
    X = np.array(kat["data"])
    for init_rows in ([0, 2, 4, 6], [1, 2, 5, 7], [0, 3, 4, 7]):   # inits inside the basin the reference reaches
        C, cost, it, lab = ko.lloyd(X, X[init_rows], 1e-4, kat["max_iter"], return_labels=True)
        for a, b in kat["expect"]["pairs"]:
            assert lab[a] == lab[b]
        assert len(set(lab.tolist())) == 4
        got = C[np.lexsort((C[:, 1], C[:, 0]))]
        want = np.array(kat["expect"]["centers"])
        want = want[np.lexsort((want[:, 1], want[:, 0]))]
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-12)


def test_kat_array_input_float_vs_double():
    # IntelKMeansSuite.scala:204-218: cost equal within abs 1e-6 for Double vs Float input
    kat = KATS["kat_array_input"]
    X = np.array(kat["data"])
    init = X[[0, 1, 2, 3, 4]]
    _, c64, _ = ko.lloyd(X, init, 1e-4, kat["max_iter"])
    _, c32, _ = ko.lloyd(X.astype(np.float32), init, 1e-4, kat["max_iter"])
    assert abs(c64 - c32) <= kat["expect"]["cost_abs_tol"]


def test_max_iter_zero_returns_init():
    X = np.array(KATS["kat_sample_data"]["data"])
    C, cost, it = ko.lloyd(X, X[[0, 3]], 1e-4, 0)
    assert it == 0 and cost == 0.0
    np.testing.assert_array_equal(C, X[[0, 3]])


def test_first_minimum_wins_ties():
    X = np.array([[0.0, 0.0], [2.0, 0.0], [1.0, 0.0]])        # row 2 is equidistant
    lab, _, _ = ko.assign(X, np.array([[0.0, 0.0], [2.0, 0.0]]))
    assert lab.tolist() == [0, 1, 0]
    lab, _, _ = ko.assign(X, np.array([[2.0, 0.0], [0.0, 0.0]]))
    assert lab.tolist() == [1, 0, 0]


def test_empty_cluster_takes_farthest_candidate():
    X = np.array([[0.0, 0], [1, 0], [0, 1], [10, 10], [11, 10]])
    init = np.array([[0.0, 0], [10, 10], [100, 100]])
    C, obj, counts, _ = ko.lloyd_step([X], init)
    assert counts.tolist() == [3, 2, 0]
    np.testing.assert_allclose(C[2], [1.0, 0.0])               # farthest row (dist 1, lowest index)
    full = sum(((x - c) ** 2).sum() for x, c in zip(X, init[[0, 0, 0, 1, 1]]))
    assert abs(obj - (full - 1.0)) < 1e-12


def test_convergence_is_all_centres_within_tol():
    old = np.zeros((2, 3))
    new = np.array([[1e-4 / np.sqrt(3)] * 3, [0, 0, 0]]) * 0.999
    assert ko.centers_converged(old, new, 1e-4)
    assert not ko.centers_converged(old, new * 1.01, 1e-4)


@pytest.mark.parametrize("n,d,k,ranks", [(3000, 8, 5, 1), (5000, 33, 17, 2), (2000, 64, 40, 3)])
def test_numpy_and_c_restatements_agree(oracle_c, n, d, k, ranks):
    X, _, _ = ko.gaussian_blobs(n, d, k, seed=n)
    init = ko.row_init_centers(X, k)
    a = ko.lloyd(X, init, 1e-4, 25, n_ranks=ranks, return_labels=True)
    rc, C, cost, it, lab = c_oracle_lloyd(oracle_c, X, init, 1e-4, 25, ranks)
    assert rc == 0 and it == a[2]
    np.testing.assert_allclose(C, a[0], rtol=1e-12, atol=1e-12)
    assert abs(cost - a[1]) <= 1e-9 * abs(a[1])
    assert (lab != a[3]).sum() == 0


def test_multi_rank_merge_equals_single_rank():
    X, _, _ = ko.gaussian_blobs(4000, 12, 6, seed=3)
    init = ko.row_init_centers(X, 6)
    a = ko.lloyd(X, init, 1e-4, 20, n_ranks=1)
    b = ko.lloyd(X, init, 1e-4, 20, n_ranks=4)
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-12)
    assert abs(a[1] - b[1]) <= 1e-10 * a[1]


def test_oracle_reproduces_committed_fixtures():
    """tests/golden/oracle_fixtures.npz pins the oracle: same seeded inputs, same centres / cost / iterations."""
    import os
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "oracle_fixtures.npz"))
    for i, (n, d, k, seed, iters) in enumerate(fx["cases"].tolist()):
        X, _, _ = ko.gaussian_blobs(n, d, k, seed=seed)
        np.testing.assert_allclose(X.astype(np.float64).sum(0), fx[f"xsum_{i}"], rtol=1e-12)
        C, cost, it = ko.lloyd(X.astype(np.float64), ko.row_init_centers(X, k), 1e-4, iters)
        assert it == int(fx[f"iters_{i}"])
        np.testing.assert_allclose(C, fx[f"centers_{i}"], rtol=1e-10, atol=1e-12)
        assert abs(cost - float(fx[f"cost_{i}"])) <= 1e-10 * cost


def test_dense_rows_restates_the_reference_conversion():
    """OneDAL.vectorsToMergedNumericTables densifies every row (row.toArray) into a row-major FP64 table: the oracle's
    restatement must give the same table from dense rows, from columns and from CSR (with empty rows)."""
    rng = np.random.default_rng(4)
    X = rng.normal(size=(37, 9))
    X[np.abs(X) < 0.7] = 0.0
    X[5] = 0.0

    class Csr:   # the three arrays of a SparseVector batch, without needing scipy
        def __init__(self, dense):
            self.shape = dense.shape
            self.indptr = np.zeros(dense.shape[0] + 1, np.int64)
            idx, val = [], []
            for r, row in enumerate(dense):
                nz = np.flatnonzero(row)
                idx.extend(nz.tolist())
                val.extend(row[nz].tolist())
                self.indptr[r + 1] = len(idx)
            self.indices, self.data = np.array(idx, np.int32), np.array(val, np.float64)

    assert np.array_equal(ko.dense_rows(X), X)
    assert np.array_equal(ko.dense_rows(tuple(np.ascontiguousarray(X[:, c]) for c in range(X.shape[1]))), X)
    assert np.array_equal(ko.dense_rows(Csr(X)), X)
    assert ko.dense_rows(X.astype(np.float32)).dtype == np.float64
