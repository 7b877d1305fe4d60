"""The PCA oracle (oracle/pca_oracle.py) against the reference's own PCA fixture, and its two routes against each other.

The fixture is the 3 x 5 matrix of IntelPCASuite.scala:41-45 (= examples/data/pca_data.csv, PCAExample.scala:37-41).  The
reference's test compares with Spark's RowMatrix PCA at absTol 1e-5 on absolute values; the numbers below are that
example's projection as printed in Spark's ML guide ("PCA" under "Extracting, transforming and selecting features") and
the explained variance of the two non-null axes.  The third axis belongs to a zero eigenvalue and is not defined."""
import numpy as np
import pytest

from oracle import pca_oracle as po

import json
import os

with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pca_fixture.json")) as _f:
    _FX = json.load(_f)
FIXTURE = np.array(_FX["rows"])
SPARK_PROJECTION = np.array(_FX["spark_projection_first_two_axes"])
SPARK_EXPLAINED = np.array(_FX["spark_explained_variance_first_two"])


@pytest.mark.parametrize("route", ["svd_dense", "gram"])
@pytest.mark.parametrize("split", [(3,), (2, 1), (1, 1, 1)])
def test_reference_fixture(route, split):
    parts, r = [], 0
    for m in split:
        parts.append(FIXTURE[r:r + m])
        r += m
    ev, axes = (po.pca_svd_dense if route == "svd_dense" else po.pca_via_gram)(parts)
    pc, explained = po.model(ev, axes, 3)
    assert np.allclose(np.abs(FIXTURE @ pc[:, :2]), np.abs(SPARK_PROJECTION), rtol=0, atol=1e-12)
    assert np.allclose(explained[:2], SPARK_EXPLAINED, rtol=0, atol=1e-13)
    assert abs(explained[2]) < 1e-12


def test_routes_agree_on_random_tables():
    rng = np.random.default_rng(3)
    for n, d in [(500, 7), (2000, 33), (300, 64)]:
        basis = np.linalg.qr(rng.standard_normal((d, d)))[0]
        X = (rng.standard_normal((n, d)) * (2.0 ** -np.arange(d) * 10 + 0.01)) @ basis + rng.uniform(-5, 5, d)
        X = X.astype(np.float32)
        parts = [X[:n // 3], X[n // 3:]]
        ev1, ax1 = po.pca_svd_dense(parts)
        ev2, ax2 = po.pca_via_gram(parts)
        assert np.allclose(ev1, ev2, rtol=1e-9, atol=1e-12 * ev1[0])
        lead = min(5, d)
        assert np.allclose(ax1[:lead], ax2[:lead], rtol=0, atol=1e-7)
        cov = np.cov(X.astype(np.float64), rowvar=False)
        assert np.allclose(np.sort(np.linalg.eigvalsh(cov))[::-1], ev1, rtol=1e-9, atol=1e-12 * ev1[0])


def test_without_centring_is_the_raw_gram():
    rng = np.random.default_rng(4)
    X = rng.standard_normal((400, 6)) + 3.0
    ev, axes = po.pca_svd_dense([X], center=False)
    w = np.sort(np.linalg.eigvalsh(X.T @ X / 399.0))[::-1]
    assert np.allclose(ev, w, rtol=1e-10)


def test_sign_rule():
    a = np.array([[0.1, -0.9, 0.3], [0.5, 0.2, -0.1]])
    f = po.fix_signs(a)
    assert np.array_equal(f[0], -a[0]) and np.array_equal(f[1], a[1])
