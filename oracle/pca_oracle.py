"""FP64 restatement of the reference's PCA path  --  TEST INFRASTRUCTURE ONLY (never imported by oap_mllib_b200).

What the reference does:

* ``PCADALImpl.train`` (/root/reference/mllib-dal/src/main/scala/org/apache/spark/ml/feature/PCADALImpl.scala:35-98)
  centres the rows on the JVM (``normalizeData``, :100-106: ``StandardScaler(withMean = true, withStd = false)``),
  builds one numeric table per executor and calls ``cPCATrainDAL`` on every rank;
* ``cPCATrainDAL`` (/root/reference/mllib-dal/src/main/native/PCADALImpl.cpp:38-190) runs oneDAL's
  ``pca::Distributed<step1Local, double, pca::svdDense>`` on the local block (flagged "already normalised", :50-51, so
  oneDAL neither centres nor scales), all-gathers the serialised partial results (:78-101) and runs
  ``pca::Distributed<step2Master, double, pca::svdDense>`` on rank 0 (:111-136); the result tables are ``eigenvectors``
  (one principal axis per ROW) and ``eigenvalues`` (1 x d, descending);
* back in Scala, ``getPrincipleComponentsFromDAL`` (:108-127) copies the first k rows into the columns of a d x k
  matrix and ``getExplainedVarianceFromDAL`` (:129-136) divides the first k eigenvalues by the sum of ALL d of them.

oneDAL (2021.2, the version the reference's pom pins) is not in /root/reference; its published ``svdDense`` method is
restated here: step 1 reduces each block to the d x d triangular factor of its QR decomposition, step 2 takes the SVD of
the stacked factors; ``eigenvalues = sigma^2 / (n - 1)`` and the eigenvectors are the right singular vectors.  The same
numbers follow from the eigen-decomposition of the centred Gram matrix ``X^T X / (n - 1)`` (``pca_via_gram``: the route the
GPU path takes), which is how Spark's own ``RowMatrix.computePrincipalComponentsAndExplainedVariance`` -- the expected
value in the reference's IntelPCASuite.scala:40-85 -- computes them.

PINNED by: the reference's only PCA fixture, the 3 x 5 matrix of IntelPCASuite.scala:41-45 / examples/data/pca_data.csv,
whose projection onto the first three axes is printed in Spark's documentation of this very example
(tests/test_pca_oracle.py holds the numbers); both routes must reproduce it.  The SIGN of an axis is not defined by the
reference (it is whatever LAPACK returns; the reference's test compares absolute values, IntelPCASuite.scala:75-84):
``fix_signs`` makes the entry of largest magnitude positive, and the GPU path follows the same rule.
"""
from __future__ import annotations

import numpy as np


def fix_signs(axes: np.ndarray) -> np.ndarray:
    """Rows = axes.  Entry of largest magnitude positive (lowest column on exact ties)."""
    out = np.array(axes, dtype=np.float64, copy=True)
    for i in range(out.shape[0]):
        j = int(np.argmax(np.abs(out[i])))
        if out[i, j] < 0:
            out[i] = -out[i]
    return out


def column_means(parts) -> tuple[np.ndarray, int]:
    n = sum(int(p.shape[0]) for p in parts)
    s = np.zeros(parts[0].shape[1], np.float64)
    for p in parts:
        s += np.asarray(p, np.float64).sum(axis=0)
    return s / n, n


def pca_svd_dense(parts, center: bool = True):
    """The reference's route.  parts: one [rows_i x d] array per rank.  Returns (eigenvalues[d] descending,
    eigenvectors[d x d], one axis per row, signs fixed).  With fewer than d rows in total the trailing eigenvalues are 0."""
    d = parts[0].shape[1]
    mean, n = column_means(parts)
    if not center:
        mean = np.zeros(d)
    factors = []
    for p in parts:   # step1Local: the block's triangular factor
        y = np.asarray(p, np.float64) - mean
        r = np.linalg.qr(y, mode="r")
        factors.append(r)
    stacked = np.vstack(factors)   # step2Master: SVD of the stacked factors
    _, sigma, vt = np.linalg.svd(stacked, full_matrices=True)
    ev = np.zeros(d)
    ev[:len(sigma)] = sigma ** 2 / max(n - 1, 1)
    return ev, fix_signs(vt[:d])


def centred_gram(parts, center: bool = True):
    """sum_i (x_i - mean)(x_i - mean)^T in FP64, the mean, the row count."""
    d = parts[0].shape[1]
    mean, n = column_means(parts)
    if not center:
        mean = np.zeros(d)
    g = np.zeros((d, d))
    for p in parts:
        y = np.asarray(p, np.float64) - mean
        for r0 in range(0, y.shape[0], 1 << 16):
            blk = y[r0:r0 + (1 << 16)]
            g += blk.T @ blk
    return g, mean, n


def eig_desc(sym: np.ndarray):
    """Eigenvalues descending, eigenvectors as rows with fixed signs."""
    w, v = np.linalg.eigh(sym)
    order = np.argsort(-w, kind="stable")
    return np.maximum(w[order], 0.0), fix_signs(v[:, order].T)


def pca_via_gram(parts, center: bool = True):
    g, _, n = centred_gram(parts, center)
    w, axes = eig_desc(g)
    return w / max(n - 1, 1), axes


def model(eigenvalues: np.ndarray, eigenvectors: np.ndarray, k: int):
    """What PCADALImpl.train hands to MLlibPCAModel (PCADALImpl.scala:108-136): pc [d x k], explainedVariance [k]."""
    assert k <= eigenvectors.shape[0], "k should be less or equal to row number"
    total = eigenvalues.sum()
    return eigenvectors[:k].T.copy(), eigenvalues[:k] / total


def subspace_gap(a_rows: np.ndarray, b_rows: np.ndarray) -> float:
    """Largest principal angle (as a sine) between the spans of two sets of orthonormal rows."""
    s = np.linalg.svd(a_rows @ b_rows.T, compute_uv=False)
    return float(np.sqrt(max(0.0, 1.0 - min(1.0, s.min()) ** 2)))
