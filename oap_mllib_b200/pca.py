"""Host-side mirror of the reference's PCA path: ``ml.feature.PCA`` ->
``PCADALImpl.train`` -> one native call per executor.

Reference: spark-3.1.1/ml/feature/PCA.scala:84-116 (``fit``: k <= numFeatures, DAL path below 65535 features),
ml/feature/PCADALImpl.scala:35-136 (``train``, ``getPrincipleComponentsFromDAL``, ``getExplainedVarianceFromDAL``),
native/PCADALImpl.cpp:38-190 (``cPCATrainDAL``).  The compute is ``kmb200_pca_train`` of libkmeans_b200.so; there is
no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from ._lib import check, lib
from .kmeans import OneCCL, OneDAL, Table, _ptr


@dataclass
class PCAResult:  # java/org/apache/spark/ml/feature/PCAResult.java: the two result tables of the native call
    pcNumericTable: Optional[np.ndarray] = None                  # d x d, one principal axis per ROW
    explainedVarianceNumericTable: Optional[np.ndarray] = None   # 1 x d eigenvalues, descending


@dataclass
class MLlibPCAModel:  # what PCADALImpl.train returns (PCADALImpl.scala:91-96)
    k: int
    pc: np.ndarray                 # d x k, one principal component per COLUMN
    explainedVariance: np.ndarray  # k


class PCADALImpl:
    """Mirror of ``PCADALImpl`` (PCADALImpl.scala:30-147): same constructor arguments; ``train(data)`` runs one native
    call per executor and returns rank 0's model.  ``normalizeData`` (:100-106, two JVM passes that subtract the column
    means) is done inside the native call (``center = True``); pass ``center=False`` for rows that were centred already."""

    def __init__(self, k: int, executorNum: int = 1, executorCores: int = 1):
        self.k, self.executorNum, self.executorCores = int(k), int(executorNum), int(executorCores)

    def cPCATrainDAL(self, table: Table, k: int, executor_num: int, executor_cores: int, result: PCAResult,
                     center: bool = True) -> int:
        """The native entry (PCADALImpl.scala:139-145 / native/PCADALImpl.cpp:38-42).  Fills ``result`` on rank 0
        only, like the reference (:150-186); ``k`` and ``executor_cores`` are unused by the native side there too."""
        d = table.cols
        ev = np.empty(d, np.float64)
        axes = np.empty((d, d), np.float64)
        check(lib().kmb200_pca_train(table.handle, OneCCL.comm(), 1 if center else 0, _ptr(ev), _ptr(axes)))
        if OneCCL.isRoot() or OneCCL.comm() is None:
            result.pcNumericTable = axes
            result.explainedVarianceNumericTable = ev.reshape(1, d)
        return 0

    @staticmethod
    def getPrincipleComponentsFromDAL(table: np.ndarray, k: int) -> np.ndarray:
        """PCADALImpl.scala:108-127: the first k rows become the columns of a numCols x k matrix."""
        if k > table.shape[0]:
            raise ValueError("k should be less or equal to row number")
        return np.ascontiguousarray(table[:k].T)

    @staticmethod
    def getExplainedVarianceFromDAL(table_1xn: np.ndarray, k: int) -> np.ndarray:
        """PCADALImpl.scala:129-136: the first k eigenvalues over the sum of ALL of them."""
        data = np.asarray(table_1xn, np.float64).reshape(-1)
        return data[:k] / data.sum()

    def train(self, data, rank: int = 0, unique_id: Optional[bytes] = None, center: bool = True) -> Optional[MLlibPCAModel]:
        """data: this executor's rows (array / list of partitions) or an existing ``Table``.  With executorNum > 1
        every rank calls train(rank=..., unique_id=...) collectively; rank 0 gets the model, the others ``None``."""
        own = not isinstance(data, Table)
        table = OneDAL.vectorsToMergedNumericTables(data, self.executorNum) if own else data
        try:
            if self.executorNum > 1:
                if unique_id is None:
                    raise ValueError("executorNum > 1 needs the unique id created by OneCCL.getUniqueId()")
                OneCCL.init(self.executorNum, rank, unique_id)
            else:
                OneCCL.init(1, 0, b"")
            result = PCAResult()
            self.cPCATrainDAL(table, self.k, self.executorNum, self.executorCores, result, center)
            ret = None
            if result.pcNumericTable is not None:
                ret = MLlibPCAModel(self.k, self.getPrincipleComponentsFromDAL(result.pcNumericTable, self.k),
                                    self.getExplainedVarianceFromDAL(result.explainedVarianceNumericTable, self.k))
        finally:
            if self.executorNum > 1:
                OneCCL.cleanup()
            if own:
                OneDAL.releaseNumericTables(table)
        return ret


class PCAModel:
    """ml.feature.PCAModel (PCA.scala:132-175): ``pc`` d x k, ``explainedVariance`` k, ``transform`` = rows . pc."""

    def __init__(self, pc: np.ndarray, explainedVariance: np.ndarray):
        self.pc, self.explainedVariance = pc, explainedVariance

    def transform(self, rows) -> np.ndarray:
        return np.asarray(rows, np.float64) @ self.pc


class PCA:
    """ml.feature.PCA (PCA.scala:62-116): ``setK`` / ``fit``."""

    def __init__(self, k: int = 1):
        self.k = int(k)

    def setK(self, v):
        self.k = int(v)
        return self

    def fit(self, X) -> PCAModel:
        own = not isinstance(X, Table)
        table = OneDAL.vectorsToMergedNumericTables(X) if own else X
        try:
            if self.k > table.cols:   # PCA.scala:93-94
                raise ValueError(f"source vector size {table.cols} must be no less than k={self.k}")
            m = PCADALImpl(self.k, 1, 1).train(table)
        finally:
            if own:
                OneDAL.releaseNumericTables(table)
        return PCAModel(m.pc, m.explainedVariance)


def gram(table: Table, center: bool = True):
    """The (centred) Gram matrix, the column means and the global row count (``kmb200_pca_gram``)."""
    d = table.cols
    g = np.empty((d, d), np.float64)
    mean = np.empty(d, np.float64)
    n = C.c_int64(0)
    check(lib().kmb200_pca_gram(table.handle, OneCCL.comm(), 1 if center else 0, _ptr(g), _ptr(mean), C.byref(n)))
    return g, mean, n.value
