"""Host-side mirror of the reference's Scala interface for the K-means path, over the C ABI.

Same names, argument meaning and error behaviour as the reference (bobjiang82/oap-mllib, paths
relative to mllib-dal/src/main/):

* ``OneDAL``         scala/org/apache/spark/ml/util/OneDAL.scala:37-181   (data conversion)
* ``OneCCL``         scala/org/apache/spark/ml/util/OneCCL.scala:26-64    (rendezvous)
* ``KMeansResult``   java/org/apache/spark/ml/clustering/KMeansResult.java:19-22
* ``KMeansDALImpl``  scala/org/apache/spark/ml/clustering/KMeansDALImpl.scala:26-111
* ``KMeans`` / ``KMeansModel`` / ``KMeansSummary``
                     scala/org/apache/spark-3.1.1/ml/clustering/KMeans.scala:134-181,285-371

The Scala toolchain is absent from this image, so this mirror is Python; INTEGRATION.md shows the
Scala/JNI binding a maintainer adds.  Nothing here computes distances or sums on the CPU: all
arithmetic of the Lloyd iteration happens in libkmeans_b200.so on the GPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import Kmb200Error, check, lib


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Table:
    """Device-resident row partition of one executor (replaces the merged HomogenNumericTable
    whose address OneDAL.vectorsToMergedNumericTables hands to the native call)."""

    def __init__(self, rows: int, cols: int):
        h = C.c_void_p()
        check(lib().kmb200_table_create(int(rows), int(cols), C.byref(h)))
        self._h = h
        self.rows, self.cols = int(rows), int(cols)

    @property
    def handle(self):
        if self._h is None:
            raise Kmb200Error(-1, "table has been released")
        return self._h

    def append(self, row0: int, rows: np.ndarray) -> None:
        """Batched upload (replaces one cSetDoubleBatch JNI call per row, OneDAL.scala:135-139)."""
        rows = np.asarray(rows)
        if rows.ndim != 2:
            raise ValueError("rows must be 2-D")
        if rows.dtype == np.float32:
            rows = np.ascontiguousarray(rows)
            check(lib().kmb200_table_append_f32(self.handle, int(row0), _ptr(rows), rows.shape[0], rows.shape[1]))
        else:
            rows = np.ascontiguousarray(rows, dtype=np.float64)
            check(lib().kmb200_table_append_f64(self.handle, int(row0), _ptr(rows), rows.shape[0], rows.shape[1]))

    def append_columns(self, row0: int, columns) -> None:
        """Columnar upload: one contiguous 1-D array per column (Arrow / ColumnVector layout)."""
        cols = [np.asarray(c) for c in columns]
        if not cols or any(c.ndim != 1 or c.shape[0] != cols[0].shape[0] for c in cols):
            raise ValueError("columns must be 1-D arrays of equal length")
        f32 = all(c.dtype == np.float32 for c in cols)
        cols = [np.ascontiguousarray(c, dtype=np.float32 if f32 else np.float64) for c in cols]
        ptrs = (C.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
        fn = lib().kmb200_table_append_columns_f32 if f32 else lib().kmb200_table_append_columns_f64
        check(fn(self.handle, int(row0), C.cast(ptrs, C.c_void_p), cols[0].shape[0], len(cols)))

    def append_csr(self, row0: int, indptr, indices, values) -> None:
        """Sparse rows (CSR), densified on the device (the reference calls Vector.toArray per row on the JVM)."""
        indptr = np.ascontiguousarray(indptr, dtype=np.int64)
        indices = np.ascontiguousarray(indices, dtype=np.int32)
        values = np.ascontiguousarray(values, dtype=np.float64)
        check(lib().kmb200_table_append_csr_f64(self.handle, int(row0), indptr.shape[0] - 1, _ptr(indptr), _ptr(indices),
                                                _ptr(values)))

    def truncate(self, rows: int) -> None:
        """Cut a table created with an upper bound to the rows actually appended (no counting pass)."""
        check(lib().kmb200_table_truncate(self.handle, int(rows)))
        self.rows = int(rows)

    def read_rows(self, row0: int, nrows: int) -> np.ndarray:
        out = np.empty((int(nrows), self.cols), dtype=np.float32)
        check(lib().kmb200_table_read_rows(self.handle, int(row0), int(nrows), _ptr(out)))
        return out

    def fill_blobs(self, centres: np.ndarray, seed: int = 42, row_offset: int = 0) -> None:
        centres = np.ascontiguousarray(centres, dtype=np.float32)
        assert centres.shape[1] == self.cols
        check(lib().kmb200_table_fill_blobs(self.handle, _ptr(centres), centres.shape[0], int(seed), int(row_offset)))

    def release(self) -> None:
        if self._h is not None:
            lib().kmb200_table_free(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.release()
        except Exception:
            pass


class OneDAL:
    """Mirror of the Scala ``OneDAL`` object (OneDAL.scala)."""

    @staticmethod
    def vectorsToMergedNumericTables(data, executorNum: int = 1, batch_rows: int = 1 << 20) -> Table:
        """rows (2-D array, a CSR matrix, a tuple of 1-D columns, or an iterable of such partitions) -> ONE device
        table for this executor
        (OneDAL.scala:92-166: per-partition tables merged per executor).  Empty partitions are
        skipped like the reference's filter at :108-113."""
        def is_csr(p):
            return hasattr(p, "indptr") and hasattr(p, "indices") and hasattr(p, "data") and hasattr(p, "shape")

        def is_columns(p):   # a tuple of 1-D columns (Arrow-style); lists stay "partitions of rows"
            return isinstance(p, tuple) and len(p) > 0 and all(np.ndim(c) == 1 for c in p)

        parts = [data] if isinstance(data, np.ndarray) or is_csr(data) or is_columns(data) else list(data)
        parts = [p if is_csr(p) or is_columns(p) else np.asarray(p) for p in parts]

        def nrows(p):
            return len(p[0]) if is_columns(p) else p.shape[0]

        parts = [p for p in parts if nrows(p) > 0]
        if not parts:
            raise ValueError("no rows")
        cols = len(parts[0]) if is_columns(parts[0]) else parts[0].shape[1]
        table = Table(sum(nrows(p) for p in parts), cols)
        r0 = 0
        for p in parts:
            if is_csr(p):        # scipy.sparse.csr_matrix-like: ml.linalg.SparseVector rows
                table.append_csr(r0, p.indptr, p.indices, p.data)
            elif is_columns(p):  # columnar batch
                table.append_columns(r0, p)
            else:
                for b in range(0, p.shape[0], batch_rows):
                    table.append(r0 + b, p[b:b + batch_rows])
            r0 += nrows(p)
        return table

    @staticmethod
    def makeNumericTable(centers: Sequence[Sequence[float]]) -> np.ndarray:
        """k x d FP64 row-major centres (OneDAL.scala:62-79 builds the same table by k*d JNI calls)."""
        c = np.ascontiguousarray(centers, dtype=np.float64)
        if c.ndim != 2:
            raise ValueError("centers must be k x d")
        return c

    @staticmethod
    def numericTableToVectors(table: np.ndarray):
        return [np.array(r, dtype=np.float64) for r in table]  # OneDAL.scala:37-52

    @staticmethod
    def releaseNumericTables(*tables: Table) -> None:
        for t in tables:  # OneDAL.scala:81-90 -> cFreeDataMemory
            t.release()

    @staticmethod
    def cCheckPlatformCompatibility() -> bool:
        """native/OneDAL.cpp:96-102 asked 'Intel CPU?'; the B200 build asks 'sm_100 GPU?'."""
        return bool(lib().kmb200_device_ok())


@dataclass
class CCLParam:  # java/org/apache/spark/ml/util/CCLParam.java:19-22
    commSize: int = 0
    rankId: int = 0


class OneCCL:
    """Mirror of the Scala ``OneCCL`` object (OneCCL.scala:26-64), one communicator per process."""

    cclParam = CCLParam()
    _comm = None

    @staticmethod
    def getUniqueId() -> bytes:
        """Called once on the driver (or rank 0); the 128 bytes ride in the task closure in place
        of the "ip_port" string of the oneCCL KVS (KMeansDALImpl.scala:39-46)."""
        buf = (C.c_uint8 * 128)()
        check(lib().kmb200_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def init(cls, executor_num: int, rank: int, unique_id: bytes) -> None:
        if executor_num <= 1:
            cls._comm, cls.cclParam = None, CCLParam(1, 0)
            return
        h = C.c_void_p()
        ident = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        check(lib().kmb200_comm_init(int(executor_num), int(rank), ident, C.byref(h)))
        cls._comm = h
        cls.cclParam = CCLParam(lib().kmb200_comm_size(h), lib().kmb200_comm_rank(h))
        # OneCCL.scala:42
        assert executor_num == cls.cclParam.commSize, "executor number should equal to comm world size"

    @classmethod
    def comm(cls):
        return cls._comm

    @classmethod
    def isRoot(cls) -> bool:
        return cls.cclParam.rankId == 0

    @classmethod
    def rankID(cls) -> int:
        return cls.cclParam.rankId

    @classmethod
    def cleanup(cls) -> None:
        if cls._comm is not None:
            lib().kmb200_comm_destroy(cls._comm)
        cls._comm, cls.cclParam = None, CCLParam()


@dataclass
class KMeansResult:  # KMeansResult.java:19-22
    iterationNum: int = 0
    totalCost: float = 0.0


@dataclass
class MLlibKMeansModel:  # what KMeansDALImpl.train returns (KMeansDALImpl.scala:97-99)
    clusterCenters: np.ndarray
    distanceMeasure: str
    trainingCost: float
    numIter: int


class KMeansDALImpl:
    """Mirror of ``KMeansDALImpl`` (KMeansDALImpl.scala:26-111): same constructor arguments,
    ``train(data)`` runs ONE native call per executor and returns rank 0's result."""

    def __init__(self, nClusters: int, maxIterations: int, tolerance: float, distanceMeasure: str,
                 centers, executorNum: int = 1, executorCores: int = 1):
        self.nClusters, self.maxIterations, self.tolerance = int(nClusters), int(maxIterations), float(tolerance)
        self.distanceMeasure = distanceMeasure
        self.centers = OneDAL.makeNumericTable(centers)
        self.executorNum, self.executorCores = int(executorNum), int(executorCores)
        if distanceMeasure != "euclidean":
            # the reference only dispatches euclidean to the DAL path (KMeans.scala:349-351)
            raise ValueError("KMeansDALImpl supports the euclidean distance only")

    def cKMeansDALComputeWithInitCenters(self, table: Table, centers: np.ndarray, cluster_num: int,
                                         tolerance: float, iteration_num: int, executor_num: int,
                                         executor_cores: int, result: KMeansResult) -> Optional[np.ndarray]:
        """The native entry (KMeansDALImpl.scala:105-111 / native/KMeansDALImpl.cpp:175-249).
        Returns the centres on rank 0 and ``None`` elsewhere; fills ``result`` on rank 0 only.
        ``cluster_num`` follows ``centers`` (the reference passes getK next to a
        possibly shorter table); ``executor_cores`` is ignored on the GPU."""
        k, d = centers.shape
        out = np.empty((k, d), dtype=np.float64)
        cost, iters = C.c_double(0.0), C.c_int32(0)
        check(lib().kmb200_fit(table.handle, OneCCL.comm(), _ptr(centers), k, d, float(tolerance),
                               int(iteration_num), _ptr(out), C.byref(cost), C.byref(iters)))
        if OneCCL.isRoot() or OneCCL.comm() is None:
            result.iterationNum, result.totalCost = iters.value, cost.value
            return out
        return None

    def train(self, data, rank: int = 0, unique_id: Optional[bytes] = None) -> Optional[MLlibKMeansModel]:
        """data: this executor's rows (array / list of partitions) or an existing ``Table``.
        With executorNum > 1 every rank calls train(rank=..., unique_id=...) collectively; rank 0
        gets the model, the others ``None`` (the reference collects rank 0's tuple, :65-79)."""
        own = not isinstance(data, Table)
        table = OneDAL.vectorsToMergedNumericTables(data, self.executorNum) if own else data
        try:
            if self.executorNum > 1:
                if unique_id is None:
                    raise ValueError("executorNum > 1 needs the unique id created by OneCCL.getUniqueId()")
                OneCCL.init(self.executorNum, rank, unique_id)
            else:
                OneCCL.init(1, 0, b"")
            result = KMeansResult()
            centers = self.cKMeansDALComputeWithInitCenters(
                table, self.centers, self.nClusters, self.tolerance, self.maxIterations,
                self.executorNum, self.executorCores, result)
            ret = None
            if centers is not None:
                ret = MLlibKMeansModel(centers, self.distanceMeasure, result.totalCost, result.iterationNum)
        finally:
            if self.executorNum > 1:
                OneCCL.cleanup()
            if own:
                OneDAL.releaseNumericTables(table)
        return ret


# ---------------------------------------------------------------------------------------------
# the user-facing estimator (ml.clustering.KMeans) -- parameter names and defaults of the
# reference's shim (spark-3.1.1/ml/clustering/KMeans.scala:45-91,285-327)
# ---------------------------------------------------------------------------------------------
@dataclass
class KMeansSummary:
    predictions: np.ndarray
    clusterSizes: np.ndarray
    k: int
    numIter: int
    trainingCost: float


class KMeansModel:
    def __init__(self, clusterCenters: np.ndarray, trainingCost: float, numIter: int):
        self._centers = np.asarray(clusterCenters, dtype=np.float64)
        self.trainingCost, self.numIter = float(trainingCost), int(numIter)
        self._summary: Optional[KMeansSummary] = None

    def clusterCenters(self):
        return [c.copy() for c in self._centers]

    @property
    def hasSummary(self) -> bool:
        return self._summary is not None

    @property
    def summary(self) -> KMeansSummary:
        if self._summary is None:
            raise RuntimeError("No training summary available for this KMeansModel")
        return self._summary

    def _assign(self, table: Table):
        labels = np.empty(table.rows, dtype=np.int32)
        k, d = self._centers.shape
        sizes = np.empty(k, dtype=np.int64)
        check(lib().kmb200_assign(table.handle, None, _ptr(self._centers), k, d, _ptr(labels), _ptr(sizes), None))
        return labels, sizes

    def transform(self, data) -> np.ndarray:
        """Cluster index of every row, computed on the GPU (kmb200_assign) instead of the JVM UDF
        pass of KMeans.scala:134-142."""
        own = not isinstance(data, Table)
        table = OneDAL.vectorsToMergedNumericTables(data) if own else data
        try:
            labels, _ = self._assign(table)
        finally:
            if own:
                table.release()
        return labels

    def predict(self, features) -> int:
        return int(self.transform(np.asarray(features, dtype=np.float64).reshape(1, -1))[0])


class KMeans:
    """Estimator with the reference's parameter surface.  ``fit`` = KMeans.fit -> trainWithDAL
    (KMeans.scala:330-439): initial centres with their row passes on the GPU (init.py), Lloyd iterations
    in libkmeans_b200.so, summary pass on the GPU."""

    def __init__(self, k: int = 2, maxIter: int = 20, tol: float = 1e-4, initMode: str = "k-means||",
                 initSteps: int = 2, seed: int = 1, distanceMeasure: str = "euclidean"):
        self.k, self.maxIter, self.tol = k, maxIter, tol
        self.initMode, self.initSteps, self.seed, self.distanceMeasure = initMode, initSteps, seed, distanceMeasure
        self.weightCol = None

    # Spark-style setters
    def setK(self, v): self.k = int(v); return self
    def setMaxIter(self, v): self.maxIter = int(v); return self
    def setTol(self, v): self.tol = float(v); return self
    def setSeed(self, v): self.seed = int(v); return self
    def setInitSteps(self, v): self.initSteps = int(v); return self

    def setInitMode(self, v):
        if v not in ("random", "k-means||"):
            raise ValueError(f"initMode must be 'random' or 'k-means||', got {v!r}")
        self.initMode = v
        return self

    def setDistanceMeasure(self, v):
        if v not in ("euclidean", "cosine"):
            raise ValueError(f"unsupported distance measure {v!r}")
        self.distanceMeasure = v
        return self

    def setWeightCol(self, v): self.weightCol = v; return self

    def _init_centers(self, table: Table) -> np.ndarray:
        """Initial centres (KMeans.scala:388-408 -> mllib initRandom / initKMeansParallel): every pass over the rows
        runs on the GPU (oap_mllib_b200/init.py), the driver-side logic on the host."""
        from . import init as _init
        if self.initMode == "random":
            return _init.init_random(table, self.k, self.seed)
        centers, _ = _init.init_kmeans_parallel(table, self.k, self.initSteps, self.seed)
        return centers

    def fit(self, X, initialCenters=None) -> KMeansModel:
        if self.distanceMeasure != "euclidean" or self.weightCol is not None:
            # KMeans.scala:351: the reference sends these to vanilla Spark; this package only
            # holds the accelerated path, so it refuses instead of computing on the CPU.
            raise NotImplementedError("only the euclidean, unweighted DAL path is implemented (no CPU fallback)")
        if not OneDAL.cCheckPlatformCompatibility():
            raise Kmb200Error(-2, "no sm_100 GPU: the accelerated path is unavailable and there is no CPU fallback")
        X = np.asarray(X)
        table = OneDAL.vectorsToMergedNumericTables(X)
        try:
            centers = np.asarray(initialCenters, dtype=np.float64) if initialCenters is not None else self._init_centers(table)
            impl = KMeansDALImpl(self.k, self.maxIter, self.tol, self.distanceMeasure, centers, 1, 1)
            parent = impl.train(table)
            model = KMeansModel(parent.clusterCenters, parent.trainingCost, parent.numIter)
            labels, sizes = model._assign(table)     # summary pass on the GPU (labels + cluster sizes)
            model._summary = KMeansSummary(labels, sizes, centers.shape[0], parent.numIter, parent.trainingCost)
        finally:
            table.release()
        return model


def profile() -> dict:
    return _lib.profile()
