"""CPU baselines for bench.py  --  TEST / MEASUREMENT INFRASTRUCTURE ONLY (never imported by oap_mllib_b200).

The reference's K-means step is oneDAL's ``kmeans::Distributed<step1Local,double>`` on ``executor_cores`` TBB
threads (/root/reference/mllib-dal/src/main/native/KMeansDALImpl.cpp:70-77,188): 512-row blocks, one MKL dgemm
per block, a fused argmin / scatter-add loop over the block, thread-local partial results merged at the end.
oneDAL cannot be built or run in this image, so bench.py times stand-ins on the box's host cores
and reports the FASTEST of them next to the GPU number:

* ``oracle``      oracle/kmeans_oracle.py -- the FP64 checker.  Written for clarity (whole-table temporaries, a
                  lexsort for the candidates): a good oracle and a slow baseline; timed for reference only.
* ``blocked``     this file: the oneDAL shape restated for speed -- row blocks handed to a thread pool, one
                  single-threaded BLAS dgemm per block (X_blk . (-2 C)^T + ||c||^2), argmin, expanded-form
                  objective (min + ||x||^2, as oneDAL), per-block sort + segmented sums into thread-local FP64
                  partials; no n x d temporaries.  FP64 or FP32 GEMM.
* ``sklearn``     ``sklearn.cluster.KMeans(algorithm="lloyd", init=<same centres>, n_init=1, tol=0)``: an
                  independent, optimised (Cython + BLAS + OpenMP over row chunks) Lloyd; FP64 and FP32.  Its per-
                  iteration time is taken as (t(max_iter=3) - t(max_iter=1)) / 2 so that the fixed costs of fit()
                  (validation, the final labelling pass) cancel.
"""
from __future__ import annotations

import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def blocked_step(X: np.ndarray, C: np.ndarray, threads: int, block: int = 2048):
    """One Lloyd step (sums, counts, objective, new centres) in the dtype of X; row blocks over a thread pool."""
    from threadpoolctl import threadpool_limits
    n, d = X.shape
    k = C.shape[0]
    Cw = np.ascontiguousarray(C, dtype=X.dtype)
    Ct2 = np.ascontiguousarray((-2.0 * Cw).T)
    cn = np.einsum("ij,ij->i", Cw, Cw)
    nblk = (n + block - 1) // block
    per = (nblk + threads - 1) // threads

    def work(t):
        sums = np.zeros((k, d), np.float64)
        counts = np.zeros(k, np.int64)
        obj = 0.0
        G = np.empty((block, k), X.dtype)
        for b in range(t * per, min(nblk, (t + 1) * per)):
            xb = X[b * block:(b + 1) * block]
            g = G[:xb.shape[0]]
            np.matmul(xb, Ct2, out=g)
            g += cn
            lab = np.argmin(g, axis=1)
            mins = g[np.arange(xb.shape[0]), lab]
            obj += float(mins.sum(dtype=np.float64) + np.einsum("ij,ij->", xb, xb, dtype=np.float64))
            order = np.argsort(lab, kind="stable")
            sl = lab[order]
            starts = np.flatnonzero(np.r_[True, sl[1:] != sl[:-1]])
            sums[sl[starts]] += np.add.reduceat(xb[order], starts, axis=0, dtype=np.float64)
            counts += np.bincount(lab, minlength=k)
        return sums, counts, obj

    with threadpool_limits(limits=1):          # one BLAS thread per block: the pool provides the parallelism
        with ThreadPoolExecutor(threads) as ex:
            parts = list(ex.map(work, range(threads)))
    sums = sum(p[0] for p in parts)
    counts = sum(p[1] for p in parts)
    obj = sum(p[2] for p in parts)
    new = np.array(C, dtype=np.float64, copy=True)
    nz = counts > 0
    new[nz] = sums[nz] * (1.0 / counts[nz])[:, None]
    return new, obj, counts


def time_blocked(X, C, threads, iters=1):
    blocked_step(X[:4096], C, threads)   # thread pool / BLAS warm-up
    t0 = time.perf_counter()
    for _ in range(iters):
        C, _, _ = blocked_step(X, C, threads)
    return (time.perf_counter() - t0) / iters


def time_oracle(X64, C, iters=1):
    from oracle import kmeans_oracle as ko
    t0 = time.perf_counter()
    for _ in range(iters):
        C, _, _, _ = ko.lloyd_step([X64], C)
    return (time.perf_counter() - t0) / iters


def time_sklearn(X, C):
    import warnings

    from sklearn.cluster import KMeans

    def fit(m):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            t0 = time.perf_counter()
            KMeans(n_clusters=C.shape[0], init=np.asarray(C, X.dtype), n_init=1, tol=0.0, max_iter=m, algorithm="lloyd",
                   copy_x=False).fit(X)
            return time.perf_counter() - t0
    t1 = fit(1)
    t3 = fit(3)
    return max((t3 - t1) / 2.0, 1e-9)


def run_all(d, k, sample_rows, seed=42, which=("blocked_f64", "sklearn_f64", "sklearn_f32", "blocked_f32", "oracle")):
    """Seconds per Lloyd iteration over `sample_rows` rows of d columns, k centres, for each baseline.  Returns
    {name: seconds}, cores."""
    cores = host_cores()
    rng = np.random.default_rng(seed)
    centres = rng.uniform(-10, 10, (k, d)).astype(np.float32)
    X32 = centres[rng.integers(0, k, sample_rows)]
    X32 += rng.standard_normal((sample_rows, d), dtype=np.float32)
    X64 = X32.astype(np.float64)          # the reference's path is FP64 end to end (KMeansDALImpl.cpp:32)
    idx = (np.arange(k, dtype=np.int64) * sample_rows) // k
    C0 = X64[idx].copy()
    out = {}
    for name in which:
        try:
            if name == "blocked_f64":
                out[name] = time_blocked(X64, C0, cores)
            elif name == "blocked_f32":
                out[name] = time_blocked(X32, C0, cores)
            elif name == "sklearn_f64":
                out[name] = time_sklearn(X64, C0)
            elif name == "sklearn_f32":
                out[name] = time_sklearn(X32, C0)
            elif name == "oracle":
                out[name] = time_oracle(X64, C0)
        except Exception as e:   # a missing optional package must not take the bench line down
            out[name] = None
            out[name + "_error"] = repr(e)[:200]
    return out, cores
