"""Initial centres with the row passes on the GPU.

Mirror of what the reference runs on vanilla Spark before its DAL iterations, mllib/clustering/KMeans.scala (paths
relative to mllib-dal/src/main/scala/org/apache/spark-3.1.1/): ``initKMeansParallel`` :370-432 and ``initRandom``
:354-359.  Every pass over the rows -- the cost update of each step (:395-401), the sampling filter (:406-410), the
candidate weights (:421-424) -- runs in libkmeans_b200.so (kmb200_init_costs, kmb200_init_sample, kmb200_assign); what
stays here is what Spark also runs on the driver: the loop over the steps, the de-duplication and the weighted local
k-means++ over the ~2k candidates (Spark's LocalKMeans.kMeansPlusPlus, java.util.Random seeded with 0).

Stated deviation: ``RDD.takeSample`` (the first centre; ``initRandom``) is Spark's partition-wise sampler plus a shuffle
on the driver and is not reproduced: the first centre is row floor(u N), u the first nextDouble of XORShiftRandom(seed),
and ``init_random`` is a partial Fisher-Yates draw of k distinct rows with the same generator."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import check, lib


class JavaRandom:
    """java.util.Random (48-bit LCG) -- what LocalKMeans.kMeansPlusPlus draws from."""

    def __init__(self, seed: int):
        self.seed = (seed ^ 0x5DEECE66D) & ((1 << 48) - 1)

    def next(self, bits: int) -> int:
        self.seed = (self.seed * 0x5DEECE66D + 0xB) & ((1 << 48) - 1)
        v = self.seed >> (48 - bits)
        return v - (1 << 32) if bits == 32 and v >= 1 << 31 else v

    def nextInt(self, bound: int) -> int:
        r = self.next(31)
        m = bound - 1
        if bound & m == 0:
            return (bound * r) >> 31
        u = r
        while True:
            r = u % bound
            if u - r + m < (1 << 31):
                return r
            u = self.next(31)

    def nextDouble(self) -> float:
        return float((self.next(26) << 27) + self.next(27)) * 2.0 ** -53


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _distinct(rows):
    seen, out = set(), []
    for r in rows:
        key = np.asarray(r, np.float64).tobytes()
        if key not in seen:      # Scala's `distinct` keeps first occurrences, in order
            seen.add(key)
            out.append(np.asarray(r, np.float64))
    return np.array(out)


def local_kmeans_pp(seed: int, points: np.ndarray, weights: np.ndarray, k: int, max_iterations: int = 30) -> np.ndarray:
    """LocalKMeans.kMeansPlusPlus over the (few thousand) weighted candidates: driver-side in Spark, host-side here."""
    rand = JavaRandom(seed)
    n = points.shape[0]
    centers = np.empty((k, points.shape[1]))

    def sq(c):
        return ((points - c) ** 2).sum(1)

    r = rand.nextDouble() * float(np.sum(weights))
    i, cur = 0, 0.0
    while i < n and cur < r:
        cur += weights[i]
        i += 1
    centers[0] = points[i - 1]
    cost = sq(centers[0])
    for c in range(1, k):
        r = rand.nextDouble() * float(np.sum(cost * weights))
        cum, j = 0.0, 0
        while j < n and cum < r:
            cum += weights[j] * cost[j]
            j += 1
        centers[c] = points[0] if j == 0 else points[j - 1]
        cost = np.minimum(cost, sq(centers[c]))
    old = np.full(n, -1)
    it, moved = 0, True
    while moved and it < max_iterations:
        closest = np.stack([sq(c) for c in centers], axis=1).argmin(1)
        moved = bool(np.any(closest != old))
        old = closest
        for j in range(k):
            sel = closest == j
            cnt = float(weights[sel].sum())
            if cnt == 0.0:
                centers[j] = points[rand.nextInt(n)]
            else:
                centers[j] = (weights[sel][:, None] * points[sel]).sum(0) * (1.0 / cnt)
        it += 1
    return centers


def first_center_row(seed_int: int, n: int) -> int:
    return min(n - 1, int(lib().kmb200_xorshift_next_double(seed_int, 0) * n))


def init_kmeans_parallel(table, k: int, init_steps: int, seed: int, comm=None, part_rows=None, part_index0: int = 0):
    """k-means|| over one executor's table.  Returns (centres, [rows chosen in every step])."""
    n, d = table.rows, table.cols
    part = np.ascontiguousarray([0, n] if part_rows is None else part_rows, dtype=np.int64)
    seed_int = lib().kmb200_xorshift_next_int(int(seed))
    centers = [table.read_rows(first_center_row(seed_int, n), 1)[0].astype(np.float64)]
    new = np.ascontiguousarray(np.array(centers))
    total = C.c_double(0.0)
    chosen = []
    for step in range(init_steps):
        if new.shape[0] > 0:     # (no new centres: the costs, and their sum, stay what they were)
            check(lib().kmb200_init_costs(table.handle, comm, _ptr(new), new.shape[0], d, 1 if step == 0 else 0,
                                          C.byref(total)))
        cap = max(16 * k, 1024)
        while True:
            rows = np.empty(cap, np.int64)
            count = C.c_int64(0)
            check(lib().kmb200_init_sample(table.handle, seed_int, step, k, total.value, _ptr(part), len(part) - 1,
                                           part_index0, _ptr(rows), cap, C.byref(count)))
            if count.value <= cap:
                break
            cap = count.value
        rows = rows[:count.value]
        chosen.append(rows.tolist())
        new = np.ascontiguousarray(np.array([table.read_rows(int(r), 1)[0] for r in rows], dtype=np.float64).reshape(-1, d))
        centers.extend(new)
    distinct = _distinct(centers)
    if distinct.shape[0] <= k:
        return distinct, chosen
    # weights = how many rows map to each candidate (:421-424, `countByValue` of findClosest): the summary pass
    sizes = np.empty(distinct.shape[0], np.int64)
    cand = np.ascontiguousarray(distinct)
    check(lib().kmb200_assign(table.handle, comm, _ptr(cand), cand.shape[0], d, None, _ptr(sizes), None))
    return local_kmeans_pp(0, distinct, sizes.astype(np.float64), k, 30), chosen


def init_random(table, k: int, seed: int):
    n = table.rows
    seed_int = lib().kmb200_xorshift_next_int(int(seed))
    # (one generator, advanced by reading its successive doubles: u_i -> i + floor(u_i (n - i)); see the module docstring)
    rows, taken = [], {}
    for i in range(min(k, n)):
        u = lib().kmb200_xorshift_next_double(seed_int, i)
        j = i + min(n - i - 1, int(u * (n - i)))
        vi, vj = taken.get(i, i), taken.get(j, j)
        taken[i], taken[j] = vj, vi
        rows.append(vj)
    return _distinct([table.read_rows(int(r), 1)[0] for r in rows])
