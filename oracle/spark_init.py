"""CPU restatement of the initialisation the reference runs on vanilla Spark before its DAL iterations  --  TEST
INFRASTRUCTURE ONLY.

Follows /root/reference/mllib-dal/src/main/scala/org/apache/spark-3.1.1/mllib/clustering/KMeans.scala:
``initKMeansParallel`` :370-432 (k-means|| by Bahmani et al.: a first random centre, ``initializationSteps`` rounds that
keep every row with probability 2 k cost / sumCosts, de-duplication, and -- when more than k candidates remain -- a
weighted local k-means++ over them) and ``initRandom`` :354-359.  Two pieces live in Spark itself, not in the reference
tree, and are restated here from Spark's published source:

* ``org.apache.spark.util.random.XORShiftRandom`` (seed hashed by two MurmurHash3.bytesHash rounds, xorshift 21/35/4 in
  ``next(bits)``, the rest inherited from java.util.Random) -- **unpinned**: neither the file nor vectors for it are under
  /root/reference;
* ``org.apache.spark.mllib.clustering.LocalKMeans.kMeansPlusPlus`` with ``java.util.Random`` (the 48-bit LCG; pinned by
  its well-known first values, tests/test_init.py).

Deviation, stated: ``RDD.takeSample`` (first centre; ``initRandom``) goes through Spark's partition-wise Bernoulli / gap
samplers and a Fisher-Yates shuffle on the driver; that machinery is not restated.  The first centre here is row
floor(u N) with u the first ``nextDouble`` of ``XORShiftRandom(seed)``, and ``init_random`` draws k distinct rows by
a partial Fisher-Yates with the same generator.  Distances are plain FP64 ||x - c||^2 (Spark's fastSquaredDistance is an
FP64 evaluation of the same quantity)."""
from __future__ import annotations

import numpy as np

M32, M64 = (1 << 32) - 1, (1 << 64) - 1


def _rotl32(x, r):
    return ((x << r) | (x >> (32 - r))) & M32


def murmur3_bytes_hash(data: bytes, seed: int) -> int:
    """scala.util.hashing.MurmurHash3.bytesHash"""
    def mix_last(h, k):
        k = (k * 0xcc9e2d51) & M32
        k = _rotl32(k, 15)
        k = (k * 0x1b873593) & M32
        return h ^ k
    h, i, left = seed & M32, 0, len(data)
    while left >= 4:
        k = data[i] | (data[i + 1] << 8) | (data[i + 2] << 16) | (data[i + 3] << 24)
        h = mix_last(h, k)
        h = _rotl32(h, 13)
        h = (h * 5 + 0xe6546b64) & M32
        i += 4
        left -= 4
    k = 0
    if left == 3:
        k ^= data[i + 2] << 16
    if left >= 2:
        k ^= data[i + 1] << 8
    if left >= 1:
        k ^= data[i]
        h = mix_last(h, k)
    h ^= len(data)
    h ^= h >> 16
    h = (h * 0x85ebca6b) & M32
    h ^= h >> 13
    h = (h * 0xc2b2ae35) & M32
    h ^= h >> 16
    return h


def _to_i32(x):
    x &= M32
    return x - (1 << 32) if x >= 1 << 31 else x


class JavaRandom:
    """java.util.Random"""

    def __init__(self, seed: int):
        self.seed = (seed ^ 0x5DEECE66D) & ((1 << 48) - 1)

    def next(self, bits: int) -> int:
        self.seed = (self.seed * 0x5DEECE66D + 0xB) & ((1 << 48) - 1)
        return _to_i32(self.seed >> (48 - bits))

    def next_int(self, bound: int | None = None) -> int:
        if bound is None:
            return self.next(32)
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

    def next_double(self) -> float:
        return float(((self.next(26) << 27) + self.next(27))) * 2.0 ** -53


class XORShiftRandom(JavaRandom):
    """org.apache.spark.util.random.XORShiftRandom"""

    def __init__(self, init: int):
        b = (init & M64).to_bytes(8, "big")
        low = murmur3_bytes_hash(b, 0x3c074a61)
        high = murmur3_bytes_hash(b, low)
        self.seed = (high << 32) | low

    def next(self, bits: int) -> int:
        x = self.seed ^ ((self.seed << 21) & M64)
        x ^= x >> 35
        x ^= (x << 4) & M64
        self.seed = x
        return _to_i32(x & ((1 << bits) - 1)) if bits == 32 else x & ((1 << bits) - 1)


def _sq_dists(X, C):
    return ((X[:, None, :] - C[None, :, :]) ** 2).sum(2) if X.shape[0] * C.shape[0] * X.shape[1] < 2e7 else \
        np.stack([((X - c) ** 2).sum(1) for c in C], axis=1)


def _distinct(rows):
    seen, out = set(), []
    for r in rows:
        key = np.asarray(r, np.float64).tobytes()
        if key not in seen:
            seen.add(key)
            out.append(np.asarray(r, np.float64))
    return np.array(out)


def local_kmeans_pp(seed: int, points: np.ndarray, weights: np.ndarray, k: int, max_iterations: int = 30) -> np.ndarray:
    """LocalKMeans.kMeansPlusPlus (weighted k-means++ then up to max_iterations Lloyd steps over the candidates)."""
    rand = JavaRandom(seed)
    n = points.shape[0]
    centers = np.empty((k, points.shape[1]))
    r = rand.next_double() * float(np.sum(weights))
    i, cur = 0, 0.0
    while i < n and cur < r:
        cur += weights[i]
        i += 1
    centers[0] = points[i - 1]
    cost = ((points - centers[0]) ** 2).sum(1)
    for c in range(1, k):
        total = float(np.sum(cost * weights))
        r = rand.next_double() * total
        cum, j = 0.0, 0
        while j < n and cum < r:
            cum += weights[j] * cost[j]
            j += 1
        centers[c] = points[0] if j == 0 else points[j - 1]
        cost = np.minimum(cost, ((points - centers[c]) ** 2).sum(1))
    old = np.full(n, -1)
    it, moved = 0, True
    while moved and it < max_iterations:
        closest = _sq_dists(points, centers).argmin(1)
        moved = bool(np.any(closest != old))
        old = closest
        for j in range(k):
            sel = closest == j
            cnt = float(weights[sel].sum())
            if cnt == 0.0:
                centers[j] = points[rand.next_int(n)]
            else:
                centers[j] = (weights[sel][:, None] * points[sel]).sum(0) * (1.0 / cnt)
        it += 1
    return centers


def first_center_row(seed_int: int, n: int) -> int:
    return min(n - 1, int(XORShiftRandom(seed_int).next_double() * n))


def init_kmeans_parallel(X, k: int, init_steps: int, seed: int, part_rows=None, part_index0: int = 0):
    """-> (centres, rows chosen per step).  X: all rows (FP64); part_rows: offsets of the Spark partitions."""
    X = np.asarray(X, np.float64)
    n = X.shape[0]
    part_rows = [0, n] if part_rows is None else list(part_rows)
    seed_int = XORShiftRandom(seed).next_int()
    centers = [X[first_center_row(seed_int, n)]]
    new = np.array(centers)
    costs = np.full(n, np.inf)
    chosen_rows = []
    for step in range(init_steps):
        costs = np.minimum(costs, _sq_dists(X, new).min(1))
        total = float(np.sum(costs))
        rows = []
        for p in range(len(part_rows) - 1):
            rand = XORShiftRandom(_to_i32(seed_int ^ (step << 16) ^ (part_index0 + p)))
            for r in range(part_rows[p], part_rows[p + 1]):
                if rand.next_double() < 2.0 * costs[r] * k / total:
                    rows.append(r)
        chosen_rows.append(rows)
        new = X[rows] if rows else np.empty((0, X.shape[1]))
        centers.extend(new)
        if not rows:
            new = np.array(centers[:1])
    distinct = _distinct(centers)
    if distinct.shape[0] <= k:
        return distinct, chosen_rows
    owner = _sq_dists(X, distinct).argmin(1)
    weights = np.bincount(owner, minlength=distinct.shape[0]).astype(np.float64)
    return local_kmeans_pp(0, distinct, weights, k, 30), chosen_rows


def init_random(X, k: int, seed: int):
    X = np.asarray(X, np.float64)
    n = X.shape[0]
    rand = XORShiftRandom(XORShiftRandom(seed).next_int())
    rows, taken = [], {}
    for i in range(min(k, n)):            # partial Fisher-Yates: u_i -> i + floor(u_i (n - i))
        j = i + min(n - i - 1, int(rand.next_double() * (n - i)))
        vi, vj = taken.get(i, i), taken.get(j, j)
        taken[i], taken[j] = vj, vi
        rows.append(vj)
    return _distinct(X[rows])
