/* Plain-C FP64 restatement of the OAP MLlib K-means Lloyd path -- TEST INFRASTRUCTURE ONLY.
 *
 * The checker, never the product: only tests/, __graft_entry__.smoke() and bench.py's CPU
 * baseline may load the library built from this file (oracle/Makefile -> libkmeans_oracle.so).
 *
 * Follows (reference = /root/reference, read-only):
 *   - driver loop / iteration count / returned cost:
 *       mllib-dal/src/main/native/KMeansDALImpl.cpp:195-248
 *   - convergence test (every centre: sequential sum_d (new-old)^2 <= tol*tol):
 *       mllib-dal/src/main/native/KMeansDALImpl.cpp:135-168
 *   - per-iteration arithmetic = Intel oneDAL 2021.2.0 lloydDense (not in /root/reference;
 *     called at KMeansDALImpl.cpp:70-77 and :103-130).  Restated from its published
 *     algorithm: G[j] = ||c_j||^2 - 2 x.c_j, strict '>' scan (lowest index wins), objective
 *     += min + ||x||^2 (EXPANDED form, as oneDAL), sums[min] += x, count[min]++, centre =
 *     sum * (1/count), empty cluster <- farthest unused candidate row, objective -= its
 *     distance.  The numpy oracle (kmeans_oracle.py) states the same thing with dgemm and a
 *     direct-form objective; tests check that the two agree.
 *
 * Parity pin: tests/golden/reference_kats.json (IntelKMeansSuite.scala / sample data).
 * Tie-breaking, empty-cluster rule and multi-rank merge order are NOT pinned by any
 * reference test: parity unpinned for those.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* one rank's step1Local over rows [r0, r1) */
static void step1_local(const double *X, int64_t r0, int64_t r1, int d, const double *C,
                        const double *cn, int k, int64_t *counts, double *sums,
                        double *objective, double *cand_dist, int64_t *cand_row,
                        int32_t *labels) {
    memset(counts, 0, sizeof(int64_t) * (size_t)k);
    memset(sums, 0, sizeof(double) * (size_t)k * d);
    for (int j = 0; j < k; j++) { cand_dist[j] = -1.0; cand_row[j] = -1; }
    double obj = 0.0;
    for (int64_t i = r0; i < r1; i++) {
        const double *x = X + (size_t)i * d;
        double xn = 0.0;
        for (int c = 0; c < d; c++) xn += x[c] * x[c];
        double min = 0.0; int minIdx = 0;
        for (int j = 0; j < k; j++) {
            const double *cj = C + (size_t)j * d;
            double dot = 0.0;
            for (int c = 0; c < d; c++) dot += x[c] * cj[c];
            double g = cn[j] - 2.0 * dot;
            if (j == 0 || min > g) { min = g; minIdx = j; }
        }
        double dist = min + xn;
        obj += dist;
        counts[minIdx]++;
        double *s = sums + (size_t)minIdx * d;
        for (int c = 0; c < d; c++) s[c] += x[c];
        if (labels) labels[i] = minIdx;
        /* keep the k largest distances, descending, earlier row wins ties */
        if (dist > cand_dist[k - 1]) {
            int p = k - 1;
            while (p > 0 && dist > cand_dist[p - 1]) {
                cand_dist[p] = cand_dist[p - 1]; cand_row[p] = cand_row[p - 1]; p--;
            }
            cand_dist[p] = dist; cand_row[p] = i;
        }
    }
    *objective = obj;
}

/* Whole cKMeansDALComputeWithInitCenters.  Returns 0, or -1 when there are more empty
 * clusters than positive-distance candidates.  labels (may be NULL) = last assignment. */
int oracle_lloyd(const double *X, int64_t n, int d, const double *init, int k, double tol,
                 int max_iter, int n_ranks, double *out_centers, double *out_cost,
                 int32_t *out_iters, int32_t *labels) {
    size_t kd = (size_t)k * d;
    double *C = (double *)malloc(sizeof(double) * kd);
    double *newC = (double *)malloc(sizeof(double) * kd);
    double *cn = (double *)malloc(sizeof(double) * k);
    int64_t *counts = (int64_t *)calloc((size_t)k, sizeof(int64_t));
    double *sums = (double *)calloc(kd, sizeof(double));
    int64_t *pc = (int64_t *)malloc(sizeof(int64_t) * (size_t)k * n_ranks);
    double *ps = (double *)malloc(sizeof(double) * kd * n_ranks);
    double *pobj = (double *)malloc(sizeof(double) * n_ranks);
    double *pcd = (double *)malloc(sizeof(double) * (size_t)k * n_ranks);
    int64_t *pcr = (int64_t *)malloc(sizeof(int64_t) * (size_t)k * n_ranks);
    memcpy(C, init, sizeof(double) * kd);
    double cost = 0.0;
    int it = 0, converged = 0, rc = 0;
    for (it = 0; it < max_iter && !converged; it++) {
        for (int j = 0; j < k; j++) {
            double s = 0.0;
            for (int c = 0; c < d; c++) s += C[(size_t)j * d + c] * C[(size_t)j * d + c];
            cn[j] = s;
        }
#pragma omp parallel for schedule(static)
        for (int r = 0; r < n_ranks; r++) {
            int64_t r0 = (int64_t)((double)n * r / n_ranks), r1 = (int64_t)((double)n * (r + 1) / n_ranks);
            step1_local(X, r0, r1, d, C, cn, k, pc + (size_t)r * k, ps + (size_t)r * kd, pobj + r,
                        pcd + (size_t)r * k, pcr + (size_t)r * k, labels);
        }
        /* step2Master.compute(): merge in rank order */
        memset(counts, 0, sizeof(int64_t) * (size_t)k);
        memset(sums, 0, sizeof(double) * kd);
        double obj = 0.0;
        for (int r = 0; r < n_ranks; r++) {
            for (int j = 0; j < k; j++) counts[j] += pc[(size_t)r * k + j];
            for (size_t e = 0; e < kd; e++) sums[e] += ps[(size_t)r * kd + e];
            obj += pobj[r];
        }
        /* finalizeCompute() */
        int used_total = 0;
        char *used = (char *)calloc((size_t)k * n_ranks, 1);
        for (int j = 0; j < k; j++) {
            if (counts[j] > 0) {
                double inv = 1.0 / (double)counts[j];
                for (int c = 0; c < d; c++) newC[(size_t)j * d + c] = sums[(size_t)j * d + c] * inv;
            } else {
                /* next unused candidate: largest distance, ties -> lower rank / earlier row */
                int bi = -1; double bd = 0.0;
                for (int e = 0; e < k * n_ranks; e++)
                    if (!used[e] && pcr[e] >= 0 && pcd[e] > bd) { bd = pcd[e]; bi = e; }
                if (bi < 0 || used_total >= k) { rc = -1; free(used); goto done; }
                used[bi] = 1; used_total++;
                memcpy(newC + (size_t)j * d, X + (size_t)pcr[bi] * d, sizeof(double) * d);
                obj -= bd;
            }
        }
        free(used);
        cost = obj;
        /* areAllCentersConverged */
        converged = 1;
        for (int j = 0; j < k && converged; j++) {
            double s = 0.0;
            for (int c = 0; c < d; c++) {
                double df = newC[(size_t)j * d + c] - C[(size_t)j * d + c];
                s += df * df;
            }
            if (!(s <= tol * tol)) converged = 0;
        }
        memcpy(C, newC, sizeof(double) * kd);
    }
done:
    memcpy(out_centers, C, sizeof(double) * kd);
    *out_cost = cost;
    *out_iters = it;
    free(C); free(newC); free(cn); free(counts); free(sums);
    free(pc); free(ps); free(pobj); free(pcd); free(pcr);
    return rc;
}
