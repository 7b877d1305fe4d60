// Implementation of sym_eig.h.  A translation unit of its own, compiled by the host compiler, so that the O(n^3) loops
// (all stride-1 thanks to the column-major work array) get an x86-64-v3 (AVX2 + FMA) clone next to the portable one: the dynamic
// linker picks by CPU (GNU ifunc).  3x faster than the baseline x86-64 code at n = 256.
#include "sym_eig.h"

namespace kmb {

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
__attribute__((target_clones("arch=x86-64-v3", "default")))
#endif
bool sym_eig_desc(const double *a, int n, double *w, double *v) {
    if (n <= 0) return true;
    std::vector<double> zbuf((size_t)n * n), dbuf(n), ebuf(n);
    double *__restrict__ z = zbuf.data(), *__restrict__ d = dbuf.data(), *__restrict__ e = ebuf.data();
#define Z(i, j) z[(size_t)(j) * n + (i)]   /* column-major: the inner loops run down columns */
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) Z(i, j) = 0.5 * (a[(size_t)i * n + j] + a[(size_t)j * n + i]);
    // ---- Householder tridiagonalisation, accumulating the transformation in z
    for (int j = 0; j < n; ++j) d[j] = Z(n - 1, j);
    for (int i = n - 1; i > 0; --i) {
        double scale = 0.0, h = 0.0;
        for (int k = 0; k < i; ++k) scale += std::fabs(d[k]);
        if (scale == 0.0) {
            e[i] = d[i - 1];
            for (int j = 0; j < i; ++j) {
                d[j] = Z(i - 1, j);
                Z(i, j) = 0.0;
                Z(j, i) = 0.0;
            }
        } else {
            for (int k = 0; k < i; ++k) {
                d[k] /= scale;
                h += d[k] * d[k];
            }
            double f = d[i - 1];
            double g = std::sqrt(h);
            if (f > 0) g = -g;
            e[i] = scale * g;
            h -= f * g;
            d[i - 1] = f - g;
            for (int j = 0; j < i; ++j) e[j] = 0.0;
            for (int j = 0; j < i; ++j) {
                f = d[j];
                Z(j, i) = f;
                g = e[j] + Z(j, j) * f;
                for (int k = j + 1; k <= i - 1; ++k) {
                    g += Z(k, j) * d[k];
                    e[k] += Z(k, j) * f;
                }
                e[j] = g;
            }
            f = 0.0;
            for (int j = 0; j < i; ++j) {
                e[j] /= h;
                f += e[j] * d[j];
            }
            const double hh = f / (h + h);
            for (int j = 0; j < i; ++j) e[j] -= hh * d[j];
            for (int j = 0; j < i; ++j) {
                f = d[j];
                g = e[j];
                for (int k = j; k <= i - 1; ++k) Z(k, j) -= (f * e[k] + g * d[k]);
                d[j] = Z(i - 1, j);
                Z(i, j) = 0.0;
            }
        }
        d[i] = h;
    }
    for (int i = 0; i < n - 1; ++i) {
        Z(n - 1, i) = Z(i, i);
        Z(i, i) = 1.0;
        const double h = d[i + 1];
        if (h != 0.0) {
            for (int k = 0; k <= i; ++k) d[k] = Z(k, i + 1) / h;
            for (int j = 0; j <= i; ++j) {
                double g = 0.0;
                for (int k = 0; k <= i; ++k) g += Z(k, i + 1) * Z(k, j);
                for (int k = 0; k <= i; ++k) Z(k, j) -= g * d[k];
            }
        }
        for (int k = 0; k <= i; ++k) Z(k, i + 1) = 0.0;
    }
    for (int j = 0; j < n; ++j) {
        d[j] = Z(n - 1, j);
        Z(n - 1, j) = 0.0;
    }
    Z(n - 1, n - 1) = 1.0;
    e[0] = 0.0;
    // ---- implicit QL on the tridiagonal matrix (d: diagonal, e: sub-diagonal), rotating the columns of z
    for (int i = 1; i < n; ++i) e[i - 1] = e[i];
    e[n - 1] = 0.0;
    double f = 0.0, tst1 = 0.0;
    const double eps = std::ldexp(1.0, -52);
    for (int l = 0; l < n; ++l) {
        tst1 = std::max(tst1, std::fabs(d[l]) + std::fabs(e[l]));
        int m = l;
        while (m < n) {
            if (std::fabs(e[m]) <= eps * tst1) break;
            ++m;
        }
        if (m >= n) m = n - 1;
        if (m > l) {
            int iter = 0;
            do {
                if (++iter > 200) return false;
                double g = d[l];
                double p = (d[l + 1] - g) / (2.0 * e[l]);
                double r = std::hypot(p, 1.0);
                if (p < 0) r = -r;
                d[l] = e[l] / (p + r);
                d[l + 1] = e[l] * (p + r);
                const double dl1 = d[l + 1];
                double h = g - d[l];
                for (int i = l + 2; i < n; ++i) d[i] -= h;
                f += h;
                p = d[m];
                double c = 1.0, c2 = c, c3 = c;
                const double el1 = e[l + 1];
                double s = 0.0, s2 = 0.0;
                for (int i = m - 1; i >= l; --i) {
                    c3 = c2;
                    c2 = c;
                    s2 = s;
                    g = c * e[i];
                    h = c * p;
                    r = std::hypot(p, e[i]);
                    e[i + 1] = s * r;
                    s = e[i] / r;
                    c = p / r;
                    p = c * d[i] - s * g;
                    d[i + 1] = h + s * (c * g + s * d[i]);
                    for (int k = 0; k < n; ++k) {
                        h = Z(k, i + 1);
                        Z(k, i + 1) = s * Z(k, i) + c * h;
                        Z(k, i) = c * Z(k, i) - s * h;
                    }
                }
                p = -s * s2 * c3 * el1 * e[l] / dl1;
                e[l] = s * p;
                d[l] = c * p;
            } while (std::fabs(e[l]) > eps * tst1);
        }
        d[l] += f;
        e[l] = 0.0;
    }
    // ---- descending order (stable: equal eigenvalues keep their order), eigenvectors into the rows of v
    std::vector<int> order(n);
    for (int i = 0; i < n; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [d](int x, int y) { return d[x] > d[y]; });
    for (int r = 0; r < n; ++r) {
        const int c = order[r];
        w[r] = d[c];
        for (int k = 0; k < n; ++k) v[(size_t)r * n + k] = Z(k, c);
    }
    return true;
#undef Z
}

}  // namespace kmb
