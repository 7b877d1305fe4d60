// Eigen-decomposition of a real symmetric matrix on the host, FP64: Householder reduction to tridiagonal form followed
// by the implicit QL iteration (the classic EISPACK tred2 / tql2 pair, restated).  This is the "master step" of the
// PCA path: the reference runs oneDAL's pca::Distributed<step2Master> on rank 0's host cores for the same d x d problem
// (mllib-dal/src/main/native/PCADALImpl.cpp:111-136).  O(d^3).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

namespace kmb {

// a: n x n symmetric, row-major (only read).  On success: w[0..n) eigenvalues DESCENDING, rows of v (n x n, row-major)
// the matching orthonormal eigenvectors.  Returns false if the QL iteration did not converge (never seen).
bool sym_eig_desc(const double *a, int n, double *w, double *v);

}  // namespace kmb
