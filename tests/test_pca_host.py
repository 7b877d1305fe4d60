"""Host side of the PCA path without a GPU: the eigen-solver of csrc/sym_eig.h through the C ABI (kmb200_sym_eig) against
LAPACK (numpy.linalg.eigh), and the Python mirror's table handling (PCADALImpl.scala:108-136)."""
import ctypes as C

import numpy as np
import pytest


def sym_eig(lib, a):
    n = a.shape[0]
    a = np.ascontiguousarray(a, np.float64)
    w, v = np.empty(n), np.empty((n, n))
    rc = lib.kmb200_sym_eig(a.ctypes.data_as(C.c_void_p), n, w.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return w, v


@pytest.mark.parametrize("n", [1, 2, 3, 8, 33, 100, 256])
def test_eigen_solver_against_lapack(built_lib, n):
    import oap_mllib_b200 as m
    lib = m.lib()
    rng = np.random.default_rng(n)
    b = rng.standard_normal((n, n))
    a = b @ b.T * 10.0 ** rng.uniform(-3, 3) + np.diag(rng.uniform(0, 1, n))
    w, v = sym_eig(lib, a)
    ref = np.sort(np.linalg.eigvalsh(a))[::-1]
    assert np.abs(w - ref).max() <= 1e-13 * abs(ref[0]) * max(n, 8)
    assert np.all(np.diff(w) <= 0)
    assert np.abs(v @ v.T - np.eye(n)).max() <= 1e-13 * max(n, 8)
    assert np.abs(v @ a @ v.T - np.diag(w)).max() <= 1e-13 * abs(ref[0]) * max(n, 8)


def test_eigen_solver_degenerate_inputs(built_lib):
    import oap_mllib_b200 as m
    lib = m.lib()
    w, v = sym_eig(lib, np.zeros((5, 5)))                       # all zero
    assert np.array_equal(w, np.zeros(5)) and np.abs(v @ v.T - np.eye(5)).max() < 1e-15
    w, v = sym_eig(lib, np.diag([3.0, 3.0, 1.0, 3.0]))          # repeated eigenvalue, already diagonal
    assert np.allclose(w, [3, 3, 3, 1]) and np.abs(v @ v.T - np.eye(4)).max() < 1e-15
    a = np.ones((6, 6))                                         # rank one
    w, v = sym_eig(lib, a)
    assert abs(w[0] - 6.0) < 1e-13 and np.abs(w[1:]).max() < 1e-13
    assert np.allclose(np.abs(v[0]), 1 / np.sqrt(6.0))
    assert lib.kmb200_sym_eig(None, 3, None, None) != 0         # argument check, no crash


def test_mirror_table_handling():
    from oap_mllib_b200.pca import PCADALImpl
    axes = np.arange(12.0).reshape(3, 4)                        # three axes (rows) of dimension 4
    pc = PCADALImpl.getPrincipleComponentsFromDAL(axes, 2)
    assert pc.shape == (4, 2) and np.array_equal(pc[:, 1], axes[1])
    with pytest.raises(ValueError):
        PCADALImpl.getPrincipleComponentsFromDAL(axes, 4)       # "k should be less or equal to row number"
    ev = PCADALImpl.getExplainedVarianceFromDAL(np.array([[4.0, 3.0, 2.0, 1.0]]), 2)
    assert np.allclose(ev, [0.4, 0.3])
