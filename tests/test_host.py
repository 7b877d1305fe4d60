"""CPU tests of the host logic and of the drop-in boundary: the C-ABI library loads without a GPU,
exports every symbol include/kmb200.h declares, and fails loudly (no CPU fallback) on compute."""
import ctypes
import os
import re

import numpy as np
import pytest

from helpers import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "kmb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(kmb200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built_lib):
    handle = ctypes.CDLL(built_lib)
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(handle, n), f"{n} declared in include/kmb200.h but not exported"


def test_python_binding_covers_the_header(built_lib):
    from oap_mllib_b200 import _lib
    assert sorted(_lib.SIGNATURES) == header_symbols()


def test_jni_symbols_exported(built_lib):
    handle = ctypes.CDLL(built_lib)
    for n in ["Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALComputeWithInitCenters",
              "Java_org_apache_spark_ml_util_OneDAL_00024_cSetDoubleBatch",
              "Java_org_apache_spark_ml_util_OneDAL_00024_setNumericTableValue",
              "Java_org_apache_spark_ml_util_OneDAL_00024_cFreeDataMemory",
              "Java_org_apache_spark_ml_util_OneDAL_00024_cCheckPlatformCompatibility",
              "Java_org_apache_spark_ml_util_OneDAL_00024_cNewTable",
              "Java_org_apache_spark_ml_util_OneDAL_00024_cNewCenters",
              "Java_org_apache_spark_ml_util_OneDAL_00024_cGetCenters",
              "Java_org_apache_spark_ml_util_OneCCL_00024_c_1getUniqueId",
              "Java_org_apache_spark_ml_util_OneCCL_00024_c_1init",
              "Java_org_apache_spark_ml_util_OneCCL_00024_c_1cleanup",
              "Java_org_apache_spark_ml_util_OneCCL_00024_isRoot",
              "Java_org_apache_spark_ml_util_OneCCL_00024_rankID",
              "Java_org_apache_spark_ml_util_OneCCL_00024_setEnv",
              "Java_org_apache_spark_ml_util_OneCCL_00024_c_1setDevice",
              "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALAssign",
              "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansInitCosts",
              "Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansInitSample",
              "Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDAL",            # native/javah/org_apache_spark_ml_feature_PCADALImpl.h
              "Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDALCentered"]:
        assert hasattr(handle, n), n


def test_no_cpu_fallback_without_gpu(built_lib):
    import oap_mllib_b200 as m
    if m.lib().kmb200_device_ok():
        pytest.skip("a GPU is present; the no-fallback check is for CPU boxes")
    with pytest.raises(m.Kmb200Error):
        m.Table(8, 3)
    with pytest.raises(m.Kmb200Error):
        m.KMeans(k=2).fit(np.zeros((4, 2)))
    with pytest.raises(m.Kmb200Error):
        m.PCA(k=1).fit(np.zeros((4, 2)))
    assert m.OneDAL.cCheckPlatformCompatibility() is False


def test_missing_library_raises(monkeypatch, tmp_path):
    from oap_mllib_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setenv("KMB200_LIBRARY", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.Kmb200Error):
        _lib.lib()
    monkeypatch.setattr(_lib, "_lib", None)


def test_estimator_parameter_surface():
    from oap_mllib_b200 import KMeans
    from helpers import load_kats
    d = load_kats()["kat_defaults"]
    km = KMeans()
    assert (km.k, km.maxIter, km.initMode, km.initSteps, km.tol, km.distanceMeasure) == \
        (d["k"], d["maxIter"], d["initMode"], d["initSteps"], d["tol"], d["distanceMeasure"])
    with pytest.raises(ValueError):
        KMeans().setDistanceMeasure("no_such_a_measure")     # IntelKMeansSuite.scala:105-107
    with pytest.raises(ValueError):
        KMeans().setInitMode("no_such_a_mode")
    km.setK(9).setMaxIter(33).setTol(1e-3).setSeed(7).setInitMode("random").setInitSteps(3)
    assert (km.k, km.maxIter, km.tol, km.seed, km.initMode, km.initSteps) == (9, 33, 1e-3, 7, "random", 3)


def test_dal_impl_rejects_non_euclidean():
    from oap_mllib_b200 import KMeansDALImpl
    with pytest.raises(ValueError):
        KMeansDALImpl(2, 5, 1e-4, "cosine", np.zeros((2, 3)))
