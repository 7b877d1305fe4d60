"""oap_mllib_b200 -- B200-native K-means Lloyd iteration behind OAP MLlib's KMeans API.

The compute lives in ``libkmeans_b200.so`` (hand-written sm_100a CUDA behind the C ABI of
``include/kmb200.h``).  This package is the host-side mirror of the reference's Scala
objects for this path (``KMeansDALImpl``, ``OneDAL``, ``OneCCL``, ``KMeansResult``) -- there
is no JVM/Scala toolchain in this image, see INTEGRATION.md for the Scala/JNI binding.
Importing the package never needs a GPU; any compute call without the CUDA library or an
sm_100 device raises (there is no CPU fallback).
"""
from ._lib import Kmb200Error, lib, library_path  # noqa: F401
from .kmeans import (KMeans, KMeansDALImpl, KMeansModel, KMeansResult, KMeansSummary,  # noqa: F401
                     OneCCL, OneDAL, Table)

from .pca import PCA, PCADALImpl, PCAModel, PCAResult  # noqa: F401

__all__ = ["PCA", "PCADALImpl", "PCAModel", "PCAResult", "KMeans", "KMeansDALImpl", "KMeansModel", "KMeansResult", "KMeansSummary", "OneCCL",
           "OneDAL", "Table", "Kmb200Error", "lib", "library_path"]
