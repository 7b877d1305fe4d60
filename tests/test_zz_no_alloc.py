"""kmb200_cluster_sizes_noalloc: cluster sizes computed without any device allocation inside the
library (its scratch is a fixed array in the library image).  It exists for hosts that cannot
cudaMalloc when they call in -- an executor whose GPU memory belongs entirely to a pool allocator,
or a latency-bound loop that must not trigger the device synchronisation an allocation implies.

Collected last on purpose (file name): a run with -x has then already executed every other test.
The check: its output against the sizes the regular, allocating entry point kmb200_assign returns
for the same labels -- exact, they are integers -- on two calls in a row.  The second call uses
fewer clusters and different labels, so counts left in the shared scratch by the first call would
show up as wrong sizes."""
import ctypes as C

import numpy as np
import pytest

from oracle import kmeans_oracle as ko

pytestmark = pytest.mark.gpu


def _assign(m, X, centers):
    k, d = centers.shape
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        labels = np.empty(t.rows, np.int32)
        sizes = np.empty(k, np.int64)
        cen = np.ascontiguousarray(centers, np.float64)
        m._lib.check(m.lib().kmb200_assign(t.handle, None, cen.ctypes.data_as(C.c_void_p), k, d,
                                           labels.ctypes.data_as(C.c_void_p), sizes.ctypes.data_as(C.c_void_p), None))
    finally:
        t.release()
    return labels, sizes


def test_no_alloc_sizes_match_assign_twice_in_a_row(built_lib):
    import torch

    import oap_mllib_b200 as m
    assert m.lib().kmb200_device_ok() == 1
    m._lib.check(m.lib().kmb200_set_device(0))
    torch.cuda.set_device(0)
    X, _, _ = ko.gaussian_blobs(60000, 24, 40, seed=21)
    cases = [(40, ko.row_init_centers(X, 40)), (13, ko.row_init_centers(X[::-1].copy(), 13))]
    stream = torch.cuda.Stream()
    for k, centers in cases:                         # two calls in a row on the same device
        labels, want = _assign(m, X, centers)        # the existing entry point (allocates its workspace)
        lab_dev = torch.from_numpy(labels).cuda()    # the caller owns inputs and outputs
        out_dev = torch.full((k,), -1, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        rc = m.lib().kmb200_cluster_sizes_noalloc(C.c_void_p(lab_dev.data_ptr()), labels.size, k,
                                                  C.c_void_p(out_dev.data_ptr()), C.c_void_p(stream.cuda_stream))
        m._lib.check(rc)
        stream.synchronize()
        got = out_dev.cpu().numpy()
        assert np.array_equal(got, want), (k, got[:8], want[:8])
        assert got.sum() == labels.size
    # beyond the built-in scratch: refused with a clear message, nothing launched
    big = 65536 + 1
    out_dev = torch.zeros(big, dtype=torch.int64, device="cuda")
    rc = m.lib().kmb200_cluster_sizes_noalloc(C.c_void_p(lab_dev.data_ptr()), 10, big,
                                              C.c_void_p(out_dev.data_ptr()), None)
    assert rc == -6
    assert b"65536" in m.lib().kmb200_last_error()
    # the largest accepted k works too (global-atomic path of the histogram kernel), n = 0 gives zeros
    out_dev = torch.full((65536,), -1, dtype=torch.int64, device="cuda")
    m._lib.check(m.lib().kmb200_cluster_sizes_noalloc(None, 0, 65536, C.c_void_p(out_dev.data_ptr()), None))
    torch.cuda.synchronize()
    assert int(out_dev.abs().sum().item()) == 0
