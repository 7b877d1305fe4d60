"""First-contact diagnostics for the tcgen05 assign kernel (run on the GPU box).  Prints the score
error of the tensor-core path against FP64 for a few shapes, with enough detail to localise a
descriptor / swizzle / TMEM-layout mistake.  Test infrastructure: uses the oracle helpers."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oap_mllib_b200 as m  # noqa: E402
from oracle import kmeans_oracle as ko  # noqa: E402


def scores(X, Cn):
    n, d = X.shape
    k = Cn.shape[0]
    t = m.OneDAL.vectorsToMergedNumericTables(X)
    try:
        kpad = C.c_int32()
        Cc = np.ascontiguousarray(Cn, np.float64)
        m._lib.check(m.lib().kmb200_debug_scores(t.handle, Cc.ctypes.data_as(C.c_void_p), k, d, None, C.byref(kpad)))
        sc = np.empty((n, kpad.value), np.float32)
        t0 = time.time()
        m._lib.check(m.lib().kmb200_debug_scores(t.handle, Cc.ctypes.data_as(C.c_void_p), k, d,
                                                 sc.ctypes.data_as(C.c_void_p), C.byref(kpad)))
        dt = time.time() - t0
    finally:
        t.release()
    return sc, dt


def main():
    shapes = [(256, 32, 16), (300, 64, 100), (2048, 256, 1000), (1024, 512, 300)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
    print("device_ok", m.lib().kmb200_device_ok(), m.lib().kmb200_version())
    for passes in (3, 1):
        m._lib.check(m.lib().kmb200_set_option(b"passes", float(passes)))
        for n, d, k in shapes:
            X, _, _ = ko.gaussian_blobs(n, d, k, seed=d + k)
            Cn = ko.row_init_centers(X, k)
            X64 = X.astype(np.float64)
            sc, dt = scores(X, Cn)
            exact = (Cn ** 2).sum(1)[None, :] - 2.0 * X64 @ Cn.T
            cmax = np.sqrt((Cn ** 2).sum(1).max())
            scale = 2.0 * np.sqrt((X64 ** 2).sum(1))[:, None] * cmax + cmax ** 2
            err = np.abs(sc[:, :k].astype(np.float64) - exact) / scale
            lab_tc = sc[:, :k].argmin(1)
            lab_ex = exact.argmin(1)
            print(f"passes={passes} shape={n}x{d} k={k} kpad={sc.shape[1]}: max_norm_err={err.max():.3e} "
                  f"mean={err.mean():.3e} argmin_mismatch={(lab_tc != lab_ex).sum()} pad_inf={np.isinf(sc[:, k:]).all()} "
                  f"({dt*1e3:.1f} ms)")
            if err.max() > 1e-3:
                print("  sample tc   :", sc[:2, :6])
                print("  sample exact:", exact[:2, :6])
                bad_rows = np.unique(np.nonzero(err > 1e-3)[0])
                bad_cols = np.unique(np.nonzero(err > 1e-3)[1])
                print("  bad rows (first 16):", bad_rows[:16], "count", len(bad_rows))
                print("  bad cols (first 16):", bad_cols[:16], "count", len(bad_cols))
    m._lib.check(m.lib().kmb200_set_option(b"passes", 3.0))


if __name__ == "__main__":
    main()
