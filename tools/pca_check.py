"""Gram matrix / PCA of the CUDA path against the FP64 oracle on a list of shapes (diagnostics; the tests proper are
tests/test_gpu_pca.py).

    python tools/pca_check.py [--swap 0|1] [--flush 1024] [--big]

Prints, per shape, the largest entry error of the centred Gram matrix relative to sqrt(G_ii G_jj), the largest relative
eigenvalue error, and the time of the call."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--swap", type=int, default=0)
    ap.add_argument("--flush", type=int, default=128)
    ap.add_argument("--big", action="store_true")
    ap.add_argument("--shapes", default="")
    ap.add_argument("--knock", action="store_true", help="timing experiments on the big table (results invalid)")
    args = ap.parse_args()
    import oap_mllib_b200 as m
    from oap_mllib_b200 import _lib, pca
    from oracle import pca_oracle as po
    lib = m.lib()
    _lib.check(lib.kmb200_set_device(0))
    _lib.check(lib.kmb200_set_option(b"pca_desc_swap", float(args.swap)))
    _lib.check(lib.kmb200_set_option(b"pca_flush_rows", float(args.flush)))
    shapes = [(64, 256), (40, 256), (5000, 256), (3000, 100), (2000, 32), (70000, 256), (4000, 300), (3000, 512), (2500, 700)]
    if args.shapes:
        shapes = [tuple(int(v) for v in s.split("x")) for s in args.shapes.split(",")]
    rng = np.random.default_rng(7)
    for n, d in shapes:
        basis = np.linalg.qr(rng.standard_normal((d, d)))[0]
        X = ((rng.standard_normal((n, d)) * (10.0 * 0.97 ** np.arange(d) + 0.05)) @ basis + rng.uniform(-3, 3, d)).astype(np.float32)
        t = m.OneDAL.vectorsToMergedNumericTables(X)
        for center in (True, False):
            t0 = time.time()
            g, mean, cnt = pca.gram(t, center)
            dt = time.time() - t0
            gr, mr, nr = po.centred_gram([X], center)
            dg = np.sqrt(np.outer(np.diag(gr), np.diag(gr)))
            err = np.abs(g - gr) / dg
            i, j = np.unravel_index(np.argmax(err), err.shape)
            w = np.sort(np.linalg.eigvalsh(g))[::-1]
            wr = np.sort(np.linalg.eigvalsh(gr))[::-1]
            evr, axr = po.eig_desc(gr)
            evg, axg = po.eig_desc(g)
            lead = min(8, d)
            print(f"[pca] {n} x {d} center={int(center)} swap={args.swap} flush={args.flush}: rows {cnt} gram err max {err.max():.3e} at ({i},{j}) "
                  f"median {np.median(err):.3e}; eigenvalue rel err {np.abs(w - wr).max() / wr[0]:.3e}; explained-variance abs err "
                  f"{np.abs(evg / evg.sum() - evr / evr.sum()).max():.3e}; first {lead} axes abs err {np.abs(axg[:lead] - axr[:lead]).max():.3e}; "
                  f"mean err {np.abs(mean - mr).max():.2e}; {dt * 1e3:.1f} ms", flush=True)
            if err.max() > 1e-3:
                bad = err > 1e-3
                print(f"      {bad.sum()} of {bad.size} entries off; rows with an error: {np.flatnonzero(bad.any(1))[:12].tolist()} ... "
                      f"cols: {np.flatnonzero(bad.any(0))[:12].tolist()}; g[0,:4] {g[0, :4]} ref {gr[0, :4]}", flush=True)
        t.release()
    if args.big:
        n, d = 20_000_000, 256
        centres = rng.uniform(-10, 10, (1000, d)).astype(np.float32)
        t = m.Table(n, d)
        t.fill_blobs(centres, seed=3)
        for flush in (128, 64, 256, 1024, 16384):
            _lib.check(lib.kmb200_set_option(b"pca_flush_rows", float(flush)))
            pca.gram(t, True)
            t0 = time.time()
            g, mean, cnt = pca.gram(t, True)
            dt = time.time() - t0
            if flush == 128:
                g_ref = g
            print(f"[pca] big {n} x {d} flush {flush}: {dt * 1e3:.1f} ms per call; max rel diff to flush=128: "
                  f"{(np.abs(g - g_ref) / np.sqrt(np.outer(np.diag(g_ref), np.diag(g_ref)))).max():.3e}", flush=True)
        if args.knock:
            def opt(name, v):
                _lib.check(lib.kmb200_set_option(name.encode(), float(v)))
            for rows_per_stage, ls, hs in [(64, 2, 4), (64, 4, 2), (32, 4, 8)]:
                for flush in (128, 256, 512, 16384):
                    opt("pca_rows", rows_per_stage), opt("pca_ls", ls), opt("pca_hs", hs), opt("pca_flush_rows", flush)
                    pca.gram(t, True)
                    pca.gram(t, True)
                    print(f"[pca] timing rows/stage {rows_per_stage} rings {ls}/{hs} flush {flush}: gram {_lib.profile()['assign_ms']:.2f} ms", flush=True)
            opt("pca_role_prof", 1)
            for rows_per_stage, ls, hs in [(64, 2, 4), (32, 4, 8)]:
                opt("pca_rows", rows_per_stage), opt("pca_ls", ls), opt("pca_hs", hs), opt("pca_flush_rows", 256)
                print(f"[pca] role profile, rows/stage {rows_per_stage}, rings {ls}/{hs}, flush 256:", flush=True)
                pca.gram(t, True)
            opt("pca_role_prof", 0), opt("pca_ls", 0), opt("pca_hs", 0), opt("pca_rows", 64)
        if args.knock and False:
            for rows_per_stage, flush, ko in [(64, 128, 0), (64, 128, 1), (64, 128, 2), (64, 128, 4), (64, 128, 8), (64, 128, 15), (64, 16384, 0),
                                              (64, 256, 0), (32, 128, 0), (32, 16384, 0)]:
                if True:
                    _lib.check(lib.kmb200_set_option(b"pca_rows", float(rows_per_stage)))
                    _lib.check(lib.kmb200_set_option(b"pca_flush_rows", float(flush)))
                    _lib.check(lib.kmb200_set_option(b"pca_knockout", float(ko)))
                    pca.gram(t, True)
                    pca.gram(t, True)
                    pr = _lib.profile()
                    print(f"[pca] knock rows/stage {rows_per_stage} flush {flush} knockout {ko}: gram {pr['assign_ms']:.2f} ms, stats {pr['accumulate_ms']:.2f} ms", flush=True)
            _lib.check(lib.kmb200_set_option(b"pca_knockout", 0.0))
        t.release()


if __name__ == "__main__":
    main()
