"""Stress loop for the Gram kernel of the PCA path (gram_pair_kernel: a warp-specialised barrier-ring kernel like the
assign kernel, so it gets the same treatment as tools/stress.py).

Part 1: a d = 256 table of --rows rows from the device generator; --calls Gram matrices of it, which must all be
bit-identical (every sum has a fixed order); in between, calls with other ring depths / rows per stage / segment lengths,
whose results must agree with the first to the accuracy of the scheme.
Part 2: random shapes (1 .. 40 000 rows, 1 .. 600 columns, with and without centring) against the FP64 oracle until
--sweep-seconds have passed.

Prints one JSON line per part; exits non-zero on the first failure."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=12_500_000)
    ap.add_argument("--calls", type=int, default=300)
    ap.add_argument("--sweep-seconds", type=float, default=90.0)
    args = ap.parse_args()
    import oap_mllib_b200 as m
    from oap_mllib_b200 import _lib, pca
    from oracle import pca_oracle as po
    lib = m.lib()
    _lib.check(lib.kmb200_set_device(0))

    def opt(name, v):
        _lib.check(lib.kmb200_set_option(name.encode(), float(v)))

    d = 256
    centres = np.random.default_rng(42).uniform(-10, 10, (1000, d)).astype(np.float32)
    t = m.Table(args.rows, d)
    t.fill_blobs(centres, seed=42)
    t0 = time.time()
    first = pca.gram(t, True)[0]
    scale = np.sqrt(np.outer(np.diag(first), np.diag(first)))
    variants = [dict(pca_rows=32), dict(pca_ls=4, pca_hs=2), dict(pca_flush_rows=64), dict(pca_flush_rows=1024, pca_promote=3),
                dict(pca_rows=32, pca_ls=8, pca_hs=4, pca_flush_rows=32)]
    worst = 0.0
    for c in range(args.calls):
        g = pca.gram(t, True)[0]
        if not np.array_equal(g, first):
            print(f"[pca stress] call {c} differs from call 0: max rel diff {(np.abs(g - first) / scale).max():.3e}", flush=True)
            sys.exit(1)
        if c % 20 == 0:
            v = variants[(c // 20) % len(variants)]
            for k, val in v.items():
                opt(k, val)
            g = pca.gram(t, True)[0]
            for k in v:
                opt(k, {"pca_rows": 64, "pca_flush_rows": 256, "pca_promote": 16}.get(k, 0))
            diff = (np.abs(g - first) / scale).max()
            worst = max(worst, diff)
            if not diff <= 1e-6 + 6e-9 * v.get("pca_flush_rows", 256):   # the truncation of the level-1 adds grows with the segment
                print(f"[pca stress] variant {v} disagrees with the default: {diff:.3e}", flush=True)
                sys.exit(1)
    t.release()
    print(json.dumps({"part": "repeated calls", "rows": args.rows, "d": d, "calls": args.calls, "bit_identical": True,
                      "largest_difference_between_schemes": worst, "seconds": round(time.time() - t0, 1)}), flush=True)

    rng = np.random.default_rng(99)
    t0 = time.time()
    done, worst = 0, 0.0
    while time.time() - t0 < args.sweep_seconds:
        n = int(rng.choice([1, 2, 31, 33, 63, 65, 127, 129, rng.integers(130, 40_000)]))
        dd = int(rng.choice([1, 3, 31, 32, 33, 64, 100, 128, 129, 255, 256, 257, rng.integers(258, 600)]))
        center = bool(rng.integers(0, 2)) and n > 1
        X = (rng.standard_normal((n, dd)) * 10.0 ** rng.uniform(-3, 3, dd) + rng.standard_normal(dd) * 10.0 ** rng.uniform(-3, 3)).astype(np.float32)
        tb = m.OneDAL.vectorsToMergedNumericTables(X)
        g, mean, cnt = pca.gram(tb, center)
        tb.release()
        g_ref, _, _ = po.centred_gram([X], center)
        sc = np.sqrt(np.outer(np.diag(g_ref), np.diag(g_ref)))
        sc[sc == 0] = 1.0
        err = float((np.abs(g - g_ref) / sc).max())
        worst = max(worst, err)
        # (without centring, a column whose mean dwarfs its spread has the SAME split residual in every row: the 2^-22
        #  terms then add up instead of averaging out)
        if not (err <= (2e-6 if center else 4e-6) and cnt == n):
            print(f"[pca stress] shape {n} x {dd} center={center}: error {err:.3e}", flush=True)
            sys.exit(1)
        done += 1
    print(json.dumps({"part": "shape sweep", "shapes": done, "largest_error": worst, "seconds": round(time.time() - t0, 1)}), flush=True)


if __name__ == "__main__":
    main()
