"""Shared helpers for the parity tests (test infrastructure; may use oracle/)."""
import ctypes
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_kats():
    with open(os.path.join(ROOT, "tests", "golden", "reference_kats.json")) as f:
        return json.load(f)


def c_oracle_lloyd(lib, X, init, tol, max_iter, n_ranks=1):
    X = np.ascontiguousarray(X, np.float64)
    init = np.ascontiguousarray(init, np.float64)
    n, d = X.shape
    k = init.shape[0]
    out = np.empty((k, d))
    cost, it = ctypes.c_double(), ctypes.c_int32()
    lab = np.empty(n, np.int32)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = lib.oracle_lloyd(P(X), ctypes.c_int64(n), d, P(init), k, ctypes.c_double(tol), max_iter, n_ranks,
                          P(out), ctypes.byref(cost), ctypes.byref(it), P(lab))
    return rc, out, cost.value, it.value, lab


# parity tolerances from BASELINE.json's north_star
CENTER_RTOL = 1e-5
COST_RTOL = 1e-5
NEAR_TIE_REL = 1e-6


def assert_centers_close(got, want, rtol=CENTER_RTOL):
    got, want = np.asarray(got), np.asarray(want)
    scale = np.maximum(np.linalg.norm(want, axis=1, keepdims=True), 1e-30)
    err = np.abs(got - want) / scale
    assert err.max() <= rtol, f"centres differ: max relative (to centre norm) error {err.max():.3e} > {rtol}"


def assert_labels_match(got, X64, C64, rel=NEAR_TIE_REL):
    """Labels identical to the FP64 oracle except rows whose best/second-best squared distances are
    within `rel` (relative) of each other; for those the GPU label must be one of the tied pair."""
    from oracle import kmeans_oracle as ko
    want, best, second = ko.assign(X64, C64)
    bad = np.nonzero(got != want)[0]
    if len(bad) == 0:
        return 0
    tie = ko.near_tie_mask(best, second, rel)
    hard = bad[~tie[bad]]
    assert len(hard) == 0, (f"{len(hard)} of {len(got)} labels differ outside near-ties; first rows {hard[:5]}, "
                            f"gpu {got[hard[:5]]}, oracle {want[hard[:5]]}, "
                            f"rel gaps {((second - best) / np.maximum(second, 1e-300))[hard[:5]]}")
    # a tied row must still be assigned to a centre at (almost) the minimum distance
    d_got = ((X64[bad] - C64[got[bad]]) ** 2).sum(1)
    assert np.all(d_got <= best[bad] * (1 + 4 * rel) + 1e-300)
    return len(bad)
