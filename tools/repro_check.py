"""Run-to-run reproducibility of the assignment step at the bench shard's geometry (12.5M x 256, k = 1000).

Every scheme of the FP16-split kernel (cascade / two-level / single pass) is FP64-faithful and, with the stable
bucketing, every reduction has a fixed order: the same call twice must give the same labels and the same bytes.  Any
difference is a bug (a race inside a kernel shows up here long before it shows up as a fault).

Phase A: labels of kmb200_assign at fixed centres, R times per scheme, compared with the first run; rows that differ are
re-scored on the host in FP64 and printed with their true top-2 gap.
Phase B: single Lloyd iterations from the same centres, R times per iteration index, compared bit for bit."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=12_500_000)
    ap.add_argument("--d", type=int, default=256)
    ap.add_argument("--k", type=int, default=1000)
    ap.add_argument("--repeats", type=int, default=300)
    ap.add_argument("--fit-repeats", type=int, default=40)
    ap.add_argument("--schemes", default="cascade,two_level,single_pass")
    args = ap.parse_args()
    import oap_mllib_b200 as m
    from oap_mllib_b200 import _lib
    lib = m.lib()
    _lib.check(lib.kmb200_set_device(0))
    n, d, k = args.rows, args.d, args.k
    centres = np.random.default_rng(42).uniform(-10, 10, (k, d)).astype(np.float32)
    table = m.Table(n, d)
    table.fill_blobs(centres, seed=42)
    init = np.ascontiguousarray(np.vstack([table.read_rows(int(i), 1) for i in (np.arange(k, dtype=np.int64) * n) // k]).astype(np.float64))

    def fit(cen, iters):
        out = np.empty((k, d))
        cost, it = C.c_double(), C.c_int32()
        _lib.check(lib.kmb200_fit(table.handle, None, cen.ctypes.data_as(C.c_void_p), k, d, -1.0, iters,
                                  out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
        return out, cost.value

    def assign(cen):
        labels = np.empty(n, np.int32)
        sizes = np.empty(k, np.int64)
        cost = C.c_double()
        _lib.check(lib.kmb200_assign(table.handle, None, cen.ctypes.data_as(C.c_void_p), k, d,
                                     labels.ctypes.data_as(C.c_void_p), sizes.ctypes.data_as(C.c_void_p), C.byref(cost)))
        return labels, sizes, cost.value

    def exact(rows, cen):
        X = np.vstack([table.read_rows(int(r), 1) for r in rows]).astype(np.float64)
        G = (cen ** 2).sum(1)[None, :] - 2.0 * X @ cen.T
        order = np.argsort(G, axis=1, kind="stable")
        return order[:, 0], order[:, 1], G[np.arange(len(rows)), order[:, 1]] - G[np.arange(len(rows)), order[:, 0]], G

    def dump(nm, offset, count, dtype):
        out = np.empty(count, dtype)
        _lib.check(lib.kmb200_debug_dump(table.handle, nm.encode(), offset, out.ctypes.data_as(C.c_void_p), out.nbytes))
        return out

    def diagnose(rows, labels, ref_labels, cen):
        """Two-level scheme: where did the wrong rows go?  (position in the near-tie copy, what the full-bookkeeping pass
        recorded for them, what the copy holds)"""
        fc = dump("flag_count", 0, 4, np.int32)
        print(f"[diag] counts: pairs {fc[0]} full {fc[1]} near-tie copy {fc[2]} coarse copy {fc[3]}", flush=True)
        cand = dump("cand_list", 0, int(fc[2]), np.int32)
        labc = dump("labels_c", 0, int(fc[2]), np.int32)
        pairs = dump("flag_list", 0, 2 * int(fc[0]), np.int32).reshape(-1, 2)
        fulls = dump("flag_list", (n - int(fc[1])) * 8, 2 * int(fc[1]), np.int32).reshape(-1, 2)
        pos_of = {int(r): i for i, r in enumerate(cand.tolist())}
        pair_of = {int(x): int(y) for x, y in pairs.tolist()}
        full_set = set(fulls[:, 0].tolist())
        idxs = sorted(pos_of[int(r)] for r in rows if int(r) in pos_of)
        print(f"[diag] {len(idxs)} of {len(rows)} wrong rows are in the near-tie copy; positions {idxs[:6]} ... {idxs[-6:]}; "
              f"row tiles {sorted(set(i // 128 for i in idxs))}", flush=True)
        for r in rows[:12]:
            i = pos_of.get(int(r), -1)
            if i < 0:
                print(f"[diag] row {r}: not in the near-tie copy", flush=True)
                continue
            e = pair_of.get(i)
            rec = "full list" if i in full_set else ("not flagged" if e is None else f"pair winner {(e >> 16) & 0xffff} runner {e & 0xffff}")
            xcrow = dump("xc", i * d * 4, d, np.float32)
            same = np.array_equal(xcrow, table.read_rows(int(r), 1)[0])
            b, s2, gap, G = exact([r], cen)
            print(f"[diag] row {r}: copy position {i} (tile {i // 128}, lane {i % 128}) labels_c {labc[i]} final {labels[r]} "
                  f"first-run {ref_labels[r]} exact best {b[0]} second {s2[0]} gap {gap[0]:.4f}; {rec}; copy row == table row: {same}",
                  flush=True)
        # how many rows of the affected tiles are flagged at all / flagged rows in a healthy tile
        tiles = sorted(set(i // 128 for i in idxs))
        for tl in tiles[:4] + [0]:
            members = range(tl * 128, min((tl + 1) * 128, len(cand)))
            fl = sum(1 for i in members if i in pair_of or i in full_set)
            print(f"[diag] copy tile {tl}: {fl} of {len(members)} rows flagged by the full-bookkeeping pass", flush=True)

    cen, _ = fit(init, 3)
    report = {"rows": n, "d": d, "k": k, "phase_a": {}, "phase_b": {}}
    all_schemes = {"cascade": {}, "two_level": {"cascade": 0}, "single_pass": {"two_level": 0}}
    schemes = [(nm, all_schemes[nm]) for nm in args.schemes.split(",") if nm]
    for name, opts in schemes:
        for o, v in opts.items():
            _lib.check(lib.kmb200_set_option(o.encode(), float(v)))
        ref_labels, ref_sizes, ref_cost = assign(cen)
        variant = _lib.profile()["assign_variant"]
        bad_runs, shown = 0, 0
        for r in range(args.repeats):
            labels, sizes, cost = assign(cen)
            if cost != ref_cost or not np.array_equal(labels, ref_labels):
                bad_runs += 1
                rows = np.flatnonzero(labels != ref_labels)
                if shown < 5:
                    shown += 1
                    b, s, gap, G = exact(rows[:8], cen)
                    print(f"[repro] {name}: run {r} differs in {len(rows)} rows (cost {cost!r} vs {ref_cost!r}); first rows "
                          f"{rows[:8].tolist()} labels now {labels[rows[:8]].tolist()} first run {ref_labels[rows[:8]].tolist()} "
                          f"exact best {b.tolist()} second {s.tolist()} gap {gap.tolist()}", flush=True)
                if shown <= 2 and name == "two_level":
                    diagnose(rows, labels, ref_labels, cen)
        for o in opts:
            _lib.check(lib.kmb200_set_option(o.encode(), 1.0))
        report["phase_a"][name] = {"assign_variant": variant, "runs": args.repeats, "runs_that_differ": bad_runs}
        print(f"[repro] phase A {name}: {bad_runs} of {args.repeats} runs differ", flush=True)
    # the three schemes must also agree with each other
    cur = init
    for it in range(5):
        ref = fit(cur, 1)
        bad = 0
        for r in range(args.fit_repeats):
            got = fit(cur, 1)
            if got[1] != ref[1] or not np.array_equal(got[0], ref[0]):
                bad += 1
                if bad <= 3:
                    print(f"[repro] phase B iteration {it}: run {r}: cost {got[1]!r} vs {ref[1]!r}, centres max abs diff "
                          f"{np.abs(got[0] - ref[0]).max():.3e}, empty-cluster events {_lib.profile()['empty_cluster_events']}", flush=True)
        report["phase_b"][f"iteration_{it}"] = {"runs": args.fit_repeats, "runs_that_differ": bad,
                                                "empty_cluster_events": _lib.profile()["empty_cluster_events"]}
        cur = ref[0]
    table.release()
    print(json.dumps(report))
    bad_total = sum(v["runs_that_differ"] for v in report["phase_a"].values()) + sum(v["runs_that_differ"] for v in report["phase_b"].values())
    sys.exit(1 if bad_total else 0)


if __name__ == "__main__":
    main()
