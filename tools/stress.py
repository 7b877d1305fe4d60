"""Stress loop for the production assign kernel (assign_pair_h16_kernel and its cascade), one process per GPU.

    python tools/stress.py [--rows 12500000] [--fits 200] [--iters 23] [--sweep-seconds 120]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... tools/stress.py ...

Part 1 ("bench shard"): a d = 256, k = 1000 table of --rows rows per GPU (12.5M = the shard of the headline config on
8 GPUs) filled by the device generator; --fits consecutive fits of --iters iterations each, every one from the same
initial centres.  A fit must return status 0; its cost and cluster sizes must equal the first fit's (cost to 1e-12
relative: the row order inside a cluster is fixed, the sums are FP64).  With N ranks every fit is collective (one
NCCL all-reduce per iteration), which is the configuration the one unexplained launch failure of round 1 was seen in.

Part 2 ("shape sweep", rank-local, no collective): short fits over the shapes that drive the kernel's barrier rings
through their corner cases -- one centroid tile per row tile (NT = 1), NT <= the number of resident B groups (the
B producer idles), NT > groups, odd and even numbers of row tiles per CTA pair, a dummy last tile, one / two A
buffers (ld <= 128 / 256), every BN -- each checked against the single-pass bookkeeping of the same kernel (cluster
sizes equal, cost to 1e-12).  Runs until --sweep-seconds have passed.

Prints one JSON summary line per part on rank 0; exits non-zero on the first failure."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=12_500_000)
    ap.add_argument("--fits", type=int, default=200)
    ap.add_argument("--iters", type=int, default=23)
    ap.add_argument("--sweep-seconds", type=float, default=120.0)
    args = ap.parse_args()
    import torch

    import oap_mllib_b200 as m
    from oap_mllib_b200 import _lib
    rank, world, local = (int(os.environ.get(v, d)) for v, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    _lib.check(m.lib().kmb200_set_device(local))
    dist = None
    comm = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ident = [m.OneCCL.getUniqueId() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        m.OneCCL.init(world, rank, ident[0])
        comm = m.OneCCL.comm()

    def fit(table, init, iters, cm=None):
        k, d = init.shape
        out = np.empty((k, d))
        cost, it = C.c_double(), C.c_int32()
        _lib.check(m.lib().kmb200_fit(table.handle, cm, init.ctypes.data_as(C.c_void_p), k, d, -1.0, iters,
                                      out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
        assert it.value == iters
        return out, cost.value

    def sizes_cost(table, cen, cm=None):
        k, d = cen.shape
        sizes = np.empty(k, np.int64)
        cost = C.c_double()
        cen = np.ascontiguousarray(cen)
        _lib.check(m.lib().kmb200_assign(table.handle, cm, cen.ctypes.data_as(C.c_void_p), k, d, None,
                                         sizes.ctypes.data_as(C.c_void_p), C.byref(cost)))
        return sizes, cost.value

    # ------------------------------------------------------------------ part 1: the bench shard, over and over
    d, k = 256, 1000
    n = args.rows
    centres = np.random.default_rng(42).uniform(-10, 10, (k, d)).astype(np.float32)
    table = m.Table(n, d)
    table.fill_blobs(centres, seed=42, row_offset=rank * n)
    init = np.zeros((k, d))
    for j, gi in enumerate(((np.arange(k, dtype=np.int64) * n * world) // k).tolist()):
        if rank * n <= gi < (rank + 1) * n:
            init[j] = table.read_rows(gi - rank * n, 1)[0]
    if dist is not None:
        t = torch.from_numpy(init).cuda()
        dist.all_reduce(t)
        init = np.ascontiguousarray(t.cpu().numpy())
    t0 = time.time()
    first = None
    bit_identical = True
    variants = set()
    for f in range(args.fits):
        cen, cost = fit(table, init, args.iters, comm)
        variants.add(_lib.profile()["assign_variant"])
        if first is None:
            first = (cen, cost)
        else:
            bit_identical &= cost == first[1] and np.array_equal(cen, first[0])
            if not abs(cost - first[1]) <= 1e-12 * abs(first[1]) or not np.allclose(cen, first[0], rtol=1e-12, atol=0):
                print(f"[stress] rank {rank}: fit {f} differs from fit 0: cost {cost!r} vs {first[1]!r}", flush=True)
                sys.exit(1)
    table.release()
    if rank == 0:
        print(json.dumps({"part": "bench shard", "ranks": world, "rows_per_rank": n, "d": d, "k": k, "fits": args.fits,
                          "iterations_per_fit": args.iters, "assign_variants_seen": sorted(variants),
                          "identical_results": True, "bit_identical": bool(bit_identical), "final_cost": first[1], "seconds": round(time.time() - t0, 1)}),
              flush=True)
    if dist is not None:
        dist.barrier()

    # ------------------------------------------------------------------ part 2: barrier-ring corner cases
    # (rows, d, k): 148 SMs -> 74 CTA pairs; rows / 128 row tiles are dealt to the pairs two at a time
    shapes = []
    for dd in (32, 64, 128, 160, 256):
        for kk in (20, 40, 100, 128, 300, 512, 640, 1000, 1300):
            for rows in (2 * 128 + 5, 148 * 128, 149 * 128 + 1, 296 * 128, 297 * 128 + 77, 445 * 128 + 3, 270_000):
                if kk * 6 <= rows:   # (tiny tables with many centres only produce empty clusters)
                    shapes.append((rows, dd, kk))
    rng = np.random.default_rng(1234 + rank)
    rng.shuffle(shapes)
    t0 = time.time()
    done = 0
    for rows, dd, kk in shapes:
        if time.time() - t0 > args.sweep_seconds:
            break
        cen0 = rng.uniform(-10, 10, (kk, dd)).astype(np.float32)
        tb = m.Table(rows, dd)
        tb.fill_blobs(cen0, seed=rows + dd + kk)
        ini = np.ascontiguousarray(np.vstack([tb.read_rows(int(i), 1) for i in (np.arange(kk) * rows) // kk]).astype(np.float64))
        res = {}
        _lib.check(m.lib().kmb200_set_option(b"two_level", 2.0))   # cascade + two-level forced on, whatever the row count
        cen, _ = fit(tb, ini, 3)
        for two_level in (2, 0):   # ... against the single pass, at the same centres
            _lib.check(m.lib().kmb200_set_option(b"two_level", float(two_level)))
            res[two_level] = sizes_cost(tb, cen)
        _lib.check(m.lib().kmb200_set_option(b"two_level", 1.0))
        tb.release()
        if not np.array_equal(res[2][0], res[0][0]) or not abs(res[2][1] - res[0][1]) <= 1e-12 * abs(res[0][1]):
            print(f"[stress] rank {rank}: shape {rows} x {dd}, k = {kk}: cascade and single pass disagree "
                  f"({res[2][1]!r} vs {res[0][1]!r})", flush=True)
            sys.exit(1)
        done += 1
    if rank == 0:
        print(json.dumps({"part": "shape sweep", "shapes_run": done, "of": len(shapes), "seconds": round(time.time() - t0, 1),
                          "all_agree": True}), flush=True)
    if dist is not None:
        m.OneCCL.cleanup()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
