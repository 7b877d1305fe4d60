"""Multi-GPU parity check, one process per GPU (launch with torch.distributed.run): every rank
uploads ITS shard of a seeded data set, all ranks call KMeansDALImpl.train collectively (NCCL
unique id from rank 0, one all-reduce per iteration), rank 0 compares with the FP64 oracle run with
the same number of ranks.  Also covers the empty-cluster candidate exchange across ranks."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist

    import oap_mllib_b200 as m
    from oap_mllib_b200.sharding import shard_bounds
    from oracle import kmeans_oracle as ko
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    m._lib.check(m.lib().kmb200_set_device(local))
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    failures = []
    cases = [("blobs", 60000, 64, 100, 8), ("blobs", 30011, 256, 300, 6), ("blobs", 200000, 32, 16, 6), ("empty", 0, 0, 0, 5),
             ("blobs2", 30011, 256, 300, 6),   # blobs2: two-level bookkeeping forced on (fast pass + near-tie pass)
             ("many_empty", 60000, 16, 160, 4),  # 60 clusters empty at once: candidates from every rank, merged
             ("blobs", 600000, 256, 1000, 4)]    # >= 2^18 rows per rank at 2 ranks: the cascade with its default options
    for kind, n, d, k, iters in cases:
        m._lib.check(m.lib().kmb200_set_option(b"two_level", 2.0 if kind == "blobs2" else 1.0))
        if kind == "blobs2":
            kind = "blobs"
        if kind == "blobs":
            X, _, _ = ko.gaussian_blobs(n, d, k, seed=n + k)
            init = ko.row_init_centers(X, k)
        elif kind == "many_empty":
            X, _, _ = ko.gaussian_blobs(n, d, k - 60, seed=n + k)
            init = ko.row_init_centers(X, k)
            init[::2][:60] = 1.0e3 + np.arange(60)[:, None]
        else:
            X = np.array([[0.0, 0], [1, 0], [0, 1], [10, 10], [11, 10], [0.5, 0.5], [10.5, 10.2], [3, 3]], np.float32)
            init = np.array([[0.0, 0], [10, 10], [100, 100], [-50, -50]])
            n, d, k = X.shape[0], 2, 4
        r0, r1 = shard_bounds(n, world)[rank]
        ident = [m.OneCCL.getUniqueId() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        impl = m.KMeansDALImpl(k, iters, 1e-4, "euclidean", init, world, 1)
        model = impl.train(X[r0:r1], rank=rank, unique_id=ident[0])
        if rank == 0:
            want_c, want_cost, want_it = ko.lloyd(X.astype(np.float64), init, 1e-4, iters, n_ranks=world)
            scale = np.maximum(np.linalg.norm(want_c, axis=1, keepdims=True), 1e-30)
            err = (np.abs(model.clusterCenters - want_c) / scale).max()
            ok = model.numIter == want_it and err <= 1e-5 and abs(model.trainingCost - want_cost) <= 1e-5 * abs(want_cost) + 1e-12
            print(f"[multigpu] {kind} n={n} d={d} k={k} world={world}: iters {model.numIter}/{want_it} "
                  f"centre err {err:.2e} cost {model.trainingCost:.6f}/{want_cost:.6f} {'OK' if ok else 'FAIL'}", flush=True)
            if not ok:
                failures.append((kind, n, d, k))
        else:
            assert model is None
        dist.barrier()
    # ---- PCA (section 8 row f4): every rank's shard in its own table, PCADALImpl.train collectively (one all-reduce of the
    # column statistics, one of the d x d Gram matrix), rank 0 against the oracle's distributed route (one QR factor per rank)
    from oracle import pca_oracle as po
    for n, d, kk in [(50001, 96, 8), (300000, 256, 16), (20000, 300, 5)]:
        rng = np.random.default_rng(n + d)
        basis = np.linalg.qr(rng.standard_normal((d, d)))[0]
        X = ((rng.standard_normal((n, d)) * (10.0 * 0.93 ** np.arange(d) + 0.05)) @ basis + rng.uniform(-5, 5, d)).astype(np.float32)
        bounds = shard_bounds(n, world)
        r0, r1 = bounds[rank]
        ident = [m.OneCCL.getUniqueId() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        model = m.PCADALImpl(kk, world, 1).train(X[r0:r1], rank=rank, unique_id=ident[0])
        if rank == 0:
            ev, axes = po.pca_svd_dense([X[a:b] for a, b in bounds])
            pc, expl = po.model(ev, axes, kk)
            e_pc, e_ex = np.abs(model.pc - pc).max(), np.abs(model.explainedVariance - expl).max()
            ok = e_pc <= 1e-6 and e_ex <= 1e-6
            print(f"[multigpu] pca n={n} d={d} k={kk} world={world}: component err {e_pc:.2e} explained-variance err {e_ex:.2e} "
                  f"{'OK' if ok else 'FAIL'}", flush=True)
            if not ok:
                failures.append(("pca", n, d, kk))
        else:
            assert model is None
        dist.barrier()
    dist.destroy_process_group()
    if failures:
        sys.exit(1)


if __name__ == "__main__":
    main()
