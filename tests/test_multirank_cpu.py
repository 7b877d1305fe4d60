"""World-size-2 tests of the N > 1 host logic on CPU (gloo): rendezvous of the 128-byte unique id,
row sharding, and the one-all-reduce-per-iteration protocol -- every rank reduces the packed
[sums | counts | cost] payload and finalizes identically, which must equal the single-rank oracle."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oap_mllib_b200 import sharding
    from oracle import kmeans_oracle as ko
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. the unique id is created once (rank 0 = the Spark driver) and reaches every rank
        ident = [bytes(range(128)) if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        assert ident[0] == bytes(range(128))
        # 2. shard + iterate with ONE all-reduce per iteration
        n, d, k = 6001, 20, 7
        X, _, _ = ko.gaussian_blobs(n, d, k, seed=5)
        X = X.astype(np.float64)
        r0, r1 = sharding.shard_bounds(n, world)[rank]
        C = ko.row_init_centers(X, k)
        iters, cost, converged = 0, 0.0, False
        while iters < 20 and not converged:
            part = ko.partial_result(X[r0:r1], C)
            buf = torch.from_numpy(sharding.pack_partial(part["sums"], part["counts"], part["objective"]))
            dist.all_reduce(buf)
            sums, counts, cost = sharding.unpack_partial(buf.numpy(), k, d)
            assert counts.min() > 0
            new = sums * (1.0 / counts[:, None])
            converged = ko.centers_converged(C, new, 1e-4)   # same flag on every rank, no broadcast
            C = new
            iters += 1
        q.put((rank, iters, cost, C))
    finally:
        dist.destroy_process_group()


def test_sharding_is_a_partition():
    from oap_mllib_b200 import sharding
    for rows, world in [(0, 1), (1, 2), (7, 3), (100_000_000, 8), (12345, 8)]:
        b = sharding.shard_bounds(rows, world)
        assert b[0][0] == 0 and b[-1][1] == rows
        assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
        sizes = [e - s for s, e in b]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(10, 0)


def test_payload_roundtrip():
    from oap_mllib_b200 import sharding
    rng = np.random.default_rng(0)
    sums, counts = rng.standard_normal((5, 33)), np.array([3, 0, 7, 2 ** 40, 1])
    buf = sharding.pack_partial(sums, counts, 12.5)
    assert buf.size == sharding.payload_size(5, 33) == 5 * 64 + 5 + 1
    s2, c2, cost = sharding.unpack_partial(buf, 5, 33)
    np.testing.assert_array_equal(s2, sums)
    np.testing.assert_array_equal(c2, counts)
    assert cost == 12.5


def test_two_ranks_gloo_equal_single_rank_oracle():
    import torch.multiprocessing as mp
    from oracle import kmeans_oracle as ko
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    X, _, _ = ko.gaussian_blobs(6001, 20, 7, seed=5)
    want_c, want_cost, want_it = ko.lloyd(X.astype(np.float64), ko.row_init_centers(X.astype(np.float64), 7), 1e-4, 20)
    for rank, iters, cost, C in out:
        assert iters == want_it
        np.testing.assert_allclose(C, want_c, rtol=1e-12)
        assert abs(cost - want_cost) <= 1e-10 * want_cost
    np.testing.assert_array_equal(out[0][3], out[1][3])          # bit-identical centres on both ranks


def _pca_worker(rank, world, port, q):
    """The PCA protocol of csrc/pca_host.inl on two gloo ranks: all-reduce of [column sums | row count] (sum) and of the
    column maxima (max), a shared FP32-rounded mean, the local Gram matrix of the rows centred with it, one all-reduce of
    d x d doubles, and the exact correction  G - N ybar ybar^T  with ybar = sum / N - mean_f32."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oap_mllib_b200 import sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, d = 5003, 12
        rng = np.random.default_rng(8)
        X = (rng.standard_normal((n, d)) * np.linspace(5, 0.2, d) + rng.uniform(-30, 30, d)).astype(np.float32)
        r0, r1 = sharding.shard_bounds(n, world)[rank]
        mine = X[r0:r1].astype(np.float64)
        stats = torch.from_numpy(np.concatenate([mine.sum(axis=0), [float(r1 - r0)]]))
        dist.all_reduce(stats)
        mx = torch.from_numpy(np.abs(mine).max(axis=0))
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        s1, ntot = stats.numpy()[:d], stats.numpy()[d]
        mean_f = (s1 / ntot).astype(np.float32).astype(np.float64)      # what the kernel subtracts
        y = mine - mean_f
        g = torch.from_numpy(y.T @ y)
        dist.all_reduce(g)
        ybar = s1 / ntot - mean_f
        q.put((rank, g.numpy() - ntot * np.outer(ybar, ybar), int(ntot), mx.numpy()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_pca_partials_equal_the_oracle():
    import torch.multiprocessing as mp
    from oracle import pca_oracle as po
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_pca_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n, d = 5003, 12
    rng = np.random.default_rng(8)
    X = (rng.standard_normal((n, d)) * np.linspace(5, 0.2, d) + rng.uniform(-30, 30, d)).astype(np.float32)
    g_ref, _, n_ref = po.centred_gram([X[:2501], X[2501:]], True)
    for rank, g, ntot, mx in out:
        assert ntot == n_ref == n
        np.testing.assert_allclose(g, g_ref, rtol=1e-10, atol=1e-9 * np.abs(g_ref).max())
        np.testing.assert_array_equal(mx, np.abs(X.astype(np.float64)).max(axis=0))
    np.testing.assert_array_equal(out[0][1], out[1][1])      # the same matrix on both ranks
    ev, axes = po.eig_desc(out[0][1])
    ev_ref, axes_ref = po.pca_svd_dense([X[:2501], X[2501:]])
    np.testing.assert_allclose(ev / (n - 1), ev_ref, rtol=1e-9)
    np.testing.assert_allclose(axes[:4], axes_ref[:4], atol=1e-8)
