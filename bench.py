#!/usr/bin/env python
"""bench.py -- K-means Lloyd iterations/s on B200 (BASELINE.json's metric) for libkmeans_b200.so.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one Lloyd iteration (assign + accumulate + all-reduce + finalize) over the whole
workload.  Default workload = BASELINE.json's headline config C3: synthetic Gaussian blobs
100M x 256 FP32, k = 1000 (102.4 GB: fits one B200).  With N GPUs the SAME 100M rows are sharded
N ways (strong scaling, one NCCL all-reduce of [k*d sums | k counts | cost] per iteration).

Prints ONE JSON line (rank 0).  `value` = iterations/s with the rows already resident in HBM,
timed on the device (CUDA events on the library's stream, max over ranks).  `e2e` = the same
metric through the reference-facing call KMeansDALImpl.train(host rows): upload of this rank's
rows from PINNED host memory + K iterations + read-back of centres/cost, all inside the timed
region.  `cpu_baseline` / `--impl reference` = the fastest of the CPU stand-ins for the reference's
oneDAL step (oracle/cpu_baselines.py: scikit-learn's Lloyd in FP64 and FP32, a blocked dgemm +
fused argmin / segmented-sum loop over a thread pool, the FP64 oracle itself) on all host cores,
each timed for one iteration on a bounded row sample and extrapolated linearly in N -- the
reference's own oneDAL build cannot run in this image.  The initial centres are rows
floor(j N / k) of the WHOLE table, so every GPU count solves the same problem and `final_cost`
must agree across N.  `--workload C2|C4|C5` runs BASELINE.json's other configurations; `--workload PCA` the PCA
path over C3's table (bench_pca.py: its own metric, same contract).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time


def _reference_arm_requested() -> bool:
    a = sys.argv[1:]
    return any(x == "--impl=reference" for x in a) or any(x == "--impl" and i + 1 < len(a) and a[i + 1] == "reference"
                                                          for i, x in enumerate(a))


if _reference_arm_requested():
    # torch.distributed.run exports OMP_NUM_THREADS=1 to its workers when nproc > 1; the CPU arm needs every host
    # core (set before NumPy / OpenBLAS / OpenMP load)
    _n = str(len(os.sched_getaffinity(0)))
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = _n

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (rows, d, k, description)
    "C2": (10_000_000, 64, 100, "C2 synthetic dense 10Mx64 fp32, k=100"),
    "C3": (100_000_000, 256, 1000, "C3 synthetic dense 100Mx256 fp32, k=1000 (headline)"),
    "C4": (20_000_000, 512, 10_000, "C4 synthetic dense 20Mx512 fp32, k=10000"),
    "C5": (400_000_000, 32, 16, "C5 synthetic dense 400Mx32 fp32, k=16"),
}
METRIC = "kmeans_lloyd_iterations_per_sec"
UNIT = "iter/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"],
                "bf16_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------- CPU arm
def make_config(args, world):
    """The `config` object, identical in both arms (the GPU arm and --impl reference)."""
    rows, d, k, desc = WORKLOADS[args.workload]
    if args.rows:
        rows = args.rows
    return {"workload": desc, "rows_total": rows, "rows_per_gpu": rows // world, "d": d, "k": k,
            "parallelism": f"rows sharded over {world} GPU(s), 1 NCCL all-reduce / iteration",
            "l2": "inputs (rows_per_gpu*d*4 bytes) far larger than the 126 MB L2; no flush needed",
            "init": "rows floor(j*N/k) of the whole table (the same centres for every GPU count)",
            "tol": "disabled (exactly K iterations)",
            "data": f"Gaussian blobs, sigma = 1, k_true = k centres ~ U(-{args.centre_range:g}, {args.centre_range:g})^d"}


def cpu_baseline(d, k, rows_full, sample_rows=0, which=None):
    """Lloyd iterations/s of the CPU stand-ins for the reference's oneDAL step (oracle/cpu_baselines.py) on a
    bounded row sample, extrapolated linearly in N to the full workload; `value` is the FASTEST of them."""
    from oracle import cpu_baselines as cb
    if not sample_rows:
        sample_rows = int(0.8e12 / (2.0 * d * k))                    # ~0.8 TFLOP of distance work per iteration
        sample_rows = min(sample_rows, int(4e9 / (12.0 * d)))        # and <= 4 GB of host rows (FP32 + FP64 copies)
    sample_rows = max(20_000, min(sample_rows, rows_full))
    kw = {"which": which} if which else {}
    secs, cores = cb.run_all(d, k, sample_rows, **kw)
    flops = 2.0 * sample_rows * d * k
    cand = {}
    for name, t in secs.items():
        if name.endswith("_error"):
            cand[name] = t
        elif t:
            cand[name] = {"seconds_per_sample_iteration": t, "iter_per_s_full_workload": (sample_rows / rows_full) / t,
                          "gflops": flops / t / 1e9, "gbytes_per_s": sample_rows * d * (4 if name.endswith("f32") else 8) / t / 1e9}
    timed = {n: c for n, c in cand.items() if isinstance(c, dict)}
    best = min(timed, key=lambda n: timed[n]["seconds_per_sample_iteration"])
    return {"value": timed[best]["iter_per_s_full_workload"], "unit": UNIT, "cores": cores, "kind": "port",
            "which": best, "gflops": timed[best]["gflops"],
            "sample": f"{sample_rows} of {rows_full} rows x {d}, k={k}: one Lloyd iteration per stand-in, extrapolated "
                      f"linearly in N; fastest = {best} ({timed[best]['seconds_per_sample_iteration']:.2f} s per sample "
                      f"iteration on {cores} cores). oneDAL itself cannot run in this image (no oneDAL/oneCCL/JVM)",
            "candidates": cand}


def run_reference(args, rank, world):
    if rank != 0:
        return
    rows, d, k, desc = WORKLOADS[args.workload]
    if args.rows:
        rows = args.rows
    # bounded whatever --steps says: one sample iteration per stand-in (sklearn: 1 + 3), a few minutes at most
    res = cpu_baseline(d, k, rows, args.cpu_sample_rows, which=("sklearn_f64", "sklearn_f32", "blocked_f64"))
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 / res["value"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32" if res["which"].endswith("f32") else "f64", "data": "synthetic",
            "config": make_config(args, world),
            "cpu_baseline": res,
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "oneDAL/oneCCL/JVM are absent from this image: the reference's own build cannot run; this is "
                    "the fastest CPU stand-in for its algorithm on the box's host cores (each timed for a bounded "
                    "number of iterations on a row sample, whatever --steps says)"}
    emit(line)


# ------------------------------------------------------------------------------------- GPU arm
def run_ours(args, rank, local_rank, world):
    import ctypes as C

    import torch

    import oap_mllib_b200 as m
    from oap_mllib_b200 import _lib

    rows, d, k, desc = WORKLOADS[args.workload]
    if args.rows:
        rows = args.rows
    torch.cuda.set_device(local_rank)
    _lib.check(m.lib().kmb200_set_device(local_rank))
    for opt in args.option:
        name, val = opt.split("=")
        _lib.check(m.lib().kmb200_set_option(name.encode(), float(val)))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- shard the rows (strong scaling: the same total N on every GPU count) -----------------
    from oap_mllib_b200.sharding import shard_bounds
    row0, row1 = shard_bounds(rows, world)[rank]
    n_local = row1 - row0
    rng = np.random.default_rng(42)
    centres = rng.uniform(-args.centre_range, args.centre_range, (k, d)).astype(np.float32)
    table = m.Table(n_local, d)
    table.fill_blobs(centres, seed=42, row_offset=row0)

    # ---- communicator: unique id from rank 0 (the "Spark driver"), one executor per GPU ---------
    comm = None
    if world > 1:
        ident = [m.OneCCL.getUniqueId() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        m.OneCCL.init(world, rank, ident[0])
        comm = m.OneCCL.comm()

    # ---- initial centres: rows floor(j*N/k) of the WHOLE table (each rank reads the ones it holds, the rest
    # arrive by all-reduce), so every GPU count solves the same problem and final_cost must agree across N
    init = np.zeros((k, d), np.float64)
    for j, gi in enumerate(((np.arange(k, dtype=np.int64) * rows) // k).tolist()):
        if row0 <= gi < row1:
            init[j] = table.read_rows(gi - row0, 1)[0]
    if dist is not None:
        t = torch.from_numpy(init).cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        init = t.cpu().numpy()
    init = np.ascontiguousarray(init)

    def fit(tbl, centers, iters):
        out = np.empty((k, d), np.float64)
        cost, it = C.c_double(), C.c_int32()
        _lib.check(m.lib().kmb200_fit(tbl.handle, comm, centers.ctypes.data_as(C.c_void_p), k, d, -1.0, iters,
                                      out.ctypes.data_as(C.c_void_p), C.byref(cost), C.byref(it)))
        return out, cost.value, it.value

    # ---- device-resident measurement ------------------------------------------------------------
    cur, _, _ = fit(table, init, args.warmup)          # W untimed warm-up iterations
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    cur2, cost, iters = fit(table, cur, args.steps)    # EXACTLY K timed iterations (tol < 0)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop()
    prof = _lib.profile()
    assert iters == args.steps
    if any(o.startswith("role_prof") for o in args.option):
        _lib.check(m.lib().kmb200_set_option(b"role_prof", 0.0))   # prints the MMA-warp cycle counters
    dev_ms = max_over_ranks(prof["total_ms"])
    wall_ms = max_over_ranks(wall * 1e3)
    ms_per_step = dev_ms / args.steps
    value = 1000.0 / ms_per_step
    assign_ms = max_over_ranks(prof["assign_ms"]) / args.steps
    # the dominant kernel alone (CUDA events around its launches on the library's stream): with the two-level
    # scheme the assign phase also holds the candidate gather and the short second pass over the near-tie rows
    main_ms = max_over_ranks(prof.get("assign_main_ms", 0.0)) / args.steps
    if main_ms <= 0.0:
        main_ms = assign_ms
    launches = prof["kernel_launches"]

    peaks = load_peaks()
    variant = prof.get("assign_variant", -1)
    acc_ms = max_over_ranks(prof["accumulate_ms"]) / args.steps
    flops_per_launch = 2.0 * n_local * d * k           # ALGORITHMIC flops of one pass over this GPU's rows
    bytes_per_launch = 4.0 * n_local * d               # ALGORITHMIC bytes of one pass
    tf32_sus, tf32_burst = 0.5 * peaks["bf16_sustained"], 0.5 * peaks["bf16_burst"]   # TF32 = half the bf16 rate
    # which roofline bounds this workload: the slower of 2ndk at the TF32 rate and 4nd at HBM bandwidth
    bound = "tensor" if flops_per_launch / (tf32_sus * 1e12) >= bytes_per_launch / (peaks["hbm_gbs"] * 1e9) else "hbm"
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tr = json.load(f)
        key = f"{args.workload}_n{world}"
        if key in tr and tr[key].get("rows") == n_local:
            traffic = tr[key].get("dram_bytes_per_launch", tr[key].get("assign_dram_bytes_per_launch"))
    names = {0: "fused_smallk_kernel (one read of X: distances, arg-min, FP64 sums and cost on the CUDA cores)",
             1: "assign_pair_h16_kernel<BN, 2, 1> (coarse pass of the cascade: tcgen05 cta_group::2, kind::f16, one product)",
             2: "assign_pair_h16_kernel<BN, 2, 3> (tcgen05 cta_group::2, kind::f16, three products, fast bookkeeping)",
             3: "assign_pair_h16_kernel<BN, 0, 3> (tcgen05 cta_group::2, kind::f16, three products)",
             4: "assign_pair_ss_kernel (tcgen05 cta_group::2, TF32 + BF16 corrections)",
             5: "assign_tc2_kernel (tcgen05, 3xTF32, resident X)", 6: "assign_tc_kernel (tcgen05, 3xTF32, streaming)",
             7: "recheck_kernel (FP64, CUDA cores)",
             8: "assign_pair_h16_kernel<BN, 2, 1>, 256 < d <= 512 (coarse pass of the cascade: tcgen05 cta_group::2, kind::f16, one "
                "product; the undecided rows go through the streaming 3xTF32 kernel, inside assign_phase_ms)"}
    if variant != 0 and acc_ms > main_ms:
        kern, kern_ms = "segment_sums_kernel (per-cluster FP64 sums and cost; second read of X)", acc_ms
    else:
        kern, kern_ms = names.get(variant, "assign kernel"), main_ms
    if bound == "hbm":
        achieved, peak, unit = bytes_per_launch / (kern_ms * 1e-3) / 1e9, peaks["hbm_gbs"], "GB/s"
        peak_source = f"hbm_gbs of {peaks['source']} MEASURED_PEAKS.json"
    else:
        achieved, unit = flops_per_launch / (kern_ms * 1e-3) / 1e12, "TFLOP/s"
        if variant in (1, 8):
            # one FP16 product per score: the ceiling is the dense FP16/BF16 rate itself
            peak, peak_source = peaks["bf16_sustained"], (f"bf16_tflops_sustained of {peaks['source']} MEASURED_PEAKS.json "
                                                          "(FP16 and BF16 MMAs run at the same rate); kernel timed inside a long step")
        else:
            peak, peak_source = tf32_sus, (f"0.5 x bf16_tflops_sustained of {peaks['source']} MEASURED_PEAKS.json "
                                           "(TF32 = half the bf16 rate); kernel timed inside a long step")
    step_tflops_gpu = flops_per_launch / (ms_per_step * 1e-3) / 1e12        # per GPU, whole step, algorithmic
    step_gbs_gpu = bytes_per_launch / (ms_per_step * 1e-3) / 1e9
    roofline = {"bound": bound, "kernel": kern, "achieved": achieved, "peak": peak, "unit": unit,
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_source,
                "algorithmic_flops_per_launch": flops_per_launch, "algorithmic_bytes_per_launch": bytes_per_launch,
                "kernel_ms_per_launch": kern_ms, "main_assign_kernel_ms_per_launch": main_ms,
                "assign_phase_ms": assign_ms,
                # the whole step (every kernel, the all-reduce, the host round trips), per GPU, algorithmic work only
                "step_tflops_per_gpu": step_tflops_gpu,
                "step_frac_of_fp16_sustained": step_tflops_gpu / peaks["bf16_sustained"],
                "step_frac_of_tf32_proxy_sustained": step_tflops_gpu / tf32_sus,
                "step_frac_of_tf32_proxy_burst": step_tflops_gpu / tf32_burst,
                "step_gbytes_per_s_per_gpu": step_gbs_gpu,
                "step_frac_of_hbm": step_gbs_gpu / peaks["hbm_gbs"]}
    if variant in (1, 8):
        roofline["coarse_undecided_fraction"] = prof["coarse_undecided_rows"] / args.steps / n_local
    t32 = tf32_measured(achieved)
    if t32:
        roofline["tf32_cublas_sustained_measured"] = t32["tf32_cublas_sustained_measured"]
    phases = {kk: max_over_ranks(prof[kk]) / args.steps for kk in
              ("assign_ms", "recheck_ms", "bucket_ms", "accumulate_ms", "allreduce_ms", "finalize_ms")}

    # ---- end to end: host rows (pinned) -> upload -> K iterations -> read back ---------------------
    e2e = None
    if not args.no_e2e:
        try:
            import psutil
            avail = psutil.virtual_memory().available
        except Exception:
            avail = 0
        need = n_local * d * 4
        e2e_rows = n_local
        if args.e2e_rows:
            e2e_rows = min(n_local, args.e2e_rows)
        elif avail < 1.6 * need * min(world, 8):
            e2e_rows = int(max(1, 0.5 * avail / (d * 4) / min(world, 8)))
        host = torch.empty((e2e_rows, d), dtype=torch.float32, pin_memory=True)
        hnp = host.numpy()
        for r0 in range(0, e2e_rows, 4_000_000):          # synthetic rows come from the device generator
            nr = min(4_000_000, e2e_rows - r0)
            _lib.check(m.lib().kmb200_table_read_rows(table.handle, r0, nr, hnp[r0:r0 + nr].ctypes.data_as(C.c_void_p)))
        table.release()
        barrier()
        t0 = time.perf_counter()
        t2 = m.Table(e2e_rows, d)                           # inside the timed region: allocation too
        for r0 in range(0, e2e_rows, 8_000_000):
            t2.append(r0, hnp[r0:r0 + 8_000_000])
        out, cost_e, it_e = fit(t2, init, args.steps)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        barrier()
        e2e_s = max_over_ranks(e2e_s)
        t2.release()
        scale = n_local / e2e_rows                          # 1.0 when the full shard fits in host RAM
        e2e = {"value": args.steps / (e2e_s * scale), "unit": UNIT,
               "h2d_bytes_per_step": (e2e_rows * d * 4 + k * d * 8) / args.steps,
               "d2h_bytes_per_step": (k * d * 8 + 16) / args.steps + 24,
               "seconds": e2e_s, "rows_per_rank": e2e_rows,
               "what": "KMeansDALImpl-style fit through the C ABI: table_create + table_append_f32 from pinned host "
                       "rows + kmb200_fit(K iterations) + centres/cost read-back; upload happens once per fit"
                       + ("" if scale == 1.0 else f"; host RAM held only {e2e_rows} rows per rank, time scaled x{scale:.2f}")}
        del host
    else:
        table.release()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(d, k, rows, args.cpu_sample_rows)

    if world > 1:
        m.OneCCL.cleanup()
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None,
            "dtype": ("f32 distances on the CUDA cores, f64 near-tie re-check / sums / cost" if variant == 0 else
                      "f16 two-term split products (fp32 accumulate), f64 re-check / sums / cost"), "data": "synthetic",
            "config": make_config(args, world),
            "tflops_algorithmic": 2.0 * rows * d * k / (ms_per_step * 1e-3) / 1e12,
            "wall_ms_per_step": wall_ms / args.steps, "phases_ms_per_step": phases,
            "rechecked_rows_per_step": prof["rechecked_rows"] / args.steps,
            "full_rechecks_per_step": prof["full_rechecks"] / args.steps,
            "coarse_undecided_rows_per_step": prof.get("coarse_undecided_rows", 0) / args.steps,
            "assign_variant": variant, "cascade_dropped_at": prof.get("cascade_dropped_at", -1),
            "empty_cluster_events": prof.get("empty_cluster_events", 0),
            "final_cost": cost, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "roofline": roofline, "cpu_baseline": cpu}
    emit(line)


_JSON_FD = None


def tf32_measured(achieved):
    """cuBLAS TF32 on this pool's B200, measured by tools/measure_tf32_peak.py with MEASURED_PEAKS.json's protocol
    (the driver's file holds bf16 only, so `peak` above stays the 0.5 x bf16 proxy; this is the direct figure)."""
    path = os.path.join(ROOT, "profiles", "tf32_peak.json")
    if not os.path.exists(path):
        return {}
    with open(path) as f:
        t = json.load(f)
    return {"tf32_cublas_sustained_measured": t["tf32_tflops_sustained"],
            "frac_of_measured_tf32_sustained": achieved / t["tf32_tflops_sustained"]}


def emit(line):
    """The ONE JSON line of the contract, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # Libraries loaded below print to stdout on their own (NCCL writes its version banner there): keep the real
    # stdout for the JSON line and send everything else to stderr.
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS) + ["PCA"])
    ap.add_argument("--rows", type=int, default=0, help="override the total row count (debugging)")
    ap.add_argument("--pca-d", type=int, default=0, help="--workload PCA: columns of the table (default 256; other shapes are not the bench line)")
    ap.add_argument("--centre-range", type=float, default=10.0,
                    help="blob centres ~ U(-R, R)^d with unit noise: 10 = BASELINE's well-separated blobs, 1 = heavy overlap")
    ap.add_argument("--e2e-rows", type=int, default=0)
    ap.add_argument("--cpu-sample-rows", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--option", action="append", default=[], help="library option name=value (experiments)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.workload == "PCA":   # the PCA path under the same contract: bench_pca.py
        import bench_pca
        me = sys.modules[__name__]
        if args.impl == "reference":
            bench_pca.run_reference(args, rank, world, me)
        else:
            bench_pca.run_ours(args, rank, local_rank, world, me)
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
