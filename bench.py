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
region.  `cpu_baseline` / `--impl reference` = the FP64 CPU restatement of the reference's oneDAL
path (oracle/kmeans_oracle.py, OpenBLAS dgemm on all host cores) on a bounded row sample -- the
reference's own oneDAL build cannot run in this image.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

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
def cpu_lloyd_steps(d, k, rows_full, steps, warmup, sample_rows):
    """Times `steps` Lloyd iterations of the FP64 oracle on a `sample_rows` sample of the workload
    and returns iterations/s extrapolated linearly in N to the full workload."""
    from oracle import kmeans_oracle as ko
    try:
        from threadpoolctl import threadpool_info
        threads = max([p.get("num_threads", 1) for p in threadpool_info()] + [1])
    except Exception:
        threads = os.cpu_count() or 1
    cores = len(os.sched_getaffinity(0))
    rng = np.random.default_rng(42)
    centres = rng.uniform(-10, 10, (k, d)).astype(np.float32)
    lab = rng.integers(0, k, sample_rows)
    X = centres[lab]
    X += rng.standard_normal((sample_rows, d), dtype=np.float32)
    X = X.astype(np.float64)          # the reference's path is FP64 end to end (KMeansDALImpl.cpp:32)
    C = ko.row_init_centers(X, k)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        C, _, _, _ = ko.lloyd_step([X], C)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    t_step = float(np.mean(times))
    value = (sample_rows / rows_full) / t_step
    return {"value": value, "unit": UNIT, "cores": min(cores, threads) if threads else cores, "kind": "port",
            "sample": f"{sample_rows} of {rows_full} rows x {d}, k={k}, FP64 numpy/OpenBLAS restatement of the "
                      f"oneDAL step (oracle/kmeans_oracle.py), {steps} iteration(s) of {t_step:.2f} s, "
                      f"extrapolated linearly in N",
            "seconds_per_sample_iteration": t_step}


def run_reference(args, rank, world):
    if rank != 0:
        return
    rows, d, k, desc = WORKLOADS[args.workload]
    if args.rows:
        rows = args.rows
    sample = args.cpu_sample_rows or max(50_000, int(2.0e12 / (2.0 * d * k)))   # ~2 TFLOP of dgemm per step
    sample = min(sample, rows)
    res = cpu_lloyd_steps(d, k, rows, max(1, args.steps), min(args.warmup, 1), sample)
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 / res["value"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "rows": rows, "d": d, "k": k},
            "cpu_baseline": res,
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "oneDAL/oneCCL/JVM are absent from this image: the reference's own build cannot run; this is "
                    "the CPU restatement of its algorithm on the box's host cores"}
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
    centres = rng.uniform(-10, 10, (k, d)).astype(np.float32)
    table = m.Table(n_local, d)
    table.fill_blobs(centres, seed=42, row_offset=row0)

    # ---- communicator: unique id from rank 0 (the "Spark driver"), one executor per GPU ---------
    comm = None
    if world > 1:
        ident = [m.OneCCL.getUniqueId() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        m.OneCCL.init(world, rank, ident[0])
        comm = m.OneCCL.comm()

    # ---- initial centres: rows floor(j*n/k) of rank 0's shard, broadcast ------------------------
    if rank == 0:
        idx = (np.arange(k, dtype=np.int64) * n_local) // k
        init = np.vstack([table.read_rows(int(i), 1) for i in idx]).astype(np.float64)
    else:
        init = np.empty((k, d), np.float64)
    if dist is not None:
        t = torch.from_numpy(init).cuda()
        dist.broadcast(t, src=0)
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
    flops_per_launch = 2.0 * n_local * d * k           # ALGORITHMIC flops of one assign launch (one pass)
    tf32_peak = 0.5 * peaks["bf16_sustained"]          # TF32 runs at half the bf16 tensor rate
    achieved = flops_per_launch / (main_ms * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tr = json.load(f)
        key = f"{args.workload}_n{world}"
        if key in tr and tr[key].get("rows") == n_local:
            traffic = tr[key]["assign_dram_bytes_per_launch"]
    h16 = d <= 256 and k > 16
    cascade = h16 and prof.get("coarse_undecided_rows", 0) > 0
    phase_achieved = flops_per_launch / (assign_ms * 1e-3) / 1e12
    if cascade:
        # The dominant kernel is the coarse pass: ONE FP16 product per score over all rows (the rows it cannot
        # decide -- coarse_undecided_rows -- go through the three-product kernel afterwards, inside assign_phase_ms).
        # Its hardware ceiling is the dense FP16/BF16 rate, i.e. MEASURED_PEAKS.json's own figure (no TF32 proxy).
        peak = peaks["bf16_sustained"]
        roofline = {"bound": "tensor",
                    "kernel": "assign_pair_h16_kernel<128, 2, 1> (coarse pass of the cascade: tcgen05 cta_group::2, kind::f16, one product)",
                    "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                    "peak_source": f"bf16_tflops_sustained of {peaks['source']} MEASURED_PEAKS.json (FP16 and BF16 MMAs run at the "
                                   "same rate); kernel timed inside a long step",
                    "algorithmic_flops_per_launch": flops_per_launch,
                    "main_kernel_ms_per_launch": main_ms,
                    "assign_phase_ms": assign_ms,
                    "assign_phase_achieved": phase_achieved,
                    "assign_phase_frac_of_tf32_proxy": phase_achieved / tf32_peak,
                    "coarse_undecided_fraction": prof["coarse_undecided_rows"] / args.steps / n_local,
                    "hbm_frac_of_step": (n_local * d * 4.0 / (ms_per_step * 1e-3) / 1e9) / peaks["hbm_gbs"]}
        t32 = tf32_measured(phase_achieved)
        if t32:
            roofline["tf32_cublas_sustained_measured"] = t32["tf32_cublas_sustained_measured"]
            roofline["assign_phase_frac_of_measured_tf32_sustained"] = t32["frac_of_measured_tf32_sustained"]
    else:
        roofline = {"bound": "tensor",
                    "kernel": "assign_pair_h16_kernel<128, 2, 3> (tcgen05 cta_group::2, kind::f16, three products)" if h16 else "assign_tc_kernel (3xTF32)",
                    "achieved": achieved, "peak": tf32_peak,
                    "unit": "TFLOP/s", "frac": achieved / tf32_peak, "traffic": traffic,
                    "peak_source": f"0.5 x bf16_tflops_sustained of {peaks['source']} MEASURED_PEAKS.json "
                                   "(TF32 = half the bf16 rate); kernel timed inside a long step",
                    "algorithmic_flops_per_launch": flops_per_launch,
                    "tensor_time_in_tf32_passes": 1.5 if h16 else 3.0,   # three FP16 products = 1.5 TF32 passes of tensor time; 3xTF32 on the fallback kernel
                    "main_kernel_ms_per_launch": main_ms,
                    **tf32_measured(achieved),
                    "assign_phase_ms": assign_ms,
                    "hbm_frac_of_step": (n_local * d * 4.0 / (ms_per_step * 1e-3) / 1e9) / peaks["hbm_gbs"]}
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
        sample = args.cpu_sample_rows or max(50_000, int(2.0e12 / (2.0 * d * k)))
        cpu = cpu_lloyd_steps(d, k, rows, 1, 0, min(sample, rows))

    if world > 1:
        m.OneCCL.cleanup()
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16 two-term split products (fp32 accumulate), f64 re-check / sums / cost", "data": "synthetic",
            "config": {"workload": desc, "rows_total": rows, "rows_per_gpu": n_local, "d": d, "k": k,
                       "parallelism": f"rows sharded over {world} GPU(s), 1 NCCL all-reduce / iteration",
                       "l2": "inputs (rows_per_gpu*d*4 bytes) far larger than the 126 MB L2; no flush needed",
                       "init": "rows floor(j*n/k) of rank 0", "tol": "disabled (exactly K iterations)"},
            "tflops_algorithmic": 2.0 * rows * d * k / (ms_per_step * 1e-3) / 1e12,
            "wall_ms_per_step": wall_ms / args.steps, "phases_ms_per_step": phases,
            "rechecked_rows_per_step": prof["rechecked_rows"] / args.steps,
            "full_rechecks_per_step": prof["full_rechecks"] / args.steps,
            "coarse_undecided_rows_per_step": prof.get("coarse_undecided_rows", 0) / args.steps,
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
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override the total row count (debugging)")
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
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
