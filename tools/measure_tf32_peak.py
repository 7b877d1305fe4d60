"""Measures the dense TF32 GEMM rate of this GPU with the protocol of MEASURED_PEAKS.json (torch.matmul 8192^3,
best of 10 = burst, back to back for 4 s = sustained), for the BASELINE.md row the builder was asked to fill in."""
import json
import time

import torch

torch.backends.cuda.matmul.allow_tf32 = True
n = 8192
a = torch.randn(n, n, device="cuda")
b = torch.randn(n, n, device="cuda")
for _ in range(3):
    a @ b
torch.cuda.synchronize()
best = 0.0
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    a @ b
    e1.record()
    torch.cuda.synchronize()
    best = max(best, 2.0 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 0
t0 = time.perf_counter()
e0.record()
while time.perf_counter() - t0 < 4.0:
    for _ in range(20):
        a @ b
    reps += 20
    torch.cuda.synchronize()
e1.record()
torch.cuda.synchronize()
sustained = reps * 2.0 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
# the same for bf16, in the same process, so that the ratio is taken under identical conditions
a16, b16 = a.bfloat16(), b.bfloat16()
for _ in range(3):
    a16 @ b16
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps16 = 0
t0 = time.perf_counter()
e0.record()
while time.perf_counter() - t0 < 4.0:
    for _ in range(20):
        a16 @ b16
    reps16 += 20
    torch.cuda.synchronize()
e1.record()
torch.cuda.synchronize()
sustained16 = reps16 * 2.0 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
print(json.dumps({"tf32_tflops_burst": best, "tf32_tflops_sustained": sustained, "bf16_tflops_sustained_same_run": sustained16,
                  "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), best of 10 / back to back for 4 s"}))
