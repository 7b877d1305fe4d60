"""Upload throughput of the input layouts, host bytes per second, next to the oracle's
restatement of the reference's conversion (densify every row, store row-major FP64) on the host cores.
usage: python tools/ingest_bench.py [rows] [cols]"""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import oap_mllib_b200 as m
from oracle import kmeans_oracle as ko

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 256
rng = np.random.default_rng(0)


def pinned(shape, dtype):
    p = C.c_void_p()
    nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
    m._lib.check(m.lib().kmb200_pinned_alloc(nbytes, C.byref(p)))
    ctype = {np.float32: C.c_float, np.float64: C.c_double}[dtype]
    return p, np.ctypeslib.as_array(C.cast(p, C.POINTER(ctype)), shape=shape)


def timed(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


rows32 = rng.standard_normal((n, d), dtype=np.float32)
table = m.Table(n, d)
lines = []


def report(name, nbytes, sec):
    lines.append(f"| {name} | {nbytes / 1e9:.2f} | {sec * 1e3:.0f} | {nbytes / sec / 1e9:.1f} |")


p32, h32 = pinned((n, d), np.float32)
h32[:] = rows32
report("row-major FP32, pinned (`kmb200_pinned_alloc`) -> `append_f32`", h32.nbytes, timed(lambda: table.append(0, h32)))
report("row-major FP32, pageable -> `append_f32`", rows32.nbytes, timed(lambda: table.append(0, rows32)))
m._lib.check(m.lib().kmb200_pinned_free(p32))
p64, h64 = pinned((n // 2, d), np.float64)
h64[:] = rows32[: n // 2]
report("row-major FP64, pinned -> `append_f64` (staged, rounded on the device)", h64.nbytes, timed(lambda: table.append(0, h64)))
cols = tuple(np.ascontiguousarray(h64[:, c]) for c in range(d))
report("columnar FP64 (one buffer per column), pageable -> `append_columns_f64`", h64.nbytes,
       timed(lambda: table.append_columns(0, cols)))
np.testing.assert_array_equal(table.read_rows(0, 1000), h64[:1000].astype(np.float32))
try:
    import scipy.sparse as sp
    dense = h64[: n // 4].copy()
    dense[np.abs(dense) < 1.6449] = 0.0          # ~10 % nonzeros
    csr = sp.csr_matrix(dense)
    nb = csr.data.nbytes + csr.indices.nbytes + csr.indptr.nbytes
    sec = timed(lambda: table.append_csr(0, csr.indptr, csr.indices, csr.data))
    lines.append(f"| CSR FP64, {100.0 * csr.nnz / dense.size:.0f} % nonzeros -> `append_csr_f64` | {nb / 1e9:.2f} | {sec * 1e3:.0f} | "
                 f"{nb / sec / 1e9:.1f} ({dense.nbytes / sec / 1e9:.1f} as dense FP64) |")
    np.testing.assert_array_equal(table.read_rows(0, 1000), dense[:1000].astype(np.float32))
    sample = csr[: 200_000]
    sec = timed(lambda: ko.dense_rows(sample), reps=1)
    lines.append(f"| host restatement of the reference (`row.toArray` + row-major FP64 copy), CSR sample of 200 000 rows | "
                 f"{200_000 * d * 8 / 1e9:.2f} | {sec * 1e3:.0f} | {200_000 * d * 8 / sec / 1e9:.2f} (dense FP64 produced) |")
except ImportError:
    pass
sample = h64[: 1_000_000]
sec = timed(lambda: np.ascontiguousarray(sample, dtype=np.float64).copy(), reps=1)
lines.append(f"| host restatement of the reference (memcpy of dense rows into the table, 1 thread), 1 000 000 rows | "
             f"{sample.nbytes / 1e9:.2f} | {sec * 1e3:.0f} | {sample.nbytes / sec / 1e9:.1f} |")
m._lib.check(m.lib().kmb200_pinned_free(p64))
table.release()
print(f"# Ingest throughput ({n} x {d}; best of 3; includes ||x||^2 and the final stream sync)\n")
print("| path | host GB | ms | GB/s |\n|---|---:|---:|---:|")
print("\n".join(lines))
