/* kmb200.h -- C ABI of libkmeans_b200.so: a B200-native (sm_100a) K-means Lloyd iteration that
 * drops in behind OAP MLlib's org.apache.spark.ml.clustering.KMeans.fit() path.
 *
 * Every entry point replaces a piece of the reference's JNI surface (reference =
 * bobjiang82/oap-mllib, paths below are relative to mllib-dal/src/main/).  The JNI symbols the
 * Scala side binds are thin wrappers over these functions (csrc/jni_shim.cpp).
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * kmb200_status; kmb200_last_error() gives the thread-local message.  No function falls back to a
 * CPU implementation: without a usable sm_100 GPU every compute call fails with KMB200_ERR_DEVICE.
 * Handles are not thread-safe; different handles may be used from different threads.  Compute
 * calls (fit / assign) on tables of the SAME device run one at a time (a per-device lock inside the
 * library); calls on different devices run concurrently.  kmb200_set_option is process-wide: set
 * options before starting threads.
 */
#ifndef KMB200_H_
#define KMB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    KMB200_OK = 0,
    KMB200_ERR_ARG = -1,      /* bad argument (null, negative size, shape mismatch)            */
    KMB200_ERR_DEVICE = -2,   /* no sm_100 device / CUDA runtime error                         */
    KMB200_ERR_MEMORY = -3,   /* device or pinned-host allocation failed                       */
    KMB200_ERR_COMM = -4,     /* NCCL unavailable or returned an error                         */
    KMB200_ERR_EMPTY = -5,    /* more empty clusters than positive-distance candidate rows     */
    KMB200_ERR_LIMIT = -6     /* size beyond a documented limit                                */
} kmb200_status;

typedef struct kmb200_table kmb200_table; /* one executor's row partition, FP32, resident in HBM */
typedef struct kmb200_comm kmb200_comm;   /* one rank of the NCCL communicator                    */

/* ---- platform gate ----------------------------------------------------------------------------
 * Replaces Java_org_apache_spark_ml_util_OneDAL_00024_cCheckPlatformCompatibility
 * (native/OneDAL.cpp:96-102, "is this an Intel CPU") with "is there at least one sm_100 GPU".
 * Returns 1 / 0 (never an error code). */
int kmb200_device_ok(void);
/* Number of visible CUDA devices (0 when the driver is absent). */
int kmb200_device_count(void);
/* Selects the CUDA device used by the calling thread for subsequent calls (one executor = one
 * GPU = one rank; S/ml/clustering/KMeansDALImpl.scala:48-50 uses the partition index as rank). */
int kmb200_set_device(int32_t ordinal);

/* ---- data table ---------------------------------------------------------------------------------
 * Replaces the oneDAL HomogenNumericTable<double> conversion: DALMatrix allocation
 * (scala/.../ml/util/OneDAL.scala:125-127), one JNI cSetDoubleBatch per row (:135-139 ->
 * native/OneDAL.cpp:50-60), RowMergedNumericTable + cAddNumericTable (:155-163 ->
 * native/OneDAL.cpp:67-76).  Here: one device table of `rows` x `cols` FP32 (leading dimension
 * padded to 32 floats), filled by BATCHED row uploads from host memory (pinned or pageable). */
int kmb200_table_create(int64_t rows, int32_t cols, kmb200_table **out);
/* Copies nrows x ncols row-major values to table rows [row0, row0+nrows).  ncols must equal the
 * table's cols.  f64 input is rounded to FP32 on the device (the reference's own test accepts FP32
 * input at abs 1e-6: IntelKMeansSuite.scala:204-218). */
int kmb200_table_append_f64(kmb200_table *t, int64_t row0, const double *rows, int64_t nrows, int32_t ncols);
int kmb200_table_append_f32(kmb200_table *t, int64_t row0, const float *rows, int64_t nrows, int32_t ncols);
/* ---- other input layouts (what feeds the path) ------------------------
 * The reference accepts whatever `ml.linalg.Vector` rows the DataFrame holds and turns each one into a dense
 * double[] on the JVM (`row.toArray`, scala/.../ml/util/OneDAL.scala:135-139) before its per-row JNI copy.
 * Here the conversion happens on the device:
 *   append_columns_*  columns[c] points to nrows contiguous values of column c (an Arrow primitive array's data
 *                     buffer, a Spark ColumnVector, a NumPy column): staged column by column, transposed and
 *                     rounded to FP32 by a tiled kernel;
 *   append_csr_f64    sparse rows (ml.linalg.SparseVector): indptr has nrows + 1 entries, indices / values are
 *                     addressed by indptr's values; absent entries are 0; an index outside [0, cols) is an error.
 * Same row-range rules and FP32 rounding as append_f64. */
int kmb200_table_append_columns_f64(kmb200_table *t, int64_t row0, const double *const *columns, int64_t nrows,
                                    int32_t ncols);
int kmb200_table_append_columns_f32(kmb200_table *t, int64_t row0, const float *const *columns, int64_t nrows,
                                    int32_t ncols);
int kmb200_table_append_csr_f64(kmb200_table *t, int64_t row0, int64_t nrows, const int64_t *indptr,
                                const int32_t *indices, const double *values);
/* Replaces the reference's counting pass (Utils.getPartitionDims, scala/.../ml/util/Utils.scala:28-49: a whole
 * Spark job that only counts rows per partition so that tables can be sized): create the table with an upper
 * bound, append, then cut it to the rows actually appended (1 <= rows <= the rows given at creation).  Drops the
 * per-(table, k) workspace. */
int kmb200_table_truncate(kmb200_table *t, int64_t rows);
/* Page-locked host memory for the caller's staging buffers (the JVM side wraps it in a direct ByteBuffer and
 * fills it instead of a double[]; replaces the GetPrimitiveArrayCritical + memcpy of native/OneDAL.cpp:50-60):
 * appends from it run at the PCIe link rate. */
int kmb200_pinned_alloc(size_t bytes, void **out);
int kmb200_pinned_free(void *p);
/* Reads rows back as FP32 (diagnostics / tests; the reference has numericTableToVectors,
 * OneDAL.scala:37-52, for the same purpose). */
int kmb200_table_read_rows(const kmb200_table *t, int64_t row0, int64_t nrows, float *out);
int64_t kmb200_table_rows(const kmb200_table *t);
int32_t kmb200_table_cols(const kmb200_table *t);
/* Benchmark utility (no reference counterpart): fills rows [0, rows) with synthetic Gaussian
 * blobs ON THE DEVICE: row i belongs to blob hash(seed, row_offset+i) mod k_true and equals
 * centres[blob] + N(0,1)^cols (counter-based RNG).  centres is k_true x cols FP32 on the host. */
int kmb200_table_fill_blobs(kmb200_table *t, const float *centres, int32_t k_true, uint64_t seed,
                            int64_t row_offset);
/* Replaces cFreeDataMemory (native/OneDAL.cpp:83-89) / OneDAL.releaseNumericTables
 * (OneDAL.scala:81-90). */
int kmb200_table_free(kmb200_table *t);

/* ---- communicator -------------------------------------------------------------------------------
 * Replaces the oneCCL KVS rendezvous: OneCCL.init -> c_init (native/OneCCL.cpp:47-86), cleanup
 * (:93-99), isRoot / rankID (:106-120), and drops the ip/port probing (c_getAvailPort :207-247).
 * The Spark driver (or rank 0) calls kmb200_comm_unique_id once; the 128 bytes travel inside the
 * task closure; each executor calls kmb200_comm_init(world, rank, id). */
#define KMB200_UNIQUE_ID_BYTES 128
int kmb200_comm_unique_id(uint8_t id[KMB200_UNIQUE_ID_BYTES]);
int kmb200_comm_init(int32_t world, int32_t rank, const uint8_t id[KMB200_UNIQUE_ID_BYTES], kmb200_comm **out);
int32_t kmb200_comm_rank(const kmb200_comm *c);
int32_t kmb200_comm_size(const kmb200_comm *c);
int kmb200_comm_destroy(kmb200_comm *c);

/* ---- the Lloyd driver ---------------------------------------------------------------------------
 * Replaces Java_org_apache_spark_ml_clustering_KMeansDALImpl_cKMeansDALComputeWithInitCenters
 * (native/KMeansDALImpl.cpp:175-249): the loop `for (it = 0; it < maxIter && !converged; it++)`,
 * oneDAL step1Local (:70-77), the gather + step2Master merge/finalize (:80-130, here ONE in-place
 * ncclAllReduce of [k*d sums | k counts | cost]), and the convergence test (:135-168).
 *
 *   table        this rank's rows;  comm NULL => single GPU
 *   init_centers k x d row-major FP64 (what OneDAL.makeNumericTable(centers) built, OneDAL.scala:62-79)
 *   tol          absolute Euclidean distance per centre; every centre must move <= tol.
 *                tol < 0 disables the test (exactly max_iter iterations; benchmark use).
 *   out_centers  k x d FP64, written on EVERY rank (the reference returns them on rank 0 only)
 *   out_cost     objective of the LAST executed assignment step (reference :125-127,:243)
 *   out_iters    executed iterations (reference :241).  max_iter == 0 returns the initial centres
 *                with cost 0.0 (the reference leaves the cost uninitialised, :195).
 * Collective: every rank of comm must call it with the same k, d, tol, max_iter, init_centers. */
int kmb200_fit(kmb200_table *table, kmb200_comm *comm, const double *init_centers, int32_t k, int32_t d,
               double tol, int32_t max_iter, double *out_centers, double *out_cost, int32_t *out_iters);

/* Post-fit assignment pass (replaces the JVM pass behind KMeansSummary:
 * spark-3.1.1/ml/clustering/KMeans.scala:134-142,359-368).  labels_host (may be NULL) receives the
 * cluster index of every local row; sizes_host (may be NULL, k entries) the cluster sizes summed
 * over all ranks; cost_host (may be NULL) the summed squared distances. */
int kmb200_assign(kmb200_table *table, kmb200_comm *comm, const double *centers, int32_t k, int32_t d,
                  int32_t *labels_host, int64_t *sizes_host, double *cost_host);

/* ---- instrumentation (replaces the integer-second prints at native/KMeansDALImpl.cpp:218-222) ---
 * Device-side timings (CUDA events on the library's stream) of the last kmb200_fit /
 * kmb200_assign on this thread, summed over iterations, in milliseconds. */
typedef struct {
    int32_t iterations;
    int32_t kernel_launches;   /* kernels of this library launched                           */
    int64_t rechecked_rows;    /* rows whose top-2 gap was inside the error bound (FP64 re-check) */
    int32_t empty_cluster_events;
    int64_t full_rechecks;     /* of those, rows that needed all k centres re-scored (third centre inside the bound) */
    float total_ms;            /* first kernel of the first iteration -> end of the last         */
    float assign_ms;           /* tcgen05 distance GEMM + argmin epilogue (all launches of the phase) */
    float recheck_ms;          /* FP64 re-evaluation of near-tie rows                            */
    float bucket_ms;           /* label histogram + scan + row bucketing                         */
    float accumulate_ms;       /* per-cluster FP64 sums / cost                                   */
    float allreduce_ms;        /* NCCL all-reduce (0 for a single GPU)                           */
    float finalize_ms;         /* centres, convergence flag, TF32 split of the new centres      */
    float assign_main_ms;      /* of assign_ms: the main assign kernel alone (without the candidate gather and the
                                  second pass over the near-tie rows of the two-level scheme)    */
    int64_t coarse_undecided_rows; /* cascade: rows the one-product pass could not decide (summed over iterations;
                                      0 when the cascade is off)                                  */
    int32_t cascade_dropped_at;    /* iteration (0-based) in which the coarse pass left more than cand1_frac of the
                                      rows undecided and the fit went on without it; -1 = never     */
    int32_t two_level_dropped_at;  /* iteration in which the near-tie copy overflowed and the fit went on with the
                                      single-pass bookkeeping; -1 = never                            */
    int32_t assign_variant;        /* what the LAST assignment step ran: 0 = one-pass small-k kernel (CUDA cores),
                                      1 = FP16-split kernel, cascade, 2 = FP16-split, two-level, 3 = FP16-split, single
                                      pass, 4 = CTA-pair TF32 kernels, 5 = single-CTA resident-X kernel, 6 = streaming
                                      3xTF32 kernel, 7 = FP64 CUDA-core kernel ("force_simt")       */
} kmb200_profile;
int kmb200_get_profile(kmb200_profile *out);

/* Tunables (process-wide; for experiments and tests).  name: "eps_scale" (multiplier on the
 * near-tie error bound, default 1.0), "passes" (3 = full-accuracy tensor-core schemes [default],
 * 1 = single TF32 pass + wider re-check), "force_simt" (1 = use the CUDA-core FP64 assign kernel
 * instead of tcgen05; debugging).  Kernel selection (all default to the production choice; every
 * variant gives the same fit, tests/test_gpu_parity.py): "h16" (1 = FP16-split CTA-pair kernel for
 * d <= 256), "two_level" (1 = fast pass + full bookkeeping over the near-tie rows for tables of
 * >= 2^18 rows, 2 = always, 0 = never), "cand_cap" (rows of the near-tie copy, 0 = rows/16),
 * "cascade" (1 = coarse one-product pass first, wherever the two-level scheme runs and a row tile carries enough tensor
 * work for it -- at least four 64-value units over all its centroid tiles; 2 = regardless of that; 0 = never), "cand1_frac"
 * (rows of the coarse pass's copy as a fraction of the table, default 0.3), "stage_mb" (device
 * staging buffer of the columnar / sparse / FP64 uploads, default 256), "dbg_products" (1 = the
 * score dump of kmb200_debug_scores shows the coarse pass),
 * "cluster", "kernel", "fused", "rb_cap" (fallback kernels), and the diagnostics
 * "knockout" / "role_prof" (timing experiments; results are invalid with knockout != 0).
 * PCA: "pca_flush_rows" (rows a tensor-core FP32 accumulator holds before it is added into the next level, default 256:
 * its truncating adds cost 4e-9 of a diagonal entry per row), "pca_promote" (such segments per promotion to FP64,
 * default 16), "pca_rows" (rows per stage on the diagonal panels, 64 or 32), "pca_ls" / "pca_hs" (landing / operand
 * ring depths, even, 0 = default), and the diagnostics "pca_role_prof", "pca_knockout", "pca_desc_swap". */
int kmb200_set_option(const char *name, double value);

/* ---- cluster sizes without any device allocation ------------------------------------------------
 * No counterpart in the reference (its summary pass counts on the JVM).  Histogram of a label
 * vector: sizes_dev[j] = number of i in [0, n) with labels_dev[i] == j.  Every other entry point of
 * this library cudaMallocs its scratch; this one does not call cudaMalloc, cudaFree or any other
 * allocating routine, so it is usable by hosts that cannot allocate on the device when they call
 * in: executors whose GPU memory has been given entirely to a pool allocator (cudaMalloc then fails
 * even though the pool has room) and latency-bound loops that must avoid the device-wide
 * synchronisation an allocation or a free implies.
 *
 *   labels_dev  DEVICE pointer, n int32 values, each in [0, k) (anything else is undefined behaviour)
 *   sizes_dev   DEVICE pointer, k int64 values, overwritten
 *   cuda_stream a cudaStream_t (NULL = the default stream); the work is only ENQUEUED on it, the
 *               result is valid once that stream has run it
 *   limits      k <= 65536: the k int32 counters of scratch sit in a fixed 256 KiB array inside
 *               the library image (each GPU that loads the library holds it until the process
 *               exits); n <= 2^31 - 1, because those counters are 32-bit and one cluster may hold
 *               every row.  Anything larger is refused with KMB200_ERR_LIMIT
 *
 * The first call on a device looks up that array's device address (one driver query per device).
 * With CUDA's lazy module loading that query can be what loads this library's device code, array
 * included, onto the GPU -- the driver, not this function, takes memory for it then.  So make the
 * first call at start-up, before the GPU's memory is handed to the pool, and before starting any
 * stream capture.  All calls on one device
 * share the array, which keeps the previous call's counts until the next call clears the part it
 * needs; two calls whose work overlaps on the GPU (different streams, no dependency) race on it and
 * both results are then garbage -- order them with an event or use one stream. */
int kmb200_cluster_sizes_noalloc(const int32_t *labels_dev, int64_t n, int32_t k, int64_t *sizes_dev,
                                 void *cuda_stream);

/* Test hook (no reference counterpart): raw tensor-core scores ||c_j||^2 - 2 x.c_j (FP32) of every
 * row against every padded centre column (rows x kpad, row-major).  scores_host may be NULL to
 * query kpad only.  Lets the tests measure the 3xTF32 error against FP64. */
int kmb200_debug_scores(kmb200_table *table, const double *centers, int32_t k, int32_t d, float *scores_host,
                        int32_t *kpad_out);

/* ---- k-means|| initialisation passes --------------------------------------------
 * Replace the JVM passes over every row in initKMeansParallel (scala/org/apache/spark-3.1.1/mllib/clustering/
 * KMeans.scala:370-432); the driver-side logic (the loop over initSteps, de-duplication, the weighted local k-means++
 * over the ~2k candidates) stays with the caller.
 *   init_costs   :395-401  cost_i = min(cost_i, min_j ||x_i - c_j||^2) over the m centres added in the last step (first_step
 *                != 0: the costs start at +infinity), kept on the device per table; *sum_costs = the sum over all
 *                ranks (`costs.sum()`).  Collective when comm is given.
 *   init_sample  :406-410  the rows kept by `rand.nextDouble() < 2.0 * c * k / sumCosts` with
 *                rand = XORShiftRandom(seed ^ (step << 16) ^ partitionIndex), one draw per row in row order;
 *                part_rows[0..nparts] are the row offsets of the Spark partitions inside this table (part_rows[0] = 0,
 *                part_rows[nparts] = rows), part_index0 the Spark index of the first.  Up to cap row numbers are
 *                written to rows_out in ascending order, *count_out is the number kept (it may exceed cap).
 * xorshift_next_int / next_double: Spark's XORShiftRandom(seed).nextInt() and the (skip + 1)-th nextDouble(), for the
 * caller's own seed derivation (`new XORShiftRandom(this.seed).nextInt()`, :377) and for tests.  XORShiftRandom is
 * restated from Spark's published source (it is not in the reference tree, nor are vectors for it): unpinned. */
int kmb200_init_costs(kmb200_table *table, kmb200_comm *comm, const double *new_centers, int32_t m, int32_t d,
                      int32_t first_step, double *sum_costs);
int kmb200_init_sample(kmb200_table *table, int32_t seed, int32_t step, int32_t k, double sum_costs,
                       const int64_t *part_rows, int32_t nparts, int32_t part_index0, int64_t *rows_out, int64_t cap,
                       int64_t *count_out);
int32_t kmb200_xorshift_next_int(int64_t seed);
double kmb200_xorshift_next_double(int64_t seed, int32_t skip);

/* ---- PCA -------------------------------------------------------------------------
 * Replaces the body of Java_org_apache_spark_ml_feature_PCADALImpl_cPCATrainDAL (mllib-dal/src/main/native/
 * PCADALImpl.cpp:38-190): oneDAL's pca::Distributed<step1Local / step2Master, double, svdDense> over the executors'
 * tables with an all-gather of the partial results in between.  Here: one pass of column statistics, the Gram matrix of
 * the resident FP32 table on the tensor cores (FP16-split products, FP64 partial sums), one all-reduce of d x d doubles,
 * and the eigen-decomposition on the host -- the same eigenvalues (sigma^2 / (N - 1)) and principal axes.
 *   center != 0: the column means are subtracted on the device (what PCADALImpl.scala:100-106 `normalizeData` does with
 *                two JVM passes before the table is built); center == 0: the table is taken as it is, which is what the
 *                reference's native code does with the rows it is given (it flags them "already normalised", :50-51).
 *   pca_train  : eigenvalues[d] descending; eigenvectors[d x d] row-major, ONE PRINCIPAL AXIS PER ROW like the
 *                reference's result table (PCADALImpl.scala:108-127 copies the first k rows into the columns of pc).
 *                The sign of an axis is not defined by the reference (LAPACK's choice; IntelPCASuite.scala:75-84
 *                compares absolute values): here the entry of largest magnitude is positive.  Collective when comm is
 *                given; every rank receives the result.
 *   pca_gram   : the (centred) Gram matrix sum_i (x_i - mean)(x_i - mean)^T [d x d], the means [d] (zeros when
 *                center == 0) and the global row count -- tests and callers that want the covariance itself.
 *   sym_eig    : the host eigen-solver alone (symmetric n x n row-major in; eigenvalues descending, eigenvectors as rows);
 *                needs no device.
 * d is not limited; rows per GPU as for every table.  Profile of a PCA call (kmb200_get_profile): accumulate_ms = the
 * column-statistics pass, assign_ms = the Gram kernels, allreduce_ms, total_ms = first kernel to the last copy (device
 * side), finalize_ms = the eigen-solver on the host (wall clock). */
int kmb200_pca_train(kmb200_table *table, kmb200_comm *comm, int32_t center, double *eigenvalues, double *eigenvectors);
int kmb200_pca_gram(kmb200_table *table, kmb200_comm *comm, int32_t center, double *gram, double *mean, int64_t *n_total);
int kmb200_sym_eig(const double *a, int32_t n, double *eigenvalues, double *eigenvectors);

/* Diagnostics (no reference counterpart): copies `bytes` bytes, starting `offset` bytes in, of one of the device
 * arrays the last fit / assign on this table left behind ("labels", "flag_list", "flag_count", "cand_list",
 * "labels_c", "cand1_list", "labels_c1", "xc", "xc1") to host memory.  Used by tools/repro_check.py. */
int kmb200_debug_dump(kmb200_table *table, const char *name, size_t offset, void *out, size_t bytes);

const char *kmb200_last_error(void);
const char *kmb200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* KMB200_H_ */
