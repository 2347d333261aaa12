/* slq.h -- C ABI of libslq, a B200 (sm_100a) stochastic-Lanczos-quadrature engine.
 *
 * This is the drop-in boundary for ONE hot path of peekxc/primate (citations are relative to the reference
 * checkout): probe generation -> probe-blocked Lanczos -> Golub-Welsch quadrature -> f(theta).tau -> running
 * mean/variance.  It replaces the pybind11 module `primate._lanczos` (src/primate/_lanczos.cpp:86-112) and
 * the per-probe Python loop wrapped around it (src/primate/operators.py:126-151, src/primate/trace.py:107-116).
 *
 * Conventions
 *  - Plain pointers and sizes only.  Every `*_dev` pointer is CUDA device memory owned by the CALLER (the
 *    Python host side allocates it as torch tensors); the library never allocates on the hot path.
 *  - A "probe block" is a dense n x k panel stored ROW-MAJOR (probe-minor): element (row r, probe p) lives at
 *    X[r*k + p], so the k values of one row are one contiguous, 32-byte-aligned vector (device pointers of probe blocks must be 32-byte aligned).  k must be a power of
 *    two in [1, 64].  (The reference stores probes F-ordered (N,k), src/primate/random.py:74-78.)
 *  - `stream` is a cudaStream_t passed as void*; all work is enqueued on it and is stream-ordered.  No entry
 *    point synchronises the device unless documented.
 *  - Every function returns SLQ_OK (0) or a negative error code; slq_last_error() returns a thread-local
 *    message.  There is NO CPU fallback: without a CUDA device every compute entry fails with SLQ_ERR_CUDA.
 */
#ifndef SLQ_H_
#define SLQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the entry points below are the only exported symbols of libslq.so */
#endif

#define SLQ_OK 0
#define SLQ_ERR_ARG (-1)         /* invalid argument (reference: TypeError / ValueError / assert)           */
#define SLQ_ERR_CUDA (-2)        /* CUDA runtime error or no device                                         */
#define SLQ_ERR_UNSUPPORTED (-3) /* valid request the library does not implement                            */
#define SLQ_ERR_CALLBACK (-4)    /* a user operator callback returned non-zero (reference: Python exception) */

enum slq_dtype { SLQ_F32 = 0, SLQ_F64 = 1 };

/* Built-in spectral functions; mirror of primate.special._BUILTIN_MATRIX_FUNCTIONS and param_callable
 * (src/primate/special.py:7,78-107).  params[] meaning: EXP: t; SMOOTHSTEP: a, b; NUMRANK: threshold;
 * SOFTSIGN: q (integer valued). */
enum slq_fun {
  SLQ_FUN_IDENTITY = 0,
  SLQ_FUN_ABS = 1,
  SLQ_FUN_SQRT = 2,
  SLQ_FUN_LOG = 3, /* log(max(x, DBL_EPSILON)), special.py:90 */
  SLQ_FUN_INV = 4,
  SLQ_FUN_EXP = 5,
  SLQ_FUN_SMOOTHSTEP = 6,
  SLQ_FUN_NUMRANK = 7,
  SLQ_FUN_SOFTSIGN = 8,
  SLQ_FUN_NONE = -1 /* do not evaluate f: only theta/tau are produced */
};

typedef struct slq_op slq_op; /* opaque symmetric linear operator, device resident */
typedef void* slq_stream;     /* cudaStream_t */

/* User operator: Y[n x k] = A * X[n x k] on device panels in the block layout above, enqueued on `stream`
 * (or completed before returning).  Replaces the C++20 `LinearOperator`/`SupportsMatrixMult` concepts
 * (src/primate/include/linear_operator.h:25-29,59-62) and PyLinearOperator (include/pylinop.h:17-40). */
typedef int (*slq_matmat_fn)(void* ctx, const void* X_dev, void* Y_dev, int64_t n, int k, int dtype,
                             slq_stream stream);

/* ---- library ------------------------------------------------------------------------------------------ */
int slq_version(void);
const char* slq_last_error(void);
/* SM count and compute capability of the current device (fails with SLQ_ERR_CUDA without one). */
int slq_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* Number of CUDA kernels this library has launched in this process (bench.py's `gpu_launches`). */
long long slq_launch_count(void);
/* Optional timing of the library's own kernels with CUDA events on the launching stream, per kernel class:
 * 0 = fused SpMM + recurrence kernel (Lanczos "step A"), 1 = axpy/norm/re-projection kernels, 2 = quadrature,
 * 3 = probe generation.  enable(max) records at most `max` launches per class from now on (0 disables);
 * read() synchronises on the recorded events and returns their summed duration, the number of launches timed and
 * the number seen.  The reference has only dead `matvec_time` counters (include/eigen_operators.h:21,25,29). */
int slq_profile_enable(int max_launches_per_class);
int slq_profile_read(int kernel_class, double* total_ms, long long* timed_launches, long long* seen_launches);

/* ---- operators (replace the per-call matrix wrappers: include/eigen_operators.h:18-104, _lanczos.cpp:93) - */
/* CSR with int32 indices; arrays stay owned by the caller and must outlive the handle.  A must be symmetric
 * (as in the reference, which feeds the same symmetric matrix in CSC form). Column indices sorted per row is
 * recommended (keeps the summation order of Eigen's column sweep) but not required. */
int slq_csr_create(int64_t n, int64_t nnz, const int32_t* rowptr_dev, const int32_t* colidx_dev,
                   const void* vals_dev, int dtype, slq_op** out);
/* As slq_csr_create, with the longest row given by the caller (it selects the SpMM variant: rows of more than 64
 * non-zeros take the non-zero-balanced kernel).  max_row_nnz >= 0: trusted, the call does no device work and never
 * synchronises (a host-ingested matrix knows its row lengths).  max_row_nnz < 0: computed by a kernel enqueued on
 * `stream` -- after whatever is still producing the CSR arrays there -- and `stream` is synchronised to read it back;
 * this is the ONLY entry point of the library that synchronises.  slq_csr_create == slq_csr_create_ex(.., -1, NULL). */
int slq_csr_create_ex(int64_t n, int64_t nnz, const int32_t* rowptr_dev, const int32_t* colidx_dev,
                      const void* vals_dev, int dtype, int64_t max_row_nnz, slq_stream stream, slq_op** out);
/* Dense symmetric n x n, leading dimension ld (row- or column-major is the same for a symmetric matrix). */
int slq_dense_create(int64_t n, const void* A_dev, int64_t ld, int dtype, slq_op** out);
/* As slq_dense_create for an operator whose order is not a multiple of 16: A_dev is an npad x ld array (npad = n rounded
 * up to 16, ld >= npad) whose rows and columns >= n are ZERO.  Panels keep n rows.  The padding lets every n take the
 * bulk-copy / tensor-core dense kernels (whole 16-byte units) instead of the generic fallback. */
int slq_dense_create_padded(int64_t n, int64_t npad, const void* A_dev, int64_t ld, int dtype, slq_op** out);
/* Optional caller-owned scratch for an operator (must outlive it; NULL detaches).  Small dense operators use it to split
 * the reduction dimension of Y = A X over several CTAs per row tile (partial slabs of n x k, summed in a fixed order):
 * 16 n k sizeof(T) bytes allow the full split.  Calls that use the scratch must not overlap on different streams. */
int slq_op_set_scratch(slq_op* op, void* scratch_dev, size_t bytes);
/* CSR operators, orth = 0, rows of <= 64 non-zeros: slq_lanczos runs ONE kernel per Lanczos step (the update
 * v = w - alpha q of the previous step is applied on the fly inside the next step's SpMM, DESIGN 3 "fused step") instead
 * of two (lanczos.h:123-143 as step A + axpy).  It moves 4 panel passes per step instead of 6 but reads TWO rows per
 * gathered non-zero, so it pays when most non-zeros are local to a row tile of 120 rows (a mesh in the locality order:
 * +20..30 % on a 2-D 5-point Laplacian) and loses when more than about two non-zeros per row leave the tile (3-D 7-point
 * Laplacian: -6 %).  enable = 0 keeps the two-kernel step for this operator; the default is 1.  Results agree to
 * rounding (alpha, beta <= 1e-11 relative; same early-break steps). */
int slq_op_set_fused_step(slq_op* op, int enable);
/* Optional band-major copy of an fp64 dense operator for the tensor-core product kernel (the caller owns the buffer; it
 * must outlive the operator or be detached with buf_dev = NULL).  slq_dense_tiled_bytes reports the size (0 when the
 * operator does not take that kernel: fp32, order not a multiple of 16 and not created padded); slq_dense_set_tiled
 * fills the buffer on `stream` (one pass over A) and switches Y = A X to it.  Row-major A makes every pipeline stage 16
 * strided lines of 1.8 KB and DRAM page conflicts hold the product at 4.2 TB/s (so is cuBLAS DGEMM on the same operand);
 * the copy stores each stage image contiguously, one bulk copy per stage, 148 sequential streams.  Results are bit-identical
 * with and without it.  Replaces the per-call column-major copy of include/eigen_operators.h:24-30 (made once here). */
int slq_dense_tiled_bytes(const slq_op* op, size_t* bytes);
int slq_dense_set_tiled(slq_op* op, void* buf_dev, size_t bytes, slq_stream stream);
/* Matrix-free operator driven by a callback (device panels in, device panels out). */
int slq_callback_create(int64_t n, int dtype, slq_matmat_fn fn, void* ctx, slq_op** out);
int slq_op_destroy(slq_op* op);
int64_t slq_op_rows(const slq_op* op);
int slq_op_dtype(const slq_op* op);
int64_t slq_op_nnz(const slq_op* op);
/* Y = A X for a probe block (the SpMM on its own; eigen_operators.h:66-99 `matvec`/`matmat`). */
int slq_op_matmat(slq_op* op, const void* X_dev, void* Y_dev, int k, slq_stream stream);

/* ---- operator ingestion: locality ordering (no reference counterpart; the reference re-converts the matrix to CSC on
 * every call, src/primate/_lanczos.cpp:89-93) ------------------------------------------------------------------ */
/* HOST arrays.  perm_out[new] = old: a greedy breadth-first clustering of the symmetric sparsity graph into groups of
 * `tile` consecutive rows, so that most non-zeros of a row tile point into the tile itself (L1 hits in the SpMM instead
 * of repeated L2 fetches).  P A P^T has the same spectrum and, started from P x, the same Lanczos coefficients. */
int slq_csr_cluster_order(int64_t n, const int32_t* rowptr_host, const int32_t* colidx_host, int tile, int32_t* perm_out_host);
/* HOST arrays.  B = P A P^T for perm[new] = old; output arrays sized like the input; column indices sorted per row. */
int slq_csr_permute(int64_t n, const int32_t* rowptr_host, const int32_t* colidx_host, const void* vals_host, int dtype,
                    const int32_t* perm_host, int32_t* rowptr_out_host, int32_t* colidx_out_host, void* vals_out_host);
/* Device probe blocks: out[r][:] = in[idx[r]][:] (gather) and out[idx[r]][:] = in[r][:] (scatter); in != out. */
int slq_block_gather_rows(const void* in_dev, const int32_t* idx_dev, int64_t n, int k, int dtype, void* out_dev, slq_stream stream);
int slq_block_scatter_rows(const void* in_dev, const int32_t* idx_dev, int64_t n, int k, int dtype, void* out_dev, slq_stream stream);

/* ---- probes (src/primate/random.py:22-29,47-80 on NumPy's PCG64) ---------------------------------------- */
/* Rademacher block: X[r*k + p] = +1 if the MSB of PCG64 raw output #((first_probe+p)*n + r) is set else -1,
 * for the generator whose 128-bit (state, inc) BEFORE any draw are given as {hi, lo} pairs.  Bit-exact with
 * primate.random.isotropic(pdf="rademacher") drawn column by column from that generator. */
int slq_probes_rademacher(const uint64_t state_hi_lo[2], const uint64_t inc_hi_lo[2], uint64_t first_probe,
                          int k, int64_t n, int dtype, void* X_dev, slq_stream stream);
/* ---- normal / sphere probes (src/primate/random.py:32-41 on NumPy's ziggurat, distributions.c random_standard_normal) --- */
/* NumPy version whose ziggurat tables are compiled in (bit-exactness is pinned against that NumPy). */
const char* slq_normals_numpy_version(void);
/* Scratch for a scan of `nraw` raw draws (slq_normals_scan / slq_probes_normal). */
size_t slq_normals_workspace_bytes(int64_t nraw);
/* A normal variate consumes a data-dependent number of raw draws, so "where does probe i start" needs a scan.  Scans
 * `nraw` raw draws from raw offset `raw_base` of the stream (a sample start) on the device: *total_dev = number of
 * samples that start inside; if 0 <= target_ordinal < that, *raw_out_dev = raw position (relative to raw_base) where
 * sample #target_ordinal starts.  Used to give every rank O(1)-communication random access to its probe range. */
int slq_normals_scan(const uint64_t state_hi_lo[2], const uint64_t inc_hi_lo[2], uint64_t raw_base, int64_t nraw,
                     int64_t target_ordinal, int64_t* total_dev, uint64_t* raw_out_dev, void* workspace_dev,
                     size_t workspace_bytes, slq_stream stream);
/* Standard-normal probe block, bit-exact with Generator.standard_normal drawn column by column: X[r*k + p] (float64)
 * = sample #(p*n + r) of the stream starting at raw offset raw_base (a sample start).  nraw >= n*k raw positions are
 * scanned (about 1.022 n k are needed); *end_raw_dev = raw draws the block consumed, or UINT64_MAX if nraw was too
 * small (retry with more).  The ~2.6e-4 tail samples are also listed -- tail_ord[i] = (ordinal << 1) | negative,
 * tail_raw[i] = raw draw U1 -- so that the caller can recompute  +-(r + (-log1p(-U1) / r))  with the host's libm (the
 * one NumPy calls) and overwrite them: every other sample does not depend on libm's last bit. */
int slq_probes_normal(const uint64_t state_hi_lo[2], const uint64_t inc_hi_lo[2], uint64_t raw_base, int64_t nraw, int k,
                      int64_t n, double* X_dev, uint64_t* end_raw_dev, uint64_t* tail_count_dev, int64_t tail_cap,
                      int64_t* tail_ord_dev, uint64_t* tail_raw_dev, void* workspace_dev, size_t workspace_bytes,
                      slq_stream stream);
/* In-place sphere scaling X[:, p] *= sqrt(n) / ||X[:, p]|| of a float64 block, the norms summed in NumPy's pairwise
 * order (np.sum(W**2, axis=0): blocks of <= 128 with 8 accumulators, halved at multiples of 8).  leaf_start/leaf_len
 * list the blocks, reduce_after[i] the number of pair-additions after block i; S_dev nleaves*k doubles, norm2_dev k. */
int slq_probes_sphere_scale(double* X_dev, int64_t n, int k, const int64_t* leaf_start_dev, const int32_t* leaf_len_dev,
                            const int32_t* reduce_after_dev, int64_t nleaves, double* S_dev, double* norm2_dev,
                            slq_stream stream);
/* Raw PCG64 outputs #offset.. (testing aid for the generator itself). */
int slq_pcg64_raw(const uint64_t state_hi_lo[2], const uint64_t inc_hi_lo[2], uint64_t offset, int64_t count,
                  uint64_t* out_dev, slq_stream stream);
/* Layout helpers between the reference's F-ordered (n,k) panels (k contiguous columns) and probe blocks. */
int slq_block_from_columns(const void* cols_dev, int64_t n, int k, int dtype, void* X_dev, slq_stream stream);
int slq_block_to_columns(const void* X_dev, int64_t n, int k, int dtype, void* cols_dev, slq_stream stream);

/* ---- Lanczos (src/primate/include/lanczos.h:92-149 + orth_vector :43-66, blocked over k probes) --------- */
/* Bytes of caller-provided scratch needed by slq_lanczos. */
size_t slq_lanczos_workspace_bytes(int64_t n, int k, int ncv, int dtype);
/* Runs `deg` steps of Paige-A27 Lanczos independently for each of the k columns of X.
 *   X_dev      n x k probe block, start vectors; CLOBBERED (the reference clobbers `v`, lanczos.h:114,127).
 *   rtol       early-break threshold is sqrt(n)*rtol per probe (lanczos.h:110,140).
 *   orth       number of stored Lanczos vectors re-projected per step, walking back cyclically from the
 *              current one (lanczos.h:133-136); fresh-basis semantics (all ncv columns start at zero).
 *   ncv        number of Lanczos vectors kept (>= 2); the reference's Q.shape[1] (_lanczos.cpp:94).
 *   alpha_dev  k x (deg+1), probe-major; alpha[p][j], j < deg.  Written from zero (tails stay 0 after a break).
 *   beta_dev   k x (deg+1); beta[p][0] = 0, beta[p][j+1] = ||v_j||.
 *   steps_dev  k int32: number of steps executed per probe (deg unless the early break fired).  May be NULL.
 *   xnorm2_dev k doubles: ||x_p||^2 of the start vectors (operators.py:147).  May be NULL.
 *   Q_dev      NULL, or n x ncv x k: the normalised Lanczos vectors, Q[(c*n + r)*k + p] (for k == 1 this is
 *              exactly the reference's F-ordered (n, ncv) basis, lanczos.py:103).
 */
int slq_lanczos(slq_op* op, void* X_dev, int k, int deg, double rtol, int orth, int ncv, void* alpha_dev,
                void* beta_dev, int32_t* steps_dev, double* xnorm2_dev, void* Q_dev, void* workspace_dev,
                size_t workspace_bytes, slq_stream stream);

/* ---- quadrature (src/primate/integrate.py:58-64, tridiag.py:7-43, tqli.py:15-90) ------------------------ */
/* For each of k Jacobi matrices T_p = tridiag(alpha[p][0:deg], beta[p][1:deg]) (row stride ld, dtype as the
 * Lanczos run): Ritz values theta (ascending) and Golub-Welsch weights tau = (first eigenvector row)^2 by an
 * implicit-shift QL sweep, one warp per matrix; then y[p] = scale[p] * sum_i f(theta_i) tau_i fused in the
 * same kernel (operators.py:150).  Any of theta_dev/tau_dev (k x deg doubles), y_dev (k doubles),
 * scale_dev (k doubles; NULL = 1) and info_dev (k int32; #eigenvalues that hit the iteration cap) may be NULL.
 * fun = SLQ_FUN_NONE skips f.  params: 4 doubles (host). */
int slq_quadrature(const void* alpha_dev, const void* beta_dev, int dtype, int k, int deg, int ld, int fun,
                   const double* params, const double* scale_dev, double* theta_dev, double* tau_dev,
                   double* y_dev, int32_t* info_dev, slq_stream stream);
/* As slq_quadrature but also returns c_p = Y f(Theta) Y^T e_1 (k x deg doubles), the Krylov coefficients of
 * f(A)x (operators.py:122-124).  scratch_dev: (deg*deg + deg)*k doubles; on return it holds, per matrix p, the
 * eigenvector matrix Zt[p][col][row] (columns in QL order) followed at offset k*deg*deg by perm[p][i] = the Zt
 * column of the i-th smallest Ritz value (so eigh_tridiag's Y[:, i] = Zt[p][perm[p][i]][:], tridiag.py:25-43). */
int slq_funm_coeffs(const void* alpha_dev, const void* beta_dev, int dtype, int k, int deg, int ld, int fun,
                    const double* params, double* coeffs_dev, double* theta_dev, double* tau_dev,
                    double* scratch_dev, int32_t* info_dev, slq_stream stream);
/* Batched in-place implicit-shift QL with the contract of the reference's Pythran kernel `tqli(d, e, Z, max_iter)`
 * (src/primate/tqli.py:15-90): d_dev k x m diagonals -> eigenvalues in the order the sweep leaves them (NOT sorted);
 * e_dev k x m off-diagonals with a leading zero (e[i] couples i-1 and i) -> the negligible remainders; Zt_dev NULL, or
 * k x m x m with Zt[p][col][row] = Z[row][col] (the reference's Z, transposed): on return Z <- Z_in * eigenvectors.
 * At most max_iter sweeps per eigenvalue (tqli.py:54-56 tolerates running out; info_dev, if given, counts them). */
int slq_tqli(double* d_dev, double* e_dev, double* Zt_dev, int k, int m, int max_iter, int32_t* info_dev, slq_stream stream);
/* y[p] = scale[p] * sum_i f(theta[p][i]) * tau[p][i]   (fused evaluation on caller-provided nodes). */
int slq_apply_fun(const double* theta_dev, const double* tau_dev, int k, int deg, int fun, const double* params,
                  const double* scale_dev, double* y_dev, slq_stream stream);
/* FTTR weights (src/primate/fttr.py:5-29) for caller-provided nodes: theta k x deg -> tau k x deg, using the
 * first `nterms` orthonormal polynomials of each Jacobi matrix (the reference's n = len(alpha)). */
int slq_fttr(const double* theta_dev, const void* alpha_dev, const void* beta_dev, int dtype, int k, int deg,
             int nterms, int ld, double* tau_dev, slq_stream stream);

/* ---- running statistics (src/primate/stats.py:66-86, estimators.py:128-141) ----------------------------- */
/* Chan/Welford merge of a batch of m samples into state_dev = {count, mean, M2} (3 doubles, device). */
int slq_stats_update(const double* y_dev, int m, double* state_dev, slq_stream stream);

/* ---- f(A) X on the same engine (src/primate/operators.py:102-124) --------------------------------------- */
/* Y[n x k] = sum_j Q_j * (coeffs[p][j] * xnorm[p])  with Q as produced by slq_lanczos (n x ncv x k, ncv>=deg). */
int slq_basis_combine(const void* Q_dev, int64_t n, int k, int deg, int dtype, const double* coeffs_dev,
                      const double* xnorm2_dev, void* Y_dev, slq_stream stream);

/* ---- basis-free f(A) X: two-pass Lanczos (src/primate/operators.py:102-124 without the n x deg basis `_Q`) ------------ */
/* Y[n x k] = f(A) X for a built-in spectral function, orth = 0.  Pass 1 runs the Lanczos recurrence on a copy of X and
 * keeps only alpha/beta; the Krylov coefficients c_p = Y f(Theta) Y^T e_1 follow on the device; pass 2 re-runs the
 * recurrence from the intact X with the RECORDED alpha/beta -- no grid-wide reductions, so the whole step (SpMM, both
 * subtractions, Y += c_{j+1} ||x|| q_{j+1}) is ONE kernel for CSR operators -- producing bit-identical Lanczos vectors.
 * Memory is five panels whatever deg is (the reference holds n x deg: 10.7 GB per probe at n = 16.7 M, deg = 80).
 * X_dev is left intact; X_dev != Y_dev.  Workspace: slq_funm_workspace_bytes, 256-byte aligned. */
size_t slq_funm_workspace_bytes(int64_t n, int k, int deg, int dtype);
int slq_funm_matmat(slq_op* op, const void* X_dev, int k, int deg, double rtol, int fun, const double* params, void* Y_dev,
                    void* workspace_dev, size_t workspace_bytes, slq_stream stream);
/* Pass 2 on its own, for coefficients computed elsewhere (a Python callable evaluated on the host on the Ritz values):
 * alpha/beta (k x (deg+1), the run's dtype), steps (k) and xnorm2 (k) as slq_lanczos(orth = 0, ncv = 2) left them for the
 * SAME start block X_dev; coeffs k x deg doubles.  Y = sum_m coeffs[p][m] ||x_p|| q_m. */
size_t slq_replay_workspace_bytes(int64_t n, int k, int deg, int dtype);
int slq_lanczos_replay(slq_op* op, const void* X_dev, int k, int deg, const void* alpha_dev, const void* beta_dev,
                       const int32_t* steps_dev, const double* xnorm2_dev, const double* coeffs_dev, void* Y_dev,
                       void* workspace_dev, size_t workspace_bytes, slq_stream stream);

/* ---- tall-skinny panel algebra for the block drivers (replaces NumPy/LAPACK on host panels: np.linalg.qr, Q.T @ G,
 * Q @ S, einsum row/column dots -- src/primate/trace.py:158-175,201-227, diagonal.py:115-137) -------------------------------
 * Panels are n x m row-major device arrays with a row stride ld (elements); m x m results are fp64.  Deterministic. */
/* G[a x b] (doubles, row-major) = X[row0:row1, :a]^T Y[row0:row1, :b]  -- Gram matrices and projections; a row range
 * so that ranks can each reduce a slice of replicated panels and all-reduce a*b doubles. */
size_t slq_panel_gram_workspace_bytes(int a, int b);
int slq_panel_gram(const void* X_dev, int a, int64_t ldx, const void* Y_dev, int b, int64_t ldy, int64_t row0, int64_t row1,
                   int dtype, double* G_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream);
/* Y[n x b] = alpha X[n x a] S[a x b] + beta Y  (S: doubles, row-major; X != Y): Q = Y R^-1, G -= Q (Q^T G), Q S, ... */
int slq_panel_matmul(const void* X_dev, int a, int64_t ldx, const double* S_dev, int b, double alpha, double beta, void* Y_dev,
                     int64_t ldy, int64_t n, int dtype, slq_stream stream);
/* out[r] = sum_j X[r][j] Y[r][j]  (n doubles): the diagonal of X Y^T. */
int slq_panel_rowdot(const void* X_dev, int64_t ldx, const void* Y_dev, int64_t ldy, int m, int64_t n, int dtype, double* out_dev,
                     slq_stream stream);
/* out[j] = sum over rows [row0, row1) of X[r][j] Y[r][j]  (m doubles): the diagonal of X^T Y. */
size_t slq_panel_coldot_workspace_bytes(int m);
int slq_panel_coldot(const void* X_dev, int64_t ldx, const void* Y_dev, int64_t ldy, int m, int64_t row0, int64_t row1, int dtype,
                     double* out_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream);

/* Allocation-free variant of slq_panel_coldot.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same kernels, same result (bit-identical to slq_panel_coldot: same grid, same summation order),
 * but the per-CTA partial table lives in a static __device__ array inside the library instead of a caller workspace:
 * 512 x 256 doubles = 1 MiB held by every GPU that loads the library until the process exits.  Limit:
 * m <= SLQ_COLDOT_STATIC_MAX_M (256); a larger m is rejected with SLQ_ERR_ARG and a message naming the limit.  The
 * call allocates and frees nothing (the first call on a device looks the array's device address up with
 * cudaGetSymbolAddress).  ALL calls on a device share the one array and it keeps the last call's
 * contents: nothing needs to start at zero (every entry read is written first by the same call), but two calls that
 * overlap on the GPU (different streams, no ordering between them) RACE on it -- serialise them. */
#define SLQ_COLDOT_STATIC_MAX_M 256
int slq_panel_coldot_static(const void* X_dev, int64_t ldx, const void* Y_dev, int64_t ldy, int m, int64_t row0, int64_t row1,
                            int dtype, double* out_dev, slq_stream stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* SLQ_H_ */
