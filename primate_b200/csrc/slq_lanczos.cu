// C entry points of the Lanczos driver: operator handles, workspace sizing, slq_lanczos / slq_op_matmat dispatch on the
// scalar type (the kernels live in slq_lanczos_impl.cuh), and the probe-block layout helpers.
#include "slq_lanczos_common.cuh"

namespace slq {

// ---- layout helpers ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void k_cols_to_block(const T* __restrict__ cols, int64_t n, int k, T* __restrict__ X) {
  // cols: k contiguous columns of length n (F-ordered (n,k)); X: n x k row-major.  32x32 smem transpose tiles.
  __shared__ T tile[32][33];
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int p0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int p = p0 + i;
    const int64_t r = r0 + threadIdx.x;
    tile[i][threadIdx.x] = (p < k && r < n) ? cols[(int64_t)p * n + r] : (T)0;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int64_t r = r0 + i;
    const int p = p0 + threadIdx.x;
    if (r < n && p < k) X[r * k + p] = tile[threadIdx.x][i];
  }
}

template <typename T>
__global__ void k_block_to_cols(const T* __restrict__ X, int64_t n, int k, T* __restrict__ cols) {
  __shared__ T tile[32][33];
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int p0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int64_t r = r0 + i;
    const int p = p0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < n && p < k) ? X[r * k + p] : (T)0;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int p = p0 + i;
    const int64_t r = r0 + threadIdx.x;
    if (p < k && r < n) cols[(int64_t)p * n + r] = tile[threadIdx.x][i];
  }
}

// longest row of a CSR operator (slq_csr_create_ex with max_row_nnz < 0): result in a caller-provided device int
static __global__ void k_max_row_nnz(int64_t n, const int32_t* __restrict__ rowptr, int* __restrict__ out) {
  int m = 0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x)
    m = max(m, __ldg(rowptr + r + 1) - __ldg(rowptr + r));
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

}  // namespace slq

using namespace slq;

extern "C" {

int slq_csr_create_ex(int64_t n, int64_t nnz, const int32_t* rowptr_dev, const int32_t* colidx_dev, const void* vals_dev, int dtype,
                      int64_t max_row_nnz, slq_stream stream, slq_op** out) {
  SLQ_REQUIRE(out != nullptr, "out is NULL");
  SLQ_REQUIRE(n > 0 && nnz >= 0, "invalid shape n=%lld nnz=%lld", (long long)n, (long long)nnz);
  SLQ_REQUIRE(nnz < (int64_t)INT32_MAX, "nnz=%lld does not fit int32 row pointers", (long long)nnz);
  SLQ_REQUIRE(n < (int64_t)INT32_MAX, "n=%lld does not fit int32 column indices", (long long)n);
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  SLQ_REQUIRE(rowptr_dev && (nnz == 0 || (colidx_dev && vals_dev)), "NULL CSR array");
  int mx = INT32_MAX;  // unknown => the step-A variant that can share long rows
  if (max_row_nnz >= 0) {
    mx = (int)(max_row_nnz < (int64_t)INT32_MAX ? max_row_nnz : (int64_t)INT32_MAX);  // trusted: no device work at all
  } else if (sm_count() > 0) {
    // computed on the caller's stream (the CSR arrays may still be in flight on it) and read back: this is the one
    // documented synchronisation of the library -- pass max_row_nnz >= 0 to avoid it
    cudaStream_t st = (cudaStream_t)stream;
    int* d_mx = nullptr;
    SLQ_CUDA(cudaMalloc(&d_mx, sizeof(int)));
    cudaError_t e = cudaMemsetAsync(d_mx, 0, sizeof(int), st);
    if (e == cudaSuccess) {
      const int g = (int)((n + NT - 1) / NT < 4096 ? (n + NT - 1) / NT : 4096);
      k_max_row_nnz<<<g, NT, 0, st>>>(n, rowptr_dev, d_mx);
      count_launch();
      int h = 0;
      e = cudaMemcpyAsync(&h, d_mx, sizeof(int), cudaMemcpyDeviceToHost, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e == cudaSuccess) mx = h;
    }
    cudaFree(d_mx);
    if (e != cudaSuccess) return fail(SLQ_ERR_CUDA, "slq_csr_create_ex: %s", cudaGetErrorString(e));
  }
  slq_op* op = new slq_op();
  op->kind = OP_CSR;
  op->dtype = dtype;
  op->n = n;
  op->nnz = nnz;
  op->rowptr = rowptr_dev;
  op->colidx = colidx_dev;
  op->vals = vals_dev;
  op->max_row_nnz = mx;
  *out = op;
  return SLQ_OK;
}

int slq_csr_create(int64_t n, int64_t nnz, const int32_t* rowptr_dev, const int32_t* colidx_dev, const void* vals_dev, int dtype,
                   slq_op** out) {
  return slq_csr_create_ex(n, nnz, rowptr_dev, colidx_dev, vals_dev, dtype, -1, nullptr, out);
}

int slq_dense_create(int64_t n, const void* A_dev, int64_t ld, int dtype, slq_op** out) {
  SLQ_REQUIRE(out != nullptr, "out is NULL");
  SLQ_REQUIRE(n > 0 && ld >= n && A_dev, "invalid dense operator (n=%lld, ld=%lld)", (long long)n, (long long)ld);
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  slq_op* op = new slq_op();
  op->kind = OP_DENSE;
  op->dtype = dtype;
  op->n = n;
  op->nnz = n * n;
  op->dense = A_dev;
  op->ld = ld;
  *out = op;
  return SLQ_OK;
}

int slq_dense_create_padded(int64_t n, int64_t npad, const void* A_dev, int64_t ld, int dtype, slq_op** out) {
  SLQ_REQUIRE(npad >= n && npad % 16 == 0 && npad - n < 16 && ld >= npad, "npad must be n rounded up to a multiple of 16 and ld >= npad");
  int rc = slq_dense_create(n, A_dev, ld, dtype, out);
  if (rc == SLQ_OK) (*out)->npad = npad;
  return rc;
}

int slq_op_set_fused_step(slq_op* op, int enable) {
  SLQ_REQUIRE(op != nullptr, "op is NULL");
  op->no_fused_step = enable ? 0 : 1;
  return SLQ_OK;
}

int slq_op_set_scratch(slq_op* op, void* scratch_dev, size_t bytes) {
  SLQ_REQUIRE(op != nullptr, "op is NULL");
  SLQ_REQUIRE(((uintptr_t)scratch_dev & 31) == 0, "scratch must be 32-byte aligned");
  op->scratch = scratch_dev;
  op->scratch_bytes = scratch_dev ? bytes : 0;
  return SLQ_OK;
}

namespace slq {
// one CTA per (chunk, band): the 16 x TRP stage image of the tensor-core kernel, read from 16 contiguous matrix rows
__global__ void __launch_bounds__(256) k_dense_tile(const double* __restrict__ A, int64_t ld, int64_t npad, int tr, double* __restrict__ out) {
  const int trp = tr + 4;
  const int64_t c = blockIdx.x, b = blockIdx.y;
  const int64_t row0 = b * tr;
  double* dst = out + ((size_t)b * (size_t)(npad / DENSE_KC) + (size_t)c) * (size_t)(DENSE_KC * trp);
  for (int idx = threadIdx.x; idx < DENSE_KC * trp; idx += 256) {
    const int kk = idx / trp, r = idx % trp;
    double v = 0.0;
    if (r < tr && row0 + r < npad) v = A[(size_t)(c * DENSE_KC + kk) * ld + row0 + r];
    dst[idx] = v;
  }
}

static bool dense_tiling_applies(const slq_op* op) {
  if (!op || op->kind != OP_DENSE || op->dtype != SLQ_F64 || !dense_use_dmma()) return false;
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  return npad % DENSE_KC == 0 && npad >= DENSE_KC && npad < (int64_t)INT32_MAX && op->ld >= npad;
}
}  // namespace slq

int slq_dense_tiled_bytes(const slq_op* op, size_t* bytes) {
  SLQ_REQUIRE(op != nullptr && bytes != nullptr, "NULL argument");
  *bytes = 0;
  if (!slq::dense_tiling_applies(op)) return SLQ_OK;
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  *bytes = slq::dense_tiled_elems(npad, slq::dense_pick_rpt(npad, true)) * sizeof(double);
  return SLQ_OK;
}

int slq_dense_set_tiled(slq_op* op, void* buf_dev, size_t bytes, slq_stream stream) {
  SLQ_REQUIRE(op != nullptr, "op is NULL");
  if (!buf_dev) {  // detach
    op->tiled = nullptr;
    op->tiled_rpt = 0;
    return SLQ_OK;
  }
  SLQ_REQUIRE(slq::dense_tiling_applies(op), "the band-major copy applies to fp64 dense operators whose (padded) order is a multiple of 16");
  SLQ_REQUIRE(((uintptr_t)buf_dev & 127) == 0, "buffer must be 128-byte aligned");
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  const int rpt = slq::dense_pick_rpt(npad, true);
  const size_t need = slq::dense_tiled_elems(npad, rpt) * sizeof(double);
  SLQ_REQUIRE(bytes >= need, "buffer too small: %zu bytes needed (slq_dense_tiled_bytes)", need);
  const int64_t tr = 16 * rpt;
  dim3 grid((unsigned)(npad / slq::DENSE_KC), (unsigned)((npad + tr - 1) / tr));
  slq::k_dense_tile<<<grid, 256, 0, (cudaStream_t)stream>>>((const double*)op->dense, op->ld, npad, (int)tr, (double*)buf_dev);
  SLQ_LAUNCH_CHECK();
  op->tiled = buf_dev;
  op->tiled_rpt = rpt;
  return SLQ_OK;
}

int slq_callback_create(int64_t n, int dtype, slq_matmat_fn fn, void* ctx, slq_op** out) {
  SLQ_REQUIRE(out != nullptr, "out is NULL");
  SLQ_REQUIRE(n > 0, "invalid shape");
  SLQ_REQUIRE(fn != nullptr, "Supplied object is missing 'matvec' attribute.");  // pylinop.h:24
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  slq_op* op = new slq_op();
  op->kind = OP_CALLBACK;
  op->dtype = dtype;
  op->n = n;
  op->nnz = 0;
  op->fn = fn;
  op->ctx = ctx;
  *out = op;
  return SLQ_OK;
}

int slq_op_destroy(slq_op* op) {
  delete op;
  return SLQ_OK;
}
int64_t slq_op_rows(const slq_op* op) { return op ? op->n : -1; }
int slq_op_dtype(const slq_op* op) { return op ? op->dtype : -1; }
int64_t slq_op_nnz(const slq_op* op) { return op ? op->nnz : -1; }

int slq_op_matmat(slq_op* op, const void* X_dev, void* Y_dev, int k, slq_stream stream) {
  SLQ_REQUIRE(op && X_dev && Y_dev, "NULL argument");
  SLQ_REQUIRE(X_dev != Y_dev, "X and Y must not alias");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  cudaStream_t st = (cudaStream_t)stream;
  return op->dtype == SLQ_F64 ? matmat_dispatch<double>(op, X_dev, Y_dev, k, st) : matmat_dispatch<float>(op, X_dev, Y_dev, k, st);
}

size_t slq_lanczos_workspace_bytes(int64_t n, int k, int ncv, int dtype) {
  if (n <= 0 || k <= 0 || ncv < 2) return 0;
  return ws_layout(n, k, ncv, dtype, nullptr, nullptr);
}

int slq_lanczos(slq_op* op, void* X_dev, int k, int deg, double rtol, int orth, int ncv, void* alpha_dev, void* beta_dev,
                int32_t* steps_dev, double* xnorm2_dev, void* Q_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(op && X_dev && alpha_dev && beta_dev && workspace_dev, "NULL argument");
  SLQ_REQUIRE(deg > 0, "Number of steps must be positive!");  // lanczos.py:81
  SLQ_REQUIRE(ncv >= 2, "ncv must be at least 2 (lanczos.h:101)");
  SLQ_REQUIRE(orth >= 0, "orth must be non-negative at the C boundary (resolve the policy on the host)");
  SLQ_REQUIRE(is_pow2(k) && k <= MAXK, "block width k=%d must be a power of two in [1,64]", k);
  SLQ_REQUIRE(((uintptr_t)X_dev & 31) == 0 && ((uintptr_t)workspace_dev & 255) == 0, "X must be 32-byte and workspace 256-byte aligned");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  size_t need = ws_layout(op->n, k, ncv, op->dtype, nullptr, nullptr);
  SLQ_REQUIRE(workspace_bytes >= need, "workspace too small: %zu < %zu", workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  return op->dtype == SLQ_F64
             ? lanczos_dispatch<double>(op, X_dev, k, deg, rtol, orth, ncv, alpha_dev, beta_dev, steps_dev, xnorm2_dev, Q_dev,
                                        (char*)workspace_dev, st)
             : lanczos_dispatch<float>(op, X_dev, k, deg, rtol, orth, ncv, alpha_dev, beta_dev, steps_dev, xnorm2_dev, Q_dev,
                                       (char*)workspace_dev, st);
}

// ---- basis-free f(A) X (two-pass Lanczos) -------------------------------------------------------------------------------
namespace {
struct ReplayWs {
  size_t panel, table_off, y0_off, total;
};
ReplayWs replay_layout(int64_t n, int k, int deg, int dtype) {
  ReplayWs w;
  w.panel = align_up((size_t)n * k * dtype_size(dtype), 256);
  w.table_off = 3 * w.panel;
  w.y0_off = w.table_off + align_up(sizeof(double) * 4 * MAXK * (size_t)(deg > 0 ? deg : 1), 256);
  w.total = w.y0_off + align_up(sizeof(double) * MAXK, 256);
  return w;
}
struct FunmWs {
  size_t panel, lz_off, table_off, y0_off, alpha_off, beta_off, steps_off, xn_off, coef_off, theta_off, tau_off, scr_off, info_off, total;
};
FunmWs funm_layout(int64_t n, int k, int deg, int dtype) {
  FunmWs w;
  const size_t s = dtype_size(dtype);
  w.panel = align_up((size_t)n * k * s, 256);
  size_t off = w.panel;  // [0, panel): clobberable copy of X for the first pass = replay panel 0
  w.lz_off = off;        // Lanczos workspace (ncv = 2): its two panels double as replay panels 1 and 2
  off += align_up(ws_layout(n, k, 2, dtype, nullptr, nullptr), 256);
  w.table_off = off;
  off += align_up(sizeof(double) * 4 * MAXK * (size_t)deg, 256);
  w.y0_off = off;
  off += align_up(sizeof(double) * MAXK, 256);
  w.alpha_off = off;
  off += align_up(s * (size_t)k * (deg + 1), 256);
  w.beta_off = off;
  off += align_up(s * (size_t)k * (deg + 1), 256);
  w.steps_off = off;
  off += align_up(sizeof(int32_t) * MAXK, 256);
  w.xn_off = off;
  off += align_up(sizeof(double) * MAXK, 256);
  w.coef_off = off;
  off += align_up(sizeof(double) * (size_t)k * deg, 256);
  w.theta_off = off;
  off += align_up(sizeof(double) * (size_t)k * deg, 256);
  w.tau_off = off;
  off += align_up(sizeof(double) * (size_t)k * deg, 256);
  w.scr_off = off;
  off += align_up(sizeof(double) * ((size_t)deg * deg + deg) * k, 256);
  w.info_off = off;
  off += align_up(sizeof(int32_t) * MAXK, 256);
  w.total = off;
  return w;
}
}  // namespace

size_t slq_replay_workspace_bytes(int64_t n, int k, int deg, int dtype) {
  if (n <= 0 || k <= 0 || deg <= 0) return 0;
  return replay_layout(n, k, deg, dtype).total;
}

int slq_lanczos_replay(slq_op* op, const void* X_dev, int k, int deg, const void* alpha_dev, const void* beta_dev,
                       const int32_t* steps_dev, const double* xnorm2_dev, const double* coeffs_dev, void* Y_dev, void* workspace_dev,
                       size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(op && X_dev && alpha_dev && beta_dev && steps_dev && xnorm2_dev && coeffs_dev && Y_dev && workspace_dev, "NULL argument");
  SLQ_REQUIRE(deg > 0, "Number of steps must be positive!");
  SLQ_REQUIRE(is_pow2(k) && k <= MAXK, "block width k=%d must be a power of two in [1,64]", k);
  SLQ_REQUIRE(X_dev != Y_dev, "X and Y must not alias");
  SLQ_REQUIRE((((uintptr_t)X_dev | (uintptr_t)Y_dev) & 31) == 0 && ((uintptr_t)workspace_dev & 255) == 0,
              "X and Y must be 32-byte and workspace 256-byte aligned");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const ReplayWs w = replay_layout(op->n, k, deg, op->dtype);
  SLQ_REQUIRE(workspace_bytes >= w.total, "workspace too small: %zu < %zu", workspace_bytes, w.total);
  char* base = (char*)workspace_dev;
  double* table = (double*)(base + w.table_off);
  double* y0 = (double*)(base + w.y0_off);
  cudaStream_t st = (cudaStream_t)stream;
  return op->dtype == SLQ_F64
             ? replay_dispatch<double>(op, X_dev, k, deg, alpha_dev, beta_dev, steps_dev, xnorm2_dev, coeffs_dev, Y_dev, table, y0, base,
                                       base + w.panel, base + 2 * w.panel, st)
             : replay_dispatch<float>(op, X_dev, k, deg, alpha_dev, beta_dev, steps_dev, xnorm2_dev, coeffs_dev, Y_dev, table, y0, base,
                                      base + w.panel, base + 2 * w.panel, st);
}

size_t slq_funm_workspace_bytes(int64_t n, int k, int deg, int dtype) {
  if (n <= 0 || k <= 0 || deg <= 0) return 0;
  return funm_layout(n, k, deg, dtype).total;
}

int slq_funm_matmat(slq_op* op, const void* X_dev, int k, int deg, double rtol, int fun, const double* params, void* Y_dev,
                    void* workspace_dev, size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(op && X_dev && Y_dev && workspace_dev, "NULL argument");
  SLQ_REQUIRE(deg > 0, "Number of steps must be positive!");
  SLQ_REQUIRE(is_pow2(k) && k <= MAXK, "block width k=%d must be a power of two in [1,64]", k);
  SLQ_REQUIRE(X_dev != Y_dev, "X and Y must not alias");
  SLQ_REQUIRE(fun >= SLQ_FUN_IDENTITY && fun <= SLQ_FUN_SOFTSIGN, "Unknown function: %d.", fun);
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const FunmWs w = funm_layout(op->n, k, deg, op->dtype);
  SLQ_REQUIRE(workspace_bytes >= w.total, "workspace too small: %zu < %zu", workspace_bytes, w.total);
  SLQ_REQUIRE(((uintptr_t)workspace_dev & 255) == 0, "workspace must be 256-byte aligned");
  char* base = (char*)workspace_dev;
  cudaStream_t st = (cudaStream_t)stream;
  void* xcopy = base;
  void* alpha = base + w.alpha_off;
  void* beta = base + w.beta_off;
  int32_t* steps = (int32_t*)(base + w.steps_off);
  double* xn = (double*)(base + w.xn_off);
  double* coeffs = (double*)(base + w.coef_off);
  // pass 1: the orth = 0 recurrence on a copy of X (the driver clobbers its start block), keeping alpha / beta only
  SLQ_CUDA(cudaMemcpyAsync(xcopy, X_dev, (size_t)op->n * k * dtype_size(op->dtype), cudaMemcpyDeviceToDevice, st));
  int rc = slq_lanczos(op, xcopy, k, deg, rtol, 0, 2, alpha, beta, steps, xn, nullptr, base + w.lz_off,
                       ws_layout(op->n, k, 2, op->dtype, nullptr, nullptr), stream);
  if (rc != SLQ_OK) return rc;
  // c = Y f(Theta) Y^T e_1 per probe (operators.py:122-124)
  rc = slq_funm_coeffs(alpha, beta, op->dtype, k, deg, deg + 1, fun, params, coeffs, (double*)(base + w.theta_off),
                       (double*)(base + w.tau_off), (double*)(base + w.scr_off), (int32_t*)(base + w.info_off), stream);
  if (rc != SLQ_OK) return rc;
  // pass 2: the same recurrence from the intact X with the recorded coefficients, Y accumulated on the fly
  double* table = (double*)(base + w.table_off);
  double* y0 = (double*)(base + w.y0_off);
  char* lzp = base + w.lz_off;  // ws_layout places the panels first
  return op->dtype == SLQ_F64
             ? replay_dispatch<double>(op, X_dev, k, deg, alpha, beta, steps, xn, coeffs, Y_dev, table, y0, xcopy, lzp, lzp + w.panel, st)
             : replay_dispatch<float>(op, X_dev, k, deg, alpha, beta, steps, xn, coeffs, Y_dev, table, y0, xcopy, lzp, lzp + w.panel, st);
}

int slq_block_from_columns(const void* cols_dev, int64_t n, int k, int dtype, void* X_dev, slq_stream stream) {
  SLQ_REQUIRE(cols_dev && X_dev && n > 0 && k > 0, "invalid argument");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  dim3 grid((unsigned)((n + 31) / 32), (unsigned)((k + 31) / 32)), block(32, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLQ_F64)
    k_cols_to_block<double><<<grid, block, 0, st>>>((const double*)cols_dev, n, k, (double*)X_dev);
  else
    k_cols_to_block<float><<<grid, block, 0, st>>>((const float*)cols_dev, n, k, (float*)X_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

int slq_block_to_columns(const void* X_dev, int64_t n, int k, int dtype, void* cols_dev, slq_stream stream) {
  SLQ_REQUIRE(cols_dev && X_dev && n > 0 && k > 0, "invalid argument");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  dim3 grid((unsigned)((n + 31) / 32), (unsigned)((k + 31) / 32)), block(32, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLQ_F64)
    k_block_to_cols<double><<<grid, block, 0, st>>>((const double*)X_dev, n, k, (double*)cols_dev);
  else
    k_block_to_cols<float><<<grid, block, 0, st>>>((const float*)X_dev, n, k, (float*)cols_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // extern "C"
