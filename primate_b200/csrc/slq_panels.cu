// Tall-skinny panel algebra for the block drivers (hutch++, XDiag, XTrace on a MatrixFunction: SURVEY 8f rank 1).
//
// The reference does this part with NumPy/LAPACK on host arrays -- np.linalg.qr of the n x m sketch, Q.T @ G, Q @ S,
// np.einsum row/column dots (src/primate/trace.py:158-175,201-227; diagonal.py:115-137).  Here the n x m panels never
// leave the GPU and every O(n m^2) / O(n m) pass over them is one of four kernels; only m x m matrices (Gram matrices,
// Cholesky factors, triangular inverses: kilobytes) travel to the host.  Thin QR is CholeskyQR2 built from them:
//     G = Y^T Y  (slq_panel_gram, row range per rank + one all-reduce of m^2 doubles)  ->  R1 = chol(G)^T on the host
//     Q1 = Y R1^-1 (slq_panel_matmul)   ->  G2 = Q1^T Q1, R2 = chol(G2)^T, Q = Q1 R2^-1, R = R2 R1.
// Panels are row-major with a row stride (ld), so column ranges of a wider buffer are panels too.  All reductions are
// deterministic (per-CTA partial tables summed in a fixed order), accumulate in fp64 and return fp64.
#include <algorithm>

#include "slq_internal.cuh"

namespace slq {

constexpr int PG_T = 16;   // thread grid 16 x 16
constexpr int PG_TILE = 128;  // output tile (rows of G x columns of G) per CTA: 8 x 8 per thread
constexpr int PG_RC = 16;  // panel rows staged per iteration

// partial[cta][tile][i][j] += sum over this CTA's rows of X[r][i0+i] * Y[r][j0+j]
template <typename T>
__global__ void __launch_bounds__(256) k_panel_gram(const T* __restrict__ X, int a, int64_t ldx, const T* __restrict__ Y, int b, int64_t ldy,
                                                     int64_t row0, int64_t row1, double* __restrict__ partial) {
  __shared__ double sX[PG_RC][PG_TILE + 1];
  __shared__ double sY[PG_RC][PG_TILE + 1];
  const int tiles_j = (b + PG_TILE - 1) / PG_TILE;
  const int i0 = (blockIdx.y / tiles_j) * PG_TILE, j0 = (blockIdx.y % tiles_j) * PG_TILE;
  const int ta = min(PG_TILE, a - i0), tb = min(PG_TILE, b - j0);
  const int tx = threadIdx.x % PG_T, ty = threadIdx.x / PG_T;
  double acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0;
  const int64_t nrows = row1 - row0;
  const int64_t chunks = (nrows + PG_RC - 1) / PG_RC;
  for (int64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
    const int64_t r0 = row0 + c * PG_RC;
    const int rc = (int)min((int64_t)PG_RC, row1 - r0);
    for (int idx = threadIdx.x; idx < PG_RC * ta; idx += 256) {
      const int r = idx / ta, i = idx % ta;
      sX[r][i] = (r < rc) ? (double)X[(r0 + r) * ldx + i0 + i] : 0.0;
    }
    for (int idx = threadIdx.x; idx < PG_RC * tb; idx += 256) {
      const int r = idx / tb, j = idx % tb;
      sY[r][j] = (r < rc) ? (double)Y[(r0 + r) * ldy + j0 + j] : 0.0;
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < PG_RC; ++r) {
      double xi[8], yj[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) xi[i] = sX[r][ty + PG_T * i];
#pragma unroll
      for (int j = 0; j < 8; ++j) yj[j] = sY[r][tx + PG_T * j];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fma(xi[i], yj[j], acc[i][j]);
    }
    __syncthreads();
  }
  // (columns beyond ta / tb read stale shared memory: those accumulators are simply never stored)
  double* out = partial + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * (PG_TILE * PG_TILE);
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ii = ty + PG_T * i, jj = tx + PG_T * j;
      if (ii < ta && jj < tb) out[ii * PG_TILE + jj] = acc[i][j];
    }
}

__global__ void k_panel_gram_reduce(const double* __restrict__ partial, int nparts, int a, int b, double* __restrict__ G) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a * b) return;
  const int i = idx / b, j = idx % b;
  const int tiles_j = (b + PG_TILE - 1) / PG_TILE, tiles_i = (a + PG_TILE - 1) / PG_TILE;
  const int tile = (i / PG_TILE) * tiles_j + (j / PG_TILE);
  const int ntiles = tiles_i * tiles_j;
  double s = 0.0;
  for (int p = 0; p < nparts; ++p) s += partial[((size_t)p * ntiles + tile) * (PG_TILE * PG_TILE) + (i % PG_TILE) * PG_TILE + (j % PG_TILE)];
  G[idx] = s;
}

// Y[r][j] = alpha * sum_i X[r][i] S[i][j] + beta * Y[r][j]     (S: a x b doubles, row-major)
constexpr int PM_RC = 32, PM_KC = 32, PM_BT = 128;
template <typename T>
__global__ void __launch_bounds__(256) k_panel_matmul(const T* __restrict__ X, int a, int64_t ldx, const double* __restrict__ S, int b,
                                                       double alpha, double beta, T* __restrict__ Y, int64_t ldy, int64_t n) {
  __shared__ double sX[PM_RC][PM_KC + 1];
  __shared__ double sS[PM_KC][PM_BT];
  const int j0 = blockIdx.y * PM_BT;
  const int tb = min(PM_BT, b - j0);
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;  // rows 2 ty, 2 ty + 1; columns tx + 16 jj
  const int64_t chunks = (n + PM_RC - 1) / PM_RC;
  for (int64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
    const int64_t r0 = c * PM_RC;
    const int rc = (int)min((int64_t)PM_RC, n - r0);
    double acc[2][8];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.0;
    for (int k0 = 0; k0 < a; k0 += PM_KC) {
      const int kc = min(PM_KC, a - k0);
      for (int idx = threadIdx.x; idx < PM_RC * PM_KC; idx += 256) {
        const int r = idx / PM_KC, kk = idx % PM_KC;
        sX[r][kk] = (r < rc && kk < kc) ? (double)X[(r0 + r) * ldx + k0 + kk] : 0.0;
      }
      for (int idx = threadIdx.x; idx < PM_KC * PM_BT; idx += 256) {
        const int kk = idx / PM_BT, j = idx % PM_BT;
        sS[kk][j] = (kk < kc && j < tb) ? S[(size_t)(k0 + kk) * b + j0 + j] : 0.0;
      }
      __syncthreads();
#pragma unroll 8
      for (int kk = 0; kk < PM_KC; ++kk) {
        const double x0 = sX[2 * ty][kk], x1 = sX[2 * ty + 1][kk];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const double sv = sS[kk][tx + 16 * j];
          acc[0][j] = fma(x0, sv, acc[0][j]);
          acc[1][j] = fma(x1, sv, acc[1][j]);
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int r = 2 * ty + i;
      if (r < rc) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int jj = tx + 16 * j;
          if (jj < tb) {
            T* y = Y + (r0 + r) * ldy + j0 + jj;
            const double old = (beta != 0.0) ? beta * (double)*y : 0.0;
            *y = (T)(alpha * acc[i][j] + old);
          }
        }
      }
    }
  }
}

// out[r] = sum_j X[r][j] * Y[r][j]   (one warp per row)
template <typename T>
__global__ void __launch_bounds__(256) k_panel_rowdot(const T* __restrict__ X, int64_t ldx, const T* __restrict__ Y, int64_t ldy, int m,
                                                       int64_t n, double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    double s = 0.0;
    for (int j = lane; j < m; j += 32) s = fma((double)X[r * ldx + j], (double)Y[r * ldy + j], s);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0) out[r] = s;
  }
}

// partial[cta][j] = sum over the CTA's rows of X[r][j] * Y[r][j]
template <typename T>
__global__ void __launch_bounds__(256) k_panel_coldot(const T* __restrict__ X, int64_t ldx, const T* __restrict__ Y, int64_t ldy, int m,
                                                       int64_t row0, int64_t row1, double* __restrict__ partial) {
  __shared__ double s_part[256];
  const int cw = m < 256 ? ((m + 31) / 32) * 32 : 256;  // column lanes
  const int rl = 256 / cw;                               // row lanes
  const int c = threadIdx.x % cw, l = threadIdx.x / cw;
  const int64_t nrows = row1 - row0;
  const int64_t per = (nrows + gridDim.x - 1) / gridDim.x;
  const int64_t a = row0 + (int64_t)blockIdx.x * per, b = min(row1, a + per);
  for (int j0 = 0; j0 < m; j0 += cw) {
    const int j = j0 + c;
    double s = 0.0;
    if (j < m && l < rl)
      for (int64_t r = a + l; r < b; r += rl) s = fma((double)X[r * ldx + j], (double)Y[r * ldy + j], s);
    s_part[threadIdx.x] = s;
    __syncthreads();
    if (l == 0 && j < m) {
      double t = 0.0;
      for (int q = 0; q < rl; ++q) t += s_part[q * cw + c];
      partial[(size_t)blockIdx.x * m + j] = t;
    }
    __syncthreads();
  }
}

__global__ void k_coldot_reduce(const double* __restrict__ partial, int nparts, int m, double* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  double s = 0.0;
  for (int p = 0; p < nparts; ++p) s += partial[(size_t)p * m + j];
  out[j] = s;
}

static int gram_grid(int64_t nrows) {
  const int sms = sm_count() > 0 ? sm_count() : 148;
  const int64_t chunks = (nrows + PG_RC - 1) / PG_RC;
  return (int)std::max<int64_t>(1, std::min<int64_t>(chunks, 2 * (int64_t)sms));
}

}  // namespace slq

using namespace slq;

extern "C" {

size_t slq_panel_gram_workspace_bytes(int a, int b) {
  if (a <= 0 || b <= 0) return 0;
  const int sms = sm_count() > 0 ? sm_count() : 148;
  const size_t tiles = (size_t)((a + PG_TILE - 1) / PG_TILE) * ((b + PG_TILE - 1) / PG_TILE);
  return align_up(sizeof(double) * 2 * (size_t)sms * tiles * PG_TILE * PG_TILE, 256);
}

int slq_panel_gram(const void* X_dev, int a, int64_t ldx, const void* Y_dev, int b, int64_t ldy, int64_t row0, int64_t row1, int dtype,
                   double* G_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(X_dev && Y_dev && G_dev && workspace_dev, "NULL argument");
  SLQ_REQUIRE(a > 0 && b > 0 && ldx >= a && ldy >= b && row0 >= 0 && row1 >= row0, "invalid panel shape");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  SLQ_REQUIRE(workspace_bytes >= slq_panel_gram_workspace_bytes(a, b), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (row1 == row0) {
    SLQ_CUDA(cudaMemsetAsync(G_dev, 0, sizeof(double) * (size_t)a * b, st));
    return SLQ_OK;
  }
  const int tiles = ((a + PG_TILE - 1) / PG_TILE) * ((b + PG_TILE - 1) / PG_TILE);
  const int gx = gram_grid(row1 - row0);
  dim3 grid((unsigned)gx, (unsigned)tiles);
  double* partial = (double*)workspace_dev;
  if (dtype == SLQ_F64)
    k_panel_gram<double><<<grid, 256, 0, st>>>((const double*)X_dev, a, ldx, (const double*)Y_dev, b, ldy, row0, row1, partial);
  else
    k_panel_gram<float><<<grid, 256, 0, st>>>((const float*)X_dev, a, ldx, (const float*)Y_dev, b, ldy, row0, row1, partial);
  SLQ_LAUNCH_CHECK();
  k_panel_gram_reduce<<<(a * b + 255) / 256, 256, 0, st>>>(partial, gx, a, b, G_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

int slq_panel_matmul(const void* X_dev, int a, int64_t ldx, const double* S_dev, int b, double alpha, double beta, void* Y_dev,
                     int64_t ldy, int64_t n, int dtype, slq_stream stream) {
  SLQ_REQUIRE(X_dev && S_dev && Y_dev, "NULL argument");
  SLQ_REQUIRE(a > 0 && b > 0 && n > 0 && ldx >= a && ldy >= b, "invalid panel shape");
  SLQ_REQUIRE(X_dev != Y_dev, "X and Y must not alias");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  const int sms = sm_count();
  if (sms <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const int64_t chunks = (n + PM_RC - 1) / PM_RC;
  dim3 grid((unsigned)std::min<int64_t>(chunks, 4 * (int64_t)sms), (unsigned)((b + PM_BT - 1) / PM_BT));
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLQ_F64)
    k_panel_matmul<double><<<grid, 256, 0, st>>>((const double*)X_dev, a, ldx, S_dev, b, alpha, beta, (double*)Y_dev, ldy, n);
  else
    k_panel_matmul<float><<<grid, 256, 0, st>>>((const float*)X_dev, a, ldx, S_dev, b, alpha, beta, (float*)Y_dev, ldy, n);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

int slq_panel_rowdot(const void* X_dev, int64_t ldx, const void* Y_dev, int64_t ldy, int m, int64_t n, int dtype, double* out_dev,
                     slq_stream stream) {
  SLQ_REQUIRE(X_dev && Y_dev && out_dev && m > 0 && n > 0 && ldx >= m && ldy >= m, "invalid argument");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  const int sms = sm_count();
  if (sms <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const unsigned g = (unsigned)std::min<int64_t>((n + 7) / 8, 16 * (int64_t)sms);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLQ_F64)
    k_panel_rowdot<double><<<g, 256, 0, st>>>((const double*)X_dev, ldx, (const double*)Y_dev, ldy, m, n, out_dev);
  else
    k_panel_rowdot<float><<<g, 256, 0, st>>>((const float*)X_dev, ldx, (const float*)Y_dev, ldy, m, n, out_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

size_t slq_panel_coldot_workspace_bytes(int m) {
  const int sms = sm_count() > 0 ? sm_count() : 148;
  return m > 0 ? align_up(sizeof(double) * 2 * (size_t)sms * m, 256) : 0;
}

int slq_panel_coldot(const void* X_dev, int64_t ldx, const void* Y_dev, int64_t ldy, int m, int64_t row0, int64_t row1, int dtype,
                     double* out_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(X_dev && Y_dev && out_dev && workspace_dev && m > 0 && row1 >= row0 && ldx >= m && ldy >= m, "invalid argument");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  const int sms = sm_count();
  if (sms <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  SLQ_REQUIRE(workspace_bytes >= slq_panel_coldot_workspace_bytes(m), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (row1 == row0) {
    SLQ_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double) * m, st));
    return SLQ_OK;
  }
  const int g = (int)std::max<int64_t>(1, std::min<int64_t>((row1 - row0 + 63) / 64, 2 * (int64_t)sms));
  double* partial = (double*)workspace_dev;
  if (dtype == SLQ_F64)
    k_panel_coldot<double><<<g, 256, 0, st>>>((const double*)X_dev, ldx, (const double*)Y_dev, ldy, m, row0, row1, partial);
  else
    k_panel_coldot<float><<<g, 256, 0, st>>>((const float*)X_dev, ldx, (const float*)Y_dev, ldy, m, row0, row1, partial);
  SLQ_LAUNCH_CHECK();
  k_coldot_reduce<<<(m + 127) / 128, 128, 0, st>>>(partial, g, m, out_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // extern "C"
