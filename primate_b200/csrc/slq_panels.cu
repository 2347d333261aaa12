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
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "slq_internal.cuh"

namespace slq {

__device__ __forceinline__ void panel_dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

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

// fp64 tensor-core variant of the Gram-type reduction (a % 8 == 0, b % 8 == 0, a, b <= 96): G = X^T Y is an (a x b)
// sum of rank-4 updates, one DMMA m8n8k4 per 8 x 8 output tile and 4 panel rows.  The 8 warps of a CTA form a
// WI x WJ grid over the output tiles times a WR-way split of the rows; the A / B fragments (4 rows x 8 columns of X / Y,
// one double per lane, 64-byte row segments) come straight from global memory.  Every (CTA, row split) pair writes
// its partial tiles to the partial table; the reduce kernel sums them in a fixed order.
template <int TI, int TJ, int WI, int WJ>
__global__ void __launch_bounds__(256) k_panel_gram_dmma(const double* __restrict__ X, int a, int64_t ldx, const double* __restrict__ Y,
                                                          int b, int64_t ldy, int64_t row0, int64_t row1, double* __restrict__ partial) {
  constexpr int WR = 8 / (WI * WJ);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int wi = warp % WI, wj = (warp / WI) % WJ, wr = warp / (WI * WJ);
  const int ib0 = wi * TI, jb0 = wj * TJ;  // first 8-blocks of this warp's output rectangle
  double acc[TI][TJ][2];
#pragma unroll
  for (int i = 0; i < TI; ++i)
#pragma unroll
    for (int j = 0; j < TJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  // rows in groups of 4; group index q = (CTA-strided chunk) interleaved over the WR row splits
  const int64_t groups = (row1 - row0 + 3) / 4;
  for (int64_t q = (int64_t)blockIdx.x * WR + wr; q < groups; q += (int64_t)gridDim.x * WR) {
    const int64_t r = row0 + q * 4 + t;
    const bool ok = r < row1;
    const double* xr = X + (ok ? r : row0) * ldx + g;
    const double* yr = Y + (ok ? r : row0) * ldy + g;
    double av[TI], bv[TJ];
#pragma unroll
    for (int i = 0; i < TI; ++i) av[i] = (ok && (ib0 + i) * 8 < a) ? __ldg(xr + (ib0 + i) * 8) : 0.0;
#pragma unroll
    for (int j = 0; j < TJ; ++j) bv[j] = (ok && (jb0 + j) * 8 < b) ? __ldg(yr + (jb0 + j) * 8) : 0.0;
#pragma unroll
    for (int i = 0; i < TI; ++i)
#pragma unroll
      for (int j = 0; j < TJ; ++j) panel_dmma(acc[i][j], av[i], bv[j]);
  }
  // partial table entry of (CTA, row split): an a x b row-major matrix
  double* out = partial + ((size_t)blockIdx.x * WR + wr) * ((size_t)a * b);
#pragma unroll
  for (int i = 0; i < TI; ++i)
#pragma unroll
    for (int j = 0; j < TJ; ++j) {
      const int gi = (ib0 + i) * 8 + g, gj = (jb0 + j) * 8 + 2 * t;
      if (gi < a && gj < b) {
        out[(size_t)gi * b + gj] = acc[i][j][0];
        out[(size_t)gi * b + gj + 1] = acc[i][j][1];
      }
    }
}

__global__ void k_panel_gram_reduce_flat(const double* __restrict__ partial, int nparts, int ab, double* __restrict__ G) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= ab) return;
  double s = 0.0;
  for (int p = 0; p < nparts; ++p) s += partial[(size_t)p * ab + idx];
  G[idx] = s;
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

// fp64 tensor-core variant of the panel product (a % 4 == 0, b % 8 == 0, b <= 96 per grid.y tile): every warp owns 16
// rows of Y; the A fragments (8 rows x 4 columns of X: one double per lane, 32-byte row segments) come straight from
// global memory -- each sector of X is read exactly once over the k loop -- the B fragments from S staged in shared
// memory (row stride = 4 (mod 16) doubles: conflict-free), the product accumulates in DMMA m8n8k4 tiles.  The panel
// passes of the block drivers are then HBM-bound (read X, write Y) instead of FMA-issue-bound (measured 2 TFLOP/s with
// the scalar kernel at m = 32: 17 ms per pass at n = 16.7 M against 1.4 ms of memory time).
constexpr int PD_BT = 96;
template <int NB>  // 8-column blocks of this tile
__global__ void __launch_bounds__(256) k_panel_matmul_dmma(const double* __restrict__ X, int a, int64_t ldx, const double* __restrict__ S,
                                                            int b, double alpha, double beta, double* __restrict__ Y, int64_t ldy,
                                                            int64_t n) {
  extern __shared__ double sS[];  // a x (8 NB + 4)
  constexpr int BP = 8 * NB + 4;
  const int j0 = blockIdx.y * PD_BT;
  for (int idx = threadIdx.x; idx < a * 8 * NB; idx += 256) {
    const int kk = idx / (8 * NB), j = idx % (8 * NB);
    sS[kk * BP + j] = (j0 + j < b) ? S[(size_t)kk * b + j0 + j] : 0.0;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int64_t slabs = (n + 15) / 16;
  for (int64_t sl = (int64_t)blockIdx.x * 8 + warp; sl < slabs; sl += (int64_t)gridDim.x * 8) {
    const int64_t r0 = sl * 16;
    double acc[2][NB][2];
#pragma unroll
    for (int mb = 0; mb < 2; ++mb)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) acc[mb][nb][0] = acc[mb][nb][1] = 0.0;
    const int64_t ra = r0 + g, rb = r0 + 8 + g;
    const double* xa = X + (ra < n ? ra : n - 1) * ldx + t;  // rows past the end read the last row (never stored)
    const double* xb = X + (rb < n ? rb : n - 1) * ldx + t;
#pragma unroll 4
    for (int k0 = 0; k0 < a; k0 += 4) {
      const double a0 = __ldg(xa + k0), a1 = __ldg(xb + k0);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        const double bv = sS[(k0 + t) * BP + nb * 8 + g];
        panel_dmma(acc[0][nb], a0, bv);
        panel_dmma(acc[1][nb], a1, bv);
      }
    }
#pragma unroll
    for (int mb = 0; mb < 2; ++mb) {
      const int64_t r = r0 + mb * 8 + g;
      if (r < n) {
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          const int j = j0 + nb * 8 + 2 * t;
          if (j < b) {  // b % 8 == 0: the pair (j, j + 1) is inside or outside together
            double* y = Y + r * ldy + j;
            double o0 = alpha * acc[mb][nb][0], o1 = alpha * acc[mb][nb][1];
            if (beta != 0.0) {
              o0 += beta * y[0];
              o1 += beta * y[1];
            }
            y[0] = o0;
            y[1] = o1;
          }
        }
      }
    }
  }
}

template <int NB>
static int launch_panel_matmul_dmma(const double* X, int a, int64_t ldx, const double* S, int b, double alpha, double beta, double* Y,
                                    int64_t ldy, int64_t n, dim3 grid, cudaStream_t st) {
  const size_t smem = sizeof(double) * (size_t)a * (8 * NB + 4);
  static DeviceOnce once;
  if (smem > 48 * 1024 && once.needed()) {
    SLQ_CUDA(cudaFuncSetAttribute(k_panel_matmul_dmma<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    once.done();
  }
  k_panel_matmul_dmma<NB><<<grid, 256, smem, st>>>(X, a, ldx, S, b, alpha, beta, Y, ldy, n);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
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

// ---- allocation-free column dots (slq_panel_coldot_static) -------------------------------------------------------------
// The per-CTA partial table of k_panel_coldot lives in a file-scope device array instead of a caller workspace: 1 MiB
// that every GPU loading the library holds until the process exits.  In HOST code `g_coldot_static` is only a stand-in
// for the symbol, not a device address: the address is looked up once per device (cudaGetSymbolAddress with that
// device current) and passed to the kernels as an ordinary pointer.
constexpr int COLDOT_STATIC_MAX_M = SLQ_COLDOT_STATIC_MAX_M;  // widest panel the static table serves (include/slq.h)
constexpr int COLDOT_STATIC_MAX_PARTS = 512;  // most CTAs (partial rows) of one call
static __device__ __align__(256) double g_coldot_static[(size_t)COLDOT_STATIC_MAX_PARTS * COLDOT_STATIC_MAX_M];

static int coldot_static_ptr(double** out) {
  static std::atomic<double*> addr[MAX_DEVICES];  // zero-initialised: nullptr = not looked up on that device yet
  const int dev = current_device();
  if (dev < 0 || dev >= MAX_DEVICES) return fail(SLQ_ERR_CUDA, "no current CUDA device (or ordinal >= %d)", MAX_DEVICES);
  double* p = addr[dev].load(std::memory_order_acquire);
  if (!p) {  // two threads may both look it up: same answer
    void* sym = nullptr;
    SLQ_CUDA(cudaGetSymbolAddress(&sym, g_coldot_static));
    if (!sym) return fail(SLQ_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    p = (double*)sym;
    addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return SLQ_OK;
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
  const size_t fma = sizeof(double) * 2 * (size_t)sms * tiles * PG_TILE * PG_TILE;
  const size_t dmma = sizeof(double) * 2 * (size_t)sms * 8 * (size_t)a * b;  // (CTA, row split) partials of the tensor-core variant
  return align_up(fma > dmma ? fma : dmma, 256);
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
  double* partial = (double*)workspace_dev;
  static const bool fma_only = [] { const char* e = getenv("SLQ_PANEL"); return e && strcmp(e, "fma") == 0; }();
  if (dtype == SLQ_F64 && !fma_only && a % 8 == 0 && b % 8 == 0 && a <= 96 && b <= 96) {
    const int sms = sm_count();
    const int64_t groups = (row1 - row0 + 3) / 4;
    const int mx = a > b ? a : b;
    const int wr = mx <= 32 ? 8 : (mx <= 64 ? 4 : 2);  // row splits per CTA (see the warp grids below)
    const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((groups + wr * 64 - 1) / (wr * 64), 2 * (int64_t)sms));
    const double* Xd = (const double*)X_dev;
    const double* Yd = (const double*)Y_dev;
    if (mx <= 32)
      k_panel_gram_dmma<4, 4, 1, 1><<<gx, 256, 0, st>>>(Xd, a, ldx, Yd, b, ldy, row0, row1, partial);
    else if (mx <= 64)
      k_panel_gram_dmma<4, 8, 2, 1><<<gx, 256, 0, st>>>(Xd, a, ldx, Yd, b, ldy, row0, row1, partial);
    else
      k_panel_gram_dmma<6, 6, 2, 2><<<gx, 256, 0, st>>>(Xd, a, ldx, Yd, b, ldy, row0, row1, partial);
    SLQ_LAUNCH_CHECK();
    k_panel_gram_reduce_flat<<<(a * b + 255) / 256, 256, 0, st>>>(partial, gx * wr, a * b, G_dev);
    SLQ_LAUNCH_CHECK();
    return SLQ_OK;
  }
  const int tiles = ((a + PG_TILE - 1) / PG_TILE) * ((b + PG_TILE - 1) / PG_TILE);
  const int gx = gram_grid(row1 - row0);
  dim3 grid((unsigned)gx, (unsigned)tiles);
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
  static const bool fma_only = [] { const char* e = getenv("SLQ_PANEL"); return e && strcmp(e, "fma") == 0; }();
  if (dtype == SLQ_F64 && !fma_only && a % 4 == 0 && b % 8 == 0 && (size_t)a * (PD_BT + 4) * sizeof(double) <= 200 * 1024) {
    // fp64 tensor-core tiles: column tiles of at most 96, 16-row slabs per warp
    const int tiles = (b + PD_BT - 1) / PD_BT;
    const int64_t slabs = (n + 15) / 16;
    dim3 gd((unsigned)std::max<int64_t>(1, std::min<int64_t>((slabs + 7) / 8, 8 * (int64_t)sms)), (unsigned)tiles);
    const double* Xd = (const double*)X_dev;
    double* Yd = (double*)Y_dev;
    const int nb_last = ((b - (tiles - 1) * PD_BT) + 7) / 8;  // 8-column blocks in the last (or only) tile
    if (tiles > 1) {  // full tiles first, then the remainder tile with its own width
      dim3 gfull(gd.x, (unsigned)(tiles - 1));
      int rc = launch_panel_matmul_dmma<12>(Xd, a, ldx, S_dev, b, alpha, beta, Yd, ldy, n, gfull, st);
      if (rc != SLQ_OK) return rc;
    }
    // the remainder tile starts at column (tiles - 1) * 96: shift the pointers, keep S's row stride via the b argument
    const int joff = (tiles - 1) * PD_BT;
    dim3 g1(gd.x, 1);
    auto go = [&](auto tag) {
      constexpr int NBv = decltype(tag)::value;
      // blockIdx.y == 0 in this launch, so the column offset is applied by hand
      return launch_panel_matmul_dmma<NBv>(Xd, a, ldx, S_dev + joff, b, alpha, beta, Yd + joff, ldy, n, g1, st);
    };
    switch (nb_last) {
      case 1: return go(std::integral_constant<int, 1>{});
      case 2: return go(std::integral_constant<int, 2>{});
      case 3: return go(std::integral_constant<int, 3>{});
      case 4: return go(std::integral_constant<int, 4>{});
      case 5: return go(std::integral_constant<int, 5>{});
      case 6: return go(std::integral_constant<int, 6>{});
      case 7: return go(std::integral_constant<int, 7>{});
      case 8: return go(std::integral_constant<int, 8>{});
      case 9: return go(std::integral_constant<int, 9>{});
      case 10: return go(std::integral_constant<int, 10>{});
      case 11: return go(std::integral_constant<int, 11>{});
      default: return go(std::integral_constant<int, 12>{});
    }
  }
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

int slq_panel_coldot_static(const void* X_dev, int64_t ldx, const void* Y_dev, int64_t ldy, int m, int64_t row0, int64_t row1, int dtype,
                            double* out_dev, slq_stream stream) {
  SLQ_REQUIRE(X_dev && Y_dev && out_dev && m > 0 && row1 >= row0 && ldx >= m && ldy >= m, "invalid argument");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  SLQ_REQUIRE(m <= COLDOT_STATIC_MAX_M, "m = %d exceeds the static scratch of slq_panel_coldot_static (at most %d columns); "
              "use slq_panel_coldot with a caller workspace", m, COLDOT_STATIC_MAX_M);
  const int sms = sm_count();
  if (sms <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  cudaStream_t st = (cudaStream_t)stream;
  if (row1 == row0) {
    SLQ_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double) * m, st));
    return SLQ_OK;
  }
  double* partial = nullptr;
  int rc = coldot_static_ptr(&partial);
  if (rc != SLQ_OK) return rc;
  // same grid rule as slq_panel_coldot (so the two entry points sum in the same order), capped by the table's rows;
  // every partial entry the reduction reads is written by this call's first kernel, so nothing needs clearing
  int g = (int)std::max<int64_t>(1, std::min<int64_t>((row1 - row0 + 63) / 64, 2 * (int64_t)sms));
  if (g > COLDOT_STATIC_MAX_PARTS) g = COLDOT_STATIC_MAX_PARTS;
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
