// Batched Golub-Welsch quadrature for sm_100a: one warp per Jacobi matrix.
//
// Replaces, for a whole block of probes in one launch, the reference's per-probe host calls
//   quadrature(alpha, beta, deg, "gw")  ->  eigh_tridiag -> SciPy eigh_tridiagonal (LAPACK dstemr)
//   (src/primate/integrate.py:58-64, tridiag.py:7-43; in-tree fallback: tqli.py:15-90)
// and the fused evaluation  y = sum_i f(theta_i) tau_i * ||x||^2  (operators.py:150), plus the batched
// Welford/Chan update of the running mean/variance (stats.py:66-86).
//
// Eigen-solver: implicit-shift QL with Wilkinson shifts on (d, e) held in shared memory.  The rotation chain is
// inherently serial, so every lane of the warp executes it redundantly on the shared (d, e); what the warp
// parallelises is (i) the search for a negligible off-diagonal, (ii) when eigenvectors are wanted, the update of
// the eigenvector matrix (each lane owns rows lane, lane+32, ... and applies every plane rotation to its rows),
// (iii) the rank-sort of the Ritz values and (iv) the f(theta).tau reduction.  For the quadrature alone only the
// first eigenvector row is carried (tau = row^2), exactly what integrate.py:64 keeps of LAPACK's output.
#include <cfloat>
#include <cmath>

#include "slq_internal.cuh"

namespace slq {

struct FunParams {
  double p[4];
};

__device__ __forceinline__ double eval_fun(int fun, double x, const FunParams& prm) {
  switch (fun) {
    case SLQ_FUN_IDENTITY: return x;
    case SLQ_FUN_ABS: return fabs(x);
    case SLQ_FUN_SQRT: return sqrt(x);
    case SLQ_FUN_LOG: return log(fmax(x, DBL_EPSILON));  // special.py:90
    case SLQ_FUN_INV: return 1.0 / x;
    case SLQ_FUN_EXP: return exp(prm.p[0] * x);  // special.py:62-66
    case SLQ_FUN_SMOOTHSTEP: {                     // special.py:46-49
      const double a = prm.p[0], b = prm.p[1];
      const double d = (a != b) ? (b - a) : 1.0;
      double y = (x - a) / d;
      y = fmin(fmax(y, 0.0), 1.0);
      return 3.0 * y * y - 2.0 * y * y * y;
    }
    case SLQ_FUN_NUMRANK: return (fabs(x) < prm.p[0]) ? 0.0 : 1.0;  // special.py:69-74 with nonnegative=True
    case SLQ_FUN_SOFTSIGN: {                                         // special.py:20-26
      const int q = (int)prm.p[0];
      const double xc = fmin(fmax(x, -1.0), 1.0);
      const double base = 1.0 - xc * xc;
      double pw = 1.0, coef = 1.0, s = 0.0;
      for (int i = 0; i <= q; ++i) {
        if (i > 0) {
          pw *= base;
          coef *= (2.0 * i - 1.0) / (2.0 * i);
        }
        s += xc * pw * coef;
      }
      return s;
    }
    default: return x;
  }
}

__device__ __forceinline__ double warp_sum(double x) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
  return x;
}

constexpr int QL_MAXIT = 60;

// Implicit QL on the warp's shared (d, e, z).  e[i] couples rows i and i+1 (e[m-1] unused, kept 0).
// If Zt != nullptr the full eigenvector matrix is accumulated: Zt[col*m + row] (column-contiguous so the lanes'
// rows are adjacent in memory).  Returns the number of eigenvalues that ran out of iterations (`maxit` per
// eigenvalue, the reference's `max_iter`, tqli.py:15,54-56).
// Every lane executes the scalar chain redundantly on the SHARED (d, e, z).  Each rotation therefore reads everything
// it needs, meets at a __syncwarp, writes, and meets again after the eigenvector update: under independent thread
// scheduling a lane that owns no eigenvector row (m % 32 != 0) could otherwise run ahead and overwrite e[i], d[i],
// z[i] before the others have read them.
__device__ int warp_ql(double* d, double* e, double* z, int m, double* Zt, int maxit) {
  const int lane = threadIdx.x & 31;
  int nfail = 0;
  for (int l = 0; l < m; ++l) {
    int iter = 0;
    for (;;) {
      __syncwarp();
      // (i) first negligible off-diagonal at or after l, searched 32 entries at a time
      int split = m - 1;
      for (int base = l; base < m - 1; base += 32) {
        const int i = base + lane;
        bool small = false;
        if (i < m - 1) {
          const double dd = fabs(d[i]) + fabs(d[i + 1]);
          small = fabs(e[i]) <= DBL_EPSILON * dd;
        }
        const unsigned mask = __ballot_sync(0xffffffffu, small);
        if (mask) {
          split = base + __ffs(mask) - 1;
          break;
        }
      }
      if (split == l) break;  // d[l] is an eigenvalue
      if (iter++ >= maxit) {
        ++nfail;
        break;
      }
      // Wilkinson shift from the 2x2 at the top of the unreduced block [l, split]
      const double dl = d[l], el = e[l];
      double g = (d[l + 1] - dl) / (2.0 * el);
      double r = hypot(g, 1.0);
      g = d[split] - dl + el / (g + copysign(r, g));
      double s = 1.0, c = 1.0, p = 0.0;
      bool deflated = false;
      __syncwarp();
      for (int i = split - 1; i >= l; --i) {
        // ---- reads
        const double ei = e[i], di = d[i], dip = d[i + 1], z1 = z[i + 1], z0 = z[i];
        __syncwarp();
        // ---- arithmetic and writes (every lane writes the same values)
        const double f = s * ei, b = c * ei;
        r = hypot(f, g);
        e[i + 1] = r;
        if (r == 0.0) {  // exact underflow of the bulge: split here and restart
          d[i + 1] = dip - p;
          e[split] = 0.0;
          deflated = true;
          break;
        }
        s = f / r;
        c = g / r;
        const double di1 = dip - p;
        r = (di - di1) * s + 2.0 * c * b;
        p = s * r;
        d[i + 1] = di1 + p;
        g = c * r - b;
        // plane rotation (c, s) acting on eigenvector columns i, i+1
        z[i + 1] = s * z0 + c * z1;
        z[i] = c * z0 - s * z1;
        if (Zt) {
          for (int row = lane; row < m; row += 32) {
            const double y1 = Zt[(size_t)(i + 1) * m + row], y0 = Zt[(size_t)i * m + row];
            Zt[(size_t)(i + 1) * m + row] = s * y0 + c * y1;
            Zt[(size_t)i * m + row] = c * y0 - s * y1;
          }
        }
        __syncwarp();
      }
      __syncwarp();
      if (deflated) continue;
      const double dl2 = d[l];
      __syncwarp();
      d[l] = dl2 - p;
      e[l] = g;
      e[split] = 0.0;
    }
  }
  __syncwarp();
  return nfail;
}

// In-place batched implicit QL with the contract of the reference's `tqli(d, e, Z, max_iter)` (tqli.py:15-90): on
// return d holds the eigenvalues in the order the QL sweep left them (NOT sorted), e the (negligible) off-diagonals,
// and, when Zt is given, Zt <- (incoming Z) x (eigenvectors of T), stored column-contiguous.  e has a leading zero on
// input (e[i] couples i-1 and i) as in the reference.
__global__ void k_tqli(double* __restrict__ d_io, double* __restrict__ e_io, double* __restrict__ Zt_io, int k, int m, int maxit,
                       int32_t* __restrict__ info) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * (blockDim.x >> 5) + warp;
  if (p >= k) return;
  double* d = sm + (size_t)warp * 3 * m;
  double* e = d + m;
  double* z = e + m;
  for (int i = lane; i < m; i += 32) {
    d[i] = d_io[(size_t)p * m + i];
    e[i] = (i + 1 < m) ? e_io[(size_t)p * m + i + 1] : 0.0;  // shift: e[i] couples i and i+1
    z[i] = (i == 0) ? 1.0 : 0.0;
  }
  __syncwarp();
  const int nfail = warp_ql(d, e, z, m, Zt_io ? Zt_io + (size_t)p * m * m : nullptr, maxit);
  if (info && lane == 0) info[p] = nfail;
  for (int i = lane; i < m; i += 32) {
    d_io[(size_t)p * m + i] = d[i];
    e_io[(size_t)p * m + i] = (i == 0) ? 0.0 : e[i - 1];
  }
}

// One warp per matrix.  Dynamic shared memory: 5*deg doubles per warp (d, e, z, sorted theta, sorted tau/g).
template <typename T>
__global__ void k_quadrature(const T* __restrict__ alpha, const T* __restrict__ beta, int k, int deg, int ld, int fun, FunParams prm,
                             const double* __restrict__ scale, double* __restrict__ theta_out, double* __restrict__ tau_out,
                             double* __restrict__ y_out, int32_t* __restrict__ info, double* __restrict__ Zscratch,
                             double* __restrict__ coeffs) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * (blockDim.x >> 5) + warp;
  if (p >= k) return;
  double* d = sm + (size_t)warp * 5 * deg;
  double* e = d + deg;
  double* z = e + deg;
  double* ts = z + deg;
  double* ws = ts + deg;
  // T_p = tridiag(alpha[p][0:deg], beta[p][1:deg])   (integrate.py:63: d[:deg], e[:deg] with e[0] = 0)
  for (int i = lane; i < deg; i += 32) {
    d[i] = (double)alpha[(size_t)p * ld + i];
    e[i] = (i + 1 < deg) ? (double)beta[(size_t)p * ld + i + 1] : 0.0;
    z[i] = (i == 0) ? 1.0 : 0.0;
  }
  double* Zt = nullptr;
  if (Zscratch) {
    Zt = Zscratch + (size_t)p * deg * deg;
    for (int idx = lane; idx < deg * deg; idx += 32) Zt[idx] = (idx / deg == idx % deg) ? 1.0 : 0.0;
  }
  __syncwarp();
  const int nfail = warp_ql(d, e, z, deg, Zt, QL_MAXIT);
  if (info && lane == 0) info[p] = nfail;

  // rank-sort ascending (LAPACK order, so nodes/weights line up with the reference's)
  for (int i = lane; i < deg; i += 32) {
    const double ti = d[i];
    int rank = 0;
    for (int j = 0; j < deg; ++j) {
      const double tj = d[j];
      rank += (tj < ti) || (tj == ti && j < i);
    }
    ts[rank] = ti;
    ws[rank] = z[i] * z[i];
    e[rank] = (double)i;  // e is free now: remember the permutation
    if (Zscratch) Zscratch[(size_t)k * deg * deg + (size_t)p * deg + rank] = (double)i;  // sorted index -> Zt column
  }
  __syncwarp();
  const double sc = scale ? scale[p] : 1.0;
  double part = 0.0;
  for (int i = lane; i < deg; i += 32) {
    const double th = ts[i], ta = ws[i];
    if (theta_out) theta_out[(size_t)p * deg + i] = th;
    if (tau_out) tau_out[(size_t)p * deg + i] = ta;
    if (fun != SLQ_FUN_NONE) part += eval_fun(fun, th, prm) * ta;
  }
  if (y_out && fun != SLQ_FUN_NONE) {
    part = warp_sum(part);
    if (lane == 0) y_out[p] = part * sc;
  }
  if (coeffs && Zt) {
    // c = Y (f(theta) o Y[0,:])   (operators.py:124); column `orig` of Zt belongs to sorted index i
    __syncwarp();
    for (int row = lane; row < deg; row += 32) {
      double acc = 0.0;
      for (int i = 0; i < deg; ++i) {
        const int orig = (int)e[i];
        const double fv = (fun != SLQ_FUN_NONE) ? eval_fun(fun, ts[i], prm) : ts[i];
        acc += Zt[(size_t)orig * deg + row] * (fv * z[orig]);
      }
      coeffs[(size_t)p * deg + row] = acc;
    }
  }
}

__global__ void k_apply_fun(const double* __restrict__ theta, const double* __restrict__ tau, int k, int deg, int fun, FunParams prm,
                            const double* __restrict__ scale, double* __restrict__ y) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * (blockDim.x >> 5) + warp;
  if (p >= k) return;
  double part = 0.0;
  for (int i = lane; i < deg; i += 32) part += eval_fun(fun, theta[(size_t)p * deg + i], prm) * tau[(size_t)p * deg + i];
  part = warp_sum(part);
  if (lane == 0) y[p] = part * (scale ? scale[p] : 1.0);
}

// FTTR weights (fttr.py:5-29): w_i = 1 / sum_j p_j(theta_i)^2 with the orthonormal polynomials of the Jacobi
// matrix evaluated by the forward three-term recurrence, p_0 = 1/sqrt(mu_0), mu_0 = sum |theta|.  One thread per node.
template <typename T>
__global__ void k_fttr(const double* __restrict__ theta, const T* __restrict__ alpha, const T* __restrict__ beta, int k, int deg,
                       int nterms, int ld, double* __restrict__ tau) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * (blockDim.x >> 5) + warp;
  if (p >= k) return;
  double mu = 0.0;
  for (int i = lane; i < deg; i += 32) mu += fabs(theta[(size_t)p * deg + i]);
  mu = warp_sum(mu);
  const double mu_sqrt_rec = 1.0 / sqrt(mu);
  const T* a = alpha + (size_t)p * ld;
  const T* b = beta + (size_t)p * ld;  // b[0] = 0, b[j] couples j-1 and j
  for (int i = lane; i < deg; i += 32) {
    const double x = theta[(size_t)p * deg + i];
    double pm1 = mu_sqrt_rec;                                          // p_0
    double pc = (nterms > 1) ? (x - (double)a[0]) * pm1 / (double)b[1] : 0.0;  // p_1
    double ss = pm1 * pm1 + ((nterms > 1) ? pc * pc : 0.0);
    for (int j = 2; j < nterms; ++j) {
      const double s = (x - (double)a[j - 1]) / (double)b[j];
      const double t = -(double)b[j - 1] / (double)b[j];
      const double pn = s * pc + t * pm1;
      pm1 = pc;
      pc = pn;
      ss += pn * pn;
    }
    tau[(size_t)p * deg + i] = (1.0 / ss) / mu;  // fttr.py:28-29
  }
}

// Chan et al. merge of a batch into (count, mean, M2)   (stats.py:66-86 restated for dim = 1).
__global__ void k_stats_update(const double* __restrict__ y, int m, double* __restrict__ state) {
  const int lane = threadIdx.x;
  double s = 0.0;
  for (int i = lane; i < m; i += 32) s += y[i];
  s = warp_sum(s);
  const double bmean = s / (double)m;
  double q = 0.0;
  for (int i = lane; i < m; i += 32) {
    const double c = y[i] - bmean;
    q += c * c;
  }
  q = warp_sum(q);
  if (lane == 0) {
    const double n0 = state[0], mu0 = state[1], S0 = state[2];
    const double n1 = n0 + (double)m;
    const double delta = bmean - mu0;
    state[0] = n1;
    state[1] = mu0 + ((double)m / n1) * delta;
    state[2] = S0 + q + (n0 * (double)m / n1) * (delta * delta);
  }
}

static int quad_smem_and_warps(int deg, int* warps_per_cta, size_t* smem) {
  const size_t per_warp = sizeof(double) * 5 * (size_t)deg;
  int w = 4;
  while (w > 1 && per_warp * w > 200 * 1024) w >>= 1;
  if (per_warp * w > 220 * 1024) return fail(SLQ_ERR_UNSUPPORTED, "deg=%d too large for the shared-memory eigensolver", deg);
  *warps_per_cta = w;
  *smem = per_warp * w;
  return SLQ_OK;
}

template <typename T>
static int quadrature_launch(const void* alpha, const void* beta, int k, int deg, int ld, int fun, const FunParams& prm,
                             const double* scale, double* theta, double* tau, double* y, int32_t* info, double* Z, double* coeffs,
                             cudaStream_t st) {
  int w;
  size_t smem;
  int rc = quad_smem_and_warps(deg, &w, &smem);
  if (rc != SLQ_OK) return rc;
  if (smem > 48 * 1024) SLQ_CUDA(cudaFuncSetAttribute(k_quadrature<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned blocks = (unsigned)((k + w - 1) / w);
  k_quadrature<T><<<blocks, w * 32, smem, st>>>((const T*)alpha, (const T*)beta, k, deg, ld, fun, prm, scale, theta, tau, y, info, Z,
                                               coeffs);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

static FunParams make_params(const double* params) {
  FunParams p;
  for (int i = 0; i < 4; ++i) p.p[i] = params ? params[i] : 0.0;
  return p;
}

}  // namespace slq

using namespace slq;

extern "C" {

int slq_quadrature(const void* alpha_dev, const void* beta_dev, int dtype, int k, int deg, int ld, int fun, const double* params,
                   const double* scale_dev, double* theta_dev, double* tau_dev, double* y_dev, int32_t* info_dev, slq_stream stream) {
  SLQ_REQUIRE(alpha_dev && beta_dev, "NULL argument");
  SLQ_REQUIRE(k > 0 && deg > 0 && ld >= deg, "invalid shape k=%d deg=%d ld=%d", k, deg, ld);
  SLQ_REQUIRE(fun >= SLQ_FUN_NONE && fun <= SLQ_FUN_SOFTSIGN, "Unknown function: %d.", fun);
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  FunParams prm = make_params(params);
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == SLQ_F64 ? quadrature_launch<double>(alpha_dev, beta_dev, k, deg, ld, fun, prm, scale_dev, theta_dev, tau_dev, y_dev,
                                                      info_dev, nullptr, nullptr, st)
                          : quadrature_launch<float>(alpha_dev, beta_dev, k, deg, ld, fun, prm, scale_dev, theta_dev, tau_dev, y_dev,
                                                     info_dev, nullptr, nullptr, st);
}

int slq_funm_coeffs(const void* alpha_dev, const void* beta_dev, int dtype, int k, int deg, int ld, int fun, const double* params,
                    double* coeffs_dev, double* theta_dev, double* tau_dev, double* scratch_dev, int32_t* info_dev, slq_stream stream) {
  SLQ_REQUIRE(alpha_dev && beta_dev && coeffs_dev && scratch_dev, "NULL argument");
  SLQ_REQUIRE(k > 0 && deg > 0 && ld >= deg, "invalid shape k=%d deg=%d ld=%d", k, deg, ld);
  SLQ_REQUIRE(fun >= SLQ_FUN_NONE && fun <= SLQ_FUN_SOFTSIGN, "Unknown function: %d.", fun);
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  FunParams prm = make_params(params);
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == SLQ_F64 ? quadrature_launch<double>(alpha_dev, beta_dev, k, deg, ld, fun, prm, nullptr, theta_dev, tau_dev, nullptr,
                                                      info_dev, scratch_dev, coeffs_dev, st)
                          : quadrature_launch<float>(alpha_dev, beta_dev, k, deg, ld, fun, prm, nullptr, theta_dev, tau_dev, nullptr,
                                                     info_dev, scratch_dev, coeffs_dev, st);
}

int slq_tqli(double* d_dev, double* e_dev, double* Zt_dev, int k, int m, int max_iter, int32_t* info_dev, slq_stream stream) {
  SLQ_REQUIRE(d_dev && e_dev, "NULL argument");
  SLQ_REQUIRE(k > 0 && m > 0 && max_iter >= 0, "invalid shape k=%d m=%d max_iter=%d", k, m, max_iter);
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const size_t per_warp = sizeof(double) * 3 * (size_t)m;
  int w = 4;
  while (w > 1 && per_warp * w > 200 * 1024) w >>= 1;
  if (per_warp * w > 220 * 1024) return fail(SLQ_ERR_UNSUPPORTED, "m=%d too large for the shared-memory eigensolver", m);
  const size_t smem = per_warp * w;
  if (smem > 48 * 1024) SLQ_CUDA(cudaFuncSetAttribute(k_tqli, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_tqli<<<(unsigned)((k + w - 1) / w), w * 32, smem, (cudaStream_t)stream>>>(d_dev, e_dev, Zt_dev, k, m, max_iter, info_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

int slq_apply_fun(const double* theta_dev, const double* tau_dev, int k, int deg, int fun, const double* params, const double* scale_dev,
                  double* y_dev, slq_stream stream) {
  SLQ_REQUIRE(theta_dev && tau_dev && y_dev, "NULL argument");
  SLQ_REQUIRE(k > 0 && deg > 0, "invalid shape");
  SLQ_REQUIRE(fun >= SLQ_FUN_IDENTITY && fun <= SLQ_FUN_SOFTSIGN, "Unknown function: %d.", fun);
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  k_apply_fun<<<(unsigned)((k + 3) / 4), 128, 0, (cudaStream_t)stream>>>(theta_dev, tau_dev, k, deg, fun, make_params(params), scale_dev,
                                                                         y_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

int slq_fttr(const double* theta_dev, const void* alpha_dev, const void* beta_dev, int dtype, int k, int deg, int nterms, int ld,
             double* tau_dev, slq_stream stream) {
  SLQ_REQUIRE(theta_dev && alpha_dev && beta_dev && tau_dev, "NULL argument");
  SLQ_REQUIRE(k > 0 && deg > 0 && nterms > 0 && ld >= nterms, "invalid shape");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SLQ_F64)
    k_fttr<double><<<(unsigned)((k + 3) / 4), 128, 0, st>>>(theta_dev, (const double*)alpha_dev, (const double*)beta_dev, k, deg, nterms, ld, tau_dev);
  else
    k_fttr<float><<<(unsigned)((k + 3) / 4), 128, 0, st>>>(theta_dev, (const float*)alpha_dev, (const float*)beta_dev, k, deg, nterms, ld, tau_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

int slq_stats_update(const double* y_dev, int m, double* state_dev, slq_stream stream) {
  SLQ_REQUIRE(y_dev && state_dev && m > 0, "invalid argument");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  k_stats_update<<<1, 32, 0, (cudaStream_t)stream>>>(y_dev, m, state_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // extern "C"
