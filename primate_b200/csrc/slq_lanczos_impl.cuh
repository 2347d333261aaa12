#pragma once
// Probe-blocked Lanczos driver for sm_100a.
//
// Restates, for a block of k probes at once, the reference's Paige-A27 recurrence
// (src/primate/include/lanczos.h:92-149) and its modified-Gram-Schmidt re-projection (lanczos.h:43-66):
//
//   w  = A q_c - beta_j q_p          alpha_j = <q_c, w>          (kernel "A": CSR SpMM fused with both)
//   v  = w - alpha_j q_c             [MGS against stored q's]    beta_{j+1} = ||v||   (kernel(s) "axpy")
//   q_n = v / beta_{j+1}                                          (never materialised: see below)
//
// Layout: every Lanczos vector block is an n x K row-major panel (probe-minor).  Vectors are kept
// UN-normalised (u = beta * q) with a per-probe reciprocal scale, so the division by beta_{j+1} -- which
// needs a grid-wide reduction first -- costs no extra pass over HBM: consumers multiply by the scale on load.
// Grid-wide reductions (alpha, beta, MGS coefficients) are deterministic: per-CTA partials in a table, summed
// in fixed order by the last CTA to finish, which also runs the scalar recurrence bookkeeping (early break
// mask, lanczos.h:140).  Dots and norms accumulate in fp64 for both fp32 and fp64 panels.
#include <cooperative_groups.h>
#include <type_traits>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#include <mutex>

#include "slq_lanczos_common.cuh"

// tuning knobs of the staged CSR kernel: resident CTAs per SM asked of ptxas, gathers kept in flight per row
#ifndef SLQ_STEPA_MINB
#define SLQ_STEPA_MINB 4
#endif
#ifndef SLQ_GATHER_U
#define SLQ_GATHER_U 5
#endif


namespace slq {

// Sum column `t` of the per-CTA partial table in a fixed order using the whole CTA.
template <int NS>
__device__ __forceinline__ void reduce_partials(const double* partial, int ncta, double* smem /* NT */, double* result /* NS, smem */) {
  constexpr int NCH = NT / NS > 0 ? NT / NS : 1;
  const int t = threadIdx.x;
  if (NS <= NT) {
    const int slot = t % NS, ch = t / NS;
    double s = 0.0;
    if (ch < NCH)
      for (int c = ch; c < ncta; c += NCH) s += __ldcg(partial + (size_t)c * NS + slot);
    smem[t] = s;  // one slot per thread (CTAs wider than NT threads pass a scratch of blockDim.x doubles)
    __syncthreads();
    if (t < NS) {
      double r = 0.0;
#pragma unroll
      for (int c = 0; c < NCH; ++c) r += smem[c * NS + t];
      result[t] = r;
    }
    __syncthreads();
  }
}

template <typename T>
__device__ __forceinline__ double round_to(double x) {
  return (double)(T)x;
}

// ---- finalisers (run by every thread of the last CTA) ---------------------------------------------------
enum Fin { FIN_INIT = 0, FIN_ALPHA = 1, FIN_BETA = 2, FIN_COEF = 3, FIN_FUSED = 4 };

template <typename T, int K, int FIN>
__device__ __forceinline__ void finalize(const double* partial, LzScal* sc, double* scale_a, const double* scale_b,
                                         const StepOut& so, double* smem) {
  constexpr int NACC = (FIN == FIN_COEF) ? 2 : (FIN == FIN_FUSED ? 3 : 1);
  constexpr int NS = NACC * K;
  __shared__ double s_res[3 * MAXK];
  reduce_partials<NS>(partial, gridDim.x, smem, s_res);
  const int t = threadIdx.x;
  if (t < K) {
    if (FIN == FIN_INIT) {
      // q_0 = x / ||x||  (Eigen normalized(): untouched when the norm is zero), lanczos.h:120
      double z = s_res[t];
      sc->xnorm2[t] = z;
      double nrm = round_to<T>(sqrt(round_to<T>(z)));
      scale_a[t] = (z > 0.0) ? round_to<T>(1.0 / nrm) : 1.0;
      sc->beta[t] = 0.0;
      sc->coef[t] = 0.0;
      sc->active[t] = 1;
      sc->steps[t] = 0;
    } else if (FIN == FIN_ALPHA) {
      // alpha_j = <q_c, w>; pending update v = w - alpha_j q_c  (lanczos.h:129-130)
      double a = round_to<T>(s_res[t]);
      if (sc->active[t]) reinterpret_cast<T*>(so.alpha)[(size_t)t * so.ld + so.j] = (T)a;
      sc->coef[t] = round_to<T>(a * scale_b[t]);
    } else if (FIN == FIN_BETA) {
      // beta_{j+1} = ||v||; early break when < sqrt(n) rtol or at the last step (lanczos.h:139-143)
      double b = round_to<T>(sqrt(s_res[t]));
      if (sc->active[t]) {
        reinterpret_cast<T*>(so.beta)[(size_t)t * so.ld + so.j + 1] = (T)b;
        sc->steps[t] = so.j + 1;
        if (b < so.tol || so.j + 1 == so.deg) sc->active[t] = 0;
      }
      sc->beta[t] = b;
      scale_a[t] = round_to<T>(1.0 / b);
      sc->coef[t] = 0.0;
    } else if (FIN == FIN_FUSED) {
      // Fused step (k_stepA_tma<STEP_FUSED>): alpha_j = <q_c, w> AND a provisional beta_{j+1} = ||w - alpha_j q_c|| in one
      // finaliser, the norm from the three sums the step kernel accumulated in fp64 -- S_a = <q_c, w>, S_w = <w, w>,
      // S_v = <v_c, v_c> -- so that the vector v_n = w - cf v_c need not exist before the next step starts (it is
      // formed on the fly there):   ||v_n||^2 = S_w - 2 cf <v_c, w> + cf^2 S_v,  <v_c, w> = S_a / r_c,  cf = alpha_j r_c.
      // The subtraction loses log10(S_w / ||v_n||^2) of the 16 digits carried, so the provisional value only SCALES the
      // next vector (vectors are stored un-normalised: a scale off by 1e-15 is a normalisation rounding error).  What
      // is RECORDED as beta_j, and what the early-break test of lanczos.h:139-143 looks at, is the exact norm
      // sqrt(S_v) of the formed vector, which arrives one step later (here, for the current vector): a breakdown is
      // therefore noticed one kernel late and the step that should not have run is simply not recorded.
      const double r = scale_b[t];
      bool act = sc->active[t] != 0;
      if (act && so.j > 0) {
        const double bex = round_to<T>(sqrt(s_res[2 * K + t]));
        reinterpret_cast<T*>(so.beta)[(size_t)t * so.ld + so.j] = (T)bex;
        if (bex < so.tol) {  // the reference stopped after the previous step: steps stays j
          act = false;
          sc->active[t] = 0;
        }
      }
      const double a = round_to<T>(s_res[t]);
      const double cf = round_to<T>(a * r);
      sc->coef[t] = cf;
      const double vw = (r != 0.0) ? s_res[t] / r : 0.0;
      const double b2 = s_res[K + t] - 2.0 * cf * vw + cf * cf * s_res[2 * K + t];
      const double b = round_to<T>(sqrt(fmax(b2, 0.0)));
      if (act) {
        reinterpret_cast<T*>(so.alpha)[(size_t)t * so.ld + so.j] = (T)a;
        reinterpret_cast<T*>(so.beta)[(size_t)t * so.ld + so.j + 1] = (T)b;  // provisional (final only for j + 1 == deg)
        sc->steps[t] = so.j + 1;
        if (so.j + 1 == so.deg) sc->active[t] = 0;
      }
      sc->beta[t] = b;
      scale_a[t] = (b > 0.0) ? round_to<T>(1.0 / b) : 0.0;
    } else {  // FIN_COEF: one modified-Gram-Schmidt projection coefficient (lanczos.h:59-64)
      double r = scale_b[t];
      double u_norm = round_to<T>(s_res[t] * r * r);
      double s_proj = round_to<T>(s_res[K + t] * r);
      bool take = (u_norm > so.orth_tol) && (fabs(s_proj) > so.orth_tol);
      sc->coef[t] = take ? round_to<T>(round_to<T>(s_proj / u_norm) * r) : 0.0;
    }
  }
}

// ---- kernel: ||x||^2 of the start block --------------------------------------------------------------------
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_init(int64_t n, const T* __restrict__ x, LzScal* sc, double* scale0, double* partial,
                                              StepOut so) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  __shared__ double smem[NT > NWARP * K ? NT : NWARP * K];
  double acc[1][V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc[0][i] = 0.0;
  const int64_t nvec = n * LPR;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < nvec; e += (int64_t)gridDim.x * NT) {
    Pack<T, V> xv = ld_pack<T, V, LD_DEFAULT>(x + e * V);
#pragma unroll
    for (int i = 0; i < V; ++i) acc[0][i] += (double)xv.v[i] * (double)xv.v[i];
  }
  cta_reduce_slots<K, V, LPR, 1>(acc, partial + (size_t)blockIdx.x * K, smem);
  if (elect_last_cta(&sc->counter)) finalize<T, K, FIN_INIT>(partial, sc, scale0, nullptr, so, smem);
}

// ---- kernel A: w = (A u_c) r_c - beta_j (u_p r_p);  alpha partials ----------------------------------------
// CSR SpMM over the probe block fused with the first half of the three-term recurrence.  One thread owns V
// adjacent probes of a row (32-byte vector accesses); LPR threads cover a row; a CTA sweeps tiles of
// consecutive rows assigned round-robin so that all CTAs advance through the matrix as one front (keeps the
// gathered rows of u_c resident in L1/L2 while neighbours re-use them).
template <typename T, int K, bool FIRST, bool FUSED_CSR>
__global__ void __launch_bounds__(NT)
    k_stepA(int64_t n, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, const T* __restrict__ vals,
            const T* __restrict__ uc, const T* __restrict__ up, T* __restrict__ w, LzScal* sc, const double* scale_c,
            const double* scale_p, double* partial, StepOut so) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  constexpr int RPC = NT / LPR;  // rows per CTA pass
  constexpr int PASSES = 4;
  constexpr int TILE = RPC * PASSES;
  __shared__ double smem[NT > NWARP * K ? NT : NWARP * K];

  const int sub = threadIdx.x % LPR, rloc = threadIdx.x / LPR;
  const int slot = sub * V;
  T rc[V], brp[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    rc[i] = (T)scale_c[slot + i];
    brp[i] = FIRST ? (T)0 : (T)(sc->beta[slot + i]) * (T)scale_p[slot + i];
  }
  double acc_a[1][V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc_a[0][i] = 0.0;

  const int64_t ntiles = (n + TILE - 1) / TILE;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t row0 = tile * TILE;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
      const int64_t r = row0 + ps * RPC + rloc;
      if (r < n) {
        T acc[V];
        if (FUSED_CSR) {
#pragma unroll
          for (int i = 0; i < V; ++i) acc[i] = (T)0;
          const int s = __ldg(rowptr + r), e = __ldg(rowptr + r + 1);
#pragma unroll 4
          for (int i = s; i < e; ++i) {
            const int64_t c = __ldg(colidx + i);
            const T a = __ldg(vals + i);
            Pack<T, V> xv = ld_pack<T, V, LD_READONLY>(uc + c * K + slot);
#pragma unroll
            for (int q = 0; q < V; ++q) acc[q] = fma(a, xv.v[q], acc[q]);
          }
        } else {
          Pack<T, V> yv = ld_pack<T, V, LD_DEFAULT>(w + r * K + slot);  // y = A u_c computed by the operator
#pragma unroll
          for (int q = 0; q < V; ++q) acc[q] = yv.v[q];
        }
        Pack<T, V> qc = ld_pack<T, V, LD_READONLY>(uc + r * K + slot);
        Pack<T, V> wv;
        if (FIRST) {
#pragma unroll
          for (int q = 0; q < V; ++q) wv.v[q] = mul_rn(acc[q], rc[q]);
        } else {
          Pack<T, V> pv = ld_pack<T, V, LD_STREAM>(up + r * K + slot);
#pragma unroll
          for (int q = 0; q < V; ++q) wv.v[q] = fnma_rn(brp[q], pv.v[q], mul_rn(acc[q], rc[q]));
        }
        st_pack<T, V, false>(w + r * K + slot, wv);
#pragma unroll
        for (int q = 0; q < V; ++q) acc_a[0][q] += (double)(qc.v[q] * rc[q]) * (double)wv.v[q];
      }
    }
  }
  cta_reduce_slots<K, V, LPR, 1>(acc_a, partial + (size_t)blockIdx.x * K, smem);
  if (elect_last_cta(&sc->counter)) finalize<T, K, FIN_ALPHA>(partial, sc, nullptr, scale_c, so, smem);
}

// ---- CSR row pipeline ------------------------------------------------------------------------------------------
// A thread owns V adjacent probes of a row (one 16-byte slice of the probe block per gathered row); the LPR threads
// of a row group sit in one warp.  The un-pipelined loop serialises rowptr -> (colidx, vals) -> X per row: three to
// five DRAM latencies per row, which is what bounds it (36% of HBM peak measured).  Here the index stream is
// software-pipelined in registers, with no shared memory and no barriers:
//   two rows ahead : the row's [s, e) bounds are loaded;
//   one row ahead  : lane `sub` of the group loads colidx[s + ch*LPR + sub] / vals[...] -- one coalesced load covers
//                    LPR consecutive non-zeros of the row (CH such loads per lane when LPR is small);
//   current row    : indices and values are broadcast inside the group by warp shuffles and up to GATHER_U
//                    16-byte probe-row gathers are issued back to back before the FMAs consume them, in storage
//                    (column) order.
// Rows longer than the prefetched chunk fetch their remaining chunks in line.
template <typename T, int K>
struct CsrTiling {
  // gathers use 16-byte slices (the 32-byte packs of the streaming kernels cost too many registers per gather in flight)
  static constexpr int VMAX = 16 / (int)sizeof(T);
  static constexpr int V = K < VMAX ? K : VMAX;
  static constexpr int LPR = K / V;
  static constexpr int RPC = NT / LPR;  // rows per CTA pass
  static constexpr int PASSES = RPC >= 256 ? 1 : (RPC >= 128 ? 2 : 4);
  static constexpr int TILE = RPC * PASSES;
  static constexpr int CH = LPR >= 8 ? 1 : 8 / LPR;  // (index, value) pairs held per lane
  static constexpr int CS = CH * LPR;                // non-zeros per prefetched chunk
  static constexpr int U = LPR < SLQ_GATHER_U ? LPR : SLQ_GATHER_U;
};

template <typename T>
__device__ __forceinline__ T shfl_group(T v, int src, int width);
template <>
__device__ __forceinline__ int shfl_group<int>(int v, int src, int width) {
  return __shfl_sync(0xffffffffu, v, src, width);
}
template <>
__device__ __forceinline__ float shfl_group<float>(float v, int src, int width) {
  return __shfl_sync(0xffffffffu, v, src, width);
}
template <>
__device__ __forceinline__ double shfl_group<double>(double v, int src, int width) {
  return __shfl_sync(0xffffffffu, v, src, width);
}


template <typename T, int CH>
struct RowChunk {
  int c[CH];
  T a[CH];
};

template <typename T, int K>
__device__ __forceinline__ void chunk_load(RowChunk<T, CsrTiling<T, K>::CH>& ck, const int32_t* __restrict__ colidx,
                                           const T* __restrict__ vals, int first, int e, int sub) {
  using C = CsrTiling<T, K>;
#pragma unroll
  for (int ch = 0; ch < C::CH; ++ch) {
    const int i = first + ch * C::LPR + sub;
    const bool on = i < e;
    ck.c[ch] = on ? __ldg(colidx + i) : 0;
    ck.a[ch] = on ? __ldg(vals + i) : (T)0;
  }
}

// acc += sum over the chunk's first `cnt` non-zeros (cnt_max: warp-uniform upper bound of cnt)
template <typename T, int K, int V>
__device__ __forceinline__ void chunk_gather(T (&acc)[V], const RowChunk<T, CsrTiling<T, K>::CH>& ck, const T* __restrict__ X, int slot,
                                             int cnt, int cnt_max) {
  using C = CsrTiling<T, K>;
#pragma unroll
  for (int ch = 0; ch < C::CH; ++ch) {
#pragma unroll
    for (int l = 0; l < C::LPR; l += C::U) {
      if (ch * C::LPR + l < cnt_max) {  // warp-uniform
        Pack<T, V> xv[C::U];
#pragma unroll
        for (int u = 0; u < C::U; ++u) {
          if (l + u < C::LPR) {
            const int c = shfl_group<int>(ck.c[ch], l + u, C::LPR);
            if (ch * C::LPR + l + u < cnt) {
              xv[u] = ld_pack<T, V, LD_READONLY>(X + (int64_t)c * K + slot);
            } else {
#pragma unroll
              for (int q = 0; q < V; ++q) xv[u].v[q] = (T)0;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < C::U; ++u) {
          if (l + u < C::LPR) {
            const T a = shfl_group<T>(ck.a[ch], l + u, C::LPR);  // value broadcast only now: keeps registers free for the gathers
            if (ch * C::LPR + l + u < cnt) {
#pragma unroll
              for (int q = 0; q < V; ++q) acc[q] = fma(a, xv[u].v[q], acc[q]);
            }
          }
        }
      }
    }
  }
}

// PLAIN: w = A u_c only (slq_op_matmat).  REPLAY: the whole Lanczos step with RECORDED alpha/beta (second pass of the
// basis-free f(A) X): v = (A u_c) r_c - beta r_p u_p - alpha r_c u_c and Y += c_{j+1} ||x|| r_n v in one kernel.
// FUSED (bulk-copy kernel, rows of <= 64 non-zeros, orth = 0): one kernel per Lanczos step.  The update v = w - cf u of the
// PREVIOUS step is applied on the fly (own rows: written out; gathered neighbours: recomputed from the two staged tiles)
// and the step's norm comes out of the finaliser (FIN_FUSED), so the separate k_axpy pass over three panels goes away.
enum StepMode { STEP_FIRST = 0, STEP_NEXT = 1, STEP_PLAIN = 2, STEP_REPLAY = 3, STEP_FUSED = 4 };

// Per-step scalars of the replay (one row of the table built by k_replay_table): rc | brp | cf | yc, MAXK doubles each.
struct ReplayArgs {
  const double* row = nullptr;
  const double* y0 = nullptr;  // coefficient of the start vector (read when first != 0)
  void* Y = nullptr;           // n x K accumulator panel (STEP_FUSED: the panel the formed vector is written to, or NULL)
  int first = 0;
  double* scale_out = nullptr;  // STEP_FUSED: where the finaliser stores the scale 1 / beta of the NEXT vector
};

// ---- kernel A (CSR operators): w = (A u_c) r_c - beta_j (u_p r_p); alpha partials ------------------------------------
template <typename T, int K, int MODE>
__global__ void __launch_bounds__(NT, SLQ_STEPA_MINB)
    k_stepA_csr(int n, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, const T* __restrict__ vals,
                const T* __restrict__ uc, const T* __restrict__ up, T* __restrict__ w, LzScal* sc, const double* scale_c,
                const double* scale_p, double* partial, StepOut so) {
  using C = CsrTiling<T, K>;
  constexpr int V = C::V, LPR = C::LPR, RPC = C::RPC, PASSES = C::PASSES, TILE = C::TILE, CS = C::CS;
  constexpr bool FIRST = (MODE == STEP_FIRST), PLAIN = (MODE == STEP_PLAIN);
  __shared__ double smem[NT > NWARP * K ? NT : NWARP * K];
  __shared__ T s_rc[K], s_brp[K];  // per-probe scales: read per row from shared memory instead of living in registers

  const int sub = threadIdx.x % LPR, rloc = threadIdx.x / LPR;
  const int slot = sub * V;
  if (!PLAIN) {
    for (int t = threadIdx.x; t < K; t += NT) {
      s_rc[t] = (T)scale_c[t];
      s_brp[t] = FIRST ? (T)0 : (T)(sc->beta[t]) * (T)scale_p[t];
    }
    __syncthreads();
  }
  double acc_a[1][V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc_a[0][i] = 0.0;

  const int ntiles = (n + TILE - 1) / TILE;
  const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int nsteps = my_tiles * PASSES;  // CTA-uniform trip count (shuffles need whole warps)
  auto row_of = [&](int m) -> int {      // the m-th row this thread visits, or -1
    if (m >= nsteps) return -1;
    const long long r = ((long long)blockIdx.x + (long long)(m / PASSES) * gridDim.x) * TILE + (m % PASSES) * RPC + rloc;
    return r < n ? (int)r : -1;
  };
  auto bounds = [&](int r, int& s, int& e) {
    s = e = 0;
    if (r >= 0) {
      s = __ldg(rowptr + r);
      e = __ldg(rowptr + r + 1);
    }
  };

  // pipeline prologue: bounds of rows 0 and 1, chunk of row 0
  int s0, e0, s1, e1;
  int r0 = row_of(0), r1 = row_of(1);
  bounds(r0, s0, e0);
  bounds(r1, s1, e1);
  RowChunk<T, C::CH> ck0, ck1;
  chunk_load<T, K>(ck0, colidx, vals, s0, e0, sub);

  for (int m = 0; m < nsteps; ++m) {
    // look-ahead: chunk of row m+1 (its bounds arrived during the previous iteration), bounds of row m+2
    chunk_load<T, K>(ck1, colidx, vals, s1, e1, sub);
    const int r2 = row_of(m + 2);
    int s2, e2;
    bounds(r2, s2, e2);

    const size_t off = (size_t)(r0 >= 0 ? r0 : 0) * K + slot;
    Pack<T, V> pv;
    if (!FIRST && !PLAIN && r0 >= 0) pv = ld_pack<T, V, LD_STREAM>(up + off);
    T acc[V];
#pragma unroll
    for (int q = 0; q < V; ++q) acc[q] = (T)0;
    const int len = e0 - s0;
    const int len_max = (int)__reduce_max_sync(0xffffffffu, (unsigned)len);
    chunk_gather<T, K, V>(acc, ck0, uc, slot, len < CS ? len : CS, len_max < CS ? len_max : CS);
    for (int base = CS; base < len_max; base += CS) {  // long rows: remaining chunks in line (warp-uniform loop)
      RowChunk<T, C::CH> ck;
      chunk_load<T, K>(ck, colidx, vals, s0 + base, e0, sub);
      int cnt = len - base;
      cnt = cnt < 0 ? 0 : (cnt < CS ? cnt : CS);
      const int rest = len_max - base;
      chunk_gather<T, K, V>(acc, ck, uc, slot, cnt, rest < CS ? rest : CS);
    }
    if (r0 >= 0) {
      Pack<T, V> wv;
      if (PLAIN) {
#pragma unroll
        for (int q = 0; q < V; ++q) wv.v[q] = acc[q];
        st_pack<T, V, true>(w + off, wv);
      } else {
        Pack<T, V> qc = ld_pack<T, V, LD_READONLY>(uc + off);  // L1/L2 hit: the row was just gathered by a neighbour
#pragma unroll
        for (int q = 0; q < V; ++q) {
          const T rcq = s_rc[slot + q];
          wv.v[q] = FIRST ? mul_rn(acc[q], rcq) : fnma_rn(s_brp[slot + q], pv.v[q], mul_rn(acc[q], rcq));
          acc_a[0][q] += (double)(qc.v[q] * rcq) * (double)wv.v[q];
        }
        st_pack<T, V, false>(w + off, wv);
      }
    }
    // rotate the pipeline
    r0 = r1, s0 = s1, e0 = e1, ck0 = ck1;
    r1 = r2, s1 = s2, e1 = e2;
  }
  if (!PLAIN) {
    cta_reduce_slots<K, V, LPR, 1>(acc_a, partial + (size_t)blockIdx.x * K, smem);
    if (elect_last_cta(&sc->counter)) finalize<T, K, FIN_ALPHA>(partial, sc, nullptr, scale_c, so, smem);
  }
}

// Launch helper: persistent grid = resident CTAs per SM (register limited) x SM count.
template <typename T, int K, int MODE>
static int launch_stepA_csr(const slq_op* op, const T* uc, const T* up, T* w, LzScal* sc, const double* scale_c, const double* scale_p,
                            double* partial, const StepOut& so, cudaStream_t st) {
  using C = CsrTiling<T, K>;
  const int64_t n = op->n;
  const int sms = sm_count();
  static std::atomic<int> resident_dev[MAX_DEVICES];  // per template instantiation and device (0 = not queried yet)
  const int dev = current_device();
  int resident = (dev >= 0 && dev < MAX_DEVICES) ? resident_dev[dev].load(std::memory_order_relaxed) : 0;
  if (resident == 0) {
    SLQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, k_stepA_csr<T, K, MODE>, NT, 0));
    if (resident < 1) resident = 1;
    if (resident > 8) resident = 8;  // the partial table holds 8 CTAs per SM
    if (dev >= 0 && dev < MAX_DEVICES) resident_dev[dev].store(resident, std::memory_order_relaxed);
  }
  int64_t ntiles = (n + C::TILE - 1) / C::TILE;
  int64_t g = ntiles < (int64_t)sms * resident ? ntiles : (int64_t)sms * resident;
  k_stepA_csr<T, K, MODE><<<(unsigned)g, NT, 0, st>>>((int)n, op->rowptr, op->colidx, (const T*)op->vals, uc, up, w, sc, scale_c, scale_p,
                                                      partial, so);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // namespace slq
#include "slq_stepa_tma.cuh"
#include "slq_dense.cuh"
namespace slq {

// Step A for a CSR operator: the TMA-staged kernel where the block width allows it (row slices of 64..512 bytes and
// 16-byte aligned CSR arrays), the register-pipelined kernel otherwise.  SLQ_STEPA=csr forces the latter (A/B timing).
static bool use_tma_path() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("SLQ_STEPA");
    cached = (e && strcmp(e, "csr") == 0) ? 0 : 1;
  }
  return cached == 1;
}

// SLQ_FUSED=0 keeps the two-kernel step (step A + axpy) everywhere (A/B timing, parity of the two paths).
static bool use_fused_step() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("SLQ_FUSED");
    cached = (e && strcmp(e, "0") == 0) ? 0 : 1;
  }
  return cached == 1;
}

template <typename T, int K, int MODE>
static int launch_stepA(const slq_op* op, const T* uc, const T* up, T* w, LzScal* sc, const double* scale_c, const double* scale_p,
                        double* partial, const StepOut& so, cudaStream_t st) {
  if constexpr (tma_supported<T, K>()) {
    const bool aligned = (((uintptr_t)op->colidx | (uintptr_t)op->vals | (uintptr_t)uc | (uintptr_t)up | (uintptr_t)w) & 15) == 0;
    if (aligned && use_tma_path()) return launch_stepA_tma<T, K, MODE>(op, uc, up, w, sc, scale_c, scale_p, partial, so, st);
  }
  return launch_stepA_csr<T, K, MODE>(op, uc, up, w, sc, scale_c, scale_p, partial, so, st);
}

// ---- kernel "axpy": v -= coef * u1, then either ||v||^2 or the next projection's dots ------------------------
enum AxpyMode { AX_NORM = 0, AX_DOT_SAME = 1, AX_DOT_OTHER = 2 };

template <typename T, int K, int MODE>
__global__ void __launch_bounds__(NT) k_axpy(int64_t n, T* __restrict__ v, const T* __restrict__ u1, const T* __restrict__ u2,
                                              LzScal* sc, double* scale_v, const double* scale_u2, double* partial, StepOut so) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  constexpr int NACC = (MODE == AX_NORM) ? 1 : 2;
  __shared__ double smem[NT > NWARP * NACC * K ? NT : NWARP * NACC * K];
  const int slot = (threadIdx.x % LPR) * V;
  T cf[V];
#pragma unroll
  for (int i = 0; i < V; ++i) cf[i] = (T)sc->coef[slot + i];
  double acc[NACC][V];
#pragma unroll
  for (int a = 0; a < NACC; ++a)
#pragma unroll
    for (int i = 0; i < V; ++i) acc[a][i] = 0.0;

  const int64_t nvec = n * LPR;
  const int64_t stride = (int64_t)gridDim.x * NT;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < nvec; e += stride) {
    Pack<T, V> vv = ld_pack<T, V, LD_STREAM>(v + e * V);
    Pack<T, V> a1 = ld_pack<T, V, LD_DEFAULT>(u1 + e * V);
    Pack<T, V> a2;
    if (MODE == AX_DOT_OTHER) a2 = ld_pack<T, V, LD_DEFAULT>(u2 + e * V);
#pragma unroll
    for (int q = 0; q < V; ++q) vv.v[q] = fnma_rn(cf[q], a1.v[q], vv.v[q]);
    st_pack<T, V, false>(v + e * V, vv);
#pragma unroll
    for (int q = 0; q < V; ++q) {
      if (MODE == AX_NORM) {
        acc[0][q] += (double)vv.v[q] * (double)vv.v[q];
      } else {
        const T uu = (MODE == AX_DOT_SAME) ? a1.v[q] : a2.v[q];
        acc[0][q] += (double)uu * (double)uu;
        acc[1][q] += (double)vv.v[q] * (double)uu;
      }
    }
  }
  cta_reduce_slots<K, V, LPR, NACC>(acc, partial + (size_t)blockIdx.x * NACC * K, smem);
  if (elect_last_cta(&sc->counter)) {
    if (MODE == AX_NORM)
      finalize<T, K, FIN_BETA>(partial, sc, scale_v, nullptr, so, smem);
    else
      finalize<T, K, FIN_COEF>(partial, sc, nullptr, scale_u2, so, smem);
  }
}

// ---- kernel "re-orthogonalisation chain": the whole modified-Gram-Schmidt sweep of one Lanczos step ------------
// The k_axpy chain above costs one launch per visited basis vector, and at small n (an 8 MB panel at n = 32 768)
// each of them is launch-latency-bound (profiles/r01f_ncu_c4_reorth_launches.txt).  When the panel of v fits in the
// shared memory of one CTA per SM, this cooperative kernel keeps every CTA's slice of v on chip for the whole sweep:
//   pass m : v -= coef_{m-1} u_{m-1};  partial dots <u_m,u_m>, <v,u_m>  ->  grid barrier  ->  every CTA sums the
//            per-CTA table in the same fixed order and forms coef_m (the FIN_COEF arithmetic, lanczos.h:59-64);
//   end    : v -= coef_last u_last, v is written back once, ||v||^2 partials, last CTA records beta (FIN_BETA).
// HBM traffic per visited vector drops from (read v, u_{m-1}, u_m + write v) to one read of u_m (u_{m-1} is an L2
// hit), and the per-vector cost from a launch to a grid barrier.  The partial tables alternate between two buffers
// so that a CTA already in pass m+1 never overwrites a table a slower CTA is still summing.
constexpr int CHAIN_MAX_VISIT = 1024;
struct ChainVisit {
  int count;
  uint16_t panel[CHAIN_MAX_VISIT];  // physical panel index of each visited column (0 = caller's X)
};

template <typename T>
__device__ __forceinline__ const T* chain_panel(const T* x0, const char* base, size_t panel_bytes, int ph) {
  return ph == 0 ? x0 : reinterpret_cast<const T*>(base + panel_bytes * (size_t)(ph - 1));
}

template <typename T, int K>
__global__ void __launch_bounds__(NT, 1)
    k_reorth_chain(int64_t n, int64_t chunk /* packs per CTA, multiple of NT */, T* __restrict__ v, const T* x0, const char* base,
                   size_t panel_bytes, const double* scales, LzScal* sc, double* scale_v, double* partial, StepOut so,
                   const __grid_constant__ ChainVisit visit) {
  namespace cg = cooperative_groups;
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  constexpr int NS = 2 * K;
  constexpr int NCH = NT / NS > 0 ? NT / NS : 1;
  extern __shared__ __align__(32) unsigned char s_dyn[];
  Pack<T, V>* s_v = reinterpret_cast<Pack<T, V>*>(s_dyn);
  __shared__ double smem[NT > NWARP * 2 * K ? NT : NWARP * 2 * K];
  __shared__ double s_coef[MAXK];
  cg::grid_group grid = cg::this_grid();

  const int t = threadIdx.x;
  const int slot = (t % LPR) * V;
  const int64_t nvec = n * LPR;
  const int64_t e0 = (int64_t)blockIdx.x * chunk;
  int64_t e1 = e0 + chunk;
  if (e1 > nvec) e1 = nvec;
  const int cnt = e1 > e0 ? (int)(e1 - e0) : 0;  // packs of this CTA; local pack l <-> global pack e0 + l
  const size_t table = (size_t)gridDim.x * NS;

  T cf[V];
#pragma unroll
  for (int i = 0; i < V; ++i) cf[i] = (T)sc->coef[slot + i];  // alpha_j r_c, left by FIN_ALPHA

  const T* u_prev = chain_panel<T>(x0, base, panel_bytes, visit.panel[0]);
  for (int m = 0; m < visit.count; ++m) {
    const int ph = visit.panel[m];
    const T* u_cur = chain_panel<T>(x0, base, panel_bytes, ph);
    double acc[2][V];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int i = 0; i < V; ++i) acc[a][i] = 0.0;
#pragma unroll 4
    for (int l = t; l < cnt; l += NT) {
      const int64_t e = e0 + l;
      Pack<T, V> vv = (m == 0) ? ld_pack<T, V, LD_STREAM>(v + e * V) : s_v[l];
      Pack<T, V> a1 = ld_pack<T, V, LD_DEFAULT>(u_prev + e * V);
      Pack<T, V> a2 = (m == 0) ? a1 : ld_pack<T, V, LD_DEFAULT>(u_cur + e * V);
#pragma unroll
      for (int q = 0; q < V; ++q) vv.v[q] = fnma_rn(cf[q], a1.v[q], vv.v[q]);
      s_v[l] = vv;
#pragma unroll
      for (int q = 0; q < V; ++q) {
        acc[0][q] += (double)a2.v[q] * (double)a2.v[q];
        acc[1][q] += (double)vv.v[q] * (double)a2.v[q];
      }
    }
    double* tab = partial + (size_t)(m & 1) * table;
    cta_reduce_slots<K, V, LPR, 2>(acc, tab + (size_t)blockIdx.x * NS, smem);
    grid.sync();
    // every CTA: column sums of the table in the order reduce_partials uses (chunks of CTAs, then the chunk sums)
    {
      double s = 0.0;
      if (NS <= NT) {
        const int col = t % NS, ch = t / NS;
        if (ch < NCH) {
          int c = ch;
          constexpr int U = 10;  // table loads in flight per thread (148 CTAs / NCH = 37 rows per thread at K = 32)
          for (; c + (U - 1) * NCH < (int)gridDim.x; c += U * NCH) {
            double p[U];
#pragma unroll
            for (int i = 0; i < U; ++i) p[i] = __ldcg(tab + (size_t)(c + i * NCH) * NS + col);
#pragma unroll
            for (int i = 0; i < U; ++i) s += p[i];
          }
          for (; c < (int)gridDim.x; c += NCH) s += __ldcg(tab + (size_t)c * NS + col);
        }
      }
      __syncthreads();  // cta_reduce_slots' readers are done with smem
      smem[t] = s;
      __syncthreads();
      if (t < K) {
        double uu = 0.0, vu = 0.0;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          uu += smem[c * NS + t];
          vu += smem[c * NS + K + t];
        }
        const double r = scales[(size_t)ph * MAXK + t];
        const double u_norm = round_to<T>(uu * r * r);
        const double s_proj = round_to<T>(vu * r);
        const bool take = (u_norm > so.orth_tol) && (fabs(s_proj) > so.orth_tol);
        s_coef[t] = take ? round_to<T>(round_to<T>(s_proj / u_norm) * r) : 0.0;
      }
      __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < V; ++i) cf[i] = (T)s_coef[slot + i];
    u_prev = u_cur;
  }

  // v -= coef_last u_last; write v back; ||v||^2
  double accn[1][V];
#pragma unroll
  for (int i = 0; i < V; ++i) accn[0][i] = 0.0;
#pragma unroll 4
  for (int l = t; l < cnt; l += NT) {
    const int64_t e = e0 + l;
    Pack<T, V> vv = s_v[l];
    Pack<T, V> a1 = ld_pack<T, V, LD_DEFAULT>(u_prev + e * V);
#pragma unroll
    for (int q = 0; q < V; ++q) vv.v[q] = fnma_rn(cf[q], a1.v[q], vv.v[q]);
    st_pack<T, V, false>(v + e * V, vv);
#pragma unroll
    for (int q = 0; q < V; ++q) accn[0][q] += (double)vv.v[q] * (double)vv.v[q];
  }
  __syncthreads();
  double* tabn = partial + 2 * table;  // a third table: slower CTAs may still be summing either pass table
  cta_reduce_slots<K, V, LPR, 1>(accn, tabn + (size_t)blockIdx.x * K, smem);
  if (elect_last_cta(&sc->counter)) finalize<T, K, FIN_BETA>(tabn, sc, scale_v, nullptr, so, smem);
}

// ---- kernel: export one normalised Lanczos vector block into the caller's basis ----------------------------
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_export(int64_t n, const T* __restrict__ u, const double* scale, const LzScal* sc, int m,
                                                T* __restrict__ out) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  const int slot = (threadIdx.x % LPR) * V;
  T r[V];
#pragma unroll
  for (int i = 0; i < V; ++i) r[i] = (m < sc->steps[slot + i] || m == 0) ? (T)scale[slot + i] : (T)0;
  const int64_t nvec = n * LPR;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < nvec; e += (int64_t)gridDim.x * NT) {
    Pack<T, V> x = ld_pack<T, V, LD_STREAM>(u + e * V);
#pragma unroll
    for (int q = 0; q < V; ++q) x.v[q] = (r[q] != (T)0) ? x.v[q] * r[q] : (T)0;
    st_pack<T, V, true>(out + e * V, x);
  }
}

static __global__ void k_copy_steps(const LzScal* sc, int k, int32_t* steps, double* xnorm2) {
  int t = threadIdx.x;
  if (t < k) {
    if (steps) steps[t] = sc->steps[t];
    if (xnorm2) xnorm2[t] = sc->xnorm2[t];
  }
}

// ---- dense symmetric operator: Y = A X ------------------------------------------------------------------------
// Shared-memory tiled panel product (A tile TR x TC staged once per CTA, X chunk TC x K staged alongside).
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_dense_matmat(int64_t n, const T* __restrict__ A, int64_t ld, const T* __restrict__ X,
                                                      T* __restrict__ Y) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  constexpr int RPC = NT / LPR;
  constexpr int RR = RPC >= 128 ? 1 : (128 / RPC > 8 ? 8 : 128 / RPC);
  constexpr int TR = RPC * RR;
  constexpr int TC = 16;
  __shared__ T As[TR][TC + 1];
  __shared__ T Xs[TC][K];
  const int sub = threadIdx.x % LPR, rloc = threadIdx.x / LPR;
  const int slot = sub * V;
  const int64_t row0 = (int64_t)blockIdx.x * TR;
  T acc[RR][V];
#pragma unroll
  for (int a = 0; a < RR; ++a)
#pragma unroll
    for (int q = 0; q < V; ++q) acc[a][q] = (T)0;
  for (int64_t c0 = 0; c0 < n; c0 += TC) {
    for (int idx = threadIdx.x; idx < TR * TC; idx += NT) {
      const int rr = idx / TC, cc = idx % TC;
      const int64_t r = row0 + rr, c = c0 + cc;
      As[rr][cc] = (r < n && c < n) ? A[r * ld + c] : (T)0;
    }
    for (int idx = threadIdx.x; idx < TC * K; idx += NT) {
      const int cc = idx / K, p = idx % K;
      const int64_t c = c0 + cc;
      Xs[cc][p] = (c < n) ? X[c * K + p] : (T)0;
    }
    __syncthreads();
#pragma unroll
    for (int cc = 0; cc < TC; ++cc) {
      T xv[V];
#pragma unroll
      for (int q = 0; q < V; ++q) xv[q] = Xs[cc][slot + q];
#pragma unroll
      for (int a = 0; a < RR; ++a) {
        const T av = As[rloc + a * RPC][cc];
#pragma unroll
        for (int q = 0; q < V; ++q) acc[a][q] = fma(av, xv[q], acc[a][q]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < RR; ++a) {
    const int64_t r = row0 + rloc + a * RPC;
    if (r < n) {
      Pack<T, V> o;
#pragma unroll
      for (int q = 0; q < V; ++q) o.v[q] = acc[a][q];
      st_pack<T, V, false>(Y + r * K + slot, o);
    }
  }
}

// ---- plain CSR SpMM: Y = A X (slq_op_matmat; the roofline probe for the gather itself) -----------------------
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_csr_matmat(int64_t n, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx,
                                                    const T* __restrict__ vals, const T* __restrict__ X, T* __restrict__ Y) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  constexpr int RPC = NT / LPR;
  constexpr int PASSES = 4;
  constexpr int TILE = RPC * PASSES;
  const int sub = threadIdx.x % LPR, rloc = threadIdx.x / LPR;
  const int slot = sub * V;
  const int64_t ntiles = (n + TILE - 1) / TILE;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
      const int64_t r = tile * TILE + ps * RPC + rloc;
      if (r < n) {
        T acc[V];
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = (T)0;
        const int s = __ldg(rowptr + r), e = __ldg(rowptr + r + 1);
#pragma unroll 4
        for (int i = s; i < e; ++i) {
          const int64_t c = __ldg(colidx + i);
          const T a = __ldg(vals + i);
          Pack<T, V> xv = ld_pack<T, V, LD_READONLY>(X + c * K + slot);
#pragma unroll
          for (int q = 0; q < V; ++q) acc[q] = fma(a, xv.v[q], acc[q]);
        }
        Pack<T, V> o;
#pragma unroll
        for (int q = 0; q < V; ++q) o.v[q] = acc[q];
        st_pack<T, V, true>(Y + r * K + slot, o);
      }
    }
  }
}

// ---- host side ------------------------------------------------------------------------------------------------------
struct Grid {
  int stream_ctas;  // persistent grid for the streaming kernels
};

static int grid_for(int64_t work_items /* vector elements or rows */, int per_cta) {
  int sms = sm_count();
  if (sms <= 0) return -1;
  int64_t need = (work_items + per_cta - 1) / per_cta;
  int64_t cap = (int64_t)sms * 4;  // 4 resident CTAs of 256 threads per SM
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

template <typename T, int K>
static int op_apply(slq_op* op, const T* X, T* Y, cudaStream_t st) {
  const int64_t n = op->n;
  if (op->kind == OP_CSR) {
    StepOut so{};
    int rc = launch_stepA<T, K, STEP_PLAIN>(op, X, X, Y, nullptr, nullptr, nullptr, nullptr, so, st);
    if (rc != SLQ_OK) return rc;
  } else if (op->kind == OP_DENSE) {
    if constexpr (dense_tma_supported<T, K>()) {
      if (dense_tma_ok<T, K>(op, X, Y) && use_tma_path()) return launch_dense_tma<T, K>(op, X, Y, st);
    }
    constexpr int RPC = NT / BlockMap<T, K>::LPR;
    constexpr int RR = RPC >= 128 ? 1 : (128 / RPC > 8 ? 8 : 128 / RPC);
    int64_t g = (n + RPC * RR - 1) / (RPC * RR);
    k_dense_matmat<T, K><<<(unsigned)g, NT, 0, st>>>(n, (const T*)op->dense, op->ld, X, Y);
    SLQ_LAUNCH_CHECK();
  } else {
    int rc = op->fn(op->ctx, X, Y, n, K, op->dtype, (slq_stream)st);
    if (rc != 0) return fail(SLQ_ERR_CALLBACK, "operator callback returned %d", rc);
  }
  return SLQ_OK;
}

// SLQ_REORTH=chain|launches forces one of the two re-orthogonalisation paths (A/B timing); default: the chain
// whenever the panel fits on chip.
static int reorth_mode() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("SLQ_REORTH");
    cached = (e && strcmp(e, "launches") == 0) ? 0 : 1;
  }
  return cached;
}

// Returns 1 when the chain ran, 0 when the caller must take the per-vector launches, < 0 on error.
template <typename T, int K>
static int try_reorth_chain(int64_t n, T* v, const T* x0, const Workspace& ws, const std::vector<int>& visit, double* scale_v,
                            const StepOut& so, cudaStream_t st) {
  using M = BlockMap<T, K>;
  if (reorth_mode() == 0 || visit.empty() || (int)visit.size() > CHAIN_MAX_VISIT || ws.npanel + 1 > 65535) return 0;
  const int sms = sm_count();
  if (sms <= 0) return 0;
  const int64_t nvec = n * M::LPR;
  const int64_t per = (nvec + sms - 1) / sms;
  int64_t chunk = (per + NT - 1) / NT * NT;
  if (chunk < NT) chunk = NT;
  const size_t pack_bytes = sizeof(T) * M::V;
  const size_t dyn = (size_t)chunk * pack_bytes;
  // partial tables: two pass tables of grid x 2K doubles and one of grid x K (ws_layout holds 8 * sms x 2 MAXK)
  // Launch configuration cached per (T, K) instantiation AND per device, under a lock (one process may drive
  // several GPUs from several threads).
  struct ChainCfg {
    int max_dyn = -1, coop = -1, occ = 0;
    size_t occ_dyn = (size_t)-1;
    bool configured = false;
  };
  static ChainCfg cfgs[MAX_DEVICES];
  static std::mutex cfg_mu;
  const int dev = current_device();
  if (dev < 0 || dev >= MAX_DEVICES) return 0;
  const int grid = (int)((nvec + chunk - 1) / chunk);  // <= sms
  int occ = 0;
  {
    std::lock_guard<std::mutex> lk(cfg_mu);
    ChainCfg& cfg = cfgs[dev];
    if (cfg.max_dyn < 0) {
      int optin = 0, c = 0;
      cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
      cudaDeviceGetAttribute(&c, cudaDevAttrCooperativeLaunch, dev);
      cudaFuncAttributes fa;
      if (cudaFuncGetAttributes(&fa, k_reorth_chain<T, K>) != cudaSuccess) return 0;
      cfg.coop = c;
      cfg.max_dyn = optin - (int)fa.sharedSizeBytes - 1024;
    }
    if (!cfg.coop || (int64_t)dyn > (int64_t)cfg.max_dyn) return 0;
    if (!cfg.configured) {
      SLQ_CUDA(cudaFuncSetAttribute(k_reorth_chain<T, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, cfg.max_dyn));
      cfg.configured = true;
    }
    if (cfg.occ_dyn != dyn) {  // co-residency of the whole grid is what grid.sync() needs
      SLQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cfg.occ, k_reorth_chain<T, K>, NT, dyn));
      cfg.occ_dyn = dyn;
    }
    occ = cfg.occ;
  }
  if (occ < 1) return 0;
  ChainVisit cv;
  cv.count = (int)visit.size();
  for (size_t i = 0; i < visit.size(); ++i) cv.panel[i] = (uint16_t)visit[i];
  int64_t n_ = n, chunk_ = chunk;
  const char* base = ws.base;
  size_t panel_bytes = ws.panel_bytes;
  const double* scales = ws.scales;
  LzScal* sc = ws.sc;
  double* partial = ws.partial;
  StepOut so_ = so;
  void* args[] = {&n_, &chunk_, &v, &x0, &base, &panel_bytes, &scales, &sc, &scale_v, &partial, &so_, &cv};
  SLQ_CUDA(cudaLaunchCooperativeKernel((const void*)k_reorth_chain<T, K>, dim3(grid), dim3(NT), args, dyn, st));
  SLQ_LAUNCH_CHECK();
  return 1;
}

template <typename T, int K>
static int lanczos_impl(slq_op* op, T* X, int deg, double rtol, int orth, int ncv, T* alpha, T* beta, int32_t* steps,
                        double* xnorm2, T* Qout, char* wsbase, cudaStream_t st) {
  const int64_t n = op->n;
  using M = BlockMap<T, K>;
  Workspace ws;
  ws_layout(n, K, ncv, op->dtype, &ws, wsbase);
  const int ld = deg + 1;
  SLQ_CUDA(cudaMemsetAsync(alpha, 0, sizeof(T) * (size_t)K * ld, st));
  SLQ_CUDA(cudaMemsetAsync(beta, 0, sizeof(T) * (size_t)K * ld, st));
  SLQ_CUDA(cudaMemsetAsync(ws.sc, 0, sizeof(LzScal), st));

  // physical panels: 0 = caller's X, 1..ncv = workspace
  std::vector<T*> phys(ncv + 1);
  phys[0] = X;
  for (int i = 0; i < ncv; ++i) phys[i + 1] = reinterpret_cast<T*>(ws.base + ws.panel_bytes * (size_t)i);
  auto scale_of = [&](int ph) { return ws.scales + (size_t)ph * MAXK; };
  std::vector<int> slot_phys(ncv, -1), slot_vec(ncv, -1);  // logical column -> physical panel, Lanczos vector index
  std::vector<int> free_list;
  for (int i = ncv; i >= 1; --i) free_list.push_back(i);
  slot_phys[0] = 0;
  slot_vec[0] = 0;

  StepOut so;
  so.alpha = alpha;
  so.beta = beta;
  so.ld = ld;
  so.j = 0;
  so.deg = deg;
  so.tol = (double)(T)(std::sqrt((double)n) * (double)(T)rtol);                          // lanczos.h:110 (F residual_tol)
  so.orth_tol = 2.0 * (double)std::numeric_limits<T>::epsilon() * std::sqrt((double)n);  // lanczos.h:53 (a double)

  const int g_vec = grid_for(n * M::LPR, NT * 4);
  const int g_row = grid_for(n, (NT / M::LPR) * 4);
  if (g_vec <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");

  k_init<T, K><<<g_vec, NT, 0, st>>>(n, X, ws.sc, scale_of(0), ws.partial, so);
  SLQ_LAUNCH_CHECK();

  // ---- fused path (orth = 0, CSR rows of <= 64 non-zeros, no basis wanted): ONE kernel per step -------------------------
  // Panels: the pending w of the previous step, its vector u, and this step's two outputs (the formed vector, the new w):
  // the caller's X plus three workspace panels (ws_layout reserves the third when ncv == 2).  Scales rotate through
  // three rows of ws.scales: row j % 3 belongs to vector j.
  if constexpr (tma_supported<T, K>()) {
    const bool aligned = (((uintptr_t)op->colidx | (uintptr_t)op->vals | (uintptr_t)X | (uintptr_t)ws.base) & 15) == 0;
    if (orth == 0 && !Qout && op->kind == OP_CSR && op->max_row_nnz <= TmaTiling<T, K>::LONG && ws.npanel >= 3 && aligned &&
        use_tma_path() && use_fused_step() && !op->no_fused_step) {
      T* P[4] = {X, reinterpret_cast<T*>(ws.base), reinterpret_cast<T*>(ws.base + ws.panel_bytes),
                 reinterpret_cast<T*>(ws.base + 2 * ws.panel_bytes)};
      std::vector<int> freep = {3, 2, 1};
      int iW = 0, iU = 0;  // first step: "pending w" = the start block with cf = 0 (k_init), u = the start block
      for (int j = 0; j < deg; ++j) {
        so.j = j;
        const int iWn = freep.back();
        freep.pop_back();
        int iV = -1;
        if (j > 0) {
          iV = freep.back();
          freep.pop_back();
        }
        const int pslot = prof_begin(PROF_STEPA, st);
        // (first step: beta = 0 multiplies the "previous" scale, which must be a finite number, not an unwritten row)
        int rc = launch_stepA_fused<T, K>(op, P[iW], P[iU], P[iWn], iV >= 0 ? P[iV] : nullptr, ws.sc, scale_of(j % 3),
                                          j > 0 ? scale_of((j + 2) % 3) : scale_of(0), scale_of((j + 1) % 3), ws.partial, so, st);
        if (rc != SLQ_OK) return rc;
        prof_end(PROF_STEPA, pslot, st);
        if (j > 0) {
          freep.push_back(iW);
          freep.push_back(iU);
          iU = iV;
        }
        iW = iWn;
      }
      if (steps || xnorm2) {
        k_copy_steps<<<1, MAXK, 0, st>>>(ws.sc, K, steps, xnorm2);
        SLQ_LAUNCH_CHECK();
      }
      return SLQ_OK;
    }
  }

  int pos[3] = {ncv - 1, 0, 1 % ncv};
  int F = free_list.back();
  free_list.pop_back();
  for (int j = 0; j < deg; ++j) {
    const int p = pos[0], c = pos[1], nx = pos[2];
    so.j = j;
    const int pc = slot_phys[c];
    const int pp = (j == 0) ? -1 : slot_phys[p];
    T* w = phys[F];
    const bool first = (pp < 0);
    const T* upp = first ? phys[pc] : phys[pp];
    const double* scp = first ? scale_of(pc) : scale_of(pp);
    const int pslot = prof_begin(PROF_STEPA, st);
    if (op->kind == OP_CSR) {
      int rc = first ? launch_stepA<T, K, STEP_FIRST>(op, phys[pc], upp, w, ws.sc, scale_of(pc), scp, ws.partial, so, st)
                     : launch_stepA<T, K, STEP_NEXT>(op, phys[pc], upp, w, ws.sc, scale_of(pc), scp, ws.partial, so, st);
      if (rc != SLQ_OK) return rc;
    } else {
      int rc = op_apply<T, K>(op, phys[pc], w, st);
      if (rc != SLQ_OK) return rc;
      if (first)
        k_stepA<T, K, true, false><<<g_row, NT, 0, st>>>(n, nullptr, nullptr, nullptr, phys[pc], upp, w, ws.sc, scale_of(pc), scp,
                                                         ws.partial, so);
      else
        k_stepA<T, K, false, false><<<g_row, NT, 0, st>>>(n, nullptr, nullptr, nullptr, phys[pc], upp, w, ws.sc, scale_of(pc), scp,
                                                          ws.partial, so);
      SLQ_LAUNCH_CHECK();
    }
    prof_end(PROF_STEPA, pslot, st);

    // columns visited by the re-projection: c, c-1, ... cyclically, `orth` of them; unwritten (zero) columns are
    // skipped exactly as the reference's u_norm > tol guard skips them (lanczos.h:58-64, fresh basis).
    std::vector<int> visit;
    for (int t = 0, i = c; t < orth; ++t, i = ((i - 1) % ncv + ncv) % ncv)
      if (slot_phys[i] >= 0) visit.push_back(slot_phys[i]);
    if (visit.empty()) {
      const int bslot = prof_begin(PROF_AXPY, st);
      k_axpy<T, K, AX_NORM><<<g_vec, NT, 0, st>>>(n, w, phys[pc], nullptr, ws.sc, scale_of(F), nullptr, ws.partial, so);
      SLQ_LAUNCH_CHECK();
      prof_end(PROF_AXPY, bslot, st);
    } else if (int ran = try_reorth_chain<T, K>(n, w, X, ws, visit, scale_of(F), so, st); ran != 0) {
      if (ran < 0) return ran;  // the whole sweep ran as one cooperative kernel (v stays in shared memory)
    } else {
      // visit[0] is always the current column: fuse "v = w - alpha q_c" with the first projection's dots
      k_axpy<T, K, AX_DOT_SAME><<<g_vec, NT, 0, st>>>(n, w, phys[pc], nullptr, ws.sc, nullptr, scale_of(visit[0]), ws.partial, so);
      SLQ_LAUNCH_CHECK();
      for (size_t m = 1; m < visit.size(); ++m) {
        k_axpy<T, K, AX_DOT_OTHER><<<g_vec, NT, 0, st>>>(n, w, phys[visit[m - 1]], phys[visit[m]], ws.sc, nullptr,
                                                         scale_of(visit[m]), ws.partial, so);
        SLQ_LAUNCH_CHECK();
      }
      k_axpy<T, K, AX_NORM><<<g_vec, NT, 0, st>>>(n, w, phys[visit.back()], nullptr, ws.sc, scale_of(F), nullptr, ws.partial, so);
      SLQ_LAUNCH_CHECK();
    }
    if (j + 1 < deg) {  // Q.col(n) = v / beta  (lanczos.h:143): the panel just becomes column `nx`
      const int old = slot_phys[nx];
      slot_phys[nx] = F;
      slot_vec[nx] = j + 1;
      if (old >= 0) {
        F = old;
      } else {
        if (free_list.empty()) return fail(SLQ_ERR_ARG, "internal: out of Lanczos panels");
        F = free_list.back();
        free_list.pop_back();
      }
      pos[0] = c;
      pos[1] = nx;
      pos[2] = (j + 2) % ncv;
    }
  }
  if (Qout) {
    SLQ_CUDA(cudaMemsetAsync(Qout, 0, sizeof(T) * (size_t)n * K * ncv, st));
    for (int s = 0; s < ncv; ++s) {
      if (slot_phys[s] < 0) continue;
      k_export<T, K><<<g_vec, NT, 0, st>>>(n, phys[slot_phys[s]], scale_of(slot_phys[s]), ws.sc, slot_vec[s],
                                           Qout + (size_t)s * n * K);
      SLQ_LAUNCH_CHECK();
    }
  }
  if (steps || xnorm2) {
    k_copy_steps<<<1, MAXK, 0, st>>>(ws.sc, K, steps, xnorm2);
    SLQ_LAUNCH_CHECK();
  }
  return SLQ_OK;
}

// ---- basis-free f(A) X: second ("replay") pass ---------------------------------------------------------------------
// Table of the per-step, per-probe scalars of a recorded Lanczos run, in exactly the roundings the first pass used
// (finalize<>: FIN_INIT / FIN_ALPHA / FIN_BETA), so that the replayed vectors are bit-identical to the first pass':
//   row j:  rc  = scale of u_j            (1/||x|| for j = 0, 1/beta_j after)
//           brp = beta_j * scale of u_{j-1}
//           cf  = alpha_j * rc
//           yc  = c_{j+1} ||x|| * scale of u_{j+1}   (0 when vector j+1 does not exist: early break, lanczos.h:140)
//   y0 = c_0 ||x|| * scale of u_0.   Steps a probe never executed get all-zero scalars: its vectors stay 0.
template <typename T>
__global__ void k_replay_table(const T* __restrict__ alpha, const T* __restrict__ beta, int ld, int deg, int k,
                               const int32_t* __restrict__ steps, const double* __restrict__ xnorm2, const double* __restrict__ coeffs,
                               double* __restrict__ table, double* __restrict__ y0) {
  const int j = blockIdx.x, t = threadIdx.x;
  if (t >= MAXK) return;
  double* row = table + (size_t)j * 4 * MAXK;
  if (t >= k) {
    row[t] = row[MAXK + t] = row[2 * MAXK + t] = row[3 * MAXK + t] = 0.0;
    if (j == 0) y0[t] = 0.0;
    return;
  }
  const double z = xnorm2[t];
  const int nst = steps[t];
  auto scale_of = [&](int m) -> double {  // reciprocal scale of the un-normalised Lanczos vector u_m
    if (m == 0) return (z > 0.0) ? round_to<T>(1.0 / round_to<T>(sqrt(round_to<T>(z)))) : 1.0;
    return round_to<T>(1.0 / (double)beta[(size_t)t * ld + m]);
  };
  const double xn = sqrt(z);
  if (j == 0) y0[t] = coeffs[(size_t)t * deg] * xn * scale_of(0);
  if (j >= nst) {
    row[t] = row[MAXK + t] = row[2 * MAXK + t] = row[3 * MAXK + t] = 0.0;
    return;
  }
  const double rc = scale_of(j);
  row[t] = rc;
  row[MAXK + t] = (j == 0) ? 0.0 : (double)((T)beta[(size_t)t * ld + j] * (T)scale_of(j - 1));
  row[2 * MAXK + t] = round_to<T>((double)alpha[(size_t)t * ld + j] * rc);
  row[3 * MAXK + t] = (j + 1 < deg && j + 1 < nst) ? coeffs[(size_t)t * deg + j + 1] * xn * scale_of(j + 1) : 0.0;
}

// Generic replay step (operators without the fused kernel: dense, callback, narrow or unaligned CSR): `w` holds A u_c on
// entry (the operator's product); v and Y are formed in the arithmetic of k_stepA<.., FUSED_CSR = false> + k_axpy.
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_replay_axpy(int64_t n, T* __restrict__ w, const T* __restrict__ uc, const T* __restrict__ up,
                                                     T* __restrict__ Y, ReplayArgs rp) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  const int slot = (threadIdx.x % LPR) * V;
  T rc[V], brp[V], cf[V], yc[V], y0[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    rc[i] = (T)rp.row[slot + i];
    brp[i] = (T)rp.row[MAXK + slot + i];
    cf[i] = (T)rp.row[2 * MAXK + slot + i];
    yc[i] = (T)rp.row[3 * MAXK + slot + i];
    y0[i] = rp.first ? (T)rp.y0[slot + i] : (T)0;
  }
  const int64_t nvec = n * LPR;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < nvec; e += (int64_t)gridDim.x * NT) {
    Pack<T, V> wv = ld_pack<T, V, LD_STREAM>(w + e * V);
    const Pack<T, V> qc = ld_pack<T, V, LD_DEFAULT>(uc + e * V);
    Pack<T, V> pv, yv;
    if (!rp.first) {
      pv = ld_pack<T, V, LD_STREAM>(up + e * V);
      yv = ld_pack<T, V, LD_STREAM>(Y + e * V);
    }
#pragma unroll
    for (int q = 0; q < V; ++q) {
      const T t1 = rp.first ? mul_rn(wv.v[q], rc[q]) : fnma_rn(brp[q], pv.v[q], mul_rn(wv.v[q], rc[q]));
      wv.v[q] = fnma_rn(cf[q], qc.v[q], t1);
      const T y_old = rp.first ? mul_rn(y0[q], qc.v[q]) : yv.v[q];
      yv.v[q] = fma_rn(yc[q], wv.v[q], y_old);
    }
    st_pack<T, V, false>(w + e * V, wv);
    st_pack<T, V, true>(Y + e * V, yv);
  }
}

// Y = y0 * X (deg == 1: only the start vector contributes)
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_replay_first_only(int64_t n, const T* __restrict__ x, T* __restrict__ Y, const double* y0) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  const int slot = (threadIdx.x % LPR) * V;
  const int64_t nvec = n * LPR;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < nvec; e += (int64_t)gridDim.x * NT) {
    Pack<T, V> xv = ld_pack<T, V, LD_STREAM>(x + e * V);
#pragma unroll
    for (int q = 0; q < V; ++q) xv.v[q] = mul_rn((T)y0[slot + q], xv.v[q]);
    st_pack<T, V, true>(Y + e * V, xv);
  }
}

// Replays steps 0 .. deg-2 of the recorded run from the intact start block X, accumulating Y = sum_m c_m ||x|| q_m
// (operators.py:124 without Q).  P0..P2: three scratch panels; `table`/`y0` from k_replay_table.
template <typename T, int K>
static int replay_impl(slq_op* op, const T* X, int deg, const double* table, const double* y0, T* Y, T* P0, T* P1, T* P2,
                       cudaStream_t st) {
  const int64_t n = op->n;
  using M = BlockMap<T, K>;
  const int g_vec = grid_for(n * M::LPR, NT * 4);
  if (g_vec <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  if (deg <= 1) {
    k_replay_first_only<T, K><<<g_vec, NT, 0, st>>>(n, X, Y, y0);
    SLQ_LAUNCH_CHECK();
    return SLQ_OK;
  }
  T* P[3] = {P0, P1, P2};
  bool fused = false;
  if constexpr (tma_supported<T, K>()) {
    fused = op->kind == OP_CSR && use_tma_path() &&
            (((uintptr_t)op->colidx | (uintptr_t)op->vals | (uintptr_t)X | (uintptr_t)Y | (uintptr_t)P0 | (uintptr_t)P1 | (uintptr_t)P2) & 15) == 0;
  }
  StepOut so{};
  for (int j = 0; j + 1 < deg; ++j) {
    const T* uc = (j == 0) ? X : P[(j - 1) % 3];
    const T* up = (j <= 1) ? X : P[(j - 2) % 3];  // unused at j = 0 (brp = 0): any finite panel
    T* w = P[j % 3];
    ReplayArgs rp;
    rp.row = table + (size_t)j * 4 * MAXK;
    rp.y0 = y0;
    rp.Y = Y;
    rp.first = (j == 0);
    const int pslot = prof_begin(PROF_STEPA, st);
    if (fused) {
      if constexpr (tma_supported<T, K>()) {
        int rc = launch_stepA_tma<T, K, STEP_REPLAY>(op, uc, up, w, nullptr, nullptr, nullptr, nullptr, so, st, rp);
        if (rc != SLQ_OK) return rc;
      }
    } else {
      int rc = op_apply<T, K>(op, uc, w, st);
      if (rc != SLQ_OK) return rc;
      k_replay_axpy<T, K><<<g_vec, NT, 0, st>>>(n, w, uc, up, Y, rp);
      SLQ_LAUNCH_CHECK();
    }
    prof_end(PROF_STEPA, pslot, st);
  }
  return SLQ_OK;
}

#define SLQ_DISPATCH_K(T, k, CALL)                        \
  switch (k) {                                            \
    case 1: { constexpr int KK = 1; return CALL; }        \
    case 2: { constexpr int KK = 2; return CALL; }        \
    case 4: { constexpr int KK = 4; return CALL; }        \
    case 8: { constexpr int KK = 8; return CALL; }        \
    case 16: { constexpr int KK = 16; return CALL; }      \
    case 32: { constexpr int KK = 32; return CALL; }      \
    case 64: { constexpr int KK = 64; return CALL; }      \
    default: return fail(SLQ_ERR_ARG, "block width k=%d must be a power of two in [1,64]", k); \
  }

template <typename T>
int lanczos_dispatch(slq_op* op, void* X, int k, int deg, double rtol, int orth, int ncv, void* alpha, void* beta,
                            int32_t* steps, double* xnorm2, void* Q, char* ws, cudaStream_t st) {
  SLQ_DISPATCH_K(T, k, (lanczos_impl<T, KK>(op, (T*)X, deg, rtol, orth, ncv, (T*)alpha, (T*)beta, steps, xnorm2, (T*)Q, ws, st)));
}

template <typename T>
int matmat_dispatch(slq_op* op, const void* X, void* Y, int k, cudaStream_t st) {
  SLQ_DISPATCH_K(T, k, (op_apply<T, KK>(op, (const T*)X, (T*)Y, st)));
}

template <typename T>
int replay_dispatch(slq_op* op, const void* X, int k, int deg, const void* alpha, const void* beta, const int32_t* steps,
                    const double* xnorm2, const double* coeffs, void* Y, double* table, double* y0, void* P0, void* P1, void* P2,
                    cudaStream_t st) {
  k_replay_table<T><<<deg, MAXK, 0, st>>>((const T*)alpha, (const T*)beta, deg + 1, deg, k, steps, xnorm2, coeffs, table, y0);
  SLQ_LAUNCH_CHECK();
  SLQ_DISPATCH_K(T, k, (replay_impl<T, KK>(op, (const T*)X, deg, table, y0, (T*)Y, (T*)P0, (T*)P1, (T*)P2, st)));
}


}  // namespace slq
