// Declarations shared by the Lanczos translation units (driver templates are instantiated per scalar type in
// slq_lanczos_f64.cu / slq_lanczos_f32.cu so that the two halves compile in parallel).
#pragma once
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#include "slq_internal.cuh"

namespace slq {

constexpr int MAXK = 64;

// Per-block scalar state in device memory (one per workspace).
struct LzScal {
  double coef[MAXK];   // pending "v -= coef * u" coefficient (already multiplied by the column's scale)
  double beta[MAXK];   // beta_j of the current step (rounded to the panel dtype)
  double xnorm2[MAXK];
  int active[MAXK];    // 0 once the probe hit the early break (lanczos.h:140) or finished
  int steps[MAXK];
  unsigned int counter;
};

struct StepOut {   // where the finalising CTA records the tridiagonal entries
  void* alpha;     // k x ld
  void* beta;      // k x ld
  int ld;
  int j;           // step index
  int deg;
  double tol;      // sqrt(n) * rtol
  double orth_tol; // 2 eps sqrt(n)
};

struct Workspace {
  char* base;
  size_t panel_bytes;
  int npanel;
  LzScal* sc;
  double* scales;   // (npanel + 1) x MAXK
  double* partial;  // max_ctas x 2 x MAXK
};

inline size_t ws_layout(int64_t n, int k, int ncv, int dtype, Workspace* ws, char* base) {
  size_t panel = align_up((size_t)n * k * dtype_size(dtype), 256);
  // ncv == 2 (orth <= 2): one panel more than the recurrence needs, for the fused step's second output (lanczos_impl)
  const int np = ncv + (ncv == 2 ? 1 : 0);
  size_t off = 0;
  size_t panels_off = off;
  off += panel * (size_t)np;
  size_t sc_off = off;
  off += align_up(sizeof(LzScal), 256);
  size_t scales_off = off;
  off += align_up(sizeof(double) * MAXK * (size_t)(np + 1), 256);
  size_t partial_off = off;
  int sms = sm_count();
  if (sms <= 0) sms = 148;
  off += align_up(sizeof(double) * 2 * MAXK * (size_t)sms * 8, 256);
  if (ws) {
    ws->base = base + panels_off;
    ws->panel_bytes = panel;
    ws->npanel = np;
    ws->sc = reinterpret_cast<LzScal*>(base + sc_off);
    ws->scales = reinterpret_cast<double*>(base + scales_off);
    ws->partial = reinterpret_cast<double*>(base + partial_off);
  }
  return off;
}

// ---- dense operator tiling shared by the kernels (slq_dense.cuh) and the band-major copy (slq_dense_set_tiled) --------
constexpr int DENSE_KC = 16;  // columns of A (rows of X) per pipeline stage

// SLQ_DENSE=fma keeps the scalar-FMA kernel for fp64 too (A/B timing and parity of the two summation orders).
inline bool dense_use_dmma() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("SLQ_DENSE");
    cached = (e && strcmp(e, "fma") == 0) ? 0 : 1;
  }
  return cached == 1;
}

// CTA height of the dense kernels in units of 16 rows.  The SM with the most row tiles sets the kernel time, so minimise
// (tiles per SM, rounded up) x (rows per tile) over the compiled heights 6 / 7 / 8 (two CTAs per SM); ties go to the
// taller tile (fewer X slabs re-read from L2).  The fp64 tensor-core kernel also has a 224-row CTA (one per SM, 1.8 KB
// contiguous per matrix row) which is taken when it balances at least as well.  SLQ_DENSE_RPT overrides (A/B timing).
inline int dense_pick_rpt(int64_t npad, bool dmma) {
  const int sms = sm_count() > 0 ? sm_count() : 148;
  int best = 8;
  int64_t best_cost = INT64_MAX;
  for (int rpt = 8; rpt >= 6; --rpt) {
    const int64_t tiles = (npad + 16 * rpt - 1) / (16 * rpt);
    const int64_t cost = ((tiles + sms - 1) / sms) * rpt;
    if (cost < best_cost) {
      best_cost = cost;
      best = rpt;
    }
  }
  if (dmma) {
    const int64_t tall = (npad + 223) / 224;
    if (((tall + sms - 1) / sms) * 14 <= best_cost) best = 14;
  }
  const char* e = getenv("SLQ_DENSE_RPT");
  if (e) {
    const int v = atoi(e);
    if ((v >= 6 && v <= 8) || (dmma && v == 14)) best = v;
  }
  return best;
}

// Band-major copy: band b = rows [b TR, (b + 1) TR) of the (symmetric) operator, TR = 16 rpt; within a band the chunks
// c = columns [16 c, 16 c + 16) follow each other, each stored as the shared-memory image of the tensor-core kernel's
// stage: 16 lines of TRP = TR + 4 doubles, line kk = A[16 c + kk][b TR .. b TR + TR) (zeros past npad and in the pad).
inline size_t dense_tiled_elems(int64_t npad, int rpt) {
  const int64_t tr = 16 * rpt;
  return (size_t)((npad + tr - 1) / tr) * (size_t)(npad / DENSE_KC) * (size_t)DENSE_KC * (size_t)(tr + 4);
}

// instantiated for float and double in slq_lanczos_f32.cu / slq_lanczos_f64.cu
template <typename T>
int lanczos_dispatch(slq_op* op, void* X, int k, int deg, double rtol, int orth, int ncv, void* alpha, void* beta, int32_t* steps,
                     double* xnorm2, void* Q, char* ws, cudaStream_t st);
template <typename T>
int matmat_dispatch(slq_op* op, const void* X, void* Y, int k, cudaStream_t st);
template <typename T>
int replay_dispatch(slq_op* op, const void* X, int k, int deg, const void* alpha, const void* beta, const int32_t* steps,
                    const double* xnorm2, const double* coeffs, void* Y, double* table, double* y0, void* P0, void* P1, void* P2,
                    cudaStream_t st);

}  // namespace slq
