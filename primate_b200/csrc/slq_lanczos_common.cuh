// Declarations shared by the Lanczos translation units (driver templates are instantiated per scalar type in
// slq_lanczos_f64.cu / slq_lanczos_f32.cu so that the two halves compile in parallel).
#pragma once
#include <cmath>
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
  size_t off = 0;
  size_t panels_off = off;
  off += panel * (size_t)ncv;
  size_t sc_off = off;
  off += align_up(sizeof(LzScal), 256);
  size_t scales_off = off;
  off += align_up(sizeof(double) * MAXK * (size_t)(ncv + 1), 256);
  size_t partial_off = off;
  int sms = sm_count();
  if (sms <= 0) sms = 148;
  off += align_up(sizeof(double) * 2 * MAXK * (size_t)sms * 8, 256);
  if (ws) {
    ws->base = base + panels_off;
    ws->panel_bytes = panel;
    ws->npanel = ncv;
    ws->sc = reinterpret_cast<LzScal*>(base + sc_off);
    ws->scales = reinterpret_cast<double*>(base + scales_off);
    ws->partial = reinterpret_cast<double*>(base + partial_off);
  }
  return off;
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
