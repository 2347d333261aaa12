/*
 * slq_oracle.c -- CPU restatement of primate's stochastic-Lanczos-quadrature hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker for the CUDA library in
 * primate_b200/csrc.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product package never links, imports or calls it.
 *
 * Every function cites the reference file:line (relative to /root/reference) whose algorithm it
 * restates.  Nothing here is copied: Eigen expressions are re-expressed as plain scalar loops
 * (sequential summation; the reference's Eigen reductions use a different, vectorised order, so
 * agreement with the compiled reference is ~1e-13 relative, cf. SURVEY.md A.4 -- checked in
 * tests/test_oracle.py against oracle/_ref and against tests/golden/).
 *
 * Third-party arithmetic restated from its published algorithm (not present under /root/reference):
 *   - NumPy (container: 2.3.5; reference pins bare "numpy", pyproject.toml:28-31) Generator(PCG64):
 *     PCG XSL-RR 128/64 (O'Neill 2014), multiplier 0x2360ED051FC65DA44385DF649FCCF645, "step then
 *     output", double = (raw >> 11) * 2^-53.  Call sites: src/primate/random.py:23, trace.py:86.
 *
 * Build: gcc -O2 -ffp-contract=off -std=c11 -shared -fPIC slq_oracle.c -o _build/liboracle_slq.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------------------------------------
 * PCG64 (NumPy's default BitGenerator) -- used by src/primate/random.py:22-29 (`_rademacher_inplace`)
 * ---------------------------------------------------------------------------------------------- */
static const uint64_t PCG_MULT_HI = 0x2360ED051FC65DA4ULL;
static const uint64_t PCG_MULT_LO = 0x4385DF649FCCF645ULL;

static inline u128 mk128(uint64_t hi, uint64_t lo) { return ((u128)hi << 64) | (u128)lo; }

static inline uint64_t pcg_output(u128 s) {
	const uint64_t hi = (uint64_t)(s >> 64), lo = (uint64_t)s;
	const uint64_t x = hi ^ lo;
	const unsigned rot = (unsigned)(hi >> 58);
	return (x >> rot) | (x << ((64u - rot) & 63u));
}

/* Brown's O(log delta) LCG jump-ahead, as in PCG64.advance(delta). */
static u128 pcg_advance(u128 state, u128 inc, u128 delta) {
	u128 cur_mult = mk128(PCG_MULT_HI, PCG_MULT_LO), cur_plus = inc;
	u128 acc_mult = 1, acc_plus = 0;
	while (delta > 0) {
		if (delta & 1) {
			acc_mult *= cur_mult;
			acc_plus = acc_plus * cur_mult + cur_plus;
		}
		cur_plus = (cur_mult + 1) * cur_plus;
		cur_mult *= cur_mult;
		delta >>= 1;
	}
	return acc_mult * state + acc_plus;
}

/* state/inc are given as (hi, lo) 64-bit halves of PCG64().state["state"]["state"|"inc"].
 * Writes raw outputs #offset .. #offset+count-1 of that stream (0-based: output #0 is the first
 * value Generator.random_raw() would return). */
void orc_pcg64_raw(const uint64_t st[2], const uint64_t inc[2], uint64_t off_hi, uint64_t off_lo,
                   int64_t count, uint64_t* out) {
	const u128 mult = mk128(PCG_MULT_HI, PCG_MULT_LO), c = mk128(inc[0], inc[1]);
	u128 s = pcg_advance(mk128(st[0], st[1]), c, mk128(off_hi, off_lo));
	for (int64_t i = 0; i < count; ++i) {
		s = s * mult + c; /* step ... */
		out[i] = pcg_output(s); /* ... then output */
	}
}

/* random.py:22-29: u = rng.random(); out = floor(2u)*2 - 1  (u = (raw>>11) * 2^-53). */
void orc_rademacher(const uint64_t st[2], const uint64_t inc[2], uint64_t off_hi, uint64_t off_lo,
                    int64_t count, double* out) {
	const u128 mult = mk128(PCG_MULT_HI, PCG_MULT_LO), c = mk128(inc[0], inc[1]);
	u128 s = pcg_advance(mk128(st[0], st[1]), c, mk128(off_hi, off_lo));
	for (int64_t i = 0; i < count; ++i) {
		s = s * mult + c;
		const double u = (double)(pcg_output(s) >> 11) * (1.0 / 9007199254740992.0);
		out[i] = floor(u * 2.0) * 2.0 - 1.0;
	}
}

/* ------------------------------------------------------------------------------------------------
 * Operators (include/eigen_operators.h)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
	int kind; /* 0 = sparse CSC (eigen_operators.h:59-104), 1 = dense col-major (:18-55) */
	int64_t n;
	const int64_t* indptr; /* CSC column pointers (n+1) */
	const int32_t* indices; /* CSC row indices */
	const void* data; /* values (sparse) or n*n col-major (dense) */
} orc_op;

/* eigen_operators.h:66-77 -> Eigen col-major sparse * dense vector
 * (extern/eigen/Eigen/src/SparseCore/SparseDenseProduct.h:97-118): y = 0; for each column j,
 * y[row] += val * x[j] in storage order. */
#define DEFINE_MATVEC(F, SUF)                                                                      \
	static void matvec_##SUF(const orc_op* A, const F* x, F* y) {                                  \
		const int64_t n = A->n;                                                                    \
		if (A->kind == 0) {                                                                        \
			const F* val = (const F*)A->data;                                                      \
			for (int64_t i = 0; i < n; ++i) y[i] = (F)0;                                           \
			for (int64_t j = 0; j < n; ++j) {                                                      \
				const F xj = x[j];                                                                 \
				for (int64_t p = A->indptr[j]; p < A->indptr[j + 1]; ++p)                          \
					y[A->indices[p]] += val[p] * xj;                                               \
			}                                                                                      \
		} else { /* eigen_operators.h:24-30: y = A x, dense; column-major axpy order */            \
			const F* M = (const F*)A->data;                                                        \
			for (int64_t i = 0; i < n; ++i) y[i] = (F)0;                                           \
			for (int64_t j = 0; j < n; ++j) {                                                      \
				const F xj = x[j];                                                                 \
				const F* col = M + j * n;                                                          \
				for (int64_t i = 0; i < n; ++i) y[i] += col[i] * xj;                               \
			}                                                                                      \
		}                                                                                          \
	}
DEFINE_MATVEC(double, f64)
DEFINE_MATVEC(float, f32)

/* ------------------------------------------------------------------------------------------------
 * orth_vector (include/lanczos.h:43-66) and lanczos_recurrence (include/lanczos.h:92-149)
 * ---------------------------------------------------------------------------------------------- */
static inline int pymod(int a, int b) { return (b + (a % b)) % b; } /* lanczos.h:34-36 */

#define DEFINE_LANCZOS(F, SUF, EPS, SQRT, FABS)                                                    \
	/* lanczos.h:43-66: MGS of v against p columns of U (n x m, col-major) from start_idx. */      \
	static void orth_vector_##SUF(F* v, const F* U, int64_t n, int m, int start_idx, int p,       \
	                              int reverse) {                                                   \
		const F tol = (F)2 * (F)EPS * SQRT((F)n); /* :53 */                                       \
		const int diff = reverse ? -1 : 1;                                                         \
		int i = pymod(start_idx, m);                                                               \
		for (int c = 0; c < p; ++c, i = pymod(i + diff, m)) { /* :58 */                            \
			const F* u = U + (int64_t)i * n;                                                       \
			F u_norm = 0, s_proj = 0;                                                              \
			for (int64_t r = 0; r < n; ++r) u_norm += u[r] * u[r]; /* :59 */                       \
			for (int64_t r = 0; r < n; ++r) s_proj += v[r] * u[r]; /* :60 */                       \
			if (u_norm > tol && FABS(s_proj) > tol) { /* :62 */                                    \
				const F coef = s_proj / u_norm;                                                    \
				for (int64_t r = 0; r < n; ++r) v[r] -= coef * u[r]; /* :63 */                     \
			}                                                                                      \
		}                                                                                          \
	}                                                                                              \
	/* lanczos.h:92-149.  q (n) is workspace and is clobbered; alpha/beta have deg+1 entries;      \
	 * V is n x ncv col-major.  Returns the number of steps j performed (break index + 1). */      \
	int orc_lanczos_##SUF(const orc_op* A, F* q, int deg, F rtol, int orth, F* alpha, F* beta,     \
	                      F* V, int ncv) {                                                         \
		const int64_t n = A->n;                                                                    \
		const F residual_tol = SQRT((F)n) * rtol; /* :110 */                                       \
		int pos[3] = {ncv - 1, 0, 1}; /* :117 */                                                   \
		F* v = q;                                                                                  \
		for (int64_t r = 0; r < n; ++r) V[(int64_t)pos[0] * n + r] = (F)0; /* :118 */              \
		{                                                                                          \
			F nrm = 0; /* :119  Q.col(0) = v.normalized() */                                      \
			for (int64_t r = 0; r < n; ++r) nrm += v[r] * v[r];                                    \
			nrm = SQRT(nrm);                                                                       \
			/* Eigen's normalized(): v / norm if squaredNorm > 0 else v (Core/Dot.h:123) */         \
			for (int64_t r = 0; r < n; ++r) V[r] = (nrm > (F)0) ? v[r] / nrm : v[r];              \
		}                                                                                          \
		beta[0] = (F)0; /* :120 */                                                                 \
		int j = 0;                                                                                 \
		for (; j < deg; ++j) {                                                                     \
			const int p = pos[0], c = pos[1], nx = pos[2];                                         \
			const F *Qp = V + (int64_t)p * n, *Qc = V + (int64_t)c * n;                            \
			matvec_##SUF(A, Qc, v); /* :127 */                                                     \
			for (int64_t r = 0; r < n; ++r) v[r] -= beta[j] * Qp[r]; /* :128 */                    \
			F a = 0;                                                                               \
			for (int64_t r = 0; r < n; ++r) a += Qc[r] * v[r]; /* :129 */                          \
			alpha[j] = a;                                                                          \
			for (int64_t r = 0; r < n; ++r) v[r] -= a * Qc[r]; /* :130 */                          \
			if (orth > 0) orth_vector_##SUF(v, V, n, ncv, c, orth, 1); /* :133-136 */              \
			F b = 0;                                                                               \
			for (int64_t r = 0; r < n; ++r) b += v[r] * v[r];                                      \
			b = SQRT(b);                                                                           \
			beta[j + 1] = b; /* :139 */                                                            \
			if (b < residual_tol || (j + 1) == deg) { ++j; break; } /* :140-142 */                 \
			F* Qn = V + (int64_t)nx * n;                                                           \
			for (int64_t r = 0; r < n; ++r) Qn[r] = v[r] / b; /* :143 */                           \
			pos[0] = c; pos[1] = nx; pos[2] = pymod(j + 2, ncv); /* :146-147 */                    \
		}                                                                                          \
		return j;                                                                                  \
	}
DEFINE_LANCZOS(double, f64, 2.220446049250313e-16, sqrt, fabs)
DEFINE_LANCZOS(float, f32, 1.1920928955078125e-07f, sqrtf, fabsf)

/* ------------------------------------------------------------------------------------------------
 * tqli (src/primate/tqli.py:15-90) incl. its non-textbook `sign` (:5-7): int(b>1) - int(a<0) + 1
 * d, e: length n (e[0] == 0 on entry); Z: n x n row-major or NULL.  In place.
 * ---------------------------------------------------------------------------------------------- */
static inline double tqli_sign(double a, double b) { return (double)((b > 1.0) - (a < 0.0) + 1); }

void orc_tqli(double* d, double* e, double* Z, int n, int max_iter) {
	for (int i = 0; i + 1 < n; ++i) e[i] = e[i + 1]; /* :38 */
	e[n - 1] = 0.0; /* :39 */
	for (int it = 0; it < n; ++it) { /* :40 */
		int ii = 0, m = it;
		for (;;) {
			for (int ml = it; ml < n - 1; ++ml) { /* :45-51 */
				m = ml;
				const double dd = fabs(d[m]) + fabs(d[m + 1]);
				if (fabs(e[m]) + dd == dd) break;
				m += 1;
			}
			if (m != it) {
				if (ii > max_iter || e[it] == 0.0) break; /* :57-59 */
				ii += 1;
				double g = (d[it + 1] - d[it]) / (2.0 * e[it]); /* :61 */
				double r = hypot(g, 1.0);
				g = d[m] - d[it]; /* :63 */
				g += e[it] / (g + tqli_sign(r, g)); /* :64 */
				double s = 1.0, c = 1.0, p = 0.0;
				int i = m - 1;
				int broke = 0;
				for (; i >= it; --i) { /* :67 */
					double f = s * e[i], b = c * e[i];
					e[i + 1] = r = hypot(f, g);
					if (r == 0.0) { /* :71-75 */
						d[i + 1] -= p;
						e[m] = 0.0;
						broke = 1;
						break;
					}
					s = f / r;
					c = g / r;
					g = d[i + 1] - p;
					r = (d[i] - g) * s + 2.0 * c * b;
					p = s * r;
					d[i + 1] = g + p;
					g = c * r - b;
					if (Z != NULL) { /* :82-86 */
						for (int k = 0; k < n; ++k) {
							f = Z[k * n + i + 1];
							Z[k * n + i + 1] = s * Z[k * n + i] + c * f;
							Z[k * n + i] = c * Z[k * n + i] - s * f;
						}
					}
				}
				if (!broke) i = it; /* Python's loop variable keeps its last value */
				if (r == 0.0 && i >= it) continue; /* :87-88 */
				d[it] -= p;
				e[it] = g;
				e[m] = 0.0;
			}
			if (m == it) break; /* :92-93 */
		}
	}
}

/* ------------------------------------------------------------------------------------------------
 * fttr / ortho_poly (src/primate/fttr.py:5-29)
 * ---------------------------------------------------------------------------------------------- */
static void ortho_poly(double x, double mu_sqrt_rec, const double* a, const double* b, double* z, int n) {
	z[0] = mu_sqrt_rec;
	if (n > 1) z[1] = (x - a[0]) * z[0] / b[1];
	for (int i = 2; i < n; ++i) {
		const double s = (x - a[i - 1]) / b[i];
		const double t = -b[i - 1] / b[i];
		z[i] = s * z[i - 1] + t * z[i - 2];
	}
}

void orc_fttr(const double* theta, const double* alpha, const double* beta, int n, int k, double* weights) {
	double mu_0 = 0.0;
	for (int i = 0; i < k; ++i) mu_0 += fabs(theta[i]); /* :24 */
	const double mu_sqrt_rec = 1.0 / sqrt(mu_0);
	double* p = (double*)calloc((size_t)n, sizeof(double));
	for (int i = 0; i < k; ++i) {
		ortho_poly(theta[i], mu_sqrt_rec, alpha, beta, p, n);
		double ss = 0.0;
		for (int j = 0; j < n; ++j) ss += p[j] * p[j];
		weights[i] = (1.0 / ss) / mu_0; /* :28-29 */
	}
	free(p);
}

/* ------------------------------------------------------------------------------------------------
 * One SLQ sample exactly as MatrixFunction.quad drives it (src/primate/operators.py:145-150) but with
 * the tridiagonal eigenproblem solved by tqli (the reference's fallback, tridiag.py:19-20) so that this
 * C port is self-contained (no LAPACK).  The Python side of the oracle uses SciPy/stemr like the
 * reference's default path; this entry is what bench.py times as the "port" baseline.
 * fun_id: 0 identity, 1 abs, 2 sqrt, 3 log(max(x,eps)), 4 inv, 5 exp(t x), 6 smoothstep(a,b),
 *         7 numrank(threshold)   (src/primate/special.py:78-107)
 * ---------------------------------------------------------------------------------------------- */
static double apply_fun(int fun_id, double x, const double* prm) {
	switch (fun_id) {
	case 0: return x;
	case 1: return fabs(x);
	case 2: return sqrt(x);
	case 3: return log(x > 2.220446049250313e-16 ? x : 2.220446049250313e-16);
	case 4: return 1.0 / x;
	case 5: return exp(prm[0] * x);
	case 6: {
		const double a = prm[0], b = prm[1], dd = (a != b) ? (b - a) : 1.0;
		double y = (x - a) / dd;
		y = y < 0.0 ? 0.0 : (y > 1.0 ? 1.0 : y);
		return 3.0 * y * y - 2.0 * y * y * y;
	}
	case 7: return (fabs(x) < prm[0]) ? 0.0 : 1.0;
	default: return NAN;
	}
}

double orc_slq_quad_f64(const orc_op* A, const double* x, int deg, double rtol, int orth, int ncv,
                        int fun_id, const double* prm, double* alpha, double* beta, double* theta,
                        double* tau) {
	const int64_t n = A->n;
	double* q = (double*)malloc((size_t)n * sizeof(double));
	double* V = (double*)calloc((size_t)n * (size_t)ncv, sizeof(double)); /* fresh-Q semantics */
	double* Z = (double*)calloc((size_t)deg * (size_t)deg, sizeof(double));
	double* e = (double*)malloc((size_t)deg * sizeof(double));
	memcpy(q, x, (size_t)n * sizeof(double));
	double nrm2 = 0.0;
	for (int64_t r = 0; r < n; ++r) nrm2 += x[r] * x[r];
	const double nrm = sqrt(nrm2);
	memset(alpha, 0, (size_t)(deg + 1) * sizeof(double));
	memset(beta, 0, (size_t)(deg + 1) * sizeof(double));
	orc_lanczos_f64(A, q, deg, rtol, orth, alpha, beta, V, ncv);
	for (int i = 0; i < deg; ++i) { theta[i] = alpha[i]; e[i] = beta[i]; Z[i * deg + i] = 1.0; }
	orc_tqli(theta, e, Z, deg, 30);
	double acc = 0.0;
	for (int i = 0; i < deg; ++i) {
		tau[i] = Z[i] * Z[i]; /* integrate.py:64: first row of the eigenvector matrix, squared */
		acc += apply_fun(fun_id, theta[i], prm) * tau[i];
	}
	free(q); free(V); free(Z); free(e);
	return acc * (nrm * nrm); /* operators.py:147,150: np.linalg.norm(xc)**2 */
}
