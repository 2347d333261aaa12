// Device-side NORMAL and SPHERE probes, bit-exact with NumPy's Generator.standard_normal as the reference uses it
// (src/primate/random.py:32-41: rng.standard_normal(out=W) on an F-ordered (N,k) array, filled in memory order;
// sphere = normals * sqrt(N) / ||column||).
//
// NumPy's sampler (numpy/random/src/distributions/distributions.c: random_standard_normal, pinned to NumPy 2.3.5) is a
// 256-layer ziggurat on the raw PCG64 stream:
//     r = next_uint64; idx = r & 0xff; sign = (r >> 8) & 1; rabs = (r >> 9) & (2^52 - 1); x = +-rabs * wi[idx]
//     rabs < ki[idx]  -> x                                   (one raw draw: 97.8% of the samples)
//     idx == 0        -> tail: loop { xx = -log1p(-U1)/r; yy = -log1p(-U2) } until 2 yy > xx^2;  +-(r + xx)
//     otherwise       -> wedge: accept x if (fi[idx-1] - fi[idx]) U + fi[idx] < exp(-x^2/2), else start over
// so the NUMBER of raw draws a sample consumes is data dependent, and the stream position of sample i is a
// sequential quantity.  It is parallelised here with a self-synchronisation property of that chain:
//
//   if the 32 raw positions before p are all "fast" (rabs < ki[idx]), then p is the start of a sample
//
// (any sample started in that window consumes exactly one draw, and a sample started before the window cannot reach
// past it unless it consumed more than 32 draws: probability < 1e-40).  Every thread owns L consecutive raw positions,
// looks back until it finds such an anchor, walks the sample chain from there to its own range, and then
//   pass 1 (k_zig_count): counts the samples that START in its range;  an exclusive scan gives sample ordinals;
//   pass 2 (k_zig_fill):  repeats the walk and writes value #ord to X[(ord % n) * k + ord / n].
// Arithmetic is NumPy's: the tables are NumPy's own (slq_ziggurat_tables.h, read from the installed NumPy by
// scripts/make_ziggurat_tables.py), products/sums are explicitly rounded (no FMA contraction: NumPy's baseline build
// has none).  exp/log1p only decide accept/reject (a last-bit difference flips a decision with probability ~1e-16),
// EXCEPT in the tail, where -log1p(-U1)/r is returned: tail samples (2.6e-4 of all) are listed (ordinal, U1) so the
// host can recompute them with the very libm NumPy calls and patch them in (primate_b200/random.py).
#include <cfloat>
#include <cmath>
#include <vector>

#include "slq_pcg64.cuh"
#include "slq_ziggurat_tables.h"

namespace slq {

__constant__ uint64_t c_zig_ki[256];
__constant__ uint64_t c_zig_wi[256];
__constant__ uint64_t c_zig_fi[256];

constexpr int ZIG_L = 512;      // raw positions owned by one thread
constexpr int ZIG_WINDOW = 32;  // consecutive fast positions that prove a sample start
constexpr int ZIG_NT = 128;
constexpr uint64_t ZIG_MASK52 = 0x000fffffffffffffull;

struct ZigTables {  // shared-memory copies (random per-thread indexing would serialise on the constant cache)
  uint64_t ki[256];
  double wi[256];
  double fi[256];
};

__device__ __forceinline__ void zig_load_tables(ZigTables& t) {
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    t.ki[i] = c_zig_ki[i];
    t.wi[i] = __longlong_as_double((long long)c_zig_wi[i]);
    t.fi[i] = __longlong_as_double((long long)c_zig_fi[i]);
  }
  __syncthreads();
}

__device__ __forceinline__ double raw_to_double(uint64_t r) { return (double)(r >> 11) * (1.0 / 9007199254740992.0); }

struct ZigSample {
  double x;
  int consumed;       // raw draws used
  int tail;           // 1: tail sample (value to be recomputed on the host with libm's log1p)
  uint64_t tail_raw;  // the accepted iteration's first raw draw (U1)
  int tail_neg;
};

// One sample starting at the generator's current position (state `s` is advanced past everything it consumed).
__device__ __forceinline__ ZigSample zig_sample(u128& s, u128 inc, const ZigTables& t) {
  ZigSample out;
  out.consumed = 0;
  out.tail = 0;
  out.tail_raw = 0;
  out.tail_neg = 0;
  for (;;) {
    uint64_t r = pcg_step_output(s, inc);
    ++out.consumed;
    const int idx = (int)(r & 0xff);
    r >>= 8;
    const int sign = (int)(r & 1);
    const uint64_t rabs = (r >> 1) & ZIG_MASK52;
    double x = __dmul_rn((double)rabs, t.wi[idx]);
    if (sign) x = -x;
    if (rabs < t.ki[idx]) {
      out.x = x;
      return out;
    }
    if (idx == 0) {
      for (;;) {
        const uint64_t r1 = pcg_step_output(s, inc), r2 = pcg_step_output(s, inc);
        out.consumed += 2;
        const double xx = __dmul_rn(-SLQ_ZIG_NOR_INV_R, log1p(-raw_to_double(r1)));
        const double yy = -log1p(-raw_to_double(r2));
        if (__dadd_rn(yy, yy) > __dmul_rn(xx, xx)) {
          out.tail = 1;
          out.tail_raw = r1;
          out.tail_neg = (int)((rabs >> 8) & 1);
          out.x = out.tail_neg ? -__dadd_rn(SLQ_ZIG_NOR_R, xx) : __dadd_rn(SLQ_ZIG_NOR_R, xx);
          return out;
        }
      }
    } else {
      const uint64_t ru = pcg_step_output(s, inc);
      ++out.consumed;
      const double lhs = __dadd_rn(__dmul_rn(__dadd_rn(t.fi[idx - 1], -t.fi[idx]), raw_to_double(ru)), t.fi[idx]);
      if (lhs < exp(__dmul_rn(__dmul_rn(-0.5, x), x))) {
        out.x = x;
        return out;
      }
    }
  }
}

__device__ __forceinline__ bool zig_is_fast(uint64_t r, const ZigTables& t) {
  const int idx = (int)(r & 0xff);
  const uint64_t rabs = (r >> 9) & ZIG_MASK52;
  return rabs < t.ki[idx];
}

// Positions are relative to the scan base (raw #0 of the scan = a known sample start).  Returns the first sample start
// >= t0 and leaves `s` positioned there.  `base_state` is the generator state at relative position 0.
__device__ __forceinline__ int64_t zig_entry(int64_t t0, u128 base_state, u128 inc, const JumpTable& jt, const ZigTables& t, u128& s) {
  if (t0 == 0) {
    s = base_state;
    return 0;
  }
  int64_t look = 96;
  for (;;) {
    int64_t from = t0 - look;
    if (from < 0) from = 0;
    s = dev_advance(base_state, jt, (uint64_t)from);
    int64_t pos = from;
    int64_t anchor = (from == 0) ? 0 : -1;  // relative position 0 is a sample start by definition
    int run = 0;
    if (anchor < 0) {
      while (pos < t0) {
        const uint64_t r = pcg_step_output(s, inc);
        ++pos;
        run = zig_is_fast(r, t) ? run + 1 : 0;
        if (run >= ZIG_WINDOW) {
          anchor = pos;  // positions pos-32 .. pos-1 are fast  =>  pos starts a sample
          break;
        }
      }
    }
    if (anchor >= 0) {
      // walk the chain from the anchor to the first start >= t0 (s is positioned at `anchor`)
      int64_t p = anchor;
      while (p < t0) {
        const ZigSample sm = zig_sample(s, inc, t);
        p += sm.consumed;
      }
      return p;
    }
    look *= 4;  // no anchor in the window (rare): look further back, down to the scan base
  }
}

// pass 1: samples starting in [t0, t0 + L) per thread (threads beyond `nraw` count nothing)
__global__ void __launch_bounds__(ZIG_NT) k_zig_count(u128 base_state, u128 inc, __grid_constant__ const JumpTable jt, int64_t nraw,
                                                       uint32_t* __restrict__ counts, int64_t nthreads) {
  __shared__ ZigTables t;
  zig_load_tables(t);
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nthreads) return;
  const int64_t t0 = gid * ZIG_L;
  int64_t t1 = t0 + ZIG_L;
  if (t1 > nraw) t1 = nraw;
  u128 s;
  int64_t p = zig_entry(t0, base_state, inc, jt, t, s);
  uint32_t c = 0;
  while (p < t1) {
    const ZigSample sm = zig_sample(s, inc, t);
    p += sm.consumed;
    ++c;
  }
  counts[gid] = c;
}

// exclusive scan of the per-thread counts (one CTA; counts are few: nraw / 512), total in *total_out
__global__ void __launch_bounds__(1024) k_scan_counts(const uint32_t* __restrict__ counts, int64_t m, int64_t* __restrict__ offsets,
                                                       int64_t* __restrict__ total_out) {
  __shared__ int64_t s_part[1024];
  const int t = threadIdx.x;
  const int64_t per = (m + 1023) / 1024;
  const int64_t a = (int64_t)t * per, b = (a + per < m) ? a + per : m;
  int64_t sum = 0;
  for (int64_t i = a; i < b; ++i) sum += counts[i];
  s_part[t] = sum;
  __syncthreads();
  if (t == 0) {
    int64_t run = 0;
    for (int i = 0; i < 1024; ++i) {
      const int64_t v = s_part[i];
      s_part[i] = run;
      run += v;
    }
    if (total_out) *total_out = run;
  }
  __syncthreads();
  int64_t run = s_part[t];
  for (int64_t i = a; i < b; ++i) {
    offsets[i] = run;
    run += counts[i];
  }
}

// pass 2: write sample #ord (ord < nsamples) to X[(ord % n) * k + ord / n]; record where sample #nsamples starts
// (the raw draws the block consumed) and list the tail samples for the host patch.
__global__ void __launch_bounds__(ZIG_NT) k_zig_fill(u128 base_state, u128 inc, __grid_constant__ const JumpTable jt, int64_t nraw,
                                                      const int64_t* __restrict__ offsets, int64_t nthreads, int64_t n, int k,
                                                      int64_t nsamples, double* __restrict__ X, unsigned long long* __restrict__ end_raw,
                                                      unsigned long long* __restrict__ tail_count, int64_t tail_cap,
                                                      int64_t* __restrict__ tail_ord, uint64_t* __restrict__ tail_raw) {
  __shared__ ZigTables t;
  zig_load_tables(t);
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nthreads) return;
  const int64_t t0 = gid * ZIG_L;
  int64_t t1 = t0 + ZIG_L;
  if (t1 > nraw) t1 = nraw;
  int64_t ord = offsets[gid];
  if (ord > nsamples) return;
  u128 s;
  int64_t p = zig_entry(t0, base_state, inc, jt, t, s);
  while (p < t1) {
    if (ord == nsamples) {
      *end_raw = (unsigned long long)p;  // exactly one thread owns this ordinal
      return;
    }
    const ZigSample sm = zig_sample(s, inc, t);
    X[(ord % n) * k + ord / n] = sm.x;
    if (sm.tail) {
      const unsigned long long slot = atomicAdd(tail_count, 1ull);
      if ((int64_t)slot < tail_cap) {
        tail_ord[slot] = (ord << 1) | (int64_t)sm.tail_neg;
        tail_raw[slot] = sm.tail_raw;
      }
    }
    p += sm.consumed;
    ++ord;
  }
}

// seek: first thread range whose samples reach ordinal `target`; writes the raw position where sample #target starts
__global__ void __launch_bounds__(ZIG_NT) k_zig_locate(u128 base_state, u128 inc, __grid_constant__ const JumpTable jt, int64_t nraw,
                                                        const int64_t* __restrict__ offsets, const uint32_t* __restrict__ counts,
                                                        int64_t nthreads, int64_t target, unsigned long long* __restrict__ raw_out) {
  __shared__ ZigTables t;
  zig_load_tables(t);
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nthreads) return;
  const int64_t first = offsets[gid];
  if (!(first <= target && target < first + (int64_t)counts[gid])) return;
  const int64_t t0 = gid * ZIG_L;
  u128 s;
  int64_t p = zig_entry(t0, base_state, inc, jt, t, s);
  for (int64_t ord = first; ord < target; ++ord) {
    const ZigSample sm = zig_sample(s, inc, t);
    p += sm.consumed;
  }
  *raw_out = (unsigned long long)p;
}

// ---- sphere: column norms in NumPy's pairwise summation order, then x / norm * sqrt(n) ---------------------------------
// np.sum(W**2, axis=0) on an F-ordered (n, k) array is numpy's pairwise sum per column (blocks of <= 128 elements with
// 8 interleaved accumulators, combined by recursive halving at multiples of 8): the leaves are listed by the host
// (leaf_start, leaf_len) together with, per leaf, the number of pair-additions that follow it on a stack machine.
__global__ void __launch_bounds__(256) k_leaf_sums(const double* __restrict__ X, int k, const int64_t* __restrict__ leaf_start,
                                                    const int32_t* __restrict__ leaf_len, int64_t nleaves, double* __restrict__ S) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t leaf = gid / k;
  const int p = (int)(gid % k);
  if (leaf >= nleaves) return;
  const int64_t a = leaf_start[leaf];
  const int len = leaf_len[leaf];
  const double* x = X + a * k + p;
  auto sq = [&](int i) {
    const double v = x[(int64_t)i * k];
    return __dmul_rn(v, v);
  };
  double res;
  if (len < 8) {
    res = 0.0;
    for (int i = 0; i < len; ++i) res = __dadd_rn(res, sq(i));
  } else {
    double r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = sq(j);
    int i = 8;
    for (; i < len - (len % 8); i += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], sq(i + j));
    }
    res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])), __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < len; ++i) res = __dadd_rn(res, sq(i));
  }
  S[leaf * k + p] = res;
}

__global__ void k_combine_leaves(const double* __restrict__ S, int k, int64_t nleaves, const int32_t* __restrict__ reduce_after,
                                 double* __restrict__ norm2) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= k) return;
  double stack[64];
  int top = 0;
  for (int64_t leaf = 0; leaf < nleaves; ++leaf) {
    stack[top++] = S[leaf * k + p];
    for (int r = reduce_after[leaf]; r > 0; --r) {
      const double b = stack[--top], a = stack[--top];
      stack[top++] = __dadd_rn(a, b);
    }
  }
  norm2[p] = stack[0];
}

__global__ void __launch_bounds__(256) k_sphere_scale(double* __restrict__ X, int64_t n, int k, const double* __restrict__ norm2,
                                                       double sqrt_n) {
  const int64_t total = n * k;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const double nrm = __dsqrt_rn(norm2[e % k]);
    X[e] = __dmul_rn(__ddiv_rn(X[e], nrm), sqrt_n);  // np.divide then np.multiply (random.py:40-41): two roundings
  }
}

static DeviceOnce g_tables_once;
static int upload_tables() {
  if (g_tables_once.needed()) {
    SLQ_CUDA(cudaMemcpyToSymbol(c_zig_ki, slq_zig_ki, sizeof(slq_zig_ki)));
    SLQ_CUDA(cudaMemcpyToSymbol(c_zig_wi, slq_zig_wi_bits, sizeof(slq_zig_wi_bits)));
    SLQ_CUDA(cudaMemcpyToSymbol(c_zig_fi, slq_zig_fi_bits, sizeof(slq_zig_fi_bits)));
    g_tables_once.done();
  }
  return SLQ_OK;
}

struct NormalWs {
  size_t counts_off, offsets_off, total_off, total;
  int64_t nthreads;
};
static NormalWs normal_layout(int64_t nraw) {
  NormalWs w;
  w.nthreads = (nraw + ZIG_L - 1) / ZIG_L;
  w.counts_off = 0;
  w.offsets_off = align_up(sizeof(uint32_t) * (size_t)w.nthreads, 256);
  w.total_off = w.offsets_off + align_up(sizeof(int64_t) * (size_t)w.nthreads, 256);
  w.total = w.total_off + 256;
  return w;
}

}  // namespace slq

using namespace slq;

extern "C" {

const char* slq_normals_numpy_version(void) { return SLQ_ZIG_NUMPY_VERSION; }

size_t slq_normals_workspace_bytes(int64_t nraw) { return nraw > 0 ? normal_layout(nraw).total : 0; }

// Scans `nraw` raw draws starting at raw offset `raw_base` of the PCG64 stream (state, inc) -- where a sample starts --
// and counts the samples that start inside; if target_ordinal >= 0 and fewer than the count, also locates the raw
// position (relative to raw_base) where sample #target_ordinal starts.  Device outputs: total_dev (int64 count),
// raw_out_dev (uint64).  No synchronisation.
int slq_normals_scan(const uint64_t state_hi_lo[2], const uint64_t inc_hi_lo[2], uint64_t raw_base, int64_t nraw, int64_t target_ordinal,
                     int64_t* total_dev, uint64_t* raw_out_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(state_hi_lo && inc_hi_lo && total_dev && workspace_dev && nraw > 0, "invalid argument");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const NormalWs w = normal_layout(nraw);
  SLQ_REQUIRE(workspace_bytes >= w.total, "workspace too small: %zu < %zu", workspace_bytes, w.total);
  int rc = upload_tables();
  if (rc != SLQ_OK) return rc;
  u128 inc = {inc_hi_lo[0], inc_hi_lo[1]}, st = {state_hi_lo[0], state_hi_lo[1]};
  JumpTable jt;
  build_jump_table(inc, &jt);
  st = host_advance(st, jt, raw_base);
  cudaStream_t cs = (cudaStream_t)stream;
  char* base = (char*)workspace_dev;
  uint32_t* counts = (uint32_t*)(base + w.counts_off);
  int64_t* offsets = (int64_t*)(base + w.offsets_off);
  const unsigned blocks = (unsigned)((w.nthreads + ZIG_NT - 1) / ZIG_NT);
  k_zig_count<<<blocks, ZIG_NT, 0, cs>>>(st, inc, jt, nraw, counts, w.nthreads);
  SLQ_LAUNCH_CHECK();
  k_scan_counts<<<1, 1024, 0, cs>>>(counts, w.nthreads, offsets, total_dev);
  SLQ_LAUNCH_CHECK();
  if (target_ordinal >= 0 && raw_out_dev) {
    k_zig_locate<<<blocks, ZIG_NT, 0, cs>>>(st, inc, jt, nraw, offsets, counts, w.nthreads, target_ordinal,
                                            (unsigned long long*)raw_out_dev);
    SLQ_LAUNCH_CHECK();
  }
  return SLQ_OK;
}

// Standard-normal probe block: X[r*k + p] (float64) = sample #(p*n + r) of the stream that starts at raw offset
// `raw_base` (which must be a sample start: 0, or an `end_raw` of an earlier call).  `nraw` raw positions are scanned
// (n*k*1.03 + 8192 is ample; if the block needs more, *end_raw_dev stays UINT64_MAX and the caller retries with more).
// end_raw_dev: raw draws consumed by the block (relative to raw_base).  Tail samples are listed for the host patch:
// tail_ord[i] = (ordinal << 1) | negative, tail_raw[i] = the raw draw U1 of the accepted tail iteration.
int slq_probes_normal(const uint64_t state_hi_lo[2], const uint64_t inc_hi_lo[2], uint64_t raw_base, int64_t nraw, int k, int64_t n,
                      double* X_dev, uint64_t* end_raw_dev, uint64_t* tail_count_dev, int64_t tail_cap, int64_t* tail_ord_dev,
                      uint64_t* tail_raw_dev, void* workspace_dev, size_t workspace_bytes, slq_stream stream) {
  SLQ_REQUIRE(state_hi_lo && inc_hi_lo && X_dev && end_raw_dev && tail_count_dev && tail_ord_dev && tail_raw_dev && workspace_dev,
              "NULL argument");
  SLQ_REQUIRE(n > 0 && k > 0 && nraw >= n * k, "invalid probe block shape (%lld, %d) / scan length", (long long)n, k);
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  const NormalWs w = normal_layout(nraw);
  SLQ_REQUIRE(workspace_bytes >= w.total, "workspace too small: %zu < %zu", workspace_bytes, w.total);
  int rc = upload_tables();
  if (rc != SLQ_OK) return rc;
  u128 inc = {inc_hi_lo[0], inc_hi_lo[1]}, st = {state_hi_lo[0], state_hi_lo[1]};
  JumpTable jt;
  build_jump_table(inc, &jt);
  st = host_advance(st, jt, raw_base);
  cudaStream_t cs = (cudaStream_t)stream;
  char* base = (char*)workspace_dev;
  uint32_t* counts = (uint32_t*)(base + w.counts_off);
  int64_t* offsets = (int64_t*)(base + w.offsets_off);
  int64_t* total = (int64_t*)(base + w.total_off);
  SLQ_CUDA(cudaMemsetAsync(end_raw_dev, 0xff, sizeof(uint64_t), cs));
  SLQ_CUDA(cudaMemsetAsync(tail_count_dev, 0, sizeof(uint64_t), cs));
  const unsigned blocks = (unsigned)((w.nthreads + ZIG_NT - 1) / ZIG_NT);
  k_zig_count<<<blocks, ZIG_NT, 0, cs>>>(st, inc, jt, nraw, counts, w.nthreads);
  SLQ_LAUNCH_CHECK();
  k_scan_counts<<<1, 1024, 0, cs>>>(counts, w.nthreads, offsets, total);
  SLQ_LAUNCH_CHECK();
  k_zig_fill<<<blocks, ZIG_NT, 0, cs>>>(st, inc, jt, nraw, offsets, w.nthreads, n, k, n * (int64_t)k, X_dev,
                                        (unsigned long long*)end_raw_dev, (unsigned long long*)tail_count_dev, tail_cap, tail_ord_dev,
                                        tail_raw_dev);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

// Sphere scaling of a float64 normal block in place: X[:, p] *= sqrt(n) / ||X[:, p]|| with the column norms summed in
// NumPy's pairwise order (src/primate/random.py:36-41).  leaf_* / reduce_after describe that order for this n
// (primate_b200.random._pairwise_plan); S_dev: nleaves * k doubles of scratch; norm2_dev: k doubles.
int slq_probes_sphere_scale(double* X_dev, int64_t n, int k, const int64_t* leaf_start_dev, const int32_t* leaf_len_dev,
                            const int32_t* reduce_after_dev, int64_t nleaves, double* S_dev, double* norm2_dev, slq_stream stream) {
  SLQ_REQUIRE(X_dev && leaf_start_dev && leaf_len_dev && reduce_after_dev && S_dev && norm2_dev, "NULL argument");
  SLQ_REQUIRE(n > 0 && k > 0 && nleaves > 0, "invalid shape");
  const int sms = sm_count();
  if (sms <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  cudaStream_t cs = (cudaStream_t)stream;
  const int64_t work = nleaves * k;
  k_leaf_sums<<<(unsigned)((work + 255) / 256), 256, 0, cs>>>(X_dev, k, leaf_start_dev, leaf_len_dev, nleaves, S_dev);
  SLQ_LAUNCH_CHECK();
  k_combine_leaves<<<(k + 63) / 64, 64, 0, cs>>>(S_dev, k, nleaves, reduce_after_dev, norm2_dev);
  SLQ_LAUNCH_CHECK();
  int64_t g = (n * k + 255) / 256;
  if (g > (int64_t)sms * 16) g = (int64_t)sms * 16;
  k_sphere_scale<<<(unsigned)g, 256, 0, cs>>>(X_dev, n, k, norm2_dev, std::sqrt((double)n));
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // extern "C"
