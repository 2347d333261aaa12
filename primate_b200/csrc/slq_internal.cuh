// Internal helpers shared by the libslq translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "slq.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libslq is written for sm_100a (B200) only"
#endif

namespace slq {

// ---- error reporting -------------------------------------------------------------------------------------
char* err_buf();  // thread-local, 512 bytes
int fail(int code, const char* fmt, ...);

#define SLQ_CUDA(expr)                                                                                 \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess)                                                                             \
      return ::slq::fail(SLQ_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define SLQ_LAUNCH_CHECK()                                                                             \
  do {                                                                                                 \
    ::slq::count_launch();                                                                             \
    cudaError_t _e = cudaGetLastError();                                                               \
    if (_e != cudaSuccess)                                                                             \
      return ::slq::fail(SLQ_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define SLQ_REQUIRE(cond, ...) \
  do {                         \
    if (!(cond)) return ::slq::fail(SLQ_ERR_ARG, __VA_ARGS__); \
  } while (0)

int sm_count();  // SM count of the CURRENT device (cached per device ordinal); <= 0 if no device

// Per-device "done once" bookkeeping for cudaFuncSetAttribute and friends: a process may drive several GPUs (one
// handle per device, SURVEY 8b "thread-compatible per handle"), so nothing is cached per process.  Usage:
//   static DeviceOnce once;  if (once.needed()) { ...configure...; once.done(); }
// Two threads may both configure (idempotent calls); the bit is only set AFTER the configuration succeeded, so no
// thread can launch on a device whose attributes are still unset.
constexpr int MAX_DEVICES = 64;
int current_device();  // cudaGetDevice, or -1
struct DeviceOnce {
  std::atomic<unsigned long long> mask{0ull};
  bool needed() const {
    const int d = current_device();
    return d < 0 || d >= MAX_DEVICES || !((mask.load(std::memory_order_acquire) >> d) & 1ull);
  }
  void done() {
    const int d = current_device();
    if (d >= 0 && d < MAX_DEVICES) mask.fetch_or(1ull << d, std::memory_order_release);
  }
};
void count_launch();  // host-side counter of kernels launched by the library (slq_launch_count)

// Optional per-kernel-class timing with CUDA events on the launching stream (slq_profile_*): used by bench.py to
// get the dominant kernel's average launch duration inside the timed region.  No-ops unless enabled.
enum ProfClass { PROF_STEPA = 0, PROF_AXPY = 1, PROF_QUAD = 2, PROF_PROBES = 3, PROF_NCLASS = 4 };
int prof_begin(int cls, cudaStream_t st);             // returns a slot (>= 0) or -1 when not recording
void prof_end(int cls, int slot, cudaStream_t st);

inline bool is_pow2(int k) { return k > 0 && (k & (k - 1)) == 0; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline size_t dtype_size(int dtype) { return dtype == SLQ_F64 ? 8 : 4; }

// ---- operator handle -------------------------------------------------------------------------------------
enum OpKind { OP_CSR = 0, OP_DENSE = 1, OP_CALLBACK = 2 };

}  // namespace slq

struct slq_op {
  int kind;
  int dtype;
  int64_t n;
  int64_t nnz;
  // CSR
  const int32_t* rowptr;
  const int32_t* colidx;
  const void* vals;
  int max_row_nnz;  // longest CSR row (computed once at creation; selects the step-A variant with the long-row path)
  // dense
  const void* dense;
  int64_t ld;
  void* scratch;         // optional caller-owned scratch (slq_op_set_scratch): split-K slabs of small dense operators
  size_t scratch_bytes;
  int64_t npad;  // 0, or n rounded up to 16: the array is npad x ld with zero rows / columns past n (slq_dense_create_padded)
  int no_fused_step;  // 1: keep the two-kernel Lanczos step even where the fused one applies (slq_op_set_fused_step)
  const void* tiled;  // optional band-major copy of the fp64 operator for the tensor-core kernel (slq_dense_set_tiled)
  int tiled_rpt;      // CTA height (16 rpt rows) the copy was laid out for
  // callback
  slq_matmat_fn fn;
  void* ctx;
};

namespace slq {

// ---- vector packs: V scalars moved with one 4/8/16/32-byte access -------------------------------------------
template <typename T, int V>
struct alignas(sizeof(T) * V) Pack {
  T v[V];
};

enum CachePolicy { LD_DEFAULT = 0, LD_STREAM = 1, LD_READONLY = 2 };

struct alignas(32) Raw256 {
  unsigned long long x, y, z, w;
};

template <int BYTES>
struct RawBytes;
template <>
struct RawBytes<4> {
  using type = unsigned int;
};
template <>
struct RawBytes<8> {
  using type = uint2;
};
template <>
struct RawBytes<16> {
  using type = uint4;
};
template <>
struct RawBytes<32> {
  using type = Raw256;
};

// 256-bit global accesses (sm_100: LDG.E.256 / STG.E.256) with the three cache policies used below.
template <int POLICY>
__device__ __forceinline__ Raw256 ld256(const void* p) {
  Raw256 r;
  if (POLICY == LD_STREAM)
    asm volatile("ld.global.cs.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(p));
  else if (POLICY == LD_READONLY)
    asm volatile("ld.global.nc.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(p));
  else
    asm volatile("ld.global.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(p) : "memory");
  return r;
}
template <bool STREAM>
__device__ __forceinline__ void st256(void* p, const Raw256& r) {
  if (STREAM)
    asm volatile("st.global.cs.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(r.x), "l"(r.y), "l"(r.z), "l"(r.w) : "memory");
  else
    asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(r.x), "l"(r.y), "l"(r.z), "l"(r.w) : "memory");
}

// Vector load of V scalars with a cache policy: LD_STREAM = evict-first (read once), LD_READONLY = non-coherent
// path (re-used panels: the gathered probe rows), LD_DEFAULT = plain.
template <typename T, int V, int POLICY>
__device__ __forceinline__ Pack<T, V> ld_pack(const T* p) {
  using R = typename RawBytes<sizeof(T) * V>::type;
  Pack<T, V> out;
  if constexpr (sizeof(T) * V == 32) {
    *reinterpret_cast<Raw256*>(&out) = ld256<POLICY>(p);
  } else {
    R raw;
    if (POLICY == LD_STREAM)
      raw = __ldcs(reinterpret_cast<const R*>(p));
    else if (POLICY == LD_READONLY)
      raw = __ldg(reinterpret_cast<const R*>(p));
    else
      raw = *reinterpret_cast<const R*>(p);
    *reinterpret_cast<R*>(&out) = raw;
  }
  return out;
}

template <typename T, int V, bool STREAM>
__device__ __forceinline__ void st_pack(T* p, const Pack<T, V>& x) {
  using R = typename RawBytes<sizeof(T) * V>::type;
  if constexpr (sizeof(T) * V == 32) {
    st256<STREAM>(p, *reinterpret_cast<const Raw256*>(&x));
  } else {
    R raw = *reinterpret_cast<const R*>(&x);
    if (STREAM)
      __stcs(reinterpret_cast<R*>(p), raw);
    else
      *reinterpret_cast<R*>(p) = raw;
  }
}

// Explicitly rounded products and fused multiply-subtracts for the recurrence arithmetic.  The three-term recurrence is
// evaluated by several kernels (the fused SpMM epilogues, the streaming axpy, the basis-free replay of f(A) X): spelling
// the contraction out -- instead of leaving `a * b - c * d` to the compiler's per-kernel choice -- makes every one of
// them produce bit-identical Lanczos vectors, which the two-pass f(A) X engine relies on.
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
// c - a * b, one rounding
__device__ __forceinline__ double fnma_rn(double a, double b, double c) { return __fma_rn(-a, b, c); }
__device__ __forceinline__ float fnma_rn(float a, float b, float c) { return __fmaf_rn(-a, b, c); }
// a * b + c, one rounding
__device__ __forceinline__ double fma_rn(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// Probe-block thread mapping for scalar type T and block width K: each thread owns V adjacent probes of a row (one
// 32-byte access); LPR threads cover one row.
template <typename T, int K>
struct BlockMap {
  static constexpr int VMAX = 32 / (int)sizeof(T);  // 256-bit accesses
  static constexpr int V = K < VMAX ? K : VMAX;
  static constexpr int LPR = K / V;  // threads per row (<= 32 for K <= 64 fp64 / 128 fp32)
  static_assert(LPR <= 32, "block width too large");
};

constexpr int NT = 256;  // threads per CTA for all streaming kernels
constexpr int NWARP = NT / 32;

// Deterministic per-CTA reduction of per-thread accumulators that belong to probe slots (tid % LPR)*V + i.
// Result: for NACC accumulators, part[a*K + probe] written to `out` (one row of the per-CTA partial table).
template <int K, int V, int LPR, int NACC>
__device__ __forceinline__ void cta_reduce_slots(double (&acc)[NACC][V], double* out, double* smem /* NWARP*NACC*K */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < NACC; ++a)
#pragma unroll
    for (int i = 0; i < V; ++i) {
      double x = acc[a][i];
#pragma unroll
      for (int off = 16; off >= LPR; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
      acc[a][i] = x;
    }
  if (lane < LPR) {
#pragma unroll
    for (int a = 0; a < NACC; ++a)
#pragma unroll
      for (int i = 0; i < V; ++i) smem[(warp * NACC + a) * K + lane * V + i] = acc[a][i];
  }
  __syncthreads();
  for (int t = threadIdx.x; t < NACC * K; t += NT) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < NWARP; ++w) s += smem[w * NACC * K + t];
    out[t] = s;
  }
}

// "Last CTA done" election.  All threads must call; returns true in every thread of the last CTA to arrive.
// The partial table written before the call is visible to the elected CTA.
__device__ __forceinline__ bool elect_last_cta(unsigned int* counter) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(counter, 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) *counter = 0u;  // re-arm for the next launch (stream ordered)
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0;
}

}  // namespace slq
