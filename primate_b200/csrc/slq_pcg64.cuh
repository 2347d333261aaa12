// PCG64 (128-bit LCG + XSL-RR 128/64 output) with O(log d) jump-ahead, shared by the probe generators.
// PCG64 = 128-bit LCG (state = state * M + inc) with the XSL-RR output function applied AFTER the step (NumPy's PCG64).
// Random access into the stream uses  s_{t+d} = A_d s_t + C_d  with (A_{2^b}, C_{2^b}) tabulated on the host for b < 64.
#pragma once
#include "slq_internal.cuh"

namespace slq {


struct u128 {
  uint64_t hi, lo;
};

__host__ __device__ __forceinline__ u128 mul128(u128 a, u128 b) {
  u128 r;
  r.lo = a.lo * b.lo;
#ifdef __CUDA_ARCH__
  r.hi = __umul64hi(a.lo, b.lo) + a.hi * b.lo + a.lo * b.hi;
#else
  r.hi = (uint64_t)(((unsigned __int128)a.lo * b.lo) >> 64) + a.hi * b.lo + a.lo * b.hi;
#endif
  return r;
}
__host__ __device__ __forceinline__ u128 add128(u128 a, u128 b) {
  u128 r;
  r.lo = a.lo + b.lo;
  r.hi = a.hi + b.hi + (r.lo < a.lo ? 1u : 0u);
  return r;
}

// PCG_DEFAULT_MULTIPLIER_128
#define PCG_MULT_HI 0x2360ED051FC65DA4ull
#define PCG_MULT_LO 0x4385DF649FCCF645ull

struct JumpTable {
  u128 mult[64];
  u128 plus[64];
};

static void build_jump_table(u128 inc, JumpTable* jt) {
  u128 m = {PCG_MULT_HI, PCG_MULT_LO}, p = inc;
  for (int b = 0; b < 64; ++b) {
    jt->mult[b] = m;
    jt->plus[b] = p;
    u128 m1 = add128(m, u128{0, 1});
    p = mul128(m1, p);
    m = mul128(m, m);
  }
}

static u128 host_advance(u128 s, const JumpTable& jt, uint64_t d) {
  for (int b = 0; b < 64 && d; ++b, d >>= 1)
    if (d & 1) s = add128(mul128(jt.mult[b], s), jt.plus[b]);
  return s;
}

__device__ __forceinline__ u128 dev_advance(u128 s, const JumpTable& jt, uint64_t d) {
  for (int b = 0; d; ++b, d >>= 1)
    if (d & 1) s = add128(mul128(jt.mult[b], s), jt.plus[b]);
  return s;
}

__device__ __forceinline__ uint64_t pcg_step_output(u128& s, u128 inc) {
  s = add128(mul128(s, u128{PCG_MULT_HI, PCG_MULT_LO}), inc);
  const uint64_t x = s.hi ^ s.lo;
  const unsigned rot = (unsigned)(s.hi >> 58);
  return (x >> rot) | (x << ((64u - rot) & 63u));
}


}  // namespace slq
