// Step-A kernel for CSR operators, TMA / mbarrier edition (sm_100a).  Included by slq_lanczos.cu.
//
//   w = (A u_c) r_c - beta_j r_p u_p ;  alpha_j partials = <q_c, w>          (lanczos.h:123-129, eigen_operators.h:66-77)
//
// One persistent CTA per SM: 15 consumer warps + 1 producer warp.  The producer walks the CTA's row tiles
// (round-robin over the grid, so all CTAs sweep the matrix as one front and gathered neighbour rows stay in L2) and
// fills a 3-stage shared-memory ring with bulk async copies (cp.async.bulk -> SASS UBLKCP, completion on an mbarrier):
//     * the tile's slice of colidx / vals  (CSR bytes: read exactly once, fully coalesced, no registers, no shuffles),
//     * the tile's own rows of u_c         (the dense probe block of the current Lanczos vector, 30 KB per stage),
//     * the tile's rows of u_p             (the previous Lanczos vector, 30 KB per stage),
//     * the tile's row pointers            (plain loads by the producer lanes, one tile ahead of their use).
// Consumers then run a lean loop per row: index/value from shared memory (broadcast inside the row group); a non-zero
// whose column lies inside the tile reads its 16-byte probe-row slice from the staged u_c tile, any other column is
// gathered from global memory through the read-only path; FMAs in storage order (so a sorted CSR row is summed in the
// order of the reference's column sweep).  R rows x U non-zeros of gathers are in flight per thread.  With the locality
// ordering of slq_csr_cluster_order ~85% of the non-zeros of a grid operator are tile-local, so DRAM latency is paid by
// the bulk copies (three stages deep), not by the consumers.
// Tiles whose non-zero slice exceeds the stage capacity are processed in several chunks (accumulators persist in
// registers), so arbitrarily long rows (power-law graphs) are handled by the same code.
#pragma once

namespace slq {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk async copy global -> shared, completion counted in bytes on `bar` (16-byte aligned addresses and size).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// 16-byte asynchronous gather global -> shared (L2 only: the gathered probe rows are not re-used from L1)
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

#ifndef SLQ_TMA_U
#define SLQ_TMA_U 5  // gathers per row per round
#endif
#ifndef SLQ_TMA_HOIST
#define SLQ_TMA_HOIST 1  // per-column scales in registers: 0 none, 1 beta*scale_p (+1% on C2), 2 both (spills at 128 registers: -17%)
#endif
#ifndef SLQ_TMA_UB
#define SLQ_TMA_UB 8  // gathers in flight per thread, non-zero-balanced variant
#endif
#ifndef SLQ_TMA_R
#define SLQ_TMA_R 2  // rows per thread processed together
#endif
#ifndef SLQ_TMA_RF
#define SLQ_TMA_RF 1
#endif
#ifndef SLQ_TMA_FUSED_STREAM_ST
#define SLQ_TMA_FUSED_STREAM_ST 1  // the fused step's two output rows are written with streaming (evict-first) stores: +3.6 % at k = 64 (L2 is left to the off-tile gathers)
#endif
#ifndef SLQ_TMA_RF512
#define SLQ_TMA_RF512 1
#endif
#ifndef SLQ_TMA_PREFETCH
#define SLQ_TMA_PREFETCH 0  // 1: the producer warp also prefetches the off-tile probe rows of the last issued stage into L2
#endif
#ifndef SLQ_TMA_STAGES
#define SLQ_TMA_STAGES 3
#endif

// BAL: layout of the non-zero-balanced variant (irregular operators): larger index chunks, fewer stages, row sums in smem.
#ifndef SLQ_TMA_CWF
#define SLQ_TMA_CWF 15  // consumer warps of the fused step at 512-byte rows (fp64, k = 64); 10 warps x 2 rows in flight measured -6%
#endif
template <typename T, int K, bool BAL = false, bool FUSEDT = false>
struct TmaTiling {
  // consumer warps (+1 producer).  Lean: 15 (512 threads, 128 registers each, one CTA per SM).  Balanced: 11 (384
  // threads) and TWO CTAs per SM -- its consumers meet at two CTA barriers per chunk (the slice hand-over of rows that
  // span slices), and a second, independent CTA fills the issue slots the first one leaves idle while it waits there.
  // (Tuning knob of the fused step at 512-byte rows: fewer warps with more registers each, so that a thread keeps two
  // rows' gathers in flight.  Measured slower than 15 warps x 1 row: the shared-memory pipe, 71 % busy, is the limit.)
  static constexpr int CW = BAL ? 11 : ((FUSEDT && K * (int)sizeof(T) == 512) ? SLQ_TMA_CWF : 15);
  static constexpr int CTAS_PER_SM = BAL ? 2 : 1;
  static constexpr int CT = CW * 32;                   // consumer threads
  static constexpr int NTHREADS = CT + 32;             // + producer warp
  static constexpr int V = 16 / (int)sizeof(T);        // 16-byte slices
  static constexpr int LPR = K / V;                    // lanes per row
  static constexpr int RPP = CT / LPR;                 // rows per pass
  static constexpr int TILE_MAX = 30720 / (K * (int)sizeof(T));  // staged panels of at most 30 KB each
  // rows per tile: lean = the cluster size of the locality ordering (120); balanced = 4 passes of the row groups
  static constexpr int TILE = BAL ? 4 * RPP : (TILE_MAX < 120 ? TILE_MAX : 120);
  static constexpr int PASSES = TILE / RPP;
  static constexpr int R = PASSES % SLQ_TMA_R == 0 ? SLQ_TMA_R : 1;  // rows per thread processed together
  static constexpr int RFW = (FUSEDT && K * (int)sizeof(T) == 512) ? SLQ_TMA_RF512 : SLQ_TMA_RF;
  static constexpr int RF = PASSES % RFW == 0 ? RFW : 1;  // ... in the fused step (two loads per non-zero)
  static constexpr size_t PANEL = (size_t)TILE * K * sizeof(T);
  // Balanced variant (irregular operators): no staged u_c / u_p tiles (on a power-law graph next to no column is
  // tile-local); the shared memory goes to a ring of RING 16-byte slots PER CONSUMER THREAD instead, filled by
  // cp.async gathers that run RING non-zeros ahead of the FMAs -- bytes in flight are bounded by shared memory
  // (RING x 480 x 16 B = 120 KB per SM), not by the 128-register budget (8 x 16 B per thread before).
  static constexpr int RING = 8;
  static constexpr size_t RING_BYTES = (size_t)RING * CT * 16;
  static constexpr size_t BAL_EXTRA = PANEL + RING_BYTES + (size_t)RPP * 4 + 32;  // row sums + gather ring + continuation rows
  static constexpr size_t BAL_BUDGET = 100 * 1024;  // two CTAs per SM
  static constexpr int BAL_STAGES = 2;
  // the chunk is the unit of the per-chunk barriers and of the slice bookkeeping: as large as the budget allows
  static constexpr int BAL_CAP = (int)((((BAL_BUDGET - BAL_EXTRA) / BAL_STAGES - 1024) / (sizeof(T) + 4)) / 128 * 128);
  static constexpr int CAP = BAL ? BAL_CAP : 960;      // non-zeros per stage (lean: 8 per row; longer slices go in chunks)
  static constexpr int STAGES = SLQ_TMA_STAGES;
  static constexpr int NRP = (TILE + 1 + 31) / 32;     // row pointers per producer lane
  static constexpr int RP_LEN = NRP * 32;
  // stage layout (bytes)
  static constexpr size_t STAGED_PANEL = BAL ? 0 : (size_t)TILE * K * sizeof(T);
  static constexpr size_t OFF_UC = 0;
  static constexpr size_t OFF_UP = OFF_UC + STAGED_PANEL;
  static constexpr size_t OFF_VA = OFF_UP + STAGED_PANEL;
  static constexpr size_t OFF_CI = OFF_VA + (size_t)CAP * sizeof(T);
  static constexpr size_t OFF_RP = OFF_CI + (size_t)CAP * 4;
  static constexpr size_t OFF_HDR = OFF_RP + (size_t)RP_LEN * 4;
  static constexpr int LONG = 64;                      // a row with more non-zeros than this inside one chunk is shared by all row groups
  static constexpr int MAXLONG = CAP / LONG + 1;
  static constexpr size_t STAGE_BYTES = (OFF_HDR + 16 + 4 * MAXLONG + 127) / 128 * 128;  // header: lo, hi, last, nlong, long rows
  static constexpr size_t SMEM_BYTES = STAGE_BYTES * STAGES;
  static constexpr size_t SMEM_BAL = STAGE_BYTES * BAL_STAGES + BAL_EXTRA;
  static_assert(!BAL || (SMEM_BAL <= 101 * 1024 && BAL_CAP >= 512), "balanced variant does not fit in shared memory");
  static_assert(K % V == 0 && LPR >= 4 && LPR <= 32, "block width not supported by the TMA kernel");
  static_assert(PASSES >= 1 && PASSES * RPP == TILE, "tile must be a whole number of passes");
};

template <typename T, int K>
constexpr bool tma_supported() {
  return (K * (int)sizeof(T)) >= 64 && (K * (int)sizeof(T)) <= 512;
}

// Probe-row slice of column c: from the staged tile when c lies in [row0, row0 + TILE), from global memory otherwise.
template <typename T, int K, int V, int TILE>
__device__ __forceinline__ Pack<T, V> tma_fetch(int c, int row0, const T* s_uc, const T* __restrict__ X, int slot) {
  const unsigned loc = (unsigned)(c - row0);
  if (loc < (unsigned)TILE) return *reinterpret_cast<const Pack<T, V>*>(s_uc + (size_t)loc * K + slot);
  return ld_pack<T, V, LD_READONLY>(X + (size_t)c * K + slot);
}

// R rows at once, U gathers per row per round; non-zeros [s[r], e[r]) index the stage's ci/va arrays.
template <typename T, int K, int V, int U, int R, int TILE>
__device__ __forceinline__ void tma_gather_rows(T (&acc)[R][V], const int (&s)[R], const int (&e)[R], const int* __restrict__ ci,
                                                const T* __restrict__ va, const T* s_uc, int row0, const T* __restrict__ X,
                                                int slot) {
  int len = 0, lmin = 0x7fffffff;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    len = max(len, e[r] - s[r]);
    lmin = min(lmin, e[r] - s[r]);
  }
  int i = 0;
  for (; i + U <= lmin; i += U) {  // full rounds: no predicates
    Pack<T, V> xv[R][U];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int u = 0; u < U; ++u) xv[r][u] = tma_fetch<T, K, V, TILE>(ci[s[r] + i + u], row0, s_uc, X, slot);
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const T a = va[s[r] + i + u];
#pragma unroll
        for (int q = 0; q < V; ++q) acc[r][q] = fma(a, xv[r][u].v[q], acc[r][q]);
      }
  }
  for (; i < len; i += U) {  // ragged rounds
    Pack<T, V> xv[R][U];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (s[r] + i + u < e[r]) xv[r][u] = tma_fetch<T, K, V, TILE>(ci[s[r] + i + u], row0, s_uc, X, slot);
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (s[r] + i + u < e[r]) {
          const T a = va[s[r] + i + u];
#pragma unroll
          for (int q = 0; q < V; ++q) acc[r][q] = fma(a, xv[r][u].v[q], acc[r][q]);
        }
  }
}

// Fused step: the gathered vector does not exist yet -- entry (c, slot) of it is w[c] - cf u[c], formed from the two
// staged tiles (in-tile columns) or two global rows, in exactly the arithmetic the row's owner uses when it writes the
// vector out (fnma_rn), so every CTA sees the same values.  Both loads of all R x U non-zeros are issued before use.
#ifndef SLQ_TMA_UF
#define SLQ_TMA_UF 5  // gathers (pairs of loads) per row per round, fused step
#endif
template <typename T, int K, int V, int U, int R, int TILE>
__device__ __forceinline__ void tma_gather_rows_fused(T (&acc)[R][V], const int (&s)[R], const int (&e)[R], const int* __restrict__ ci,
                                                      const T* __restrict__ va, const T* s_w, const T* s_u, int row0,
                                                      const T* __restrict__ W, const T* __restrict__ Uv, int slot, const T (&cf)[V]) {
  // Branch-free: an in-tile column and an off-tile one differ only in the (generic) address the two rows are loaded
  // from -- the staged tiles or the global panels, each pair a constant distance apart -- and a slot past the end of a
  // short row re-reads entry 0 of the stage with a zero matrix value (fma(0, x, acc) == acc).  The per-non-zero
  // branches of the first version cost more issue slots than the loads and FMAs together (profiles/r02d_fused_step.md).
  const ptrdiff_t d_smem = s_u - s_w, d_glob = Uv - W;
  int len = 0, lmin = 0x7fffffff;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    len = max(len, e[r] - s[r]);
    lmin = min(lmin, e[r] - s[r]);
  }
  auto round = [&](int i, auto ragged) {
    constexpr bool RAGGED = decltype(ragged)::value;
    Pack<T, V> wv[R][U], uv[R][U];
    T a[R][U];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        int idx = s[r] + i + u;
        bool ok = true;
        if (RAGGED) {
          ok = idx < e[r];
          idx = ok ? idx : 0;
        }
        const int c = ci[idx];
        a[r][u] = ok ? va[idx] : (T)0;
        const unsigned loc = (unsigned)(c - row0);
        const bool in = loc < (unsigned)TILE;
        const T* pw = in ? s_w + (size_t)loc * K + slot : W + (size_t)c * K + slot;
        const T* pu = pw + (in ? d_smem : d_glob);
        wv[r][u] = *reinterpret_cast<const Pack<T, V>*>(pw);
        uv[r][u] = *reinterpret_cast<const Pack<T, V>*>(pu);
      }
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int q = 0; q < V; ++q) acc[r][q] = fma(a[r][u], fnma_rn(cf[q], uv[r][u].v[q], wv[r][u].v[q]), acc[r][q]);
  };
  int i = 0;
  for (; i + U <= lmin; i += U) round(i, std::false_type{});
  for (; i < len; i += U) round(i, std::true_type{});
}

// LONGROWS = false: compiled without the shared long-row path (operators whose longest row has <= LONG non-zeros; the
// extra live state costs the lean kernel registers -- it spilled and ran 2.4x slower with the path compiled in).
template <typename T, int K, int MODE, bool LONGROWS>
__global__ void __launch_bounds__(TmaTiling<T, K, LONGROWS, MODE == STEP_FUSED>::NTHREADS, TmaTiling<T, K, LONGROWS, MODE == STEP_FUSED>::CTAS_PER_SM)
    k_stepA_tma(int n, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, const T* __restrict__ vals,
                const T* __restrict__ uc, const T* __restrict__ up, T* __restrict__ w, LzScal* sc, const double* scale_c,
                const double* scale_p, double* partial, StepOut so, ReplayArgs rp) {
  using C = TmaTiling<T, K, LONGROWS, MODE == STEP_FUSED>;
  constexpr int V = C::V, LPR = C::LPR, RPP = C::RPP, PASSES = C::PASSES, TILE = C::TILE, CAP = C::CAP;
  constexpr int STAGES = LONGROWS ? C::BAL_STAGES : C::STAGES;  // the balanced variant gives stages up for the row accumulators
  constexpr int UB = SLQ_TMA_UB;
  constexpr int U = SLQ_TMA_U, R = (MODE == STEP_FUSED) ? C::RF : C::R;
  constexpr bool FIRST = (MODE == STEP_FIRST), PLAIN = (MODE == STEP_PLAIN), REPLAY = (MODE == STEP_REPLAY);
  constexpr bool FUSED = (MODE == STEP_FUSED);  // uc = the pending w of the previous step, up = its vector u (see StepMode)
  static_assert(!(FUSED && LONGROWS), "the fused step exists for the lean variant only");
  constexpr bool HAS_UP = !FIRST && !PLAIN;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bar_full[STAGES], bar_empty[STAGES];
  constexpr int RED = C::CW * K > C::NTHREADS ? C::CW * K : C::NTHREADS;
  constexpr int SCR = RED * 8 > C::CT * 16 ? RED * 8 : C::CT * 16;
  __shared__ __align__(16) unsigned char s_scratch[SCR];
  double* s_red = reinterpret_cast<double*>(s_scratch);  // alpha partials per consumer warp; re-used by the finaliser
  T* s_part = reinterpret_cast<T*>(s_scratch);           // long rows: one partial probe-row per row group
  __shared__ T s_rc[K], s_brp[K];
  __shared__ T s_cf[(REPLAY || FUSED) ? K : 1], s_yc[REPLAY ? K : 1], s_y0[REPLAY ? K : 1];  // alpha r_c (replay: of this step; fused: of the previous one), Y coefficients

  unsigned char* const base = smem_raw;  // bulk copies need 16-byte alignment only
  T* const s_acc = reinterpret_cast<T*>(smem_raw + (size_t)STAGES * C::STAGE_BYTES);  // balanced variant: TILE x K row sums
  int* const s_crow = reinterpret_cast<int*>(s_acc + (size_t)TILE * K);                // ... and one continuation row per group
  unsigned char* const s_ring = reinterpret_cast<unsigned char*>(s_crow) + (((size_t)RPP * 4 + 16 + 15) / 16) * 16;  // gather ring
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], C::CW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (LONGROWS) {
    for (int t = tid; t < TILE * K; t += C::NTHREADS) s_acc[t] = (T)0;
  }
  if (REPLAY) {
    // recorded recurrence: every per-probe scalar of this step comes from the replay table (no grid-wide reductions)
    for (int t = tid; t < K; t += C::NTHREADS) {
      s_rc[t] = (T)rp.row[t];
      s_brp[t] = (T)rp.row[MAXK + t];
      s_cf[REPLAY ? t : 0] = (T)rp.row[2 * MAXK + t];
      s_yc[REPLAY ? t : 0] = (T)rp.row[3 * MAXK + t];
      s_y0[REPLAY ? t : 0] = rp.first ? (T)rp.y0[t] : (T)0;
    }
  } else if (!PLAIN) {
    for (int t = tid; t < K; t += C::NTHREADS) {
      s_rc[t] = (T)scale_c[t];
      s_brp[t] = FIRST ? (T)0 : (T)(sc->beta[t]) * (T)scale_p[t];
      if (FUSED) s_cf[FUSED ? t : 0] = (T)sc->coef[t];  // pending update of the previous step (0 at the first)
    }
  }
  __syncthreads();

  const int ntiles = (n + TILE - 1) / TILE;
  double acc_a[V], acc_w[FUSED ? V : 1], acc_v[FUSED ? V : 1];  // <q_c, w>; fused: <w, w>, <v_c, v_c>
#pragma unroll
  for (int q = 0; q < V; ++q) acc_a[q] = 0.0;
#pragma unroll
  for (int q = 0; q < (FUSED ? V : 1); ++q) acc_w[q] = acc_v[q] = 0.0;

  if (warp == C::CW) {
    // ------------------------------------------------ producer warp ------------------------------------------------
    int stage = 0;
    uint32_t phase = 0;
    int pf_stage = -1, pf_cnt = 0, pf_row0 = 0;  // the stage issued one iteration ago (optional L2 prefetcher)
    uint32_t pf_phase = 0;
    // Once the index slice of a stage has landed, touch the probe rows it points at with prefetch.global.L2: the one
    // first-touch (DRAM) gather per matrix row then costs the consumers an L2 hit instead of a full DRAM round trip.
    auto prefetch_stage = [&]() {
      // (Round 1 ran this for the irregular variant; a folded-column experiment -- every gather an L2 hit, same kernel
      // time -- showed the single producer warp spending ~30k cycles per chunk issuing 3840 scattered prefetches and
      // thereby gating the whole CTA: profiles/r02_c3_experiments.md.  The consumers' cp.async ring hides that latency.)
      if (pf_stage < 0 || !SLQ_TMA_PREFETCH) return;
      mbar_wait(&bar_full[pf_stage], pf_phase);
      const int* p_ci = reinterpret_cast<const int*>(base + (size_t)pf_stage * C::STAGE_BYTES + C::OFF_CI);
      constexpr int LINES = (K * (int)sizeof(T) + 127) / 128;
      for (int i = lane; i < pf_cnt * LINES; i += 32) {
        const int c = p_ci[i / LINES];
        if (!LONGROWS && (unsigned)(c - pf_row0) < (unsigned)TILE) continue;  // staged with the tile
        const char* a = reinterpret_cast<const char*>(uc + (size_t)c * K) + (i % LINES) * 128;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
      }
    };
    int rpn[C::NRP];  // row pointers of the NEXT tile: loaded one tile ahead so their latency is off the issue path
    auto load_rowptr = [&](int tile) {
      const int r0 = tile * TILE;
      const int nr = min(TILE, n - r0);
#pragma unroll
      for (int j = 0; j < C::NRP; ++j) rpn[j] = (tile < ntiles) ? __ldg(rowptr + r0 + min(lane + 32 * j, nr)) : 0;
    };
    load_rowptr(blockIdx.x);
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int row0 = tile * TILE;
      const int nrows = min(TILE, n - row0);
      int rpv[C::NRP];
#pragma unroll
      for (int j = 0; j < C::NRP; ++j) rpv[j] = rpn[j];
      load_rowptr(tile + gridDim.x);
      const int segA = __shfl_sync(0xffffffffu, rpv[0], 0);
      const int segB = __shfl_sync(0xffffffffu, rpv[C::NRP - 1], 31);  // index >= TILE + 1 > nrows: clamped to rowptr[row0+nrows]
      const int lo0 = segA & ~3;                                      // 16-byte aligned start of the index slice
      const int nch = max(1, (segB - lo0 + CAP - 1) / CAP);
      for (int ch = 0; ch < nch; ++ch) {
        mbar_wait(&bar_empty[stage], phase ^ 1u);
        unsigned char* S = base + (size_t)stage * C::STAGE_BYTES;
        int* s_rp = reinterpret_cast<int*>(S + C::OFF_RP);
        int* s_ci = reinterpret_cast<int*>(S + C::OFF_CI);
        T* s_va = reinterpret_cast<T*>(S + C::OFF_VA);
        int* s_hdr = reinterpret_cast<int*>(S + C::OFF_HDR);
#pragma unroll
        for (int j = 0; j < C::NRP; ++j) s_rp[lane + 32 * j] = rpv[j];
        const int lo = lo0 + ch * CAP;
        const int hi = min(lo + CAP, segB);
        const int cnt = max(hi - lo, 0);
        const int nb = cnt & ~3;  // elements moved by the bulk copies; the (< 4) tail goes through the lanes
        if (lane < cnt - nb) {
          s_ci[nb + lane] = __ldg(colidx + lo + nb + lane);
          s_va[nb + lane] = __ldg(vals + lo + nb + lane);
        }
        const bool last = (ch == nch - 1);
        __syncwarp();
        if (lane == 0) {
          s_hdr[0] = lo;
          s_hdr[1] = hi;
          s_hdr[2] = last ? 1 : 0;
        }
        __syncwarp();
        if (lane == 0) {
          const uint32_t b_ci = (uint32_t)nb * 4u, b_va = (uint32_t)nb * (uint32_t)sizeof(T);
          // (fused: in-tile columns are formed from BOTH tiles, so u_p travels with every chunk like u_c)
          const uint32_t b_up = (!LONGROWS && HAS_UP && (last || FUSED)) ? (uint32_t)nrows * K * (uint32_t)sizeof(T) : 0u;
          const uint32_t b_uc = LONGROWS ? 0u : (uint32_t)nrows * K * (uint32_t)sizeof(T);  // every chunk: in-tile columns read it
          const uint32_t total = b_ci + b_va + b_up + b_uc;
          if (total > 0) {
            mbar_arrive_expect_tx(&bar_full[stage], total);
            if (nb > 0) {
              bulk_g2s(s_ci, colidx + lo, b_ci, &bar_full[stage]);
              bulk_g2s(s_va, vals + lo, b_va, &bar_full[stage]);
            }
            if (b_uc) bulk_g2s(S + C::OFF_UC, uc + (size_t)row0 * K, b_uc, &bar_full[stage]);
            if (b_up) bulk_g2s(S + C::OFF_UP, up + (size_t)row0 * K, b_up, &bar_full[stage]);
          } else {
            mbar_arrive(&bar_full[stage]);
          }
        }
        __syncwarp();
        prefetch_stage();  // for the stage issued one iteration ago (its copies have landed by now, or nearly)
        pf_stage = stage;
        pf_phase = phase;
        pf_cnt = cnt;
        pf_row0 = row0;
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
    prefetch_stage();
  } else {
    // ------------------------------------------------ consumer warps -----------------------------------------------
    const int sub = tid % LPR, rloc = tid / LPR;
    const int slot = sub * V;
    int stage = 0;
    uint32_t phase = 0;
    // this thread's probe columns never change: keep their scales in registers (the per-row shared-memory reads were
    // a fifth of the kernel's L1/shared-memory wavefronts, and that pipe is the co-bottleneck next to HBM)
    constexpr int HOIST = sizeof(T) == 8 ? SLQ_TMA_HOIST : 0;  // the fp32 k = 64 variant has no registers to spare
    T rc_r[V], brp_r[V], cf_r[V];
#pragma unroll
    for (int q = 0; q < V; ++q) {
      rc_r[q] = (PLAIN || REPLAY || HOIST < 2) ? (T)1 : s_rc[slot + q];
      brp_r[q] = (PLAIN || REPLAY || HOIST < 1) ? (T)0 : s_brp[slot + q];
      cf_r[q] = FUSED ? s_cf[FUSED ? slot + q : 0] : (T)0;
    }
    if constexpr (!LONGROWS) {
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int row0 = tile * TILE;
      T acc[PASSES][V];
#pragma unroll
      for (int ps = 0; ps < PASSES; ++ps)
#pragma unroll
        for (int q = 0; q < V; ++q) acc[ps][q] = (T)0;
      bool last;
      do {
        mbar_wait(&bar_full[stage], phase);
        const unsigned char* S = base + (size_t)stage * C::STAGE_BYTES;
        const int* s_rp = reinterpret_cast<const int*>(S + C::OFF_RP);
        const int* s_ci = reinterpret_cast<const int*>(S + C::OFF_CI);
        const T* s_va = reinterpret_cast<const T*>(S + C::OFF_VA);
        const T* s_up = reinterpret_cast<const T*>(S + C::OFF_UP);
        const T* s_uc = reinterpret_cast<const T*>(S + C::OFF_UC);
        const int* s_hdr = reinterpret_cast<const int*>(S + C::OFF_HDR);
        const int lo = s_hdr[0], hi = s_hdr[1];
        last = s_hdr[2] != 0;
#pragma unroll
        for (int g = 0; g < PASSES / R; ++g) {
          int s[R], e[R];
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const int rl = (g * R + r) * RPP + rloc;
            s[r] = min(max(s_rp[rl], lo), hi) - lo;
            e[r] = min(max(s_rp[rl + 1], lo), hi) - lo;
          }
          T a2[R][V];
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int q = 0; q < V; ++q) a2[r][q] = acc[g * R + r][q];
          // replay: the Y rows this thread updates are requested BEFORE the gathers, so their DRAM latency hides behind them
          Pack<T, V> yrow[REPLAY ? R : 1];
          if constexpr (REPLAY) {
            if (last && !rp.first) {
#pragma unroll
              for (int r = 0; r < R; ++r) {
                const int rl = (g * R + r) * RPP + rloc;
                if (row0 + rl < n) yrow[r] = ld_pack<T, V, LD_STREAM>(reinterpret_cast<const T*>(rp.Y) + (size_t)(row0 + rl) * K + slot);
              }
            }
          }
          if constexpr (FUSED)
            tma_gather_rows_fused<T, K, V, SLQ_TMA_UF, R, TILE>(a2, s, e, s_ci, s_va, s_uc, s_up, row0, uc, up, slot, cf_r);
          else
            tma_gather_rows<T, K, V, U, R, TILE>(a2, s, e, s_ci, s_va, s_uc, row0, uc, slot);
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int q = 0; q < V; ++q) acc[g * R + r][q] = a2[r][q];
          if constexpr (FUSED) {
            if (last) {
#pragma unroll
              for (int r = 0; r < R; ++r) {
                const int rl = (g * R + r) * RPP + rloc;
                if (row0 + rl < n) {
                  const size_t off = (size_t)(row0 + rl) * K + slot;
                  const Pack<T, V> pw = *reinterpret_cast<const Pack<T, V>*>(s_uc + (size_t)rl * K + slot);  // pending w of the previous step
                  const Pack<T, V> pu = *reinterpret_cast<const Pack<T, V>*>(s_up + (size_t)rl * K + slot);  // its vector u
                  Pack<T, V> vn, wv;
#pragma unroll
                  for (int q = 0; q < V; ++q) {
                    const T rcq = HOIST >= 2 ? rc_r[q] : s_rc[slot + q];
                    const T brq = HOIST >= 1 ? brp_r[q] : s_brp[slot + q];
                    vn.v[q] = fnma_rn(cf_r[q], pu.v[q], pw.v[q]);                      // the current vector, formed at last
                    wv.v[q] = fnma_rn(brq, pu.v[q], mul_rn(a2[r][q], rcq));            // w = (A v) r_c - beta r_p u
                    acc_a[q] += (double)(vn.v[q] * rcq) * (double)wv.v[q];
                    acc_w[FUSED ? q : 0] += (double)wv.v[q] * (double)wv.v[q];
                    acc_v[FUSED ? q : 0] += (double)vn.v[q] * (double)vn.v[q];
                  }
                  if (rp.Y) st_pack<T, V, SLQ_TMA_FUSED_STREAM_ST != 0>(reinterpret_cast<T*>(rp.Y) + off, vn);
                  st_pack<T, V, SLQ_TMA_FUSED_STREAM_ST != 0>(w + off, wv);
                }
              }
            }
          } else
          if (last) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
              const int rl = (g * R + r) * RPP + rloc;
              if (row0 + rl < n) {
                const size_t off = (size_t)(row0 + rl) * K + slot;
                Pack<T, V> wv;
                if (PLAIN) {
#pragma unroll
                  for (int q = 0; q < V; ++q) wv.v[q] = a2[r][q];
                  st_pack<T, V, true>(w + off, wv);
                } else {
                  Pack<T, V> pv;
                  if (HAS_UP) pv = *reinterpret_cast<const Pack<T, V>*>(s_up + (size_t)rl * K + slot);
                  const Pack<T, V> qcr = *reinterpret_cast<const Pack<T, V>*>(s_uc + (size_t)rl * K + slot);
                  if constexpr (REPLAY) {
                    // v = (A u_c) r_c - beta_j r_p u_p - alpha_j r_c u_c in the arithmetic of step A + axpy;  Y += yc v
                    Pack<T, V> yv = yrow[REPLAY ? r : 0];
#pragma unroll
                    for (int q = 0; q < V; ++q) {
                      const T t1 = fnma_rn(s_brp[slot + q], pv.v[q], mul_rn(a2[r][q], s_rc[slot + q]));
                      wv.v[q] = fnma_rn(s_cf[slot + q], qcr.v[q], t1);
                      const T y_old = rp.first ? mul_rn(s_y0[slot + q], qcr.v[q]) : yv.v[q];
                      yv.v[q] = fma_rn(s_yc[slot + q], wv.v[q], y_old);
                    }
                    st_pack<T, V, false>(w + off, wv);
                    st_pack<T, V, true>(reinterpret_cast<T*>(rp.Y) + off, yv);
                  } else {
#pragma unroll
                  for (int q = 0; q < V; ++q) {
                    const T rcq = HOIST >= 2 ? rc_r[q] : s_rc[slot + q];
                    const T brq = HOIST >= 1 ? brp_r[q] : s_brp[slot + q];
                    wv.v[q] = FIRST ? mul_rn(a2[r][q], rcq) : fnma_rn(brq, pv.v[q], mul_rn(a2[r][q], rcq));
                    acc_a[q] += (double)(qcr.v[q] * rcq) * (double)wv.v[q];
                  }
                  st_pack<T, V, false>(w + off, wv);
                  }
                }
              }
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_empty[stage]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      } while (!last);
    }
    } else {
      // Irregular operators (some row longer than LONG): non-zero-balanced processing.  Rows of any length cost the
      // same per non-zero; row sums are accumulated in shared memory (s_acc) in a fixed order, hence deterministic.
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int row0 = tile * TILE;
      bool last;
      do {
        mbar_wait(&bar_full[stage], phase);
        const unsigned char* S = base + (size_t)stage * C::STAGE_BYTES;
        const int* s_rp = reinterpret_cast<const int*>(S + C::OFF_RP);
        const int* s_ci = reinterpret_cast<const int*>(S + C::OFF_CI);
        const T* s_va = reinterpret_cast<const T*>(S + C::OFF_VA);
        const T* s_up = reinterpret_cast<const T*>(S + C::OFF_UP);
        const T* s_uc = reinterpret_cast<const T*>(S + C::OFF_UC);
        const int* s_hdr = reinterpret_cast<const int*>(S + C::OFF_HDR);
        const int lo = s_hdr[0], hi = s_hdr[1];
        last = s_hdr[2] != 0;
        // ---- phase 1: every row group takes an equal slice of the chunk's non-zeros, whatever rows they belong to ----
        const int vs = max(s_rp[0] - lo, 0), cnt = max(hi - lo, vs);  // [vs, cnt): the chunk's non-zeros of this tile
        const int per = ((cnt - vs + RPP - 1) / RPP) | 1;  // odd: the row groups of a warp read different smem banks
        int i = min(vs + rloc * per, cnt);
        const int end = min(i + per, cnt);
        int cont_row = -1, r_last = -1;
        bool last_is_cont = false;
        if (i < end) {
          int a = 0, b = TILE;  // largest r with rowptr[r] <= lo + i
          while (b - a > 1) {
            const int m = (a + b) >> 1;
            if (s_rp[m] <= lo + i) a = m; else b = m;
          }
          int r = a;
          int row_end = min(s_rp[r + 1], hi) - lo;
          bool cont = rloc > 0 && (s_rp[r] - lo) < i;  // the first row began in an earlier slice
          T part[V];
#pragma unroll
          for (int q = 0; q < V; ++q) part[q] = (T)0;
          // close the row segment held in `part`: a continuation goes to this group's slot of s_part, anything else is
          // added to the row's sum (only this group adds to a row that begins inside its slice: no race)
          auto flush = [&]() {
            Pack<T, V>* dst = reinterpret_cast<Pack<T, V>*>((cont ? s_part + (size_t)rloc * K : s_acc + (size_t)r * K) + slot);
            Pack<T, V> x;
            if (cont) {
#pragma unroll
              for (int q = 0; q < V; ++q) x.v[q] = part[q];
              cont_row = r;
            } else {
              x = *dst;
#pragma unroll
              for (int q = 0; q < V; ++q) x.v[q] += part[q];
            }
            *dst = x;
          };
          // The slice is walked as a flat list.  Gathers run RING non-zeros ahead of the FMAs through this thread's
          // private ring in shared memory: slot d of thread tid is the 16-byte word (d * CT + tid), so the lanes of a row
          // group write and read adjacent words (conflict-free) and nobody else ever touches them (no barrier needed,
          // cp.async.wait_group orders a thread's own copies).  One commit group per non-zero, empty past the end.
          constexpr int RING = C::RING, UW = 4;  // UW non-zeros per loop trip and per commit group
          static_assert(RING % UW == 0, "ring depth must be a multiple of the unroll");
          unsigned char* const my_ring = s_ring + (size_t)tid * 16;
          auto slot_ptr = [&](int d) { return my_ring + (size_t)d * C::CT * 16; };
          int issued = i;
#pragma unroll
          for (int d = 0; d < RING; d += UW) {
#pragma unroll
            for (int u = 0; u < UW; ++u)
              if (issued + u < end) cp_async16(slot_ptr(d + u), uc + (size_t)s_ci[issued + u] * K + slot);
            cp_async_commit();
            issued += UW;
          }
          int rs = 0;
          for (; i < end; i += UW) {
            cp_async_wait<RING / UW - 1>();  // the oldest group (non-zeros i .. i+UW-1) has landed
            Pack<T, V> xv[UW];
            T av[UW];
#pragma unroll
            for (int u = 0; u < UW; ++u) {
              xv[u] = *reinterpret_cast<const Pack<T, V>*>(slot_ptr(rs + u));
              av[u] = (i + u < end) ? s_va[i + u] : (T)0;
            }
            if (i + UW <= row_end && i + UW <= end) {  // common case: the whole group belongs to the open row
#pragma unroll
              for (int u = 0; u < UW; ++u)
#pragma unroll
                for (int q = 0; q < V; ++q) part[q] = fma(av[u], xv[u].v[q], part[q]);
            } else {
#pragma unroll
              for (int u = 0; u < UW; ++u)
                if (i + u < end) {
                  if (i + u >= row_end) {
                    flush();
                    cont = false;
#pragma unroll
                    for (int q = 0; q < V; ++q) part[q] = (T)0;
                    do {
                      ++r;
                      row_end = min(s_rp[r + 1], hi) - lo;
                    } while (i + u >= row_end && r < TILE - 1);  // skip empty rows
                  }
#pragma unroll
                  for (int q = 0; q < V; ++q) part[q] = fma(av[u], xv[u].v[q], part[q]);
                }
            }
#pragma unroll
            for (int u = 0; u < UW; ++u)  // refill the slots just read
              if (issued + u < end) cp_async16(slot_ptr(rs + u), uc + (size_t)s_ci[issued + u] * K + slot);
            cp_async_commit();
            issued += UW;
            rs = (rs + UW == RING) ? 0 : rs + UW;
          }
          cp_async_wait<0>();
          flush();
          r_last = r;
          last_is_cont = cont;
        }
        if (sub == 0) s_crow[rloc] = cont_row;
        asm volatile("bar.sync 1, %0;" ::"n"(C::CT) : "memory");
        // ---- phase 2: the group in whose slice a row begins adds the later slices' partials of that row, in order ----
        if (r_last >= 0 && !last_is_cont && min(s_rp[r_last + 1], hi) - lo > end) {
          Pack<T, V>* dst = reinterpret_cast<Pack<T, V>*>(s_acc + (size_t)r_last * K + slot);
          Pack<T, V> x = *dst;
          for (int g2 = rloc + 1; g2 < RPP && s_crow[g2] == r_last; ++g2) {
            const Pack<T, V> y = *reinterpret_cast<const Pack<T, V>*>(s_part + (size_t)g2 * K + slot);
#pragma unroll
            for (int q = 0; q < V; ++q) x.v[q] += y.v[q];
          }
          *dst = x;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(C::CT) : "memory");
        if (last) {
          Pack<T, V> ytile[REPLAY ? PASSES : 1];
          if constexpr (REPLAY) {
            if (!rp.first) {
#pragma unroll
              for (int ps = 0; ps < PASSES; ++ps) {
                const int rl = ps * RPP + rloc;
                if (row0 + rl < n) ytile[ps] = ld_pack<T, V, LD_STREAM>(reinterpret_cast<const T*>(rp.Y) + (size_t)(row0 + rl) * K + slot);
              }
            }
          }
#pragma unroll
          for (int ps = 0; ps < PASSES; ++ps) {
            const int rl = ps * RPP + rloc;
            Pack<T, V>* ap = reinterpret_cast<Pack<T, V>*>(s_acc + (size_t)rl * K + slot);
            const Pack<T, V> av = *ap;
            Pack<T, V> zero;
#pragma unroll
            for (int q = 0; q < V; ++q) zero.v[q] = (T)0;
            *ap = zero;  // ready for the next tile
            if (row0 + rl < n) {
              const size_t off = (size_t)(row0 + rl) * K + slot;
              Pack<T, V> wv;
              if (PLAIN) {
                st_pack<T, V, true>(w + off, av);
              } else {
                Pack<T, V> pv;
                if (HAS_UP) pv = ld_pack<T, V, LD_STREAM>(up + off);
                const Pack<T, V> qcr = ld_pack<T, V, LD_READONLY>(uc + off);
                if constexpr (REPLAY) {
                  Pack<T, V> yv = ytile[REPLAY ? ps : 0];
#pragma unroll
                  for (int q = 0; q < V; ++q) {
                    const T t1 = fnma_rn(s_brp[slot + q], pv.v[q], mul_rn(av.v[q], s_rc[slot + q]));
                    wv.v[q] = fnma_rn(s_cf[slot + q], qcr.v[q], t1);
                    const T y_old = rp.first ? mul_rn(s_y0[slot + q], qcr.v[q]) : yv.v[q];
                    yv.v[q] = fma_rn(s_yc[slot + q], wv.v[q], y_old);
                  }
                  st_pack<T, V, false>(w + off, wv);
                  st_pack<T, V, true>(reinterpret_cast<T*>(rp.Y) + off, yv);
                } else {
#pragma unroll
                for (int q = 0; q < V; ++q) {
                  const T rcq = HOIST >= 2 ? rc_r[q] : s_rc[slot + q];
                  const T brq = HOIST >= 1 ? brp_r[q] : s_brp[slot + q];
                  wv.v[q] = FIRST ? mul_rn(av.v[q], rcq) : fnma_rn(brq, pv.v[q], mul_rn(av.v[q], rcq));
                  acc_a[q] += (double)(qcr.v[q] * rcq) * (double)wv.v[q];
                }
                st_pack<T, V, false>(w + off, wv);
                }
              }
            }
          }
          asm volatile("bar.sync 1, %0;" ::"n"(C::CT) : "memory");  // rows are zeroed before anyone adds to them again
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_empty[stage]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      } while (!last);
    }
    }
  }

  if constexpr (FUSED) {
    // three sums per probe column, reduced one after the other through the same scratch: partial[cta][which][column]
#pragma unroll
    for (int which = 0; which < 3; ++which) {
      if (warp < C::CW) {
#pragma unroll
        for (int q = 0; q < V; ++q) {
          double x = which == 0 ? acc_a[q] : (which == 1 ? acc_w[FUSED ? q : 0] : acc_v[FUSED ? q : 0]);
#pragma unroll
          for (int off = 16; off >= LPR; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
          if (lane < LPR) s_red[warp * K + lane * V + q] = x;
        }
      }
      __syncthreads();
      if (tid < K) {
        double sum = 0.0;
#pragma unroll
        for (int wi = 0; wi < C::CW; ++wi) sum += s_red[wi * K + tid];
        partial[((size_t)blockIdx.x * 3 + which) * K + tid] = sum;
      }
      __syncthreads();
    }
    if (elect_last_cta(&sc->counter)) finalize<T, K, FIN_FUSED>(partial, sc, rp.scale_out, scale_c, so, s_red);
  } else
  if (!PLAIN && !REPLAY) {
    // deterministic CTA reduction of the alpha partials: row groups inside a warp by shuffles, warps through shared memory
    if (warp < C::CW) {
#pragma unroll
      for (int q = 0; q < V; ++q) {
        double x = acc_a[q];
#pragma unroll
        for (int off = 16; off >= LPR; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
        acc_a[q] = x;
      }
      if (lane < LPR) {
#pragma unroll
        for (int q = 0; q < V; ++q) s_red[warp * K + lane * V + q] = acc_a[q];
      }
    }
    __syncthreads();
    if (tid < K) {
      double s = 0.0;
#pragma unroll
      for (int wi = 0; wi < C::CW; ++wi) s += s_red[wi * K + tid];
      partial[(size_t)blockIdx.x * K + tid] = s;
    }
    if (elect_last_cta(&sc->counter)) finalize<T, K, FIN_ALPHA>(partial, sc, nullptr, scale_c, so, s_red);
  }
}

template <typename T, int K, int MODE>
static int launch_stepA_tma(const slq_op* op, const T* uc, const T* up, T* w, LzScal* sc, const double* scale_c, const double* scale_p,
                            double* partial, const StepOut& so, cudaStream_t st, const ReplayArgs& rp = ReplayArgs{}) {
  using C = TmaTiling<T, K>;
  using CB = TmaTiling<T, K, true>;
  static DeviceOnce once;  // per (T, K, MODE) instantiation, tracked per device
  if (once.needed()) {
    SLQ_CUDA(cudaFuncSetAttribute(k_stepA_tma<T, K, MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES));
    SLQ_CUDA(cudaFuncSetAttribute(k_stepA_tma<T, K, MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CB::SMEM_BAL));
    once.done();
  }
  const int n = (int)op->n;
  const int ntiles = (n + C::TILE - 1) / C::TILE;
  const int g = ntiles < sm_count() ? ntiles : sm_count();
  const int ntiles_b = (n + CB::TILE - 1) / CB::TILE;
  const int gb = ntiles_b < CB::CTAS_PER_SM * sm_count() ? ntiles_b : CB::CTAS_PER_SM * sm_count();
  if (op->max_row_nnz > C::LONG)
    k_stepA_tma<T, K, MODE, true><<<gb, CB::NTHREADS, CB::SMEM_BAL, st>>>(n, op->rowptr, op->colidx, (const T*)op->vals, uc, up, w, sc,
                                                                         scale_c, scale_p, partial, so, rp);
  else
    k_stepA_tma<T, K, MODE, false><<<g, C::NTHREADS, C::SMEM_BYTES, st>>>(n, op->rowptr, op->colidx, (const T*)op->vals, uc, up, w, sc,
                                                                          scale_c, scale_p, partial, so, rp);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

// Fused step (lean variant only: the caller checks op->max_row_nnz <= LONG).  W / U: pending w and vector u of the previous
// step (the start block twice at the first step, where cf = 0); w_out, v_out: this step's w and the formed vector.
template <typename T, int K>
static int launch_stepA_fused(const slq_op* op, const T* W, const T* U, T* w_out, T* v_out, LzScal* sc, const double* scale_c,
                              const double* scale_p, double* scale_out, double* partial, const StepOut& so, cudaStream_t st) {
  using C = TmaTiling<T, K, false, true>;
  static DeviceOnce once;
  if (once.needed()) {
    SLQ_CUDA(cudaFuncSetAttribute(k_stepA_tma<T, K, STEP_FUSED, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES));
    once.done();
  }
  const int n = (int)op->n;
  const int ntiles = (n + C::TILE - 1) / C::TILE;
  const int g = ntiles < sm_count() ? ntiles : sm_count();
  ReplayArgs rp;
  rp.Y = v_out;
  rp.scale_out = scale_out;
  k_stepA_tma<T, K, STEP_FUSED, false><<<g, C::NTHREADS, C::SMEM_BYTES, st>>>(n, op->rowptr, op->colidx, (const T*)op->vals, W, U, w_out, sc,
                                                                              scale_c, scale_p, partial, so, rp);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // namespace slq
