// Dense symmetric operator: Y[n x K] = A[n x n] X[n x K] on a probe block (sm_100a).  Included by slq_lanczos.cu.
//
// Replaces the reference's Eigen GEMV on a per-call column-major copy of A (include/eigen_operators.h:24-30), blocked
// over the K probes: A (n^2 s bytes, 8.6 GB at config C4) is read ONCE per block of probes instead of once per probe.
// At K = 32 fp64 the product needs 8 flop per byte of A, i.e. it sits at the crossover of the HBM roofline
// (8.6 GB / 6.5 TB/s = 1.3 ms) and the fp64 FMA roofline (69 GFLOP / ~37 TFLOP/s = 1.9 ms); below K ~ 24 it is HBM-bound.
//
// One CTA owns 16 RPT rows of Y (128, 112 or 96).  Because A is symmetric, the 128 x KC tile A[row0 .. row0+127][c0 .. c0+KC-1] is staged
// TRANSPOSED for free: for every kk the slice A[c0+kk][row0 .. row0+127] is one contiguous (<= 1 KB) line, fetched by
// one cp.async.bulk; the matching KC x K slab of X is one contiguous bulk copy.  A 4-stage mbarrier ring keeps three
// chunks in flight.  Each of the 128 consumer threads keeps an RPT x (K/8) register tile (RPT = 8, 7 or 6 rows); per kk
// it reads its RPT A values and K/8 X values from shared memory with conflict-free 16-byte loads (rows and columns are
// interleaved in pairs across the threads) and issues RPT K/8 FMAs.
#pragma once

namespace slq {

template <typename T, int K, int RPT = 8>
struct DenseTiling {
  static constexpr int TR = 16 * RPT;  // rows of Y per CTA (RPT rows per thread: 128, 112 or 96)
  static constexpr int KC = 16;       // columns of A (rows of X) per stage
  static constexpr size_t A_BYTES = (size_t)KC * TR * sizeof(T);
  static constexpr size_t X_BYTES = (size_t)KC * K * sizeof(T);
  static constexpr size_t STAGE_BYTES = A_BYTES + X_BYTES;
  // as many stages as fit beside a second resident CTA (2 x 110 KB of the SM's 227 KB), between 3 and 8
  static constexpr int STAGES_FIT = (int)((110 * 1024) / STAGE_BYTES);
  static constexpr int STAGES = STAGES_FIT > 8 ? 8 : (STAGES_FIT < 3 ? 3 : STAGES_FIT);
  static constexpr int NCONS = 128;   // consumer threads: 8 (column groups) x 16 (row groups)
  static constexpr int NTD = NCONS + 32;  // + one producer warp
  static constexpr int CW = K / 8;    // columns per thread
  static constexpr size_t SMEM_BYTES = STAGE_BYTES * STAGES;
};

template <typename T, int K>
constexpr bool dense_tma_supported() {
  return K >= 8 && K <= 64;
}

// Warp-specialised: warp 4 is the producer (waits for a stage to be released by the four consumer warps, then issues
// its KC + 1 bulk copies); the consumer warps never meet at a CTA barrier, they only wait on the stage's "full"
// mbarrier and release it through the "empty" one, so they drift apart by up to STAGES - 1 chunks instead of stalling
// on each other every 16 columns (ncu before: 24% of the issue stalls were the per-chunk __syncthreads, the fp64 pipe
// 60% busy).  RPT rows per thread sets the CTA height: the host picks the one that balances the row tiles over the
// SMs (n = 32 768 on 148 SMs at two CTAs per SM: 293 tiles of 112 rows, at most 224 rows per SM instead of 256).
template <typename T, int K, int RPT>
__global__ void __launch_bounds__(DenseTiling<T, K, RPT>::NTD, 2)
    k_dense_tma(int n, int npad, const T* __restrict__ A, int64_t ld, const T* __restrict__ X, T* __restrict__ Y) {
  using D = DenseTiling<T, K, RPT>;
  constexpr int TR = D::TR, KC = D::KC, STAGES = D::STAGES, CW = D::CW;
  constexpr int CP = CW >= 2 ? CW / 2 : 1;  // column pairs (or the single column) per thread
  constexpr int CV = CW >= 2 ? 2 : 1;
  constexpr int NP = RPT / 2, ODD = RPT % 2;  // row pairs per thread (+ one single row when RPT is odd)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bar_full[STAGES], bar_empty[STAGES];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row0 = blockIdx.x * TR;
  // A is npad x npad (npad = n rounded up to 16, rows / columns >= n zero: slq_dense_create_padded), the panels have n
  // rows: tiles and chunks are whole 16-byte units of the padded operator, the last X slab is short, stores stop at n.
  const int rows = min(TR, npad - row0);
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], D::NCONS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (npad != n) {  // rows of the last X slab past n are never copied: make sure they hold zeros, not stale bits
    for (int s = 0; s < STAGES; ++s) {
      T* Xz = reinterpret_cast<T*>(smem_raw + (size_t)s * D::STAGE_BYTES + D::A_BYTES);
      for (int i = tid; i < KC * K; i += D::NTD) Xz[i] = (T)0;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy zeros before async-proxy (bulk copy) writes
  }
  __syncthreads();
  // split-K: blockIdx.y owns the chunk range [c0, c0 + nchunks) and writes its partial product to slab blockIdx.y of Y
  const int per_split = (npad / KC + gridDim.y - 1) / gridDim.y;
  const int c0 = blockIdx.y * per_split;
  const int nchunks = max(0, min(npad / KC, c0 + per_split) - c0);
  Y += (size_t)blockIdx.y * (size_t)n * K;
  if (warp == D::NCONS / 32) {  // ---- producer warp
    const uint32_t a_line = (uint32_t)(rows * sizeof(T));
    for (int ci = 0; ci < nchunks; ++ci) {
      const int c = c0 + ci;
      const int s = ci % STAGES;
      if (ci >= STAGES) mbar_wait(&bar_empty[s], (uint32_t)(((ci / STAGES) - 1) & 1));
      unsigned char* S = smem_raw + (size_t)s * D::STAGE_BYTES;
      const uint32_t x_bytes = (uint32_t)(min(KC, n - c * KC) * K * sizeof(T));
      if (lane == 0) mbar_arrive_expect_tx(&bar_full[s], (uint32_t)(KC * a_line + x_bytes));
      __syncwarp();
      if (lane < KC) bulk_g2s(S + (size_t)lane * TR * sizeof(T), A + (size_t)(c * KC + lane) * ld + row0, a_line, &bar_full[s]);
      if (lane == KC) bulk_g2s(S + D::A_BYTES, X + (size_t)c * KC * K, x_bytes, &bar_full[s]);
    }
    return;
  }
  // ---- consumer warps
  const int tx = tid & 7, ty = tid >> 3;  // ty = 0..15
  T acc[RPT][CP][CV];
#pragma unroll
  for (int i = 0; i < RPT; ++i)
#pragma unroll
    for (int p = 0; p < CP; ++p)
#pragma unroll
      for (int v = 0; v < CV; ++v) acc[i][p][v] = (T)0;

  for (int ci = 0; ci < nchunks; ++ci) {
    const int s = ci % STAGES;
    mbar_wait(&bar_full[s], (uint32_t)((ci / STAGES) & 1));
    const T* As = reinterpret_cast<const T*>(smem_raw + (size_t)s * D::STAGE_BYTES);
    const T* Xs = reinterpret_cast<const T*>(smem_raw + (size_t)s * D::STAGE_BYTES + D::A_BYTES);
#pragma unroll
    for (int kk = 0; kk < KC; ++kk) {
      T a[RPT];
#pragma unroll
      for (int i = 0; i < NP; ++i) {
        const Pack<T, 2> ap = *reinterpret_cast<const Pack<T, 2>*>(As + kk * TR + i * 32 + ty * 2);
        a[2 * i] = ap.v[0];
        a[2 * i + 1] = ap.v[1];
      }
      if (ODD) a[RPT - 1] = As[kk * TR + NP * 32 + ty];
      Pack<T, CV> b[CP];
#pragma unroll
      for (int p = 0; p < CP; ++p) b[p] = *reinterpret_cast<const Pack<T, CV>*>(Xs + kk * K + p * 16 + tx * CV);
#pragma unroll
      for (int i = 0; i < RPT; ++i)
#pragma unroll
        for (int p = 0; p < CP; ++p)
#pragma unroll
          for (int v = 0; v < CV; ++v) acc[i][p][v] = fma(a[i], b[p].v[v], acc[i][p][v]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&bar_empty[s]);  // this warp is done with slot s
  }
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int rl = (i < 2 * NP) ? (i / 2) * 32 + ty * 2 + (i & 1) : NP * 32 + ty;
    if (row0 + rl < n) {
      const int r = row0 + rl;
#pragma unroll
      for (int p = 0; p < CP; ++p) {
        Pack<T, CV> o;
#pragma unroll
        for (int v = 0; v < CV; ++v) o.v[v] = acc[i][p][v];
        *reinterpret_cast<Pack<T, CV>*>(Y + (size_t)r * K + p * 16 + tx * CV) = o;
      }
    }
  }
}

// ---- fp64 tensor-core variant (DMMA) -----------------------------------------------------------------------------------
// Same producer/consumer ring, but the consumers issue mma.sync.m8n8k4.f64 (SASS DMMA) instead of scalar FMAs: the only
// GEMM-shaped fp64 work on the path (north_star: "the only place fp64 DMMA tensor-core tiles are considered").
// Each consumer warp owns 16 rows x K columns of Y: per k4 step 2 A fragments (As[kk][m], transposed tile = the
// layout DMMA's row-major A fragment wants) + K/8 B fragments (the X slab) feed 2 K/8 DMMAs.  Shared-memory rows are
// padded by 4 doubles: a 64-bit warp load is served one half-warp (lanes 0-15: g = 0..3, t = 0..3) at a time, and
// with a row stride = 4 (mod 16) doubles the sixteen 8-byte words of a half-warp, t * stride + g, fall into sixteen
// different bank pairs (2 wavefronts per warp load: the minimum; a pad of 8 measured 2x that, ncu r02g).  The fragment-to-lane maps are PTX's for m8n8k4 f64:
//   A (8x4, row): lane l holds A[l / 4][l % 4];  B (4x8, col): lane l holds B[l % 4][l / 4];
//   C (8x8): lane l holds C[l / 4][2 (l % 4)], C[l / 4][2 (l % 4) + 1].
template <int K, int RPT>
struct DmmaTiling {
  static constexpr int TR = 16 * RPT;   // rows of Y per CTA: one 16-row slab per consumer warp
  static constexpr int KC = 16;         // columns of A (rows of X) per stage
  static constexpr int TRP = TR + 4;    // padded shared-memory row strides (doubles): = 4 (mod 16)
  static constexpr int KP = K + 4;
  static constexpr int CTAS = RPT > 8 ? 1 : 2;  // resident CTAs per SM the tiling is sized for
  static constexpr size_t A_BYTES = (size_t)KC * TRP * sizeof(double);
  static constexpr size_t X_BYTES = (size_t)KC * KP * sizeof(double);
  static constexpr size_t STAGE_BYTES = A_BYTES + X_BYTES;
  static constexpr int STAGES_FIT = (int)(((CTAS == 1 ? 216 : 110) * 1024) / STAGE_BYTES);
  static constexpr int STAGES = STAGES_FIT > 8 ? 8 : (STAGES_FIT < 3 ? 3 : STAGES_FIT);
  static constexpr int NB = K / 8;      // 8-column blocks of Y
  // Consumer warp grid WM (row groups) x WN (column groups); a warp owns MB 8-row blocks x NBW 8-column blocks of Y.
  // Default: one 16-row slab x all columns per warp (WM = RPT).  A warp runs on sub-partition (warp id % 4) of the SM,
  // and each sub-partition has its own fp64 tensor pipe: the 14 slab warps of the 224-row CTA land 4 / 4 / 3 / 3 on
  // the four pipes, so the kernel cannot pass 87.5 % of the DMMA peak (measured: 84 % busy, r02c_ncu_dense_tiled.txt).
  // The 224-row CTA therefore uses 4 x 2 warps of 56 rows x K / 2 columns: two consumer warps on every pipe.
  static constexpr bool GRID8 = (RPT == 14 && NB >= 2);
  static constexpr int WM = GRID8 ? 4 : RPT;
  static constexpr int WN = GRID8 ? 2 : 1;
  static constexpr int MB = TR / 8 / WM;
  static constexpr int NBW = NB / WN;
  static constexpr int NCONS = 32 * WM * WN;
  static constexpr int NTD = NCONS + 32;
  static_assert(MB * WM * 8 == TR && NBW * WN == NB, "warp grid must tile the CTA's block of Y");
  static_assert(TRP % 16 == 4 && (KP % 16 == 4 || KP % 16 == 12), "row strides must spread the four k-rows over the bank pairs");
  static constexpr size_t SMEM_BYTES = STAGE_BYTES * STAGES;
};

__device__ __forceinline__ void dmma_884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

//
// TILED: A points at the band-major copy made by slq_dense_set_tiled (dense_tiled_elems): the stage image of chunk c of
// this CTA's band is ONE contiguous 16 x TRP block, fetched by one bulk copy, and a CTA walks its band front to back --
// 148 sequential streams through HBM.  Reading the row-major operator instead costs KC strided 1.8 KB lines per stage,
// ~112 matrix rows open per CTA at a time: DRAM page conflicts hold that form (and cuBLAS DGEMM on the same operand) at
// 4.2 TB/s where a stream reads 6.9 TB/s (profiles/r02c_dense_tiled.md).  Same arithmetic in the same order either way.
template <int K, int RPT, bool TILED>
__global__ void __launch_bounds__(DmmaTiling<K, RPT>::NTD, DmmaTiling<K, RPT>::CTAS)
    k_dense_dmma(int n, int npad, const double* __restrict__ A, int64_t ld, const double* __restrict__ X, double* __restrict__ Y) {
  using D = DmmaTiling<K, RPT>;
  constexpr int TR = D::TR, KC = D::KC, STAGES = D::STAGES, TRP = D::TRP, KP = D::KP;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bar_full[STAGES], bar_empty[STAGES];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row0 = blockIdx.x * TR;
  const int rows = min(TR, npad - row0);  // npad is a multiple of 16: whole 16-row warp slabs, whole 16-byte units
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], D::NCONS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (npad != n) {  // X rows past n are never copied (the matching columns of the padded A are zero): keep them finite
    for (int s = 0; s < STAGES; ++s) {
      double* Xz = reinterpret_cast<double*>(smem_raw + (size_t)s * D::STAGE_BYTES + D::A_BYTES);
      for (int i = tid; i < KC * KP; i += D::NTD) Xz[i] = 0.0;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  // split-K: blockIdx.y owns the chunk range [c0, c0 + nchunks) and writes its partial product to slab blockIdx.y of Y
  const int per_split = (npad / KC + gridDim.y - 1) / gridDim.y;
  const int c0 = blockIdx.y * per_split;
  const int nchunks = max(0, min(npad / KC, c0 + per_split) - c0);
  Y += (size_t)blockIdx.y * (size_t)n * K;
  if (warp == D::NCONS / 32) {  // ---- producer warp: KC slices of A (one per column, contiguous by symmetry) + KC rows of X
    const uint32_t a_line = (uint32_t)(rows * sizeof(double));
    constexpr uint32_t x_line = (uint32_t)(K * sizeof(double));
    for (int ci = 0; ci < nchunks; ++ci) {
      const int c = c0 + ci;
      const int s = ci % STAGES;
      if (ci >= STAGES) mbar_wait(&bar_empty[s], (uint32_t)(((ci / STAGES) - 1) & 1));
      unsigned char* S = smem_raw + (size_t)s * D::STAGE_BYTES;
      const int xr = min(KC, n - c * KC);  // rows of X in this slab (short only in the last chunk of a padded operator)
      if constexpr (TILED) {
        if (lane == 0) {
          mbar_arrive_expect_tx(&bar_full[s], (uint32_t)(D::A_BYTES + xr * x_line));
          bulk_g2s(S, A + ((size_t)blockIdx.x * (size_t)(npad / KC) + (size_t)c) * (size_t)(KC * TRP), (uint32_t)D::A_BYTES, &bar_full[s]);
        }
        __syncwarp();
        if (lane >= KC && lane - KC < xr)
          bulk_g2s(S + D::A_BYTES + (size_t)(lane - KC) * KP * sizeof(double), X + (size_t)(c * KC + lane - KC) * K, x_line, &bar_full[s]);
      } else {
      if (lane == 0) mbar_arrive_expect_tx(&bar_full[s], (uint32_t)(KC * a_line + xr * x_line));
      __syncwarp();
      if (lane < KC)
        bulk_g2s(S + (size_t)lane * TRP * sizeof(double), A + (size_t)(c * KC + lane) * ld + row0, a_line, &bar_full[s]);
      else if (lane - KC < xr)
        bulk_g2s(S + D::A_BYTES + (size_t)(lane - KC) * KP * sizeof(double), X + (size_t)(c * KC + lane - KC) * K, x_line, &bar_full[s]);
      }
    }
    return;
  }
  // ---- consumer warps: MB 8-row blocks from row rbase of the tile x NBW 8-column blocks from column block cb0
  constexpr int MB = D::MB, NBW = D::NBW;
  const int g = lane >> 2, t = lane & 3;
  const int rbase = (warp % D::WM) * (MB * 8), cb0 = (warp / D::WM) * NBW;
  const bool active = rbase < rows;  // the last tile may hold fewer rows (stale smem rows are computed on but never stored)
  double acc[MB][NBW][2];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb)
#pragma unroll
    for (int nb = 0; nb < NBW; ++nb) acc[mb][nb][0] = acc[mb][nb][1] = 0.0;
  for (int ci = 0; ci < nchunks; ++ci) {
    const int s = ci % STAGES;
    mbar_wait(&bar_full[s], (uint32_t)((ci / STAGES) & 1));
    const double* As = reinterpret_cast<const double*>(smem_raw + (size_t)s * D::STAGE_BYTES);
    const double* Xs = reinterpret_cast<const double*>(smem_raw + (size_t)s * D::STAGE_BYTES + D::A_BYTES);
    if (active) {
#pragma unroll
      for (int k0 = 0; k0 < KC; k0 += 4) {
        double a[MB], b[NBW];
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) a[mb] = As[(k0 + t) * TRP + rbase + mb * 8 + g];
#pragma unroll
        for (int nb = 0; nb < NBW; ++nb) b[nb] = Xs[(k0 + t) * KP + (cb0 + nb) * 8 + g];
#pragma unroll
        for (int mb = 0; mb < MB; ++mb)
#pragma unroll
          for (int nb = 0; nb < NBW; ++nb) dmma_884(acc[mb][nb], a[mb], b[nb]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&bar_empty[s]);
  }
  if (active) {
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) {
      const int r = row0 + rbase + mb * 8 + g;
      if (r >= n) continue;
#pragma unroll
      for (int nb = 0; nb < NBW; ++nb) {
        Pack<double, 2> o;
        o.v[0] = acc[mb][nb][0];
        o.v[1] = acc[mb][nb][1];
        *reinterpret_cast<Pack<double, 2>*>(Y + (size_t)r * K + (cb0 + nb) * 8 + 2 * t) = o;
      }
    }
  }
}

// Y = sum over the split-K slabs (fixed order: deterministic)
template <typename T>
__global__ void __launch_bounds__(256) k_sum_slabs(const T* __restrict__ part, int nslab, int64_t len, T* __restrict__ Y) {
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < len; e += (int64_t)gridDim.x * 256) {
    T s = part[e];
    for (int j = 1; j < nslab; ++j) s += part[(size_t)j * len + e];
    Y[e] = s;
  }
}

// Small operators leave most SMs idle (n = 1000: 11 row tiles) and every tile walks all n / 16 chunks one bulk-copy
// latency after the other.  With caller-provided scratch (slq_op_set_scratch) the chunk range is split over gridDim.y
// CTAs per tile, each writing a partial slab, and k_sum_slabs adds the slabs in order.
template <typename T, int K>
static int dense_splits(const slq_op* op, int tile_rows) {
  if (!op->scratch) return 1;
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  const int64_t tiles = (npad + tile_rows - 1) / tile_rows;
  const int sms = sm_count() > 0 ? sm_count() : 148;
  int64_t s = (2 * (int64_t)sms) / tiles;
  const int64_t nch = npad / 16;
  if (s > 16) s = 16;
  if (s > nch / 2) s = nch / 2;  // at least two chunks per split
  const int64_t fit = (int64_t)(op->scratch_bytes / ((size_t)op->n * K * sizeof(T)));
  if (s > fit) s = fit;
  return s < 2 ? 1 : (int)s;
}

template <int K, int RPT>
static int launch_dense_dmma_rpt(const slq_op* op, const double* X, double* Y, cudaStream_t st) {
  using D = DmmaTiling<K, RPT>;
  static DeviceOnce once;
  if (once.needed()) {
    SLQ_CUDA(cudaFuncSetAttribute(k_dense_dmma<K, RPT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)D::SMEM_BYTES));
    SLQ_CUDA(cudaFuncSetAttribute(k_dense_dmma<K, RPT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)D::SMEM_BYTES));
    once.done();
  }
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  const int splits = dense_splits<double, K>(op, D::TR);
  dim3 grid((unsigned)((npad + D::TR - 1) / D::TR), (unsigned)splits);
  if (op->tiled && op->tiled_rpt == RPT)
    k_dense_dmma<K, RPT, true><<<grid, D::NTD, D::SMEM_BYTES, st>>>((int)op->n, (int)npad, (const double*)op->tiled, op->ld, X,
                                                                    splits > 1 ? (double*)op->scratch : Y);
  else
    k_dense_dmma<K, RPT, false><<<grid, D::NTD, D::SMEM_BYTES, st>>>((int)op->n, (int)npad, (const double*)op->dense, op->ld, X,
                                                                     splits > 1 ? (double*)op->scratch : Y);
  SLQ_LAUNCH_CHECK();
  if (splits > 1) {
    const int64_t len = op->n * K;
    k_sum_slabs<double><<<(unsigned)std::min<int64_t>((len + 255) / 256, 1024), 256, 0, st>>>((const double*)op->scratch, splits, len, Y);
    SLQ_LAUNCH_CHECK();
  }
  return SLQ_OK;
}

template <typename T, int K>
static bool dense_tma_ok(const slq_op* op, const T* X, const T* Y) {
  using D = DenseTiling<T, K>;
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  return npad % D::KC == 0 && npad >= D::KC && op->ld >= npad && (op->ld * (int64_t)sizeof(T)) % 16 == 0 && npad < (int64_t)INT32_MAX &&
         (((uintptr_t)op->dense | (uintptr_t)X | (uintptr_t)Y) & 15) == 0;
}

template <typename T, int K, int RPT>
static int launch_dense_tma_rpt(const slq_op* op, const T* X, T* Y, cudaStream_t st) {
  using D = DenseTiling<T, K, RPT>;
  static DeviceOnce once;
  if (once.needed()) {
    SLQ_CUDA(cudaFuncSetAttribute(k_dense_tma<T, K, RPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)D::SMEM_BYTES));
    once.done();
  }
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  const int splits = dense_splits<T, K>(op, D::TR);
  dim3 grid((unsigned)((npad + D::TR - 1) / D::TR), (unsigned)splits);
  k_dense_tma<T, K, RPT><<<grid, D::NTD, D::SMEM_BYTES, st>>>((int)op->n, (int)npad, (const T*)op->dense, op->ld, X,
                                                             splits > 1 ? (T*)op->scratch : Y);
  SLQ_LAUNCH_CHECK();
  if (splits > 1) {
    const int64_t len = op->n * K;
    k_sum_slabs<T><<<(unsigned)std::min<int64_t>((len + 255) / 256, 1024), 256, 0, st>>>((const T*)op->scratch, splits, len, Y);
    SLQ_LAUNCH_CHECK();
  }
  return SLQ_OK;
}

// CTA height: dense_pick_rpt (slq_lanczos_common.cuh), the same rule the band-major copy was laid out with.
template <typename T, int K>
static int launch_dense_tma(const slq_op* op, const T* X, T* Y, cudaStream_t st) {
  const int64_t npad = op->npad > 0 ? op->npad : op->n;
  if constexpr (sizeof(T) == 8) {
    if (dense_use_dmma()) {  // fp64: tensor-core tiles, one 16-row slab per consumer warp (TR = 16 RPT)
      switch (dense_pick_rpt(npad, true)) {
        case 14: return launch_dense_dmma_rpt<K, 14>(op, (const double*)X, (double*)Y, st);
        case 6: return launch_dense_dmma_rpt<K, 6>(op, (const double*)X, (double*)Y, st);
        case 7: return launch_dense_dmma_rpt<K, 7>(op, (const double*)X, (double*)Y, st);
        default: return launch_dense_dmma_rpt<K, 8>(op, (const double*)X, (double*)Y, st);
      }
    }
  }
  switch (dense_pick_rpt(npad, false)) {
    case 6: return launch_dense_tma_rpt<T, K, 6>(op, X, Y, st);
    case 7: return launch_dense_tma_rpt<T, K, 7>(op, X, Y, st);
    default: return launch_dense_tma_rpt<T, K, 8>(op, X, Y, st);
  }
}

}  // namespace slq
