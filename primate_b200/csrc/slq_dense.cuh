// Dense symmetric operator: Y[n x K] = A[n x n] X[n x K] on a probe block (sm_100a).  Included by slq_lanczos.cu.
//
// Replaces the reference's Eigen GEMV on a per-call column-major copy of A (include/eigen_operators.h:24-30), blocked
// over the K probes: A (n^2 s bytes, 8.6 GB at config C4) is read ONCE per block of probes instead of once per probe.
// At K = 32 fp64 the product needs 8 flop per byte of A, i.e. it sits at the crossover of the HBM roofline
// (8.6 GB / 6.5 TB/s = 1.3 ms) and the fp64 FMA roofline (69 GFLOP / ~37 TFLOP/s = 1.9 ms); below K ~ 24 it is HBM-bound.
//
// One CTA owns 128 rows of Y.  Because A is symmetric, the 128 x KC tile A[row0 .. row0+127][c0 .. c0+KC-1] is staged
// TRANSPOSED for free: for every kk the slice A[c0+kk][row0 .. row0+127] is one contiguous 1 KB (fp64) line, fetched by
// one cp.async.bulk; the matching KC x K slab of X is one contiguous bulk copy.  A 4-stage mbarrier ring keeps three
// chunks in flight.  Each of the 128 threads keeps an 8 x (K/8) register tile; per kk it reads its 8 A values and K/8
// X values from shared memory with conflict-free 16-byte loads (rows and columns are interleaved in pairs across the
// threads) and issues 8 K/8 FMAs.
#pragma once

namespace slq {

template <typename T, int K>
struct DenseTiling {
  static constexpr int TR = 128;      // rows of Y per CTA
  static constexpr int KC = 16;       // columns of A (rows of X) per stage
  static constexpr int STAGES = 4;
  static constexpr int NTD = 128;     // threads: 8 (column groups) x 16 (row groups)
  static constexpr int CW = K / 8;    // columns per thread
  static constexpr size_t A_BYTES = (size_t)KC * TR * sizeof(T);
  static constexpr size_t X_BYTES = (size_t)KC * K * sizeof(T);
  static constexpr size_t STAGE_BYTES = A_BYTES + X_BYTES;
  static constexpr size_t SMEM_BYTES = STAGE_BYTES * STAGES;
};

template <typename T, int K>
constexpr bool dense_tma_supported() {
  return K >= 8 && K <= 64;
}

template <typename T, int K>
__global__ void __launch_bounds__(DenseTiling<T, K>::NTD, 2)
    k_dense_tma(int n, const T* __restrict__ A, int64_t ld, const T* __restrict__ X, T* __restrict__ Y) {
  using D = DenseTiling<T, K>;
  constexpr int TR = D::TR, KC = D::KC, STAGES = D::STAGES, CW = D::CW;
  constexpr int CP = CW >= 2 ? CW / 2 : 1;  // column pairs (or the single column) per thread
  constexpr int CV = CW >= 2 ? 2 : 1;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bar_full[STAGES];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tx = tid & 7, ty = tid >> 3;  // ty = 0..15
  const int row0 = blockIdx.x * TR;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&bar_full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = n / KC;
  auto issue = [&](int c, int s) {  // warp 0 only
    unsigned char* S = smem_raw + (size_t)s * D::STAGE_BYTES;
    if (lane == 0) mbar_arrive_expect_tx(&bar_full[s], (uint32_t)D::STAGE_BYTES);
    __syncwarp();
    if (lane < KC) bulk_g2s(S + (size_t)lane * TR * sizeof(T), A + (size_t)(c * KC + lane) * ld + row0, TR * sizeof(T), &bar_full[s]);
    if (lane == KC) bulk_g2s(S + D::A_BYTES, X + (size_t)c * KC * K, (uint32_t)D::X_BYTES, &bar_full[s]);
  };
  if (warp == 0) {
    for (int c = 0; c < STAGES - 1 && c < nchunks; ++c) issue(c, c);
  }
  T acc[4][2][CP][CV];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int p = 0; p < CP; ++p)
#pragma unroll
        for (int v = 0; v < CV; ++v) acc[i][j][p][v] = (T)0;

  for (int c = 0; c < nchunks; ++c) {
    const int s = c % STAGES;
    // refill the slot freed by the previous iteration (all threads passed its __syncthreads)
    if (warp == 0 && c + STAGES - 1 < nchunks) issue(c + STAGES - 1, (c + STAGES - 1) % STAGES);
    mbar_wait(&bar_full[s], (uint32_t)((c / STAGES) & 1));
    const T* As = reinterpret_cast<const T*>(smem_raw + (size_t)s * D::STAGE_BYTES);
    const T* Xs = reinterpret_cast<const T*>(smem_raw + (size_t)s * D::STAGE_BYTES + D::A_BYTES);
#pragma unroll
    for (int kk = 0; kk < KC; ++kk) {
      Pack<T, 2> a[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const Pack<T, 2>*>(As + kk * TR + i * 32 + ty * 2);
      Pack<T, CV> b[CP];
#pragma unroll
      for (int p = 0; p < CP; ++p) b[p] = *reinterpret_cast<const Pack<T, CV>*>(Xs + kk * K + p * 16 + tx * CV);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int p = 0; p < CP; ++p)
#pragma unroll
            for (int v = 0; v < CV; ++v) acc[i][j][p][v] = fma(a[i].v[j], b[p].v[v], acc[i][j][p][v]);
    }
    __syncthreads();  // everyone is done with slot s before warp 0 refills it at the top of the next iteration
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int r = row0 + i * 32 + ty * 2 + j;
#pragma unroll
      for (int p = 0; p < CP; ++p) {
        Pack<T, CV> o;
#pragma unroll
        for (int v = 0; v < CV; ++v) o.v[v] = acc[i][j][p][v];
        *reinterpret_cast<Pack<T, CV>*>(Y + (size_t)r * K + p * 16 + tx * CV) = o;
      }
    }
}

template <typename T, int K>
static bool dense_tma_ok(const slq_op* op, const T* X, const T* Y) {
  using D = DenseTiling<T, K>;
  return op->n % D::TR == 0 && op->n % D::KC == 0 && (op->ld * (int64_t)sizeof(T)) % 16 == 0 && op->n < (int64_t)INT32_MAX &&
         (((uintptr_t)op->dense | (uintptr_t)X | (uintptr_t)Y) & 15) == 0;
}

template <typename T, int K>
static int launch_dense_tma(const slq_op* op, const T* X, T* Y, cudaStream_t st) {
  using D = DenseTiling<T, K>;
  static bool configured = false;
  if (!configured) {
    SLQ_CUDA(cudaFuncSetAttribute(k_dense_tma<T, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)D::SMEM_BYTES));
    configured = true;
  }
  k_dense_tma<T, K><<<(unsigned)(op->n / D::TR), D::NTD, D::SMEM_BYTES, st>>>((int)op->n, (const T*)op->dense, op->ld, X, Y);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // namespace slq
