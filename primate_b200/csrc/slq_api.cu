// Library-level entry points of libslq: error reporting, device info, and the Krylov-basis combination used by
// f(A) X (src/primate/operators.py:124: y = ||x|| * Q @ (Y (f(theta) o Y[0,:]))).
#include <atomic>
#include <cstring>
#include <mutex>
#include <vector>

#include "slq_internal.cuh"

namespace slq {

char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return dev;
}

int sm_count() {
  static std::atomic<int> cached[MAX_DEVICES];  // zero-initialised: 0 = not queried yet
  const int dev = current_device();
  if (dev < 0) return -1;  // nothing cached: a device may be selected later
  if (dev < MAX_DEVICES) {
    const int c = cached[dev].load(std::memory_order_relaxed);
    if (c > 0) return c;
  }
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  if (dev < MAX_DEVICES) cached[dev].store(sms, std::memory_order_relaxed);
  return sms;
}

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- event profiling (off by default) -----------------------------------------------------------------------
struct ProfState {
  std::mutex mu;
  bool enabled = false;
  int cap = 0;
  std::vector<cudaEvent_t> ev[PROF_NCLASS];  // 2 events per slot
  int used[PROF_NCLASS] = {0, 0, 0, 0};
  long long seen[PROF_NCLASS] = {0, 0, 0, 0};
};
static ProfState g_prof;

int prof_begin(int cls, cudaStream_t st) {
  if (!g_prof.enabled) return -1;
  std::lock_guard<std::mutex> lk(g_prof.mu);
  g_prof.seen[cls]++;
  if (g_prof.used[cls] >= g_prof.cap) return -1;
  const int slot = g_prof.used[cls]++;
  if ((int)g_prof.ev[cls].size() < 2 * (slot + 1)) {
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) {
      g_prof.used[cls]--;
      return -1;
    }
    g_prof.ev[cls].push_back(a);
    g_prof.ev[cls].push_back(b);
  }
  cudaEventRecord(g_prof.ev[cls][2 * slot], st);
  return slot;
}

void prof_end(int cls, int slot, cudaStream_t st) {
  if (slot < 0) return;
  cudaEventRecord(g_prof.ev[cls][2 * slot + 1], st);
}

// Y[r][p] = sum_j Q[j][r][p] * (coef[p][j] * xnorm[p])
template <typename T, int K>
__global__ void __launch_bounds__(NT) k_basis_combine(const T* __restrict__ Q, int64_t n, int deg, const double* __restrict__ coeffs,
                                                       const double* __restrict__ xnorm2, T* __restrict__ Y) {
  using M = BlockMap<T, K>;
  constexpr int V = M::V, LPR = M::LPR;
  extern __shared__ double s_c[];  // deg x K, scaled coefficients
  for (int idx = threadIdx.x; idx < deg * K; idx += NT) {
    const int j = idx / K, p = idx % K;
    s_c[idx] = coeffs[(size_t)p * deg + j] * sqrt(xnorm2[p]);
  }
  __syncthreads();
  const int slot = (threadIdx.x % LPR) * V;
  const int64_t nvec = n * LPR;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < nvec; e += (int64_t)gridDim.x * NT) {
    double acc[V];
#pragma unroll
    for (int q = 0; q < V; ++q) acc[q] = 0.0;
    for (int j = 0; j < deg; ++j) {
      Pack<T, V> x = ld_pack<T, V, LD_STREAM>(Q + (size_t)j * n * K + e * V);
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] += (double)x.v[q] * s_c[j * K + slot + q];
    }
    Pack<T, V> o;
#pragma unroll
    for (int q = 0; q < V; ++q) o.v[q] = (T)acc[q];
    st_pack<T, V, false>(Y + e * V, o);
  }
}

template <typename T, int K>
static int combine_launch(const void* Q, int64_t n, int deg, const double* coeffs, const double* xnorm2, void* Y, cudaStream_t st) {
  const size_t smem = sizeof(double) * (size_t)deg * K;
  if (smem > 200 * 1024) return fail(SLQ_ERR_UNSUPPORTED, "deg*k too large for slq_basis_combine");
  if (smem > 48 * 1024) SLQ_CUDA(cudaFuncSetAttribute(k_basis_combine<T, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t need = (n * BlockMap<T, K>::LPR + NT - 1) / NT;
  int64_t cap = (int64_t)sm_count() * 4;
  int g = (int)(need < cap ? need : cap);
  if (g < 1) g = 1;
  k_basis_combine<T, K><<<g, NT, smem, st>>>((const T*)Q, n, deg, coeffs, xnorm2, (T*)Y);
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

template <typename T>
static int combine_dispatch(const void* Q, int64_t n, int k, int deg, const double* coeffs, const double* xnorm2, void* Y, cudaStream_t st) {
  switch (k) {
    case 1: return combine_launch<T, 1>(Q, n, deg, coeffs, xnorm2, Y, st);
    case 2: return combine_launch<T, 2>(Q, n, deg, coeffs, xnorm2, Y, st);
    case 4: return combine_launch<T, 4>(Q, n, deg, coeffs, xnorm2, Y, st);
    case 8: return combine_launch<T, 8>(Q, n, deg, coeffs, xnorm2, Y, st);
    case 16: return combine_launch<T, 16>(Q, n, deg, coeffs, xnorm2, Y, st);
    case 32: return combine_launch<T, 32>(Q, n, deg, coeffs, xnorm2, Y, st);
    case 64: return combine_launch<T, 64>(Q, n, deg, coeffs, xnorm2, Y, st);
    default: return fail(SLQ_ERR_ARG, "block width k=%d must be a power of two in [1,64]", k);
  }
}

}  // namespace slq

using namespace slq;

extern "C" {

int slq_version(void) { return 100; }

const char* slq_last_error(void) { return err_buf(); }

int slq_device_info(int* sm, int* cc_major, int* cc_minor) {
  int dev = 0;
  SLQ_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  SLQ_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (sm) *sm = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  return SLQ_OK;
}

long long slq_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int slq_profile_enable(int max_launches_per_class) {
  std::lock_guard<std::mutex> lk(g_prof.mu);
  g_prof.enabled = max_launches_per_class > 0;
  g_prof.cap = max_launches_per_class > 0 ? max_launches_per_class : 0;
  for (int c = 0; c < PROF_NCLASS; ++c) g_prof.used[c] = 0, g_prof.seen[c] = 0;
  return SLQ_OK;
}

int slq_profile_read(int cls, double* total_ms, long long* timed_launches, long long* seen_launches) {
  SLQ_REQUIRE(cls >= 0 && cls < PROF_NCLASS, "invalid profile class %d", cls);
  std::lock_guard<std::mutex> lk(g_prof.mu);
  double tot = 0.0;
  for (int s = 0; s < g_prof.used[cls]; ++s) {
    SLQ_CUDA(cudaEventSynchronize(g_prof.ev[cls][2 * s + 1]));
    float ms = 0.f;
    SLQ_CUDA(cudaEventElapsedTime(&ms, g_prof.ev[cls][2 * s], g_prof.ev[cls][2 * s + 1]));
    tot += ms;
  }
  if (total_ms) *total_ms = tot;
  if (timed_launches) *timed_launches = g_prof.used[cls];
  if (seen_launches) *seen_launches = g_prof.seen[cls];
  return SLQ_OK;
}

int slq_basis_combine(const void* Q_dev, int64_t n, int k, int deg, int dtype, const double* coeffs_dev, const double* xnorm2_dev,
                      void* Y_dev, slq_stream stream) {
  SLQ_REQUIRE(Q_dev && coeffs_dev && xnorm2_dev && Y_dev, "NULL argument");
  SLQ_REQUIRE(n > 0 && deg > 0, "invalid shape");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  if (sm_count() <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == SLQ_F64 ? combine_dispatch<double>(Q_dev, n, k, deg, coeffs_dev, xnorm2_dev, Y_dev, st)
                          : combine_dispatch<float>(Q_dev, n, k, deg, coeffs_dev, xnorm2_dev, Y_dev, st);
}

}  // extern "C"
