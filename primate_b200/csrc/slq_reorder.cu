// Operator ingestion helpers: a locality ordering for CSR operators, and the row gather that moves probe blocks
// between the caller's row order and the operator's.
//
// Why: on B200 the step-A kernel is bound by L2<->SM traffic, not by DRAM (ncu: 1.11 GB of L2->L1 reads per launch on
// the 1000x1000 Laplacian against 0.58 GB of DRAM reads; the sum of L2 reads and writes sits at the same ~6.4 TB/s as
// a plain copy).  In the natural (lexicographic) order of a grid operator only the +-1 neighbours of a row live in the
// same row tile; the +-g neighbours are fetched from L2 again by every tile that needs them (3.0 distinct probe rows
// fetched per row of the tile).  Grouping rows into breadth-first clusters of about one tile makes most neighbours
// tile-local (1.4 distinct rows per row), so they are L1 hits.  A symmetric permutation P A P^T has the same spectrum
// and, started from P x, the same Lanczos coefficients (only the summation order of the dot products changes, ~1e-16).
// The reference has no counterpart: it converts the matrix to CSC on every call (src/primate/_lanczos.cpp:89-93).
#include <algorithm>
#include <cstring>
#include <cstdlib>
#include <deque>
#include <thread>
#include <vector>

#include "slq_internal.cuh"

namespace slq {

// out[r][:] = in[idx[r]][:]   (rows of k scalars; one 16-byte slice per thread where the row allows it)
template <typename T, int V>
__global__ void __launch_bounds__(NT) k_gather_rows(const T* __restrict__ in, const int32_t* __restrict__ idx, int64_t n, int k,
                                                     T* __restrict__ out) {
  const int lpr = k / V;  // threads per row
  const int64_t total = n * lpr;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < total; e += (int64_t)gridDim.x * NT) {
    const int64_t r = e / lpr;
    const int s = (int)(e - r * lpr) * V;
    const int64_t src = __ldg(idx + r);
    Pack<T, V> x = ld_pack<T, V, LD_STREAM>(in + src * k + s);
    st_pack<T, V, true>(out + r * k + s, x);
  }
}

// out[idx[r]][:] = in[r][:]
template <typename T, int V>
__global__ void __launch_bounds__(NT) k_scatter_rows(const T* __restrict__ in, const int32_t* __restrict__ idx, int64_t n, int k,
                                                      T* __restrict__ out) {
  const int lpr = k / V;
  const int64_t total = n * lpr;
  for (int64_t e = (int64_t)blockIdx.x * NT + threadIdx.x; e < total; e += (int64_t)gridDim.x * NT) {
    const int64_t r = e / lpr;
    const int s = (int)(e - r * lpr) * V;
    const int64_t dst = __ldg(idx + r);
    Pack<T, V> x = ld_pack<T, V, LD_STREAM>(in + r * k + s);
    st_pack<T, V, true>(out + dst * k + s, x);
  }
}

template <typename T>
static int move_rows(const void* in, const int32_t* idx, int64_t n, int k, void* out, bool scatter, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return fail(SLQ_ERR_CUDA, "no CUDA device");
  constexpr int VMAX = 16 / (int)sizeof(T);
  const int64_t work = n * (k >= VMAX ? k / VMAX : k);
  int64_t g = (work + NT - 1) / NT;
  if (g > (int64_t)sms * 8) g = (int64_t)sms * 8;
  if (g < 1) g = 1;
  if (k % VMAX == 0) {
    if (scatter)
      k_scatter_rows<T, VMAX><<<(unsigned)g, NT, 0, st>>>((const T*)in, idx, n, k, (T*)out);
    else
      k_gather_rows<T, VMAX><<<(unsigned)g, NT, 0, st>>>((const T*)in, idx, n, k, (T*)out);
  } else {
    if (scatter)
      k_scatter_rows<T, 1><<<(unsigned)g, NT, 0, st>>>((const T*)in, idx, n, k, (T*)out);
    else
      k_gather_rows<T, 1><<<(unsigned)g, NT, 0, st>>>((const T*)in, idx, n, k, (T*)out);
  }
  SLQ_LAUNCH_CHECK();
  return SLQ_OK;
}

}  // namespace slq

using namespace slq;

extern "C" {

int slq_block_gather_rows(const void* in_dev, const int32_t* idx_dev, int64_t n, int k, int dtype, void* out_dev, slq_stream stream) {
  SLQ_REQUIRE(in_dev && idx_dev && out_dev && in_dev != out_dev, "invalid argument");
  SLQ_REQUIRE(n > 0 && k > 0, "invalid shape");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  return dtype == SLQ_F64 ? move_rows<double>(in_dev, idx_dev, n, k, out_dev, false, (cudaStream_t)stream)
                          : move_rows<float>(in_dev, idx_dev, n, k, out_dev, false, (cudaStream_t)stream);
}

int slq_block_scatter_rows(const void* in_dev, const int32_t* idx_dev, int64_t n, int k, int dtype, void* out_dev, slq_stream stream) {
  SLQ_REQUIRE(in_dev && idx_dev && out_dev && in_dev != out_dev, "invalid argument");
  SLQ_REQUIRE(n > 0 && k > 0, "invalid shape");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  return dtype == SLQ_F64 ? move_rows<double>(in_dev, idx_dev, n, k, out_dev, true, (cudaStream_t)stream)
                          : move_rows<float>(in_dev, idx_dev, n, k, out_dev, true, (cudaStream_t)stream);
}

}  // extern "C"

// Host threads for the ingestion helpers below (SLQ_HOST_THREADS overrides; small inputs stay on one thread).
static int host_threads(int64_t work_items) {
  int t = (int)std::thread::hardware_concurrency();
  if (const char* e = getenv("SLQ_HOST_THREADS")) t = atoi(e);
  if (t > 32) t = 32;
  const int64_t by_work = work_items / 65536 + 1;
  if (t > by_work) t = (int)by_work;
  return t < 1 ? 1 : t;
}

template <typename F>
static void parallel_ranges(int nthreads, F&& body) {  // body(thread index); thread 0 runs on the caller
  if (nthreads <= 1) {
    body(0);
    return;
  }
  std::vector<std::thread> pool;
  pool.reserve(nthreads - 1);
  for (int t = 1; t < nthreads; ++t) pool.emplace_back([&body, t]() { body(t); });
  body(0);
  for (auto& th : pool) th.join();
}

// Greedy breadth-first clustering of rows [r0, r1) of the (symmetric) sparsity graph into groups of `tile` consecutive
// new indices (neighbours outside the range are ignored).  Each cluster is a BFS ball grown from a seed on the boundary
// of the clusters already made (FIFO of left-over frontier vertices), so consecutive clusters are adjacent as well.
static void cluster_range(int64_t r0, int64_t r1, const int32_t* rowptr, const int32_t* colidx, int tile, unsigned char* state,
                          int32_t* perm_out) {
  std::vector<int32_t> queue;  // BFS queue of the current cluster
  queue.reserve((size_t)tile * 8);
  std::deque<int32_t> pending;  // left-over frontier: preferred seeds
  int64_t pos = r0, scan = r0;
  while (pos < r1) {
    int32_t seed = -1;
    while (!pending.empty()) {
      const int32_t c = pending.front();
      pending.pop_front();
      if (state[c] == 0) {
        seed = c;
        break;
      }
    }
    if (seed < 0) {
      while (state[scan] != 0) ++scan;
      seed = (int32_t)scan;
    }
    queue.clear();
    queue.push_back(seed);
    state[seed] = 1;
    size_t head = 0;
    int cnt = 0;
    while (head < queue.size() && cnt < tile) {
      const int32_t v = queue[head++];
      perm_out[pos++] = v;
      state[v] = 2;
      ++cnt;
      for (int32_t i = rowptr[v]; i < rowptr[v + 1]; ++i) {
        const int32_t u = colidx[i];
        if (u >= r0 && u < r1 && state[u] == 0) {
          state[u] = 1;
          queue.push_back(u);
        }
      }
    }
    for (; head < queue.size(); ++head) {  // not placed: free again, and first in line as seeds
      const int32_t u = queue[head];
      state[u] = 0;
      pending.push_back(u);
    }
    if (pending.size() > (size_t)(r1 - r0)) {  // keep the seed list bounded: drop stale entries
      std::deque<int32_t> keep;
      for (int32_t u : pending)
        if (state[u] == 0) keep.push_back(u);
      pending.swap(keep);
    }
  }
}

extern "C" {

// O(n + nnz) on the host.  Large operators are cut into one contiguous row range per host thread (range starts are
// multiples of `tile`, so clusters still coincide with the SpMM's row tiles) and the ranges are clustered concurrently:
// the ordering is a per-operator ingestion cost, paid again by every process that ingests the matrix.
int slq_csr_cluster_order(int64_t n, const int32_t* rowptr, const int32_t* colidx, int tile, int32_t* perm_out) {
  SLQ_REQUIRE(n > 0 && rowptr && perm_out && tile > 0, "invalid argument");
  SLQ_REQUIRE(n < (int64_t)INT32_MAX, "n does not fit int32");
  std::vector<unsigned char> state((size_t)n, 0);  // 0 free, 1 queued in the current cluster, 2 placed
  const int nt = host_threads(n / 4);
  const int64_t tiles = (n + tile - 1) / tile;
  const int64_t per = ((tiles + nt - 1) / nt) * tile;  // rows per range
  parallel_ranges(nt, [&](int t) {
    const int64_t r0 = std::min<int64_t>(n, (int64_t)t * per), r1 = std::min<int64_t>(n, r0 + per);
    if (r1 > r0) cluster_range(r0, r1, rowptr, colidx, tile, state.data(), perm_out);
  });
  return SLQ_OK;
}

// B = P A P^T for perm[new] = old, CSR in / CSR out with sorted column indices (host arrays; out arrays sized like in).
// Rows are independent: contiguous ranges of new rows, balanced by their non-zeros, go to the host threads.
int slq_csr_permute(int64_t n, const int32_t* rowptr, const int32_t* colidx, const void* vals, int dtype, const int32_t* perm,
                    int32_t* rowptr_out, int32_t* colidx_out, void* vals_out) {
  SLQ_REQUIRE(n > 0 && rowptr && perm && rowptr_out, "invalid argument");
  SLQ_REQUIRE(dtype == SLQ_F32 || dtype == SLQ_F64, "Only 32- or 64-bit floats are supported.");
  const size_t s = dtype_size(dtype);
  std::vector<int32_t> inv((size_t)n, -1);
  for (int64_t i = 0; i < n; ++i) {
    SLQ_REQUIRE(perm[i] >= 0 && perm[i] < n, "perm is not a permutation of 0..n-1");
    inv[perm[i]] = (int32_t)i;
  }
  rowptr_out[0] = 0;
  for (int64_t i = 0; i < n; ++i) rowptr_out[i + 1] = rowptr_out[i] + (rowptr[perm[i] + 1] - rowptr[perm[i]]);
  const int64_t nnz = rowptr_out[n];
  const int nt = host_threads(nnz / 8);
  const char* vin = (const char*)vals;
  char* vout = (char*)vals_out;
  parallel_ranges(nt, [&](int t) {
    // new rows [i0, i1): the range whose non-zeros are the t-th share of nnz
    const int64_t lo = nnz * t / nt, hi = nnz * (t + 1) / nt;
    const int64_t i0 = std::lower_bound(rowptr_out, rowptr_out + n, (int32_t)lo) - rowptr_out;
    const int64_t i1 = (t + 1 == nt) ? n : std::lower_bound(rowptr_out, rowptr_out + n, (int32_t)hi) - rowptr_out;
    std::vector<std::pair<int32_t, int32_t>> row;  // (new column, source position)
    for (int64_t i = i0; i < i1; ++i) {
      const int32_t old = perm[i];
      const int32_t b = rowptr[old], e = rowptr[old + 1];
      row.resize((size_t)(e - b));
      for (int32_t p = b; p < e; ++p) row[(size_t)(p - b)] = {inv[colidx[p]], p};
      if (e - b <= 24) {  // stencil-sized rows: insertion sort
        for (int a = 1; a < e - b; ++a) {
          const auto x = row[(size_t)a];
          int c = a - 1;
          for (; c >= 0 && row[(size_t)c].first > x.first; --c) row[(size_t)c + 1] = row[(size_t)c];
          row[(size_t)c + 1] = x;
        }
      } else {
        std::sort(row.begin(), row.end());
      }
      int32_t o = rowptr_out[i];
      if (vin && vout && s == 8) {
        const double* vi = (const double*)vin;
        double* vo = (double*)vout;
        for (const auto& el : row) {
          colidx_out[o] = el.first;
          vo[o++] = vi[el.second];
        }
      } else if (vin && vout) {
        const float* vi = (const float*)vin;
        float* vo = (float*)vout;
        for (const auto& el : row) {
          colidx_out[o] = el.first;
          vo[o++] = vi[el.second];
        }
      } else {
        for (const auto& el : row) colidx_out[o++] = el.first;
      }
    }
  });
  return SLQ_OK;
}

}  // extern "C"
