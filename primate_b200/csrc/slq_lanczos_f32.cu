// Explicit instantiation of the probe-blocked Lanczos driver for float panels (see slq_lanczos_impl.cuh).
#include "slq_lanczos_impl.cuh"

namespace slq {
template int lanczos_dispatch<float>(slq_op*, void*, int, int, double, int, int, void*, void*, int32_t*, double*, void*, char*, cudaStream_t);
template int matmat_dispatch<float>(slq_op*, const void*, void*, int, cudaStream_t);
template int replay_dispatch<float>(slq_op*, const void*, int, int, const void*, const void*, const int32_t*, const double*, const double*,
                                 void*, double*, double*, void*, void*, void*, cudaStream_t);
}  // namespace slq
