// the IRN random walk of src/indexing.py:116-150.
//   to_transition_matrix: T0 = A^beta with every column scaled to sum 1 (:117-119), then `times` squarings T <- T T (:120-121)
//   propagate (second half of propagate_to_edge, :144-148): rw = (x * (1 - edge)) T
// The element-wise half is fused here (power, column sums and scaling in two passes over A, HBM-bound: the
// reference makes four: pow, sum, divide, and the copy `trans_mat` starts as); the squarings are plain dense
// fp32 GEMMs of an [hw, hw] matrix with itself — library work (cuBLAS through torch.matmul in the Python mirror, as
// the task allows for plain GEMMs), because the reference's result is defined in fp32 and eight chained bf16
// squarings of a stochastic matrix would not reproduce it.
#include "host_util.cuh"

using namespace mb200;

namespace {

// colsum[j] += sum_i a[i][j]^beta over this block's rows
__global__ void __launch_bounds__(256) irn_colsum_kernel(const float* __restrict__ a, int n, float beta, int rows_per_block,
                                                         float* __restrict__ colsum) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= n) return;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(n, r0 + rows_per_block);
  float s = 0.f;
  for (int i = r0; i < r1; ++i) s += powf(__ldg(a + static_cast<size_t>(i) * n + j), beta);
  atomicAdd(colsum + j, s);
}
__global__ void __launch_bounds__(256) irn_scale_kernel(const float* __restrict__ a, int n, float beta,
                                                        const float* __restrict__ colsum, float* __restrict__ t) {
  const size_t total = static_cast<size_t>(n) * n;
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < total; i += static_cast<size_t>(gridDim.x) * 256)
    t[i] = powf(__ldg(a + i), beta) / __ldg(colsum + i % n);
}
// y[c][p] = x[c][p] * (1 - edge[p])
__global__ void __launch_bounds__(256) irn_mask_kernel(const float* __restrict__ x, const float* __restrict__ edge, int planes,
                                                       int hw, float* __restrict__ y) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < planes * hw) y[i] = x[i] * (1.f - __ldg(edge + i % hw));
}

int need_sm100_irn() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}

}  // namespace

extern "C" {

int muscle_irn_transition_init(const float* affinity, int n, float beta, float* colsum_scratch, float* trans_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100_irn()) return rc;
  if (!affinity || !colsum_scratch || !trans_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n <= 0 || n > 46340) return fail(MUSCLE_ERR_BAD_ARG, "matrix size out of range");
  MB_CUDA(cudaMemsetAsync(colsum_scratch, 0, static_cast<size_t>(n) * sizeof(float), stream));
  const int rpb = 64;
  irn_colsum_kernel<<<dim3((n + 255) / 256, (n + rpb - 1) / rpb), 256, 0, stream>>>(affinity, n, beta, rpb, colsum_scratch);
  MB_LAUNCH_CHECK();
  const size_t total = static_cast<size_t>(n) * n;
  const int grid = static_cast<int>(std::min<size_t>((total + 255) / 256, static_cast<size_t>(32) * device_info().sm_count));
  irn_scale_kernel<<<grid, 256, 0, stream>>>(affinity, n, beta, colsum_scratch, trans_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_irn_mask_edge(const float* x, const float* edge, int planes, int hw, float* y, void* stream_) {
  if (int rc = need_sm100_irn()) return rc;
  if (!x || !edge || !y) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (planes <= 0 || hw <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  irn_mask_kernel<<<(planes * hw + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream_)>>>(x, edge, planes, hw, y);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // extern "C"
