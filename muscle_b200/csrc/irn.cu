// the IRN random walk of src/indexing.py:116-150.
//   to_transition_matrix: T0 = A^beta with every column scaled to sum 1 (:117-119), then `times` squarings T <- T T (:120-121)
//   propagate (second half of propagate_to_edge, :144-148): rw = (x * (1 - edge)) T
// The element-wise half is fused here (power, column sums and scaling in two passes over A, HBM-bound: the
// reference makes four: pow, sum, divide, and the copy `trans_mat` starts as); the squarings are plain dense
// fp32 GEMMs of an [hw, hw] matrix with itself — library work (cuBLAS through torch.matmul in the Python mirror, as
// the task allows for plain GEMMs), because the reference's result is defined in fp32 and eight chained bf16
// squarings of a stochastic matrix would not reproduce it.
#include <cuda_bf16.h>

#include "host_util.cuh"
#include "ptx.cuh"

using namespace mb200;

namespace {

// ==========================================================================================================
// fp32-accurate products on the tensor pipe: x = h + m + l with three bf16 parts (24 mantissa bits), and
//   A B  ~=  Ah Bh + Ah Bm + Am Bh + Ah Bl + Al Bh + Am Bm      (the dropped terms are <= 2^-24 relative)
// accumulated in fp32 in TMEM.  Every entry of a transition matrix is >= 0, so there is no cancellation and the
// relative error of an output entry is that of its terms.  One CTA per 128 x 128 output tile, K in blocks of 64:
// per block three A tiles [128 x 64] (K-major as TMA delivers row-major rows) and three B tiles [64 x 128] (the
// row-major rows of B are its K index: read MN-major, exactly like the gradient pass of the InfoNCE reads its keys),
// 24 tcgen05.mma (six part pairs x four K = 16 steps).  The epilogue writes the fp32 tile and / or its three bf16 parts
// -- the operands of the next squaring -- so a chain of squarings never re-splits in a pass of its own.
constexpr int IG_BK = 64;
constexpr int IG_NST = 2;
constexpr int IG_TILE = 128 * 128;                    // bytes of one [128 x 64] or [64 x 128] bf16 tile
constexpr int IG_STAGE = 6 * IG_TILE;                 // Ah Am Al Bh Bm Bl
constexpr int IG_SMEM = IG_NST * IG_STAGE + 256 + 1024;
constexpr int IG_THREADS = 192;

struct IgMaps {
  CUtensorMap a[3], b[3];
};

__device__ __forceinline__ void split3(float x, __nv_bfloat16& h, __nv_bfloat16& m, __nv_bfloat16& l) {
  h = __float2bfloat16_rn(x);
  const float r1 = x - __bfloat162float(h);
  m = __float2bfloat16_rn(r1);
  l = __float2bfloat16_rn(r1 - __bfloat162float(m));
}

// parts[q][r][c] (q = 0 hi, 1 mid, 2 lo; [rows_pad][cols_pad], zero outside rows x cols)
__global__ void __launch_bounds__(256) irn_split3_kernel(const float* __restrict__ x, int rows, int cols, int rows_pad,
                                                         int cols_pad, __nv_bfloat16* __restrict__ parts) {
  const size_t plane = static_cast<size_t>(rows_pad) * cols_pad;
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < plane; i += static_cast<size_t>(gridDim.x) * 256) {
    const int r = static_cast<int>(i / cols_pad), c = static_cast<int>(i - static_cast<size_t>(r) * cols_pad);
    const float v = (r < rows && c < cols) ? __ldg(x + static_cast<size_t>(r) * cols + c) : 0.f;
    __nv_bfloat16 h, m, l;
    split3(v, h, m, l);
    parts[i] = h;
    parts[plane + i] = m;
    parts[2 * plane + i] = l;
  }
}

__global__ void __launch_bounds__(IG_THREADS, 1)
irn_matmul3_kernel(const __grid_constant__ IgMaps maps, int k_blocks, int m, int n, int m_pad, int n_pad,
                   float* __restrict__ c_out, __nv_bfloat16* __restrict__ c_parts) {
  extern __shared__ uint8_t ig_smem_raw[];
  const uint32_t base = (smem_u32(ig_smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = ig_smem_raw + (base - smem_u32(ig_smem_raw));
  const uint32_t bar0 = base + IG_NST * IG_STAGE;
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (IG_NST + s); };
  auto acc_full = [&](int a) { return bar0 + 8u * (2 * IG_NST + a); };
  auto acc_empty = [&](int a) { return bar0 + 8u * (2 * IG_NST + 2 + a); };
  volatile uint32_t* tptr = reinterpret_cast<volatile uint32_t*>(gen + IG_NST * IG_STAGE + 8 * (2 * IG_NST + 4));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.y * 128, col0 = blockIdx.x * 128;

  if (threadIdx.x == 0) {
    for (int s = 0; s < IG_NST; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(acc_full(a), 1); mbar_init(acc_empty(a), 4); }
    fence_barrier_init();
  }
  if (warp == 4) { tmem_alloc(smem_u32(const_cast<uint32_t*>(tptr)), 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tptr;

  if (warp == 5) {
    if (lane == 0) {
      for (int q = 0; q < 3; ++q) { tma_prefetch_desc(&maps.a[q]); tma_prefetch_desc(&maps.b[q]); }
      for (int kb = 0; kb < k_blocks; ++kb) {
        const int st = kb % IG_NST;
        mbar_wait_spin(empty(st), ((kb / IG_NST) & 1) ^ 1);
        mbar_arrive_expect_tx(full(st), IG_STAGE);
        const uint32_t sa = base + st * IG_STAGE;
        for (int q = 0; q < 3; ++q) {
          tma_load_2d(sa + q * IG_TILE, &maps.a[q], full(st), kb * IG_BK, row0);                  // [128 rows x 64 k]
          tma_load_2d(sa + (3 + q) * IG_TILE, &maps.b[q], full(st), col0, kb * IG_BK);            // [64 k x 64 n], n 0..63
          tma_load_2d(sa + (3 + q) * IG_TILE + 8192, &maps.b[q], full(st), col0 + 64, kb * IG_BK);  // n 64..127
        }
      }
    }
  } else if (warp == 4) {
    if (elect_one()) {
      constexpr uint32_t IDESC = umma_idesc_bf16(128, 128, 0, 1);      // A K-major, B MN-major
      // part pairs, small terms first: (m,m) (l,h) (h,l) (m,h) (h,m) (h,h)
      const int pa[6] = {1, 2, 0, 1, 0, 0}, pb[6] = {1, 0, 2, 0, 1, 0};
      for (int kb = 0; kb < k_blocks; ++kb) {
        const int st = kb % IG_NST, acc = kb & 1;
        mbar_wait_spin(acc_empty(acc), ((kb >> 1) & 1) ^ 1);
        mbar_wait_spin(full(st), (kb / IG_NST) & 1);
        tc_fence_after();
        const uint32_t sa = base + st * IG_STAGE;
#pragma unroll
        for (int t = 0; t < 6; ++t) {
          const uint32_t a_tile = sa + pa[t] * IG_TILE, b_tile = sa + (3 + pb[t]) * IG_TILE;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            mma_ss(tmem + acc * 128, umma_desc_sw128(a_tile + kk * 32, 16, 1024), umma_desc_sw128(b_tile + kk * 2048, 8192, 1024),
                   IDESC, (t | kk) != 0);
        }
        tc_commit(empty(st));
        tc_commit(acc_full(acc));
      }
    }
  } else {
    // The tensor pipe adds into its fp32 accumulator with truncation: over a long K the bias grows with the number of
    // additions (measured: 1e-4 relative per squaring at K = 4096 when the whole K was accumulated in TMEM).  So TMEM
    // holds ONE K block (24 products) at a time, in two alternating accumulators, and the running sum lives in
    // registers, added with round-to-nearest -- the scheme of the error-corrected tensor-core GEMMs in the literature.
    const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
    float sum[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) sum[i] = 0.f;
    for (int kb = 0; kb < k_blocks; ++kb) {
      const int acc = kb & 1;
      mbar_wait_warp(acc_full(acc), (kb >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_addr + acc * 128 + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) sum[c * 32 + i] += __uint_as_float(v[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(acc));
    }
    const int r = row0 + warp * 32 + lane;                 // < m_pad
    const size_t plane = static_cast<size_t>(m_pad) * n_pad;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int c0 = c * 32;
      if (c_out != nullptr && r < m) {
        float* dst = c_out + static_cast<size_t>(r) * n + col0 + c0;
        if ((n & 3) == 0 && col0 + c0 + 32 <= n) {
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<float4*>(dst + i) = make_float4(sum[c0 + i], sum[c0 + i + 1], sum[c0 + i + 2], sum[c0 + i + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (col0 + c0 + i < n) dst[i] = sum[c0 + i];
        }
      }
      if (c_parts != nullptr) {
        uint32_t ph[16], pm[16], pl[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          __nv_bfloat16 h0, m0, l0, h1, m1, l1;
          split3(sum[c0 + 2 * i], h0, m0, l0);
          split3(sum[c0 + 2 * i + 1], h1, m1, l1);
          ph[i] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
          pm[i] = static_cast<uint32_t>(__bfloat16_as_ushort(m0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(m1)) << 16);
          pl[i] = static_cast<uint32_t>(__bfloat16_as_ushort(l0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(l1)) << 16);
        }
        __nv_bfloat16* dst = c_parts + static_cast<size_t>(r) * n_pad + col0 + c0;       // 64-byte aligned
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
          *reinterpret_cast<uint4*>(dst + 2 * i) = make_uint4(ph[i], ph[i + 1], ph[i + 2], ph[i + 3]);
          *reinterpret_cast<uint4*>(dst + plane + 2 * i) = make_uint4(pm[i], pm[i + 1], pm[i + 2], pm[i + 3]);
          *reinterpret_cast<uint4*>(dst + 2 * plane + 2 * i) = make_uint4(pl[i], pl[i + 1], pl[i + 2], pl[i + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 256);
}

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

size_t muscle_irn_split3_bytes(int rows_pad, int cols_pad) {
  if (rows_pad <= 0 || cols_pad <= 0) return 0;
  return static_cast<size_t>(3) * rows_pad * cols_pad * sizeof(__nv_bfloat16);
}

int muscle_irn_split3(const float* x, int rows, int cols, int rows_pad, int cols_pad, void* parts, void* stream_) {
  if (int rc = need_sm100_irn()) return rc;
  if (!x || !parts) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (rows <= 0 || cols <= 0 || rows_pad < rows || cols_pad < cols || rows_pad % 128 || cols_pad % 128)
    return fail(MUSCLE_ERR_BAD_ARG, "padded sizes must be multiples of 128 that cover the matrix");
  const size_t plane = static_cast<size_t>(rows_pad) * cols_pad;
  const int grid = static_cast<int>(std::min<size_t>((plane + 255) / 256, static_cast<size_t>(32) * device_info().sm_count));
  irn_split3_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(x, rows, cols, rows_pad, cols_pad,
                                                                             static_cast<__nv_bfloat16*>(parts));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_irn_matmul_split3(const void* a_parts, const void* b_parts, int m_pad, int k_pad, int n_pad, int m, int n,
                             float* c_out, void* c_parts_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100_irn()) return rc;
  if (!a_parts || !b_parts || (!c_out && !c_parts_out)) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (m_pad <= 0 || k_pad <= 0 || n_pad <= 0 || m_pad % 128 || k_pad % 128 || n_pad % 128)
    return fail(MUSCLE_ERR_BAD_ARG, "padded sizes must be positive multiples of 128");
  if (m <= 0 || n <= 0 || m > m_pad || n > n_pad) return fail(MUSCLE_ERR_BAD_ARG, "m, n must lie inside the padded sizes");
  if (c_parts_out == a_parts || c_parts_out == b_parts) return fail(MUSCLE_ERR_BAD_ARG, "the product cannot be formed in place");
  IgMaps maps;
  const auto* a = static_cast<const __nv_bfloat16*>(a_parts);
  const auto* b = static_cast<const __nv_bfloat16*>(b_parts);
  for (int q = 0; q < 3; ++q) {
    if (int rc = make_bf16_box_map(&maps.a[q], a + static_cast<size_t>(q) * m_pad * k_pad, m_pad, k_pad, 128)) return rc;
    if (int rc = make_bf16_box_map(&maps.b[q], b + static_cast<size_t>(q) * k_pad * n_pad, k_pad, n_pad, 64)) return rc;
  }
  if (int rc = ensure_dynamic_smem<irn_matmul3_kernel>(IG_SMEM)) return rc;
  irn_matmul3_kernel<<<dim3(n_pad / 128, m_pad / 128), IG_THREADS, IG_SMEM, stream>>>(
      maps, k_pad / IG_BK, m, n, m_pad, n_pad, c_out, static_cast<__nv_bfloat16*>(c_parts_out));
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
