// Unit-test kernel for the tcgen05 building blocks of pair_kernel.cuh, one CTA, one tile:
//   S = A B^T           (TMA SW128 loads, K-major SS descriptors, tcgen05.ld 32x32b)
//   G = bf16(S) B       (tcgen05.st of packed bf16 into the S columns, TS MMA with the MN-major
//                        view of the same B tile)
// tests/ compares both against torch on the same bf16 inputs, so a wrong descriptor bit shows up
// here and not as a silently wrong loss.
#include "host_util.cuh"
#include "ptx.cuh"

using namespace mb200;

namespace {

template <int D>
__global__ void __launch_bounds__(128, 1)
debug_tile_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                  float* s_out, float* g_out) {
  constexpr int TILE_BYTES = 128 * D * 2;
  constexpr int PANEL_BYTES = 128 * 128;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_smem = base, b_smem = base + TILE_BYTES;
  const uint32_t bar_load = base + 2 * TILE_BYTES, bar_mma = bar_load + 8, bar_t = bar_load + 16;
  volatile uint32_t* tptr = reinterpret_cast<volatile uint32_t*>(gen + 2 * TILE_BYTES + 32);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    mbar_init(bar_t, 4);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(const_cast<uint32_t*>(tptr)), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tptr;
  constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
  constexpr uint32_t IDESC_G = umma_idesc_bf16(128, D, 0, 1);

  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_load, 2 * TILE_BYTES);
    for (int a = 0; a < D / 64; ++a) {
      tma_load_2d(a_smem + a * PANEL_BYTES, &tm_a, bar_load, a * 64, 0);
      tma_load_2d(b_smem + a * PANEL_BYTES, &tm_b, bar_load, a * 64, 0);
    }
    mbar_wait(bar_load, 0);
    tc_fence_after();
    for (int kk = 0; kk < D / 16; ++kk) {
      const uint32_t off = (kk >> 2) * PANEL_BYTES + (kk & 3) * 32;
      mma_ss(tmem, umma_desc_sw128(a_smem + off, 16, 1024), umma_desc_sw128(b_smem + off, 16, 1024),
             IDESC_S, kk > 0);
    }
    tc_commit(bar_mma);
  }
  __syncwarp();
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
  const int row = warp * 32 + lane;
  for (int c = 0; c < 4; ++c) {
    uint32_t v[32];
    tmem_ld32(tmem + lane_addr + c * 32, v);
    tmem_wait_ld();
    uint32_t packed[16];
    for (int i = 0; i < 32; ++i) s_out[row * 128 + c * 32 + i] = __uint_as_float(v[i]);
    for (int i = 0; i < 16; ++i)
      packed[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
    tmem_st16(tmem + lane_addr + c * 16, packed);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(bar_t);
  if (threadIdx.x == 0) {
    mbar_wait(bar_t, 0);
    tc_fence_after();
    for (int kk = 0; kk < 8; ++kk)
      mma_ts(tmem + 256, tmem + kk * 8, umma_desc_sw128(b_smem + kk * 2048, PANEL_BYTES, 1024),
             IDESC_G, kk > 0);
    tc_commit(bar_mma);
  }
  __syncwarp();
  mbar_wait(bar_mma, 1);
  tc_fence_after();
  for (int c = 0; c < D / 32; ++c) {
    uint32_t v[32];
    tmem_ld32(tmem + lane_addr + 256 + c * 32, v);
    tmem_wait_ld();
    for (int i = 0; i < 32; ++i) g_out[row * D + c * 32 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int D>
int launch_debug(const void* a, const void* b, float* s_out, float* g_out, cudaStream_t stream) {
  CUtensorMap tm_a, tm_b;
  if (int rc = make_bf16_tile_map(&tm_a, a, 128, D)) return rc;
  if (int rc = make_bf16_tile_map(&tm_b, b, 128, D)) return rc;
  const int smem = 2 * 128 * D * 2 + 64 + 1024;
  MB_CUDA(cudaFuncSetAttribute(debug_tile_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  debug_tile_kernel<D><<<1, 128, smem, stream>>>(tm_a, tm_b, s_out, g_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // namespace

extern "C" int muscle_debug_tile(const void* a_bf16, const void* b_bf16, int dim, float* s_out,
                                 float* g_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a_bf16 || !b_bf16 || !s_out || !g_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (device_info().cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "needs an sm_100 GPU");
  switch (dim) {
    case 64: return launch_debug<64>(a_bf16, b_bf16, s_out, g_out, stream);
    case 128: return launch_debug<128>(a_bf16, b_bf16, s_out, g_out, stream);
    case 256: return launch_debug<256>(a_bf16, b_bf16, s_out, g_out, stream);
  }
  return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
}
