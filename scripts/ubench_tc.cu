// Micro-benchmarks of the sm_100a building blocks the pair kernel is made of (one CTA per SM):
//   tcgen05.mma issue/throughput for the shapes used (SS N=64, SS N=128, TS N=128),
//   tcgen05.ld bandwidth with 4 / 8 warps, MUFU ex2 and FFMA2 rates.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_tc scripts/ubench_tc.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../muscle_b200/csrc/ptx.cuh"

using namespace mb200;

// mode 0: SS M128 N64 K16 ; 1: SS M128 N128 ; 2: TS M128 N128 ; 3: SS N=256
template <int mode>
__global__ void __launch_bounds__(128, 1) mma_bench(int reps, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tptr;
  __shared__ uint64_t bar_storage;
  const uint32_t bar = smem_u32(&bar_storage);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tptr;
  // operands: zero-filled smem is fine for timing
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  fence_proxy_async_smem();
  __syncthreads();
  if (warp == 1) {
    if (elect_one()) {
      const int N = (mode == 0 || (mode >= 4 && mode <= 6)) ? 64 : (mode == 3 ? 256 : 128);
      const uint32_t idesc_ts64 = umma_idesc_bf16(128, 64, 0, 0);
      const uint32_t idesc_ss = umma_idesc_bf16(128, N, 0, 0);
      const uint32_t idesc_ts = umma_idesc_bf16(128, 128, 0, 1);
      const uint32_t a_smem = base, b_smem = base + 32768;
      long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          if (mode == 2)
            mma_ts(tmem + 256, tmem + kk * 8, umma_desc_sw128(b_smem + kk * 2048, 16384, 1024), idesc_ts, 1);
          else if (mode == 7)   // column sums: A = e tile MN-major (M contiguous) from smem, B = ones K-major, N = 16
            mma_ss(tmem + 384, umma_desc_sw128(a_smem + kk * 2048, 16384, 1024), umma_desc_sw128(b_smem, 16, 1024), umma_idesc_bf16(128, 16, 1, 0), kk > 0);
          else if (mode == 8)   // the same with N = 64
            mma_ss(tmem + 384, umma_desc_sw128(a_smem + kk * 2048, 16384, 1024), umma_desc_sw128(b_smem, 16, 1024), umma_idesc_bf16(128, 64, 1, 0), kk > 0);
          else if (mode == 9)   // column sums the other way round: A = ones K-major, B = e tile MN-major, N = 128
            mma_ss(tmem + 384, umma_desc_sw128(b_smem, 16, 1024), umma_desc_sw128(a_smem + kk * 2048, 16384, 1024), umma_idesc_bf16(128, 128, 0, 1), kk > 0);
          else if (mode == 10)  // A = ones K-major M = 64, B = e tile MN-major, N = 128
            mma_ss(tmem + 384, umma_desc_sw128(b_smem, 16, 1024), umma_desc_sw128(a_smem + kk * 2048, 16384, 1024), umma_idesc_bf16(64, 128, 0, 1), kk > 0);
          else if (mode == 11)  // K-major A, N = 16 (is a narrow N slow by itself?)
            mma_ss(tmem + 384, umma_desc_sw128(a_smem + off, 16, 1024), umma_desc_sw128(b_smem, 16, 1024), umma_idesc_bf16(128, 16, 0, 0), kk > 0);
          else if (mode == 12)  // A from TMEM (ones), B = e tile MN-major N = 128
            mma_ts(tmem + 384, tmem + 8, umma_desc_sw128(a_smem + kk * 2048, 16384, 1024), umma_idesc_bf16(128, 128, 0, 1), kk > 0);
          else if (mode == 4)   // S from a TMEM-resident row tile: A in TMEM, B K-major N=64
            mma_ts(tmem + 256, tmem + kk * 8, umma_desc_sw128(b_smem + off, 16, 1024), idesc_ts64, kk > 0);
          else if (mode == 5)   // SS N=64, consecutive MMAs on alternating accumulators
            mma_ss(tmem + (kk & 1) * 64, umma_desc_sw128(a_smem + off, 16, 1024), umma_desc_sw128(b_smem + off, 16, 1024), idesc_ss, kk > 1);
          else if (mode == 6)   // TS N=64 K-major B followed by TS N=128 MN-major B (the gradient pass mix)
            { mma_ts(tmem + 256, tmem + kk * 8, umma_desc_sw128(b_smem + off, 16, 1024), idesc_ts64, kk > 0);
              if (kk < 4) mma_ts(tmem + 384, tmem + 64 + kk * 8, umma_desc_sw128(b_smem + kk * 2048, 16384, 1024), idesc_ts, 1); }
          else
            mma_ss(tmem, umma_desc_sw128(a_smem + off, 16, 1024), umma_desc_sw128(b_smem + off, 16, 1024), idesc_ss, kk > 0);
        }
      }
      long long t1 = clock64();   // issue time
      tc_commit(bar);
      mbar_wait(bar, 0);
      long long t2 = clock64();   // completion time
      if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// TMEM read bandwidth: nwarps warps each read `reps` x (32 lanes x 32 cols) from their quadrant
__global__ void __launch_bounds__(256, 1) ldtm_bench(int reps, long long* out) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tptr;
  const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
  uint32_t acc = 0;
  __syncthreads();
  long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    uint32_t v[32], u[32];
    tmem_ld32(tmem + lane_addr + ((r * 64) & 255) + (warp >> 2) * 256, v);
    tmem_ld32(tmem + lane_addr + ((r * 64 + 32) & 255) + (warp >> 2) * 256, u);
    tmem_wait_ld();
#pragma unroll
    for (int i = 0; i < 32; ++i) acc ^= v[i] ^ u[i];
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  if (acc == 0x12345u) out[3] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// MUFU ex2 and FFMA2 throughput, 8 warps per SM
__global__ void __launch_bounds__(256, 1) alu_bench(int mode, int reps, float seed, long long* out, float* sink) {
  float x[8];
  for (int i = 0; i < 8; ++i) x[i] = seed + i * 0.001f + threadIdx.x * 1e-6f;
  __syncthreads();
  long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    if (mode == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = ex2_approx(x[i]);
    } else {
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        uint64_t p = fma2(pack2(x[i], x[i + 1]), pack2(0.999f, 0.999f), pack2(1e-3f, 1e-3f));
        unpack2(p, x[i], x[i + 1]);
      }
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  float s = 0; for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 123.456f) sink[0] = s;
}

int main() {
  long long* out; float* sink;
  cudaMallocManaged(&out, 64); cudaMalloc(&sink, 64);
  cudaFuncSetAttribute(mma_bench<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<11>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(mma_bench<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* names[13] = {"SS M128 N64 K16", "SS M128 N128 K16", "TS M128 N128 K16", "SS M128 N256 K16",
                          "TS M128 N64 Kmaj", "SS N64 alt-acc", "TS64x8+TS128x4 /8", "SS A-MNmaj N16", "SS A-MNmaj N64",
                          "SS ones x B-MNmaj N128", "SS M64 ones x B-MN N128", "SS A-Kmaj N16", "TS ones x B-MNmaj N128"};
  for (int mode = 0; mode < 13; ++mode) {
    for (int grid : {1, 148}) {
      const int reps = 512;
      switch (mode) {
        case 0: mma_bench<0><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 1: mma_bench<1><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 2: mma_bench<2><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 3: mma_bench<3><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 4: mma_bench<4><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 5: mma_bench<5><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 6: mma_bench<6><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 7: mma_bench<7><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 8: mma_bench<8><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 9: mma_bench<9><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 10: mma_bench<10><<<grid, 128, 100 * 1024>>>(reps, out); break;
        case 11: mma_bench<11><<<grid, 128, 100 * 1024>>>(reps, out); break;
        default: mma_bench<12><<<grid, 128, 100 * 1024>>>(reps, out); break;
      }
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      printf("%-18s grid %3d: issue %.1f cyc/mma, complete %.1f cyc/mma\n", names[mode], grid,
             double(out[0]) / (reps * 8), double(out[1]) / (reps * 8));
    }
  }
  for (int threads : {128, 256}) {
    const int reps = 256;
    ldtm_bench<<<148, threads>>>(reps, out);
    cudaDeviceSynchronize();
    const double bytes = double(threads) * 64 * 4 * reps;
    printf("tcgen05.ld x32, %d warps: %.1f B/clk/SM (%.0f cycles per 128x128 fp32 tile)\n", threads / 32,
           bytes / out[0], 65536.0 / (bytes / out[0]));
  }
  for (int mode = 0; mode < 2; ++mode) {
    const int reps = 512;
    alu_bench<<<148, 256>>>(mode, reps, 0.5f, out, sink);
    cudaDeviceSynchronize();
    const double ops = 256.0 * 8 * reps;
    printf("%s: %.1f lane-ops/clk/SM\n", mode == 0 ? "MUFU ex2 " : "FFMA2 (x2 elems)", ops / out[0]);
  }
  return 0;
}
