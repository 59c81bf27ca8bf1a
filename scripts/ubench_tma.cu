// Micro-benchmark of the column-tile ring of the pair kernel in isolation: one CTA per SM, a
// producer thread streams [128 x 128] bf16 tiles (two SW128 panels, 32 KB) through a 4-stage ring,
// an observer thread timestamps every completion (issue -> mbarrier phase flip), a consumer thread
// holds each stage for `delay` cycles before handing it back, and optionally a fourth thread keeps
// the tensor pipe busy with SS or TS MMAs on other shared memory (the operand-fetch traffic the
// real kernel has).  Answers: what is the TMA latency / per-SM bandwidth from L2, and does
// concurrent tcgen05 operand traffic change it?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_tma scripts/ubench_tma.cu -lcuda
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#include "../include/muscle_b200.h"
#include "../muscle_b200/csrc/host_util.cuh"
#include "../muscle_b200/csrc/ptx.cuh"

using namespace mb200;

constexpr int NST = 4;
constexpr int TILE = 32768;

// hammer: 0 none, 1 SS M128 N64, 2 TS M128 N64 (A in TMEM)
// OPS: 0 = two 2-D panel copies, 1 = + two 512-B bulk copies, 2 = one 3-D copy, 3 = one 3-D copy + one 1-KB bulk copy
template <int HAMMER, int OPS>
__global__ void __launch_bounds__(128, 1) ring_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tm3, const char* meta, int n_tiles_total,
                                                       int tiles, int delay, int stride, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bars[2 * NST + 1];
  __shared__ uint32_t tptr;
  __shared__ volatile uint32_t issue_clk[NST];
  __shared__ volatile int stop;
  auto k_full = [&](int s) { return smem_u32(&bars[s]); };
  auto k_empty = [&](int s) { return smem_u32(&bars[NST + s]); };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); }
    mbar_init(smem_u32(&bars[2 * NST]), 1);
    stop = 0;
    fence_barrier_init();
  }
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 512); tmem_relinquish(); }
  // zero the hammer operands (64 KB after the ring)
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)) + NST * TILE)[i] = 0;
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tptr;
  const uint32_t k_smem = base;
  const uint32_t ham = base + NST * TILE;

  if (warp == 0 && lane == 0) {          // producer
    const long long t0 = clock64();
    long long wait_empty = 0;
    for (int t = 0; t < tiles; ++t) {
      const int st = t % NST;
      const long long w0 = clock64();
      mbar_wait(k_empty(st), ((t / NST) & 1) ^ 1);
      wait_empty += clock64() - w0;
      const int j = static_cast<int>((static_cast<long long>(blockIdx.x) * stride + t) % n_tiles_total);
      issue_clk[st] = static_cast<uint32_t>(clock64());
      mbar_arrive_expect_tx(k_full(st), TILE + (OPS == 1 || OPS == 3 ? 1024 : 0));
      if (OPS < 2) {
        tma_load_2d(k_smem + st * TILE, &tm, k_full(st), 0, j * 128);
        tma_load_2d(k_smem + st * TILE + 16384, &tm, k_full(st), 64, j * 128);
      } else {
        tma_load_3d(k_smem + st * TILE, &tm3, k_full(st), 0, j * 128, 0);
      }
      if (OPS == 1) {
        bulk_load_1d(ham + 65536 + st * 1024, meta + j * 1024, 512, k_full(st));
        bulk_load_1d(ham + 65536 + st * 1024 + 512, meta + j * 1024 + 512, 512, k_full(st));
      }
      if (OPS == 3) bulk_load_1d(ham + 65536 + st * 1024, meta + j * 1024, 1024, k_full(st));
    }
    out[blockIdx.x * 8 + 0] = clock64() - t0;
    out[blockIdx.x * 8 + 1] = wait_empty;
  } else if (warp == 1 && lane == 0) {   // observer: completion latency
    long long lat = 0, lat_max = 0;
    for (int t = 0; t < tiles; ++t) {
      const int st = t % NST;
      mbar_wait(k_full(st), (t / NST) & 1);
      const uint32_t d = static_cast<uint32_t>(clock64()) - issue_clk[st];
      lat += d;
      if (static_cast<long long>(d) > lat_max) lat_max = d;
    }
    out[blockIdx.x * 8 + 2] = lat;
    out[blockIdx.x * 8 + 3] = lat_max;
  } else if (warp == 2 && lane == 0) {   // consumer
    const long long t0 = clock64();
    long long wait_full = 0;
    for (int t = 0; t < tiles; ++t) {
      const int st = t % NST;
      const long long w0 = clock64();
      mbar_wait(k_full(st), (t / NST) & 1);
      const long long w1 = clock64();
      wait_full += w1 - w0;
      while (clock64() - w1 < delay) {}
      mbar_arrive(k_empty(st));
    }
    out[blockIdx.x * 8 + 4] = clock64() - t0;
    out[blockIdx.x * 8 + 5] = wait_full;
    stop = 1;
  } else if (warp == 3 && HAMMER != 0) { // tensor-pipe load
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
      const uint32_t bar = smem_u32(&bars[2 * NST]);
      long long n = 0;
      uint32_t ph = 0;
      while (!stop) {
        for (int r = 0; r < 4; ++r) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
            if (HAMMER == 1)
              mma_ss(tmem, umma_desc_sw128(ham + off, 16, 1024), umma_desc_sw128(ham + 32768 + off, 16, 1024), idesc, kk > 0);
            else
              mma_ts(tmem, tmem + 256 + kk * 8, umma_desc_sw128(ham + 32768 + off, 16, 1024), idesc, kk > 0);
          }
        }
        n += 32;
        tc_commit(bar);
        mbar_wait(bar, ph);
        ph ^= 1;
      }
      out[blockIdx.x * 8 + 6] = n;
    }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int HAMMER, int OPS>
void run(const CUtensorMap& tm, const CUtensorMap& tm3, const char* meta, int n_tiles_total, long long* out, int smem) {
  cudaFuncSetAttribute(ring_kernel<HAMMER, OPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int tiles = 2000;
  for (int grid : {148, 1})
    for (int delay : {0, 600, 2000}) {
      const int stride = 97;
      cudaMemset(out, 0, 148 * 8 * 8);
      for (int rep = 0; rep < 2; ++rep)
        ring_kernel<HAMMER, OPS><<<grid, 128, smem>>>(tm, tm3, meta, n_tiles_total, tiles, delay, stride, out);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); exit(1); }
      std::vector<long long> h(148 * 8);
      cudaMemcpy(h.data(), out, 148 * 8 * 8, cudaMemcpyDeviceToHost);
      double cyc = 0, lat = 0, lmax = 0, nm = 0, busy = 0;
      for (int b = 0; b < grid; ++b) {
        cyc += (double)h[b * 8 + 4] / tiles;
        lat += (double)h[b * 8 + 2] / tiles;
        lmax = std::max(lmax, (double)h[b * 8 + 3]);
        nm += (double)h[b * 8 + 6] / tiles;
        busy += (double)(h[b * 8 + 0] - h[b * 8 + 1]) / tiles;
      }
      cyc /= grid; lat /= grid; nm /= grid; busy /= grid;
      printf("%-7d %-4d %-5d %-6d | %-9.0f %-9.0f %-9.0f %-9.1f %-9.1f %-9.0f\n", HAMMER, OPS, grid, delay, cyc, lat, lmax,
             TILE / cyc, nm, busy);
    }
}

int main(int argc, char** argv) {
  const int64_t rows = 65536;   // 16 MB of bf16 [rows][128]: L2 resident
  void* buf;
  cudaMalloc(&buf, rows * 128 * 2);
  cudaMemset(buf, 0, rows * 128 * 2);
  char* meta;
  cudaMalloc(&meta, rows / 128 * 1024);
  cudaMemset(meta, 0, rows / 128 * 1024);
  CUtensorMap tm, tm3;
  if (make_bf16_box_map(&tm, buf, rows, 128, 128)) { printf("map failed\n"); return 1; }
  if (make_bf16_panel_map(&tm3, buf, rows, 128)) { printf("panel map failed\n"); return 1; }
  long long* out;
  cudaMalloc(&out, 148 * 8 * 8);
  const int smem = NST * TILE + 65536 + NST * 1024 + 1024;
  printf("%-7s %-4s %-5s %-6s | %-9s %-9s %-9s %-9s %-9s %-9s\n", "hammer", "ops", "grid", "delay", "cyc/tile", "lat_avg",
         "lat_max", "B/clk/SM", "mma/tile", "prod_busy");
  const int nt = (int)(rows / 128);
  run<0, 0>(tm, tm3, meta, nt, out, smem);
  run<0, 1>(tm, tm3, meta, nt, out, smem);
  run<0, 2>(tm, tm3, meta, nt, out, smem);
  run<0, 3>(tm, tm3, meta, nt, out, smem);
  run<1, 0>(tm, tm3, meta, nt, out, smem);
  run<1, 1>(tm, tm3, meta, nt, out, smem);
  run<1, 3>(tm, tm3, meta, nt, out, smem);
  return 0;
}
