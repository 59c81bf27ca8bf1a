// CAM-weighted prototype pooling on the tensor pipe (include/muscle_b200.h: muscle_pool_prototypes_fwd/bwd with a bf16
// embedding matrix of 64 / 128 / 256 columns).
//
// The CUDA-core kernels of pool.cu need N*C*D FMAs per direction and run at the FMA/LSU issue rate (5 % / 13 % of HBM
// at 65 536 pixels).  The pooling is a skinny matrix product, so here the tensor pipe does the multiplying and the
// kernel only has to keep the bytes coming:
//
//   forward :  sum[c][d] = sum_p w[p][c] e[p][d]        D_acc[M = class, N = d] += A[M x K = pixel] B[K x N]
//              * B = the embedding rows as TMA delivers them ([64 pixels][64 columns] panels, 128-B swizzle), read
//                MN-major (the 64 contiguous elements are the N index), exactly like the gradient pass reads its keys
//              * A = the weights of the same 64 pixels, converted by four warps from fp32 to bf16 HI and LO parts
//                (w = hi + lo to 2^-17 relative) and written into shared memory in the swizzled layout an MN-major
//                operand expects: row = pixel, 64 contiguous M elements = [32 hi classes | 32 lo classes].  M = 128 for
//                the instruction; rows 64..127 come from a panel of zeros shared by all stages.
//              * the accumulator [64 lanes x D columns] stays in TMEM for the whole kernel; at the end lanes 0..31 (hi)
//                and 32..63 (lo) are added to sum[] with 16-byte vector reductions.  mass[c] = sum_p w[p][c] is summed
//                in fp32 by the converting threads.
//
// One persistent CTA per SM, chunks of 64 pixels taken round robin, a ring of NST stages (embedding tile + weight
// tile), warp roles: 0-3 weight conversion / epilogue (two groups of 64 threads alternate chunks, each with its next
// chunk's weights already in registers), 4 = TMA producer, 5 = MMA issuer + TMEM allocator.
#pragma once
#include <cuda_bf16.h>

#include "ptx.cuh"

namespace mb200 {

constexpr int PTC_K = 64;            // pixels per chunk (the K extent of one step: 4 MMAs of K = 16)
constexpr int PTC_MAXC = 24;         // classes (padded to 32 hi + 32 lo rows of M)
constexpr int PTC_THREADS = 192;

template <int D>
struct PoolTcCfg {
  static_assert(D == 64 || D == 128 || D == 256, "tensor-pipe pooling: 64, 128 or 256 columns");
  static constexpr int NST = D == 256 ? 4 : 6;
  static constexpr int E_PANEL = PTC_K * 128;              // [64 pixels][64 columns] bf16
  static constexpr int E_STAGE = E_PANEL * (D / 64);
  static constexpr int A_STAGE = PTC_K * 128;              // [64 pixels][32 hi | 32 lo] bf16
  static constexpr int W_STAGE = PTC_K * PTC_MAXC * 4;      // the chunk's fp32 weights as they lie in HBM (one bulk copy)
  static constexpr int OFF_A = NST * E_STAGE;
  static constexpr int OFF_ZERO = OFF_A + NST * A_STAGE;   // panel of zeros: rows 64..127 of M
  static constexpr int OFF_W32 = OFF_ZERO + A_STAGE;
  static constexpr int OFF_BAR = OFF_W32 + NST * W_STAGE;
  static constexpr int OFF_MASS = OFF_BAR + 256;           // [4 warps][32] partial masses
  static constexpr int SMEM_BYTES = OFF_MASS + 512 + 1024; // + alignment slack
  static constexpr int X_STRIDE = D + 4;                   // floats per row of the hi/lo exchange (reuses stage 0)
  static_assert(32 * X_STRIDE * 4 <= NST * E_STAGE, "exchange buffer fits the embedding ring");
  static constexpr int TMEM_COLS = D < 32 ? 32 : D;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// hi/lo bf16 parts of eight consecutive classes -> one 16-byte piece each
__device__ __forceinline__ void ptc_split8(const float (&w)[32], int c0, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float a = w[c0 + 2 * i], b = w[c0 + 2 * i + 1];
    const __nv_bfloat16 ah = __float2bfloat16_rn(a), bh = __float2bfloat16_rn(b);
    const __nv_bfloat16 al = __float2bfloat16_rn(a - __bfloat162float(ah));
    const __nv_bfloat16 bl = __float2bfloat16_rn(b - __bfloat162float(bh));
    h[i] = static_cast<uint32_t>(__bfloat16_as_ushort(ah)) | (static_cast<uint32_t>(__bfloat16_as_ushort(bh)) << 16);
    l[i] = static_cast<uint32_t>(__bfloat16_as_ushort(al)) | (static_cast<uint32_t>(__bfloat16_as_ushort(bl)) << 16);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

template <int D>
__global__ void __launch_bounds__(PTC_THREADS, 1)
pool_tc_fwd_kernel(const __grid_constant__ CUtensorMap tm_e, const float* __restrict__ w, long long n, int C, int w_bulk,
                   float* __restrict__ sum, float* __restrict__ mass, float* __restrict__ partial) {
  // w_bulk != 0 (w is 16-byte aligned): the producer brings every full chunk's fp32 weights -- 64 * C * 4 contiguous
  //   bytes of w -- into a shared-memory ring with one bulk copy on the embedding tile's barrier, NST chunks ahead;
  //   the converting threads read their pixel's row from there.  Loading them with ld.global one chunk ahead left the
  //   kernel waiting on HBM latency (1 800 cycles per chunk against ~950 at the HBM rate, profiles/r2v_*).  The last,
  //   partial chunk and unaligned w are loaded directly.
  // partial == nullptr: this CTA's sums are ADDED to sum[] / mass[] (zeroed by the caller) with reductions.
  // partial != nullptr: they are STORED to partial[blockIdx.x][C*D] and partial[gridDim.x*C*D + blockIdx.x*32 + c]
  //   (no atomics, no zeroing; pool_tc_finish_kernel adds the CTAs' parts in a fixed order: deterministic, and it does
  //   not queue 2 x 148 vector reductions on each of the 84 lines of sum[], which cost ~25 us at any size)
  using Cfg = PoolTcCfg<D>;
  constexpr int NST = Cfg::NST;
  extern __shared__ uint8_t ptc_smem_raw[];
  const uint32_t base = (smem_u32(ptc_smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = ptc_smem_raw + (base - smem_u32(ptc_smem_raw));
  const uint32_t e_smem = base, a_smem = base + Cfg::OFF_A, zero_smem = base + Cfg::OFF_ZERO;
  const uint32_t bar0 = base + Cfg::OFF_BAR;
  auto e_full = [&](int st) { return bar0 + 8u * st; };
  auto a_full = [&](int st) { return bar0 + 8u * (NST + st); };
  auto empty = [&](int st) { return bar0 + 8u * (2 * NST + st); };
  const uint32_t done_bar = bar0 + 8u * 3 * NST;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(gen + Cfg::OFF_BAR + 8 * (3 * NST + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int st = 0; st < NST; ++st) {
      mbar_init(e_full(st), 1);
      mbar_init(a_full(st), 2);       // one arrival per converting warp of the group
      mbar_init(empty(st), 1);        // the issuer's commit
    }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  // the panel of zeros (generic-proxy stores, made visible to the tensor pipe by the fence below)
  for (int i = threadIdx.x; i < Cfg::A_STAGE / 16; i += PTC_THREADS)
    reinterpret_cast<uint4*>(gen + Cfg::OFF_ZERO)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  if (warp == 5) {
    tmem_alloc(smem_u32(tmem_ptr_smem), Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  if (warp == 4 && lane == 0) tma_prefetch_desc(&tm_e);
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;
  pdl_wait();

  const long long n_chunks = (n + PTC_K - 1) / PTC_K;
  const long long first = blockIdx.x, stride = gridDim.x;
  const long long mine = first < n_chunks ? (n_chunks - first + stride - 1) / stride : 0;   // chunks of this CTA

  if (warp == 4) {
    // ------------------------------------------------------------ TMA producer: embedding tiles
    if (lane == 0) {
      for (long long k = 0; k < mine; ++k) {
        const int st = static_cast<int>(k % NST);
        mbar_wait(empty(st), static_cast<uint32_t>(((k / NST) & 1) ^ 1));
        const long long p0 = (first + k * stride) * PTC_K;
        const uint32_t wbytes = (w_bulk && p0 + PTC_K <= n) ? static_cast<uint32_t>(PTC_K * C * 4) : 0u;
        mbar_arrive_expect_tx(e_full(st), Cfg::E_STAGE + wbytes);
        tma_load_3d(e_smem + st * Cfg::E_STAGE, &tm_e, e_full(st), 0, static_cast<int>(p0), 0);
        if (wbytes) bulk_load_1d(base + Cfg::OFF_W32 + st * Cfg::W_STAGE, w + p0 * C, wbytes, e_full(st));
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      constexpr uint32_t IDESC = umma_idesc_bf16(128, D, 1, 1);     // A and B both MN-major
      for (long long k = 0; k < mine; ++k) {
        const int st = static_cast<int>(k % NST);
        const uint32_t ph = static_cast<uint32_t>((k / NST) & 1);
        mbar_wait(a_full(st), ph);
        mbar_wait(e_full(st), ph);
        tc_fence_after();
        const uint32_t a_base = a_smem + st * Cfg::A_STAGE;
        const uint32_t b_base = e_smem + st * Cfg::E_STAGE;
#pragma unroll
        for (int kk = 0; kk < PTC_K / 16; ++kk)
          mma_ss(tmem, umma_desc_sw128(a_base + kk * 2048, zero_smem - a_base, 1024),
                 umma_desc_sw128(b_base + kk * 2048, Cfg::E_PANEL, 1024), IDESC, (k > 0 || kk > 0) ? 1u : 0u);
        tc_commit(empty(st));
      }
      tc_commit(done_bar);
    }
  } else {
    // ------------------------------------------------------------ weight conversion (two groups of two warps)
    const int grp = warp >> 1;
    const int t = threadIdx.x & 63;                 // pixel of the chunk
    float macc[PTC_MAXC];
#pragma unroll
    for (int c = 0; c < PTC_MAXC; ++c) macc[c] = 0.f;
    for (long long k = grp; k < mine; k += 2) {
      const int st = static_cast<int>(k % NST);
      const long long p0 = (first + k * stride) * PTC_K;
      float wc[32];
      if (w_bulk && p0 + PTC_K <= n) {
        mbar_wait_warp(e_full(st), static_cast<uint32_t>((k / NST) & 1));
        const float* ws = reinterpret_cast<const float*>(gen + Cfg::OFF_W32 + st * Cfg::W_STAGE) + t * C;
#pragma unroll
        for (int c = 0; c < PTC_MAXC; ++c) wc[c] = c < C ? ws[c] : 0.f;
      } else {
        const long long p = p0 + t;
#pragma unroll
        for (int c = 0; c < PTC_MAXC; ++c) wc[c] = (p < n && c < C) ? __ldg(w + p * C + c) : 0.f;
      }
#pragma unroll
      for (int c = PTC_MAXC; c < 32; ++c) wc[c] = 0.f;
#pragma unroll
      for (int c = 0; c < PTC_MAXC; ++c) macc[c] += wc[c];
      mbar_wait_warp(empty(st), static_cast<uint32_t>(((k / NST) & 1) ^ 1));
      const uint32_t row = a_smem + st * Cfg::A_STAGE + t * 128;
      const uint32_t swz = static_cast<uint32_t>(t & 7);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 hi, lo;
        ptc_split8(wc, 8 * j, hi, lo);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((static_cast<uint32_t>(j) ^ swz) << 4)),
                     "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((static_cast<uint32_t>(j + 4) ^ swz) << 4)),
                     "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w) : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_full(st));
    }
    // mass: fp32 sums of the weights this thread converted -> per warp -> per CTA
    float* mass_sm = reinterpret_cast<float*>(gen + Cfg::OFF_MASS);
#pragma unroll
    for (int c = 0; c < PTC_MAXC; ++c) {
      float v = macc[c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) mass_sm[warp * 32 + c] = v;
    }
    asm volatile("bar.sync 2, 128;" ::: "memory");
    if (warp == 0 && lane < C && mine > 0) {
      const float m = (mass_sm[lane] + mass_sm[32 + lane]) + (mass_sm[64 + lane] + mass_sm[96 + lane]);
      if (partial != nullptr) partial[static_cast<size_t>(gridDim.x) * C * D + blockIdx.x * 32 + lane] = m;
      else atomicAdd(mass + lane, m);
    }
    // ---------------------------------------------------------- epilogue: TMEM lanes 0..31 = hi, 32..63 = lo
    if (warp < 2 && mine > 0) {
      mbar_wait_warp(done_bar, 0);
      tc_fence_after();
      const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
      if (partial == nullptr) {
#pragma unroll 1
        for (int c0 = 0; c0 < D; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_addr + c0, v);
          tmem_wait_ld();
          if (lane < C) {
            float* dst = sum + static_cast<size_t>(lane) * D + c0;
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              red_add_v4(dst + i, __uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]),
                         __uint_as_float(v[i + 3]));
          }
        }
      } else {
        // lo part through shared memory (the ring is idle now), hi + lo stored as this CTA's part
        float* x = reinterpret_cast<float*>(gen);             // [32][X_STRIDE]
        if (warp == 1) {
#pragma unroll 1
          for (int c0 = 0; c0 < D; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tmem + lane_addr + c0, v);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<uint4*>(x + lane * Cfg::X_STRIDE + c0 + i) = make_uint4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
        }
        asm volatile("bar.sync 1, 64;" ::: "memory");
        if (warp == 0) {
          float* dst_row = partial + static_cast<size_t>(blockIdx.x) * C * D + static_cast<size_t>(lane) * D;
#pragma unroll 1
          for (int c0 = 0; c0 < D; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tmem + lane_addr + c0, v);
            tmem_wait_ld();
            if (lane < C) {
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                const float4 l = *reinterpret_cast<const float4*>(x + lane * Cfg::X_STRIDE + c0 + i);
                *reinterpret_cast<float4*>(dst_row + c0 + i) =
                    make_float4(__uint_as_float(v[i]) + l.x, __uint_as_float(v[i + 1]) + l.y,
                                __uint_as_float(v[i + 2]) + l.z, __uint_as_float(v[i + 3]) + l.w);
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem, Cfg::TMEM_COLS);
}

// sum[i] = sum over the CTAs' parts (fixed order), mass likewise, proto = sum / max(mass, eps).  Block = 32 elements x 32
// slices of the CTA list: a warp reads 128 contiguous bytes of one part, a thread at most PTC_FIN_LOADS parts, all loads
// issued before the first add (the first version's 8 slices walked 19 dependent-latency loads each: 9 us for 1.6 MB).
constexpr int PTC_FIN_THREADS = 1024;
constexpr int PTC_FIN_LOADS = 6;           // 32 slices x 6 = 192 parts (PTC_MAX_PARTS in pool.cu)
__global__ void __launch_bounds__(PTC_FIN_THREADS) pool_tc_finish_kernel(const float* __restrict__ partial, int parts, int C,
                                                                         int dim, float eps, float* __restrict__ sum,
                                                                         float* __restrict__ mass, float* __restrict__ proto) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[32][33];
  __shared__ float mred[32];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const int total = C * dim;
  const int i = blockIdx.x * 32 + lane;
  const int c = (blockIdx.x * 32) / dim;              // dim % 32 == 0: one class per block
  float v[PTC_FIN_LOADS];
#pragma unroll
  for (int j = 0; j < PTC_FIN_LOADS; ++j) {
    const int p = slice + 32 * j;
    v[j] = (i < total && p < parts) ? __ldg(partial + static_cast<size_t>(p) * total + i) : 0.f;
  }
  // the class's mass: thread t < parts holds part t's mass (parts <= 192 < 1024)
  const int t = threadIdx.x;
  float m = (t < parts) ? __ldg(partial + static_cast<size_t>(parts) * total + t * 32 + c) : 0.f;
  float acc = 0.f;
#pragma unroll
  for (int j = 0; j < PTC_FIN_LOADS; ++j) acc += v[j];
  red[slice][lane] = acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
  if (lane == 0) mred[slice] = m;
  __syncthreads();
  if (slice == 0 && i < total) {
    float s = 0.f, mm = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) s += red[k][lane];
#pragma unroll
    for (int k = 0; k < 8; ++k) mm += mred[k];          // warps 0..5 hold parts; the rest added zeros
    sum[i] = s;
    if (i % dim == 0) mass[c] = mm;
    if (proto != nullptr) proto[i] = s / fmaxf(mm, eps);
  }
}

// ==========================================================================================================
// Backward on the tensor pipe (D = 64 / 128).  With g[c][:] = dproto[c][:] / max(mass[c], eps) split into bf16 hi + lo:
//
//   de[p][:] = sum_c w[p][c] g[c][:]      product 1: A = weight tile [128 pixels][wh(32) | wl(32)] K-major,
//                                          B = [gh ; gh ; gl] (96 rows) MN-major: (wh + wl) gh + wh gl, fp32 in TMEM
//   dw[p][c] = e[p] . g[c] + dm[c]        product 2: A = the embedding tile K-major (as TMA delivers it),
//                                          B = [gh ; gl] (64 rows) K-major: columns c and 32 + c are added in the epilogue
//
// Chunks of 128 pixels; both accumulators double-buffered in TMEM so the epilogue of chunk i (TMEM -> padded shared
// rows -> one 16-byte-aligned bulk copy per de row, a cooperative coalesced copy for dw) overlaps the loads and
// products of chunk i + 1.  Warps 0-3 epilogue, 4-7 weight conversion (thread = pixel), 8 TMA producer, 9 issuer.
constexpr int PTB_M = 128;
constexpr int PTB_THREADS = 320;

template <int D>
struct PoolTcBwdCfg {
  static_assert(D == 64 || D == 128, "tensor-pipe pooling backward: 64 or 128 columns");
  static constexpr int NST = 2;
  static constexpr int E_PANEL = PTB_M * 128;
  static constexpr int E_STAGE = E_PANEL * (D / 64);
  static constexpr int W_STAGE = PTB_M * 128;
  static constexpr int G1_PANEL = 96 * 128, G1_BYTES = G1_PANEL * (D / 64);
  static constexpr int G2_PANEL = 64 * 128, G2_BYTES = G2_PANEL * (D / 64);
  static constexpr int DE_STRIDE = (D + 4) * 4;                // bytes per staged de row (conflict-free 16-byte stores)
  static constexpr int DE_BYTES = PTB_M * DE_STRIDE;
  static constexpr int DW_BYTES = PTB_M * PTC_MAXC * 4;
  static constexpr int OFF_W = NST * E_STAGE;
  static constexpr int OFF_G1 = OFF_W + NST * W_STAGE;
  static constexpr int OFF_G2 = OFF_G1 + G1_BYTES;
  static constexpr int OFF_DE = OFF_G2 + G2_BYTES;
  static constexpr int OFF_DW = OFF_DE + DE_BYTES;
  static constexpr int OFF_BAR = OFF_DW + DW_BYTES;
  static constexpr int OFF_DM = OFF_BAR + 128;
  static constexpr int SMEM_BYTES = OFF_DM + 128 + 1024;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
  static constexpr int DE_COL = 0, DW_COL = 2 * D, TMEM_COLS = 512;
  static_assert(2 * D + 128 <= 512, "TMEM budget");
};

__device__ __forceinline__ void bulk_store_1d(void* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

template <int D>
__global__ void __launch_bounds__(PTB_THREADS, 1)
pool_tc_bwd_kernel(const __grid_constant__ CUtensorMap tm_e, const float* __restrict__ w, long long n, int C,
                   const float* __restrict__ proto, const float* __restrict__ mass, const float* __restrict__ dproto,
                   float eps, float* __restrict__ de, float* __restrict__ dw) {
  using Cfg = PoolTcBwdCfg<D>;
  constexpr int NST = Cfg::NST;
  extern __shared__ uint8_t ptc_smem_raw[];
  const uint32_t base = (smem_u32(ptc_smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = ptc_smem_raw + (base - smem_u32(ptc_smem_raw));
  const uint32_t e_smem = base, w_smem = base + Cfg::OFF_W, g1_smem = base + Cfg::OFF_G1, g2_smem = base + Cfg::OFF_G2;
  const uint32_t de_smem = base + Cfg::OFF_DE;
  float* dw_sm = reinterpret_cast<float*>(gen + Cfg::OFF_DW);
  float* dm_sm = reinterpret_cast<float*>(gen + Cfg::OFF_DM);
  const uint32_t bar0 = base + Cfg::OFF_BAR;
  auto e_full = [&](int st) { return bar0 + 8u * st; };
  auto w_full = [&](int st) { return bar0 + 8u * (2 + st); };
  auto in_empty = [&](int st) { return bar0 + 8u * (4 + st); };
  auto acc_full = [&](int b) { return bar0 + 8u * (6 + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (8 + b); };
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(gen + Cfg::OFF_BAR + 8 * 10);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool need_de = de != nullptr, need_dw = dw != nullptr;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(e_full(i), 1);
      mbar_init(w_full(i), 4);
      mbar_init(in_empty(i), 1);
      mbar_init(acc_full(i), 1);
      mbar_init(acc_empty(i), 4);
    }
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(smem_u32(tmem_ptr_smem), Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  if (warp == 8 && lane == 0) tma_prefetch_desc(&tm_e);
  pdl_trigger();
  __syncthreads();                       // barriers initialised
  pdl_wait();
  const long long n_chunks = (n + PTB_M - 1) / PTB_M;
  const long long first = blockIdx.x, stride = gridDim.x;
  const long long mine = first < n_chunks ? (n_chunks - first + stride - 1) / stride : 0;
  // the first two embedding tiles are on their way while the operands below are built
  if (warp == 8 && lane == 0 && need_dw)
    for (long long k = 0; k < mine && k < NST; ++k) {
      mbar_arrive_expect_tx(e_full(static_cast<int>(k)), Cfg::E_STAGE);
      tma_load_3d(e_smem + static_cast<int>(k) * Cfg::E_STAGE, &tm_e, e_full(static_cast<int>(k)), 0,
                  static_cast<int>((first + k * stride) * PTB_M), 0);
    }
  // ---- the two B operands (constant for the launch) and dm[c] = -(g[c] . proto[c])
  for (int i = threadIdx.x; i < (Cfg::G1_BYTES + Cfg::G2_BYTES) / 16; i += PTB_THREADS)
    reinterpret_cast<uint4*>(gen + Cfg::OFF_G1)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  // (every load of a round is issued before the first use: a thread walks up to PTB_G_ROUND elements per round, and the
  //  one-element-per-iteration form of this loop cost ~7 us of dependent HBM latency per launch, profiles/r2v_*)
  constexpr int PTB_G_ROUND = 9;
  for (int i0 = 0; i0 < C * D; i0 += PTB_G_ROUND * PTB_THREADS) {
    float gv[PTB_G_ROUND], mv[PTB_G_ROUND];
#pragma unroll
    for (int j = 0; j < PTB_G_ROUND; ++j) {
      const int i = i0 + j * PTB_THREADS + threadIdx.x;
      const bool in = i < C * D;
      gv[j] = in ? __ldg(dproto + i) : 0.f;
      mv[j] = in ? __ldg(mass + i / D) : 1.f;
    }
#pragma unroll
    for (int j = 0; j < PTB_G_ROUND; ++j) {
      const int i = i0 + j * PTB_THREADS + threadIdx.x;
      if (i < C * D) {
        const int c = i / D, d = i - c * D;
        const float g = gv[j] / fmaxf(mv[j], eps);
        const __nv_bfloat16 gh = __float2bfloat16_rn(g);
        const __nv_bfloat16 gl = __float2bfloat16_rn(g - __bfloat162float(gh));
        const int panel = d >> 6, dd = d & 63;
        auto at1 = [&](int r) { return Cfg::OFF_G1 + panel * Cfg::G1_PANEL + r * 128 + ((((dd >> 3) ^ (r & 7))) << 4) + (dd & 7) * 2; };
        auto at2 = [&](int r) { return Cfg::OFF_G2 + panel * Cfg::G2_PANEL + r * 128 + ((((dd >> 3) ^ (r & 7))) << 4) + (dd & 7) * 2; };
        *reinterpret_cast<__nv_bfloat16*>(gen + at1(c)) = gh;
        *reinterpret_cast<__nv_bfloat16*>(gen + at1(32 + c)) = gh;
        *reinterpret_cast<__nv_bfloat16*>(gen + at1(64 + c)) = gl;
        *reinterpret_cast<__nv_bfloat16*>(gen + at2(c)) = gh;
        *reinterpret_cast<__nv_bfloat16*>(gen + at2(32 + c)) = gl;
      }
    }
  }
  static_assert(D % 32 == 0 && D / 32 <= 4, "dm loop below");
  for (int c0 = 0; c0 < C; c0 += PTB_THREADS / 32) {      // warp = class (C <= 24: at most three rounds, usually three)
    const int c = c0 + warp;
    float dp[D / 32], pr[D / 32], ms = 1.f;
    if (c < C) {
      ms = __ldg(mass + c);
#pragma unroll
      for (int j = 0; j < D / 32; ++j) {
        dp[j] = __ldg(dproto + c * D + lane + 32 * j);
        pr[j] = __ldg(proto + c * D + lane + 32 * j);
      }
      const float im = 1.f / fmaxf(ms, eps);
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < D / 32; ++j) s = fmaf(dp[j] * im, pr[j], s);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) dm_sm[c] = (ms >= eps) ? -s : 0.f;
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer (embedding tiles: product 2 only)
    if (lane == 0 && need_dw) {
      for (long long k = NST; k < mine; ++k) {
        const int st = static_cast<int>(k & 1);
        mbar_wait(in_empty(st), static_cast<uint32_t>(((k >> 1) & 1) ^ 1));
        mbar_arrive_expect_tx(e_full(st), Cfg::E_STAGE);
        tma_load_3d(e_smem + st * Cfg::E_STAGE, &tm_e, e_full(st), 0, static_cast<int>((first + k * stride) * PTB_M), 0);
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      constexpr uint32_t IDESC1 = umma_idesc_bf16(128, D, 0, 1);
      constexpr uint32_t IDESC2 = umma_idesc_bf16(128, 64, 0, 0);
      for (long long k = 0; k < mine; ++k) {
        const int st = static_cast<int>(k & 1), b = st;
        const uint32_t ph = static_cast<uint32_t>((k >> 1) & 1);
        mbar_wait(acc_empty(b), ph ^ 1);
        if (need_de) mbar_wait(w_full(st), ph);
        if (need_dw) mbar_wait(e_full(st), ph);
        tc_fence_after();
        if (need_de) {
          const uint32_t a_base = w_smem + st * Cfg::W_STAGE;
#pragma unroll
          for (int kk = 0; kk < 6; ++kk)
            mma_ss(tmem + Cfg::DE_COL + b * D, umma_desc_sw128(a_base + (kk & 3) * 32, 16, 1024),
                   umma_desc_sw128(g1_smem + kk * 2048, Cfg::G1_PANEL, 1024), IDESC1, kk > 0);
        }
        if (need_dw) {
          const uint32_t a_base = e_smem + st * Cfg::E_STAGE;
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk)
            mma_ss(tmem + Cfg::DW_COL + b * 64, umma_desc_sw128(a_base + (kk >> 2) * Cfg::E_PANEL + (kk & 3) * 32, 16, 1024),
                   umma_desc_sw128(g2_smem + (kk >> 2) * Cfg::G2_PANEL + (kk & 3) * 32, 16, 1024), IDESC2, kk > 0);
        }
        tc_commit(in_empty(st));
        tc_commit(acc_full(b));
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ weight conversion: thread = pixel of the chunk
    if (need_de) {
      const int t = threadIdx.x - 128;
      float wn[PTC_MAXC];
      auto load_w = [&](long long k) {
        const long long p = (first + k * stride) * PTB_M + t;
#pragma unroll
        for (int c = 0; c < PTC_MAXC; ++c) wn[c] = (p < n && c < C) ? __ldg(w + p * C + c) : 0.f;
      };
      if (mine > 0) load_w(0);
      for (long long k = 0; k < mine; ++k) {
        float wc[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) wc[c] = c < PTC_MAXC ? wn[c < PTC_MAXC ? c : 0] : 0.f;
        if (k + 1 < mine) load_w(k + 1);
        const int st = static_cast<int>(k & 1);
        mbar_wait_warp(in_empty(st), static_cast<uint32_t>(((k >> 1) & 1) ^ 1));
        const uint32_t row = w_smem + st * Cfg::W_STAGE + t * 128;
        const uint32_t swz = static_cast<uint32_t>(t & 7);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 hi, lo;
          ptc_split8(wc, 8 * j, hi, lo);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((static_cast<uint32_t>(j) ^ swz) << 4)),
                       "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w) : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((static_cast<uint32_t>(j + 4) ^ swz) << 4)),
                       "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w) : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(w_full(st));
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue: thread = pixel row of the chunk
    const int r = threadIdx.x;                                  // 0..127 = TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t my_row = de_smem + r * Cfg::DE_STRIDE;
    for (long long k = 0; k < mine; ++k) {
      const int b = static_cast<int>(k & 1);
      const long long p0 = (first + k * stride) * PTB_M;
      const int np = static_cast<int>(n - p0 < PTB_M ? n - p0 : PTB_M);
      mbar_wait_warp(acc_full(b), static_cast<uint32_t>((k >> 1) & 1));
      tc_fence_after();
      if (need_dw) {
        uint32_t vh[32], vl[32];
        tmem_ld32(tmem + lane_addr + Cfg::DW_COL + b * 64, vh);
        tmem_ld32(tmem + lane_addr + Cfg::DW_COL + b * 64 + 32, vl);
        tmem_wait_ld();
        asm volatile("bar.sync 1, 128;" ::: "memory");          // the previous chunk's copy-out has read dw_sm
#pragma unroll
        for (int c = 0; c < PTC_MAXC; ++c)
          if (c < C) dw_sm[r * C + c] = (__uint_as_float(vh[c]) + __uint_as_float(vl[c])) + dm_sm[c];
      }
      if (need_de) {
        bulk_wait_read0();                                      // this thread's previous row copy has left shared memory
#pragma unroll 1
        for (int c0 = 0; c0 < D; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_addr + Cfg::DE_COL + b * D + c0, v);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + (c0 + i) * 4), "r"(v[i]), "r"(v[i + 1]),
                         "r"(v[i + 2]), "r"(v[i + 3]) : "memory");
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(b));
      if (need_de) {
        fence_proxy_async_smem();
        if (r < np) bulk_store_1d(de + (p0 + r) * D, my_row, D * 4);
        bulk_commit();
      }
      if (need_dw) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float* dst = dw + p0 * C;
        for (int i = r; i < np * C; i += 128) dst[i] = dw_sm[i];
      }
    }
    if (need_de) bulk_wait0();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, Cfg::TMEM_COLS);
}

}  // namespace mb200
