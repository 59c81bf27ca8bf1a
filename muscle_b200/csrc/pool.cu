// CAM-weighted (soft) prototype pooling and its gradient scatter (include/muscle_b200.h,
// muscle_pool_prototypes_*).  HBM-bound kernels: every embedding row and every weight row is read
// once, nothing else of size N is written in the forward.
//
//   forward :  sum[c][:] = sum_p w[p][c] e[p][:],  mass[c] = sum_p w[p][c]      (prototype = sum / mass)
//   backward:  de[p][:]  = sum_c w[p][c] ds[c][:],  dw[p][c] = e[p] . ds[c] + dm[c]
//              with ds = dproto / mass and dm[c] = -(dproto[c] . proto[c]) / mass[c] formed on the fly
//
// References: the positive-key mean of src/loss_multilabel.py:12 (0/1 weights), the region mean of
// src/loss_multilabel.py:255 and the CAM aggregation `cam @ aff` of src/MuSCLe.py:222 (w = cam^T, no
// division) are all instances.  Layout: e [N, D] pixel-major (bf16 = the sorted `ehat` of the InfoNCE
// path, or fp32), w [N, C] pixel-major fp32, C <= 24, D % 4 == 0, D <= 256 (generic lane mapping).
//
// Forward mapping: one warp = one pixel at a time, lane l owns the 4-channel groups l, l+32, ... of the
// row (8-byte / 16-byte coalesced loads); the lane keeps C x (its channels) partial sums in registers
// (C = 21, D = 128: 84 accumulators, updated with packed FFMA2), so a pixel costs one row load, C
// broadcast weights from shared memory and 2 C packed FMAs per lane.  At 7.9 FMA per input byte plain
// per-pixel FMAs could not keep up with HBM; with this register blocking they do.
#include <cuda_bf16.h>
#include <algorithm>
#include <cstdlib>

#include "host_util.cuh"
#include "ptx.cuh"
#include "pool_tc.cuh"

using namespace mb200;

namespace {

constexpr int POOL_THREADS = 256;
constexpr int POOL_WARPS = POOL_THREADS / 32;
constexpr int POOL_CHUNK = 64;      // backward: pixels per block
constexpr int POOL_FCHUNK = 128;    // forward: pixels whose weights are staged per step
constexpr int POOL_UNROLL = 4;      // rows in flight per warp (one row per warp left the kernels latency bound: 5 % of HBM)
constexpr int POOL_MAXC = 24;

__device__ __forceinline__ void load4(const float* row, int g, float (&v)[4]) {
  const float4 t = __ldg(reinterpret_cast<const float4*>(row) + g);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const __nv_bfloat16* row, int g, float (&v)[4]) {
  const uint2 t = __ldg(reinterpret_cast<const uint2*>(row) + g);
  v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
  v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
}

// Forward.  One accumulator set [C x D] per BLOCK (an earlier version kept one per warp: its end-of-kernel
// reduction moved as many values as the kernel read and ran at shared-memory-atomic speed, 65 us for 22 MB):
// warp w owns the classes w, w + 8, w + 16 (CM = 24: three per warp), lane l the 4-channel groups l, l + 32, ...;
// the chunk's rows are staged in shared memory once (coalesced 16-byte loads, the next chunk's in flight in
// registers while this one is consumed) and every warp reads all of them: per pixel and warp one 8/16-byte
// shared load, CPW broadcast weights and 2 CPW packed FMAs.  NG = 4-channel groups per lane (D <= 128: 1, else 2).
template <typename T, int CM, int NG>
__global__ void __launch_bounds__(POOL_THREADS) pool_fwd_kernel(const T* __restrict__ e, const float* __restrict__ w,
                                                                long long n, int dim, int C, float* __restrict__ sum,
                                                                float* __restrict__ mass) {
  constexpr int CPW = (CM + POOL_WARPS - 1) / POOL_WARPS;      // classes per warp
  constexpr int ROW16 = 128 * NG * static_cast<int>(sizeof(T)) / 16;   // 16-byte pieces per staged row (row padded to 128 NG)
  constexpr int PIECES = POOL_FCHUNK * ROW16 / POOL_THREADS;   // per thread and chunk
  __shared__ float2 w_sm[POOL_FCHUNK][CM];                     // weights duplicated into both halves (FFMA2 operand)
  extern __shared__ uint4 rows_sm[];                           // [POOL_FCHUNK][ROW16]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int groups = dim / 4;
  const int row16 = dim * static_cast<int>(sizeof(T)) / 16;    // valid 16-byte pieces per row
  uint64_t acc[CPW][NG][2];
#pragma unroll
  for (int k = 0; k < CPW; ++k)
#pragma unroll
    for (int g = 0; g < NG; ++g) acc[k][g][0] = acc[k][g][1] = 0ull;
  float macc[CPW];
#pragma unroll
  for (int k = 0; k < CPW; ++k) macc[k] = 0.f;
  const long long n_chunks = (n + POOL_FCHUNK - 1) / POOL_FCHUNK;
  uint4 pre[PIECES];
  auto fetch = [&](long long ch) {                             // the chunk's rows -> registers (coalesced)
    const long long p0 = ch * POOL_FCHUNK;
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int idx = threadIdx.x + i * POOL_THREADS;
      const int pp = idx / ROW16, q = idx - pp * ROW16;
      pre[i] = (p0 + pp < n && q < row16) ? __ldg(reinterpret_cast<const uint4*>(e + (p0 + pp) * dim) + q) : make_uint4(0u, 0u, 0u, 0u);
    }
  };
  long long ch = blockIdx.x;
  if (ch < n_chunks) fetch(ch);
  for (; ch < n_chunks; ch += gridDim.x) {
    const long long p0 = ch * POOL_FCHUNK;
    const int np = static_cast<int>(min(static_cast<long long>(POOL_FCHUNK), n - p0));
    __syncthreads();                                           // the previous chunk has been consumed
#pragma unroll
    for (int i = 0; i < PIECES; ++i) rows_sm[threadIdx.x + i * POOL_THREADS] = pre[i];
    for (int i = threadIdx.x; i < np * C; i += POOL_THREADS) {
      const float v = __ldg(w + p0 * C + i);
      w_sm[i / C][i % C] = make_float2(v, v);
    }
    __syncthreads();
    if (ch + gridDim.x < n_chunks) fetch(ch + gridDim.x);      // next chunk in flight under this one's arithmetic
    for (int pp = 0; pp < np; ++pp) {
      float v[NG][4];
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        if (sizeof(T) == 2) {
          const uint2 t = reinterpret_cast<const uint2*>(rows_sm + pp * ROW16)[lane + 32 * g];
          v[g][0] = __uint_as_float(t.x << 16); v[g][1] = __uint_as_float(t.x & 0xffff0000u);
          v[g][2] = __uint_as_float(t.y << 16); v[g][3] = __uint_as_float(t.y & 0xffff0000u);
        } else {
          const uint4 t = rows_sm[pp * ROW16 + lane + 32 * g];
          v[g][0] = __uint_as_float(t.x); v[g][1] = __uint_as_float(t.y);
          v[g][2] = __uint_as_float(t.z); v[g][3] = __uint_as_float(t.w);
        }
      }
#pragma unroll
      for (int k = 0; k < CPW; ++k) {
        const int c = warp + POOL_WARPS * k;
        if (c < C) {
          const float2 wv = w_sm[pp][c];
          const uint64_t wc = pack2(wv.x, wv.y);
          macc[k] += wv.x;
#pragma unroll
          for (int g = 0; g < NG; ++g) {
            acc[k][g][0] = fma2(pack2(v[g][0], v[g][1]), wc, acc[k][g][0]);
            acc[k][g][1] = fma2(pack2(v[g][2], v[g][3]), wc, acc[k][g][1]);
          }
        }
      }
    }
  }
  // one global atomic per (class, channel) and block
#pragma unroll
  for (int k = 0; k < CPW; ++k) {
    const int c = warp + POOL_WARPS * k;
    if (c >= C) continue;
#pragma unroll
    for (int g = 0; g < NG; ++g)
      if (lane + 32 * g < groups) {
        float a0, a1, a2, a3;
        unpack2(acc[k][g][0], a0, a1);
        unpack2(acc[k][g][1], a2, a3);
        float* dst = sum + c * dim + 4 * (lane + 32 * g);
        if (a0 != 0.f) atomicAdd(dst, a0);
        if (a1 != 0.f) atomicAdd(dst + 1, a1);
        if (a2 != 0.f) atomicAdd(dst + 2, a2);
        if (a3 != 0.f) atomicAdd(dst + 3, a3);
      }
    if (lane == 0 && macc[k] != 0.f) atomicAdd(mass + c, macc[k]);
  }
}

// proto = sum / max(mass, eps)
__global__ void pool_finish_kernel(const float* sum, const float* mass, int C, int dim, float eps, float* proto) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < C * dim) proto[i] = sum[i] / fmaxf(mass[i / dim], eps);
}

// Backward.  Block = POOL_CHUNK pixels.  Phase A (warp per pixel, as the forward): de = sum_c w_c ds_c.
// Phase B (thread = (pixel, class group)): dw = e . ds_c + dm_c from the chunk's rows staged in shared memory
// (row stride dim + 4 floats: conflict-free 16-byte reads down a column of pixels).
template <typename T, int CM, int NG>
__global__ void __launch_bounds__(POOL_THREADS) pool_bwd_kernel(const T* __restrict__ e, const float* __restrict__ w,
                                                                long long n, int dim, int C, const float* __restrict__ proto,
                                                                const float* __restrict__ mass,
                                                                const float* __restrict__ dproto, float eps,
                                                                float* __restrict__ de, float* __restrict__ dw) {
  extern __shared__ float sm[];
  float* ds = sm;                                  // [C][dim]   dproto / mass
  float* e_sm = ds + C * dim;                      // [POOL_CHUNK][dim + 4]
  __shared__ float w_sm[POOL_CHUNK][CM];
  __shared__ float dm_sm[POOL_MAXC];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int groups = dim / 4;
  const int estride = dim + 4;
  for (int i = threadIdx.x; i < C * dim; i += POOL_THREADS) ds[i] = __ldg(dproto + i) / fmaxf(__ldg(mass + i / dim), eps);
  __syncthreads();
  for (int c = warp; c < C; c += POOL_WARPS) {     // dm[c] = -(ds[c] . proto[c])   (mass below eps: no dependence)
    float s = 0.f;
    for (int d = lane; d < dim; d += 32) s = fmaf(ds[c * dim + d], __ldg(proto + c * dim + d), s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) dm_sm[c] = (__ldg(mass + c) >= eps) ? -s : 0.f;
  }
  const long long p0 = static_cast<long long>(blockIdx.x) * POOL_CHUNK;
  const int np = static_cast<int>(min(static_cast<long long>(POOL_CHUNK), n - p0));
  for (int i = threadIdx.x; i < np * C; i += POOL_THREADS) w_sm[i / C][i % C] = __ldg(w + p0 * C + i);
  __syncthreads();
  // phase A  (ds read from shared memory per class: keeping the lane's slice of ds in registers instead -- 84 of
  // them -- drops the occupancy to one block per SM and measured 111 us against 70 us at 65 536 pixels)
  for (int pp0 = warp; pp0 < np; pp0 += POOL_WARPS * POOL_UNROLL) {
    float v[POOL_UNROLL][NG][4];
#pragma unroll
    for (int u = 0; u < POOL_UNROLL; ++u) {                      // row loads first: POOL_UNROLL lines in flight per warp
      const int pp = pp0 + u * POOL_WARPS;
#pragma unroll
      for (int g = 0; g < NG; ++g)
        if (pp < np && lane + 32 * g < groups) load4(e + (p0 + pp) * dim, lane + 32 * g, v[u][g]);
    }
#pragma unroll
    for (int u = 0; u < POOL_UNROLL; ++u) {
      const int pp = pp0 + u * POOL_WARPS;
      if (pp >= np) break;
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        const int gi = lane + 32 * g;
        if (gi >= groups) continue;
        if (dw != nullptr)
          *reinterpret_cast<float4*>(e_sm + pp * estride + 4 * gi) = make_float4(v[u][g][0], v[u][g][1], v[u][g][2], v[u][g][3]);
        if (de != nullptr) {
          uint64_t a0 = 0ull, a1 = 0ull;
#pragma unroll
          for (int c = 0; c < CM; ++c) {
            if (c < C) {
              const float wc = w_sm[pp][c];
              const uint64_t wc2 = pack2(wc, wc);
              const float4 s4 = *reinterpret_cast<const float4*>(ds + c * dim + 4 * gi);
              a0 = fma2(pack2(s4.x, s4.y), wc2, a0);
              a1 = fma2(pack2(s4.z, s4.w), wc2, a1);
            }
          }
          float r0, r1, r2, r3;
          unpack2(a0, r0, r1);
          unpack2(a1, r2, r3);
          *reinterpret_cast<float4*>(de + (p0 + pp) * dim + 4 * gi) = make_float4(r0, r1, r2, r3);
        }
      }
    }
  }
  if (dw == nullptr) return;
  __syncthreads();
  // phase B: thread -> pixel pp = tid % 64, classes c = cg, cg + 4, ... (cg = tid / 64)
  const int pp = threadIdx.x & (POOL_CHUNK - 1), cg = threadIdx.x / POOL_CHUNK;
  constexpr int CPT = (CM + 3) / 4;
  if (pp < np) {
    float acc[CPT];
#pragma unroll
    for (int k = 0; k < CPT; ++k) acc[k] = 0.f;
    const float* er = e_sm + pp * estride;
    for (int d = 0; d < dim; d += 4) {
      const float4 ev = *reinterpret_cast<const float4*>(er + d);
#pragma unroll
      for (int k = 0; k < CPT; ++k) {
        const int c = cg + 4 * k;
        if (c < C) {
          const float4 s4 = *reinterpret_cast<const float4*>(ds + c * dim + d);
          acc[k] = fmaf(ev.x, s4.x, fmaf(ev.y, s4.y, fmaf(ev.z, s4.z, fmaf(ev.w, s4.w, acc[k]))));
        }
      }
    }
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      const int c = cg + 4 * k;
      if (c < C) dw[(p0 + pp) * C + c] = acc[k] + dm_sm[c];
    }
  }
}

// sum[] and mass[] start at zero (one launch instead of two memset nodes: it takes part in the programmatic launch chain)
__global__ void pool_zero_kernel(float* sum, int n_sum, float* mass, int n_mass) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_sum) sum[i] = 0.f;
  if (i < n_mass) mass[i] = 0.f;
}

// MUSCLE_POOL_TC=0 keeps the CUDA-core kernels for bf16 inputs too (A/B measurements, tests of both paths); read per call
bool pool_tc_enabled() {
  const char* v = getenv("MUSCLE_POOL_TC");
  return v == nullptr || atoi(v) != 0;
}
bool pool_tc_fwd_ok(const void* e, int e_is_bf16, int dim, const float* sum) {
  return e_is_bf16 && (dim == 64 || dim == 128 || dim == 256) && (reinterpret_cast<uintptr_t>(e) & 127) == 0 &&
         (reinterpret_cast<uintptr_t>(sum) & 15) == 0 && pool_tc_enabled();
}

constexpr int PTC_MAX_PARTS = 192;     // CTAs of the tensor-pipe forward (<= SMs): bounds the caller's scratch
static_assert(PTC_MAX_PARTS <= 32 * PTC_FIN_LOADS, "pool_tc_finish_kernel reads at most 32 x PTC_FIN_LOADS parts");

template <int D>
int launch_fwd_tc(const void* e, const float* w, int64_t n, int C, float* sum, float* mass, float* partial, int* parts,
                  cudaStream_t stream) {
  using Cfg = PoolTcCfg<D>;
  CUtensorMap tm;
  if (int rc = make_bf16_panel_map(&tm, e, n, D, PTC_K)) return rc;
  if (int rc = ensure_dynamic_smem<pool_tc_fwd_kernel<D>>(Cfg::SMEM_BYTES)) return rc;
  const long long n_chunks = (n + PTC_K - 1) / PTC_K;
  const int grid = static_cast<int>(std::min<long long>(n_chunks, std::min(device_info().sm_count, PTC_MAX_PARTS)));
  const int w_bulk = (reinterpret_cast<uintptr_t>(w) & 15) == 0 ? 1 : 0;     // bulk copies want 16-byte aligned sources
  MB_CUDA(launch_pdl(pool_tc_fwd_kernel<D>, dim3(grid), dim3(PTC_THREADS), Cfg::SMEM_BYTES, stream, tm, w,
                     static_cast<long long>(n), C, w_bulk, sum, mass, partial));
  MB_LAUNCH_CHECK();
  *parts = grid;
  return MUSCLE_OK;
}

template <int D>
int launch_bwd_tc(const void* e, const float* w, int64_t n, int C, const float* proto, const float* mass, const float* dproto,
                  float eps, float* de, float* dw, cudaStream_t stream) {
  using Cfg = PoolTcBwdCfg<D>;
  CUtensorMap tm;
  if (int rc = make_bf16_panel_map(&tm, e, n, D, PTB_M)) return rc;
  if (int rc = ensure_dynamic_smem<pool_tc_bwd_kernel<D>>(Cfg::SMEM_BYTES)) return rc;
  const long long n_chunks = (n + PTB_M - 1) / PTB_M;
  const int grid = static_cast<int>(std::min<long long>(n_chunks, device_info().sm_count));
  MB_CUDA(launch_pdl(pool_tc_bwd_kernel<D>, dim3(grid), dim3(PTB_THREADS), Cfg::SMEM_BYTES, stream, tm, w,
                     static_cast<long long>(n), C, proto, mass, dproto, eps, de, dw));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int check_device_pool() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}

int check_pool_args(int64_t n, int dim, int C) {
  if (n <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (dim % 4 != 0 || dim < 4 || dim > 256) return fail(MUSCLE_ERR_BAD_ARG, "dim must be a multiple of 4 in [4,256] (got %d)", dim);
  if (C < 1 || C > POOL_MAXC) return fail(MUSCLE_ERR_BAD_ARG, "num_classes must be in [1,%d]", POOL_MAXC);
  return MUSCLE_OK;
}

template <typename T, int CM, int NG>
int launch_fwd_cm(const T* e, const float* w, int64_t n, int dim, int C, float* sum, float* mass, cudaStream_t stream) {
  const long long n_chunks = (n + POOL_FCHUNK - 1) / POOL_FCHUNK;
  const int grid = static_cast<int>(std::min<long long>(n_chunks, 2ll * device_info().sm_count));
  const size_t smem = static_cast<size_t>(POOL_FCHUNK) * 128 * NG * sizeof(T);
  if (int rc = ensure_dynamic_smem<pool_fwd_kernel<T, CM, NG>>(static_cast<int>(smem))) return rc;
  pool_fwd_kernel<T, CM, NG><<<grid, POOL_THREADS, smem, stream>>>(e, w, n, dim, C, sum, mass);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
template <typename T, int NG>
int launch_fwd(const T* e, const float* w, int64_t n, int dim, int C, float* sum, float* mass, cudaStream_t stream) {
  if (C <= 8) return launch_fwd_cm<T, 8, NG>(e, w, n, dim, C, sum, mass, stream);
  if (C <= 16) return launch_fwd_cm<T, 16, NG>(e, w, n, dim, C, sum, mass, stream);
  return launch_fwd_cm<T, 24, NG>(e, w, n, dim, C, sum, mass, stream);
}

template <typename T, int CM, int NG>
int launch_bwd_cm(const T* e, const float* w, int64_t n, int dim, int C, const float* proto, const float* mass,
                  const float* dproto, float eps, float* de, float* dw, cudaStream_t stream) {
  const int grid = static_cast<int>((n + POOL_CHUNK - 1) / POOL_CHUNK);
  const size_t smem = (static_cast<size_t>(C) * dim + static_cast<size_t>(POOL_CHUNK) * (dim + 4)) * sizeof(float);
  if (int rc = ensure_dynamic_smem<pool_bwd_kernel<T, CM, NG>>(static_cast<int>((24 * 256 + POOL_CHUNK * 260) * sizeof(float)))) return rc;
  pool_bwd_kernel<T, CM, NG><<<grid, POOL_THREADS, smem, stream>>>(e, w, n, dim, C, proto, mass, dproto, eps, de, dw);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
template <typename T, int NG>
int launch_bwd(const T* e, const float* w, int64_t n, int dim, int C, const float* proto, const float* mass,
               const float* dproto, float eps, float* de, float* dw, cudaStream_t stream) {
  if (C <= 8) return launch_bwd_cm<T, 8, NG>(e, w, n, dim, C, proto, mass, dproto, eps, de, dw, stream);
  if (C <= 16) return launch_bwd_cm<T, 16, NG>(e, w, n, dim, C, proto, mass, dproto, eps, de, dw, stream);
  return launch_bwd_cm<T, 24, NG>(e, w, n, dim, C, proto, mass, dproto, eps, de, dw, stream);
}

}  // namespace

extern "C" {

size_t muscle_pool_prototypes_fwd_scratch_bytes(int dim, int num_classes) {
  if (dim <= 0 || num_classes <= 0) return 0;
  return static_cast<size_t>(PTC_MAX_PARTS) * (static_cast<size_t>(num_classes) * dim + 32) * sizeof(float);
}

int muscle_pool_prototypes_fwd_ws(const void* e, int e_is_bf16, const float* w, int64_t n, int dim, int num_classes,
                                  float eps, float* sum, float* mass, float* proto, void* scratch, size_t scratch_bytes,
                                  void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device_pool()) return rc;
  if (!e || !w || !sum || !mass) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (int rc = check_pool_args(n, dim, num_classes)) return rc;
  const bool ws_ok = scratch != nullptr && (reinterpret_cast<uintptr_t>(scratch) & 15) == 0 &&
                     scratch_bytes >= muscle_pool_prototypes_fwd_scratch_bytes(dim, num_classes);
  if (scratch != nullptr && !ws_ok)
    return fail(MUSCLE_ERR_BAD_ARG, "scratch must be 16-byte aligned and hold muscle_pool_prototypes_fwd_scratch_bytes() bytes");
  if (!ws_ok || !pool_tc_fwd_ok(e, e_is_bf16, dim, sum))
    return muscle_pool_prototypes_fwd(e, e_is_bf16, w, n, dim, num_classes, eps, sum, mass, proto, stream_);
  // tensor-pipe kernel with per-CTA parts in the scratch, then one kernel that adds them in a fixed order: two launches,
  // no atomics, bit-reproducible
  float* partial = static_cast<float*>(scratch);
  int parts = 0;
  int rc = dim == 64    ? launch_fwd_tc<64>(e, w, n, num_classes, sum, mass, partial, &parts, stream)
           : dim == 128 ? launch_fwd_tc<128>(e, w, n, num_classes, sum, mass, partial, &parts, stream)
                        : launch_fwd_tc<256>(e, w, n, num_classes, sum, mass, partial, &parts, stream);
  if (rc) return rc;
  const int total = num_classes * dim;
  MB_CUDA(launch_pdl(pool_tc_finish_kernel, dim3(total / 32), dim3(PTC_FIN_THREADS), 0, stream, static_cast<const float*>(partial), parts,
                     num_classes, dim, eps, sum, mass, proto));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_pool_prototypes_fwd(const void* e, int e_is_bf16, const float* w, int64_t n, int dim, int num_classes,
                               float eps, float* sum, float* mass, float* proto, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device_pool()) return rc;
  if (!e || !w || !sum || !mass) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (int rc = check_pool_args(n, dim, num_classes)) return rc;
  if (pool_tc_fwd_ok(e, e_is_bf16, dim, sum)) {      // the tensor-pipe kernel (pool_tc.cuh): zero -> pool -> finish, one chain
    const int total = num_classes * dim;
    MB_CUDA(launch_pdl(pool_zero_kernel, dim3((total + 255) / 256), dim3(256), 0, stream, sum, total, mass, num_classes));
    MB_LAUNCH_CHECK();
    int parts = 0;
    int rc = dim == 64    ? launch_fwd_tc<64>(e, w, n, num_classes, sum, mass, nullptr, &parts, stream)
             : dim == 128 ? launch_fwd_tc<128>(e, w, n, num_classes, sum, mass, nullptr, &parts, stream)
                          : launch_fwd_tc<256>(e, w, n, num_classes, sum, mass, nullptr, &parts, stream);
    if (rc) return rc;
    if (proto) {
      MB_CUDA(launch_pdl(pool_finish_kernel, dim3((total + 255) / 256), dim3(256), 0, stream, static_cast<const float*>(sum),
                         static_cast<const float*>(mass), num_classes, dim, eps, proto));
      MB_LAUNCH_CHECK();
    }
    return MUSCLE_OK;
  }
  MB_CUDA(cudaMemsetAsync(sum, 0, static_cast<size_t>(num_classes) * dim * sizeof(float), stream));
  MB_CUDA(cudaMemsetAsync(mass, 0, static_cast<size_t>(num_classes) * sizeof(float), stream));
  int rc;
  if (e_is_bf16)
    rc = dim <= 128 ? launch_fwd<__nv_bfloat16, 1>(static_cast<const __nv_bfloat16*>(e), w, n, dim, num_classes, sum, mass, stream)
                    : launch_fwd<__nv_bfloat16, 2>(static_cast<const __nv_bfloat16*>(e), w, n, dim, num_classes, sum, mass, stream);
  else
    rc = dim <= 128 ? launch_fwd<float, 1>(static_cast<const float*>(e), w, n, dim, num_classes, sum, mass, stream)
                    : launch_fwd<float, 2>(static_cast<const float*>(e), w, n, dim, num_classes, sum, mass, stream);
  if (rc) return rc;
  if (proto) {
    const int total = num_classes * dim;
    pool_finish_kernel<<<(total + 255) / 256, 256, 0, stream>>>(sum, mass, num_classes, dim, eps, proto);
    MB_LAUNCH_CHECK();
  }
  return MUSCLE_OK;
}

int muscle_pool_prototypes_bwd(const void* e, int e_is_bf16, const float* w, int64_t n, int dim, int num_classes,
                               float eps, const float* proto, const float* mass, const float* dproto, float* de,
                               float* dw, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device_pool()) return rc;
  if (!e || !w || !proto || !mass || !dproto) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!de && !dw) return MUSCLE_OK;
  if (int rc = check_pool_args(n, dim, num_classes)) return rc;
  if (e_is_bf16 && (dim == 64 || dim == 128) && (reinterpret_cast<uintptr_t>(e) & 127) == 0 &&
      (reinterpret_cast<uintptr_t>(de) & 15) == 0 && pool_tc_enabled())      // the tensor-pipe kernel (pool_tc.cuh)
    return dim == 64 ? launch_bwd_tc<64>(e, w, n, num_classes, proto, mass, dproto, eps, de, dw, stream)
                     : launch_bwd_tc<128>(e, w, n, num_classes, proto, mass, dproto, eps, de, dw, stream);
  if (e_is_bf16)
    return dim <= 128 ? launch_bwd<__nv_bfloat16, 1>(static_cast<const __nv_bfloat16*>(e), w, n, dim, num_classes, proto, mass, dproto, eps, de, dw, stream)
                      : launch_bwd<__nv_bfloat16, 2>(static_cast<const __nv_bfloat16*>(e), w, n, dim, num_classes, proto, mass, dproto, eps, de, dw, stream);
  return dim <= 128 ? launch_bwd<float, 1>(static_cast<const float*>(e), w, n, dim, num_classes, proto, mass, dproto, eps, de, dw, stream)
                    : launch_bwd<float, 2>(static_cast<const float*>(e), w, n, dim, num_classes, proto, mass, dproto, eps, de, dw, stream);
}

}  // extern "C"
