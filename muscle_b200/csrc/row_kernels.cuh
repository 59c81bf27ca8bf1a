// HBM-bound per-pixel kernels around the tcgen05 pair passes of the dense InfoNCE:
//   prepare : L2-normalise, cast to bf16 (pixel-major), class bit, per-class sums
//   pos     : pixel -> class-prototype logit, seeds the row sums
//   mid     : 1/rowsum, per-pixel loss term, prototype-gradient accumulation
//   final   : sum the column splits, prototype terms, normalisation backward, layout
// All vectors of one pixel are handled by one warp (D/32 elements per lane); the NCHW <->
// pixel-major transposition goes through a padded shared-memory tile so both sides coalesce.
#pragma once
#include <cuda_bf16.h>

#include "ptx.cuh"

namespace mb200 {

constexpr int ROW_THREADS = 256;
constexpr int ROW_WARPS = ROW_THREADS / 32;
constexpr int TILE_PX = 32;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ int class_of_bits(uint32_t b) { return 31 - __clz(b); }

// Loads the D/32 consecutive bf16 values of lane `lane` (elements lane*EPL .. +EPL-1) as floats.
template <int EPL>
__device__ __forceinline__ void load_bf16_lane(const __nv_bfloat16* row, int lane, float (&v)[EPL]) {
  if constexpr (EPL == 2) {
    const __nv_bfloat162 t = reinterpret_cast<const __nv_bfloat162*>(row)[lane];
    v[0] = __low2float(t);
    v[1] = __high2float(t);
  } else if constexpr (EPL == 4) {
    const uint2 t = reinterpret_cast<const uint2*>(row)[lane];
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&t.x);
    const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
    v[0] = __low2float(a); v[1] = __high2float(a);
    v[2] = __low2float(b); v[3] = __high2float(b);
  } else {
    static_assert(EPL == 8, "EPL");
    const uint4 t = reinterpret_cast<const uint4*>(row)[lane];
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
      v[2 * i] = __low2float(a);
      v[2 * i + 1] = __high2float(a);
    }
  }
}

template <int EPL>
__device__ __forceinline__ void store_bf16_lane(__nv_bfloat16* row, int lane, const float (&v)[EPL]) {
  uint32_t w[EPL / 2];
#pragma unroll
  for (int i = 0; i < EPL / 2; ++i) w[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
  if constexpr (EPL == 2) {
    reinterpret_cast<uint32_t*>(row)[lane] = w[0];
  } else if constexpr (EPL == 4) {
    reinterpret_cast<uint2*>(row)[lane] = make_uint2(w[0], w[1]);
  } else {
    reinterpret_cast<uint4*>(row)[lane] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// Shared-memory per-class accumulators [C][D] with a bitmask of the classes that were touched,
// flushed to global with one atomic per touched element.
struct ClassAcc {
  float* acc;         // [C*D]
  float* cnt;         // [C]
  uint32_t* present;  // [1]
  __device__ void init(int C, int D) {
    for (int i = threadIdx.x; i < C * D; i += blockDim.x) acc[i] = 0.f;
    for (int i = threadIdx.x; i < C; i += blockDim.x) cnt[i] = 0.f;
    if (threadIdx.x == 0) *present = 0u;
  }
  __device__ void flush(int C, int D, float* gsum, float* gcnt) {
    const uint32_t pm = *present;
    for (int c = 0; c < C; ++c) {
      if (!((pm >> c) & 1u)) continue;
      for (int d = threadIdx.x; d < D; d += blockDim.x) atomicAdd(gsum + c * D + d, acc[c * D + d]);
      if (gcnt != nullptr && threadIdx.x == 0) atomicAdd(gcnt + c, cnt[c]);
    }
  }
};

struct PrepareParams {
  const float* x;
  const void* cls;
  int cls64;
  int layout;
  long long n;       // pixels
  long long ppi;     // pixels per image (NCHW)
  int C;
  float eps;
  __nv_bfloat16* ehat;
  uint32_t* bits;
  float* inv_norm;
  float* class_sum;
  float* class_mass;
};

template <int D>
__global__ void __launch_bounds__(ROW_THREADS) prepare_kernel(const PrepareParams p) {
  constexpr int EPL = D / 32;
  extern __shared__ float smem_f[];
  float* xs = smem_f;                              // [TILE_PX][D+1]
  ClassAcc ca;
  ca.acc = xs + TILE_PX * (D + 1);                 // [C*D]
  ca.cnt = ca.acc + p.C * D;                       // [C]
  ca.present = reinterpret_cast<uint32_t*>(ca.cnt + p.C);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  ca.init(p.C, D);
  __syncthreads();

  const long long tiles_per_img = (p.layout == MUSCLE_LAYOUT_NCHW) ? (p.ppi + TILE_PX - 1) / TILE_PX : 0;
  const long long n_tiles = (p.layout == MUSCLE_LAYOUT_NCHW) ? (p.n / p.ppi) * tiles_per_img
                                                             : (p.n + TILE_PX - 1) / TILE_PX;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    long long n0;
    int npx;
    if (p.layout == MUSCLE_LAYOUT_NCHW) {
      const long long b = tile / tiles_per_img;
      const long long p0 = (tile - b * tiles_per_img) * TILE_PX;
      n0 = b * p.ppi + p0;
      npx = static_cast<int>(min(static_cast<long long>(TILE_PX), p.ppi - p0));
      const float* src = p.x + b * D * p.ppi + p0;
      for (int d = warp; d < D; d += ROW_WARPS)
        xs[lane * (D + 1) + d] = (lane < npx) ? __ldg(src + static_cast<long long>(d) * p.ppi + lane) : 0.f;
    } else {
      n0 = tile * TILE_PX;
      npx = static_cast<int>(min(static_cast<long long>(TILE_PX), p.n - n0));
      for (int r = warp; r < npx; r += ROW_WARPS) {
        const float* src = p.x + (n0 + r) * D;
#pragma unroll
        for (int k = 0; k < EPL; ++k) xs[r * (D + 1) + lane + 32 * k] = __ldg(src + lane + 32 * k);
      }
    }
    __syncthreads();
    for (int r = warp; r < npx; r += ROW_WARPS) {
      // lane holds elements lane*EPL .. lane*EPL+EPL-1 (what the bf16 row store wants)
      float v[EPL];
      float ss = 0.f;
#pragma unroll
      for (int k = 0; k < EPL; ++k) {
        v[k] = xs[r * (D + 1) + lane * EPL + k];
        ss += v[k] * v[k];
      }
      ss = warp_sum(ss);
      const float inv = 1.f / fmaxf(sqrtf(ss), p.eps);
      long long c = p.cls64 ? reinterpret_cast<const long long*>(p.cls)[n0 + r]
                            : static_cast<long long>(reinterpret_cast<const int*>(p.cls)[n0 + r]);
      const bool valid = (c >= 0 && c < p.C);
#pragma unroll
      for (int k = 0; k < EPL; ++k) v[k] *= inv;
      store_bf16_lane<EPL>(p.ehat + (n0 + r) * D, lane, v);
      if (lane == 0) {
        p.bits[n0 + r] = valid ? (1u << c) : 0u;
        p.inv_norm[n0 + r] = inv;
      }
      if (valid) {
        // the prototypes are means of the SAME bf16-rounded vectors the tensor cores see
#pragma unroll
        for (int k = 0; k < EPL; ++k)
          atomicAdd(&ca.acc[c * D + lane * EPL + k], __bfloat162float(__float2bfloat16_rn(v[k])));
        if (lane == 0) {
          atomicAdd(&ca.cnt[c], 1.f);
          atomicOr(ca.present, 1u << c);
        }
      }
    }
    __syncthreads();
  }
  ca.flush(p.C, D, p.class_sum, p.class_mass);
}

struct PosParams {
  const __nv_bfloat16* ehat;
  const uint32_t* bits;
  long long n_rows, rows_pad;
  const float* class_sum;
  const float* class_mass;
  const float* bound;
  float inv_tau;
  float* pos;
  float* rowsum;
};

template <int D>
__global__ void __launch_bounds__(ROW_THREADS) pos_kernel(const PosParams p) {
  constexpr int EPL = D / 32;
  const int lane = threadIdx.x & 31;
  const long long warp_g = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const float bt = __ldg(p.bound) * p.inv_tau;
  for (long long r = warp_g; r < p.rows_pad; r += n_warps) {
    const uint32_t b = (r < p.n_rows) ? p.bits[r] : 0u;
    float pos = 0.f, rs = 0.f;
    if (b) {
      const int c = class_of_bits(b);
      float v[EPL];
      load_bf16_lane<EPL>(p.ehat + r * D, lane, v);
      float dot = 0.f;
#pragma unroll
      for (int k = 0; k < EPL; ++k) dot += v[k] * __ldg(p.class_sum + c * D + lane * EPL + k);
      dot = warp_sum(dot);
      pos = dot / __ldg(p.class_mass + c) * p.inv_tau;
      rs = __expf(pos - bt);
    }
    if (lane == 0) {
      p.pos[r] = pos;
      p.rowsum[r] = rs;
    }
  }
}

struct MidParams {
  const __nv_bfloat16* ehat;
  const uint32_t* bits;
  long long n_rows;
  int C;
  const float* pos;
  const float* rowsum;
  const float* bound;
  float inv_tau;
  float* w;
  float* gpos;
  double* loss_acc;
  float* dproto;
};

template <int D>
__global__ void __launch_bounds__(ROW_THREADS) mid_kernel(const MidParams p) {
  constexpr int EPL = D / 32;
  extern __shared__ float smem_f[];
  ClassAcc ca;
  ca.acc = smem_f;
  ca.cnt = ca.acc + p.C * D;
  ca.present = reinterpret_cast<uint32_t*>(ca.cnt + p.C);
  __shared__ double loss_sh[ROW_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  ca.init(p.C, D);
  __syncthreads();
  const long long warp_g = static_cast<long long>(blockIdx.x) * ROW_WARPS + warp;
  const long long n_warps = static_cast<long long>(gridDim.x) * ROW_WARPS;
  const float bt = __ldg(p.bound) * p.inv_tau;
  double loss = 0.0;
  for (long long r = warp_g; r < p.n_rows; r += n_warps) {
    const uint32_t b = p.bits[r];
    float w = 0.f, g = 0.f;
    if (b) {
      const int c = class_of_bits(b);
      const float rs = p.rowsum[r];
      const float pos = p.pos[r];
      w = 1.f / rs;
      g = (__expf(pos - bt) * w - 1.f) * p.inv_tau;
      loss += static_cast<double>(bt + logf(rs) - pos);
      float v[EPL];
      load_bf16_lane<EPL>(p.ehat + r * D, lane, v);
#pragma unroll
      for (int k = 0; k < EPL; ++k) atomicAdd(&ca.acc[c * D + lane * EPL + k], g * v[k]);
      if (lane == 0) atomicOr(ca.present, 1u << c);
    }
    if (lane == 0) {
      p.w[r] = w;
      p.gpos[r] = g;
    }
  }
  if (lane == 0) loss_sh[warp] = loss;  // every lane of the warp carries the same sum
  __syncthreads();
  ca.flush(p.C, D, p.dproto, nullptr);
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < ROW_WARPS; ++i) t += loss_sh[i];
    if (t != 0.0) atomicAdd(p.loss_acc, t);
  }
}

struct FinalParams {
  const __nv_bfloat16* ehat;
  const uint32_t* bits;
  const float* inv_norm;
  const float* part;
  int n_split;
  long long rows_pad;
  long long n;
  long long ppi;
  int C;
  const float* class_sum;
  const float* class_mass;
  const float* gpos;
  const float* dproto;
  const double* loss_acc;
  float inv_tau;
  int layout;
  float* dx;
  float* loss_out;
};

template <int D>
__global__ void __launch_bounds__(ROW_THREADS) final_kernel(const FinalParams p) {
  constexpr int EPL = D / 32;
  extern __shared__ float smem_f[];
  float* xs = smem_f;  // [TILE_PX][D+1]
  __shared__ float n_valid_sh;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    float nv = 0.f;
    for (int c = 0; c < p.C; ++c) nv += p.class_mass[c];
    n_valid_sh = nv;
    if (blockIdx.x == 0 && p.loss_out != nullptr)
      *p.loss_out = static_cast<float>(*p.loss_acc / static_cast<double>(nv));  // 0/0 -> NaN like the reference
  }
  __syncthreads();
  if (p.dx == nullptr) return;
  const float inv_n = n_valid_sh > 0.f ? 1.f / n_valid_sh : 0.f;

  const long long tiles_per_img = (p.layout == MUSCLE_LAYOUT_NCHW) ? (p.ppi + TILE_PX - 1) / TILE_PX : 0;
  const long long n_tiles = (p.layout == MUSCLE_LAYOUT_NCHW) ? (p.n / p.ppi) * tiles_per_img
                                                             : (p.n + TILE_PX - 1) / TILE_PX;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    long long n0, b = 0, p0 = 0;
    int npx;
    if (p.layout == MUSCLE_LAYOUT_NCHW) {
      b = tile / tiles_per_img;
      p0 = (tile - b * tiles_per_img) * TILE_PX;
      n0 = b * p.ppi + p0;
      npx = static_cast<int>(min(static_cast<long long>(TILE_PX), p.ppi - p0));
    } else {
      n0 = tile * TILE_PX;
      npx = static_cast<int>(min(static_cast<long long>(TILE_PX), p.n - n0));
    }
    for (int r = warp; r < npx; r += ROW_WARPS) {
      const long long q = n0 + r;
      const uint32_t bq = p.bits[q];
      float g[EPL];
#pragma unroll
      for (int k = 0; k < EPL; ++k) g[k] = 0.f;
      if (bq) {
        const int c = class_of_bits(bq);
        for (int s = 0; s < p.n_split; ++s) {
          const float* src = p.part + (static_cast<long long>(s) * p.rows_pad + q) * D + lane * EPL;
#pragma unroll
          for (int k = 0; k < EPL; ++k) g[k] += __ldg(src + k);
        }
        const float inv_mass = 1.f / __ldg(p.class_mass + c);
        const float gq = p.gpos[q] * inv_mass;
        float e[EPL];
        load_bf16_lane<EPL>(p.ehat + q * D, lane, e);
        float dot = 0.f;
#pragma unroll
        for (int k = 0; k < EPL; ++k) {
          const int d = lane * EPL + k;
          g[k] = inv_n * (p.inv_tau * g[k] + gq * __ldg(p.class_sum + c * D + d) +
                          inv_mass * __ldg(p.dproto + c * D + d));
          dot += g[k] * e[k];
        }
        dot = warp_sum(dot);
        const float inv = p.inv_norm[q];
#pragma unroll
        for (int k = 0; k < EPL; ++k) g[k] = inv * (g[k] - e[k] * dot);
      }
      if (p.layout == MUSCLE_LAYOUT_NCHW) {
#pragma unroll
        for (int k = 0; k < EPL; ++k) xs[r * (D + 1) + lane * EPL + k] = g[k];
      } else {
        float* dst = p.dx + q * D + lane * EPL;
#pragma unroll
        for (int k = 0; k < EPL; ++k) dst[k] = g[k];
      }
    }
    if (p.layout == MUSCLE_LAYOUT_NCHW) {
      __syncthreads();
      float* dst = p.dx + b * D * p.ppi + p0;
      for (int d = warp; d < D; d += ROW_WARPS)
        if (lane < npx) dst[static_cast<long long>(d) * p.ppi + lane] = xs[lane * (D + 1) + d];
      __syncthreads();
    }
  }
}

}  // namespace mb200
