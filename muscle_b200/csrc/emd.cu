// Region matching of the MCL stage:
//   crops   : get_dynamic_crops (src/torchutils.py:217-291) — window -> bilinear resize
//             (align_corners=True) -> optional 4x4 average pool, for every window of a batch in ONE
//             launch (the reference issues one interpolate + one avg_pool2d per window);
//   EMD     : EMD.dynamic_matching (src/loss_multilabel.py:287-326) — for every (crop1, crop2) pair
//             of an image: cost = 1 - x y^T (pair_wise_cos :232-242), weights (get_weight_vector
//             :250-257), 10 log-domain Sinkhorn iterations (:201-230); keep the pair with the
//             smallest score; gradient of that pair's value through the unrolled iterations.
// One CTA per pair keeps both regions (<= 1024 pixels x 21 channels each) in shared memory and
// recomputes the cost entries on the fly, so no N x M matrix is ever stored; the reference
// materialises an [M,N,C] broadcast product and 21 N x M temporaries per pair.  fp32 throughout
// (latency/launch bound work: the point is one launch per stage and no host round trips).
#include <algorithm>
#include <cfloat>

#include "host_util.cuh"

using namespace mb200;

namespace {

constexpr int EC = 21;        // CAM channels of the regions (VOC: 20 + bg)
constexpr int ECP = 24;       // padded row (float4 broadcast loads)
constexpr int EMD_THREADS = 512;
constexpr int EMD_MAX_PIX = 1024;
constexpr float LOG2E = 1.4426950408889634f;

int need_sm100() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum_f(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
  if (warp == 0) {
    t = warp_sum_f(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_max_f(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : -FLT_MAX;
  if (warp == 0) {
    t = warp_max_f(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// ==========================================================================================
// crops
// ==========================================================================================
struct CropDesc {   // one window; mirrors the int32[10] rows the host passes
  int b, top, left, lh, lw, rh, rw, pool, oh, ow;
};

// bilinear sample of plane `src` (row stride W) window (top,left,lh,lw) at resized position (y,x) of
// an (rh,rw) grid, align_corners=True — the index arithmetic of torch's upsample_bilinear2d (fp32)
__device__ __forceinline__ void bilinear_taps(int y, int x, const CropDesc& d, int& y0, int& y1, int& x0, int& x1,
                                              float& ly, float& lx) {
  const float sy = (d.rh > 1) ? static_cast<float>(d.lh - 1) / static_cast<float>(d.rh - 1) : 0.f;
  const float sx = (d.rw > 1) ? static_cast<float>(d.lw - 1) / static_cast<float>(d.rw - 1) : 0.f;
  const float fy = sy * y, fx = sx * x;
  y0 = static_cast<int>(fy);
  x0 = static_cast<int>(fx);
  y1 = y0 + ((y0 < d.lh - 1) ? 1 : 0);
  x1 = x0 + ((x0 < d.lw - 1) ? 1 : 0);
  ly = fy - y0;
  lx = fx - x0;
}

__device__ __forceinline__ int find_crop(const long long* offs, int n_crops, long long e) {
  int lo = 0, hi = n_crops;   // offs[lo] <= e < offs[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (offs[mid] <= e) lo = mid; else hi = mid;
  }
  return lo;
}

template <bool BWD>
__global__ void __launch_bounds__(256) crops_kernel(const float* x, float* dx, int C, int H, int W,
                                                    const CropDesc* desc, const long long* offs, int n_crops,
                                                    long long total, float* flat, const float* dflat) {
  const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const int ci = find_crop(offs, n_crops, e);
  const CropDesc d = desc[ci];
  long long r = e - offs[ci];
  const int ox = static_cast<int>(r % d.ow);
  r /= d.ow;
  const int oy = static_cast<int>(r % d.oh);
  const int c = static_cast<int>(r / d.oh);
  const long long plane = (static_cast<long long>(d.b) * C + c) * H * W + static_cast<long long>(d.top) * W + d.left;
  const int reps = d.pool ? 4 : 1;
  const float wgt = d.pool ? (1.f / 16.f) : 1.f;
  float acc = 0.f;
  const float g = BWD ? dflat[e] * wgt : 0.f;
  if (BWD && g == 0.f) return;
  for (int py = 0; py < reps; ++py)
    for (int px = 0; px < reps; ++px) {
      int y0, y1, x0, x1;
      float ly, lx;
      bilinear_taps(oy * reps + py, ox * reps + px, d, y0, y1, x0, x1, ly, lx);
      const float hy = 1.f - ly, hx = 1.f - lx;
      if (!BWD) {
        const float* s = x + plane;
        acc += hy * (hx * s[static_cast<long long>(y0) * W + x0] + lx * s[static_cast<long long>(y0) * W + x1]) +
               ly * (hx * s[static_cast<long long>(y1) * W + x0] + lx * s[static_cast<long long>(y1) * W + x1]);
      } else {
        float* s = dx + plane;
        atomicAdd(s + static_cast<long long>(y0) * W + x0, g * hy * hx);
        atomicAdd(s + static_cast<long long>(y0) * W + x1, g * hy * lx);
        atomicAdd(s + static_cast<long long>(y1) * W + x0, g * ly * hx);
        atomicAdd(s + static_cast<long long>(y1) * W + x1, g * ly * lx);
      }
    }
  if (!BWD) flat[e] = acc * wgt;
}

}  // namespace

namespace {

// ==========================================================================================
// Sinkhorn EMD
// ==========================================================================================
struct PairDesc {   // int64[6] rows from the host
  long long off1, n, off2, m, image, pad;
};

// Shared-memory working set of one pair.
struct EmdSmem {
  float* xs;     // [n][ECP]
  float* ys;     // [m][ECP]
  float* u;      // [n]
  float* v;      // [m]
  float* lmu;    // [n] log(w1 + 1e-6)
  float* lnu;    // [m] log(w2 + 1e-6)
  float* red;    // [32]
  float* mean;   // [2][ECP]
};
__host__ __device__ inline size_t emd_smem_floats(long long n, long long m) { return (n + m) * (ECP + 3) + 32 + 2 * ECP; }

__device__ __forceinline__ EmdSmem emd_carve(float* base, int n, int m) {
  EmdSmem s;
  s.xs = base;
  s.ys = s.xs + static_cast<size_t>(n) * ECP;
  s.u = s.ys + static_cast<size_t>(m) * ECP;
  s.v = s.u + n;
  s.lmu = s.v + m;
  s.lnu = s.lmu + n;
  s.red = s.lnu + m;
  s.mean = s.red + 32;
  return s;
}

// dot of a register row with a shared-memory row (float4 broadcast loads)
__device__ __forceinline__ float dot_row(const float (&a)[ECP], const float* b) {
  float s = 0.f;
#pragma unroll
  for (int q = 0; q < ECP / 4; ++q) {
    const float4 t = reinterpret_cast<const float4*>(b)[q];
    s = fmaf(a[4 * q], t.x, s);
    s = fmaf(a[4 * q + 1], t.y, s);
    s = fmaf(a[4 * q + 2], t.z, s);
    s = fmaf(a[4 * q + 3], t.w, s);
  }
  return s;
}
__device__ __forceinline__ void load_row(float (&a)[ECP], const float* src) {
#pragma unroll
  for (int q = 0; q < ECP / 4; ++q) {
    const float4 t = reinterpret_cast<const float4*>(src)[q];
    a[4 * q] = t.x; a[4 * q + 1] = t.y; a[4 * q + 2] = t.z; a[4 * q + 3] = t.w;
  }
}

// regions -> smem (pixel-major, zero-padded rows), weights (get_weight_vector) and their logs
__device__ void emd_load(const float* flat1, const float* flat2, const PairDesc& d, const EmdSmem& s) {
  const int n = static_cast<int>(d.n), m = static_cast<int>(d.m);
  for (int e = threadIdx.x; e < n * ECP; e += blockDim.x) {
    const int i = e / ECP, c = e - i * ECP;
    s.xs[e] = (c < EC) ? flat1[d.off1 + static_cast<long long>(c) * n + i] : 0.f;
  }
  for (int e = threadIdx.x; e < m * ECP; e += blockDim.x) {
    const int j = e / ECP, c = e - j * ECP;
    s.ys[e] = (c < EC) ? flat2[d.off2 + static_cast<long long>(c) * m + j] : 0.f;
  }
  __syncthreads();
  // channel means of both regions: warp w handles channels w, w+16, ...
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int c = warp; c < 2 * ECP; c += nw) {
    const bool second = c >= ECP;
    const int cc = second ? c - ECP : c;
    const float* src = second ? s.ys : s.xs;
    const int cnt = second ? m : n;
    float t = 0.f;
    for (int i = lane; i < cnt; i += 32) t += src[i * ECP + cc];
    t = warp_sum_f(t);
    if (lane == 0) s.mean[c] = t / static_cast<float>(cnt);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float w = 0.f;
    for (int c = 0; c < EC; ++c) w = fmaf(s.xs[i * ECP + c], s.mean[ECP + c], w);   // x_i . mean(y)
    s.lmu[i] = logf(w + 1e-6f);
    s.u[i] = 0.f;
  }
  for (int j = threadIdx.x; j < m; j += blockDim.x) {
    float w = 0.f;
    for (int c = 0; c < EC; ++c) w = fmaf(s.ys[j * ECP + c], s.mean[c], w);        // y_j . mean(x)
    s.lnu[j] = logf(w + 1e-6f);
    s.v[j] = 0.f;
  }
  __syncthreads();
}

// `maxiter` log-domain updates; both updates of an iteration read the pre-update (u, v)
// (src/loss_multilabel.py:215-217).  With C >= 0 the shifts (v_max, u_max) bound every exponent by 0,
// so no running maximum is needed.  uv_hist (optional) receives (u_t | v_t) for t = 0..maxiter.
__device__ void emd_iterate(const EmdSmem& s, int n, int m, float reg, int maxiter, float* uv_hist) {
  const float k2 = LOG2E / reg;
  for (int it = 0; it < maxiter; ++it) {
    if (uv_hist) {
      float* h = uv_hist + static_cast<size_t>(it) * (n + m);
      for (int i = threadIdx.x; i < n; i += blockDim.x) h[i] = s.u[i];
      for (int j = threadIdx.x; j < m; j += blockDim.x) h[n + j] = s.v[j];
    }
    float mu_ = -FLT_MAX, mv_ = -FLT_MAX;
    for (int i = threadIdx.x; i < n; i += blockDim.x) mu_ = fmaxf(mu_, s.u[i]);
    for (int j = threadIdx.x; j < m; j += blockDim.x) mv_ = fmaxf(mv_, s.v[j]);
    const float umax = block_max(mu_, s.red);
    const float vmax = block_max(mv_, s.red);
    constexpr int RPT = EMD_MAX_PIX / EMD_THREADS;   // rows (columns) per thread; blockDim.x == EMD_THREADS
    float unew[RPT], vnew[RPT];
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int i = threadIdx.x + q * EMD_THREADS;
      if (i < n) {
        float xi[ECP];
        load_row(xi, s.xs + i * ECP);
        float acc = 0.f;
        for (int j = 0; j < m; ++j) {
          const float cst = 1.f - dot_row(xi, s.ys + j * ECP);
          acc += exp2f((s.v[j] - vmax - cst) * k2);
        }
        unew[q] = reg * (s.lmu[i] - logf(acc)) - vmax;
      }
    }
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int j = threadIdx.x + q * EMD_THREADS;
      if (j < m) {
        float yj[ECP];
        load_row(yj, s.ys + j * ECP);
        float acc = 0.f;
        for (int i = 0; i < n; ++i) {
          const float cst = 1.f - dot_row(yj, s.xs + i * ECP);
          acc += exp2f((s.u[i] - umax - cst) * k2);
        }
        vnew[q] = reg * (s.lnu[j] - logf(acc)) - umax;
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int i = threadIdx.x + q * EMD_THREADS;
      if (i < n) s.u[i] = unew[q];
      if (i < m) s.v[i] = vnew[q];
    }
    __syncthreads();
  }
  if (uv_hist) {
    float* h = uv_hist + static_cast<size_t>(maxiter) * (n + m);
    for (int i = threadIdx.x; i < n; i += blockDim.x) h[i] = s.u[i];
    for (int j = threadIdx.x; j < m; j += blockDim.x) h[n + j] = s.v[j];
  }
}

// sum_ij exp((-C + u_i + v_j)/reg) C_ij / (n m)     (src/loss_multilabel.py:219-221)
__device__ float emd_value(const EmdSmem& s, int n, int m, float reg) {
  const float k2 = LOG2E / reg;
  float part = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float xi[ECP];
    load_row(xi, s.xs + i * ECP);
    const float ui = s.u[i];
    for (int j = 0; j < m; ++j) {
      const float cst = 1.f - dot_row(xi, s.ys + j * ECP);
      part = fmaf(exp2f((ui + s.v[j] - cst) * k2), cst, part);
    }
  }
  return block_sum(part, s.red) / (static_cast<float>(n) * static_cast<float>(m));
}

__global__ void __launch_bounds__(EMD_THREADS) emd_score_kernel(const float* flat1, const float* flat2,
                                                                const PairDesc* pairs, float reg, int maxiter,
                                                                float* scores) {
  extern __shared__ float sm[];
  const PairDesc d = pairs[blockIdx.x];
  const int n = static_cast<int>(d.n), m = static_cast<int>(d.m);
  const EmdSmem s = emd_carve(sm, n, m);
  emd_load(flat1, flat2, d, s);
  emd_iterate(s, n, m, reg, maxiter, nullptr);
  const float val = emd_value(s, n, m, reg);
  if (threadIdx.x == 0) scores[blockIdx.x] = val;
}

// best[i] = first pair of image i with the smallest score (the reference sorts a dict keyed by the
// score, stable -> first inserted wins; :318);  loss = mean_i score[best[i]]
__global__ void __launch_bounds__(256) emd_select_kernel(const float* scores, const PairDesc* pairs, int n_pairs,
                                                         int n_images, int* best, float* loss_out) {
  __shared__ float tot;
  if (threadIdx.x == 0) tot = 0.f;
  __syncthreads();
  for (int img = threadIdx.x; img < n_images; img += blockDim.x) {
    int arg = -1;
    float mn = FLT_MAX;
    for (int p = 0; p < n_pairs; ++p)
      if (pairs[p].image == img && (arg < 0 || scores[p] < mn)) { mn = scores[p]; arg = p; }
    best[img] = arg;
    if (arg >= 0) atomicAdd(&tot, mn);
  }
  __syncthreads();
  if (threadIdx.x == 0) *loss_out = tot / static_cast<float>(n_images);
}


// ---------------------------------------------------------------------------------------------
// EMD.static_matching (src/loss_multilabel.py:259-285).  For pair i both regions are resized
// (bilinear, align_corners=True) to at most 7x7 and the value is the MEAN of 1 - x_p . y_q over all
// position pairs and the batch dim; a mean of a bilinear form factorises:
//   dist_i = 1 - (1/n) sum_{b,c} xbar[b,c] ybar[b,c],  xbar = mean over the resized grid,
// and the mean over the resized grid is a separable weighted mean of the SOURCE pixels
// (weight = wr[row] wc[col] / (oh ow), wr/wc = column sums of the interpolation matrices).  So the
// resized maps and the [M,N,C] broadcast product of the reference are never formed: one read of
// both regions, two scalars per plane.
struct StaticDesc {   // int64[12] rows from the host
  long long off1, planes, h1, w1, oh1, ow1, off2, h2, w2, oh2, ow2, batch;
};
constexpr int ST_MAX_SIDE = 1024;
constexpr int ST_THREADS = 256;

// wts[r] = sum over output index i of the bilinear weight of source index r (align_corners=True:
// src = i (in-1)/(out-1), out == 1 reads index 0), times `norm`.
__device__ void static_axis_weights(float* wts, int in, int out, float norm) {
  const float scale = out > 1 ? static_cast<float>(in - 1) / static_cast<float>(out - 1) : 0.f;
  for (int r = threadIdx.x; r < in; r += blockDim.x) {
    float acc = 0.f;
    for (int i = 0; i < out; ++i) {
      const float src = scale * static_cast<float>(i);
      const int i0 = min(static_cast<int>(src), in - 1);
      const int i1 = i0 + (i0 < in - 1 ? 1 : 0);
      const float l1 = src - static_cast<float>(i0);
      acc += (i0 == r ? 1.f - l1 : 0.f) + (i1 == r ? l1 : 0.f);
    }
    wts[r] = acc * norm;
  }
}

// grid (n_pairs, 2 sides): means[(pair*2 + side) * max_planes + plane].
__global__ void __launch_bounds__(ST_THREADS) emd_static_mean_kernel(const float* flat1, const float* flat2,
                                                                     const StaticDesc* desc, int max_planes,
                                                                     float* means) {
  __shared__ float wr[ST_MAX_SIDE], wc[ST_MAX_SIDE];
  const StaticDesc d = desc[blockIdx.x];
  const int side = blockIdx.y;
  const int h = static_cast<int>(side ? d.h2 : d.h1), w = static_cast<int>(side ? d.w2 : d.w1);
  const int oh = static_cast<int>(side ? d.oh2 : d.oh1), ow = static_cast<int>(side ? d.ow2 : d.ow1);
  const float* src = side ? flat2 + d.off2 : flat1 + d.off1;
  float* out = means + (static_cast<size_t>(blockIdx.x) * 2 + side) * max_planes;
  if (h < 1 || w < 1 || h > ST_MAX_SIDE || w > ST_MAX_SIDE || oh < 1 || ow < 1) {   // descriptor out of range: poison, loudly
    for (int p = threadIdx.x; p < d.planes; p += blockDim.x) out[p] = __int_as_float(0x7fc00000);
    return;
  }
  static_axis_weights(wr, h, oh, 1.f / static_cast<float>(oh));
  static_axis_weights(wc, w, ow, 1.f / static_cast<float>(ow));
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, hw = h * w;
  for (int p = warp; p < d.planes; p += ST_THREADS / 32) {
    const float* x = src + static_cast<size_t>(p) * hw;
    float acc = 0.f;
    for (int i = lane; i < hw; i += 32) {
      const int r = i / w;
      acc = fmaf(x[i], wr[r] * wc[i - r * w], acc);
    }
    acc = warp_sum_f(acc);
    if (lane == 0) out[p] = acc;
  }
}

// One block: dist per pair (a warp each), then the two smallest in list.sort() order (stable: the
// first of equal values wins), loss = their sum (:283-284).
__global__ void __launch_bounds__(ST_THREADS) emd_static_select_kernel(const StaticDesc* desc, const float* means,
                                                                       int n_pairs, int max_planes, float* dists,
                                                                       int* sel, float* loss_out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int p = warp; p < n_pairs; p += ST_THREADS / 32) {
    const float* xb = means + static_cast<size_t>(p) * 2 * max_planes;
    const float* yb = xb + max_planes;
    const int planes = static_cast<int>(desc[p].planes);
    float acc = 0.f;
    for (int i = lane; i < planes; i += 32) acc = fmaf(xb[i], yb[i], acc);
    acc = warp_sum_f(acc);
    if (lane == 0) dists[p] = 1.f - acc / static_cast<float>(desc[p].batch);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int a = 0;
    for (int p = 1; p < n_pairs; ++p)
      if (dists[p] < dists[a]) a = p;
    int b = -1;
    for (int p = 0; p < n_pairs; ++p)
      if (p != a && (b < 0 || dists[p] < dists[b])) b = p;
    sel[0] = a;
    sel[1] = b;
    *loss_out = dists[a] + dists[b];
  }
}

// grid (plane chunks, 2 kept pairs): d dist / d x[b,c,r,col] = -(1/n) ybar[b,c] wr[r] wc[col].
__global__ void __launch_bounds__(ST_THREADS) emd_static_grad_kernel(const StaticDesc* desc, const float* means,
                                                                     const int* sel, int max_planes, float* dflat1) {
  __shared__ float wr[ST_MAX_SIDE], wc[ST_MAX_SIDE];
  const int pair = sel[blockIdx.y];
  const StaticDesc d = desc[pair];
  const int h = static_cast<int>(d.h1), w = static_cast<int>(d.w1), hw = h * w;
  if (h < 1 || w < 1 || h > ST_MAX_SIDE || w > ST_MAX_SIDE || d.oh1 < 1 || d.ow1 < 1) return;
  static_axis_weights(wr, h, static_cast<int>(d.oh1), 1.f / static_cast<float>(d.oh1));
  static_axis_weights(wc, w, static_cast<int>(d.ow1), 1.f / static_cast<float>(d.ow1));
  __syncthreads();
  const float* yb = means + (static_cast<size_t>(pair) * 2 + 1) * max_planes;
  const float inv_n = -1.f / static_cast<float>(d.batch);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int p = blockIdx.x * (ST_THREADS / 32) + warp; p < d.planes; p += gridDim.x * (ST_THREADS / 32)) {
    const float coef = yb[p] * inv_n;
    float* g = dflat1 + d.off1 + static_cast<size_t>(p) * hw;
    for (int i = lane; i < hw; i += 32) {
      const int r = i / w;
      g[i] = coef * wr[r] * wc[i - r * w];
    }
  }
}

}  // namespace

namespace {

// Gradient of the selected pair's Sinkhorn value w.r.t. region 1, through the unrolled iterations.
// With R^t_ij = softmax_j(M^t)_ij = exp((-C_ij + u_{t+1,i} + v_{t,j})/reg) / mu_i and
//      Q^t_ij = softmax_i(M^t)_ij = exp((-C_ij + u_{t,i} + v_{t+1,j})/reg) / nu_j :
//   dC_ij  = -dM^T_ij/reg + sum_t ( du_{t+1,i} R^t_ij + dv_{t+1,j} Q^t_ij ),  dM^T = exp(M^T) C / (n m)
//   du_t,i = -sum_j dv_{t+1,j} Q^t_ij ,  dv_t,j = -sum_i du_{t+1,i} R^t_ij
//   du_T,i = sum_j dM^T_ij / reg ,       dv_T,j = sum_i dM^T_ij / reg
//   dx_i   = -sum_j dC_ij y_j            (C = 1 - x y^T; the weights and region 2 are constants)
__global__ void __launch_bounds__(EMD_THREADS) emd_grad_kernel(const float* flat1, const float* flat2,
                                                               const PairDesc* pairs, const int* best, float reg,
                                                               int maxiter, float scale, float* uv_scratch,
                                                               size_t uv_stride, float* dflat1) {
  extern __shared__ float sm[];
  const int arg = best[blockIdx.x];
  if (arg < 0) return;
  const PairDesc d = pairs[arg];
  const int n = static_cast<int>(d.n), m = static_cast<int>(d.m);
  const EmdSmem s = emd_carve(sm, n, m);
  float* du = s.mean + 2 * ECP;   // [n]   du_{t+1}
  float* dv = du + n;             // [m]   dv_{t+1}
  float* v1 = dv + m;             // [m]   v_{t+1}
  float* hist = uv_scratch + blockIdx.x * uv_stride;
  emd_load(flat1, flat2, d, s);
  emd_iterate(s, n, m, reg, maxiter, hist);
  __syncthreads();
  const float k2 = LOG2E / reg;
  const float inv_nm = 1.f / (static_cast<float>(n) * static_cast<float>(m));
  constexpr int RPT = EMD_MAX_PIX / EMD_THREADS;
  float gx[RPT][ECP];
#pragma unroll
  for (int q = 0; q < RPT; ++q)
#pragma unroll
    for (int c = 0; c < ECP; ++c) gx[q][c] = 0.f;

  // ---- final value: s.u = u_T, s.v = v_T
  {
    float du_new[RPT], dv_new[RPT];
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int i = threadIdx.x + q * EMD_THREADS;
      du_new[q] = 0.f;
      if (i < n) {
        float xi[ECP];
        load_row(xi, s.xs + i * ECP);
        const float ui = s.u[i];
        float acc = 0.f;
        for (int j = 0; j < m; ++j) {
          const float* yj = s.ys + j * ECP;
          const float cst = 1.f - dot_row(xi, yj);
          const float dm = exp2f((ui + s.v[j] - cst) * k2) * cst * inv_nm / reg;   // dM^T / reg
          acc += dm;
#pragma unroll
          for (int c = 0; c < EC; ++c) gx[q][c] = fmaf(dm, yj[c], gx[q][c]);       // -dC y = +dM/reg y
        }
        du_new[q] = acc;
      }
    }
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int j = threadIdx.x + q * EMD_THREADS;
      dv_new[q] = 0.f;
      if (j < m) {
        float yj[ECP];
        load_row(yj, s.ys + j * ECP);
        const float vj = s.v[j];
        float acc = 0.f;
        for (int i = 0; i < n; ++i) {
          const float cst = 1.f - dot_row(yj, s.xs + i * ECP);
          acc += exp2f((s.u[i] + vj - cst) * k2) * cst * inv_nm / reg;
        }
        dv_new[q] = acc;
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int i = threadIdx.x + q * EMD_THREADS;
      if (i < n) du[i] = du_new[q];
      if (i < m) { dv[i] = dv_new[q]; v1[i] = s.v[i]; }
    }
    __syncthreads();
  }
  // ---- iterations T-1 .. 0: s.u holds u_{t+1}, v1 holds v_{t+1}, s.v gets v_t
  for (int t = maxiter - 1; t >= 0; --t) {
    const float* ht = hist + static_cast<size_t>(t) * (n + m);   // (u_t | v_t)
    for (int j = threadIdx.x; j < m; j += blockDim.x) s.v[j] = ht[n + j];
    __syncthreads();
    float du_new[RPT], dv_new[RPT], u_t[RPT];
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int i = threadIdx.x + q * EMD_THREADS;
      du_new[q] = 0.f;
      u_t[q] = 0.f;
      if (i < n) {
        float xi[ECP];
        load_row(xi, s.xs + i * ECP);
        const float u1 = s.u[i], u0 = ht[i], dui = du[i], lm = s.lmu[i] * LOG2E;
        u_t[q] = u0;
        float acc = 0.f;
        for (int j = 0; j < m; ++j) {
          const float* yj = s.ys + j * ECP;
          const float cst = 1.f - dot_row(xi, yj);
          const float R = exp2f((u1 + s.v[j] - cst) * k2 - lm);
          const float Q = exp2f((u0 + v1[j] - cst) * k2 - s.lnu[j] * LOG2E);
          const float dvQ = dv[j] * Q;
          acc += dvQ;
          const float dc = dui * R + dvQ;
#pragma unroll
          for (int c = 0; c < EC; ++c) gx[q][c] = fmaf(-dc, yj[c], gx[q][c]);
        }
        du_new[q] = -acc;
      }
    }
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int j = threadIdx.x + q * EMD_THREADS;
      dv_new[q] = 0.f;
      if (j < m) {
        float yj[ECP];
        load_row(yj, s.ys + j * ECP);
        const float v0 = s.v[j];
        float acc = 0.f;
        for (int i = 0; i < n; ++i) {
          const float cst = 1.f - dot_row(yj, s.xs + i * ECP);
          acc = fmaf(du[i], exp2f((s.u[i] + v0 - cst) * k2 - s.lmu[i] * LOG2E), acc);
        }
        dv_new[q] = -acc;
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int i = threadIdx.x + q * EMD_THREADS;
      if (i < n) { du[i] = du_new[q]; s.u[i] = u_t[q]; }
      if (i < m) { dv[i] = dv_new[q]; v1[i] = s.v[i]; }
    }
    __syncthreads();
  }
#pragma unroll
  for (int q = 0; q < RPT; ++q) {
    const int i = threadIdx.x + q * EMD_THREADS;
    if (i < n)
#pragma unroll
      for (int c = 0; c < EC; ++c) dflat1[d.off1 + static_cast<long long>(c) * n + i] = scale * gx[q][c];
  }
}

size_t grad_smem_bytes(long long n, long long m) { return (emd_smem_floats(n, m) + n + 2 * m) * sizeof(float); }

}  // namespace

extern "C" {

int muscle_crops_fwd(const float* x, int channels, int height, int width, const int32_t* desc,
                     const int64_t* offsets, int n_crops, int64_t total, float* flat_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!x || !desc || !offsets || !flat_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n_crops <= 0 || total <= 0) return fail(MUSCLE_ERR_BAD_ARG, "no crops");
  crops_kernel<false><<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      x, nullptr, channels, height, width, reinterpret_cast<const CropDesc*>(desc),
      reinterpret_cast<const long long*>(offsets), n_crops, total, flat_out, nullptr);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_crops_bwd(const float* dflat, int batch, int channels, int height, int width, const int32_t* desc,
                     const int64_t* offsets, int n_crops, int64_t total, float* dx_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!dflat || !desc || !offsets || !dx_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n_crops <= 0 || total <= 0 || batch <= 0) return fail(MUSCLE_ERR_BAD_ARG, "no crops");
  MB_CUDA(cudaMemsetAsync(dx_out, 0, static_cast<size_t>(batch) * channels * height * width * sizeof(float), stream));
  crops_kernel<true><<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      nullptr, dx_out, channels, height, width, reinterpret_cast<const CropDesc*>(desc),
      reinterpret_cast<const long long*>(offsets), n_crops, total, nullptr, dflat);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

size_t muscle_emd_scratch_bytes(int n_pairs, int n_images, int64_t max_n_plus_m, int maxiter) {
  if (n_pairs <= 0 || n_images <= 0 || max_n_plus_m <= 0 || maxiter < 0) return 0;
  const size_t scores = (static_cast<size_t>(n_pairs) * 4 + 255) / 256 * 256;
  const size_t best = (static_cast<size_t>(n_images) * 4 + 255) / 256 * 256;
  const size_t hist = static_cast<size_t>(n_images) * (maxiter + 1) * max_n_plus_m * 4;
  return scores + best + hist + 256;
}

int muscle_emd_dynamic_fwd_bwd(const float* flat1, const float* flat2, const int64_t* pair_desc, int n_pairs,
                               int n_images, int channels, int64_t max_n, int64_t max_m, int64_t max_n_plus_m,
                               float reg, int maxiter, float* loss_out, float* dflat1_out, int64_t total1,
                               void* scratch, size_t scratch_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!flat1 || !flat2 || !pair_desc || !loss_out || !scratch) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n_pairs <= 0 || n_images <= 0) return fail(MUSCLE_ERR_BAD_ARG, "no region pairs");
  if (channels != EC) return fail(MUSCLE_ERR_BAD_ARG, "regions must have %d channels (got %d)", EC, channels);
  if (max_n > EMD_MAX_PIX || max_m > EMD_MAX_PIX || max_n <= 0 || max_m <= 0)
    return fail(MUSCLE_ERR_BAD_ARG, "regions must have 1..%d pixels", EMD_MAX_PIX);
  if (!(reg > 0.f) || maxiter < 0 || maxiter > 64) return fail(MUSCLE_ERR_BAD_ARG, "bad Sinkhorn parameters");
  if (scratch_bytes < muscle_emd_scratch_bytes(n_pairs, n_images, max_n_plus_m, maxiter))
    return fail(MUSCLE_ERR_WORKSPACE, "scratch too small");
  char* base = static_cast<char*>(scratch);
  float* scores = reinterpret_cast<float*>(base);
  size_t off = (static_cast<size_t>(n_pairs) * 4 + 255) / 256 * 256;
  int* best = reinterpret_cast<int*>(base + off);
  off += (static_cast<size_t>(n_images) * 4 + 255) / 256 * 256;
  float* hist = reinterpret_cast<float*>(base + off);
  const PairDesc* pairs = reinterpret_cast<const PairDesc*>(pair_desc);
  const size_t smem_s = emd_smem_floats(max_n, max_m) * sizeof(float);
  const size_t smem_g = grad_smem_bytes(max_n, max_m);
  if (smem_g > 227 * 1024) return fail(MUSCLE_ERR_BAD_ARG, "region pair too large for one SM (%zu bytes)", smem_g);
  MB_CUDA(cudaFuncSetAttribute(emd_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_s)));
  emd_score_kernel<<<n_pairs, EMD_THREADS, smem_s, stream>>>(flat1, flat2, pairs, reg, maxiter, scores);
  MB_LAUNCH_CHECK();
  emd_select_kernel<<<1, 256, 0, stream>>>(scores, pairs, n_pairs, n_images, best, loss_out);
  MB_LAUNCH_CHECK();
  if (dflat1_out) {
    MB_CUDA(cudaMemsetAsync(dflat1_out, 0, static_cast<size_t>(total1) * sizeof(float), stream));
    MB_CUDA(cudaFuncSetAttribute(emd_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_g)));
    emd_grad_kernel<<<n_images, EMD_THREADS, smem_g, stream>>>(flat1, flat2, pairs, best, reg, maxiter,
                                                               1.f / static_cast<float>(n_images), hist,
                                                               static_cast<size_t>(maxiter + 1) * max_n_plus_m, dflat1_out);
    MB_LAUNCH_CHECK();
  }
  return MUSCLE_OK;
}

size_t muscle_emd_static_scratch_bytes(int n_pairs, int max_planes) {
  if (n_pairs <= 0 || max_planes <= 0) return 0;
  const size_t means = (static_cast<size_t>(n_pairs) * 2 * max_planes * 4 + 255) / 256 * 256;
  const size_t dists = (static_cast<size_t>(n_pairs) * 4 + 255) / 256 * 256;
  return means + dists + 256;
}

int muscle_emd_static_fwd_bwd(const float* flat1, const float* flat2, const int64_t* pair_desc, int n_pairs,
                              int max_planes, float* loss_out, float* dists_out, float* dflat1_out, int64_t total1,
                              void* scratch, size_t scratch_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!flat1 || !flat2 || !pair_desc || !loss_out || !scratch) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n_pairs < 2) return fail(MUSCLE_ERR_BAD_ARG, "static matching sums the two smallest distances: needs >= 2 pairs (got %d)", n_pairs);
  if (max_planes <= 0) return fail(MUSCLE_ERR_BAD_ARG, "regions must have batch*channels >= 1 planes");
  if (scratch_bytes < muscle_emd_static_scratch_bytes(n_pairs, max_planes))
    return fail(MUSCLE_ERR_WORKSPACE, "scratch too small");
  char* base = static_cast<char*>(scratch);
  float* means = reinterpret_cast<float*>(base);
  size_t off = (static_cast<size_t>(n_pairs) * 2 * max_planes * 4 + 255) / 256 * 256;
  float* dists = reinterpret_cast<float*>(base + off);
  off += (static_cast<size_t>(n_pairs) * 4 + 255) / 256 * 256;
  int* sel = reinterpret_cast<int*>(base + off);
  const StaticDesc* desc = reinterpret_cast<const StaticDesc*>(pair_desc);
  emd_static_mean_kernel<<<dim3(n_pairs, 2), ST_THREADS, 0, stream>>>(flat1, flat2, desc, max_planes, means);
  MB_LAUNCH_CHECK();
  emd_static_select_kernel<<<1, ST_THREADS, 0, stream>>>(desc, means, n_pairs, max_planes, dists, sel, loss_out);
  MB_LAUNCH_CHECK();
  if (dists_out)
    MB_CUDA(cudaMemcpyAsync(dists_out, dists, static_cast<size_t>(n_pairs) * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  if (dflat1_out) {
    MB_CUDA(cudaMemsetAsync(dflat1_out, 0, static_cast<size_t>(total1) * sizeof(float), stream));
    const int chunks = std::min(16, (max_planes + ST_THREADS / 32 - 1) / (ST_THREADS / 32));
    emd_static_grad_kernel<<<dim3(chunks, 2), ST_THREADS, 0, stream>>>(desc, means, sel, max_planes, dflat1_out);
    MB_LAUNCH_CHECK();
  }
  return MUSCLE_OK;
}

}  // extern "C"
