// Launch-bound / HBM-bound members of the contrastive-loss path (SURVEY.md §8a rows a2, a3, a9,
// a10, a18): image-level contrast, PixPro, the CAM head, the image embedding + logits and the
// trainers' CAM normalisers.  None of these is a dense contraction worth the tensor pipe (B x B,
// 21 channels, or one pass over a feature map); the goal is one launch, no host synchronisation,
// coalesced 128-bit traffic.
#include <algorithm>
#include <cfloat>
#include <type_traits>

#include "host_util.cuh"

using namespace mb200;

namespace {

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
__device__ __forceinline__ float warp_min_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum; `red` has >= 32 floats; result valid in every thread
__device__ __forceinline__ float block_sum_f(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum_f(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (threadIdx.x < (blockDim.x + 31) / 32) ? red[threadIdx.x] : 0.f;
  if (warp == 0) {
    t = warp_sum_f(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

int need_sm100() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}

// =============================================================================================
// a2  image_level_contrast  (src/loss_multilabel.py:36-66)
// =============================================================================================
constexpr int IMC_THREADS = 256;

// ehat[i] = emb[i] / max(||emb[i]||, eps);  nrm[i] = ||emb[i]||
__global__ void __launch_bounds__(IMC_THREADS) imc_norm_kernel(const float* emb, int dim, float eps,
                                                               float* ehat, float* nrm) {
  __shared__ float red[32];
  const float* src = emb + static_cast<size_t>(blockIdx.x) * dim;
  float s = 0.f;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) s = fmaf(src[d], src[d], s);
  const float n = sqrtf(block_sum_f(s, red));
  const float inv = 1.f / fmaxf(n, eps);
  for (int d = threadIdx.x; d < dim; d += blockDim.x) ehat[static_cast<size_t>(blockIdx.x) * dim + d] = src[d] * inv;
  if (threadIdx.x == 0) nrm[blockIdx.x] = n;
}

// pair type of images (i, j) as a bit mask: bit 0 = positive (labels identical, :51), bit 1 = negative (labels
// disjoint under `long & long`, :54).  The reference tests the two with independent `if`s, so a pair of
// all-zero label rows is BOTH.  Warp-cooperative over the label width.
__device__ __forceinline__ int imc_pair_type(const float* label, int n_labels, int i, int j, int lane) {
  bool same = true;
  long long inter = 0;
  for (int l = lane; l < n_labels; l += 32) {
    const float a = label[static_cast<size_t>(i) * n_labels + l], b = label[static_cast<size_t>(j) * n_labels + l];
    same = same && (a == b);
    inter += static_cast<long long>(a) & static_cast<long long>(b);
  }
  same = __all_sync(0xffffffffu, same);
  const bool disjoint = __all_sync(0xffffffffu, inter == 0);
  return (same ? 1 : 0) | (disjoint ? 2 : 0);
}

// stat[i] = {S_pos, S_neg, n_pos, n_neg} over partners j > i (:46), sums seeded with 1e-6 (:44-45)
__global__ void __launch_bounds__(IMC_THREADS) imc_rows_kernel(const float* ehat, const float* label, int batch,
                                                               int dim, int n_labels, float inv_tau,
                                                               float4* stat) {
  extern __shared__ float ei[];  // [dim]
  __shared__ float red[4][IMC_THREADS / 32];
  const int i = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) ei[d] = ehat[static_cast<size_t>(i) * dim + d];
  __syncthreads();
  float sp = 0.f, sn = 0.f, np = 0.f, nn = 0.f;
  for (int j = i + 1 + warp; j < batch; j += IMC_THREADS / 32) {
    const int type = imc_pair_type(label, n_labels, i, j, lane);
    if (type == 0) continue;
    float dot = 0.f;
    for (int d = lane; d < dim; d += 32) dot = fmaf(ei[d], ehat[static_cast<size_t>(j) * dim + d], dot);
    dot = warp_sum_f(dot);
    const float s = expf(dot * inv_tau);
    if (type & 1) { sp += s; np += 1.f; }
    if (type & 2) { sn += s; nn += 1.f; }
  }
  if (lane == 0) { red[0][warp] = sp; red[1][warp] = sn; red[2][warp] = np; red[3][warp] = nn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 1e-6f, b = 1e-6f, c = 0.f, d = 0.f;
    for (int w = 0; w < IMC_THREADS / 32; ++w) { a += red[0][w]; b += red[1][w]; c += red[2][w]; d += red[3][w]; }
    stat[i] = make_float4(a, b, c, d);
  }
}

__device__ __forceinline__ bool imc_use(const float4 s) { return s.z > 0.f && s.w > 0.f && s.w > s.z; }  // :57

// demb[i] = d loss / d emb[i];  block 0 also writes the loss and the number of contributing images
__global__ void __launch_bounds__(IMC_THREADS) imc_grad_kernel(const float* ehat, const float* nrm,
                                                               const float* label, const float4* stat,
                                                               int batch, int dim, int n_labels, float inv_tau,
                                                               float eps, float* demb, float* loss_out,
                                                               int* n_used_out) {
  extern __shared__ float sm[];  // ei[dim] | acc[8][dim]
  float* ei = sm;
  float* acc = sm + dim;
  __shared__ float red[32];
  const int i = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = IMC_THREADS / 32;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) ei[d] = ehat[static_cast<size_t>(i) * dim + d];
  for (int d = threadIdx.x; d < NW * dim; d += blockDim.x) acc[d] = 0.f;
  __syncthreads();
  if (demb != nullptr) {
    const float inv_b = 1.f / static_cast<float>(batch);
    for (int j = warp; j < batch; j += NW) {
      if (j == i) continue;
      const int lo = min(i, j), hi = max(i, j);
      const float4 s = stat[lo];           // the pair is counted in the row of the smaller index
      if (!imc_use(s)) continue;
      const int type = imc_pair_type(label, n_labels, lo, hi, lane);
      if (type == 0) continue;
      float dot = 0.f;
      for (int d = lane; d < dim; d += 32) dot = fmaf(ei[d], ehat[static_cast<size_t>(j) * dim + d], dot);
      dot = warp_sum_f(dot);
      const float sim = expf(dot * inv_tau);
      // term = -log(S_pos) + log(S_pos + S_neg)
      const float dterm = ((type & 1) ? (1.f / (s.x + s.y) - 1.f / s.x) : 0.f) + ((type & 2) ? (1.f / (s.x + s.y)) : 0.f);
      const float coef = inv_b * dterm * sim * inv_tau;
      for (int d = lane; d < dim; d += 32)
        acc[warp * dim + d] = fmaf(coef, ehat[static_cast<size_t>(j) * dim + d], acc[warp * dim + d]);
    }
    __syncthreads();
    float gdot = 0.f;
    for (int d = threadIdx.x; d < dim; d += blockDim.x) {
      float g = 0.f;
      for (int w = 0; w < NW; ++w) g += acc[w * dim + d];
      acc[d] = g;  // column d is owned by this thread in both loops
      gdot = fmaf(g, ei[d], gdot);
    }
    gdot = block_sum_f(gdot, red);
    const float n = nrm[i];
    for (int d = threadIdx.x; d < dim; d += blockDim.x) {
      const float g = acc[d];
      demb[static_cast<size_t>(i) * dim + d] = (n >= eps) ? (g - ei[d] * gdot) / n : g / eps;
    }
  }
  if (blockIdx.x == 0) {
    float l = 0.f, u = 0.f;
    for (int k = threadIdx.x; k < batch; k += blockDim.x) {
      const float4 s = stat[k];
      if (imc_use(s)) { l += -logf(s.x / (s.x + s.y)); u += 1.f; }
    }
    l = block_sum_f(l, red);
    u = block_sum_f(u, red);
    if (threadIdx.x == 0) {
      *loss_out = l / static_cast<float>(batch);
      *n_used_out = static_cast<int>(u + 0.5f);
    }
  }
}

// =============================================================================================
// a3  PixPro  (src/loss_multilabel.py:93-105)
// =============================================================================================
constexpr int PIX_THREADS = 256;

// One thread = one pixel of image b's overlap box.  acc += cos / (h*w) ; dfm1 (pre-zeroed) gets
// -1/(B h w) d cos / d a at the box pixels of view 1.
template <int CH>
__global__ void __launch_bounds__(PIX_THREADS) pixpro_kernel(const float* fm1, const float* fm2, const int* box1,
                                                             const int* box2, int H, int W, int batch,
                                                             float eps, double* acc, float* dfm1) {
  __shared__ float red[32];
  const int b = blockIdx.y;
  const int t1 = box1[b * 4 + 0], l1 = box1[b * 4 + 1], h = box1[b * 4 + 2], w = box1[b * 4 + 3];
  const int t2 = box2[b * 4 + 0], l2 = box2[b * 4 + 1];
  const int npx = h * w;
  const size_t plane = static_cast<size_t>(H) * W;
  float part = 0.f;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npx; p += gridDim.x * blockDim.x) {
    const int y = p / w, x = p - y * w;
    const float* a = fm1 + (static_cast<size_t>(b) * CH) * plane + static_cast<size_t>(t1 + y) * W + (l1 + x);
    const float* c = fm2 + (static_cast<size_t>(b) * CH) * plane + static_cast<size_t>(t2 + y) * W + (l2 + x);
    float av[CH], cv[CH];
    float aa = 0.f, cc = 0.f, ac = 0.f;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      av[k] = a[k * plane];
      cv[k] = c[k * plane];
      aa = fmaf(av[k], av[k], aa);
      cc = fmaf(cv[k], cv[k], cc);
      ac = fmaf(av[k], cv[k], ac);
    }
    const float na = sqrtf(aa), nc = sqrtf(cc);
    const float nac = fmaxf(na, eps), ncc = fmaxf(nc, eps);
    const float cosv = ac / (nac * ncc);
    part += cosv;
    if (dfm1 != nullptr) {
      const float go = -1.f / (static_cast<float>(batch) * static_cast<float>(npx));
      const float k1 = go / (nac * ncc);
      const float k2 = (na > eps) ? go * cosv / (nac * nac) : 0.f;  // the clamp has zero slope below eps
      float* d = dfm1 + (static_cast<size_t>(b) * CH) * plane + static_cast<size_t>(t1 + y) * W + (l1 + x);
#pragma unroll
      for (int k = 0; k < CH; ++k) d[k * plane] = k1 * cv[k] - k2 * av[k];
    }
  }
  part = block_sum_f(part, red);
  if (threadIdx.x == 0 && npx > 0) atomicAdd(acc, static_cast<double>(part) / static_cast<double>(npx));
}

__global__ void pixpro_final_kernel(const double* acc, int batch, float* loss) {
  *loss = static_cast<float>(1.0 - *acc / static_cast<double>(batch));
}

// =============================================================================================
// a9  CAM head  relu(sum_ch p7 * W)  (src/MuSCLe.py:242-247),  a10 GAP + logits (:240-241)
// =============================================================================================
constexpr int CAM_PXT = 64;     // pixel-threads per block (2 pixels each -> 128 pixels per block)
constexpr int CAM_SLICES = 8;   // channel slices per block (one warp pair per slice)
constexpr int CAM_THREADS = CAM_PXT * CAM_SLICES;
constexpr int CAM_PX = 2;

// cam[b,k,p] = relu(sum_c W[k,c] p7[b,c,p]).  Thread = (2 consecutive pixels, channel slice): it
// accumulates the 21 class sums over its channels, the slices are then reduced through shared
// memory.  W is transposed in smem as [c][KP] so a channel's weights are broadcast LDS.128 runs.
// Splitting the channels (not only the pixels) over threads keeps every SM busy at feature-map
// sizes where pixels alone give < 1 warp per SM-quarter.
template <int KP>
__global__ void __launch_bounds__(CAM_THREADS) cam_head_fwd_kernel(const float* p7, const float* w, int ch, int P,
                                                                   int K, float* cam) {
  extern __shared__ float wsm[];            // [ch][KP] | red [CAM_SLICES][KP][128]
  float* red = wsm + static_cast<size_t>(ch) * KP;
  const int b = blockIdx.y;
  // W [K][ch] -> wsm [ch][KP], read in its own order (consecutive threads = consecutive channels of one class: coalesced;
  // reading it in wsm's order made every load a 32-way scattered one, ~18 us per block at 640 channels)
  for (int i = threadIdx.x; i < K * ch; i += blockDim.x) {
    const int k = i / ch, c = i - k * ch;
    wsm[c * KP + k] = __ldg(w + i);
  }
  for (int i = threadIdx.x; i < (KP - K) * ch; i += blockDim.x) {
    const int k = K + i / ch, c = i % ch;
    wsm[c * KP + k] = 0.f;
  }
  __syncthreads();
  const int pt = threadIdx.x % CAM_PXT, slice = threadIdx.x / CAM_PXT;
  const int p0 = (blockIdx.x * CAM_PXT + pt) * CAM_PX;
  const bool vec = (P % 2 == 0);
  float acc[CAM_PX][KP];
#pragma unroll
  for (int q = 0; q < CAM_PX; ++q)
#pragma unroll
    for (int k = 0; k < KP; ++k) acc[q][k] = 0.f;
  if (p0 < P) {
    const float* src = p7 + static_cast<size_t>(b) * ch * P;
    const int c_lo = (ch * slice) / CAM_SLICES, c_hi = (ch * (slice + 1)) / CAM_SLICES;
#pragma unroll 8
    for (int c = c_lo; c < c_hi; ++c) {
      float v[CAM_PX];
      if (vec) {
        const float2 t = *reinterpret_cast<const float2*>(src + static_cast<size_t>(c) * P + p0);
        v[0] = t.x; v[1] = t.y;
      } else {
        v[0] = src[static_cast<size_t>(c) * P + p0];
        v[1] = src[static_cast<size_t>(c) * P + min(p0 + 1, P - 1)];
      }
      const float4* wr = reinterpret_cast<const float4*>(wsm + c * KP);
#pragma unroll
      for (int k4 = 0; k4 < KP / 4; ++k4) {
        const float4 ww = wr[k4];
#pragma unroll
        for (int q = 0; q < CAM_PX; ++q) {
          acc[q][k4 * 4 + 0] = fmaf(v[q], ww.x, acc[q][k4 * 4 + 0]);
          acc[q][k4 * 4 + 1] = fmaf(v[q], ww.y, acc[q][k4 * 4 + 1]);
          acc[q][k4 * 4 + 2] = fmaf(v[q], ww.z, acc[q][k4 * 4 + 2]);
          acc[q][k4 * 4 + 3] = fmaf(v[q], ww.w, acc[q][k4 * 4 + 3]);
        }
      }
    }
  }
  // red[slice][k][pixel-in-block]
#pragma unroll
  for (int k = 0; k < KP; ++k)
#pragma unroll
    for (int q = 0; q < CAM_PX; ++q) red[(slice * KP + k) * (CAM_PXT * CAM_PX) + pt * CAM_PX + q] = acc[q][k];
  __syncthreads();
  float* dst = cam + static_cast<size_t>(b) * K * P;
  const int pb = blockIdx.x * CAM_PXT * CAM_PX;
  for (int i = threadIdx.x; i < K * CAM_PXT * CAM_PX; i += blockDim.x) {
    const int k = i / (CAM_PXT * CAM_PX), px = i - k * (CAM_PXT * CAM_PX);
    if (pb + px >= P) continue;
    float sum = 0.f;
#pragma unroll
    for (int sl = 0; sl < CAM_SLICES; ++sl) sum += red[(sl * KP + k) * (CAM_PXT * CAM_PX) + px];
    dst[static_cast<size_t>(k) * P + pb + px] = fmaxf(sum, 0.f);
  }
}

// dp7[b,c,p] = sum_k W[k,c] dcam[b,k,p] [cam[b,k,p] > 0]  (+ demb[b,c] / P when demb != null:
// the GAP backward, so p7's gradient is written once).  No gradient reaches W (fc.weight.data).
// Thread = (2 consecutive pixels, channel slice): the gated CAM gradients of its pixels stay in
// registers, it writes its slice of the channels.
template <int KP>
__global__ void __launch_bounds__(CAM_THREADS) cam_head_bwd_kernel(const float* dcam, const float* cam,
                                                                   const float* w, const float* demb, int ch,
                                                                   int P, int K, float* dp7) {
  extern __shared__ float wsm[];  // [ch][KP]
  const int b = blockIdx.y;
  // W [K][ch] -> wsm [ch][KP], read in its own order (consecutive threads = consecutive channels of one class: coalesced;
  // reading it in wsm's order made every load a 32-way scattered one, ~18 us per block at 640 channels)
  for (int i = threadIdx.x; i < K * ch; i += blockDim.x) {
    const int k = i / ch, c = i - k * ch;
    wsm[c * KP + k] = __ldg(w + i);
  }
  for (int i = threadIdx.x; i < (KP - K) * ch; i += blockDim.x) {
    const int k = K + i / ch, c = i % ch;
    wsm[c * KP + k] = 0.f;
  }
  __syncthreads();
  const int pt = threadIdx.x % CAM_PXT, slice = threadIdx.x / CAM_PXT;
  const int p0 = (blockIdx.x * CAM_PXT + pt) * CAM_PX;
  if (p0 >= P) return;
  const bool vec = (P % 2 == 0);
  float g[CAM_PX][KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) {
#pragma unroll
    for (int q = 0; q < CAM_PX; ++q) {
      float v = 0.f;
      if (k < K && (dcam != nullptr)) {
        const size_t idx = (static_cast<size_t>(b) * K + k) * P + min(p0 + q, P - 1);
        v = (cam[idx] > 0.f) ? dcam[idx] : 0.f;
      }
      g[q][k] = v;
    }
  }
  const float inv_p = 1.f / static_cast<float>(P);
  float* dst = dp7 + static_cast<size_t>(b) * ch * P;
  const int c_lo = (ch * slice) / CAM_SLICES, c_hi = (ch * (slice + 1)) / CAM_SLICES;
#pragma unroll 2
  for (int c = c_lo; c < c_hi; ++c) {
    const float base = demb ? demb[static_cast<size_t>(b) * ch + c] * inv_p : 0.f;
    float o[CAM_PX] = {base, base};
    const float4* wr = reinterpret_cast<const float4*>(wsm + c * KP);
#pragma unroll
    for (int k4 = 0; k4 < KP / 4; ++k4) {
      const float4 ww = wr[k4];
#pragma unroll
      for (int q = 0; q < CAM_PX; ++q) {
        o[q] = fmaf(g[q][k4 * 4 + 0], ww.x, o[q]);
        o[q] = fmaf(g[q][k4 * 4 + 1], ww.y, o[q]);
        o[q] = fmaf(g[q][k4 * 4 + 2], ww.z, o[q]);
        o[q] = fmaf(g[q][k4 * 4 + 3], ww.w, o[q]);
      }
    }
    if (vec) {
      *reinterpret_cast<float2*>(dst + static_cast<size_t>(c) * P + p0) = make_float2(o[0], o[1]);
    } else {
      dst[static_cast<size_t>(c) * P + p0] = o[0];
      if (p0 + 1 < P) dst[static_cast<size_t>(c) * P + p0 + 1] = o[1];
    }
  }
}

// emb[b,c] = mean_p p7[b,c,p] : one warp per (b,c) row
__global__ void __launch_bounds__(256) gap_kernel(const float* p7, long long rows, int P, float* emb) {
  const long long r = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* src = p7 + r * P;
  float s = 0.f;
  if (P % 4 == 0) {
    for (int p = lane * 4; p < P; p += 128) {
      const float4 t = *reinterpret_cast<const float4*>(src + p);
      s += (t.x + t.y) + (t.z + t.w);
    }
  } else {
    for (int p = lane; p < P; p += 32) s += src[p];
  }
  s = warp_sum_f(s);
  if (lane == 0) emb[r] = s / static_cast<float>(P);
}

// logits[b,k] = sum_c emb[b,c] W[k,c] : one warp per (b,k)
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float* emb, const float* w, int batch, int ch, int K,
                                                         float* logits) {
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= batch * K) return;
  const int b = r / K, k = r - b * K, lane = threadIdx.x & 31;
  float s = 0.f;
  for (int c = lane; c < ch; c += 32) s = fmaf(emb[static_cast<size_t>(b) * ch + c], w[static_cast<size_t>(k) * ch + c], s);
  s = warp_sum_f(s);
  if (lane == 0) logits[r] = s;
}
// demb[b,c] = sum_k dlogits[b,k] W[k,c] ;  dW[k,c] = sum_b dlogits[b,k] emb[b,c]
__global__ void __launch_bounds__(256) linear_bwd_kernel(const float* dlogits, const float* emb, const float* w,
                                                         int batch, int ch, int K, float* demb, float* dw) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < batch * ch) {
    const int b = idx / ch, c = idx - b * ch;
    float s = 0.f;
    for (int k = 0; k < K; ++k) s = fmaf(dlogits[b * K + k], w[static_cast<size_t>(k) * ch + c], s);
    demb[idx] = s;
  }
  if (idx < K * ch) {
    const int k = idx / ch, c = idx - k * ch;
    float s = 0.f;
    for (int b = 0; b < batch; ++b) s = fmaf(dlogits[b * K + k], emb[static_cast<size_t>(b) * ch + c], s);
    dw[idx] = s;
  }
}

// =============================================================================================
// a18  cam_softmaxnorm (train_mcl.py:30-36) and cam_maxnorm (train_mcl.py:21-28)
// =============================================================================================
constexpr int NORM_MAXC = 32;

// out[:,0] = 1 - max_k softmax_fg ; out[:,1:] = softmax over the fg channels.  Thread = pixel.
__global__ void __launch_bounds__(256) cam_softmaxnorm_fwd_kernel(const float* cam, int C, long long P, float* out) {
  const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= P) return;
  const float* src = cam + static_cast<size_t>(b) * C * P + p;
  float v[NORM_MAXC];
  float mx = -FLT_MAX;
#pragma unroll
  for (int k = 1; k < NORM_MAXC; ++k)
    if (k < C) { v[k] = src[k * P]; mx = fmaxf(mx, v[k]); }
  float s = 0.f;
#pragma unroll
  for (int k = 1; k < NORM_MAXC; ++k)
    if (k < C) { v[k] = expf(v[k] - mx); s += v[k]; }
  const float inv = 1.f / s;
  float top = 0.f;
  float* dst = out + static_cast<size_t>(b) * C * P + p;
#pragma unroll
  for (int k = 1; k < NORM_MAXC; ++k)
    if (k < C) { const float y = v[k] * inv; dst[k * P] = y; top = fmaxf(top, y); }
  dst[0] = 1.f - top;
}
// din[:,0] = 0 ; din[:,k] = y_k (g_k - sum_j g_j y_j) with g = dout[:,1:] - dout[:,0] at the (first) arg-max
__global__ void __launch_bounds__(256) cam_softmaxnorm_bwd_kernel(const float* out, const float* dout, int C,
                                                                  long long P, float* din) {
  const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= P) return;
  const size_t base = static_cast<size_t>(b) * C * P + p;
  float y[NORM_MAXC], g[NORM_MAXC];
  int arg = 1;
  float top = -1.f;
#pragma unroll
  for (int k = 1; k < NORM_MAXC; ++k)
    if (k < C) {
      y[k] = out[base + k * P];
      g[k] = dout[base + k * P];
      if (y[k] > top) { top = y[k]; arg = k; }
    }
  const float g0 = dout[base];
  float dot = 0.f;
#pragma unroll
  for (int k = 1; k < NORM_MAXC; ++k)
    if (k < C) {
      if (k == arg) g[k] -= g0;
      dot = fmaf(g[k], y[k], dot);
    }
  din[base] = 0.f;
#pragma unroll
  for (int k = 1; k < NORM_MAXC; ++k)
    if (k < C) din[base + k * P] = y[k] * (g[k] - dot);
}

// cam_maxnorm: y = relu((relu(x) - lo - 1e-6) / (hi - lo + 1e-6)), lo/hi = min/max of relu(x) over
// the plane.  One block per (b, class) plane; also records lo, hi and the first arg-min/arg-max.
__global__ void __launch_bounds__(256) cam_maxnorm_fwd_kernel(const float* cam, long long P, float* out,
                                                              float* lohi, long long* argmm) {
  __shared__ float slo[8], shi[8];
  __shared__ long long salo[8], sahi[8];
  const size_t base = static_cast<size_t>(blockIdx.x) * P;
  float lo = FLT_MAX, hi = -FLT_MAX;
  long long alo = 0, ahi = 0;
  for (long long p = threadIdx.x; p < P; p += blockDim.x) {
    const float v = fmaxf(cam[base + p], 0.f);
    if (v < lo) { lo = v; alo = p; }
    if (v > hi) { hi = v; ahi = p; }
  }
  // warp then block arg-reduction, ties -> smallest index
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float l2 = __shfl_xor_sync(0xffffffffu, lo, o);
    const long long a2 = __shfl_xor_sync(0xffffffffu, alo, o);
    if (l2 < lo || (l2 == lo && a2 < alo)) { lo = l2; alo = a2; }
    const float h2 = __shfl_xor_sync(0xffffffffu, hi, o);
    const long long b2 = __shfl_xor_sync(0xffffffffu, ahi, o);
    if (h2 > hi || (h2 == hi && b2 < ahi)) { hi = h2; ahi = b2; }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { slo[warp] = lo; shi[warp] = hi; salo[warp] = alo; sahi[warp] = ahi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) {
      if (slo[w] < slo[0] || (slo[w] == slo[0] && salo[w] < salo[0])) { slo[0] = slo[w]; salo[0] = salo[w]; }
      if (shi[w] > shi[0] || (shi[w] == shi[0] && sahi[w] < sahi[0])) { shi[0] = shi[w]; sahi[0] = sahi[w]; }
    }
    lohi[blockIdx.x * 2] = slo[0];
    lohi[blockIdx.x * 2 + 1] = shi[0];
    argmm[blockIdx.x * 2] = salo[0];
    argmm[blockIdx.x * 2 + 1] = sahi[0];
  }
  __syncthreads();
  lo = slo[0];
  hi = shi[0];
  const float inv = 1.f / (hi - lo + 1e-6f);
  for (long long p = threadIdx.x; p < P; p += blockDim.x)
    out[base + p] = fmaxf((fmaxf(cam[base + p], 0.f) - lo - 1e-6f) * inv, 0.f);
}
// din = [x>0] ( [y>0] dout / den  + at arg-min: sum_p t_p (u_p/den^2 - 1/den) + at arg-max: -sum_p t_p u_p/den^2 )
// with t = [y>0] dout, u = relu(x) - lo - 1e-6, den = hi - lo + 1e-6
__global__ void __launch_bounds__(256) cam_maxnorm_bwd_kernel(const float* cam, const float* dout, const float* lohi,
                                                              const long long* argmm, long long P, float* din) {
  __shared__ float red[32];
  const size_t base = static_cast<size_t>(blockIdx.x) * P;
  const float lo = lohi[blockIdx.x * 2], hi = lohi[blockIdx.x * 2 + 1];
  const float den = hi - lo + 1e-6f, inv = 1.f / den;
  float s_t = 0.f, s_tu = 0.f;
  for (long long p = threadIdx.x; p < P; p += blockDim.x) {
    const float x = cam[base + p];
    const float u = fmaxf(x, 0.f) - lo - 1e-6f;
    const float t = (u * inv > 0.f) ? dout[base + p] : 0.f;
    din[base + p] = (x > 0.f) ? t * inv : 0.f;
    s_t += t;
    s_tu = fmaf(t, u, s_tu);
  }
  s_t = block_sum_f(s_t, red);
  s_tu = block_sum_f(s_tu, red);
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long alo = argmm[blockIdx.x * 2], ahi = argmm[blockIdx.x * 2 + 1];
    const float dlo = s_tu * inv * inv - s_t * inv;
    const float dhi = -s_tu * inv * inv;
    if (cam[base + alo] > 0.f) din[base + alo] += dlo;
    if (cam[base + ahi] > 0.f) din[base + ahi] += dhi;
  }
}

template <typename F>
int dispatch_kp(int K, F&& f) {
  if (K <= 8) return f(std::integral_constant<int, 8>{});
  if (K <= 16) return f(std::integral_constant<int, 16>{});
  if (K <= 24) return f(std::integral_constant<int, 24>{});
  if (K <= 32) return f(std::integral_constant<int, 32>{});
  return fail(MUSCLE_ERR_BAD_ARG, "at most 32 CAM classes are supported (got %d)", K);
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------------------------------- IMC
size_t muscle_imc_workspace_bytes(int batch, int dim) {
  if (batch <= 0 || dim <= 0) return 0;
  return (static_cast<size_t>(batch) * dim * 4 + 255) / 256 * 256 + (static_cast<size_t>(batch) * 4 + 255) / 256 * 256 +
         static_cast<size_t>(batch) * 16;
}

int muscle_imc_fwd_bwd(const float* emb, const float* label, int batch, int dim, int n_labels, float temperature,
                       float* loss_out, int32_t* n_used_out, float* demb_out, void* workspace,
                       size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!emb || !label || !loss_out || !n_used_out || !workspace) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || dim <= 0 || n_labels <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (dim > 4096) return fail(MUSCLE_ERR_BAD_ARG, "embedding width %d > 4096", dim);
  if (!(temperature > 0.f)) return fail(MUSCLE_ERR_BAD_ARG, "temperature must be > 0");
  if (workspace_bytes < muscle_imc_workspace_bytes(batch, dim)) return fail(MUSCLE_ERR_WORKSPACE, "workspace too small");
  char* base = static_cast<char*>(workspace);
  float* ehat = reinterpret_cast<float*>(base);
  size_t off = (static_cast<size_t>(batch) * dim * 4 + 255) / 256 * 256;
  float* nrm = reinterpret_cast<float*>(base + off);
  off += (static_cast<size_t>(batch) * 4 + 255) / 256 * 256;
  float4* stat = reinterpret_cast<float4*>(base + off);
  const float eps = 1e-6f;  // F.normalize(emb, eps=1e-6), :41
  imc_norm_kernel<<<batch, IMC_THREADS, 0, stream>>>(emb, dim, eps, ehat, nrm);
  MB_LAUNCH_CHECK();
  imc_rows_kernel<<<batch, IMC_THREADS, dim * sizeof(float), stream>>>(ehat, label, batch, dim, n_labels,
                                                                       1.f / temperature, stat);
  MB_LAUNCH_CHECK();
  const size_t smem = static_cast<size_t>(dim) * (1 + IMC_THREADS / 32) * sizeof(float);
  MB_CUDA(cudaFuncSetAttribute(imc_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  imc_grad_kernel<<<batch, IMC_THREADS, smem, stream>>>(ehat, nrm, label, stat, batch, dim, n_labels,
                                                        1.f / temperature, eps, demb_out, loss_out, n_used_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

// ------------------------------------------------------------------------------------- PixPro
int muscle_pixpro_fwd_bwd(const float* fm1, const float* fm2, const int32_t* box1, const int32_t* box2, int batch,
                          int channels, int height, int width, int max_box_pixels, float* loss_out,
                          float* dfm1_out, double* scratch, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!fm1 || !fm2 || !box1 || !box2 || !loss_out || !scratch) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || height <= 0 || width <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (channels != 21) return fail(MUSCLE_ERR_BAD_ARG, "PixPro maps must have 21 channels (got %d)", channels);
  MB_CUDA(cudaMemsetAsync(scratch, 0, sizeof(double), stream));
  if (dfm1_out)
    MB_CUDA(cudaMemsetAsync(dfm1_out, 0, static_cast<size_t>(batch) * channels * height * width * sizeof(float), stream));
  const int gx = std::max(1, std::min((max_box_pixels + PIX_THREADS - 1) / PIX_THREADS, 1024));
  pixpro_kernel<21><<<dim3(gx, batch), PIX_THREADS, 0, stream>>>(fm1, fm2, box1, box2, height, width, batch, 1e-8f,
                                                                scratch, dfm1_out);
  MB_LAUNCH_CHECK();
  pixpro_final_kernel<<<1, 1, 0, stream>>>(scratch, batch, loss_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

// ------------------------------------------------------------------------------- CAM head etc.
int muscle_cam_head_fwd(const float* p7, const float* fc_weight, int batch, int channels, int pixels,
                        int num_classes, float* cam_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!p7 || !fc_weight || !cam_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || pixels <= 0 || num_classes <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if ((reinterpret_cast<uintptr_t>(p7) | reinterpret_cast<uintptr_t>(cam_out)) & 7)
    return fail(MUSCLE_ERR_BAD_ARG, "feature maps must be 8-byte aligned");
  const dim3 grid((pixels + CAM_PXT * CAM_PX - 1) / (CAM_PXT * CAM_PX), batch);
  return dispatch_kp(num_classes, [&](auto kp) -> int {
    constexpr int KP = decltype(kp)::value;
    const size_t smem = (static_cast<size_t>(channels) * KP + static_cast<size_t>(CAM_SLICES) * KP * CAM_PXT * CAM_PX) * sizeof(float);
    if (smem > 200 * 1024) return fail(MUSCLE_ERR_BAD_ARG, "too many channels (%d) for the CAM head", channels);
    MB_CUDA(cudaFuncSetAttribute(cam_head_fwd_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    cam_head_fwd_kernel<KP><<<grid, CAM_THREADS, smem, stream>>>(p7, fc_weight, channels, pixels, num_classes, cam_out);
    MB_LAUNCH_CHECK();
    return MUSCLE_OK;
  });
}

int muscle_cam_head_bwd(const float* dcam, const float* cam, const float* fc_weight, const float* demb, int batch,
                        int channels, int pixels, int num_classes, float* dp7_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!fc_weight || !dp7_out || (dcam && !cam)) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || pixels <= 0 || num_classes <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (reinterpret_cast<uintptr_t>(dp7_out) & 7) return fail(MUSCLE_ERR_BAD_ARG, "feature maps must be 8-byte aligned");
  const dim3 grid((pixels + CAM_PXT * CAM_PX - 1) / (CAM_PXT * CAM_PX), batch);
  return dispatch_kp(num_classes, [&](auto kp) -> int {
    constexpr int KP = decltype(kp)::value;
    const size_t smem = static_cast<size_t>(channels) * KP * sizeof(float);
    if (smem > 200 * 1024) return fail(MUSCLE_ERR_BAD_ARG, "too many channels (%d) for the CAM head", channels);
    MB_CUDA(cudaFuncSetAttribute(cam_head_bwd_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    cam_head_bwd_kernel<KP><<<grid, CAM_THREADS, smem, stream>>>(dcam, cam, fc_weight, demb, channels, pixels,
                                                                 num_classes, dp7_out);
    MB_LAUNCH_CHECK();
    return MUSCLE_OK;
  });
}

int muscle_gap_fwd(const float* p7, int batch, int channels, int pixels, float* emb_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!p7 || !emb_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || pixels <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  const long long rows = static_cast<long long>(batch) * channels;
  gap_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, stream>>>(p7, rows, pixels, emb_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_linear_fwd(const float* emb, const float* weight, int batch, int channels, int num_classes,
                      float* logits_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!emb || !weight || !logits_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || num_classes <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  linear_fwd_kernel<<<(batch * num_classes + 7) / 8, 256, 0, stream>>>(emb, weight, batch, channels, num_classes, logits_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_linear_bwd(const float* dlogits, const float* emb, const float* weight, int batch, int channels,
                      int num_classes, float* demb_out, float* dweight_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!dlogits || !emb || !weight || !demb_out || !dweight_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || num_classes <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  const int n = std::max(batch, num_classes) * channels;
  linear_bwd_kernel<<<(n + 255) / 256, 256, 0, stream>>>(dlogits, emb, weight, batch, channels, num_classes, demb_out, dweight_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

// -------------------------------------------------------------------------------- normalisers
int muscle_cam_softmaxnorm_fwd(const float* cam, int batch, int num_classes, int64_t pixels, float* out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!cam || !out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || pixels <= 0 || num_classes < 2 || num_classes > NORM_MAXC)
    return fail(MUSCLE_ERR_BAD_ARG, "need 2..%d classes", NORM_MAXC);
  cam_softmaxnorm_fwd_kernel<<<dim3(static_cast<unsigned>((pixels + 255) / 256), batch), 256, 0, stream>>>(cam, num_classes, pixels, out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
int muscle_cam_softmaxnorm_bwd(const float* out, const float* dout, int batch, int num_classes, int64_t pixels,
                               float* din, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!out || !dout || !din) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || pixels <= 0 || num_classes < 2 || num_classes > NORM_MAXC)
    return fail(MUSCLE_ERR_BAD_ARG, "need 2..%d classes", NORM_MAXC);
  cam_softmaxnorm_bwd_kernel<<<dim3(static_cast<unsigned>((pixels + 255) / 256), batch), 256, 0, stream>>>(out, dout, num_classes, pixels, din);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
int muscle_cam_maxnorm_fwd(const float* cam, int64_t planes, int64_t pixels, float* out, float* lohi, int64_t* argmm,
                           void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!cam || !out || !lohi || !argmm) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (planes <= 0 || pixels <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  cam_maxnorm_fwd_kernel<<<static_cast<unsigned>(planes), 256, 0, stream>>>(cam, pixels, out, lohi,
                                                                             reinterpret_cast<long long*>(argmm));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
int muscle_cam_maxnorm_bwd(const float* cam, const float* dout, const float* lohi, const int64_t* argmm,
                           int64_t planes, int64_t pixels, float* din, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!cam || !dout || !lohi || !argmm || !din) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (planes <= 0 || pixels <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  cam_maxnorm_bwd_kernel<<<static_cast<unsigned>(planes), 256, 0, stream>>>(cam, dout, lohi,
                                                                             reinterpret_cast<const long long*>(argmm), pixels, din);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // extern "C"
