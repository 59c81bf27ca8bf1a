// FieldLoss — the boundary-aware pseudo-mask pixel-pixel contrast of stage 2 (src/edge.py:175-384,
// SURVEY.md §8a rows a12-a17).
//
//   edges   : softmax(seg*beta) -> 20 fg planes -> Sobel 5x5 (x, y) -> * image label   (:423-440)
//   quantise: mag = sqrt(gx^2+gy^2+1e-8), 8 orientation bins of atan2(gy,gx)           (:65-89)
//   select  : pos = mag >= 0.8 max_plane & max_plane > 1 & label                        (:370-374)
//   in/out  : per boundary pixel an "outside" and an "inside" pixel index              (:195-227)
//   sample  : host draws k of each list with Python's random.sample                    (:297-304)
//   contrast: sim_out = outs^T ins over softmax(dense_ft) columns, sim_mask over softmax(mask);
//             sign tests on row/column means, BEACON terms                              (:318-341, 229-258)
//
// The reference soft-maxes the whole [n,256,H,W] feature map and materialises k x k similarity
// matrices only to take their row, column and global means.  Here the channel softmax runs at the
// <= 2k sampled pixels of each (image, class) only, and the means come from
//   mean_b' sim[a,b'] = outs[a] . (sum_b' ins[b']) / k
// so no k x k matrix and no full-resolution softmax ever exists.  All of it is fp32 (the sign tests
// compare nearly equal means; bf16 would flip them), HBM/latency bound, no tensor-pipe work.
#include <algorithm>
#include <cfloat>

#include "bilinear.cuh"
#include "host_util.cuh"

using namespace mb200;

namespace {

constexpr int NFG = 20;  // foreground classes handled per pixel (num_classes - 1 <= NFG)

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

// the reference's 5x5 "Sobel" pair, including its 1e-6 stand-ins for zero (src/edge.py:36-43)
__constant__ float c_kx[25] = {2.f, 1.f, 1e-6f, -1.f, -2.f, 3.f, 2.f, 1e-6f, -2.f, -3.f, 4.f, 3.f, 0.f, -3.f, -4.f,
                               3.f, 2.f, 1e-6f, -2.f, -3.f, 2.f, 1.f, 1e-6f, -1.f, -2.f};
__constant__ float c_ky[25] = {2.f, 3.f, 4.f, 3.f, 2.f, 1.f, 2.f, 3.f, 2.f, 1.f, 1e-6f, 1e-6f, 1e-6f, 1e-6f, 1e-6f,
                               -1.f, -2.f, -3.f, -2.f, -1.f, -2.f, -3.f, -4.f, -3.f, -2.f};

// ... and the 3x3 pair (src/edge.py:33-35)
__constant__ float c_kx3[9] = {1.f, 1e-6f, -1.f, 2.f, 1e-6f, -2.f, 1.f, 1e-6f, -1.f};
__constant__ float c_ky3[9] = {1.f, 2.f, 1.f, 1e-6f, 1e-6f, 1e-6f, -1.f, -2.f, -1.f};

// prob[b][c][p] = softmax_k(seg[b][k][p] * beta)[1 + c]
__global__ void __launch_bounds__(256) field_prob_kernel(const float* seg, float beta, int C, long long HW, float* prob) {
  const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= HW) return;
  const float* src = seg + static_cast<long long>(b) * C * HW + p;
  float v[NFG + 1];
  float mx = -FLT_MAX;
#pragma unroll
  for (int k = 0; k <= NFG; ++k)
    if (k < C) { v[k] = src[k * HW] * beta; mx = fmaxf(mx, v[k]); }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k <= NFG; ++k)
    if (k < C) { v[k] = expf(v[k] - mx); s += v[k]; }
  float* dst = prob + static_cast<long long>(b) * (C - 1) * HW + p;
#pragma unroll
  for (int k = 1; k <= NFG; ++k)
    if (k < C) dst[(k - 1) * HW] = v[k] / s;
}

// 8-bin quantisation of atan2 with the reference's constant 3.1416/8 (src/edge.py:63,72-88); bins 0..7.
// The reference compares the fp32 angle with PYTHON floats k * (3.1416/8): the product is formed in double and
// then takes the tensor's dtype, so the thresholds are fp32(k * 0.3927), not k * fp32(0.3927).
__device__ __forceinline__ int orient_bin(float gy, float gx) {
  const float ang = atan2f(gy, gx);
  constexpr double d = 3.1416 / 8;
  constexpr float t1 = static_cast<float>(1 * d), t3 = static_cast<float>(3 * d), t5 = static_cast<float>(5 * d),
                  t7 = static_cast<float>(7 * d);
  constexpr float m1 = static_cast<float>(-1 * d), m3 = static_cast<float>(-3 * d), m5 = static_cast<float>(-5 * d),
                  m7 = static_cast<float>(-7 * d);
  if (ang >= t1 && ang < t3) return 0;
  if (ang >= t3 && ang < t5) return 1;
  if (ang >= t5 && ang < t7) return 2;
  if (ang >= m7 && ang < m5) return 4;
  if (ang >= m5 && ang < m3) return 5;
  if (ang >= m3 && ang < m1) return 6;
  if (ang >= m1 && ang < t1) return 7;
  return 3;
}

// Thread = pixel, loops over the fg classes present in the image.  KS = 5 (the trainer's choice) or 3.
template <int KS>
__global__ void __launch_bounds__(256) field_sobel_kernel(const float* prob, const float* label, int C, int H, int W,
                                                          float* mag, uint8_t* orient, float* mag_sum,
                                                          unsigned int* plane_max) {
  const long long HW = static_cast<long long>(H) * W;
  const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= HW) return;
  const int y = static_cast<int>(p / W), x = static_cast<int>(p - static_cast<long long>(y) * W);
  float total = 0.f;
  for (int c = 0; c < C - 1; ++c) {
    const float lab = label[b * C + 1 + c];
    float gx = 0.f, gy = 0.f;
    if (lab != 0.f) {
      const float* pl = prob + (static_cast<long long>(b) * (C - 1) + c) * HW;
#pragma unroll
      for (int r = 0; r < KS; ++r) {
        const int yy = y + r - KS / 2;
        if (yy < 0 || yy >= H) continue;
#pragma unroll
        for (int s = 0; s < KS; ++s) {
          const int xx = x + s - KS / 2;
          if (xx < 0 || xx >= W) continue;
          const float v = pl[static_cast<long long>(yy) * W + xx];
          gx = fmaf(KS == 5 ? c_kx[r * 5 + s] : c_kx3[r * 3 + s], v, gx);
          gy = fmaf(KS == 5 ? c_ky[r * 5 + s] : c_ky3[r * 3 + s], v, gy);
        }
      }
      gx *= lab;
      gy *= lab;
    }
    const float m = sqrtf(gx * gx + gy * gy + 1e-8f);
    const long long idx = (static_cast<long long>(b) * (C - 1) + c) * HW + p;
    mag[idx] = m;
    orient[idx] = static_cast<uint8_t>(orient_bin(gy, gx));
    total += m;
    if (lab != 0.f) atomicMax(plane_max + b * (C - 1) + c, __float_as_uint(m));   // m > 0: uint order == float order
  }
  mag_sum[static_cast<long long>(b) * HW + p] = total;
}

// outside / inside pixel of boundary pixel idx with orientation bin `o` (0..7); src/edge.py:195-227:
// bins 3,7 (reference values 4,8) -> out = idx - step*W, in = idx + W; bins 1,5 -> out = idx-1, in = idx+1;
// other bins -> idx itself.  Indices with t % (W-1) in {0,1}, t <= 0 or t >= W*H-1 are dropped (:220-221).
__device__ __forceinline__ void in_out_of(long long idx, int o, int step, int W, long long HW, long long& out_i,
                                          long long& in_i, bool& keep_out, bool& keep_in) {
  const bool vert = (o == 3) || (o == 7), horiz = (o == 1) || (o == 5);
  out_i = idx - (vert ? static_cast<long long>(step) * W : 0) - (horiz ? 1 : 0);
  in_i = idx + (vert ? W : 0) + (horiz ? 1 : 0);
  auto keep = [&](long long t) {
    long long m = t % (W - 1);
    if (m < 0) m += (W - 1);   // python modulo
    return m != 0 && m != 1 && t > 0 && t < HW - 1;
  };
  keep_out = keep(out_i);
  keep_in = keep(in_i);
}

// pos flags + per-plane counts {boundary pixels, kept outside indices, kept inside indices}
__global__ void __launch_bounds__(256) field_pos_kernel(const float* mag, const uint8_t* orient, const float* label,
                                                        const unsigned int* plane_max, int C, int H, int W, int step,
                                                        uint8_t* pos, int* counts) {
  __shared__ int sh[3];
  const long long HW = static_cast<long long>(H) * W;
  const int plane = blockIdx.y;  // b * (C-1) + c
  const int b = plane / (C - 1), c = plane - b * (C - 1);
  if (threadIdx.x < 3) sh[threadIdx.x] = 0;
  __syncthreads();
  const float mx = __uint_as_float(plane_max[plane]);
  const bool active = label[b * C + 1 + c] != 0.f && mx > 1.f;
  const float thr = 0.8f * mx;
  int np = 0, no = 0, ni = 0;
  for (long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; p < HW;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long idx = static_cast<long long>(plane) * HW + p;
    const bool is = active && mag[idx] >= thr;
    pos[idx] = is ? 1 : 0;
    if (is) {
      long long o_i, i_i;
      bool ko, ki;
      in_out_of(p, orient[idx], step, W, HW, o_i, i_i, ko, ki);
      ++np;
      no += ko;
      ni += ki;
    }
  }
  np = __reduce_add_sync(0xffffffffu, np);
  no = __reduce_add_sync(0xffffffffu, no);
  ni = __reduce_add_sync(0xffffffffu, ni);
  if ((threadIdx.x & 31) == 0) { atomicAdd(&sh[0], np); atomicAdd(&sh[1], no); atomicAdd(&sh[2], ni); }
  __syncthreads();
  if (threadIdx.x < 3 && sh[threadIdx.x]) atomicAdd(counts + plane * 3 + threadIdx.x, sh[threadIdx.x]);
}

// Ordered (raster) compaction of the kept outside / inside indices of one active plane.
// One block of 1024 threads per plane; lists[off_out + j], lists[off_in + j].
__global__ void __launch_bounds__(1024) field_compact_kernel(const uint8_t* orient, const uint8_t* pos, int H, int W,
                                                             int step, const int* planes, const long long* offsets,
                                                             int* lists) {
  __shared__ int wtot[2][32];
  __shared__ int carry[2];
  const long long HW = static_cast<long long>(H) * W;
  const int plane = planes[blockIdx.x];
  const long long off_out = offsets[blockIdx.x * 2], off_in = offsets[blockIdx.x * 2 + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < 2) carry[threadIdx.x] = 0;
  __syncthreads();
  for (long long base = 0; base < HW; base += blockDim.x) {
    const long long p = base + threadIdx.x;
    long long o_i = 0, i_i = 0;
    bool ko = false, ki = false;
    if (p < HW) {
      const long long idx = static_cast<long long>(plane) * HW + p;
      if (pos[idx]) in_out_of(p, orient[idx], step, W, HW, o_i, i_i, ko, ki);
    }
    const unsigned bo = __ballot_sync(0xffffffffu, ko), bi = __ballot_sync(0xffffffffu, ki);
    const int ro = __popc(bo & ((1u << lane) - 1)), ri = __popc(bi & ((1u << lane) - 1));
    if (lane == 0) { wtot[0][warp] = __popc(bo); wtot[1][warp] = __popc(bi); }
    __syncthreads();
    int wo = 0, wi = 0;
    for (int w = 0; w < warp; ++w) { wo += wtot[0][w]; wi += wtot[1][w]; }
    if (ko) lists[off_out + carry[0] + wo + ro] = static_cast<int>(o_i);
    if (ki) lists[off_in + carry[1] + wi + ri] = static_cast<int>(i_i);
    __syncthreads();
    if (threadIdx.x == 0) {
      int to = 0, ti = 0;
      for (int w = 0; w < 32; ++w) { to += wtot[0][w]; ti += wtot[1][w]; }
      carry[0] += to;
      carry[1] += ti;
    }
    __syncthreads();
  }
}

// Warp = one sampled pixel of one active plane: channel softmax of dense_ft (ch channels) and of the
// pseudo-mask (C channels) at that pixel.  probs_d [n_active][2][k][ch], probs_m [n_active][2][k][C],
// pix [n_active][2][k] (side 0 = outside, 1 = inside).
__global__ void __launch_bounds__(256) field_gather_softmax_kernel(const float* dense, const float* mask, int ch, int C,
                                                                   long long HW, const int* planes,
                                                                   const long long* offsets, const int* sel,
                                                                   const int* lists, int k, int n_samples,
                                                                   float* probs_d, float* probs_m, int* pix) {
  const int s = blockIdx.x * 8 + (threadIdx.x >> 5);   // global sample id = (plane_i * 2 + side) * k + j
  if (s >= n_samples) return;
  const int lane = threadIdx.x & 31;
  const int pi = s / (2 * k), side = (s / k) & 1;
  const int plane = planes[pi];
  const int b = plane / (C - 1);
  const int px = lists[offsets[pi * 2 + side] + sel[s]];
  if (lane == 0) pix[s] = px;
  {
    const float* src = dense + static_cast<long long>(b) * ch * HW + px;
    float* dst = probs_d + static_cast<long long>(s) * ch;
    float mx = -FLT_MAX;
    for (int c = lane; c < ch; c += 32) { const float v = src[c * HW]; dst[c] = v; mx = fmaxf(mx, v); }
    mx = warp_max_f(mx);
    float sum = 0.f;
    for (int c = lane; c < ch; c += 32) { const float e = expf(dst[c] - mx); dst[c] = e; sum += e; }
    sum = warp_sum_f(sum);
    for (int c = lane; c < ch; c += 32) dst[c] = dst[c] / sum;
  }
  {
    const float* src = mask + static_cast<long long>(b) * C * HW + px;
    float* dst = probs_m + static_cast<long long>(s) * C;
    const float v = (lane < C) ? src[lane * HW] : -FLT_MAX;
    const float mx = warp_max_f(v);
    const float e = (lane < C) ? expf(v - mx) : 0.f;
    const float sum = warp_sum_f(e);
    if (lane < C) dst[lane] = e / sum;
  }
}

// The same when dense_ft is never materialised (src/MuSCLe.py:285): the sampled pixel's
// dense_ft column is the bilinear interpolation (align_corners=True) of four columns of the low-resolution decoder
// output p3_dec [batch, ch, hs, ws] — 4 * ch reads per sample instead of a [batch, ch, H, W] tensor.
struct UpGeom { int hs, ws, W; float sy, sx; };
__global__ void __launch_bounds__(256) field_gather_up_softmax_kernel(const float* p3, const float* mask, int ch, int C,
                                                                      long long HW, UpGeom g, const int* planes,
                                                                      const long long* offsets, const int* sel,
                                                                      const int* lists, int k, int n_samples,
                                                                      float* probs_d, float* probs_m, int* pix) {
  const int s = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (s >= n_samples) return;
  const int lane = threadIdx.x & 31;
  const int pi = s / (2 * k), side = (s / k) & 1;
  const int plane = planes[pi];
  const int b = plane / (C - 1);
  const int px = lists[offsets[pi * 2 + side] + sel[s]];
  if (lane == 0) pix[s] = px;
  {
    const Tap ty = tap_of(px / g.W, g.sy, g.hs), tx = tap_of(px % g.W, g.sx, g.ws);
    const long long hw_s = static_cast<long long>(g.hs) * g.ws;
    const float* src = p3 + static_cast<long long>(b) * ch * hw_s;
    const int o00 = ty.i0 * g.ws + tx.i0, o01 = ty.i0 * g.ws + tx.i1, o10 = ty.i1 * g.ws + tx.i0, o11 = ty.i1 * g.ws + tx.i1;
    float* dst = probs_d + static_cast<long long>(s) * ch;
    float mx = -FLT_MAX;
    for (int c = lane; c < ch; c += 32) {
      const float* pc = src + c * hw_s;
      const float v = ty.w0 * (tx.w0 * pc[o00] + tx.w1 * pc[o01]) + ty.w1 * (tx.w0 * pc[o10] + tx.w1 * pc[o11]);
      dst[c] = v;
      mx = fmaxf(mx, v);
    }
    mx = warp_max_f(mx);
    float sum = 0.f;
    for (int c = lane; c < ch; c += 32) { const float e = expf(dst[c] - mx); dst[c] = e; sum += e; }
    sum = warp_sum_f(sum);
    for (int c = lane; c < ch; c += 32) dst[c] = dst[c] / sum;
  }
  {
    const float* src = mask + static_cast<long long>(b) * C * HW + px;
    float* dst = probs_m + static_cast<long long>(s) * C;
    const float v = (lane < C) ? src[lane * HW] : -FLT_MAX;
    const float mx = warp_max_f(v);
    const float e = (lane < C) ? expf(v - mx) : 0.f;
    const float sum = warp_sum_f(e);
    if (lane < C) dst[lane] = e / sum;
  }
}

// seg_small[b][k][p] = bias[k] + sum_c W[k][c] x[b][c][p]: the 1x1 `fuse_dec` at the decoder's own resolution
// (it commutes with the bilinear up-sampling that follows, src/MuSCLe.py:285-286).  Thread = pixel, K <= 32 outputs.
__global__ void __launch_bounds__(128) conv1x1_kernel(const float* __restrict__ x, const float* __restrict__ wgt,
                                                      const float* __restrict__ bias, int ch, int K, long long hw,
                                                      float* __restrict__ y) {
  extern __shared__ float w_sm[];   // [K][ch]
  for (int i = threadIdx.x; i < K * ch; i += blockDim.x) w_sm[i] = wgt[i];
  __syncthreads();
  const long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= hw) return;
  float acc[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) acc[k] = (k < K && bias) ? bias[k] : 0.f;
  const float* src = x + static_cast<long long>(b) * ch * hw + p;
  for (int c = 0; c < ch; ++c) {
    const float v = __ldg(src + c * hw);
#pragma unroll
    for (int k = 0; k < 32; ++k)
      if (k < K) acc[k] = fmaf(w_sm[k * ch + c], v, acc[k]);
  }
#pragma unroll
  for (int k = 0; k < 32; ++k)
    if (k < K) y[(static_cast<long long>(b) * K + k) * hw + p] = acc[k];
}

// One block per active plane: means, sign tests, BEACON terms, gradient scatter.
// smem (floats): sumO[ch] sumI[ch] wI[ch] | msumO[32] msumI[32] | m1[k] m0[k] mm1[k] mm0[k] coefA[k] coefB[k]
constexpr int FP_THREADS = 256;
__global__ void __launch_bounds__(FP_THREADS) field_plane_kernel(const float* probs_d, const float* probs_m,
                                                                 const int* pix, const int* planes, int ch, int C,
                                                                 long long HW, int k, float grad_scale, double* acc,
                                                                 float* ddense, UpGeom up) {
  extern __shared__ float sm[];
  float* sumO = sm;
  float* sumI = sumO + ch;
  float* wI = sumI + ch;
  float* msumO = wI + ch;
  float* msumI = msumO + 32;
  float* m1 = msumI + 32;
  float* m0 = m1 + k;
  float* mm1 = m0 + k;
  float* mm0 = mm1 + k;
  float* coefA = mm0 + k;
  float* coefB = coefA + k;
  __shared__ float gmean[2];
  __shared__ float set_sum[2][4];
  __shared__ int set_cnt[2][4];
  const int pi = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = FP_THREADS / 32;
  const float* outs = probs_d + static_cast<long long>(pi) * 2 * k * ch;
  const float* ins = outs + static_cast<long long>(k) * ch;
  const float* outs_m = probs_m + static_cast<long long>(pi) * 2 * k * C;
  const float* ins_m = outs_m + static_cast<long long>(k) * C;
  const float inv_k = 1.f / static_cast<float>(k);

  for (int c = threadIdx.x; c < ch; c += blockDim.x) {
    float so = 0.f, si = 0.f;
    for (int j = 0; j < k; ++j) { so += outs[static_cast<long long>(j) * ch + c]; si += ins[static_cast<long long>(j) * ch + c]; }
    sumO[c] = so;
    sumI[c] = si;
  }
  if (threadIdx.x < 64) {
    const int c = threadIdx.x & 31;
    const float* src = (threadIdx.x < 32) ? outs_m : ins_m;
    float s = 0.f;
    if (c < C) for (int j = 0; j < k; ++j) s += src[j * C + c];
    ((threadIdx.x < 32) ? msumO : msumI)[c] = s;
  }
  if (threadIdx.x < 8) { (&set_sum[0][0])[threadIdx.x] = 0.f; (&set_cnt[0][0])[threadIdx.x] = 0; }
  __syncthreads();
  // row means (index a over the outside samples) and column means (index b' over the inside samples)
  for (int j = warp; j < k; j += NW) {
    float a = 0.f, b2 = 0.f;
    for (int c = lane; c < ch; c += 32) {
      a = fmaf(outs[static_cast<long long>(j) * ch + c], sumI[c], a);
      b2 = fmaf(ins[static_cast<long long>(j) * ch + c], sumO[c], b2);
    }
    a = warp_sum_f(a);
    b2 = warp_sum_f(b2);
    float am = (lane < C) ? outs_m[j * C + lane] * msumI[lane] : 0.f;
    float bm = (lane < C) ? ins_m[j * C + lane] * msumO[lane] : 0.f;
    am = warp_sum_f(am);
    bm = warp_sum_f(bm);
    if (lane == 0) { m1[j] = a * inv_k; m0[j] = b2 * inv_k; mm1[j] = am * inv_k; mm0[j] = bm * inv_k; }
  }
  if (warp == 0) {
    float g = 0.f;
    for (int c = lane; c < ch; c += 32) g = fmaf(sumO[c], sumI[c], g);
    g = warp_sum_f(g);
    float gm2 = (lane < C) ? msumO[lane] * msumI[lane] : 0.f;
    gm2 = warp_sum_f(gm2);
    if (lane == 0) { gmean[0] = g * inv_k * inv_k; gmean[1] = gm2 * inv_k * inv_k; }
  }
  __syncthreads();
  // sets: 0 = FP (mask & !dense, -), 1 = FN (!mask & dense, +), 2 = TP (!mask & !dense, +), 3 = TN (mask & dense, -)
  auto set_of = [&](float m_dense, float m_mask) {
    const bool sd = m_dense > gmean[0], smk = m_mask > gmean[1];
    return smk ? (sd ? 3 : 0) : (sd ? 1 : 2);
  };
  for (int j = threadIdx.x; j < k; j += blockDim.x) {
    const int sa = set_of(m1[j], mm1[j]), sb = set_of(m0[j], mm0[j]);
    atomicAdd(&set_cnt[0][sa], 1);
    atomicAdd(&set_sum[0][sa], m1[j]);
    atomicAdd(&set_cnt[1][sb], 1);
    atomicAdd(&set_sum[1][sb], m0[j]);
  }
  __syncthreads();
  const float sign[4] = {-1.f, 1.f, 1.f, -1.f};
  if (threadIdx.x == 0) {
    double l = 0.0;
    for (int d = 0; d < 2; ++d)
      for (int s = 0; s < 4; ++s)
        if (set_cnt[d][s] > 0) l += static_cast<double>(sign[s] * set_sum[d][s] / static_cast<float>(set_cnt[d][s]));
    atomicAdd(acc, l);
  }
  if (ddense == nullptr) return;
  for (int j = threadIdx.x; j < k; j += blockDim.x) {
    const int sa = set_of(m1[j], mm1[j]), sb = set_of(m0[j], mm0[j]);
    coefA[j] = sign[sa] / static_cast<float>(set_cnt[0][sa]);
    coefB[j] = sign[sb] / static_cast<float>(set_cnt[1][sb]);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < ch; c += blockDim.x) {
    float w = 0.f;
    for (int j = 0; j < k; ++j) w = fmaf(coefB[j], ins[static_cast<long long>(j) * ch + c], w);
    wI[c] = w;
  }
  __syncthreads();
  // d outs[a] = (coefA[a] sumI + wI) / k ; back through the channel softmax; scatter into ddense
  const int plane = planes[pi];
  const int b = plane / (C - 1);
  const int* pix_out = pix + static_cast<long long>(pi) * 2 * k;
  for (int j = warp; j < k; j += NW) {
    const float ca = coefA[j];
    float dot = 0.f;
    for (int c = lane; c < ch; c += 32) {
      const float dy = (ca * sumI[c] + wI[c]) * inv_k;
      dot = fmaf(outs[static_cast<long long>(j) * ch + c], dy, dot);
    }
    dot = warp_sum_f(dot);
    if (up.hs > 0) {   // dense_ft was interpolated from p3_dec: the gradient goes to the four source columns
      const int px = pix_out[j];
      const Tap ty = tap_of(px / up.W, up.sy, up.hs), tx = tap_of(px % up.W, up.sx, up.ws);
      const long long hw_s = static_cast<long long>(up.hs) * up.ws;
      float* dst = ddense + static_cast<long long>(b) * ch * hw_s;
      for (int c = lane; c < ch; c += 32) {
        const float dy = (ca * sumI[c] + wI[c]) * inv_k;
        const float yv = outs[static_cast<long long>(j) * ch + c];
        const float gv = grad_scale * yv * (dy - dot);
        float* dc = dst + c * hw_s;
        atomicAdd(dc + ty.i0 * up.ws + tx.i0, gv * ty.w0 * tx.w0);
        atomicAdd(dc + ty.i0 * up.ws + tx.i1, gv * ty.w0 * tx.w1);
        atomicAdd(dc + ty.i1 * up.ws + tx.i0, gv * ty.w1 * tx.w0);
        atomicAdd(dc + ty.i1 * up.ws + tx.i1, gv * ty.w1 * tx.w1);
      }
      continue;
    }
    float* dst = ddense + static_cast<long long>(b) * ch * HW + pix_out[j];
    for (int c = lane; c < ch; c += 32) {
      const float dy = (ca * sumI[c] + wI[c]) * inv_k;
      const float yv = outs[static_cast<long long>(j) * ch + c];
      atomicAdd(dst + c * HW, grad_scale * yv * (dy - dot));
    }
  }
}

__global__ void field_final_kernel(const double* acc, float scale, float* loss) { *loss = static_cast<float>(*acc * scale); }

}  // namespace

extern "C" {

int muscle_field_edges(const float* seg_map, const float* label_with_bg, int batch, int num_classes, int height,
                       int width, float beta, int step, float* prob_scratch, float* mag_scratch,
                       uint8_t* orient_out, uint8_t* pos_out, float* mag_sum_out, uint32_t* plane_max_out,
                       int32_t* counts_out, void* stream_) {
  return muscle_field_edges_sobel(seg_map, label_with_bg, batch, num_classes, height, width, beta, step, 5, prob_scratch,
                                  mag_scratch, orient_out, pos_out, mag_sum_out, plane_max_out, counts_out, stream_);
}

int muscle_field_edges_sobel(const float* seg_map, const float* label_with_bg, int batch, int num_classes, int height,
                             int width, float beta, int step, int sobel_size, float* prob_scratch, float* mag_scratch,
                             uint8_t* orient_out, uint8_t* pos_out, float* mag_sum_out, uint32_t* plane_max_out,
                             int32_t* counts_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (sobel_size != 3 && sobel_size != 5)
    return fail(MUSCLE_ERR_BAD_ARG, "sobel_size must be 3 or 5 like src/edge.py:33-43 (got %d)", sobel_size);
  if (!seg_map || !label_with_bg || !prob_scratch || !mag_scratch || !orient_out || !pos_out || !mag_sum_out ||
      !plane_max_out || !counts_out)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || height < 3 || width < 3) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (num_classes < 2 || num_classes > NFG + 1) return fail(MUSCLE_ERR_BAD_ARG, "need 2..%d classes", NFG + 1);
  const long long HW = static_cast<long long>(height) * width;
  const int planes = batch * (num_classes - 1);
  if (planes > 65535 || batch > 65535) return fail(MUSCLE_ERR_BAD_ARG, "batch too large");
  MB_CUDA(cudaMemsetAsync(plane_max_out, 0, planes * sizeof(uint32_t), stream));
  MB_CUDA(cudaMemsetAsync(counts_out, 0, planes * 3 * sizeof(int32_t), stream));
  const dim3 gpx(static_cast<unsigned>((HW + 255) / 256), batch);
  field_prob_kernel<<<gpx, 256, 0, stream>>>(seg_map, beta, num_classes, HW, prob_scratch);
  MB_LAUNCH_CHECK();
  if (sobel_size == 5)
    field_sobel_kernel<5><<<gpx, 256, 0, stream>>>(prob_scratch, label_with_bg, num_classes, height, width, mag_scratch,
                                                   orient_out, mag_sum_out, plane_max_out);
  else
    field_sobel_kernel<3><<<gpx, 256, 0, stream>>>(prob_scratch, label_with_bg, num_classes, height, width, mag_scratch,
                                                   orient_out, mag_sum_out, plane_max_out);
  MB_LAUNCH_CHECK();
  const unsigned gx = static_cast<unsigned>(std::min<long long>((HW + 255) / 256, 64));
  field_pos_kernel<<<dim3(gx, planes), 256, 0, stream>>>(mag_scratch, orient_out, label_with_bg, plane_max_out,
                                                         num_classes, height, width, step, pos_out, counts_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_field_compact(const uint8_t* orient, const uint8_t* pos, int height, int width, int step,
                         const int32_t* planes, const int64_t* offsets, int n_active, int32_t* lists,
                         void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!orient || !pos || !planes || !offsets || !lists) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n_active <= 0) return fail(MUSCLE_ERR_BAD_ARG, "no active planes");
  if (static_cast<long long>(height) * width >= (1ll << 31)) return fail(MUSCLE_ERR_BAD_ARG, "image too large");
  field_compact_kernel<<<n_active, 1024, 0, stream>>>(orient, pos, height, width, step, planes,
                                                      reinterpret_cast<const long long*>(offsets), lists);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

size_t muscle_field_loss_scratch_bytes(int n_active, int k, int channels, int num_classes) {
  if (n_active <= 0 || k <= 0 || channels <= 0 || num_classes <= 0) return 0;
  const size_t ns = static_cast<size_t>(n_active) * 2 * k;
  return (ns * channels * 4 + 255) / 256 * 256 + (ns * num_classes * 4 + 255) / 256 * 256 + (ns * 4 + 255) / 256 * 256 + 256;
}

}  // extern "C"

namespace {
// dense_ft [batch, channels, height, width] (small_h == 0) or p3_dec [batch, channels, small_h, small_w] interpolated on the fly
int field_loss_impl(const float* dense_ft, int small_h, int small_w, const float* mask, int batch, int channels,
                    int num_classes, int height, int width, const int32_t* planes, const int64_t* offsets,
                    const int32_t* sel, const int32_t* lists, int n_active, int k, float* loss_out, float* ddense_out,
                    void* scratch, size_t scratch_bytes, cudaStream_t stream) {
  if (int rc = need_sm100()) return rc;
  if (!dense_ft || !mask || !planes || !offsets || !sel || !lists || !loss_out || !scratch)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || n_active <= 0 || k <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (num_classes < 2 || num_classes > 32) return fail(MUSCLE_ERR_BAD_ARG, "need 2..32 mask channels");
  if (scratch_bytes < muscle_field_loss_scratch_bytes(n_active, k, channels, num_classes))
    return fail(MUSCLE_ERR_WORKSPACE, "scratch too small");
  const long long HW = static_cast<long long>(height) * width;
  const size_t ns = static_cast<size_t>(n_active) * 2 * k;
  char* base = static_cast<char*>(scratch);
  float* probs_d = reinterpret_cast<float*>(base);
  size_t off = (ns * channels * 4 + 255) / 256 * 256;
  float* probs_m = reinterpret_cast<float*>(base + off);
  off += (ns * num_classes * 4 + 255) / 256 * 256;
  int* pix = reinterpret_cast<int*>(base + off);
  off += (ns * 4 + 255) / 256 * 256;
  double* acc = reinterpret_cast<double*>(base + off);
  UpGeom up{small_h, small_w, width, small_h > 0 ? bilinear_scale(small_h, height) : 0.f,
            small_h > 0 ? bilinear_scale(small_w, width) : 0.f};
  const size_t grad_elems = static_cast<size_t>(batch) * channels * (small_h > 0 ? static_cast<size_t>(small_h) * small_w : static_cast<size_t>(HW));
  MB_CUDA(cudaMemsetAsync(acc, 0, sizeof(double), stream));
  if (ddense_out) MB_CUDA(cudaMemsetAsync(ddense_out, 0, grad_elems * sizeof(float), stream));
  if (small_h > 0)
    field_gather_up_softmax_kernel<<<static_cast<unsigned>((ns + 7) / 8), 256, 0, stream>>>(
        dense_ft, mask, channels, num_classes, HW, up, planes, reinterpret_cast<const long long*>(offsets), sel, lists, k,
        static_cast<int>(ns), probs_d, probs_m, pix);
  else
    field_gather_softmax_kernel<<<static_cast<unsigned>((ns + 7) / 8), 256, 0, stream>>>(
        dense_ft, mask, channels, num_classes, HW, planes, reinterpret_cast<const long long*>(offsets), sel, lists, k,
        static_cast<int>(ns), probs_d, probs_m, pix);
  MB_LAUNCH_CHECK();
  const size_t smem = (static_cast<size_t>(3) * channels + 64 + static_cast<size_t>(6) * k) * sizeof(float);
  if (smem > 200 * 1024) return fail(MUSCLE_ERR_BAD_ARG, "k or channel count too large for one block");
  if (int rc = ensure_dynamic_smem<field_plane_kernel>(200 * 1024)) return rc;
  const float inv_n = 1.f / static_cast<float>(batch);   // loss / n  (src/edge.py:359-360)
  field_plane_kernel<<<n_active, FP_THREADS, smem, stream>>>(probs_d, probs_m, pix, planes, channels, num_classes, HW, k,
                                                             inv_n, acc, ddense_out, up);
  MB_LAUNCH_CHECK();
  field_final_kernel<<<1, 1, 0, stream>>>(acc, inv_n, loss_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
}  // namespace

extern "C" {

int muscle_field_loss(const float* dense_ft, const float* mask, int batch, int channels, int num_classes,
                      int height, int width, const int32_t* planes, const int64_t* offsets, const int32_t* sel,
                      const int32_t* lists, int n_active, int k, float* loss_out, float* ddense_out, void* scratch,
                      size_t scratch_bytes, void* stream_) {
  return field_loss_impl(dense_ft, 0, 0, mask, batch, channels, num_classes, height, width, planes, offsets, sel, lists,
                         n_active, k, loss_out, ddense_out, scratch, scratch_bytes, static_cast<cudaStream_t>(stream_));
}

int muscle_field_loss_decoder(const float* p3_dec, int small_height, int small_width, const float* mask, int batch,
                              int channels, int num_classes, int height, int width, const int32_t* planes,
                              const int64_t* offsets, const int32_t* sel, const int32_t* lists, int n_active, int k,
                              float* loss_out, float* dp3_out, void* scratch, size_t scratch_bytes, void* stream_) {
  if (small_height <= 0 || small_width <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty decoder map");
  return field_loss_impl(p3_dec, small_height, small_width, mask, batch, channels, num_classes, height, width, planes,
                         offsets, sel, lists, n_active, k, loss_out, dp3_out, scratch, scratch_bytes,
                         static_cast<cudaStream_t>(stream_));
}

int muscle_conv1x1_fwd(const float* x, const float* weight, const float* bias, int batch, int channels, int out_channels,
                       int64_t pixels, float* y, void* stream_) {
  if (int rc = need_sm100()) return rc;
  if (!x || !weight || !y) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || channels <= 0 || pixels <= 0 || out_channels < 1 || out_channels > 32)
    return fail(MUSCLE_ERR_BAD_ARG, "1 <= out_channels <= 32 and a non-empty input");
  const size_t smem = static_cast<size_t>(out_channels) * channels * sizeof(float);
  if (smem > 160 * 1024) return fail(MUSCLE_ERR_BAD_ARG, "weight matrix too large for shared memory");
  if (int rc = ensure_dynamic_smem<conv1x1_kernel>(160 * 1024)) return rc;
  conv1x1_kernel<<<dim3(static_cast<unsigned>((pixels + 127) / 128), batch), 128, smem, static_cast<cudaStream_t>(stream_)>>>(
      x, weight, bias, channels, out_channels, pixels, y);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // extern "C"
