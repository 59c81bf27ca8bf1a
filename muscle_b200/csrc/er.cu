// the ER loss of train_mcl.py:175-188 and the bilinear up-sampling in front of it
// (src/MuSCLe.py:256-257), include/muscle_b200.h muscle_upsample_bilinear_* / muscle_er_loss_*.
//
// ER loss: cams = cam_softmaxnorm(raw_cams) (detached), sgcs = cam_softmaxnorm(raw_sgcs), both masked by the image
// labels (background always on); per image the k = int(0.2 * valid_channel * H * W) largest |cams - sgcs| over
// (class, pixel); the mean of all n*k values.  The reference materialises both normalised maps and sorts
// 21*H*W values per image (torch.topk); here nothing of size N is written in the forward: the k-th largest value
// is found by a 3-level radix select (11 + 11 + 10 bits) on the values as 32-bit fixed point (they lie in [0, 1]; a
// linear scale spreads them over the bins, float exponents would pile them into ~30) whose histograms also carry
// the SUM of every bin as an integer, so the sum of the top k falls out of the same three passes, exactly and in
// any order; neighbouring pixels of an up-sampled map fall into the same bin, so a warp first groups its lanes by
// bin (match.any) and issues one shared-memory atomic per group; |cams - sgcs| is recomputed per pass from the
// two raw maps (HBM-bound: 2 * 21 * 4 bytes per pixel and pass).  valid_channel is summed on the device: no host
// synchronisation (the reference reads it back, train_mcl.py:178).
// Backward: one pass, d loss / d raw_sgcs = cam_softmaxnorm' ( -sign(cams - sgcs) [in the top k] / (n k) ).
// Ties at the k-th value share the remaining weight equally (torch.topk picks arbitrary ones: equal in value).
#include <cfloat>

#include "bilinear.cuh"
#include "host_util.cuh"

using namespace mb200;

namespace {

constexpr int ER_C = 21;            // VOC classes incl. background (src: cams[:, 1:] are the 20 fg channels)
constexpr int ER_THREADS = 256;
constexpr int ER_BINS = 2048;

struct ErSel {                      // per image, device resident
  unsigned int prefix;              // bits fixed so far (top-aligned); after level 3: the k-th largest value (fixed point)
  unsigned int need;                // how many values are still to be taken from the current bin
  unsigned long long sum_above;     // fixed-point sum of the values in the bins above the chosen ones (+ the ties taken)
  unsigned long long cnt_above;     // their count
  float tie_weight;                 // weight of a value == the threshold
  unsigned int done;                // k exceeds the number of non-zero values: everything non-zero counts
  unsigned long long k;
};

// |cams - sgcs| in [0, 1] as 32-bit fixed point (monotone; exact for values >= 2^-8, 2^-32 resolution below)
__device__ __forceinline__ unsigned int fixed_of(float d) {
  return d >= 1.f ? 0xffffffffu : __float2uint_rz(d * 4294967296.f);
}

// softmax over the fg channels, bg = 1 - max fg, times the label mask -> out[0..C)
__device__ __forceinline__ void softmaxnorm_masked(const float* __restrict__ src, long long hw, const float* __restrict__ lab_fg,
                                                   float (&out)[ER_C], int* argmax_out) {
  float v[ER_C];
  float mx = -FLT_MAX;
#pragma unroll
  for (int c = 1; c < ER_C; ++c) { v[c] = __ldg(src + c * hw); mx = fmaxf(mx, v[c]); }
  float s = 0.f;
#pragma unroll
  for (int c = 1; c < ER_C; ++c) { v[c] = __expf(v[c] - mx); s += v[c]; }
  const float inv = 1.f / s;
  float pmax = 0.f;
  int am = 1;
#pragma unroll
  for (int c = 1; c < ER_C; ++c) {
    const float p = v[c] * inv;
    if (p > pmax) { pmax = p; am = c; }
    out[c] = p * lab_fg[c - 1];
    v[c] = p;
  }
  out[0] = 1.f - pmax;              // background label is always 1 (train_mcl.py:167-170)
  if (argmax_out) *argmax_out = am;
}

__device__ __forceinline__ unsigned int digit_of(unsigned int bits, int level) {
  return level == 0 ? bits >> 21 : (level == 1 ? (bits >> 10) & 0x7ffu : bits & 0x3ffu);
}
__device__ __forceinline__ bool prefix_ok(unsigned int bits, unsigned int prefix, int level) {
  return level == 0 ? true : (level == 1 ? (bits >> 21) == (prefix >> 21) : (bits >> 10) == (prefix >> 10));
}

// one level of the select: histogram (count + fixed-point sum) of the level's digit over the values that match the prefix
__global__ void __launch_bounds__(ER_THREADS, 2) er_hist_kernel(const float* __restrict__ cams, const float* __restrict__ sgcs,
                                                             const float* __restrict__ label, long long hw, int level,
                                                             const ErSel* __restrict__ sel, unsigned int* __restrict__ hist,
                                                             unsigned long long* __restrict__ hsum) {
  // per-bin count and fixed-point sum; the sum as two 32-bit halves (sum of the high 16 bits, sum of the low 16 bits):
  // the host sizes the grid so that a block sees fewer than 2^16 values, so neither half overflows, and 32-bit shared-memory
  // atomics are native where the 64-bit add was a compare-and-swap loop (level 0: 615 us, short-scoreboard bound)
  __shared__ unsigned int h_sm[ER_BINS];
  __shared__ unsigned int shi_sm[ER_BINS], slo_sm[ER_BINS];
  __shared__ float lab_sm[ER_C - 1];
  const int b = blockIdx.y;
  if (level > 0 && sel[b].done) return;
  for (int i = threadIdx.x; i < ER_BINS; i += ER_THREADS) { h_sm[i] = 0u; shi_sm[i] = 0u; slo_sm[i] = 0u; }
  if (threadIdx.x < ER_C - 1) lab_sm[threadIdx.x] = label[b * (ER_C - 1) + threadIdx.x];
  __syncthreads();
  const unsigned int prefix = level > 0 ? sel[b].prefix : 0u;
  const float* cb = cams + static_cast<long long>(b) * ER_C * hw;
  const float* sb = sgcs + static_cast<long long>(b) * ER_C * hw;
  const int lane = threadIdx.x & 31;
  const long long stride = static_cast<long long>(gridDim.x) * ER_THREADS;
  const long long p_end = (hw + stride - 1) / stride * stride;     // uniform trip count: the warp votes stay converged
  for (long long p = static_cast<long long>(blockIdx.x) * ER_THREADS + threadIdx.x; p < p_end; p += stride) {
    const bool in = p < hw;
    float a[ER_C], q[ER_C];
    if (in) {
      softmaxnorm_masked(cb + p, hw, lab_sm, a, nullptr);
      softmaxnorm_masked(sb + p, hw, lab_sm, q, nullptr);
    }
#pragma unroll
    for (int c = 0; c < ER_C; ++c) {
      // a channel whose label is off is exactly 0 on both sides (softmaxnorm_masked multiplies by the label): nothing to
      // count, and skipping it -- a block-uniform branch, one image per blockIdx.y -- skips its match.any / redux, which
      // is what this kernel spends its time on (typically 2-4 of the 21 channels are on)
      if (c > 0 && lab_sm[c - 1] == 0.f) continue;
      const unsigned int u = in ? fixed_of(fabsf(a[c] - q[c])) : 0u;
      const bool take = u != 0u && prefix_ok(u, prefix, level);
      if (!__any_sync(0xffffffffu, take)) continue;      // levels 1 and 2: almost no warp holds a value with the chosen prefix
      const unsigned int dg = take ? digit_of(u, level) : 0xffffffffu;
      const unsigned int peers = __match_any_sync(0xffffffffu, dg);
      if (take) {
        const unsigned int hi = __reduce_add_sync(peers, u >> 16), lo = __reduce_add_sync(peers, u & 0xffffu);
        if (lane == __ffs(peers) - 1) {
          atomicAdd(&h_sm[dg], static_cast<unsigned int>(__popc(peers)));
          atomicAdd(&shi_sm[dg], hi);
          atomicAdd(&slo_sm[dg], lo);
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ER_BINS; i += ER_THREADS)
    if (h_sm[i]) {
      atomicAdd(hist + static_cast<size_t>(b) * ER_BINS + i, h_sm[i]);
      atomicAdd(hsum + static_cast<size_t>(b) * ER_BINS + i, (static_cast<unsigned long long>(shi_sm[i]) << 16) + slo_sm[i]);
    }
}

// one block per image: walk the bins from the top until `need` values are covered
__global__ void __launch_bounds__(256) er_pick_kernel(const float* __restrict__ label, int n_fg, int batch, long long hw,
                                                      int height, int width, int level, unsigned int* __restrict__ hist,
                                                      unsigned long long* __restrict__ hsum, ErSel* __restrict__ sel,
                                                      float* __restrict__ loss_out) {
  const int b = blockIdx.x;
  __shared__ unsigned long long k_sh;
  if (threadIdx.x == 0) {
    if (level == 0) {
      float valid = 0.f;                       // valid_channel = int(label.sum()) over the WHOLE batch (:178)
      for (int i = 0; i < batch * n_fg; ++i) valid += label[i];
      const double kk = ((0.2 * static_cast<double>(static_cast<long long>(valid))) * height) * width;   // python's order
      ErSel s{};
      s.k = static_cast<unsigned long long>(kk);
      s.need = 0u; s.prefix = 0u; s.sum_above = 0ull; s.cnt_above = 0ull; s.done = 0u; s.tie_weight = 0.f;
      sel[b] = s;
    }
    k_sh = sel[b].k;
  }
  __syncthreads();
  // Walk from the top bin down until `need` values are covered -- as a block-wide scan: thread t owns the eight bins
  // r = 8t .. 8t+7 counted from the top (r = nb - 1 - i), an exclusive scan over the threads gives the count and the
  // fixed-point sum above its run, and the one thread whose run crosses `need` finishes the walk inside its eight bins.
  // (One thread walking up to 2 048 bins of global memory took 212 us per level: a third of the whole ER loss.)
  __shared__ unsigned int cnt_w[8];
  __shared__ unsigned long long sum_w[8];
  __shared__ int found_sh;
  if (threadIdx.x == 0) found_sh = 0;
  const bool active = !sel[b].done;            // uniform: every thread reads the same word
  if (active) {
    const unsigned int* h = hist + static_cast<size_t>(b) * ER_BINS;
    const unsigned long long* hs = hsum + static_cast<size_t>(b) * ER_BINS;
    const int nb = (level == 2) ? 1024 : ER_BINS;
    const int per = nb / 256;                  // 8 or 4 bins per thread
    unsigned int c_loc[8];
    unsigned long long s_loc[8];
    unsigned int c_tot = 0u;
    unsigned long long s_tot = 0ull;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int i = nb - 1 - (static_cast<int>(threadIdx.x) * per + q);
      const bool in = q < per;
      c_loc[q] = in ? h[i] : 0u;
      s_loc[q] = in ? hs[i] : 0ull;
      c_tot += c_loc[q];
      s_tot += s_loc[q];
    }
    // exclusive scan over the 256 threads (top run first)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int c_inc = c_tot;
    unsigned long long s_inc = s_tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned int cu = __shfl_up_sync(0xffffffffu, c_inc, o);
      const unsigned long long su = __shfl_up_sync(0xffffffffu, s_inc, o);
      if (lane >= o) { c_inc += cu; s_inc += su; }
    }
    if (lane == 31) { cnt_w[warp] = c_inc; sum_w[warp] = s_inc; }
    __syncthreads();
    unsigned long long c_before = 0ull, s_before = 0ull;     // (counts of a whole image can exceed 32 bits only in theory: 64-bit anyway)
    for (int w = 0; w < warp; ++w) { c_before += cnt_w[w]; s_before += sum_w[w]; }
    c_before += c_inc - c_tot;
    s_before += s_inc - s_tot;
    ErSel s = sel[b];
    const unsigned long long need0 = (level == 0) ? s.k : s.need;
    if (c_before < need0 && c_before + c_tot >= need0) {     // this thread's run holds the bin that completes `need`
      unsigned long long need = need0 - c_before;
      s.cnt_above += c_before;
      s.sum_above += s_before;
      int chosen = -1;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (chosen >= 0 || q >= per) continue;
        const unsigned int c = c_loc[q];
        if (c == 0u) continue;
        if (c >= need) { chosen = nb - 1 - (static_cast<int>(threadIdx.x) * per + q); continue; }
        need -= c;
        s.cnt_above += c;
        s.sum_above += s_loc[q];
      }
      s.prefix |= (level == 0) ? (static_cast<unsigned int>(chosen) << 21)
                               : (level == 1 ? (static_cast<unsigned int>(chosen) << 10) : static_cast<unsigned int>(chosen));
      s.need = static_cast<unsigned int>(need);
      if (level == 2) {
        const unsigned int ties = h[chosen];
        s.tie_weight = static_cast<float>(static_cast<double>(need) / static_cast<double>(ties));
        s.sum_above += need * static_cast<unsigned long long>(s.prefix);
      }
      sel[b] = s;
      found_sh = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0 && !found_sh) {       // fewer non-zero values than k: every one of them is taken, the rest are zeros
      unsigned long long c_all = 0ull, s_all = 0ull;
      for (int w = 0; w < 8; ++w) { c_all += cnt_w[w]; s_all += sum_w[w]; }
      s.cnt_above += c_all;
      s.sum_above += s_all;
      s.done = 1u;
      s.tie_weight = 0.f;
      sel[b] = s;
    }
  }
  // clear the image's histogram for the next level
  __syncthreads();
  for (int i = threadIdx.x; i < ER_BINS; i += blockDim.x) {
    hist[static_cast<size_t>(b) * ER_BINS + i] = 0u;
    hsum[static_cast<size_t>(b) * ER_BINS + i] = 0ull;
  }
  if (level == 2 && loss_out != nullptr) {     // loss = sum_b top-k sums / (n k): the last block to finish adds nothing
    __threadfence();                           // special: every block adds its share atomically (loss_out zeroed by the host)
    if (threadIdx.x == 0) {
      const ErSel s = sel[b];
      const double share = (s.k == 0ull) ? nan("")
                                         : (static_cast<double>(s.sum_above) / 4294967296.0) /
                                               (static_cast<double>(batch) * static_cast<double>(s.k));
      atomicAdd(loss_out, static_cast<float>(share));
    }
  }
}

// d loss / d raw_sgcs (all 21 channels written; the raw background channel gets 0: cam_softmaxnorm never reads it)
__global__ void __launch_bounds__(ER_THREADS) er_bwd_kernel(const float* __restrict__ cams, const float* __restrict__ sgcs,
                                                            const float* __restrict__ label, long long hw, int batch,
                                                            const ErSel* __restrict__ sel, const float* __restrict__ grad_scale,
                                                            float* __restrict__ dsgcs) {
  __shared__ float lab_sm[ER_C - 1];
  const int b = blockIdx.y;
  if (threadIdx.x < ER_C - 1) lab_sm[threadIdx.x] = label[b * (ER_C - 1) + threadIdx.x];
  __syncthreads();
  const ErSel s = sel[b];
  const float up = (grad_scale ? __ldg(grad_scale) : 1.f) / (static_cast<float>(batch) * static_cast<float>(s.k));
  const float* cb = cams + static_cast<long long>(b) * ER_C * hw;
  const float* sb = sgcs + static_cast<long long>(b) * ER_C * hw;
  float* db = dsgcs + static_cast<long long>(b) * ER_C * hw;
  for (long long p = static_cast<long long>(blockIdx.x) * ER_THREADS + threadIdx.x; p < hw;
       p += static_cast<long long>(gridDim.x) * ER_THREADS) {
    float a[ER_C], q[ER_C];
    int am = 1;
    softmaxnorm_masked(cb + p, hw, lab_sm, a, nullptr);
    softmaxnorm_masked(sb + p, hw, lab_sm, q, &am);
    // gradient w.r.t. the MASKED sgcs, then back through the mask
    float g[ER_C];
#pragma unroll
    for (int c = 0; c < ER_C; ++c) {
      const float diff = a[c] - q[c];
      const unsigned int u = fixed_of(fabsf(diff));
      float wgt = 0.f;
      if (u != 0u) wgt = s.done ? 1.f : (u > s.prefix ? 1.f : (u == s.prefix ? s.tie_weight : 0.f));
      const float sg = diff > 0.f ? -1.f : 1.f;     // d|a - q| / dq
      g[c] = wgt * sg * up * (c == 0 ? 1.f : lab_sm[c - 1]);
    }
    // back through cam_softmaxnorm: out_c = p_c (c >= 1), out_0 = 1 - p_argmax
    // (q holds p_c * label; the unmasked p_c is needed: recover it where the label is on, recompute otherwise)
    float pr[ER_C];
    {
      float mx = -FLT_MAX;
#pragma unroll
      for (int c = 1; c < ER_C; ++c) { pr[c] = __ldg(sb + p + c * hw); mx = fmaxf(mx, pr[c]); }
      float ssum = 0.f;
#pragma unroll
      for (int c = 1; c < ER_C; ++c) { pr[c] = __expf(pr[c] - mx); ssum += pr[c]; }
      const float inv = 1.f / ssum;
#pragma unroll
      for (int c = 1; c < ER_C; ++c) pr[c] *= inv;
    }
    float dot = 0.f;
#pragma unroll
    for (int c = 1; c < ER_C; ++c) {
      g[c] -= (c == am) ? g[0] : 0.f;
      dot = fmaf(pr[c], g[c], dot);
    }
    db[p] = 0.f;
#pragma unroll
    for (int c = 1; c < ER_C; ++c) db[p + c * hw] = pr[c] * (g[c] - dot);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// bilinear up/down-sampling, align_corners=True (F.interpolate as called at src/MuSCLe.py:256-257, :285)
// Block = UP_ROWS consecutive output rows of one plane (blockIdx.x = row group, blockIdx.y = plane, blockIdx.z = 512-pixel
// stretch of the row), four consecutive outputs per thread: the x taps of a thread's four columns are computed ONCE and
// reused for every row, a row's y taps are block-uniform, the stores are 16-byte vectors and no index needs a division.
// (History: one output per thread on a flat index, three 64-bit divisions per element: 148 us for a 135 MB map; one row
// per block with the taps recomputed per row: 85 us, 84 % issue-bound at 74 instructions per output.)
constexpr int UP_THREADS = 128;
constexpr int UP_ROWS = 8;
__global__ void __launch_bounds__(UP_THREADS) upsample_fwd_kernel(const float* __restrict__ x, long long planes, int h, int w, int H,
                                                                  int W, float sy, float sx, float* __restrict__ y) {
  const long long pl = blockIdx.y;
  const int oy0 = blockIdx.x * UP_ROWS;
  const int ox0 = (blockIdx.z * UP_THREADS + threadIdx.x) * 4;
  if (ox0 >= W) return;
  Tap tx[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) tx[q] = tap_of(ox0 + q < W ? ox0 + q : W - 1, sx, w);
  const float* src = x + pl * h * w;
  const bool vec = (W & 3) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0;
#pragma unroll 2
  for (int r = 0; r < UP_ROWS; ++r) {
    const int oy = oy0 + r;
    if (oy >= H) break;
    const Tap ty = tap_of(oy, sy, h);
    const float* r0 = src + static_cast<long long>(ty.i0) * w;
    const float* r1 = src + static_cast<long long>(ty.i1) * w;
    float o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float v00 = __ldg(r0 + tx[q].i0), v01 = __ldg(r0 + tx[q].i1);
      const float v10 = __ldg(r1 + tx[q].i0), v11 = __ldg(r1 + tx[q].i1);
      o[q] = ty.w0 * (tx[q].w0 * v00 + tx[q].w1 * v01) + ty.w1 * (tx[q].w0 * v10 + tx[q].w1 * v11);
    }
    float* dst = y + (pl * H + oy) * W;
    if (vec) {
      *reinterpret_cast<float4*>(dst + ox0) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (ox0 + q < W) dst[ox0 + q] = o[q];
    }
  }
}

// adjoint, separable.  Pass 1 (along x): tmp[pl][oy][j] = sum_ox wx(ox -> j) dy[pl][oy][ox]
__global__ void __launch_bounds__(256) upsample_bwd_x_kernel(const float* __restrict__ dy, long long rows, int w, int W,
                                                             float sx, float* __restrict__ tmp) {
  const long long total = rows * w;
  const float inv = sx > 0.f ? 1.f / sx : 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * 256) {
    const int j = static_cast<int>(i % w);
    const long long row = i / w;
    const float* src = dy + row * W;
    // outputs whose source index falls in (j-1, j+1): candidates, each tested with the forward's own arithmetic
    int lo = sx > 0.f ? static_cast<int>(floorf((static_cast<float>(j) - 1.f) * inv)) - 1 : 0;
    int hi = sx > 0.f ? static_cast<int>(ceilf((static_cast<float>(j) + 1.f) * inv)) + 1 : W - 1;
    lo = lo < 0 ? 0 : lo;
    hi = hi > W - 1 ? W - 1 : hi;
    float acc = 0.f;
    for (int o = lo; o <= hi; ++o) {
      const Tap t = tap_of(o, sx, w);
      float wgt = 0.f;
      if (t.i0 == j) wgt += t.w0;
      if (t.i1 == j) wgt += t.w1;
      if (wgt != 0.f) acc = fmaf(wgt, __ldg(src + o), acc);
    }
    tmp[i] = acc;
  }
}
// Pass 2 (along y): dx[pl][i][j] = sum_oy wy(oy -> i) tmp[pl][oy][j]
__global__ void __launch_bounds__(256) upsample_bwd_y_kernel(const float* __restrict__ tmp, long long planes, int h, int w, int H,
                                                             float sy, float* __restrict__ dx) {
  const long long total = planes * h * w;
  const float inv = sy > 0.f ? 1.f / sy : 0.f;
  for (long long idx = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * 256) {
    const int j = static_cast<int>(idx % w);
    const long long r = idx / w;
    const int i = static_cast<int>(r % h);
    const long long pl = r / h;
    const float* src = tmp + pl * H * w + j;
    int lo = sy > 0.f ? static_cast<int>(floorf((static_cast<float>(i) - 1.f) * inv)) - 1 : 0;
    int hi = sy > 0.f ? static_cast<int>(ceilf((static_cast<float>(i) + 1.f) * inv)) + 1 : H - 1;
    lo = lo < 0 ? 0 : lo;
    hi = hi > H - 1 ? H - 1 : hi;
    float acc = 0.f;
    for (int o = lo; o <= hi; ++o) {
      const Tap t = tap_of(o, sy, h);
      float wgt = 0.f;
      if (t.i0 == i) wgt += t.w0;
      if (t.i1 == i) wgt += t.w1;
      if (wgt != 0.f) acc = fmaf(wgt, __ldg(src + static_cast<long long>(o) * w), acc);
    }
    dx[idx] = acc;
  }
}

int need_sm100_er() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}
float scale_of(int n_in, int n_out) { return bilinear_scale(n_in, n_out); }
int grid_for(long long total) {
  const long long blocks = (total + 255) / 256;
  const long long cap = 32ll * device_info().sm_count;
  return static_cast<int>(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace

extern "C" {

int muscle_upsample_bilinear_fwd(const float* x, int64_t planes, int height, int width, int out_height, int out_width,
                                 float* y, void* stream_) {
  if (int rc = need_sm100_er()) return rc;
  if (!x || !y) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (planes <= 0 || height <= 0 || width <= 0 || out_height <= 0 || out_width <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  const long long total = planes * out_height * out_width;
  if (planes > 65535) return fail(MUSCLE_ERR_BAD_ARG, "too many planes (max 65535)");
  const unsigned gz = static_cast<unsigned>((out_width + UP_THREADS * 4 - 1) / (UP_THREADS * 4));
  if (gz > 65535) return fail(MUSCLE_ERR_BAD_ARG, "output too wide");
  const dim3 grid(static_cast<unsigned>((out_height + UP_ROWS - 1) / UP_ROWS), static_cast<unsigned>(planes), gz);
  upsample_fwd_kernel<<<grid, UP_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
      x, planes, height, width, out_height, out_width, scale_of(height, out_height), scale_of(width, out_width), y);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

size_t muscle_upsample_bilinear_bwd_scratch_bytes(int64_t planes, int out_height, int width) {
  return static_cast<size_t>(planes) * out_height * width * sizeof(float);
}

int muscle_upsample_bilinear_bwd(const float* dy, int64_t planes, int height, int width, int out_height, int out_width,
                                 float* dx, void* scratch, size_t scratch_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100_er()) return rc;
  if (!dy || !dx || !scratch) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (planes <= 0 || height <= 0 || width <= 0 || out_height <= 0 || out_width <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (scratch_bytes < muscle_upsample_bilinear_bwd_scratch_bytes(planes, out_height, width))
    return fail(MUSCLE_ERR_WORKSPACE, "scratch too small");
  float* tmp = static_cast<float*>(scratch);
  const long long rows = planes * out_height;
  upsample_bwd_x_kernel<<<grid_for(rows * width), 256, 0, stream>>>(dy, rows, width, out_width, scale_of(width, out_width), tmp);
  MB_LAUNCH_CHECK();
  upsample_bwd_y_kernel<<<grid_for(planes * height * width), 256, 0, stream>>>(tmp, planes, height, width, out_height,
                                                                              scale_of(height, out_height), dx);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

size_t muscle_er_loss_state_bytes(int batch) {
  return static_cast<size_t>(batch) * (sizeof(ErSel) + ER_BINS * (sizeof(unsigned int) + sizeof(unsigned long long))) + 256;
}

// raw_cams, raw_sgcs [batch, 21, height, width]; label [batch, 20]; loss_out float[1]; state: muscle_er_loss_state_bytes
// (kept for the backward).  No host synchronisation.
int muscle_er_loss_fwd(const float* raw_cams, const float* raw_sgcs, const float* label, int batch, int num_classes,
                       int height, int width, float* loss_out, void* state, size_t state_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100_er()) return rc;
  if (!raw_cams || !raw_sgcs || !label || !loss_out || !state) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (num_classes != ER_C) return fail(MUSCLE_ERR_BAD_ARG, "the ER loss is built for 21 classes (got %d)", num_classes);
  if (batch <= 0 || height <= 0 || width <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (state_bytes < muscle_er_loss_state_bytes(batch)) return fail(MUSCLE_ERR_WORKSPACE, "state too small");
  ErSel* sel = static_cast<ErSel*>(state);
  unsigned long long* hsum = reinterpret_cast<unsigned long long*>(sel + batch);     // (ErSel is a multiple of 8 bytes)
  unsigned int* hist = reinterpret_cast<unsigned int*>(hsum + static_cast<size_t>(batch) * ER_BINS);
  const long long hw = static_cast<long long>(height) * width;
  MB_CUDA(cudaMemsetAsync(state, 0, muscle_er_loss_state_bytes(batch), stream));
  MB_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float), stream));
  // blocks per image: ~4 per SM over the batch, and never so few that a block could put 2^16 values into one bin (its
  // 32-bit half sums, er_hist_kernel)
  long long gx_ll = std::min<long long>((hw + ER_THREADS - 1) / ER_THREADS, 4ll * device_info().sm_count / std::max(1, batch) + 1);
  gx_ll = std::max<long long>(gx_ll, (hw * ER_C + 64999) / 65000);
  const int gx = static_cast<int>(gx_ll);
  for (int level = 0; level < 3; ++level) {
    er_hist_kernel<<<dim3(gx, batch), ER_THREADS, 0, stream>>>(raw_cams, raw_sgcs, label, hw, level, sel, hist, hsum);
    MB_LAUNCH_CHECK();
    er_pick_kernel<<<batch, 256, 0, stream>>>(label, ER_C - 1, batch, hw, height, width, level, hist, hsum, sel,
                                              level == 2 ? loss_out : nullptr);
    MB_LAUNCH_CHECK();
  }
  return MUSCLE_OK;
}

int muscle_er_loss_bwd(const float* raw_cams, const float* raw_sgcs, const float* label, int batch, int num_classes,
                       int height, int width, const void* state, size_t state_bytes, const float* grad_scale,
                       float* dsgcs_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100_er()) return rc;
  if (!raw_cams || !raw_sgcs || !label || !state || !dsgcs_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (num_classes != ER_C) return fail(MUSCLE_ERR_BAD_ARG, "the ER loss is built for 21 classes (got %d)", num_classes);
  if (state_bytes < muscle_er_loss_state_bytes(batch)) return fail(MUSCLE_ERR_WORKSPACE, "state too small");
  const long long hw = static_cast<long long>(height) * width;
  const int gx = static_cast<int>(std::min<long long>((hw + ER_THREADS - 1) / ER_THREADS, 8ll * device_info().sm_count / std::max(1, batch) + 1));
  er_bwd_kernel<<<dim3(gx, batch), ER_THREADS, 0, stream>>>(raw_cams, raw_sgcs, label, hw, batch,
                                                            static_cast<const ErSel*>(state), grad_scale, dsgcs_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // extern "C"
