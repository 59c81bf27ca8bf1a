// Classification losses of the MCL stage in one launch (include/muscle_b200.h, muscle_cls_losses_*): FocalLoss
// (src/loss_multilabel.py:68-91), Log_Sum_Exp_Pairwise_Loss (:24-33) and nn.MultiLabelSoftMarginLoss as
// train_mcl.py:180-183 composes them.  Launch/latency-bound work on [B, 20] logits: one warp per image, lane =
// class; the reference spends ~25 elementwise launches and materialises a [B, C, C] tensor for the pairwise term,
// which factorises:  sum_{i,j} exp(neg_j - pos_i) = (sum_j exp(neg_j)) (sum_i exp(-pos_i)).
#include "host_util.cuh"

using namespace mb200;

namespace {

constexpr int CLS_THREADS = 256;

__device__ __forceinline__ float warp_sum_c(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// mode 0: `in` = logits [B, in_stride] (class c at column in_off + c): focal(sigmoid) + soft margin + LSEP(sigmoid)
// mode 1: `in` = probabilities: focal only            mode 2: `in` = predictions: LSEP rows only
// out[0..3] = total, focal, soft margin, pair (means over the batch);  rows_out [B] (mode 2: the per-image values)
// din [B, in_stride] (columns outside [in_off, in_off + C) are zeroed) = d(sum_b up_b * term_b)/d in, up_b =
// row_grad[b] (mode 2) or grad_scale / B.
__global__ void __launch_bounds__(CLS_THREADS) cls_losses_kernel(const float* __restrict__ in, const float* __restrict__ label,
                                                                 int B, int C, int in_stride, int in_off, int mode,
                                                                 float gamma, float alpha, const float* __restrict__ row_grad,
                                                                 const float* __restrict__ grad_scale,
                                                                 float* __restrict__ out, float* __restrict__ rows_out,
                                                                 float* __restrict__ din) {
  __shared__ float red[3][CLS_THREADS / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float gs = grad_scale ? __ldg(grad_scale) : 1.f;
  const float inv_b = 1.f / static_cast<float>(B);
  float s_focal = 0.f, s_soft = 0.f, s_pair = 0.f;
  for (int b = warp; b < B; b += CLS_THREADS / 32) {
    const bool on = lane < C;
    const float x = on ? __ldg(in + static_cast<size_t>(b) * in_stride + in_off + lane) : 0.f;
    const float y = on ? __ldg(label + static_cast<size_t>(b) * C + lane) : 0.f;
    const float p = (mode == 0) ? 1.f / (1.f + __expf(-x)) : x;
    float dp = 0.f, dx = 0.f;      // d / d probability, d / d logit (soft margin)
    if (mode != 2) {               // focal (:86-90)
      const float pt = y * p + (1.f - y) * (1.f - p);
      const float om = 1.f - pt;
      const float lg = logf(pt + 1e-9f);
      const float pw = (gamma == 2.f) ? om * om : powf(om, gamma);
      const float pw1 = (gamma == 2.f) ? om : powf(om, gamma - 1.f);
      const float f = on ? -alpha * pw * lg : 0.f;
      s_focal += warp_sum_c(f);
      // df/dpt = -alpha (-gamma om^(gamma-1) lg + om^gamma / (pt + 1e-9)),  dpt/dp = 2y - 1
      if (on) dp += -alpha * (-gamma * pw1 * lg + pw / (pt + 1e-9f)) * (2.f * y - 1.f);
    }
    if (mode == 0) {               // soft margin: mean over classes of -(y log s(x) + (1-y) log s(-x))
      const float sp_neg = fmaxf(-x, 0.f) + log1pf(__expf(-fabsf(x)));   // softplus(-x) = -log sigmoid(x)
      const float sp_pos = sp_neg + x;                                   // softplus(x)  = -log sigmoid(-x)
      const float t = on ? (y * sp_neg + (1.f - y) * sp_pos) / static_cast<float>(C) : 0.f;
      s_soft += warp_sum_c(t);
      if (on) dx += (p - y) / static_cast<float>(C);
    }
    float up_pair = 0.f;
    if (mode != 1) {               // LSEP (:24-33): zeros of the masked entries take part in the pairs
      const float pos = (y == 0.f) ? 0.f : p;
      const float neg = (y == 1.f) ? 0.f : p;
      const float a = warp_sum_c(on ? __expf(neg) : 0.f);
      const float bb = warp_sum_c(on ? __expf(-pos) : 0.f);
      const float cc = 1.f / (static_cast<float>(C) * static_cast<float>(C));
      const float es = a * bb * cc;
      const float val = logf(1.f + es);
      s_pair += val;
      if (rows_out != nullptr && lane == 0) rows_out[b] = val;
      up_pair = (mode == 2) ? (row_grad ? __ldg(row_grad + b) : 0.f) : gs * inv_b;
      if (on) {
        const float dneg = __expf(neg) * bb * cc / (1.f + es);
        const float dpos = -a * __expf(-pos) * cc / (1.f + es);
        const float d = ((y != 1.f) ? dneg : 0.f) + ((y != 0.f) ? dpos : 0.f);
        if (mode == 2) dp = d * up_pair;
        else dp = dp * gs * inv_b + d * up_pair;
      }
    } else if (on) {
      dp *= gs * inv_b;
    }
    if (din != nullptr) {
      for (int c = lane; c < in_stride; c += 32)       // columns outside the class range (the background logit)
        if (c < in_off || c >= in_off + C) din[static_cast<size_t>(b) * in_stride + c] = 0.f;
      if (on) {
        const float g = (mode == 0) ? dp * p * (1.f - p) + dx * gs * inv_b : dp;
        din[static_cast<size_t>(b) * in_stride + in_off + lane] = g;
      }
    }
  }
  if (lane == 0) { red[0][warp] = s_focal; red[1][warp] = s_soft; red[2][warp] = s_pair; }
  __syncthreads();
  if (threadIdx.x == 0 && out != nullptr) {
    float f = 0.f, s = 0.f, pr = 0.f;
    for (int w = 0; w < CLS_THREADS / 32; ++w) { f += red[0][w]; s += red[1][w]; pr += red[2][w]; }
    f *= inv_b; s *= inv_b; pr *= inv_b;
    out[0] = f + s + pr; out[1] = f; out[2] = s; out[3] = pr;
  }
}

int check_cls(const float* in, const float* label, int batch, int num_fg, int in_stride, int in_off) {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  if (!in || !label) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (batch <= 0 || num_fg <= 0 || num_fg > 32) return fail(MUSCLE_ERR_BAD_ARG, "batch > 0 and 1 <= classes <= 32");
  if (in_off < 0 || in_off + num_fg > in_stride) return fail(MUSCLE_ERR_BAD_ARG, "class columns outside the row");
  return MUSCLE_OK;
}

}  // namespace

extern "C" {

int muscle_cls_losses_fwd_bwd(const float* logits, const float* label, int batch, int num_fg, int logit_stride,
                              int logit_offset, float gamma, float alpha, const float* grad_scale, float* losses_out,
                              float* dlogits_out, void* stream_) {
  if (int rc = check_cls(logits, label, batch, num_fg, logit_stride, logit_offset)) return rc;
  if (!losses_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  cls_losses_kernel<<<1, CLS_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(logits, label, batch, num_fg, logit_stride,
                                                                                logit_offset, 0, gamma, alpha, nullptr,
                                                                                grad_scale, losses_out, nullptr, dlogits_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_focal_loss_fwd_bwd(const float* prob, const float* target, int batch, int num_classes, float gamma, float alpha,
                              const float* grad_scale, float* loss_out4, float* dprob_out, void* stream_) {
  if (int rc = check_cls(prob, target, batch, num_classes, num_classes, 0)) return rc;
  if (!loss_out4) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  cls_losses_kernel<<<1, CLS_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(prob, target, batch, num_classes, num_classes,
                                                                                0, 1, gamma, alpha, nullptr, grad_scale,
                                                                                loss_out4, nullptr, dprob_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_lsep_fwd_bwd(const float* pred, const float* labels, int batch, int num_classes, const float* row_grad,
                        float* rows_out, float* dpred_out, void* stream_) {
  if (int rc = check_cls(pred, labels, batch, num_classes, num_classes, 0)) return rc;
  cls_losses_kernel<<<1, CLS_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(pred, labels, batch, num_classes, num_classes,
                                                                                0, 2, 2.f, 0.5f, row_grad, nullptr, nullptr,
                                                                                rows_out, dpred_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // extern "C"
