// The InfoNCE primitive itself, src/loss_multilabel.py:9-21, with distinct query / positive /
// negative sets:  pos_q = q . mean(P),  neg_qk = q . k,  loss_q = logsumexp([pos_q, neg_q.]/T) - pos_q/T,
// and its gradients to all three operands.  Built from the staged tcgen05 passes of the dense op
// (include/muscle_b200.h, "Staged form"): queries are "class 0" rows, negative keys "class 1"
// columns, so every (query, key) pair is live; the gradient pass runs twice, once per role
//   dQ = 1/T sum_k p_qk k + gpos_q proto        (rows = queries, w_row = 1/rowsum, w_col = 0)
//   dK = 1/T sum_q p_qk q                        (rows = keys,    w_row = 0, w_col = 1/rowsum)
// Like the reference, nothing is normalised here.  The pair passes assume |q.k| <= 1 (their exponent shift), so rows
// of any norm are handled by a uniform scale: with Rq >= max |q| and Rk >= max(|k|, |mean P|) (given by the caller),
//   q.k / T = (q/Rq).(k/Rk) / T'   with T' = T / (Rq Rk),
// the staged passes run on q/Rq and k/Rk at temperature T', and the gradients are scaled back by 1/Rq resp. 1/Rk.
// (muscle_infonce_prepare divides a row by max(|row|, eps): eps = the bound makes that division the uniform scale.)
// row_weight (optional, [Q]): the gradients become those of sum_q row_weight[q] loss_q -- the backward of
// reduction = 'none' with an upstream gradient per row.
#include <algorithm>
#include <cmath>
#include <cuda_bf16.h>

#include "host_util.cuh"
#include "pair_kernel.cuh"

using namespace mb200;

namespace {

size_t al(size_t v) { return (v + 255) / 256 * 256; }

struct QkWs {
  size_t eq, ek, bq, bk, tq, tk, inq, ink, pq, pk, segq, segk, cls, stats_tmp, cs, cm, rowsum, w, gpos, zk, dproto,
      lacc, uo_qk, uo_kq, part_q, part_k, scratch, zero_lo, zero_hi, total;
  int64_t qp, kp;
};
QkWs plan(int64_t Q, int64_t K, int dim) {
  QkWs w{};
  w.qp = muscle_pad_rows(Q);
  w.kp = muscle_pad_rows(K);
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off = al(off + b); return o; };
  const size_t qp = w.qp, kp = w.kp, mx = std::max(qp, kp);
  w.eq = take(qp * dim * 2); w.ek = take(kp * dim * 2);
  w.bq = take(qp * 4); w.bk = take(kp * 4);
  w.tq = take(qp / 128 * 4); w.tk = take(kp / 128 * 4);
  w.inq = take(qp * 4); w.ink = take(kp * 4);
  w.pq = take(qp * 4); w.pk = take(kp * 4);
  w.segq = take(16); w.segk = take(16);
  w.cls = take(mx * 4);
  w.uo_qk = take((qp / 128 + 1) * 4); w.uo_kq = take((kp / 128 + 1) * 4);
  w.w = take(qp * 4); w.gpos = take(qp * 4);
  w.zero_lo = off;                       // everything from here to zero_hi is cleared by one memset
  w.stats_tmp = take((2 * dim + 2) * 4);
  w.cs = take(2 * dim * 4); w.cm = take(8);
  w.rowsum = take(qp * 4); w.zk = take(mx * 4);
  w.dproto = take(2 * dim * 4); w.lacc = take(8);
  w.zero_hi = off;
  w.part_q = take(static_cast<size_t>(PAIR_SLOTS) * qp * dim * 4);
  w.part_k = take(static_cast<size_t>(PAIR_SLOTS) * kp * dim * 4);
  w.scratch = take(std::max(muscle_infonce_prepare_scratch_bytes(Q), muscle_infonce_prepare_scratch_bytes(K)));
  w.total = off;
  return w;
}

__global__ void fill_int_kernel(int* p, int v, long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
// class_sum[0][d] = scale * sum_p pos[p][d]; class_mass[0] = P
__global__ void __launch_bounds__(256) pos_sum_kernel(const float* pos, long long P, int dim, float scale, float* cs, float* cm) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < dim) {
    float s = 0.f;
    for (long long p = 0; p < P; ++p) s += pos[p * dim + d];
    cs[d] = s * scale;
  }
  if (d == 0) cm[0] = static_cast<float>(P);
}
// Per-row upstream gradient: w[q] *= g[q], gpos[q] *= g[q] (queries are all "class 0", so the class-sorted order of the
// staged passes is the caller's order), and dproto[0][:] = sum_q gpos[q] ehat[q][:] formed again with the weighted gpos.
__global__ void __launch_bounds__(128) qk_row_weight_kernel(const float* row_weight, long long Q, int dim,
                                                            const __nv_bfloat16* ehat, float* w, float* gpos, float* dproto) {
  const long long r0 = static_cast<long long>(blockIdx.x) * 64;
  const long long r1 = r0 + 64 < Q ? r0 + 64 : Q;
  __shared__ float gs[64];
  if (threadIdx.x < 64) {
    const long long r = r0 + threadIdx.x;
    float gv = 0.f;
    if (r < Q) {
      const float wt = row_weight[r];
      w[r] *= wt;
      gv = gpos[r] * wt;
      gpos[r] = gv;
    }
    gs[threadIdx.x] = gv;
  }
  __syncthreads();
  for (int d = threadIdx.x; d < dim; d += 128) {
    float acc = 0.f;
    for (long long r = r0; r < r1; ++r) acc = fmaf(gs[r - r0], __bfloat162float(ehat[r * dim + d]), acc);
    atomicAdd(dproto + d, acc);
  }
}

struct QkFinal {
  const float* part;       // [slots][rows_pad][dim]
  const uint32_t* unit_off;
  int n_row_blocks, n_col_tiles, n_ctas, rb;
  long long rows_pad, n;
  int dim;
  const float* gpos;       // may be null (key side)
  const float* proto_sum;  // class_sum[0] (query side)
  float inv_mass, inv_tau, scale;
  float* dst;              // [n][dim]
  const float* w;          // query side: 1/rowsum -> per-row loss
  float* row_loss;         // may be null
};
// one warp per row
__global__ void __launch_bounds__(256) qk_final_kernel(const QkFinal p) {
  const long long r = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= p.n) return;
  const int lane = threadIdx.x & 31;
  const uint32_t per = pair_units_per_cta(p.unit_off[p.n_row_blocks], p.n_ctas, p.n_col_tiles);
  const int rbk = static_cast<int>(r / p.rb);
  const uint32_t U0 = p.unit_off[rbk], U1 = p.unit_off[rbk + 1];
  const int n_slots = (U1 > U0) ? static_cast<int>((U1 - 1) / per - U0 / per + 1) : 0;
  const float g = p.gpos ? p.gpos[r] * p.inv_mass : 0.f;
  for (int d = lane; d < p.dim; d += 32) {
    float s = 0.f;
    for (int k = 0; k < n_slots; ++k) s += p.part[(static_cast<long long>(k) * p.rows_pad + r) * p.dim + d];
    float v = p.inv_tau * s;
    if (p.gpos) v = fmaf(g, p.proto_sum[d], v);
    p.dst[r * p.dim + d] = p.scale * v;
  }
  if (p.row_loss && lane == 0) p.row_loss[r] = -logf(fmaxf(1.f + p.gpos[r] / p.inv_tau, 1e-30f));
}
// dpos[p][d] = scale * key_scale * dproto[0][d] / P ;  loss = loss_acc * scale
__global__ void __launch_bounds__(256) qk_pos_final_kernel(const float* dproto, const double* lacc, long long P, int dim,
                                                           float scale, float key_scale, float* dpos, float* loss) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (dpos && i < P * dim) dpos[i] = scale * key_scale * dproto[i % dim] / static_cast<float>(P);
  if (i == 0 && loss) *loss = static_cast<float>(*lacc * static_cast<double>(scale));
}

}  // namespace

extern "C" {

size_t muscle_info_nce_workspace_bytes(int64_t n_query, int64_t n_negative, int dim) {
  if (n_query <= 0 || n_negative <= 0 || !(dim == 64 || dim == 128 || dim == 256)) return 0;
  return plan(n_query, n_negative, dim).total;
}

int muscle_info_nce_fwd_bwd(const float* query, const float* positive_keys, const float* negative_keys,
                            int64_t n_query, int64_t n_positive, int64_t n_negative, int dim, float temperature,
                            int reduction_mean, float* loss_out, float* row_loss_out, float* dquery_out,
                            float* dpositive_out, float* dnegative_out, void* workspace, size_t workspace_bytes,
                            void* stream_) {
  return muscle_info_nce_fwd_bwd_ex(query, positive_keys, negative_keys, n_query, n_positive, n_negative, dim, temperature,
                                    reduction_mean, 1.f, 1.f, nullptr, loss_out, row_loss_out, dquery_out, dpositive_out,
                                    dnegative_out, workspace, workspace_bytes, stream_);
}

int muscle_info_nce_fwd_bwd_ex(const float* query, const float* positive_keys, const float* negative_keys,
                               int64_t n_query, int64_t n_positive, int64_t n_negative, int dim, float temperature,
                               int reduction_mean, float query_norm_bound, float key_norm_bound, const float* row_weight,
                               float* loss_out, float* row_loss_out, float* dquery_out, float* dpositive_out,
                               float* dnegative_out, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!(query_norm_bound > 0.f) || !(key_norm_bound > 0.f) || !isfinite(query_norm_bound) || !isfinite(key_norm_bound))
    return fail(MUSCLE_ERR_BAD_ARG, "norm bounds must be positive and finite");
  if (row_weight && row_loss_out) return fail(MUSCLE_ERR_BAD_ARG, "row_loss_out is not available together with row_weight");
  const float Rq = query_norm_bound, Rk = key_norm_bound;
  const float user_temperature = temperature;
  temperature = temperature / (Rq * Rk);      // the staged passes see q/Rq, k/Rk (|q.k| <= 1) at this temperature
  if (!(temperature >= MUSCLE_MIN_TEMPERATURE))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature %g over norm bounds %g x %g is below the supported minimum %g",
                static_cast<double>(user_temperature), static_cast<double>(Rq), static_cast<double>(Rk), static_cast<double>(MUSCLE_MIN_TEMPERATURE));
  if (!query || !positive_keys || !negative_keys || !loss_out || !workspace)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (n_query <= 0 || n_positive <= 0 || n_negative <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (!(dim == 64 || dim == 128 || dim == 256)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (!(temperature > 0.f)) return fail(MUSCLE_ERR_BAD_ARG, "temperature must be > 0");
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return fail(MUSCLE_ERR_BAD_ARG, "workspace must be 256-byte aligned");
  const QkWs W = plan(n_query, n_negative, dim);
  if (workspace_bytes < W.total) return fail(MUSCLE_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes, W.total);
  char* b = static_cast<char*>(workspace);
  auto F = [&](size_t o) { return reinterpret_cast<float*>(b + o); };
  auto U = [&](size_t o) { return reinterpret_cast<uint32_t*>(b + o); };
  int* cls = reinterpret_cast<int*>(b + W.cls);
  const int64_t Q = n_query, K = n_negative;
  const bool need_grad = dquery_out || dnegative_out || dpositive_out;
  MB_CUDA(cudaMemsetAsync(b + W.zero_lo, 0, W.zero_hi - W.zero_lo, stream));
  // queries: class 0
  MB_CUDA(cudaMemsetAsync(cls, 0, Q * 4, stream));
  int rc = muscle_infonce_prepare(query, MUSCLE_LAYOUT_PIXEL_MAJOR, 1, Q, dim, cls, 0, 2, Rq, b + W.eq, U(W.bq),
                                  F(W.inq), reinterpret_cast<int32_t*>(b + W.pq), U(W.tq), U(W.segq), F(W.stats_tmp),
                                  F(W.stats_tmp) + 2 * dim, b + W.scratch, muscle_infonce_prepare_scratch_bytes(Q), stream_);
  if (rc) return rc;
  // negative keys: class 1
  fill_int_kernel<<<static_cast<unsigned>((K + 255) / 256), 256, 0, stream>>>(cls, 1, K);
  MB_LAUNCH_CHECK();
  rc = muscle_infonce_prepare(negative_keys, MUSCLE_LAYOUT_PIXEL_MAJOR, 1, K, dim, cls, 0, 2, Rk, b + W.ek, U(W.bk),
                              F(W.ink), reinterpret_cast<int32_t*>(b + W.pk), U(W.tk), U(W.segk), F(W.stats_tmp),
                              F(W.stats_tmp) + 2 * dim, b + W.scratch, muscle_infonce_prepare_scratch_bytes(K), stream_);
  if (rc) return rc;
  pos_sum_kernel<<<(dim + 255) / 256, 256, 0, stream>>>(positive_keys, n_positive, dim, 1.f / Rk, F(W.cs), F(W.cm));
  MB_LAUNCH_CHECK();
  if ((rc = muscle_infonce_plan(U(W.tq), W.qp, U(W.tk), W.kp, dim, U(W.uo_qk), stream_))) return rc;
  if ((rc = muscle_infonce_rowsum(b + W.eq, U(W.bq), U(W.tq), W.qp, b + W.ek, U(W.bk), U(W.tk), W.kp, U(W.uo_qk), dim,
                                  temperature, F(W.rowsum), stream_)))
    return rc;
  if ((rc = muscle_infonce_mid(b + W.eq, U(W.segq), W.qp, dim, 2, F(W.cs), F(W.cm), F(W.rowsum), temperature, F(W.w),
                               F(W.gpos), reinterpret_cast<double*>(b + W.lacc), F(W.dproto), stream_)))
    return rc;
  if (row_weight) {
    MB_CUDA(cudaMemsetAsync(b + W.dproto, 0, static_cast<size_t>(dim) * 4, stream));
    qk_row_weight_kernel<<<static_cast<unsigned>((Q + 63) / 64), 128, 0, stream>>>(
        row_weight, Q, dim, reinterpret_cast<const __nv_bfloat16*>(b + W.eq), F(W.w), F(W.gpos), F(W.dproto));
    MB_LAUNCH_CHECK();
  }
  const float scale = reduction_mean ? 1.f / static_cast<float>(Q) : 1.f;
  const int rb = PAIR_RB;
  const int n_ctas = device_info().sm_count;
  if (dquery_out || row_loss_out) {
    if (dquery_out)
      if ((rc = muscle_infonce_grad(b + W.eq, U(W.bq), U(W.tq), F(W.w), W.qp, b + W.ek, U(W.bk), U(W.tk), F(W.zk), W.kp,
                                    U(W.uo_qk), dim, temperature, F(W.part_q), stream_)))
        return rc;
    QkFinal f{};
    f.part = F(W.part_q); f.unit_off = U(W.uo_qk); f.n_row_blocks = static_cast<int>(W.qp / rb);
    f.n_col_tiles = static_cast<int>(W.kp / 128); f.n_ctas = n_ctas; f.rb = rb; f.rows_pad = W.qp; f.n = Q; f.dim = dim;
    f.gpos = F(W.gpos); f.proto_sum = F(W.cs); f.inv_mass = 1.f / static_cast<float>(n_positive);
    f.inv_tau = 1.f / temperature; f.scale = scale / Rq; f.dst = dquery_out; f.w = F(W.w); f.row_loss = row_loss_out;
    if (!dquery_out) { f.dim = 0; }
    qk_final_kernel<<<static_cast<unsigned>((Q + 7) / 8), 256, 0, stream>>>(f);
    MB_LAUNCH_CHECK();
  }
  if (dnegative_out) {
    if ((rc = muscle_infonce_plan(U(W.tk), W.kp, U(W.tq), W.qp, dim, U(W.uo_kq), stream_))) return rc;
    if ((rc = muscle_infonce_grad(b + W.ek, U(W.bk), U(W.tk), F(W.zk), W.kp, b + W.eq, U(W.bq), U(W.tq), F(W.w), W.qp,
                                  U(W.uo_kq), dim, temperature, F(W.part_k), stream_)))
      return rc;
    QkFinal f{};
    f.part = F(W.part_k); f.unit_off = U(W.uo_kq); f.n_row_blocks = static_cast<int>(W.kp / rb);
    f.n_col_tiles = static_cast<int>(W.qp / 128); f.n_ctas = n_ctas; f.rb = rb; f.rows_pad = W.kp; f.n = K; f.dim = dim;
    f.gpos = nullptr; f.inv_tau = 1.f / temperature; f.scale = scale / Rk; f.dst = dnegative_out;
    qk_final_kernel<<<static_cast<unsigned>((K + 7) / 8), 256, 0, stream>>>(f);
    MB_LAUNCH_CHECK();
  }
  (void)need_grad;
  const long long np = dpositive_out ? n_positive * dim : 1;
  qk_pos_final_kernel<<<static_cast<unsigned>((np + 255) / 256), 256, 0, stream>>>(
      F(W.dproto), reinterpret_cast<const double*>(b + W.lacc), n_positive, dim, scale, 1.f / Rk, dpositive_out, loss_out);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

}  // extern "C"
