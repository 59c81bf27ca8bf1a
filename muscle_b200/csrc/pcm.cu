// Projection head + PCM (src/MuSCLe.py:181,213-223) as ReLU-kernel attention on the tensor pipe.
//
//   g = fuse(f) (1x1 conv Cin -> 128, bias)      pcm_project_kernel   D[d,px] = W[d,c] F[c,px]
//   ghat = g / (||g|| + 1e-5)                      (its epilogue: transpose through smem, bf16 rows)
//   S = relu(ghat^T ghat)  [P x P] per image       pcm_pair_kernel<FWD>  S_h = Q K_h^T      (SS MMA)
//   out = cam S / (colsum S + 1e-5)                                      O += relu(S_h) V_h (TS MMA)
// V = cam as bf16 [32 rows, P] with row 31 = 1, so the column sums of S come out of the same MMA
// as the numerator (S is symmetric: "columns" of the reference are rows here).  The P x P
// affinity never reaches HBM (the reference keeps three fp32 copies of it for backward).
//
// Backward (dOut given; A = dOut / den, r_j = sum_c out_cj A_cj):
//   dcam = A S                                     the forward kernel again with V = A, no division
//   dS_ij = sum_c cam_ci A_cj - r_j  =>  dghat_i = sum_j [z_ij > 0] (U_ij - r_i - r_j) ghat_j,
//   U = cam^T A + A^T cam                          pcm_pair_kernel<BWD>: Z_h = Q K_h^T, W_h = X Y_h^T
//     (U - r is a deviation from a weighted mean: bf16 cam/A lose it, so both are split hi + lo)
//                                                  X_i = [ch ch cl | ah al ah]_i, Y_j = [ah al ah | ch ch cl]_j
//                                                  (K = 6 x 21 -> 128);  r_i, r_j added in fp32
//                                                  T_h = [Z_h > 0] (W_h - r_i - r_j) -> bf16 -> G += T_h K_h
//   dg from dghat through the normalisation (kernel epilogue), stored d-major bf16
//   dW = dg f^T, db = dg 1                          pcm_dw_kernel (f carries a row of ones)
#include <algorithm>
#include <cstdlib>
#include <cuda_bf16.h>

#include "host_util.cuh"
#include "ptx.cuh"

using namespace mb200;

namespace {

constexpr int PD = 128;       // projection width, hard-coded in the reference (src/MuSCLe.py:181)
constexpr int VROWS = 32;     // CAM classes padded; row 31 carries the ones of the column sum
constexpr float PCM_EPS = 1e-5f;
constexpr int XSEG = 21;      // classes per segment of the split X/Y operands of the backward pass
constexpr int XK = 128;       // their K extent (6 segments of 21, 2 zero columns)
constexpr int PANEL = 128 * 128;  // bytes of one [128 rows x 64 bf16] swizzle panel

int need_sm100() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10) return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}

inline int64_t pad128(int64_t p) { return (p + 127) / 128 * 128; }
inline int cin_pad(int cin) { return (cin + 1 + 63) / 64 * 64; }  // room for the row of ones
inline size_t al256(size_t v) { return (v + 255) / 256 * 256; }

struct PcmState {   // offsets into the forward state buffer (kept for backward)
  size_t fb, wb, ghat, gnorm, camb, den, total;
};
PcmState plan_state(int batch, int cin, int64_t P) {
  const int64_t pp = pad128(P);
  const int cp = cin_pad(cin);
  PcmState s{};
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off = al256(off + b); return o; };
  s.fb = take(static_cast<size_t>(batch) * cp * pp * 2);
  s.wb = take(static_cast<size_t>(PD) * cp * 2);
  s.ghat = take(static_cast<size_t>(batch) * pp * PD * 2);
  s.gnorm = take(static_cast<size_t>(batch) * pp * 4);
  s.camb = take(static_cast<size_t>(batch) * VROWS * pp * 2);
  s.den = take(static_cast<size_t>(batch) * pp * 4);
  s.total = off;
  return s;
}
struct PcmScratch {
  size_t ab, xp, yp, rr, dgt, dwp, total;
};
PcmScratch plan_scratch(int batch, int cin, int64_t P) {
  const int64_t pp = pad128(P);
  const int cp = cin_pad(cin);
  PcmScratch s{};
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off = al256(off + b); return o; };
  s.ab = take(static_cast<size_t>(batch) * VROWS * pp * 2);
  s.xp = take(static_cast<size_t>(batch) * pp * XK * 2);
  s.yp = take(static_cast<size_t>(batch) * pp * XK * 2);
  s.rr = take(static_cast<size_t>(batch) * pp * 4);
  s.dgt = take(static_cast<size_t>(batch) * PD * pp * 2);
  s.dwp = take(static_cast<size_t>(PD) * cp * 4);
  s.total = off;
  return s;
}

// ------------------------------------------------------------------------------------------
// element-wise preparation (HBM bound)
// fb[b][c][p] = bf16(f[b][c][p]) for c < cin, 1 for c == cin (p < P), 0 elsewhere.  Eight pixels per thread (two
// 16-byte loads, one 16-byte store) when the rows of f are 16-byte aligned (P % 4 == 0).
__global__ void __launch_bounds__(256) pcm_cast_f_kernel(const float* f, int cin, int cp, int64_t P, int64_t pp,
                                                         __nv_bfloat16* fb) {
  const int64_t row = blockIdx.x;  // b * cp + c
  const int b = static_cast<int>(row / cp), c = static_cast<int>(row - static_cast<int64_t>(b) * cp);
  const int64_t p = (static_cast<int64_t>(blockIdx.y) * blockDim.x + threadIdx.x) * 8;
  if (p >= pp) return;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = 0.f;
  if (c < cin) {
    const float* src = f + (static_cast<int64_t>(b) * cin + c) * P;
    if ((P & 3) == 0 && p + 8 <= P) {
      const float4 lo = __ldg(reinterpret_cast<const float4*>(src + p));
      const float4 hi = __ldg(reinterpret_cast<const float4*>(src + p) + 1);
      v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w; v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (p + j < P) v[j] = __ldg(src + p + j);
    }
  } else if (c == cin) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (p + j < P) ? 1.f : 0.f;
  }
  *reinterpret_cast<uint4*>(fb + row * pp + p) =
      make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}
__global__ void __launch_bounds__(256) pcm_cast_w_kernel(const float* w, int cin, int cp, __nv_bfloat16* wb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= PD * cp) return;
  const int d = i / cp, c = i - d * cp;
  wb[i] = __float2bfloat16(c < cin ? w[d * cin + c] : 0.f);
}
// vb[b][k][p] = bf16(src[b][k][p] * (inv_den ? 1/den[b][p] : 1)) for k < K; row 31 = ones ? 1 : 0; pad = 0.
// grid (pixels / 256, VROWS, batch): one element per thread (the per-pixel loop over the 32 rows this replaced ran 104
// blocks and took 14 us for 3.7 MB).
__global__ void __launch_bounds__(256) pcm_cast_v_kernel(const float* src, const float* den, int K, int64_t P,
                                                         int64_t pp, int ones, __nv_bfloat16* vb) {
  const int b = blockIdx.z, k = blockIdx.y;
  const int64_t p = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (p >= pp) return;
  const bool in = p < P;
  float v = 0.f;
  if (in && k < K) {
    v = __ldg(src + (static_cast<int64_t>(b) * K + k) * P + p);
    if (den) v *= 1.f / __ldg(den + static_cast<int64_t>(b) * pp + p);
  }
  if (in && ones && k == VROWS - 1) v = 1.f;
  vb[(static_cast<int64_t>(b) * VROWS + k) * pp + p] = __float2bfloat16(v);
}
// Per pixel: A = dOut/den, r = sum_c out A; A rows -> ab (V operand of the dcam pass);
// X = [ch ch cl | ah al ah], Y = [ah al ah | ch ch cl] in segments of 21 (128 bf16 each), with
// cam = ch + cl and A = ah + al split into bf16 hi + lo parts; rr = r (fp32).
// Block = 32 pixels x 8 warps: warp w takes the classes w, w + 8, w + 16, w + 24 (lane = pixel: every global read and
// the ab store are coalesced), the X / Y rows are assembled in shared memory and leave as whole 256-byte rows.  (One
// thread per pixel -- 25 600 threads in all at config 2, 63 loads and 32 scattered 16-byte stores each -- took 20 us.)
constexpr int PREP_PX = 32;
constexpr int PREP_ROW_W = XK / 2 + 1;      // shared row stride in 32-bit words (odd: conflict-free)
__global__ void __launch_bounds__(256) pcm_bwd_prep_kernel(const float* dout, const float* out, const float* cam,
                                                           const float* den, int K, int64_t P, int64_t pp,
                                                           __nv_bfloat16* ab, __nv_bfloat16* xp, __nv_bfloat16* yp,
                                                           float* rr) {
  __shared__ uint32_t xs_w[PREP_PX * PREP_ROW_W], ys_w[PREP_PX * PREP_ROW_W];
  __shared__ float rs[8][PREP_PX];
  const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t p0 = static_cast<int64_t>(blockIdx.x) * PREP_PX, p = p0 + lane;     // pp % 128 == 0: p < pp
  const bool in = p < P;
  __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(xs_w) + lane * (2 * PREP_ROW_W);
  __nv_bfloat16* ys = reinterpret_cast<__nv_bfloat16*>(ys_w) + lane * (2 * PREP_ROW_W);
  if (warp == 0) {                                     // the two columns past the six segments
    xs[6 * XSEG] = xs[6 * XSEG + 1] = __float2bfloat16(0.f);
    ys[6 * XSEG] = ys[6 * XSEG + 1] = __float2bfloat16(0.f);
  }
  const float inv = in ? 1.f / __ldg(den + static_cast<int64_t>(b) * pp + p) : 0.f;
  float r = 0.f;
#pragma unroll
  for (int k = warp; k < VROWS; k += 8) {
    float a = 0.f, c = 0.f;
    if (in && k < K) {
      const int64_t idx = (static_cast<int64_t>(b) * K + k) * P + p;
      a = __ldg(dout + idx) * inv;
      c = __ldg(cam + idx);
      r = fmaf(__ldg(out + idx), a, r);
    }
    const __nv_bfloat16 ah = __float2bfloat16(a), ch = __float2bfloat16(c);
    ab[(static_cast<int64_t>(b) * VROWS + k) * pp + p] = ah;
    if (k < XSEG) {
      const __nv_bfloat16 al = __float2bfloat16(a - __bfloat162float(ah));
      const __nv_bfloat16 cl = __float2bfloat16(c - __bfloat162float(ch));
      xs[k] = ch; xs[XSEG + k] = ch; xs[2 * XSEG + k] = cl; xs[3 * XSEG + k] = ah; xs[4 * XSEG + k] = al; xs[5 * XSEG + k] = ah;
      ys[k] = ah; ys[XSEG + k] = al; ys[2 * XSEG + k] = ah; ys[3 * XSEG + k] = ch; ys[4 * XSEG + k] = ch; ys[5 * XSEG + k] = cl;
    }
  }
  rs[warp][lane] = r;
  __syncthreads();
  if (warp == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += rs[w][lane];
    rr[static_cast<int64_t>(b) * pp + p] = t;
  }
  // rows out: warp w writes rows w, w + 8, ...; a row is 64 words = two coalesced 128-byte stores
  for (int row = warp; row < PREP_PX; row += 8) {
    uint32_t* xd = reinterpret_cast<uint32_t*>(xp + (static_cast<int64_t>(b) * pp + p0 + row) * XK);
    uint32_t* yd = reinterpret_cast<uint32_t*>(yp + (static_cast<int64_t>(b) * pp + p0 + row) * XK);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      xd[lane + 32 * h] = xs_w[row * PREP_ROW_W + lane + 32 * h];
      yd[lane + 32 * h] = ys_w[row * PREP_ROW_W + lane + 32 * h];
    }
  }
}
// dw[d][c] = dwp[d][c] (c < cin), db[d] = dwp[d][cin]
__global__ void __launch_bounds__(256) pcm_unpad_dw_kernel(const float* dwp, int cin, int cp, float* dw, float* db) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < PD * cin) {
    const int d = i / cin, c = i - d * cin;
    dw[i] = dwp[d * cp + c];
  }
  if (i < PD) db[i] = dwp[i * cp + cin];
}

// ------------------------------------------------------------------------------------------
// Common CTA prologue for the tcgen05 kernels: 1024-aligned dynamic smem, TMEM allocation.
struct CtaCtx {
  uint32_t base;   // shared-space address of the aligned region
  uint8_t* gen;    // generic pointer to the same byte
  uint32_t tmem;
};

__device__ __forceinline__ void named_bar_sync(int id, int n) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}

// ==========================================================================================
// projection: ghat[b][p][:] = normalise(W f[b][:, p] + bias), gnorm[b][p] = ||g||
// ==========================================================================================
constexpr int PJ_NST = 4;
constexpr int PJ_STAGE = 2 * 64 * 128;  // two [64 c-rows x 64 px] panels
constexpr int PJ_TRW = 133;             // padded row of the transpose buffer (floats)
constexpr int PJ_THREADS = 192;

__host__ __device__ inline int pj_smem_bytes(int kc) {
  return kc * PANEL + PJ_NST * PJ_STAGE + 32 * PJ_TRW * 4 + 256 + 1024;
}

__global__ void __launch_bounds__(PJ_THREADS, 1)
pcm_project_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_f,
                   const float* bias, int kc, int cp, int tiles_per_image, int n_tiles, int64_t P, int64_t pp,
                   __nv_bfloat16* ghat, float* gnorm) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t w_smem = base;
  const uint32_t f_smem = base + kc * PANEL;
  const int tr_off = kc * PANEL + PJ_NST * PJ_STAGE;
  float* tr = reinterpret_cast<float*>(gen + tr_off);
  const uint32_t bar0 = base + tr_off + 32 * PJ_TRW * 4;
  const uint32_t w_full = bar0;
  auto f_full = [&](int s) { return bar0 + 8u * (1 + s); };
  auto f_empty = [&](int s) { return bar0 + 8u * (1 + PJ_NST + s); };
  auto d_full = [&](int a) { return bar0 + 8u * (1 + 2 * PJ_NST + a); };
  auto d_empty = [&](int a) { return bar0 + 8u * (3 + 2 * PJ_NST + a); };
  volatile uint32_t* tptr = reinterpret_cast<volatile uint32_t*>(gen + tr_off + 32 * PJ_TRW * 4 + 8 * (5 + 2 * PJ_NST));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(w_full, 1);
    for (int s = 0; s < PJ_NST; ++s) { mbar_init(f_full(s), 1); mbar_init(f_empty(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(d_full(a), 1); mbar_init(d_empty(a), 4); }
    fence_barrier_init();
  }
  if (warp == 4) { tmem_alloc(smem_u32(const_cast<uint32_t*>(tptr)), 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tptr;

  if (warp == 5) {
    if (lane == 0) {
      tma_prefetch_desc(&tm_w);
      tma_prefetch_desc(&tm_f);
      mbar_arrive_expect_tx(w_full, kc * PANEL);
      for (int k = 0; k < kc; ++k) tma_load_2d(w_smem + k * PANEL, &tm_w, w_full, k * 64, 0);
      uint32_t cnt = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = tile / tiles_per_image, p0 = (tile - b * tiles_per_image) * 128;
        for (int k = 0; k < kc; ++k, ++cnt) {
          const int st = cnt % PJ_NST;
          mbar_wait_spin(f_empty(st), ((cnt / PJ_NST) & 1) ^ 1);
          mbar_arrive_expect_tx(f_full(st), PJ_STAGE);
          for (int h = 0; h < 2; ++h)
            tma_load_2d(f_smem + st * PJ_STAGE + h * 8192, &tm_f, f_full(st), p0 + h * 64, b * cp + k * 64);
        }
      }
    }
  } else if (warp == 4) {
    if (elect_one()) {
      constexpr uint32_t IDESC = umma_idesc_bf16(128, 128, 0, 1);  // A = W K-major, B = f MN-major
      mbar_wait_spin(w_full, 0);
      tc_fence_after();
      uint32_t cnt = 0, it = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        mbar_wait_spin(d_empty(acc), ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int k = 0; k < kc; ++k, ++cnt) {
          const int st = cnt % PJ_NST;
          mbar_wait_spin(f_full(st), (cnt / PJ_NST) & 1);
          tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            mma_ss(tmem + acc * 128, umma_desc_sw128(w_smem + k * PANEL + kk * 32, 16, 1024),
                   umma_desc_sw128(f_smem + st * PJ_STAGE + kk * 2048, 8192, 1024), IDESC, (k | kk) != 0);
          tc_commit(f_empty(st));
        }
        tc_commit(d_full(acc));
      }
    }
  } else {
    // epilogue: thread = output channel d (TMEM lane); 32 pixels at a time through `tr`
    const int d = threadIdx.x;
    const float bd = bias[d];
    const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
    const int px_l = threadIdx.x >> 2, q = threadIdx.x & 3;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const int b = tile / tiles_per_image, p0 = (tile - b * tiles_per_image) * 128;
      mbar_wait_warp(d_full(acc), (it >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c4 = 0; c4 < 4; ++c4) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_addr + acc * 128 + c4 * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int k = 0; k < 32; ++k) tr[k * PJ_TRW + warp * 33 + lane] = __uint_as_float(v[k]) + bd;
        named_bar_sync(1, 128);
        float g[32];
        float ss = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) { g[k] = tr[px_l * PJ_TRW + q * 33 + k]; ss = fmaf(g[k], g[k], ss); }
        ss += __shfl_xor_sync(0xffffffffu, ss, 1);
        ss += __shfl_xor_sync(0xffffffffu, ss, 2);
        const float n = sqrtf(ss);
        const int64_t p = p0 + c4 * 32 + px_l;
        const bool in = p < P;
        const float inv = in ? 1.f / (n + PCM_EPS) : 0.f;
        uint32_t pk[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) pk[k] = pack_bf16x2(g[2 * k] * inv, g[2 * k + 1] * inv);
        uint4* dst = reinterpret_cast<uint4*>(ghat + (static_cast<int64_t>(b) * pp + p) * PD + q * 32);
#pragma unroll
        for (int k = 0; k < 4; ++k) dst[k] = make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
        if (q == 0) gnorm[static_cast<int64_t>(b) * pp + p] = in ? n : 0.f;
        named_bar_sync(1, 128);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(d_empty(acc));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 256);
}

// ==========================================================================================
// pair passes over one image: work item = (image, 128-pixel row tile); loops over all column tiles
//
// Every column tile is two HALF units of 64 columns.  A half unit u runs through
//   S(u) [+ W(u)]  (SS products into TMEM buffer u % HB)  ->  epilogue group u % NG: relu / mask, bf16, tcgen05.st over the
//   first half of the same buffer's columns  ->  O += T(u) V(u)  (TS product)  ->  the buffer takes S(u + HB).
// A unit is a whole column tile forward and a 64-column half backward (PpCfg::UC).  The issuer keeps HB products in
// flight and NG epilogue groups of four warps alternate units.  ncu (profiles/r2z_*): the issuer never waits for an
// epilogue (its t_full polls succeed at once) and the tcgen05.mma queue is full (stall_mio) at 32 % pipe activity: the
// passes are bound by the NUMBER of tcgen05.mma per tile (~100 cycles each here, whatever N), which is why whole-tile
// S products (8 instead of 16 per tile) gave 58 -> 44 us forward while deeper buffering alone gave nothing.  MMAs
// complete in issue order, which is what makes the parity waits safe with NG <= HB: when a group has seen S(u) it has
// seen every earlier phase of every buffer.
// ==========================================================================================
enum PcmMode { PCM_FWD = 0, PCM_BWD = 1 };
constexpr int PP_THREADS = 320;   // warps 0-3, 4-7: epilogue groups; 8: MMA issuer + TMEM allocator; 9: TMA producer

template <int MODE, int HB_ = 3, int NG_ = 2>
struct PpCfg {
  // A unit = UC columns of a column tile: the whole tile forward (one N = 128 product for S: a tcgen05.mma costs about
  // the same pipe time at N = 64 as at N = 128 here, so half as many is nearly twice as fast), a 64-column half backward
  // (S and W buffers + the 128-column accumulator leave no room for whole tiles).
  static constexpr int UC = (MODE == PCM_BWD) ? 64 : 128;
  static constexpr int UPT = 128 / UC;                      // units per column tile
  static constexpr int HB = HB_;                            // UC-column S (and W) buffers in TMEM
  static constexpr int NG = NG_;                            // epilogue groups of four warps (warps 4 NG .. 7 idle)
  static_assert(NG <= HB, "a group must have seen the previous phase of every buffer it waits on");
  static constexpr int NST = (MODE == PCM_BWD) ? 2 : 4;     // column-tile ring
  static constexpr int Q_BYTES = 2 * PANEL + (MODE == PCM_BWD ? 2 * PANEL : 0);      // ghat row tile (+ X tile)
  static constexpr int AUX_BYTES = (MODE == PCM_BWD) ? 2 * PANEL : 2 * VROWS * 128;  // Y tile | two V halves
  static constexpr int TX_BYTES = 2 * PANEL + AUX_BYTES + (MODE == PCM_BWD ? 512 : 0);  // + r of the column pixels
  static constexpr int STAGE = 2 * PANEL + AUX_BYTES + (MODE == PCM_BWD ? 1024 : 0);
  static constexpr int SMEM_K = Q_BYTES;
  static constexpr int SMEM_BAR = SMEM_K + NST * STAGE;
  static constexpr int SMEM_BYTES = SMEM_BAR + 256 + 1024;
  static constexpr int NBAR = 2 + 2 * NST + 2 * HB;
  static constexpr int W_COL = HB * UC;                                         // BWD: W buffers
  static constexpr int ACC_COL = (MODE == PCM_BWD) ? 2 * HB * UC : HB * UC;     // 384 | 384
  static constexpr int ACC_N = (MODE == PCM_BWD) ? PD : VROWS;
  static_assert(ACC_COL + ACC_N <= 512, "TMEM budget");
  static_assert(NBAR * 8 + 8 <= 256, "barrier block");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

struct PpParams {
  int tiles_per_image;
  int64_t P, pp;
  int K;                  // CAM classes
  int normalise;          // FWD: divide by (O[31] + eps) and save den
  float* out;             // FWD: [B][K][P]
  float* den;             // FWD normalise: [B][pp]
  const __nv_bfloat16* ghat;  // BWD epilogue
  const float* gnorm;         // BWD epilogue
  const float* rr;            // BWD: [B][pp] r of every pixel
  __nv_bfloat16* dgt;         // BWD: [B][128][pp]
};

template <int MODE, int HB_, int NG_>
__global__ void __launch_bounds__(PP_THREADS, 1)
pcm_pair_kernel(const __grid_constant__ CUtensorMap tm_g, const __grid_constant__ CUtensorMap tm_aux_row,
                const __grid_constant__ CUtensorMap tm_aux_col, const PpParams p) {
  using C = PpCfg<MODE, HB_, NG_>;
  constexpr int NG = NG_;
  constexpr int PP_NST = C::NST, HB = C::HB, UC = C::UC, UPT = C::UPT;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t q_smem = base, x_smem = base + 2 * PANEL, k_smem = base + C::SMEM_K;
  const uint32_t bar0 = base + C::SMEM_BAR;
  const uint32_t q_full = bar0;
  auto k_full = [&](int s) { return bar0 + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bar0 + 8u * (1 + PP_NST + s); };
  auto s_full = [&](int b) { return bar0 + 8u * (1 + 2 * PP_NST + b); };
  auto t_full = [&](int b) { return bar0 + 8u * (1 + 2 * PP_NST + HB + b); };
  const uint32_t o_full = bar0 + 8u * (1 + 2 * PP_NST + 2 * HB);
  volatile uint32_t* tptr = reinterpret_cast<volatile uint32_t*>(gen + C::SMEM_BAR + 8 * C::NBAR);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = p.tiles_per_image;
  const int U = UPT * T;                                     // units of this item
  const int b = blockIdx.x / T, rt = blockIdx.x - b * T;
  const int row0 = static_cast<int>(b * p.pp) + rt * 128;   // row of the [B*pp, *] pixel-major operands

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < PP_NST; ++s) { mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); }
    for (int h = 0; h < HB; ++h) { mbar_init(s_full(h), 1); mbar_init(t_full(h), 4); }
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 8) { tmem_alloc(smem_u32(const_cast<uint32_t*>(tptr)), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tptr;

  if (warp == 9) {
    if (lane == 0) {
      tma_prefetch_desc(&tm_g);
      tma_prefetch_desc(&tm_aux_col);
      mbar_arrive_expect_tx(q_full, C::Q_BYTES);
      tma_load_2d(q_smem, &tm_g, q_full, 0, row0);
      tma_load_2d(q_smem + PANEL, &tm_g, q_full, 64, row0);
      if (MODE == PCM_BWD) {
        tma_load_2d(x_smem, &tm_aux_row, q_full, 0, row0);
        tma_load_2d(x_smem + PANEL, &tm_aux_row, q_full, 64, row0);
      }
      for (int t = 0; t < T; ++t) {
        const int st = t % PP_NST;
        mbar_wait_spin(k_empty(st), ((t / PP_NST) & 1) ^ 1);
        mbar_arrive_expect_tx(k_full(st), C::TX_BYTES);
        const uint32_t ks = k_smem + st * C::STAGE;
        const int col_row = static_cast<int>(b * p.pp) + t * 128;
        tma_load_2d(ks, &tm_g, k_full(st), 0, col_row);
        tma_load_2d(ks + PANEL, &tm_g, k_full(st), 64, col_row);
        if (MODE == PCM_BWD) {
          tma_load_2d(ks + 2 * PANEL, &tm_aux_col, k_full(st), 0, col_row);
          tma_load_2d(ks + 3 * PANEL, &tm_aux_col, k_full(st), 64, col_row);
          bulk_load_1d(ks + 4 * PANEL, p.rr + col_row, 512, k_full(st));
        } else {
          for (int h = 0; h < 2; ++h)
            tma_load_2d(ks + 2 * PANEL + h * VROWS * 128, &tm_aux_col, k_full(st), t * 128 + h * 64, b * VROWS);
        }
      }
    }
  } else if (warp == 8) {
    if (elect_one()) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, UC, 0, 0);
      constexpr uint32_t IDESC_O = (MODE == PCM_BWD) ? umma_idesc_bf16(128, PD, 0, 1) : umma_idesc_bf16(128, VROWS, 0, 0);
      mbar_wait_spin(q_full, 0);
      tc_fence_after();
      auto issue_S = [&](int u) {
        const int t = u / UPT, h = u % UPT, st = t % PP_NST, hb = u % HB;
        mbar_wait_spin(k_full(st), (t / PP_NST) & 1);
        tc_fence_after();
        const uint32_t ks = k_smem + st * C::STAGE;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * PANEL + (kk & 3) * 32;
          mma_ss(tmem + hb * UC, umma_desc_sw128(q_smem + off, 16, 1024),
                 umma_desc_sw128(ks + h * 8192 + off, 16, 1024), IDESC_S, kk > 0);
        }
        if (MODE == PCM_BWD) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t off = (kk >> 2) * PANEL + (kk & 3) * 32;
            mma_ss(tmem + C::W_COL + hb * UC, umma_desc_sw128(x_smem + off, 16, 1024),
                   umma_desc_sw128(ks + 2 * PANEL + h * 8192 + off, 16, 1024), IDESC_S, kk > 0);
          }
        }
        tc_commit(s_full(hb));
      };
      for (int u = 0; u < HB && u < U; ++u) issue_S(u);
      for (int u = 0; u < U; ++u) {
        const int t = u / UPT, h = u % UPT, st = t % PP_NST, hb = u % HB;
        mbar_wait_spin(t_full(hb), (u / HB) & 1);
        tc_fence_after();
        const uint32_t ks = k_smem + st * C::STAGE;
#pragma unroll
        for (int kk = 0; kk < UC / 16; ++kk) {     // K = the unit's column pixels, 16 per product
          const uint64_t bdesc = (MODE == PCM_BWD)
                                     ? umma_desc_sw128(ks + (h * 4 + kk) * 2048, PANEL, 1024)
                                     : umma_desc_sw128(ks + 2 * PANEL + (kk >> 2) * VROWS * 128 + (kk & 3) * 32, 16, 1024);
          mma_ts(tmem + C::ACC_COL, tmem + hb * UC + kk * 8, bdesc, IDESC_O, (u | kk) != 0);
        }
        if (h == UPT - 1) tc_commit(k_empty(st));
        if (u + HB < U) issue_S(u + HB);     // after O(u) in the pipe: it overwrites the buffer O(u) reads T from
      }
      tc_commit(o_full);
    }
  } else if (warp < 4 * NG) {
    const int grp = warp >> 2, wq = warp & 3;                // wq = this warp's TMEM lane quarter
    const uint32_t lane_addr = static_cast<uint32_t>(wq * 32) << 16;
    float r_row = 0.f;
    if (MODE == PCM_BWD) r_row = p.rr[row0 + wq * 32 + lane];
    for (int u = grp; u < U; u += NG) {
      const int h = u % UPT, hb = u % HB;
      const int t = u / UPT, st = t % PP_NST;
      if (MODE == PCM_BWD) mbar_wait_warp(k_full(st), (t / PP_NST) & 1);   // r of the column pixels landed (bulk copy)
      mbar_wait_warp(s_full(hb), (u / HB) & 1);   // one polling lane per warp (32 lanes on one mbarrier word serialise)
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < UC / 32; ++c) {
        uint32_t z[32];
        tmem_ld32(tmem + lane_addr + hb * UC + c * 32, z);
        uint32_t packed[16];
        if (MODE == PCM_BWD) {
          uint32_t w[32];
          tmem_ld32(tmem + lane_addr + C::W_COL + hb * UC + c * 32, w);
          tmem_wait_ld();
          const float* rc = reinterpret_cast<const float*>(gen + C::SMEM_K + st * C::STAGE + 4 * PANEL) + h * 64 + c * 32;
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const float2 rj = reinterpret_cast<const float2*>(rc)[k];
            const float t0 = (__uint_as_float(z[2 * k]) > 0.f) ? __uint_as_float(w[2 * k]) - r_row - rj.x : 0.f;
            const float t1 = (__uint_as_float(z[2 * k + 1]) > 0.f) ? __uint_as_float(w[2 * k + 1]) - r_row - rj.y : 0.f;
            packed[k] = pack_bf16x2(t0, t1);
          }
        } else {
          tmem_wait_ld();
#pragma unroll
          for (int k = 0; k < 16; ++k)
            packed[k] = pack_bf16x2(fmaxf(__uint_as_float(z[2 * k]), 0.f), fmaxf(__uint_as_float(z[2 * k + 1]), 0.f));
        }
        // chunk c reads columns 32c..32c+31 and writes 16c..16c+15: never a column still to be read
        tmem_st16(tmem + lane_addr + hb * UC + c * 16, packed);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_full(hb));
    }
    // ---- item epilogue
    mbar_wait_warp(o_full, 0);
    tc_fence_after();
    const int64_t px = static_cast<int64_t>(rt) * 128 + wq * 32 + lane;  // pixel within the image
    if (MODE == PCM_FWD) {
      if (grp == 0) {
        uint32_t o[32];
        tmem_ld32(tmem + lane_addr + C::ACC_COL, o);
        tmem_wait_ld();
        if (px < p.P) {
          float scale = 1.f;
          if (p.normalise) {
            const float dn = __uint_as_float(o[VROWS - 1]) + PCM_EPS;
            p.den[static_cast<int64_t>(b) * p.pp + px] = dn;
            scale = 1.f / dn;
          }
#pragma unroll
          for (int k = 0; k < VROWS; ++k)
            if (k < p.K) p.out[(static_cast<int64_t>(b) * p.K + k) * p.P + px] = __uint_as_float(o[k]) * scale;
        }
      }
    } else {
      // dghat row in TMEM -> dg through ghat = g / (n + eps) -> bf16, d-major.  Both groups form the row's radial
      // component; group g stores columns 64g .. 64g+63.
      const int64_t row = static_cast<int64_t>(b) * p.pp + px;
      const bool in = px < p.P;
      const float n = in ? p.gnorm[row] : 0.f;
      const uint4* gsrc = reinterpret_cast<const uint4*>(p.ghat + row * PD);
      float dot = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_addr + C::ACC_COL + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const uint4 gq = gsrc[c * 4 + k4];
          const uint32_t gw[4] = {gq.x, gq.y, gq.z, gq.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float g0 = __uint_as_float(gw[j] << 16), g1 = __uint_as_float(gw[j] & 0xffff0000u);
            dot = fmaf(__uint_as_float(v[k4 * 8 + 2 * j]), g0, dot);
            dot = fmaf(__uint_as_float(v[k4 * 8 + 2 * j + 1]), g1, dot);
          }
        }
      }
      const float inv = in ? 1.f / (n + PCM_EPS) : 0.f;
      const float radial = (n > 0.f) ? dot * (n + PCM_EPS) / n : 0.f;
      __nv_bfloat16* dst = p.dgt + static_cast<int64_t>(b) * PD * p.pp + px;
#pragma unroll 1
      for (int c = (4 / NG) * grp; c < (4 / NG) * (grp + 1); ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_addr + C::ACC_COL + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const uint4 gq = gsrc[c * 4 + k4];
          const uint32_t gw[4] = {gq.x, gq.y, gq.z, gq.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float g0 = __uint_as_float(gw[j] << 16), g1 = __uint_as_float(gw[j] & 0xffff0000u);
            const int d0 = c * 32 + k4 * 8 + 2 * j;
            dst[static_cast<int64_t>(d0) * p.pp] = __float2bfloat16((__uint_as_float(v[k4 * 8 + 2 * j]) - g0 * radial) * inv);
            dst[static_cast<int64_t>(d0 + 1) * p.pp] = __float2bfloat16((__uint_as_float(v[k4 * 8 + 2 * j + 1]) - g1 * radial) * inv);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

// ==========================================================================================
// dW_pad[d][c] += sum_px dgT[b][d][px] fb[b][c][px]   (column `cin` of fb is 1 -> db)
// work item = (image, chunk of DW_CHUNK pixels)
// ==========================================================================================
constexpr int DW_NST = 3;
constexpr int DW_CHUNK = 512;     // (1024 gave 32 CTAs at config 2 and 37.6 us; the epilogue's reductions are 16-byte vectors now)
constexpr int DW_THREADS = 192;

__host__ __device__ inline int dw_stage_bytes(int cp) { return PANEL + cp * 128; }
__host__ __device__ inline int dw_smem_bytes(int cp) { return DW_NST * dw_stage_bytes(cp) + 256 + 1024; }

__global__ void __launch_bounds__(DW_THREADS, 1)
pcm_dw_kernel(const __grid_constant__ CUtensorMap tm_dg, const __grid_constant__ CUtensorMap tm_f, int cp,
              int chunks_per_image, int64_t pp, float* dwp) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const int stage = dw_stage_bytes(cp);
  const uint32_t bar0 = base + DW_NST * stage;
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (DW_NST + s); };
  const uint32_t d_full = bar0 + 8u * (2 * DW_NST);
  volatile uint32_t* tptr = reinterpret_cast<volatile uint32_t*>(gen + DW_NST * stage + 8 * (2 * DW_NST + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x / chunks_per_image, ch = blockIdx.x - b * chunks_per_image;
  const int64_t px0 = static_cast<int64_t>(ch) * DW_CHUNK;
  const int64_t left = pp - px0;
  const int steps = static_cast<int>((left < DW_CHUNK ? left : DW_CHUNK) / 64);
  const int nh = (cp <= 256) ? 1 : 2;   // N halves (N <= 256 per MMA)
  const int NH = cp / nh;

  if (threadIdx.x == 0) {
    for (int s = 0; s < DW_NST; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    mbar_init(d_full, 1);
    fence_barrier_init();
  }
  if (warp == 4) { tmem_alloc(smem_u32(const_cast<uint32_t*>(tptr)), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tptr;

  if (warp == 5) {
    if (lane == 0) {
      tma_prefetch_desc(&tm_dg);
      tma_prefetch_desc(&tm_f);
      for (int s = 0; s < steps; ++s) {
        const int st = s % DW_NST;
        mbar_wait_spin(empty(st), ((s / DW_NST) & 1) ^ 1);
        mbar_arrive_expect_tx(full(st), stage);
        const uint32_t a = base + st * stage;
        const int col = static_cast<int>(px0) + s * 64;
        tma_load_2d(a, &tm_dg, full(st), col, b * PD);
        for (int r = 0; r < cp / 64; ++r) tma_load_2d(a + PANEL + r * 8192, &tm_f, full(st), col, b * cp + r * 64);
      }
    }
  } else if (warp == 4) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, NH, 0, 0);
      for (int s = 0; s < steps; ++s) {
        const int st = s % DW_NST;
        mbar_wait_spin(full(st), (s / DW_NST) & 1);
        tc_fence_after();
        const uint32_t a = base + st * stage;
        for (int n = 0; n < nh; ++n)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            mma_ss(tmem + n * NH, umma_desc_sw128(a + kk * 32, 16, 1024),
                   umma_desc_sw128(a + PANEL + n * NH * 128 + kk * 32, 16, 1024), idesc, (s | kk) != 0);
        tc_commit(empty(st));
      }
      tc_commit(d_full);
    }
  } else {
    mbar_wait_warp(d_full, 0);
    tc_fence_after();
    const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
    const int d = threadIdx.x;
    if (steps > 0) {
      for (int c = 0; c < cp / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_addr + c * 32, v);
        tmem_wait_ld();
        float* dst = dwp + static_cast<size_t>(d) * cp + c * 32;     // 16-byte aligned: cp % 64 == 0
#pragma unroll
        for (int k = 0; k < 32; k += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + k), "r"(v[k]), "r"(v[k + 1]), "r"(v[k + 2]),
                       "r"(v[k + 3]) : "memory");
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 512);
}

int check_shapes(int batch, int num_classes, int cin, int64_t pixels) {
  if (batch <= 0 || pixels <= 0 || cin <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if (num_classes < 1 || num_classes > XSEG) return fail(MUSCLE_ERR_BAD_ARG, "PCM supports 1..%d CAM classes (got %d)", XSEG, num_classes);
  if (cin_pad(cin) > 384) return fail(MUSCLE_ERR_BAD_ARG, "PCM supports at most 383 input channels (got %d)", cin);
  if (static_cast<int64_t>(batch) * pad128(pixels) >= (1ll << 31)) return fail(MUSCLE_ERR_BAD_ARG, "too many pixels");
  return MUSCLE_OK;
}

template <int MODE, int HB_, int NG_>
int launch_pair_v(const CUtensorMap& tm_g, const CUtensorMap& tm_row, const CUtensorMap& tm_col, const PpParams& prm,
                  int items, cudaStream_t stream) {
  using C = PpCfg<MODE, HB_, NG_>;
  MB_CUDA(cudaFuncSetAttribute(pcm_pair_kernel<MODE, HB_, NG_>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
  pcm_pair_kernel<MODE, HB_, NG_><<<items, PP_THREADS, C::SMEM_BYTES, stream>>>(tm_g, tm_row, tm_col, prm);
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
// Buffers and epilogue groups per mode, from the A/B runs on B200 (profiles/r2aa_pcm_variants.txt): forward 3 whole-tile
// buffers (2 or 3 buffers, 1 or 2 groups: the same 44 us -- the pass is bound by the number of tcgen05.mma per tile, 16
// after the switch from 64- to 128-column S products, 58 us before); backward 2 half-tile buffers, 2 groups (three
// buffers were SLOWER: 137 vs 107 us).  A forward variant with ONE buffer, 256 TMEM columns and 113 KB of shared memory, so
// that two CTAs share an SM and all 200 items of config 2 are resident at once, ran at the same 42-44 us
// (profiles/r2ac_*): per SM the passes sustain one tcgen05.mma per ~100 cycles however the work is interleaved.
template <int MODE>
int launch_pair(const CUtensorMap& tm_g, const CUtensorMap& tm_row, const CUtensorMap& tm_col, const PpParams& prm,
                int items, cudaStream_t stream) {
  if constexpr (MODE == PCM_BWD) return launch_pair_v<MODE, 2, 2>(tm_g, tm_row, tm_col, prm, items, stream);
  else return launch_pair_v<MODE, 3, 2>(tm_g, tm_row, tm_col, prm, items, stream);
}

}  // namespace

extern "C" {

size_t muscle_pcm_state_bytes(int batch, int cin, int64_t pixels) {
  if (batch <= 0 || cin <= 0 || pixels <= 0) return 0;
  return plan_state(batch, cin, pixels).total;
}
size_t muscle_pcm_bwd_scratch_bytes(int batch, int cin, int64_t pixels) {
  if (batch <= 0 || cin <= 0 || pixels <= 0) return 0;
  return plan_scratch(batch, cin, pixels).total;
}

int muscle_pcm_fwd(const float* cam, const float* f, const float* fuse_w, const float* fuse_b, int batch,
                   int num_classes, int cin, int64_t pixels, float* out, void* state, size_t state_bytes,
                   void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!cam || !f || !fuse_w || !fuse_b || !out || !state) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (int rc = check_shapes(batch, num_classes, cin, pixels)) return rc;
  if (reinterpret_cast<uintptr_t>(state) & 255) return fail(MUSCLE_ERR_BAD_ARG, "state must be 256-byte aligned");
  const PcmState S = plan_state(batch, cin, pixels);
  if (state_bytes < S.total) return fail(MUSCLE_ERR_WORKSPACE, "state too small: %zu < %zu", state_bytes, S.total);
  const int64_t pp = pad128(pixels);
  const int cp = cin_pad(cin), kc = cp / 64, T = static_cast<int>(pp / 128);
  char* base = static_cast<char*>(state);
  auto* fb = reinterpret_cast<__nv_bfloat16*>(base + S.fb);
  auto* wb = reinterpret_cast<__nv_bfloat16*>(base + S.wb);
  auto* ghat = reinterpret_cast<__nv_bfloat16*>(base + S.ghat);
  float* gnorm = reinterpret_cast<float*>(base + S.gnorm);
  auto* camb = reinterpret_cast<__nv_bfloat16*>(base + S.camb);
  float* den = reinterpret_cast<float*>(base + S.den);

  pcm_cast_f_kernel<<<dim3(batch * cp, static_cast<unsigned>((pp / 8 + 255) / 256)), 256, 0, stream>>>(f, cin, cp, pixels, pp, fb);
  MB_LAUNCH_CHECK();
  pcm_cast_w_kernel<<<(PD * cp + 255) / 256, 256, 0, stream>>>(fuse_w, cin, cp, wb);
  MB_LAUNCH_CHECK();
  pcm_cast_v_kernel<<<dim3(static_cast<unsigned>((pp + 255) / 256), VROWS, batch), 256, 0, stream>>>(cam, nullptr, num_classes, pixels, pp, 1, camb);
  MB_LAUNCH_CHECK();

  CUtensorMap tm_w, tm_f, tm_g, tm_v;
  if (int rc = make_bf16_box_map(&tm_w, wb, PD, cp, 128)) return rc;
  if (int rc = make_bf16_box_map(&tm_f, fb, static_cast<int64_t>(batch) * cp, pp, 64)) return rc;
  if (int rc = make_bf16_box_map(&tm_g, ghat, static_cast<int64_t>(batch) * pp, PD, 128)) return rc;
  if (int rc = make_bf16_box_map(&tm_v, camb, static_cast<int64_t>(batch) * VROWS, pp, VROWS)) return rc;

  const int n_tiles = batch * T;
  const int smem = pj_smem_bytes(kc);
  MB_CUDA(cudaFuncSetAttribute(pcm_project_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int grid = std::min(n_tiles, device_info().sm_count);
  pcm_project_kernel<<<grid, PJ_THREADS, smem, stream>>>(tm_w, tm_f, fuse_b, kc, cp, T, n_tiles, pixels, pp, ghat, gnorm);
  MB_LAUNCH_CHECK();

  PpParams prm{};
  prm.tiles_per_image = T; prm.P = pixels; prm.pp = pp; prm.K = num_classes; prm.normalise = 1;
  prm.out = out; prm.den = den;
  return launch_pair<PCM_FWD>(tm_g, tm_v, tm_v, prm, n_tiles, stream);
}

int muscle_pcm_bwd(const float* dout, const float* cam, const float* out, int batch, int num_classes, int cin,
                   int64_t pixels, const void* state, size_t state_bytes, float* dcam, float* dfuse_w,
                   float* dfuse_b, void* scratch, size_t scratch_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = need_sm100()) return rc;
  if (!dout || !cam || !out || !state || !scratch) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if ((dfuse_w == nullptr) != (dfuse_b == nullptr)) return fail(MUSCLE_ERR_BAD_ARG, "dfuse_w and dfuse_b go together");
  if (int rc = check_shapes(batch, num_classes, cin, pixels)) return rc;
  if ((reinterpret_cast<uintptr_t>(state) | reinterpret_cast<uintptr_t>(scratch)) & 255)
    return fail(MUSCLE_ERR_BAD_ARG, "state and scratch must be 256-byte aligned");
  const PcmState S = plan_state(batch, cin, pixels);
  const PcmScratch W = plan_scratch(batch, cin, pixels);
  if (state_bytes < S.total) return fail(MUSCLE_ERR_WORKSPACE, "state too small");
  if (scratch_bytes < W.total) return fail(MUSCLE_ERR_WORKSPACE, "scratch too small: %zu < %zu", scratch_bytes, W.total);
  const int64_t pp = pad128(pixels);
  const int cp = cin_pad(cin), T = static_cast<int>(pp / 128);
  const char* sb = static_cast<const char*>(state);
  char* wb = static_cast<char*>(scratch);
  auto* fb = reinterpret_cast<const __nv_bfloat16*>(sb + S.fb);
  auto* ghat = reinterpret_cast<const __nv_bfloat16*>(sb + S.ghat);
  const float* gnorm = reinterpret_cast<const float*>(sb + S.gnorm);
  const float* den = reinterpret_cast<const float*>(sb + S.den);
  auto* ab = reinterpret_cast<__nv_bfloat16*>(wb + W.ab);
  auto* xp = reinterpret_cast<__nv_bfloat16*>(wb + W.xp);
  auto* yp = reinterpret_cast<__nv_bfloat16*>(wb + W.yp);
  float* rr = reinterpret_cast<float*>(wb + W.rr);
  auto* dgt = reinterpret_cast<__nv_bfloat16*>(wb + W.dgt);
  float* dwp = reinterpret_cast<float*>(wb + W.dwp);

  pcm_bwd_prep_kernel<<<dim3(static_cast<unsigned>(pp / PREP_PX), batch), 256, 0, stream>>>(dout, out, cam, den, num_classes, pixels, pp, ab, xp, yp, rr);
  MB_LAUNCH_CHECK();
  CUtensorMap tm_g, tm_a, tm_x, tm_y;
  if (int rc = make_bf16_box_map(&tm_g, ghat, static_cast<int64_t>(batch) * pp, PD, 128)) return rc;
  const int n_tiles = batch * T;
  if (dcam) {
    if (int rc = make_bf16_box_map(&tm_a, ab, static_cast<int64_t>(batch) * VROWS, pp, VROWS)) return rc;
    PpParams prm{};
    prm.tiles_per_image = T; prm.P = pixels; prm.pp = pp; prm.K = num_classes; prm.normalise = 0; prm.out = dcam;
    if (int rc = launch_pair<PCM_FWD>(tm_g, tm_a, tm_a, prm, n_tiles, stream)) return rc;
  }
  if (dfuse_w) {
    if (int rc = make_bf16_box_map(&tm_x, xp, static_cast<int64_t>(batch) * pp, XK, 128)) return rc;
    if (int rc = make_bf16_box_map(&tm_y, yp, static_cast<int64_t>(batch) * pp, XK, 128)) return rc;
    PpParams prm{};
    prm.tiles_per_image = T; prm.P = pixels; prm.pp = pp; prm.K = num_classes;
    prm.ghat = ghat; prm.gnorm = gnorm; prm.dgt = dgt; prm.rr = rr;
    if (int rc = launch_pair<PCM_BWD>(tm_g, tm_x, tm_y, prm, n_tiles, stream)) return rc;
    CUtensorMap tm_dg, tm_f;
    if (int rc = make_bf16_box_map(&tm_dg, dgt, static_cast<int64_t>(batch) * PD, pp, 128)) return rc;
    if (int rc = make_bf16_box_map(&tm_f, fb, static_cast<int64_t>(batch) * cp, pp, 64)) return rc;
    MB_CUDA(cudaMemsetAsync(dwp, 0, static_cast<size_t>(PD) * cp * 4, stream));
    const int chunks = static_cast<int>((pp + DW_CHUNK - 1) / DW_CHUNK);
    const int smem = dw_smem_bytes(cp);
    MB_CUDA(cudaFuncSetAttribute(pcm_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    pcm_dw_kernel<<<batch * chunks, DW_THREADS, smem, stream>>>(tm_dg, tm_f, cp, chunks, pp, dwp);
    MB_LAUNCH_CHECK();
    pcm_unpad_dw_kernel<<<(PD * cin + 255) / 256, 256, 0, stream>>>(dwp, cin, cp, dfuse_w, dfuse_b);
    MB_LAUNCH_CHECK();
  }
  return MUSCLE_OK;
}

}  // extern "C"
