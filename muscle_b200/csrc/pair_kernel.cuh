// The dense pixel-pair mainloop shared by the InfoNCE row-sum pass, the InfoNCE gradient pass
// and the PCM (ReLU-affinity) passes.
//
// One persistent CTA per SM, warp-specialised:
//   warps 0-3  epilogue group 0  (TMEM lanes 32*(warp%4) .. +31, thread = one row of a 128-row tile)
//   warps 4-7  epilogue group 1
//   warp  8    TMA producer      (row tile once per work item, column tiles through a ring)
//   warp  9    tcgen05.mma issuer (one lane) + TMEM allocator
//
// Work item = (row block, column split).  Per column tile of 128 pixels and per "slot"
// (two S buffers in TMEM, one per epilogue group):
//   S = R C^T        tcgen05.mma SS, bf16 x bf16 -> fp32 in TMEM        [128 x 128]
//   epilogue         tcgen05.ld S, e = exp2(a*s - a*bound), label mask, then
//       MODE_ROWSUM: row-sum of e kept in registers, one atomicAdd per row per item
//       MODE_GRAD  : T = e * (w_row + w_col) -> bf16 -> tcgen05.st into the S columns
//   (MODE_GRAD) G += T C   tcgen05.mma TS: A = T from TMEM, B = the same column tile, MN-major
// so the N x N logits never leave the SM.
//
// TMEM columns: [0,128) S slot 0 (T aliases its first 64), [128,256) S slot 1, [256,512) G.
// D = 64/128: two row tiles per item (slot == row tile, two accumulators of D columns);
// D = 256    : one row tile per item, the two slots alternate column tiles into one accumulator.
#pragma once
#include "ptx.cuh"

namespace mb200 {

enum PairMode { MODE_ROWSUM = 0, MODE_GRAD = 1 };

struct PairParams {
  const uint32_t* bits_row;  // [rows_pad]  1<<class, 0 = pixel not in the loss
  const uint32_t* bits_col;  // [cols_pad]
  const float* w_row;        // [rows_pad]  MODE_GRAD: 1/rowsum of the row pixel (0 = no query role)
  const float* w_col;        // [cols_pad]  MODE_GRAD: 1/rowsum of the column pixel (0 = no key->query role)
  float* rowsum;             // [rows_pad]  MODE_ROWSUM: += sum_k mask * exp((s - bound)/tau)
  float* part;               // [n_split][rows_pad][D]  MODE_GRAD: partial sum_k T_qk c_k
  const float* bound;        // device scalar: upper bound of |s| (1 for unit vectors)
  float scale_log2;          // log2(e) / tau
  int n_row_blocks;          // row blocks of RB rows
  int n_col_tiles;           // column tiles of 128
  int n_split;               // column splits per row block
  int rows_pad;
};

template <int D>
struct PairCfg {
  static_assert(D == 64 || D == 128 || D == 256, "projection dim must be 64, 128 or 256");
  static constexpr int QT = (D == 256) ? 1 : 2;          // row tiles per item
  static constexpr int RB = 128 * QT;                    // rows per item
  static constexpr int NST = (D == 256) ? 2 : (D == 128 ? 4 : 6);
  static constexpr int TILE_BYTES = 128 * D * 2;         // one 128-pixel bf16 tile
  static constexpr int PANEL_BYTES = 128 * 128;          // one [128 x 64] swizzle panel
  static constexpr int NPANEL = D / 64;
  static constexpr int SMEM_Q = 0;
  static constexpr int SMEM_K = SMEM_Q + QT * TILE_BYTES;
  static constexpr int SMEM_BITS = SMEM_K + NST * TILE_BYTES;
  static constexpr int SMEM_W = SMEM_BITS + NST * 512;
  static constexpr int SMEM_BAR = SMEM_W + NST * 512;
  static constexpr int NBAR = 2 * NST + 12;
  static constexpr int SMEM_TMEMPTR = SMEM_BAR + NBAR * 8;
  static constexpr int SMEM_BYTES = SMEM_TMEMPTR + 16 + 1024;  // +1024: manual 1024-B alignment
  static constexpr int ACC_COL = 256;
  static constexpr int THREADS = 320;
};

template <int D, int MODE>
__global__ void __launch_bounds__(320, 1)
pair_kernel(const __grid_constant__ CUtensorMap tm_row, const __grid_constant__ CUtensorMap tm_col,
            const PairParams p) {
  using C = PairCfg<D>;
  constexpr int QT = C::QT;
  constexpr int NST = C::NST;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t q_smem = smem_base + C::SMEM_Q;
  const uint32_t k_smem = smem_base + C::SMEM_K;
  const uint32_t bits_smem = smem_base + C::SMEM_BITS;
  const uint32_t w_smem = smem_base + C::SMEM_W;
  const uint32_t bar0 = smem_base + C::SMEM_BAR;
  auto k_full = [&](int s) { return bar0 + 8u * s; };
  auto k_empty = [&](int s) { return bar0 + 8u * (NST + s); };
  const uint32_t q_full = bar0 + 8u * (2 * NST);
  const uint32_t q_empty = bar0 + 8u * (2 * NST + 1);
  auto s_full = [&](int s) { return bar0 + 8u * (2 * NST + 2 + s); };
  auto s_free = [&](int s) { return bar0 + 8u * (2 * NST + 4 + s); };
  auto t_full = [&](int s) { return bar0 + 8u * (2 * NST + 6 + s); };
  auto acc_full = [&](int s) { return bar0 + 8u * (2 * NST + 8 + s); };
  auto acc_empty = [&](int s) { return bar0 + 8u * (2 * NST + 10 + s); };
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + C::SMEM_TMEMPTR);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(k_full(s), 1);
      mbar_init(k_empty(s), 1 + (QT == 2 ? 8 : 4));
    }
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(s_full(s), 1);
      mbar_init(s_free(s), 4);
      mbar_init(t_full(s), 4);
      mbar_init(acc_full(s), 1);
      mbar_init(acc_empty(s), QT == 2 ? 4 : 8);
    }
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_ptr_smem)), 512);
    tmem_relinquish();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tm_row);
    tma_prefetch_desc(&tm_col);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;

  const int n_items = p.n_row_blocks * p.n_split;
  auto item_range = [&](int item, int& rb, int& ks, int& t0, int& t1) {
    rb = item / p.n_split;
    ks = item - rb * p.n_split;
    t0 = static_cast<int>((static_cast<long long>(p.n_col_tiles) * ks) / p.n_split);
    t1 = static_cast<int>((static_cast<long long>(p.n_col_tiles) * (ks + 1)) / p.n_split);
  };

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t g = 0, it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        int rb, ks, t0, t1;
        item_range(item, rb, ks, t0, t1);
        mbar_wait(q_empty, (it & 1) ^ 1);
        mbar_arrive_expect_tx(q_full, QT * C::TILE_BYTES);
        for (int qt = 0; qt < QT; ++qt)
          for (int a = 0; a < C::NPANEL; ++a)
            tma_load_2d(q_smem + qt * C::TILE_BYTES + a * C::PANEL_BYTES, &tm_row, q_full, a * 64,
                        (rb * QT + qt) * 128);
        for (int j = t0; j < t1; ++j, ++g) {
          const int st = g % NST;
          mbar_wait(k_empty(st), ((g / NST) & 1) ^ 1);
          mbar_arrive_expect_tx(k_full(st), C::TILE_BYTES + 512 + (MODE == MODE_GRAD ? 512 : 0));
          for (int a = 0; a < C::NPANEL; ++a)
            tma_load_2d(k_smem + st * C::TILE_BYTES + a * C::PANEL_BYTES, &tm_col, k_full(st),
                        a * 64, j * 128);
          bulk_load_1d(bits_smem + st * 512, p.bits_col + static_cast<size_t>(j) * 128, 512,
                       k_full(st));
          if (MODE == MODE_GRAD)
            bulk_load_1d(w_smem + st * 512, p.w_col + static_cast<size_t>(j) * 128, 512,
                         k_full(st));
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t IDESC_G = umma_idesc_bf16(128, D, 0, 1);
      uint32_t g = 0, it = 0;
      uint32_t cnt_s[2] = {0, 0};
      uint32_t cnt_t[2] = {0, 0};
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        int rb, ks, t0, t1;
        item_range(item, rb, ks, t0, t1);
        const int nkt = t1 - t0;
        const int U = (QT == 2) ? 2 * nkt : nkt;
        mbar_wait(q_full, it & 1);
        tc_fence_after();

        auto issue_S = [&](int u) {
          const int slot = u & 1;
          const int j = (QT == 2) ? (u >> 1) : u;
          const uint32_t gg = g + j;
          const int st = gg % NST;
          if (QT == 1 || slot == 0) {
            mbar_wait(k_full(st), (gg / NST) & 1);
            tc_fence_after();
          }
          if (MODE == MODE_ROWSUM) {
            mbar_wait(s_free(slot), (cnt_s[slot] & 1) ^ 1);
            tc_fence_after();
          }
          const uint32_t a_base = q_smem + (QT == 2 ? slot * C::TILE_BYTES : 0);
          const uint32_t b_base = k_smem + st * C::TILE_BYTES;
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = (kk >> 2) * C::PANEL_BYTES + (kk & 3) * 32;
            mma_ss(tmem + slot * 128, umma_desc_sw128(a_base + off, 16, 1024),
                   umma_desc_sw128(b_base + off, 16, 1024), IDESC_S, kk > 0);
          }
          tc_commit(s_full(slot));
          ++cnt_s[slot];
          if (MODE == MODE_ROWSUM && (QT == 1 || slot == 1)) tc_commit(k_empty(st));
        };

        issue_S(0);
        if (U > 1) issue_S(1);
        if (MODE == MODE_ROWSUM) {
          for (int u = 2; u < U; ++u) issue_S(u);
        } else {
          for (int u = 0; u < U; ++u) {
            const int slot = u & 1;
            const int j = (QT == 2) ? (u >> 1) : u;
            const int st = (g + j) % NST;
            const int acc = (QT == 2) ? slot : 0;
            const bool first = (QT == 2) ? (u < 2) : (u == 0);
            const bool last = (QT == 2) ? (u >= U - 2) : (u == U - 1);
            if (first) {
              mbar_wait(acc_empty(acc), (it & 1) ^ 1);
              tc_fence_after();
            }
            mbar_wait(t_full(slot), cnt_t[slot] & 1);
            ++cnt_t[slot];
            tc_fence_after();
            const uint32_t b_base = k_smem + st * C::TILE_BYTES;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk)
              mma_ts(tmem + C::ACC_COL + acc * D, tmem + slot * 128 + kk * 8,
                     umma_desc_sw128(b_base + kk * 2048, C::PANEL_BYTES, 1024), IDESC_G,
                     (first && kk == 0) ? 0u : 1u);
            if (QT == 1 || slot == 1) tc_commit(k_empty(st));
            if (last) tc_commit(acc_full(acc));
            if (u + 2 < U) issue_S(u + 2);
          }
        }
        tc_commit(q_empty);
        g += nkt;
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue groups
    const int slot = warp >> 2;
    const int wq = warp & 3;
    const uint32_t lane_addr = static_cast<uint32_t>(wq * 32) << 16;
    const float off = p.scale_log2 * __ldg(p.bound);
    const float a = p.scale_log2;
    uint32_t g = 0, it = 0, cnt = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      int rb, ks, t0, t1;
      item_range(item, rb, ks, t0, t1);
      const int nkt = t1 - t0;
      const int row = rb * C::RB + (QT == 2 ? slot * 128 : 0) + wq * 32 + lane;
      const uint32_t rbits = __ldg(p.bits_row + row);
      const uint32_t qmask = rbits ? ~rbits : 0u;
      float wrow = 0.f;
      if (MODE == MODE_GRAD) wrow = __ldg(p.w_row + row);
      float rsum = 0.f;
      const int n_my = (QT == 2) ? nkt : ((nkt - slot + 1) >> 1);
      for (int i = 0; i < n_my; ++i) {
        const int j = (QT == 2) ? i : (2 * i + slot);
        const uint32_t gg = g + j;
        const int st = gg % NST;
        mbar_wait(k_full(st), (gg / NST) & 1);
        mbar_wait(s_full(slot), cnt & 1);
        ++cnt;
        tc_fence_after();
        const uint4* kb4 = reinterpret_cast<const uint4*>(smem_gen + C::SMEM_BITS + st * 512);
        const float4* kw4 = reinterpret_cast<const float4*>(smem_gen + C::SMEM_W + st * 512);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_addr + slot * 128 + c * 32, v);
          tmem_wait_ld();
          uint32_t packed[16];
#pragma unroll
          for (int q4 = 0; q4 < 8; ++q4) {
            const uint4 kb = kb4[c * 8 + q4];
            const uint32_t kbv[4] = {kb.x, kb.y, kb.z, kb.w};
            float kwv[4] = {0.f, 0.f, 0.f, 0.f};
            if (MODE == MODE_GRAD) {
              const float4 kw = kw4[c * 8 + q4];
              kwv[0] = kw.x; kwv[1] = kw.y; kwv[2] = kw.z; kwv[3] = kw.w;
            }
            float t[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float s = __uint_as_float(v[q4 * 4 + e]);
              const float x = ex2_approx(fmaf(s, a, -off));
              const bool m = (kbv[e] & qmask) != 0u;
              if (MODE == MODE_ROWSUM) {
                rsum += m ? x : 0.f;
              } else {
                t[e] = m ? x * (wrow + kwv[e]) : 0.f;
              }
            }
            if (MODE == MODE_GRAD) {
              packed[q4 * 2 + 0] = pack_bf16x2(t[0], t[1]);
              packed[q4 * 2 + 1] = pack_bf16x2(t[2], t[3]);
            }
          }
          if (MODE == MODE_GRAD) tmem_st16(tmem + lane_addr + slot * 128 + c * 16, packed);
        }
        if (MODE == MODE_GRAD) tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(k_empty(st));
          mbar_arrive(MODE == MODE_GRAD ? t_full(slot) : s_free(slot));
        }
      }
      if (MODE == MODE_ROWSUM) {
        atomicAdd(p.rowsum + row, rsum);
      } else {
        const int acc = (QT == 2) ? slot : 0;
        mbar_wait(acc_full(acc), it & 1);
        tc_fence_after();
        constexpr int NCH = (QT == 2) ? (D / 32) : (D / 64);
        const int col0 = (QT == 2) ? 0 : slot * (D / 2);
        float* dst = p.part + (static_cast<size_t>(ks) * p.rows_pad + row) * D + col0;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_addr + C::ACC_COL + acc * D + col0 + c * 32, v);
          tmem_wait_ld();
#pragma unroll
          for (int q4 = 0; q4 < 8; ++q4)
            reinterpret_cast<float4*>(dst + c * 32)[q4] =
                make_float4(__uint_as_float(v[q4 * 4]), __uint_as_float(v[q4 * 4 + 1]),
                            __uint_as_float(v[q4 * 4 + 2]), __uint_as_float(v[q4 * 4 + 3]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty(acc));
      }
      g += nkt;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, 512);
}

}  // namespace mb200
