// The dense pixel-pair mainloop of the InfoNCE row-sum pass and the InfoNCE gradient pass.
//
// One persistent CTA per SM, warp-specialised:
//   warps 0-3  epilogue group 0  (TMEM lanes 32*(warp%4) .. +31, thread = one row of a 128-row tile)
//   warps 4-7  epilogue group 1
//   warp  8    TMA producer      (row tiles once per work item, column tiles through a ring)
//   warp  9    tcgen05.mma issuer (one lane) + TMEM allocator
//
// Pixels arrive sorted by class (row_kernels.cuh), and a per-tile class union (`tor`) lets the
// producer drop every (row block, column tile) whose pairs are all masked (same single class)
// before anything is loaded or multiplied; the epilogue takes a mask-free path for tiles whose
// class sets are disjoint.  The LIVE (row block, column tile) units are numbered row block by row
// block (`unit_off`, pair_plan_kernel) and every CTA takes one contiguous share of that range, so
// the SMs finish together whatever the class mix; a CTA's share of one row block is a work item,
// and the item's index among the CTAs that touch the row block is its output `slot`.
//
// Two "streams" run through the tensor pipe, one per epilogue group (D <= 128: stream = row
// tile, both see every column tile; D = 256: one row tile, the streams alternate column tiles).
// Within a stream every 128 x 128 similarity tile is produced and consumed as two 64-column halves
// held in two TMEM buffers, so the group's epilogue on one half always overlaps the stream's MMAs
// on the other:
//   S_h = R C_h^T      tcgen05.mma SS, bf16 x bf16 -> fp32 in TMEM            [128 x 64]
//   epilogue           tcgen05.ld S_h, e = exp2(a*s - a), label mask, then
//       MODE_ROWSUM : row-sum of e kept in registers, one atomicAdd per row per item
//       MODE_GRAD   : T_h = e * (w_row + w_col) -> bf16 -> tcgen05.st over the S_h columns
//   (MODE_GRAD) G += T_h C_h   tcgen05.mma TS: A = T_h from TMEM, B = the same column rows, MN-major
// so the N x N logits never leave the SM.  A share of the exponentials is evaluated with a cubic on
// the FMA pipe (ptx.cuh: ex2_poly2) because MUFU alone (16/clk/SM) would cap the pass below the
// tensor pipe; all elementwise math is packed fp32x2.
//
// TMEM columns: stream g: S halves at [128g, 128g+64) and [128g+64, 128g+128), T_h aliases the
// first 32 columns of S_h; G accumulators from column 256 (D <= 128: one per stream; D = 256: one).
#pragma once
#include "ptx.cuh"

namespace mb200 {

enum PairMode { MODE_ROWSUM = 0, MODE_GRAD = 1 };

constexpr uint32_t TOR_INVALID = 0x80000000u;  // tile contains a pixel that is in no class
constexpr uint32_t TOR_CLASSES = 0x7fffffffu;

struct PairParams {
  const uint32_t* bits_row;  // [rows_pad]  1<<class, 0 = pixel not in the loss
  const uint32_t* bits_col;  // [cols_pad]
  const uint32_t* tor_row;   // [rows_pad/128]  OR of bits over the tile | TOR_INVALID
  const uint32_t* tor_col;   // [cols_pad/128]
  const uint32_t* unit_off;  // [n_row_blocks+1]  first live unit of every row block
  const float* w_row;        // [rows_pad]  MODE_GRAD: 1/rowsum of the row pixel (0 = no query role)
  const float* w_col;        // [cols_pad]  MODE_GRAD: 1/rowsum of the column pixel
  float* rowsum;             // [rows_pad]  MODE_ROWSUM: += sum_k mask * exp((s - 1)/tau)
  float* part;               // [PAIR_SLOTS][rows_pad][D]  MODE_GRAD: partial sum_k T_qk c_k
  float scale_log2;          // log2(e) / tau
  int n_row_blocks;          // row blocks of RB rows
  int n_col_tiles;           // column tiles of 128
  int rows_pad;
  unsigned long long* dbg;   // optional [gridDim.x][16] cycle counters (MUSCLE_PAIR_TIMING), else null
};

constexpr int PAIR_SLOTS = 8;             // a row block is shared by at most this many CTAs
constexpr uint32_t ITEM_END = 0xffffffffu;

// live units per CTA: an even share, but never so small that a row block spans > PAIR_SLOTS CTAs
__host__ __device__ inline uint32_t pair_units_per_cta(uint32_t total_units, int n_ctas, int n_col_tiles) {
  const uint32_t even = (total_units + n_ctas - 1) / n_ctas;
  const uint32_t floor_ = (n_col_tiles + PAIR_SLOTS - 2) / (PAIR_SLOTS - 1);
  return even > floor_ ? even : floor_;
}

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
  static constexpr int NBAR = 2 * NST + 14;
  static constexpr int SMEM_INFO = SMEM_BAR + NBAR * 8;  // item headers [2][4], stage info [NST]
  static constexpr int SMEM_TMEMPTR = SMEM_INFO + (8 + NST) * 4 + NST * 4;   // + debug: TMA issue clocks [NST]
  static constexpr int SMEM_BYTES = SMEM_TMEMPTR + 16 + 1024;  // +1024: manual 1024-B alignment
  static constexpr int ACC_COL = 256;
  static constexpr int THREADS = 320;
};

// all pairs of (row tile, column tile) masked: one side empty, or both the same single class
__device__ __forceinline__ bool tile_pair_dead(uint32_t tor_r, uint32_t tor_c) {
  const uint32_t r = tor_r & TOR_CLASSES, c = tor_c & TOR_CLASSES;
  return r == 0u || c == 0u || (r == c && (r & (r - 1u)) == 0u);
}
// no pair masked: disjoint class sets and every column pixel valid
__device__ __forceinline__ bool tile_pair_full(uint32_t tor_r, uint32_t tor_c) {
  return (tor_r & tor_c & TOR_CLASSES) == 0u && (tor_c & TOR_INVALID) == 0u;
}

// ------------------------------------------------------------------------------------------
// unit_off[rb] = number of live (row block, column tile) units before row block rb;
// unit_off[n_row_blocks] = total.  A unit is dead when every row tile of the block is dead
// against the column tile (tile_pair_dead), which depends on the column tile only through
// "empty" / "exactly class c", so a histogram of the column tiles gives every row block's count.
template <int QT>
__global__ void __launch_bounds__(1024) pair_plan_kernel(const uint32_t* tor_row, int n_row_blocks,
                                                         const uint32_t* tor_col, int n_col_tiles,
                                                         uint32_t* unit_off) {
  __shared__ uint32_t single[32];   // column tiles whose class set is exactly {c}
  __shared__ uint32_t empty_cnt;    // column tiles without any class
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x < 32) single[threadIdx.x] = 0u;
  if (threadIdx.x == 0) { empty_cnt = 0u; carry = 0u; }
  __syncthreads();
  for (int j = threadIdx.x; j < n_col_tiles; j += blockDim.x) {
    const uint32_t c = tor_col[j] & TOR_CLASSES;
    if (c == 0u) atomicAdd(&empty_cnt, 1u);
    else if ((c & (c - 1u)) == 0u) atomicAdd(&single[31 - __clz(c)], 1u);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n_row_blocks; base += blockDim.x) {
    const int rb = base + threadIdx.x;
    uint32_t live = 0;
    if (rb < n_row_blocks) {
      uint32_t uni = 0;       // union of the block's row-tile class sets
      bool one = true;        // every non-empty row tile is the same single class
      for (int qt = 0; qt < QT; ++qt) {
        const uint32_t r = tor_row[rb * QT + qt] & TOR_CLASSES;
        if (r == 0u) continue;
        if ((r & (r - 1u)) != 0u || (uni != 0u && uni != r)) one = false;
        uni |= r;
      }
      if (uni != 0u) {
        live = static_cast<uint32_t>(n_col_tiles) - empty_cnt;
        if (one) live -= single[31 - __clz(uni)];
      }
    }
    // block-wide exclusive scan of `live`
    uint32_t inc = live;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += u;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += warp_tot[w];
    const uint32_t before = carry + woff + inc - live;
    if (rb < n_row_blocks) unit_off[rb] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + live;
    __syncthreads();
  }
  if (threadIdx.x == 0) unit_off[n_row_blocks] = carry;
}

// ------------------------------------------------------------------------------------------
// Epilogue math of one 64-column half for one thread (= one row); v = the S values.
// FULL: no label mask needed.  POLY of every 4 element pairs take the FMA-pipe exponential.
template <int MODE, int POLY, bool FULL>
__device__ __forceinline__ void pair_half_math(uint32_t (&v)[2][32], uint32_t t_addr,
                                               const uint8_t* bits_sm, const uint8_t* w_sm,
                                               uint32_t qmask, float a, float wrow,
                                               uint64_t (&racc)[4]) {
  const uint64_t a2 = pack2(a, a);
  const uint64_t noff2 = pack2(-a, -a);  // the shift: exp((s - 1)/tau), |s| <= 1 for unit vectors
  const uint64_t wr2 = pack2(wrow, wrow);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint4* kb4 = reinterpret_cast<const uint4*>(bits_sm) + c * 8;
    const float4* kw4 = reinterpret_cast<const float4*>(w_sm) + c * 8;
    uint32_t packed[16];
#pragma unroll
    for (int q4 = 0; q4 < 8; ++q4) {
      uint64_t e[2];
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint64_t x = fma2(pack2u(v[c][q4 * 4 + 2 * s], v[c][q4 * 4 + 2 * s + 1]), a2, noff2);
        if (POLY < 0) {
          e[s] = x;  // measurement variant: no exponential
        } else if (((q4 * 2 + s) & 3) < POLY) {
          e[s] = ex2_poly2(x);
        } else {
          float x0, x1;
          unpack2(x, x0, x1);
          e[s] = pack2(ex2_approx(x0), ex2_approx(x1));
        }
      }
      if (!FULL) {
        const uint4 kb = kb4[q4];
        float e0, e1, e2, e3;
        unpack2(e[0], e0, e1);
        unpack2(e[1], e2, e3);
        e0 = (kb.x & qmask) ? e0 : 0.f;
        e1 = (kb.y & qmask) ? e1 : 0.f;
        e2 = (kb.z & qmask) ? e2 : 0.f;
        e3 = (kb.w & qmask) ? e3 : 0.f;
        e[0] = pack2(e0, e1);
        e[1] = pack2(e2, e3);
      }
      if (MODE == MODE_ROWSUM) {
        racc[(q4 * 2) & 3] = add2(racc[(q4 * 2) & 3], e[0]);
        racc[(q4 * 2 + 1) & 3] = add2(racc[(q4 * 2 + 1) & 3], e[1]);
      } else {
        const float4 kw = kw4[q4];
        const uint64_t t0 = mul2(e[0], add2(wr2, pack2(kw.x, kw.y)));
        const uint64_t t1 = mul2(e[1], add2(wr2, pack2(kw.z, kw.w)));
        float f0, f1, f2, f3;
        unpack2(t0, f0, f1);
        unpack2(t1, f2, f3);
        packed[q4 * 2 + 0] = pack_bf16x2(f0, f1);
        packed[q4 * 2 + 1] = pack_bf16x2(f2, f3);
      }
    }
    if (MODE == MODE_GRAD) tmem_st16(t_addr + c * 16, packed);
  }
}

template <int D, int MODE, int POLY>
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
  auto s_full = [&](int g, int h) { return bar0 + 8u * (2 * NST + 2 + g * 2 + h); };
  auto t_full = [&](int g, int h) { return bar0 + 8u * (2 * NST + 6 + g * 2 + h); };  // GRAD: T stored; ROWSUM: S consumed
  auto acc_full = [&](int s) { return bar0 + 8u * (2 * NST + 10 + s); };
  auto acc_empty = [&](int s) { return bar0 + 8u * (2 * NST + 12 + s); };
  volatile uint32_t* info = reinterpret_cast<volatile uint32_t*>(smem_gen + C::SMEM_INFO);
  // info[4*(it&1) + {0,1,2}]: live tiles, row block, slot of item `it`; info[8+st]: tor of the stage's column tile
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
    mbar_init(q_empty, 1 + 8);  // MMA issuer + every epilogue warp (after it has read the item header)
    for (int g = 0; g < 2; ++g) {
      for (int h = 0; h < 2; ++h) {
        mbar_init(s_full(g, h), 1);
        mbar_init(t_full(g, h), 4);
      }
      mbar_init(acc_full(g), 1);
      mbar_init(acc_empty(g), QT == 2 ? 4 : 8);
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

  // stream-local tile k of stream g -> index within the item's list of live column tiles
  auto n_stream = [&](int n, int g) { return (QT == 2) ? n : ((n - g + 1) >> 1); };
  auto tile_of = [&](int g, int k) { return (QT == 2) ? k : (2 * k + g); };

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer + work partition
    if (lane == 0) {
      const uint32_t total = __ldg(p.unit_off + p.n_row_blocks);
      const uint32_t per = pair_units_per_cta(total, gridDim.x, p.n_col_tiles);
      const uint64_t lo64 = static_cast<uint64_t>(per) * blockIdx.x;
      uint32_t u = lo64 < total ? static_cast<uint32_t>(lo64) : total;
      const uint32_t u1 = (lo64 + per < total) ? static_cast<uint32_t>(lo64 + per) : total;
      // largest rb with unit_off[rb] <= u
      int rb = 0;
      {
        int lo = 0, hi = p.n_row_blocks;  // unit_off[lo] <= u, unit_off[hi] > u (when u < total)
        while (hi - lo > 1) {
          const int mid = (lo + hi) >> 1;
          if (__ldg(p.unit_off + mid) <= u) lo = mid; else hi = mid;
        }
        rb = lo;
      }
      uint32_t gcnt = 0, it = 0;
      long long p_wait = 0, p_t0 = p.dbg ? clock64() : 0;
      while (u < u1) {
        const uint32_t U0 = __ldg(p.unit_off + rb), U1 = __ldg(p.unit_off + rb + 1);
        if (U1 <= u) { ++rb; continue; }   // row blocks without live units
        const uint32_t a = u - U0;
        const uint32_t e = (u1 < U1 ? u1 : U1) - U0;
        uint32_t tor_r[QT];
        for (int qt = 0; qt < QT; ++qt) tor_r[qt] = __ldg(p.tor_row + rb * QT + qt);
        auto dead = [&](uint32_t tc) {
          bool d = true;
          for (int qt = 0; qt < QT; ++qt) d = d && tile_pair_dead(tor_r[qt], tc);
          return d;
        };
        mbar_wait(q_empty, (it & 1) ^ 1);
        info[4 * (it & 1) + 0] = e - a;
        info[4 * (it & 1) + 1] = static_cast<uint32_t>(rb);
        info[4 * (it & 1) + 2] = blockIdx.x - U0 / per;
        mbar_arrive_expect_tx(q_full, QT * C::TILE_BYTES);
        for (int qt = 0; qt < QT; ++qt)
          for (int pa = 0; pa < C::NPANEL; ++pa)
            tma_load_2d(q_smem + qt * C::TILE_BYTES + pa * C::PANEL_BYTES, &tm_row, q_full, pa * 64,
                        (rb * QT + qt) * 128);
        uint32_t li = 0;  // live index of column tile j within the row block
        for (int j = 0; j < p.n_col_tiles && li < e; ++j) {
          const uint32_t tc = __ldg(p.tor_col + j);
          if (dead(tc)) continue;
          if (li++ < a) continue;
          const int st = gcnt % NST;
          { const long long w0 = p.dbg ? clock64() : 0;
            mbar_wait(k_empty(st), ((gcnt / NST) & 1) ^ 1);
            if (p.dbg) p_wait += clock64() - w0; }
          info[8 + st] = tc;
          if (p.dbg) info[8 + NST + st] = static_cast<uint32_t>(clock64());
          mbar_arrive_expect_tx(k_full(st), C::TILE_BYTES + 512 + (MODE == MODE_GRAD ? 512 : 0));
          for (int pa = 0; pa < C::NPANEL; ++pa)
            tma_load_2d(k_smem + st * C::TILE_BYTES + pa * C::PANEL_BYTES, &tm_col, k_full(st),
                        pa * 64, j * 128);
          bulk_load_1d(bits_smem + st * 512, p.bits_col + static_cast<size_t>(j) * 128, 512,
                       k_full(st));
          if (MODE == MODE_GRAD)
            bulk_load_1d(w_smem + st * 512, p.w_col + static_cast<size_t>(j) * 128, 512,
                         k_full(st));
          ++gcnt;
        }
        u = U0 + e;
        ++it;
      }
      if (p.dbg) { p.dbg[blockIdx.x * 16 + 8] = p_wait; p.dbg[blockIdx.x * 16 + 9] = clock64() - p_t0; }
      mbar_wait(q_empty, (it & 1) ^ 1);
      info[4 * (it & 1) + 0] = ITEM_END;
      mbar_arrive(q_full);
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer
    // (elect.sync, not lane == 0: ptxas then knows a single thread runs the region and feeds the
    // uniform-register operands of UTCHMMA without a per-instruction waterfall loop)
    if (elect_one()) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 64, 0, 0);
      constexpr uint32_t IDESC_G = umma_idesc_bf16(128, D, 0, 1);
      uint32_t gbase = 0;
      uint32_t cnt_t[2][2] = {{0, 0}, {0, 0}};
      long long i_kfull = 0, i_tfull = 0, i_t0 = p.dbg ? clock64() : 0;
      for (uint32_t it = 0;; ++it) {
        mbar_wait(q_full, it & 1);
        tc_fence_after();
        const uint32_t n_u = info[4 * (it & 1) + 0];
        if (n_u == ITEM_END) break;
        const int n = static_cast<int>(n_u);
        const int nk[2] = {n_stream(n, 0), n_stream(n, 1)};
        const int kmax = nk[0];  // nk[0] >= nk[1]

        auto stage_of = [&](int g, int k, uint32_t& par) {
          const uint32_t gg = gbase + tile_of(g, k);
          par = (gg / NST) & 1;
          return static_cast<int>(gg % NST);
        };
        auto issue_S = [&](int g, int k, int h) {
          uint32_t par;
          const int st = stage_of(g, k, par);
          { const long long w0 = p.dbg ? clock64() : 0;
            mbar_wait(k_full(st), par);
            if (p.dbg) i_kfull += clock64() - w0; }
          tc_fence_after();
          const uint32_t a_base = q_smem + (QT == 2 ? g * C::TILE_BYTES : 0);
          const uint32_t b_base = k_smem + st * C::TILE_BYTES + h * 8192;
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = (kk >> 2) * C::PANEL_BYTES + (kk & 3) * 32;
            mma_ss(tmem + g * 128 + h * 64, umma_desc_sw128(a_base + off, 16, 1024),
                   umma_desc_sw128(b_base + off, 16, 1024), IDESC_S, kk > 0);
          }
          tc_commit(s_full(g, h));
          // ROWSUM: the stage is free once the last S MMA that reads it has run
          if (MODE == MODE_ROWSUM && h == 1 && (QT == 1 || g == 1)) tc_commit(k_empty(st));
        };

        // prologue: both halves of every stream's first tile
        for (int h = 0; h < 2; ++h)
          for (int g = 0; g < 2; ++g)
            if (nk[g] > 0) issue_S(g, 0, h);
        if (kmax <= 1) tc_commit(q_empty);
        if (MODE == MODE_GRAD) {
          mbar_wait(acc_empty(0), (it & 1) ^ 1);
          if (QT == 2) mbar_wait(acc_empty(1), (it & 1) ^ 1);
          tc_fence_after();
        }
        for (int k = 0; k < kmax; ++k) {
          for (int h = 0; h < 2; ++h) {
            for (int g = 0; g < 2; ++g) {
              if (k >= nk[g]) continue;
              const bool more = (k + 1 < nk[g]);
              // GRAD: T_h is stored; ROWSUM: S_h has been consumed (also guards the next item's prologue)
              { const long long w0 = p.dbg ? clock64() : 0;
                mbar_wait(t_full(g, h), cnt_t[g][h] & 1);
                if (p.dbg) i_tfull += clock64() - w0; }
              tc_fence_after();
              ++cnt_t[g][h];
              if (MODE == MODE_GRAD) {
                uint32_t par;
                const int st = stage_of(g, k, par);
                const int acc = (QT == 2) ? g : 0;
                const bool first = (k == 0 && h == 0) && (QT == 2 || g == 0);
                const uint32_t b_base = k_smem + st * C::TILE_BYTES;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  mma_ts(tmem + C::ACC_COL + acc * D, tmem + g * 128 + h * 64 + kk * 8,
                         umma_desc_sw128(b_base + (h * 4 + kk) * 2048, C::PANEL_BYTES, 1024),
                         IDESC_G, (first && kk == 0) ? 0u : 1u);
                if (h == 1 && (QT == 1 || g == 1)) tc_commit(k_empty(st));
              }
              // D = 256 (two ring stages, one per stream): the stream's next tile reuses the stage
              // of its current one, which only frees after the G MMA of half 1
              if (QT == 1 && MODE == MODE_GRAD) {
                if (more && h == 1) {
                  issue_S(g, k + 1, 0);
                  issue_S(g, k + 1, 1);
                }
              } else if (more) {
                issue_S(g, k + 1, h);
              }
            }
          }
          if (k == kmax - 2) tc_commit(q_empty);  // the last S MMAs of the item are in flight
        }
        if (MODE == MODE_GRAD) {
          tc_commit(acc_full(0));
          if (QT == 2) tc_commit(acc_full(1));
        }
        gbase += n;
      }
      if (p.dbg) { p.dbg[blockIdx.x * 16 + 10] = i_kfull; p.dbg[blockIdx.x * 16 + 11] = i_tfull; p.dbg[blockIdx.x * 16 + 12] = clock64() - i_t0; }
    }
  } else {
    // ------------------------------------------------------------ epilogue groups
    const int g = warp >> 2;
    const int wq = warp & 3;
    const uint32_t lane_addr = static_cast<uint32_t>(wq * 32) << 16;
    const float a = p.scale_log2;
    uint32_t gbase = 0, cnt = 0;
    const bool timing = (p.dbg != nullptr) && warp == 0 && lane == 0;
    long long t_lat = 0, n_lat = 0, t_kfull = 0, t_sfull = 0, t_ld = 0, t_math = 0, t_q = 0, t_tail = 0, t_mark = timing ? clock64() : 0;
    auto lap = [&](long long& acc) { if (timing) { const long long t = clock64(); acc += t - t_mark; t_mark = t; } };
    for (uint32_t it = 0;; ++it) {
      mbar_wait(q_full, it & 1);
      lap(t_q);
      const uint32_t n_u = info[4 * (it & 1) + 0];
      if (n_u == ITEM_END) break;
      const int n = static_cast<int>(n_u);
      const int rb = static_cast<int>(info[4 * (it & 1) + 1]);
      const int slot = static_cast<int>(info[4 * (it & 1) + 2]);
      __syncwarp();
      if (lane == 0) mbar_arrive(q_empty);
      const int row_tile = rb * QT + (QT == 2 ? g : 0);
      const int row = row_tile * 128 + wq * 32 + lane;
      const uint32_t rbits = __ldg(p.bits_row + row);
      const uint32_t tor_r = __ldg(p.tor_row + row_tile);
      const uint32_t qmask = rbits ? ~rbits : 0u;
      float wrow = 0.f;
      if (MODE == MODE_GRAD) wrow = __ldg(p.w_row + row);
      const int my_n = n_stream(n, g);
      uint64_t racc[4] = {0ull, 0ull, 0ull, 0ull};
      for (int k = 0; k < my_n; ++k, ++cnt) {
        const uint32_t gg = gbase + tile_of(g, k);
        const int st = gg % NST;
        { const long long w0 = timing ? clock64() : 0;
          mbar_wait(k_full(st), (gg / NST) & 1);
          if (timing) { const long long w1 = clock64();
            if (w1 - w0 > 100) { t_lat += static_cast<uint32_t>(static_cast<uint32_t>(w1) - info[8 + NST + st]); ++n_lat; } } }
        lap(t_kfull);
        const uint32_t tor_c = info[8 + st];
        const bool full = tile_pair_full(tor_r, tor_c);
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          mbar_wait(s_full(g, h), cnt & 1);
          lap(t_sfull);
          tc_fence_after();
          const uint32_t s_addr = tmem + lane_addr + g * 128 + h * 64;
          const uint8_t* bsm = smem_gen + C::SMEM_BITS + st * 512 + h * 256;
          const uint8_t* wsm = smem_gen + C::SMEM_W + st * 512 + h * 256;
          uint32_t v[2][32];
          tmem_ld32(s_addr, v[0]);
          tmem_ld32(s_addr + 32, v[1]);
          tmem_wait_ld();
          lap(t_ld);
          if (MODE == MODE_ROWSUM) {  // S_h is in registers: hand the buffer back before the math
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(t_full(g, h));
          }
          if (full)
            pair_half_math<MODE, POLY, true>(v, s_addr, bsm, wsm, qmask, a, wrow, racc);
          else
            pair_half_math<MODE, POLY, false>(v, s_addr, bsm, wsm, qmask, a, wrow, racc);
          if (MODE == MODE_GRAD) {
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(t_full(g, h));
          }
          lap(t_math);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(k_empty(st));
      }
      if (MODE == MODE_ROWSUM) {
        float r0, r1, r2, r3;
        unpack2(add2(racc[0], racc[1]), r0, r1);
        unpack2(add2(racc[2], racc[3]), r2, r3);
        const float rsum = (r0 + r1) + (r2 + r3);
        if (my_n > 0 && rbits) atomicAdd(p.rowsum + row, rsum);
      } else {
        const int acc = (QT == 2) ? g : 0;
        constexpr int NCH = (QT == 2) ? (D / 32) : (D / 64);
        const int col0 = (QT == 2) ? 0 : g * (D / 2);
        float* dst = p.part + (static_cast<size_t>(slot) * p.rows_pad + row) * D + col0;
        mbar_wait(acc_full(acc), it & 1);
        tc_fence_after();
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
      lap(t_tail);
      gbase += n;
    }
    if (timing) {
      unsigned long long* d = p.dbg + blockIdx.x * 16;
      d[0] = t_q; d[1] = t_kfull; d[2] = t_sfull; d[3] = t_ld; d[4] = t_math; d[5] = t_tail; d[6] = cnt; d[13] = t_lat; d[14] = n_lat;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, 512);
}

}  // namespace mb200
