// Row-sum pass over a SYMMETRIC similarity matrix (rows == columns: the single-GPU dense InfoNCE).
//
// S = E E^T is symmetric and so is the label mask, so the masked exponential tile of the unit
// (row tile r, column tile c) is the transpose of the one of (c, r).  This pass therefore visits only
// the live units with c >= r.  For c > r the tile e = mask * exp2(a s - a) is used twice:
//   rows of r   : row sums, in registers (as in pair_kernel MODE_ROWSUM)
//   rows of c   : COLUMN sums of the same tile, on the tensor pipe: the epilogue writes e as bf16 into
//                 shared memory in the 128-B-swizzled [q rows][64 k] x 2 panel layout, and
//                       CS[k, 0..15] = sum_q e[q, k] * 1        tcgen05.mma, M = 128 (k), N = 16, K = 128 (q)
//                 A = the e tile viewed MN-major (M = k contiguous), B = a 2 KB tile of ones.  The sums land
//                 with TMEM lane = k, so every epilogue thread reads ONE word and adds it to rowsum[c*128 + k].
// Against the ordered pass this halves the S products and -- what actually bounds that pass, MUFU at
// 16 ex2/clk/SM -- the exponentials; the extra product costs 8 x 8 tensor cycles per unit.
//
// Structure as pair_kernel (one persistent CTA per SM; NG epilogue groups of four warps, one TMA producer
// warp, one MMA issuer warp), with these differences:
//   * work list = live units with c >= r, numbered by pair_plan_sym_body; a CTA takes an even share;
//   * two issuer threads (warps of their own), one for the S products and one for the column-sum products:
//     the streams are independent (S(k) needs its column tile and a free S buffer; CS(k) the group's e tile);
//   * a group drains the column sums of its unit k while it already works on unit k+NG (deferred, so
//     the column-sum product's latency is never waited for);
//   * the column-tile ring (2 stages at D = 128: the e tiles need the room) is shorter than the number of
//     groups, so a group has not seen the previous phase of the stage it waits for (PairCfg's
//     static_assert explains the hazard: a wait for phase n+1 is also satisfied while phase n-1 is the
//     last complete one).  The producer therefore publishes the unit number a stage is being filled
//     for (seq word, written before the barrier is armed) and a group first waits for ITS unit number.
//     A stage is released by its S product AND by the owning group once it has taken what it needs of
//     the class bits (measured on B200: making the stage private to producer and S issuer, with the bits
//     in a deeper ring of their own, was 8-25 % slower -- the coupling keeps the groups in step);
//   * the single-thread roles (producer, issuers) poll their barriers with test_wait instead of parking
//     in try_wait (MB_SYM_SPIN): same mean time, but without the occasional 20 % slower launch;
//   * S = R C^T takes R from TENSOR memory (tcgen05.mma with A in TMEM): group 0 copies the row tile from
//     its staging buffer into TMEM once per work item, and the S products then read only the column tile
//     from shared memory.  Shared-memory bandwidth is what bounds this pass (per unit: 32 KB TMA write,
//     32 KB e-tile stores, 36 KB column-sum operands, and 64 KB of S operands in the SS form -- 32 KB now).
// TMEM: S buffer b at [128 b, 128 b + 128), b < NB = 3; column sums of group g at [384 + 16 g, +16);
// the row tile at [384 + 16 NG, + D/2).
#pragma once
#include "pair_kernel.cuh"

// experiment switch (scripts/build_variants.sh -DMB_SYM_SPIN=n): bit 0 = producer / issuers poll with test_wait,
// bit 1 = the epilogue groups do too
#ifndef MB_SYM_SPIN
#define MB_SYM_SPIN 1
#endif

namespace mb200 {

__device__ __forceinline__ void sym_wait_role(uint32_t bar, uint32_t parity) {
  if (MB_SYM_SPIN & 1) mbar_wait_spin(bar, parity); else mbar_wait(bar, parity);
}
__device__ __forceinline__ void sym_wait_epi(uint32_t bar, uint32_t parity) {
  if (MB_SYM_SPIN & 2) mbar_wait_spin(bar, parity); else mbar_wait(bar, parity);
}

template <int D, int NG_ = 3>
struct SymCfg {
  static_assert(D == 64 || D == 128, "symmetric row-sum pass: projection dim 64 or 128");
  static_assert(NG_ == 2 || NG_ == 3, "two or three epilogue groups");
  static constexpr int NG = NG_;
  static constexpr int NB = 3;
  // shared memory decides the ring depth at D = 128: row tile 32 KB + NST x 32 KB + NG e tiles of 32 KB
  static constexpr int NST = (D == 128) ? (NG == 3 ? 2 : 3) : 4;
  static constexpr int TILE_BYTES = 128 * D * 2;
  static constexpr int PANEL_BYTES = 128 * 128;
  static constexpr int E_BYTES = 2 * PANEL_BYTES;          // bf16 [128 q][128 k] as two [128][64] panels
  static constexpr int SMEM_Q = 0;
  static constexpr int SMEM_K = SMEM_Q + TILE_BYTES;
  static constexpr int SMEM_E = SMEM_K + NST * TILE_BYTES;
  static constexpr int SMEM_ONES = SMEM_E + NG * E_BYTES;  // bf16 1.0 x 1024 (B operand of the column-sum product)
  static constexpr int NBR = NST;                          // class bits and unit record travel with the column-tile stage
  static constexpr int SMEM_BITS = SMEM_ONES + 2048;
  static constexpr int SMEM_CB = SMEM_BITS + NBR * 512;    // exponent-shift vectors [NG][2][BIAS_SLOTS][128]
  static constexpr int SMEM_CBM = SMEM_CB + 2 * NG * BIAS_SLOTS * 512;   // one all-masked vector (rows outside the loss)
  static constexpr int SMEM_BAR = SMEM_CBM + 512;
  static constexpr int NBAR = 2 * NST + 2 + 2 * NB + 2 * NG + 1;
  static constexpr int SMEM_INFO = SMEM_BAR + NBAR * 8;    // item headers [2][4], unit records [NBR][4]
  static constexpr int SMEM_TMEMPTR = SMEM_INFO + (8 + 4 * NBR) * 4;
  static constexpr int SMEM_BYTES = SMEM_TMEMPTR + 16 + 1024;
  static constexpr int CS_COL = NB * 128;
  static constexpr int R_COL = CS_COL + 16 * NG;           // the row tile as the A operand: bf16 [128 lanes][D/2 columns]
  static constexpr int WARP_PRODUCER = 4 * NG;
  static constexpr int WARP_ISSUER_S = 4 * NG + 1;
  static constexpr int WARP_ISSUER_CS = 4 * NG + 2;
  static constexpr int THREADS = (4 * NG + 4) * 32;
  // (registers are allocated to warps in fours: 14 warps cost what 16 do, so 128 registers per thread is the
  // limit with three groups -- a __maxnreg__(144) build does not launch -- and 168 with two)
  static_assert(R_COL + D / 2 <= 512, "TMEM budget");
  static_assert(NG <= NB, "a group must have seen the previous phase of its S buffer");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

// unit_off[rb] = number of live units (rb', c >= rb') with rb' < rb; unit_off[n] = total.  The tiles are
// class-sorted, so the single-class tiles of one class and the empty tiles (invalid + pad rows only) are
// contiguous runs: counts "at or after rb" come from each run's first index and length.
__device__ __forceinline__ void pair_plan_sym_body(const uint32_t* tor, int n_tiles, uint32_t* unit_off) {
  __shared__ int run_lo[33];     // [c] first single-class tile of class c; [32] first empty tile
  __shared__ int run_cnt[33];
  __shared__ uint32_t warp_tot_s[32];
  __shared__ uint32_t carry_s;
  if (threadIdx.x < 33) { run_lo[threadIdx.x] = 0x7fffffff; run_cnt[threadIdx.x] = 0; }
  if (threadIdx.x == 0) carry_s = 0u;
  __syncthreads();
  for (int j = threadIdx.x; j < n_tiles; j += blockDim.x) {
    const uint32_t c = tor[j] & TOR_CLASSES;
    int slot = -1;
    if (c == 0u) slot = 32;
    else if ((c & (c - 1u)) == 0u) slot = 31 - __clz(c);
    if (slot >= 0) { atomicMin(&run_lo[slot], j); atomicAdd(&run_cnt[slot], 1); }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n_tiles; base += blockDim.x) {
    const int rb = base + threadIdx.x;
    uint32_t live = 0;
    if (rb < n_tiles) {
      const uint32_t r = tor[rb] & TOR_CLASSES;
      if (r != 0u) {
        auto at_or_after = [&](int slot) {   // tiles of the run with index >= rb
          const int lo = run_lo[slot], cnt = run_cnt[slot];
          if (cnt == 0) return 0;
          const int first = rb > lo ? rb : lo;
          const int left = lo + cnt - first;
          return left > 0 ? left : 0;
        };
        int l = (n_tiles - rb) - at_or_after(32);
        if ((r & (r - 1u)) == 0u) l -= at_or_after(31 - __clz(r));
        live = static_cast<uint32_t>(l > 0 ? l : 0);
      }
    }
    uint32_t inc = live;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += u;
    }
    if (lane == 31) warp_tot_s[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += warp_tot_s[w];
    const uint32_t before = carry_s + woff + inc - live;
    if (rb < n_tiles) unit_off[rb] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry_s = before + live;
    __syncthreads();
  }
  if (threadIdx.x == 0) unit_off[n_tiles] = carry_s;
}

// Which rank computes the unordered tile pair {row tile r of rank g, column tile j}?  Over G ranks the pixel
// blocks form a ring: a rank takes its own block's pairs with column >= row, every pair with the next
// ceil((G-1)/2) blocks -- and, for even G, one half of the pairs with the block opposite (by the parity of the
// two local tile indices; the opposite rank takes the other half).  Every unordered pair has exactly one owner,
// and every rank owns the same number of (row tile, column tile) slots.  T = tiles per rank.
__host__ __device__ inline bool sym_owned(int r_loc, int j, int T, int g, int G) {
  const int bj = j / T, jl = j - bj * T;
  int o = bj - g;
  if (o < 0) o += G;
  if (o == 0) return jl >= r_loc;
  if (2 * o < G) return true;
  if (2 * o == G) return ((r_loc + jl) & 1) == (g < bj ? 0 : 1);
  return false;
}

// Multi-rank plan, step 1: cnt[rb] = live units the rank owns in row block rb (one warp per row block; the column
// tiles are visited from the diagonal tile on, wrapping around -- the order the producer uses).  Step 2
// (pair_plan_scan_kernel): exclusive scan -> unit_off.
__global__ void __launch_bounds__(256) pair_plan_sym_count_kernel(const uint32_t* tor_row, int n_row_blocks,
                                                                  const uint32_t* tor_col, int n_col_tiles, int g, int G,
                                                                  uint32_t* cnt) {
  pdl_trigger();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rb = blockIdx.x * 8 + warp;
  if (rb >= n_row_blocks) return;
  const uint32_t tor_r = tor_row[rb];
  uint32_t c = 0;
  for (int j = lane; j < n_col_tiles; j += 32)
    c += (sym_owned(rb, j, n_row_blocks, g, G) && !tile_pair_dead(tor_r, tor_col[j])) ? 1u : 0u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) cnt[rb] = c;
}
__global__ void __launch_bounds__(1024) pair_plan_scan_kernel(const uint32_t* cnt, int n, uint32_t* unit_off) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0u;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < n ? cnt[i] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += u;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += warp_tot[w];
    const uint32_t before = carry + woff + inc - v;
    if (i < n) unit_off[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) unit_off[n] = carry;
}

// One block of 1024 threads (the single-GPU composite runs the same body at the end of class_scan_kernel).
__global__ void __launch_bounds__(1024) pair_plan_sym_kernel(const uint32_t* tor, int n_tiles, uint32_t* unit_off) {
  pdl_trigger();
  pdl_wait();
  pair_plan_sym_body(tor, n_tiles, unit_off);
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
  uint32_t v;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
  return v;
}

// One 64-column half of a unit for one thread (= one row q): e = mask * exp2(a s + shift), row sums into
// racc, and (STORE) the bf16 tile row into the e panel at e_row (shared address of this row's 128-byte
// line; its 16-byte chunks are XOR-swizzled with swz = row & 7, the layout TMA's 128-B swizzle writes).
template <int POLY, int MASK, bool STORE>
__device__ __forceinline__ void sym_half_math(uint32_t (&v)[2][32], uint32_t e_row, uint32_t swz,
                                              const uint8_t* bits_sm, const uint8_t* bias_sm,
                                              uint32_t qmask, float a, uint64_t (&racc)[4]) {
  const uint64_t a2 = pack2(a, a);
  const uint64_t noff2 = pack2(-a, -a);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint4* kb4 = reinterpret_cast<const uint4*>(bits_sm) + c * 8;
    const float4* kc4 = reinterpret_cast<const float4*>(bias_sm) + c * 8;
    uint32_t packed[16];
#pragma unroll
    for (int q4 = 0; q4 < 8; ++q4) {
      uint64_t e[2];
      uint64_t off2[2] = {noff2, noff2};
      if (MASK == MASK_BIAS) {
        const float4 kc = kc4[q4];
        off2[0] = pack2(kc.x, kc.y);
        off2[1] = pack2(kc.z, kc.w);
      }
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint64_t x = fma2(pack2u(v[c][q4 * 4 + 2 * s], v[c][q4 * 4 + 2 * s + 1]), a2, off2[s]);
        if (((q4 * 2 + s) & 3) < POLY) {
          e[s] = ex2_poly2(x);
        } else {
          float x0, x1;
          unpack2(x, x0, x1);
          e[s] = pack2(ex2_approx(x0), ex2_approx(x1));
        }
      }
      float e0, e1, e2, e3;
      if (MASK == MASK_SELECT || STORE) {
        unpack2(e[0], e0, e1);
        unpack2(e[1], e2, e3);
      }
      if (MASK == MASK_SELECT) {
        const uint4 kb = kb4[q4];
        e0 = (kb.x & qmask) ? e0 : 0.f;
        e1 = (kb.y & qmask) ? e1 : 0.f;
        e2 = (kb.z & qmask) ? e2 : 0.f;
        e3 = (kb.w & qmask) ? e3 : 0.f;
        e[0] = pack2(e0, e1);
        e[1] = pack2(e2, e3);
      }
      racc[(q4 * 2) & 3] = add2(racc[(q4 * 2) & 3], e[0]);
      racc[(q4 * 2 + 1) & 3] = add2(racc[(q4 * 2 + 1) & 3], e[1]);
      if (STORE) {
        packed[q4 * 2 + 0] = pack_bf16x2(e0, e1);
        packed[q4 * 2 + 1] = pack_bf16x2(e2, e3);
      }
    }
    if (STORE) {
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const uint32_t chunk = static_cast<uint32_t>(c * 4 + jj);
        st_shared_v4(e_row + ((chunk ^ swz) << 4), packed[4 * jj], packed[4 * jj + 1], packed[4 * jj + 2],
                     packed[4 * jj + 3]);
      }
    }
  }
}

template <int D, int POLY, int NG_>
__global__ void __launch_bounds__((SymCfg<D, NG_>::THREADS), 1)
pair_sym_kernel(const __grid_constant__ CUtensorMap tm_row, const __grid_constant__ CUtensorMap tm_col,
                const PairParams p) {
  using C = SymCfg<D, NG_>;
  constexpr int NST = C::NST;
  constexpr int NB = C::NB;
  constexpr int NG = C::NG;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t q_smem = smem_base + C::SMEM_Q;
  const uint32_t k_smem = smem_base + C::SMEM_K;
  const uint32_t e_smem = smem_base + C::SMEM_E;
  const uint32_t ones_smem = smem_base + C::SMEM_ONES;
  const uint32_t bits_smem = smem_base + C::SMEM_BITS;
  const uint32_t bar0 = smem_base + C::SMEM_BAR;
  auto k_full = [&](int s) { return bar0 + 8u * s; };
  auto k_empty = [&](int s) { return bar0 + 8u * (NST + s); };
  const uint32_t q_full = bar0 + 8u * (2 * NST);
  const uint32_t q_empty = bar0 + 8u * (2 * NST + 1);
  auto s_full = [&](int b) { return bar0 + 8u * (2 * NST + 2 + b); };
  auto s_free = [&](int b) { return bar0 + 8u * (2 * NST + 2 + NB + b); };
  auto e_full = [&](int g) { return bar0 + 8u * (2 * NST + 2 + 2 * NB + g); };
  auto cs_full = [&](int g) { return bar0 + 8u * (2 * NST + 2 + 2 * NB + NG + g); };
  const uint32_t r_full = bar0 + 8u * (2 * NST + 2 + 2 * NB + 2 * NG);   // row tile copied into TMEM
  volatile uint32_t* info = reinterpret_cast<volatile uint32_t*>(smem_gen + C::SMEM_INFO);
  // info[4*(it&1) + {0,1,2}]: live units, row block of item `it`, first unit is the diagonal one; info[8 + 4 st + {0,1,2}]: class union and index of the stage's column tile, number of the unit it is filled for
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem_gen + C::SMEM_TMEMPTR);
  __shared__ uint2 ulist_sm[32];           // producer warp: (class union, index) of the next column tiles to load

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(k_full(s), 1);
      mbar_init(k_empty(s), 1 + 4);      // S product done + the four warps of the group that owns the unit
    }
    mbar_init(q_full, 1);
    mbar_init(q_empty, 2 + 4 * NG);      // S issuer (last S of the item) + column-sum issuer + every epilogue warp
    for (int b = 0; b < NB; ++b) {
      mbar_init(s_full(b), 1);
      mbar_init(s_free(b), 4);
    }
    for (int g = 0; g < NG; ++g) {
      mbar_init(e_full(g), 4);
      mbar_init(cs_full(g), 1);
    }
    mbar_init(r_full, 4);
    fence_barrier_init();
  }
  // constants in shared memory: the ones tile and the all-masked shift vector
  for (int i = threadIdx.x; i < 512; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem_gen + C::SMEM_ONES)[i] = 0x3f803f80u;
  for (int i = threadIdx.x; i < 128; i += blockDim.x)
    reinterpret_cast<float*>(smem_gen + C::SMEM_CBM)[i] = -p.scale_log2 - p.mask_shift;
  if (threadIdx.x < C::NBR) info[8 + 4 * threadIdx.x + 2] = 0xffffffffu;
  fence_proxy_async_smem();
  if (warp == C::WARP_ISSUER_S) {
    tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_ptr_smem)), 512);
    tmem_relinquish();
  }
  if (warp == C::WARP_PRODUCER && lane == 0) {
    tma_prefetch_desc(&tm_row);
    tma_prefetch_desc(&tm_col);
  }
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;
  pdl_wait();

  if (warp == C::WARP_PRODUCER) {
    // ------------------------------------------------------------ TMA producer + work partition
    // this launch's part of the unit list (all of it, or one of share_G equal parts: ranks that split ONE global
    // list evenly finish together whatever their own images' class mix), then this CTA's even share of that
    const uint32_t all_units = __ldg(p.unit_off + p.n_row_blocks);
    const uint32_t sG = p.share_G > 1 ? static_cast<uint32_t>(p.share_G) : 1u;
    const uint32_t sg = p.share_G > 1 ? static_cast<uint32_t>(p.share_g) : 0u;
    const uint32_t base = static_cast<uint32_t>(static_cast<uint64_t>(all_units) * sg / sG);
    const uint32_t total = static_cast<uint32_t>(static_cast<uint64_t>(all_units) * (sg + 1) / sG);
    const uint32_t per = (total - base + gridDim.x - 1) / gridDim.x;
    const uint64_t lo64 = base + static_cast<uint64_t>(per) * blockIdx.x;
    uint32_t u = lo64 < total ? static_cast<uint32_t>(lo64) : total;
    const uint32_t u1 = (lo64 + per < total) ? static_cast<uint32_t>(lo64 + per) : total;
    int rb = 0;
    {
      int lo = 0, hi = p.n_row_blocks;
      while (hi - lo > 1) {
        const int span = hi - lo;
        const int step = (span + 31) / 32;
        const int probe = lo + step * (lane + 1);
        const bool le = (probe < hi) && (__ldg(p.unit_off + probe) <= u);
        const uint32_t m = __ballot_sync(0xffffffffu, le);
        const int k = __popc(m);
        const int new_lo = lo + step * k;
        const int new_hi = lo + step * (k + 1);
        lo = new_lo;
        hi = new_hi < hi ? new_hi : hi;
      }
      rb = lo;
    }
    uint32_t gcnt = 0, it = 0;
    long long p_wait = 0, p_t0 = (kPairTiming && p.dbg) ? clock64() : 0;
    while (u < u1) {
      const uint32_t U0 = __ldg(p.unit_off + rb), U1 = __ldg(p.unit_off + rb + 1);
      if (U1 <= u) { ++rb; continue; }
      const uint32_t a = u - U0;
      const uint32_t e = (u1 < U1 ? u1 : U1) - U0;
      const uint32_t tor_r = __ldg(p.tor_row + rb);
      if (lane == 0) {
        mbar_wait(q_empty, (it & 1) ^ 1);
        info[4 * (it & 1) + 0] = e - a;
        info[4 * (it & 1) + 1] = static_cast<uint32_t>(rb);
        info[4 * (it & 1) + 2] = (a == 0u && !tile_pair_dead(tor_r, tor_r)) ? 1u : 0u;   // first unit = the diagonal one
        mbar_arrive_expect_tx(q_full, C::TILE_BYTES);
        tma_load_3d(q_smem, &tm_row, q_full, 0, rb * 128, 0);
      }
      uint32_t li = 0;
      // column tiles from the diagonal tile on, wrapping around (so the diagonal unit, if live, is the item's first)
      const int jstart = rb + p.rb_off;
      auto col_of = [&](int t) { const int j = jstart + t; return j >= p.n_col_tiles ? j - p.n_col_tiles : j; };
      uint32_t tc_next = (lane < p.n_col_tiles) ? __ldg(p.tor_col + col_of(lane)) : 0u;
      for (int t0 = 0; t0 < p.n_col_tiles && li < e; t0 += 32) {
        const uint32_t tc = tc_next;
        const int tn = t0 + 32 + lane;
        tc_next = (tn < p.n_col_tiles) ? __ldg(p.tor_col + col_of(tn)) : 0u;
        const int tme = t0 + lane;
        const int jme = tme < p.n_col_tiles ? col_of(tme) : 0;
        uint32_t live = __ballot_sync(0xffffffffu, tme < p.n_col_tiles &&
                                                       sym_owned(rb, jme, p.n_row_blocks, p.own_g, p.own_G) &&
                                                       !tile_pair_dead(tor_r, tc));
        const uint32_t cnt = __popc(live);
        if (li + cnt <= a) { li += cnt; continue; }
        // The lanes whose tile is one of the CTA's units write it into a 32-entry list (one step for up to 32
        // units); lane 0 then only waits for a free stage and issues the copies.  (Walking the ballot word bit by
        // bit on the issuing path cost ~900 cycles of this warp per unit -- it shares its scheduler with three busy
        // epilogue warps -- and with a two-stage ring that latency was exposed on every unit.)
        const uint32_t rank = __popc(live & ((1u << lane) - 1u));
        const uint32_t idx = li + rank;                         // live index of this lane's tile in the row block
        const uint32_t first = li > a ? li : a;
        const uint32_t last = (li + cnt < e) ? li + cnt : e;    // units [first, last) of the row block come from this word
        if (((live >> lane) & 1u) && idx >= first && idx < last)
          ulist_sm[idx - first] = make_uint2(tc, static_cast<uint32_t>(jme));
        __syncwarp();
        if (lane == 0) {
          for (uint32_t t = 0; t < last - first; ++t) {
            const uint2 ent = ulist_sm[t];
            const uint32_t g = gcnt + t;
            const int st = static_cast<int>(g % NST);
            { const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
              sym_wait_role(k_empty(st), ((g / NST) & 1) ^ 1);
              if (kPairTiming && p.dbg) p_wait += clock64() - w0; }
            info[8 + 4 * st] = ent.x;
            info[8 + 4 * st + 1] = ent.y;
            info[8 + 4 * st + 2] = g;
            mbar_arrive_expect_tx(k_full(st), C::TILE_BYTES + 512);
            tma_load_3d(k_smem + st * C::TILE_BYTES, &tm_col, k_full(st), 0, static_cast<int>(ent.y) * 128, 0);
            bulk_load_1d(bits_smem + st * 512, p.bits_col + static_cast<size_t>(ent.y) * 128, 512, k_full(st));
          }
        }
        gcnt += last - first;
        li += cnt;
        __syncwarp();
      }
      u = U0 + e;
      ++it;
    }
    if (lane == 0) {
      if (kPairTiming && p.dbg) { p.dbg[blockIdx.x * 16 + 8] = p_wait; p.dbg[blockIdx.x * 16 + 9] = clock64() - p_t0; p.dbg[blockIdx.x * 16 + 10] = gcnt; }
      mbar_wait(q_empty, (it & 1) ^ 1);
      info[4 * (it & 1) + 0] = ITEM_END;
      mbar_arrive(q_full);
    }
  } else if (warp == C::WARP_ISSUER_S) {
    // ------------------------------------------------------------ S = R C^T issuer
    if (elect_one()) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
      uint32_t gS = 0;
      long long i_t0 = (kPairTiming && p.dbg) ? clock64() : 0, i_wait = 0, i_wait_k = 0;
      for (uint32_t it = 0;; ++it) {
        mbar_wait(q_full, it & 1);
        const uint32_t n_u = info[4 * (it & 1) + 0];
        if (n_u == ITEM_END) break;
        const int n = static_cast<int>(n_u);
        mbar_wait(r_full, it & 1);       // group 0 has put the item's row tile into TMEM
        tc_fence_after();
        for (int k = 0; k < n; ++k, ++gS) {
          const int st = static_cast<int>(gS % NST);
          const int b = static_cast<int>(gS % NB);
          { const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
            sym_wait_role(k_full(st), (gS / NST) & 1);
            const long long w1 = (kPairTiming && p.dbg) ? clock64() : 0;
            sym_wait_role(s_free(b), ((gS / NB) & 1) ^ 1);
            if (kPairTiming && p.dbg) { i_wait += clock64() - w0; i_wait_k += w1 - w0; } }
          tc_fence_after();
          const uint32_t b_base = k_smem + st * C::TILE_BYTES;
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = (kk >> 2) * C::PANEL_BYTES + (kk & 3) * 32;
            mma_ts(tmem + b * 128, tmem + C::R_COL + kk * 8, umma_desc_sw128(b_base + off, 16, 1024), IDESC_S, kk > 0);
          }
          tc_commit(s_full(b));
          tc_commit(k_empty(st));
          if (k == n - 1) tc_commit(q_empty);   // the item's last S is in flight: the TMEM row tile may be replaced after it
        }
      }
      if (kPairTiming && p.dbg) { p.dbg[blockIdx.x * 16 + 11] = i_wait; p.dbg[blockIdx.x * 16 + 12] = clock64() - i_t0; p.dbg[blockIdx.x * 16 + 7] = i_wait_k; }
    }
  } else if (warp == C::WARP_ISSUER_CS) {
    // ------------------------------------------------------------ column-sum issuer (its own thread: the two
    // product streams are independent, and one thread polling for both spent more time on its own
    // instructions -- it shares a scheduler with three busy epilogue warps -- than the tensor pipe on the products)
    if (elect_one()) {
      constexpr uint32_t IDESC_CS = umma_idesc_bf16(128, 16, 1, 0);   // A = e tile MN-major, B = ones K-major
      uint32_t gC = 0;
      uint32_t ce[NG];                  // column-sum products issued per group
#pragma unroll
      for (int g = 0; g < NG; ++g) ce[g] = 0;
      long long c_t0 = (kPairTiming && p.dbg) ? clock64() : 0, c_wait = 0;
      for (uint32_t it = 0;; ++it) {
        mbar_wait(q_full, it & 1);
        const uint32_t n_u = info[4 * (it & 1) + 0];
        if (n_u == ITEM_END) break;
        const int n = static_cast<int>(n_u);
        const bool diag0 = info[4 * (it & 1) + 2] != 0u;      // the item's first unit is the diagonal one
        mbar_arrive(q_empty);                                 // header read
        for (int k = 0; k < n; ++k, ++gC) {
          if (k == 0 && diag0) continue;
          const int g = static_cast<int>(gC % NG);
          { const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
            sym_wait_role(e_full(g), ce[g] & 1);
            if (kPairTiming && p.dbg) c_wait += clock64() - w0; }
          tc_fence_after();
          const uint32_t a_base = e_smem + g * C::E_BYTES;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            mma_ss(tmem + C::CS_COL + 16 * g, umma_desc_sw128(a_base + kk * 2048, C::PANEL_BYTES, 1024),
                   umma_desc_sw128(ones_smem, 16, 1024), IDESC_CS, kk > 0);
          tc_commit(cs_full(g));
          ++ce[g];
        }
      }
      if (kPairTiming && p.dbg) { p.dbg[blockIdx.x * 16 + 13] = c_wait; p.dbg[blockIdx.x * 16 + 14] = clock64() - c_t0; }
    }
  } else if (warp >= C::WARP_ISSUER_CS + 1) {
    // (a fourth non-epilogue warp: registers are allocated to warps in fours, so it is free; it has nothing to do)
  } else {
    // ------------------------------------------------------------ epilogue groups
    const int g = warp >> 2;
    const int wq = warp & 3;
    const int trow = wq * 32 + lane;                       // row of the tile = TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>(wq * 32) << 16;
    const float a = p.scale_log2;
    const uint32_t e_row0 = e_smem + g * C::E_BYTES + trow * 128;
    const uint32_t swz = static_cast<uint32_t>(trow & 7);
    const uint8_t* cbm = smem_gen + C::SMEM_CBM;
    uint32_t gbase = 0, n_bias = 0, n_cs = 0;
    bool pending = false, pend_valid = false;
    int pend_j = 0;
    // cycle counters of one thread (timing builds, MUSCLE_PAIR_TIMING=1): [0] item wait, [1] column tile, [2] S,
    // [3] column-sum drain, [4] tcgen05.ld, [5] math + e-tile stores, [6] own units, [7] off-diagonal own units
    const bool timing = kPairTiming && (p.dbg != nullptr) && warp == 0 && lane == 0;
    long long tacc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // [8] e-tile fence + publish
    long long t_mark = timing ? clock64() : 0;
    auto lap = [&](int i) { if (timing) { const long long t = clock64(); tacc[i] += t - t_mark; t_mark = t; } };
    auto drain = [&]() {   // column sums of this group's previous off-diagonal unit -> rowsum of its column pixels
      sym_wait_epi(cs_full(g), n_cs & 1);
      ++n_cs;
      tc_fence_after();
      const float cs = __uint_as_float(tmem_ld1(tmem + lane_addr + C::CS_COL + 16 * g));
      tmem_wait_ld();
      tc_fence_before();
      // (a masked column keeps 2^-shift of its terms: pixels outside the loss are skipped explicitly)
      if (cs != 0.f && pend_valid) atomicAdd(p.rowsum_cols + static_cast<size_t>(pend_j) * 128 + trow, cs);
      pending = false;
    };
    for (uint32_t it = 0;; ++it) {
      mbar_wait(q_full, it & 1);
      lap(0);
      const uint32_t n_u = info[4 * (it & 1) + 0];
      if (n_u == ITEM_END) break;
      const int n = static_cast<int>(n_u);
      const int rb = static_cast<int>(info[4 * (it & 1) + 1]);
      if (g == 0) {
        // row tile: staging buffer (128-B-swizzled [128 rows][64] panels) -> TMEM, lane = row, column = k / 2
        const uint32_t src = q_smem + trow * 128;
#pragma unroll
        for (int pnl = 0; pnl < D / 64; ++pnl) {
          uint32_t rv[2][16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t addr = src + pnl * C::PANEL_BYTES + ((static_cast<uint32_t>(j) ^ swz) << 4);
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(rv[j >> 2][(j & 3) * 4]), "=r"(rv[j >> 2][(j & 3) * 4 + 1]),
                           "=r"(rv[j >> 2][(j & 3) * 4 + 2]), "=r"(rv[j >> 2][(j & 3) * 4 + 3])
                         : "r"(addr));
          }
          tmem_st16(tmem + lane_addr + C::R_COL + pnl * 32, rv[0]);
          tmem_st16(tmem + lane_addr + C::R_COL + pnl * 32 + 16, rv[1]);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(r_full);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(q_empty);
      const int row = rb * 128 + trow;
      const uint32_t rbits = __ldg(p.bits_row + row);
      const uint32_t tor_r = __ldg(p.tor_row + rb);
      const uint32_t qmask = rbits ? ~rbits : 0u;
      const bool bias_ok = __popc(tor_r & TOR_CLASSES) <= BIAS_SLOTS;
      const int my_slot = rbits ? __popc(tor_r & TOR_CLASSES & (rbits - 1u)) : 0;
      uint64_t racc[4] = {0ull, 0ull, 0ull, 0ull};
      for (int k = 0; k < n; ++k) {
        const uint32_t gk = gbase + k;
        const int st = static_cast<int>(gk % NST);
        if (gk % NG != static_cast<uint32_t>(g)) continue;
        {   // the stage is being (or has been) filled for THIS unit: only then is the phase parity unambiguous
          uint32_t spins = 0;
          while (info[8 + 4 * st + 2] != gk) {
            if (++spins > (1u << 26)) {
              printf("pair_sym seq timeout: block %d thread %d unit %u\n", blockIdx.x, threadIdx.x, gk);
              __trap();
            }
          }
        }
        sym_wait_epi(k_full(st), (gk / NST) & 1);
        lap(1);
        const int b = static_cast<int>(gk % NB);
        const uint32_t tor_c = info[8 + 4 * st];
        const int j = static_cast<int>(info[8 + 4 * st + 1]);
        const bool diag = (j == rb + p.rb_off);
        const bool col_valid = reinterpret_cast<const uint32_t*>(smem_gen + C::SMEM_BITS + st * 512)[trow] != 0u;
        // no element masked, in either direction: disjoint class sets, every column and every row in the loss
        const bool full = (tor_r & tor_c & TOR_CLASSES) == 0u && ((tor_c | tor_r) & TOR_INVALID) == 0u;
        const int mask = full ? MASK_NONE : (bias_ok ? MASK_BIAS : MASK_SELECT);
        const uint8_t* cb = smem_gen + C::SMEM_CB + ((g * 2 + (n_bias & 1)) * BIAS_SLOTS) * 512;
        if (mask == MASK_BIAS) {
          const uint32_t kb = reinterpret_cast<const uint32_t*>(smem_gen + C::SMEM_BITS + st * 512)[trow];
          float* cbw = reinterpret_cast<float*>(const_cast<uint8_t*>(cb)) + trow;
          uint32_t rem = tor_r & TOR_CLASSES;
          for (int s = 0; rem != 0u; ++s, rem &= rem - 1u)
            cbw[s * 128] = (kb & ~(rem & (0u - rem))) ? -a : -a - p.mask_shift;
          ++n_bias;
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        }
        if (mask != MASK_SELECT) {   // no class bits left to read: the stage is only the tensor pipe's now
          __syncwarp();
          if (lane == 0) mbar_arrive(k_empty(st));
        }
        sym_wait_epi(s_full(b), (gk / NB) & 1);
        lap(2);
        tc_fence_after();
        if (!diag && pending) drain();    // the e tile is about to be overwritten: its product must be done
        lap(3);
        if (timing) { ++tacc[6]; tacc[7] += diag ? 0 : 1; }
        const uint8_t* my_cb = rbits ? cb + my_slot * 512 : cbm;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          const uint32_t s_addr = tmem + lane_addr + b * 128 + h * 64;
          const uint8_t* bsm = smem_gen + C::SMEM_BITS + st * 512 + h * 256;
          const uint32_t e_row = e_row0 + h * C::PANEL_BYTES;
          uint32_t v[2][32];
          tmem_ld32(s_addr, v[0]);
          tmem_ld32(s_addr + 32, v[1]);
          tmem_wait_ld();
          lap(4);
          if (h == 1) {   // S is in registers: hand the buffer back before the math
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_free(b));
          }
          if (diag) {
            if (mask == MASK_BIAS)
              sym_half_math<POLY, MASK_BIAS, false>(v, e_row, swz, bsm, my_cb + h * 256, qmask, a, racc);
            else
              sym_half_math<POLY, MASK_SELECT, false>(v, e_row, swz, bsm, cb, qmask, a, racc);
          } else if (mask == MASK_NONE) {
            sym_half_math<POLY, MASK_NONE, true>(v, e_row, swz, bsm, cb, qmask, a, racc);
          } else if (mask == MASK_BIAS) {
            sym_half_math<POLY, MASK_BIAS, true>(v, e_row, swz, bsm, my_cb + h * 256, qmask, a, racc);
          } else {
            sym_half_math<POLY, MASK_SELECT, true>(v, e_row, swz, bsm, cb, qmask, a, racc);
          }
          lap(5);
        }
        if (!diag) {
          fence_proxy_async_smem();     // the e tile is read by the tensor pipe (async proxy)
          __syncwarp();
          if (lane == 0) mbar_arrive(e_full(g));
          lap(8);
          pending = true;
          pend_j = j;
          pend_valid = col_valid;
        }
        if (mask == MASK_SELECT) {
          __syncwarp();
          if (lane == 0) mbar_arrive(k_empty(st));
        }
      }
      {
        float r0, r1, r2, r3;
        unpack2(add2(racc[0], racc[1]), r0, r1);
        unpack2(add2(racc[2], racc[3]), r2, r3);
        const float rsum = (r0 + r1) + (r2 + r3);
        if (rbits && rsum != 0.f) atomicAdd(p.rowsum + row, rsum);
      }
      gbase += n;
    }
    if (pending) drain();
    if (timing) {
      unsigned long long* d = p.dbg + blockIdx.x * 16;
      for (int i = 0; i < 7; ++i) d[i] = static_cast<unsigned long long>(tacc[i]);
      d[15] = static_cast<unsigned long long>(tacc[8]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == C::WARP_ISSUER_S) tmem_dealloc(tmem, 512);
}

}  // namespace mb200
