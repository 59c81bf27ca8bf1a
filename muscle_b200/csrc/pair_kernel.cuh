// The dense pixel-pair mainloop of the InfoNCE row-sum pass and the InfoNCE gradient pass.
//
// One persistent CTA per SM, warp-specialised (NG epilogue groups: 3 in the row-sum pass, 2 in the
// gradient pass and for D = 256, see PairCfg):
//   warps 0 .. 4 NG-1   epilogue groups of four warps (TMEM lanes 32*(warp%4) .. +31, thread = one row
//                       of the 128-row tile); group g takes every NG-th unit
//   warp  4 NG          TMA producer      (the row tile once per work item, column tiles through a ring)
//   warp  4 NG + 1      tcgen05.mma issuer (one lane) + TMEM allocator
//
// Pixels arrive sorted by class (row_kernels.cuh), and a per-tile class union (`tor`) lets the
// producer drop every (row tile, column tile) whose pairs are all masked (same single class)
// before anything is loaded or multiplied; the epilogue takes a mask-free path for tiles whose
// class sets are disjoint.  The LIVE (row tile, column tile) units are numbered row tile by row
// tile (`unit_off`, pair_plan_kernel) and every CTA takes one contiguous share of that range, so
// the SMs finish together whatever the class mix; a CTA's share of one row tile is a work item,
// and the item's index among the CTAs that touch the row tile is its output `slot`.
//
// A unit is one 128 x 128 similarity tile.  Its S = R C^T goes into one of NB TMEM buffers of
// 128 columns (round robin over the CTA's units), the two epilogue groups take alternate units, so
// while a group works on unit k the tensor pipe fills the buffers of units k+1 .. k+NB-1:
//   S = R C^T          tcgen05.mma SS, bf16 x bf16 -> fp32 in TMEM            [128 x 128], N = 128
//   epilogue           tcgen05.ld S (64 columns at a time), e = exp2(a*s - a), label mask, then
//       MODE_ROWSUM : row-sum of e kept in registers, one atomicAdd per row per item and group
//       MODE_GRAD   : T = e * (w_row + w_col) -> bf16 -> tcgen05.st over the first 64 columns of S
//   (MODE_GRAD) G += T C   tcgen05.mma TS: A = T from TMEM, B = the same column rows, MN-major
// so the N x N logits never leave the SM.  (Measured on B200, scripts/ubench_tc.cu: an N = 128 SS
// MMA costs 64 cycles where two N = 64 ones cost 96, and one barrier round trip between the
// epilogue and the issuer is ~1000 cycles: whole tiles and NB-deep buffering hide it, the earlier
// 64-column ping-pong could not.)  A share of the exponentials is evaluated with a cubic on the
// FMA pipe (ptx.cuh: ex2_poly2) because MUFU alone (16/clk/SM) would cap the pass below the tensor
// pipe; all elementwise math is packed fp32x2.
//
// TMEM columns: S buffer b at [128 b, 128 b + 128); MODE_GRAD: G accumulator [128 x D] after them.
//
// Variants of the gradient pass that were built, verified and measured slower or equal on B200
// (config 2, 125 us as is): a third epilogue group at 128 registers (128 us); T in separate 64-column
// TMEM buffers with two S buffers, so that S(k+2) is issued as soon as S(k) has been read (125 us:
// the "T done -> G(k) -> S(k+3)" chain is not what bounds the pass); that plus both groups on every
// unit, 64 columns each, to halve the time a column tile is held between S(k) and G(k) (151 us: four
// more barrier operations per unit and group cost more than the shorter residence buys); the next
// tcgen05.ld issued under the math of the current 32-column chunk (+3 %: more registers, no gain).
#pragma once
// Cycle counters of the pair kernel's phases (printed by infonce.cu under MUSCLE_PAIR_TIMING=1) cost
// ~20 registers per epilogue thread: compiled in only with -DMB_PAIR_TIMING=1 (scripts/build_variants.sh).
#ifndef MB_PAIR_TIMING
#define MB_PAIR_TIMING 0
#endif
#include "ptx.cuh"

// -DMB_PAIR_SPIN=0: the single-thread roles (producer, issuer) park in try_wait instead of polling with test_wait
#ifndef MB_PAIR_SPIN
#define MB_PAIR_SPIN 1
#endif

namespace mb200 {

__device__ __forceinline__ void pair_wait_role(uint32_t bar, uint32_t parity) {
  if (MB_PAIR_SPIN) mbar_wait_spin(bar, parity); else mbar_wait(bar, parity);
}

constexpr bool kPairTiming = MB_PAIR_TIMING != 0;

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
  float* rowsum_cols;        // symmetric pass: [cols_pad] row sums of ALL column pixels (== rowsum on one GPU)
  int rb_off;                // symmetric pass: global tile index of this rank's row tile 0
  int own_g, own_G;          // symmetric pass: this rank and the number of ranks (unit ownership, sym_owned)
  int share_g, share_G;      // symmetric pass: the launch covers share share_g of share_G equal parts of the unit list (0 = 1)
  float* part;               // [PAIR_SLOTS][rows_pad][D]  MODE_GRAD: partial sum_k T_qk c_k
  float scale_log2;          // log2(e) / tau
  float mask_shift;          // extra exponent shift of a masked column in the MASK_BIAS form (see pair_mask_shift)
  int n_row_blocks;          // row blocks of PAIR_RB rows
  int n_col_tiles;           // column tiles of 128
  int rows_pad;
  unsigned long long* dbg;   // optional [gridDim.x][16] cycle counters (MUSCLE_PAIR_TIMING), else null
  int dbg_flags;             // experiments (MUSCLE_PAIR_FLAGS): 1 = column tiles loaded once per stage only, 2 = no S MMAs
};

constexpr int PAIR_SLOTS = 8;             // a row block is shared by at most this many CTAs
constexpr uint32_t ITEM_END = 0xffffffffu;

// live units per CTA: an even share, but never so small that a row block spans > PAIR_SLOTS CTAs
__host__ __device__ inline uint32_t pair_units_per_cta(uint32_t total_units, int n_ctas, int n_col_tiles) {
  const uint32_t even = (total_units + n_ctas - 1) / n_ctas;
  const uint32_t floor_ = (n_col_tiles + PAIR_SLOTS - 2) / (PAIR_SLOTS - 1);
  return even > floor_ ? even : floor_;
}

constexpr int BIAS_SLOTS = 4;            // distinct row classes per tile the exponent-bias mask handles
constexpr int PAIR_RB = 128;              // rows per row block of the pair passes (= one row tile)

template <int D, int MODE>
struct PairCfg {
  static_assert(D == 64 || D == 128 || D == 256, "projection dim must be 64, 128 or 256");
  // epilogue groups of 4 warps: the row-sum pass's epilogue fits 128 registers, so a third group (4
  // warps per scheduler instead of 3) hides more of its MUFU / TMEM-load latency; the gradient pass
  // needs ~156 registers and stays at two (and D = 256 has only two ring stages: see the assert below)
  static constexpr int NG = (MODE == MODE_ROWSUM && D != 256) ? 3 : 2;
  static constexpr int NB = (MODE == MODE_ROWSUM) ? 4 : (D == 256 ? 2 : 3);   // S buffers in TMEM
  static constexpr int NST = (D == 256) ? 2 : (D == 128 ? 5 : 8);             // column-tile ring
  static constexpr int TILE_BYTES = 128 * D * 2;         // one 128-pixel bf16 tile
  static constexpr int PANEL_BYTES = 128 * 128;          // one [128 x 64] swizzle panel
  static constexpr int NPANEL = D / 64;
  static constexpr int SMEM_Q = 0;
  static constexpr int SMEM_K = SMEM_Q + TILE_BYTES;
  static constexpr int SMEM_BITS = SMEM_K + NST * TILE_BYTES;
  static constexpr int SMEM_W = SMEM_BITS + NST * 512;
  static constexpr int SMEM_CB = SMEM_W + NST * 512;     // exponent-shift vectors [NG groups][2][BIAS_SLOTS][128] (MASK_BIAS)
  static constexpr int SMEM_BAR = SMEM_CB + 2 * NG * BIAS_SLOTS * 512;
  static constexpr int NBAR = 2 * NST + 2 + 2 * NB + 2;
  static constexpr int SMEM_INFO = SMEM_BAR + NBAR * 8;  // item headers [2][4], stage info [NST]
  static constexpr int SMEM_TMEMPTR = SMEM_INFO + (8 + NST) * 4;
  static constexpr int SMEM_BYTES = SMEM_TMEMPTR + 16 + 1024;  // +1024: manual 1024-B alignment
  static constexpr int ACC_COL = NB * 128;
  static constexpr int WARP_PRODUCER = 4 * NG;
  static constexpr int WARP_ISSUER = 4 * NG + 1;
  static constexpr int THREADS = (4 * NG + 2) * 32;
  static_assert(ACC_COL + (MODE == MODE_GRAD ? D : 0) <= 512, "TMEM budget");
  // A group waits on a stage's / buffer's mbarrier by phase parity.  Its previous unit is NG units
  // back; the previous user of the same stage (buffer) is NST (NB) units back.  Only if NG <= NST and
  // NG <= NB has the group itself seen that earlier phase complete -- otherwise a wait for phase n+2
  // can be satisfied by phase n (same parity) and the group runs ahead of the data.
  static_assert(NG <= NST && NG <= NB, "epilogue groups must not outnumber ring stages or S buffers");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
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
__device__ __forceinline__ void pair_plan_body(const uint32_t* tor_row, int n_row_blocks,
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

// One block of 1024 threads.  (The single-GPU composite runs the same body at the end of
// class_scan_kernel, which produces `tor`, instead of launching this.)
template <int QT>
__global__ void __launch_bounds__(1024) pair_plan_kernel(const uint32_t* tor_row, int n_row_blocks,
                                                         const uint32_t* tor_col, int n_col_tiles,
                                                         uint32_t* unit_off) {
  pdl_trigger();
  pdl_wait();
  pair_plan_body<QT>(tor_row, n_row_blocks, tor_col, n_col_tiles, unit_off);
}

// ------------------------------------------------------------------------------------------
// Epilogue math of one 64-column half for one thread (= one row); v = the S values.
// MASK_NONE: no label mask needed.  MASK_SELECT: per-element select on the class bits.  MASK_BIAS:
// the rows of the tile have at most BIAS_SLOTS distinct classes, so the mask is one vector per row
// class and column tile, folded into the exponent shift: bias_sm[k] = -a for a kept column,
// -a - mask_shift for a masked one (pair_mask_shift: the leak stays below fp32 resolution of any row sum,
// at no instruction cost); every thread reads the vector of its row's class.  POLY of every 4 element pairs take the FMA-pipe exponential.
enum { MASK_NONE = 0, MASK_SELECT = 1, MASK_BIAS = 2 };
// A masked column keeps 2^-shift of its term.  The smallest term a kept column can have is 2^(-2a)
// (s = -1), so the shift is at least 2a + 26: a masked column then leaks < 2^-26 of the smallest possible
// negative whatever the temperature (a fixed 64 leaked as much as a real negative for tau < 0.027).
// With the FMA-pipe polynomial in use the exponent must stay above -120 (ex2_poly2), which bounds
// a; without it the shift is 200 and ex2.approx.ftz returns exactly 0 for a masked column.
__host__ __device__ inline float pair_mask_shift(float a, bool poly) {
  if (!poly) return 200.f;
  const float s = 2.f * a + 26.f;
  return s > 64.f ? s : 64.f;
}
__host__ __device__ inline bool pair_poly_ok(float a) { return 2.f * a + pair_mask_shift(a, true) < 118.f; }
template <int MODE, int POLY, int MASK>
__device__ __forceinline__ void pair_half_math(uint32_t (&v)[2][32], uint32_t t_addr,
                                               const uint8_t* bits_sm, const uint8_t* w_sm,
                                               const uint8_t* bias_sm,
                                               uint32_t qmask, float a, float wrow,
                                               uint64_t (&racc)[4]) {
  const uint64_t a2 = pack2(a, a);
  const uint64_t noff2 = pack2(-a, -a);  // the shift: exp((s - 1)/tau), |s| <= 1 for unit vectors
  const uint64_t wr2 = pack2(wrow, wrow);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint4* kb4 = reinterpret_cast<const uint4*>(bits_sm) + c * 8;
    const float4* kw4 = reinterpret_cast<const float4*>(w_sm) + c * 8;
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
      if (MASK == MASK_SELECT) {
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
__global__ void __launch_bounds__((PairCfg<D, MODE>::THREADS), 1)
pair_kernel(const __grid_constant__ CUtensorMap tm_row, const __grid_constant__ CUtensorMap tm_col,
            const PairParams p) {
  using C = PairCfg<D, MODE>;
  constexpr int NST = C::NST;
  constexpr int NB = C::NB;
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
  auto s_full = [&](int b) { return bar0 + 8u * (2 * NST + 2 + b); };
  auto t_full = [&](int b) { return bar0 + 8u * (2 * NST + 2 + NB + b); };  // GRAD: T stored; ROWSUM: S consumed
  const uint32_t acc_full = bar0 + 8u * (2 * NST + 2 + 2 * NB);
  const uint32_t acc_empty = bar0 + 8u * (2 * NST + 3 + 2 * NB);
  volatile uint32_t* info = reinterpret_cast<volatile uint32_t*>(smem_gen + C::SMEM_INFO);
  // info[4*(it&1) + {0,1,2}]: live tiles, row block, slot of item `it`; info[8+st]: tor of the stage's column tile
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(smem_gen + C::SMEM_TMEMPTR);
  __shared__ uint2 ulist_sm[32];           // producer warp: (class union, index) of the next column tiles to load

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (kPairTiming && p.dbg && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.dbg[blockIdx.x * 16 + 14] = t;
  }

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(k_full(s), 1);
      mbar_init(k_empty(s), 1 + 4);   // MMA issuer + the four warps of the group that owns the unit
    }
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1 + 4 * C::NG);  // MMA issuer + every epilogue warp (after it has read the item header)
    for (int b = 0; b < NB; ++b) {
      mbar_init(s_full(b), 1);
      mbar_init(t_full(b), 4);
    }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 8);
    fence_barrier_init();
  }
  if (warp == C::WARP_ISSUER) {
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
  pdl_wait();   // barriers, TMEM and descriptors are set up under the previous kernel's tail

  if (warp == C::WARP_PRODUCER) {
    // ------------------------------------------------------------ TMA producer + work partition
    // The whole warp walks the work list (32 column-tile class unions per coalesced load, the next
    // 32 prefetched: a dependent L2 load per tile on one lane made this warp the bottleneck);
    // lane 0 alone touches barriers and issues the copies.  (A second producer warp taking
    // alternate units changed nothing: what remains of this warp's time per unit is the copy
    // engine itself, ~32 KB per unit against shared-memory bandwidth the S MMAs saturate.)
    {
      const uint32_t total = __ldg(p.unit_off + p.n_row_blocks);
      const uint32_t per = pair_units_per_cta(total, gridDim.x, p.n_col_tiles);
      const uint64_t lo64 = static_cast<uint64_t>(per) * blockIdx.x;
      uint32_t u = lo64 < total ? static_cast<uint32_t>(lo64) : total;
      const uint32_t u1 = (lo64 + per < total) ? static_cast<uint32_t>(lo64 + per) : total;
      // largest rb with unit_off[rb] <= u: a 32-ary search, one coalesced-ish probe per lane and
      // round (2 rounds up to 1024 row blocks; the 8+ dependent loads of a binary search were
      // ~5 us of every launch, cold)
      int rb = 0;
      {
        int lo = 0, hi = p.n_row_blocks;  // unit_off[lo] <= u, unit_off[hi] > u (when u < total)
        while (hi - lo > 1) {
          const int span = hi - lo;
          const int step = (span + 31) / 32;                  // probes at lo + step*(lane+1), clipped below hi
          const int probe = lo + step * (lane + 1);
          const bool le = (probe < hi) && (__ldg(p.unit_off + probe) <= u);
          const uint32_t m = __ballot_sync(0xffffffffu, le);  // monotone: a prefix of the lanes
          const int k = __popc(m);                            // probes that are <= u
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
        if (U1 <= u) { ++rb; continue; }   // row blocks without live units
        const uint32_t a = u - U0;
        const uint32_t e = (u1 < U1 ? u1 : U1) - U0;
        const uint32_t tor_r = __ldg(p.tor_row + rb);
        if (lane == 0) {
          mbar_wait(q_empty, (it & 1) ^ 1);
          info[4 * (it & 1) + 0] = e - a;
          info[4 * (it & 1) + 1] = static_cast<uint32_t>(rb);
          info[4 * (it & 1) + 2] = blockIdx.x - U0 / per;
          mbar_arrive_expect_tx(q_full, C::TILE_BYTES);
          tma_load_3d(q_smem, &tm_row, q_full, 0, rb * 128, 0);
        }
        uint32_t li = 0;  // live index of the next live column tile within the row block
        uint32_t tc_next = (lane < p.n_col_tiles) ? __ldg(p.tor_col + lane) : 0u;
        for (int j0 = 0; j0 < p.n_col_tiles && li < e; j0 += 32) {
          const uint32_t tc = tc_next;
          const int jn = j0 + 32 + lane;
          tc_next = (jn < p.n_col_tiles) ? __ldg(p.tor_col + jn) : 0u;
          uint32_t live = __ballot_sync(0xffffffffu, (j0 + lane < p.n_col_tiles) && !tile_pair_dead(tor_r, tc));
          // skip whole words that end before the share starts
          const uint32_t cnt = __popc(live);
          if (li + cnt <= a) { li += cnt; continue; }
          // The lanes whose tile is one of the CTA's units write it into a 32-entry list (one step for up to 32
          // units); lane 0 then only waits for a free stage and issues the copies.  (Walking the ballot word bit
          // by bit on the issuing path cost ~900 cycles of this warp per unit: it shares its scheduler with busy
          // epilogue warps, and the issuer waited ~450 cycles per unit for column tiles.)
          const uint32_t rank = __popc(live & ((1u << lane) - 1u));
          const uint32_t idx = li + rank;                         // live index of this lane's tile in the row block
          const uint32_t first = li > a ? li : a;
          const uint32_t last = (li + cnt < e) ? li + cnt : e;    // units [first, last) of the row block come from this word
          if (((live >> lane) & 1u) && idx >= first && idx < last)
            ulist_sm[idx - first] = make_uint2(tc, static_cast<uint32_t>(j0 + lane));
          __syncwarp();
          if (lane == 0) {
            for (uint32_t t = 0; t < last - first; ++t) {
              const uint2 ent = ulist_sm[t];
              const uint32_t g = gcnt + t;
              const int st = static_cast<int>(g % NST);
              const int j = static_cast<int>(ent.y);
              { const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
                pair_wait_role(k_empty(st), ((g / NST) & 1) ^ 1);
                if (kPairTiming && p.dbg) p_wait += clock64() - w0; }
              info[8 + st] = ent.x;
              if ((kPairTiming && (p.dbg_flags & 1)) && g >= static_cast<uint32_t>(NST)) {
                mbar_arrive(k_full(st));
              } else {
                mbar_arrive_expect_tx(k_full(st), C::TILE_BYTES + 512 + (MODE == MODE_GRAD ? 512 : 0));
                tma_load_3d(k_smem + st * C::TILE_BYTES, &tm_col, k_full(st), 0, j * 128, 0);
                bulk_load_1d(bits_smem + st * 512, p.bits_col + static_cast<size_t>(j) * 128, 512,
                             k_full(st));
                if (MODE == MODE_GRAD)
                  bulk_load_1d(w_smem + st * 512, p.w_col + static_cast<size_t>(j) * 128, 512,
                               k_full(st));
              }
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
        if (kPairTiming && p.dbg) { p.dbg[blockIdx.x * 16 + 8] = p_wait; p.dbg[blockIdx.x * 16 + 9] = clock64() - p_t0; }
        mbar_wait(q_empty, (it & 1) ^ 1);
        info[4 * (it & 1) + 0] = ITEM_END;
        mbar_arrive(q_full);
      }
    }
  } else if (warp == C::WARP_ISSUER) {
    // ------------------------------------------------------------ MMA issuer
    // (elect.sync, not lane == 0: ptxas then knows a single thread runs the region and feeds the
    // uniform-register operands of UTCHMMA without a per-instruction waterfall loop)
    if (elect_one()) {
      constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t IDESC_G = umma_idesc_bf16(128, D, 0, 1);
      uint32_t gbase = 0;
      long long i_kfull = 0, i_tfull = 0, i_t0 = (kPairTiming && p.dbg) ? clock64() : 0;
      for (uint32_t it = 0;; ++it) {
        mbar_wait(q_full, it & 1);
        tc_fence_after();
        const uint32_t n_u = info[4 * (it & 1) + 0];
        if (n_u == ITEM_END) break;
        const int n = static_cast<int>(n_u);

        // S of the item's unit k into buffer (gbase + k) % NB
        auto issue_S = [&](int k) {
          const uint32_t gk = gbase + k;
          const int st = static_cast<int>(gk % NST);
          const int b = static_cast<int>(gk % NB);
          { const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
            pair_wait_role(k_full(st), (gk / NST) & 1);
            if (kPairTiming && p.dbg) i_kfull += clock64() - w0; }
          if (MODE == MODE_ROWSUM) {   // the buffer's previous unit has been read out
            const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
            pair_wait_role(t_full(b), ((gk / NB) & 1) ^ 1);
            if (kPairTiming && p.dbg) i_tfull += clock64() - w0;
          }   // MODE_GRAD: the G MMA of that unit was issued earlier and the pipe runs in order
          tc_fence_after();
          const uint32_t b_base = k_smem + st * C::TILE_BYTES;
          if (!(kPairTiming && (p.dbg_flags & 2))) {
#pragma unroll
            for (int kk = 0; kk < D / 16; ++kk) {
              const uint32_t off = (kk >> 2) * C::PANEL_BYTES + (kk & 3) * 32;
              mma_ss(tmem + b * 128, umma_desc_sw128(q_smem + off, 16, 1024),
                     umma_desc_sw128(b_base + off, 16, 1024), IDESC_S, kk > 0);
            }
          }
          tc_commit(s_full(b));
          if (MODE == MODE_ROWSUM) tc_commit(k_empty(st));   // the stage's last reader on the tensor pipe
          if (k == n - 1) tc_commit(q_empty);                // the last S MMAs of the item are in flight
        };

        if (MODE == MODE_ROWSUM) {
          for (int k = 0; k < n; ++k) issue_S(k);
        } else {
          const int pro = n < NB ? n : NB;
          for (int k = 0; k < pro; ++k) issue_S(k);
          pair_wait_role(acc_empty, (it & 1) ^ 1);   // the previous item's G has been read out
          tc_fence_after();
          for (int k = 0; k < n; ++k) {
            const uint32_t gk = gbase + k;
            const int st = static_cast<int>(gk % NST);
            const int b = static_cast<int>(gk % NB);
            { const long long w0 = (kPairTiming && p.dbg) ? clock64() : 0;
              pair_wait_role(t_full(b), (gk / NB) & 1);
              if (kPairTiming && p.dbg) i_tfull += clock64() - w0; }
            tc_fence_after();
            const uint32_t b_base = k_smem + st * C::TILE_BYTES;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk)
              mma_ts(tmem + C::ACC_COL, tmem + b * 128 + kk * 8,
                     umma_desc_sw128(b_base + kk * 2048, C::PANEL_BYTES, 1024), IDESC_G,
                     (k == 0 && kk == 0) ? 0u : 1u);
            tc_commit(k_empty(st));
            if (k + NB < n) issue_S(k + NB);
          }
          tc_commit(acc_full);
        }
        gbase += n;
      }
      if (kPairTiming && p.dbg) { p.dbg[blockIdx.x * 16 + 10] = i_kfull; p.dbg[blockIdx.x * 16 + 11] = i_tfull; p.dbg[blockIdx.x * 16 + 12] = clock64() - i_t0; }
    }
  } else {
    // ------------------------------------------------------------ epilogue groups
    const int g = warp >> 2;
    const int wq = warp & 3;
    const uint32_t lane_addr = static_cast<uint32_t>(wq * 32) << 16;
    const float a = p.scale_log2;
    uint32_t gbase = 0, n_bias = 0;
    const bool timing = kPairTiming && (p.dbg != nullptr) && warp == 0 && lane == 0;
    long long t_kfull = 0, t_sfull = 0, t_ld = 0, t_math = 0, t_q = 0, t_tail = 0, n_mine = 0, n_sel = 0, n_bi = 0, t_mark = timing ? clock64() : 0;
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
      const int row = rb * 128 + wq * 32 + lane;
      const uint32_t rbits = __ldg(p.bits_row + row);
      const uint32_t tor_r = __ldg(p.tor_row + rb);
      const uint32_t qmask = rbits ? ~rbits : 0u;
      // few enough distinct row classes for the bias form; my_slot = rank of this row's class among
      // them (rows outside the loss read slot 0: their sums are never used)
      const bool bias_ok = __popc(tor_r & TOR_CLASSES) <= BIAS_SLOTS;
      const int my_slot = rbits ? __popc(tor_r & TOR_CLASSES & (rbits - 1u)) : 0;
      float wrow = 0.f;
      if (MODE == MODE_GRAD) wrow = __ldg(p.w_row + row);
      uint64_t racc[4] = {0ull, 0ull, 0ull, 0ull};
      // this group's units: every NG-th one of the CTA's running unit count
      for (int k = static_cast<int>((static_cast<uint32_t>(g) + C::NG - gbase % C::NG) % C::NG); k < n; k += C::NG) {
        const uint32_t gk = gbase + k;
        const int st = static_cast<int>(gk % NST);
        const int b = static_cast<int>(gk % NB);
        mbar_wait(k_full(st), (gk / NST) & 1);
        lap(t_kfull);
        const uint32_t tor_c = info[8 + st];
        const bool full = tile_pair_full(tor_r, tor_c);
        const int mask = full ? MASK_NONE : (bias_ok ? MASK_BIAS : MASK_SELECT);
        const uint8_t* cb = smem_gen + C::SMEM_CB + ((g * 2 + (n_bias & 1)) * BIAS_SLOTS) * 512;
        if (mask == MASK_BIAS) {   // one shift vector per row class of the tile
          const int t = wq * 32 + lane;
          const uint32_t kb = reinterpret_cast<const uint32_t*>(smem_gen + C::SMEM_BITS + st * 512)[t];
          float* cbw = reinterpret_cast<float*>(const_cast<uint8_t*>(cb)) + t;
          uint32_t rem = tor_r & TOR_CLASSES;
          for (int j = 0; rem != 0u; ++j, rem &= rem - 1u)
            cbw[j * 128] = (kb & ~(rem & (0u - rem))) ? -a : -a - p.mask_shift;
          ++n_bias;   // double buffered: the barrier of the next masked tile orders reuse
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        }
        const bool early = (MODE == MODE_ROWSUM) && mask != MASK_SELECT;
        if (early) {   // no class bits left to read: the stage is only the tensor pipe's now
          __syncwarp();
          if (lane == 0) mbar_arrive(k_empty(st));
        }
        mbar_wait(s_full(b), (gk / NB) & 1);
        lap(t_sfull);
        tc_fence_after();
        if (timing) { ++n_mine; n_sel += (mask == MASK_SELECT); n_bi += (mask == MASK_BIAS); }
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          const uint32_t s_addr = tmem + lane_addr + b * 128 + h * 64;
          const uint32_t t_addr = tmem + lane_addr + b * 128 + h * 32;
          const uint8_t* bsm = smem_gen + C::SMEM_BITS + st * 512 + h * 256;
          const uint8_t* wsm = smem_gen + C::SMEM_W + st * 512 + h * 256;
          uint32_t v[2][32];
          tmem_ld32(s_addr, v[0]);
          tmem_ld32(s_addr + 32, v[1]);
          tmem_wait_ld();
          lap(t_ld);
          if (MODE == MODE_ROWSUM && h == 1) {  // S is in registers: hand the buffer back before the math
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(t_full(b));
          }
          if (mask == MASK_NONE)
            pair_half_math<MODE, POLY, MASK_NONE>(v, t_addr, bsm, wsm, cb, qmask, a, wrow, racc);
          else if (mask == MASK_BIAS)
            pair_half_math<MODE, POLY, MASK_BIAS>(v, t_addr, bsm, wsm, cb + my_slot * 512 + h * 256, qmask, a, wrow, racc);
          else
            pair_half_math<MODE, POLY, MASK_SELECT>(v, t_addr, bsm, wsm, cb, qmask, a, wrow, racc);
          lap(t_math);
        }
        if (MODE == MODE_GRAD) {
          tmem_wait_st();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t_full(b));
        }
        if (!early) {
          __syncwarp();
          if (lane == 0) mbar_arrive(k_empty(st));
        }
      }
      if (MODE == MODE_ROWSUM) {
        float r0, r1, r2, r3;
        unpack2(add2(racc[0], racc[1]), r0, r1);
        unpack2(add2(racc[2], racc[3]), r2, r3);
        const float rsum = (r0 + r1) + (r2 + r3);
        if (rbits && rsum != 0.f) atomicAdd(p.rowsum + row, rsum);
      } else {
        constexpr int NCH = D / 64;          // 32-column chunks of this group's half of G
        const int col0 = g * (D / 2);
        float* dst = p.part + (static_cast<size_t>(slot) * p.rows_pad + row) * D + col0;
        mbar_wait(acc_full, it & 1);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_addr + C::ACC_COL + col0 + c * 32, v);
          tmem_wait_ld();
#pragma unroll
          for (int q4 = 0; q4 < 8; ++q4)
            reinterpret_cast<float4*>(dst + c * 32)[q4] =
                make_float4(__uint_as_float(v[q4 * 4]), __uint_as_float(v[q4 * 4 + 1]),
                            __uint_as_float(v[q4 * 4 + 2]), __uint_as_float(v[q4 * 4 + 3]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty);
      }
      lap(t_tail);
      gbase += n;
    }
    if (timing) {
      unsigned long long* d = p.dbg + blockIdx.x * 16;
      d[0] = t_q; d[1] = t_kfull; d[2] = t_sfull; d[3] = t_ld; d[4] = t_math; d[5] = t_tail; d[6] = n_mine; d[7] = n_sel; d[13] = n_bi;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == C::WARP_ISSUER) tmem_dealloc(tmem, 512);
  if (kPairTiming && p.dbg && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.dbg[blockIdx.x * 16 + 15] = t;
  }
}

}  // namespace mb200
