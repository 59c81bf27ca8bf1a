// HBM-bound per-pixel kernels around the tcgen05 pair passes of the dense InfoNCE.
//
//   class_hist / class_scan : counting sort of the pixels by class (stable, deterministic):
//                             per-chunk histograms, then per-class offsets, segment starts and the
//                             per-tile class unions the pair kernel uses to skip masked tiles
//   prepare   : L2-normalise, cast to bf16, write each pixel's row at its class-sorted position
//   class_sum : per-class column sums of the sorted bf16 rows (segmented block reduction)
//   mid       : pixel -> class-prototype logit, 1/rowsum, per-pixel loss term, prototype gradient
//   final     : sum the column splits, prototype terms, normalisation backward, un-sort, layout
//
// Layouts: source order n = b*P + p (the caller's [B,D,h,w] or [N,D]); sorted order r in
// [0, rows_pad): class 0 pixels first, ..., invalid pixels last, pad rows (zero) at the end.
#pragma once
#include <cuda_bf16.h>

#include "ptx.cuh"
#include "pair_kernel.cuh"
#include "pair_sym_kernel.cuh"

namespace mb200 {

constexpr int SORT_CHUNK = 128;   // source pixels per block in hist / prepare / final (4 warps)
constexpr int SORT_WARPS = SORT_CHUNK / 32;
constexpr int MAX_BINS = 32;      // classes (<= 31) + one bin for invalid pixels

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ int class_of_bits(uint32_t b) { return 31 - __clz(b); }

__device__ __forceinline__ int load_class(const void* cls, int cls64, long long n, int C) {
  const long long c = cls64 ? reinterpret_cast<const long long*>(cls)[n]
                            : static_cast<long long>(reinterpret_cast<const int*>(cls)[n]);
  return (c >= 0 && c < C) ? static_cast<int>(c) : C;  // bin C = invalid
}

// ------------------------------------------------------------------------------------------
// hist[blk][bin] = number of pixels of class `bin` among source pixels [blk*128, blk*128+128)
// Also zeroes the pad rows [n, rows_pad) of ehat / bits / inv_norm (zero vectors that belong to no
// class), spread over the grid: three memsets less on the launch chain.
__global__ void __launch_bounds__(SORT_CHUNK) class_hist_kernel(const void* cls, int cls64,
                                                                long long n, int C,
                                                                uint32_t* hist, long long rows_pad, int dim,
                                                                uint32_t* ehat_words, uint32_t* bits,
                                                                float* inv_norm) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint32_t h[MAX_BINS];
  if (threadIdx.x < MAX_BINS) h[threadIdx.x] = 0u;
  __syncthreads();
  const long long i = static_cast<long long>(blockIdx.x) * SORT_CHUNK + threadIdx.x;
  if (i < n) atomicAdd(&h[load_class(cls, cls64, i, C)], 1u);
  __syncthreads();
  if (threadIdx.x < MAX_BINS) hist[static_cast<size_t>(blockIdx.x) * MAX_BINS + threadIdx.x] = h[threadIdx.x];
  // pad rows: (rows_pad - n) * (dim/2 + 2) words in all
  const long long wpr = dim / 2;
  const long long pad_rows = rows_pad - n;
  const long long total = pad_rows * (wpr + 2);
  for (long long k = static_cast<long long>(blockIdx.x) * SORT_CHUNK + threadIdx.x; k < total;
       k += static_cast<long long>(gridDim.x) * SORT_CHUNK) {
    if (k < pad_rows * wpr) ehat_words[n * wpr + k] = 0u;
    else if (k < pad_rows * (wpr + 1)) bits[n + (k - pad_rows * wpr)] = 0u;
    else inv_norm[n + (k - pad_rows * (wpr + 1))] = 0.f;
  }
}

// One block of 32 warps; warp c owns bin c.  In place: hist[blk][c] becomes the sorted position
// of the first pixel of class c in chunk blk.  seg[c] = first sorted row of class c
// (seg[C] = first invalid row, seg[C+1] = n).  class_mass[c] += count.  tor[t] = class union of
// sorted tile t (rows 128t..128t+127) | TOR_INVALID when it holds invalid or pad rows.
__global__ void __launch_bounds__(1024) class_scan_kernel(uint32_t* hist, int n_blk, int C,
                                                          long long n, long long rows_pad,
                                                          uint32_t* seg, float* class_mass,
                                                          uint32_t* tor, uint32_t* unit_off,
                                                          uint32_t* unit_off_sym) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint32_t total[MAX_BINS + 1];
  __shared__ uint32_t start[MAX_BINS + 2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Lane l of warp c owns the contiguous run of chunks [l*per, l*per + per) of bin c: one round of
  // independent loads, serial adds in registers, ONE warp scan -- instead of a 5-step shuffle scan
  // per 32 chunks.
  const int per = (n_blk + 31) / 32;
  const int b0 = lane * per;
  // (the run's counts stay in registers for the rewrite below when they fit: the second pass over hist[] was `per`
  //  dependent L2 round trips, the loads not being movable across the stores to the same array)
  constexpr int SCAN_KEEP = 16;
  uint32_t kept[SCAN_KEEP];
  const bool keep = per <= SCAN_KEEP;
  uint32_t mine = 0;
  if (warp <= C) {
    if (keep) {
#pragma unroll
      for (int i = 0; i < SCAN_KEEP; ++i) {
        kept[i] = (i < per && b0 + i < n_blk) ? hist[static_cast<size_t>(b0 + i) * MAX_BINS + warp] : 0u;
        mine += kept[i];
      }
    } else {
      for (int i = 0; i < per; ++i)
        if (b0 + i < n_blk) mine += hist[static_cast<size_t>(b0 + i) * MAX_BINS + warp];
    }
  }
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += u;
  }
  if (lane == 31) total[warp] = (warp <= C) ? inc : 0u;
  __syncthreads();
  if (warp == 0) {   // exclusive prefix over the bins (lane = bin)
    const uint32_t v = (lane <= C) ? total[lane] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += u;
    }
    if (lane <= C) start[lane] = x - v;
    if (lane == C) start[C + 1] = x;  // == n
  }
  __syncthreads();
  if (threadIdx.x <= C + 1) seg[threadIdx.x] = start[threadIdx.x];
  if (threadIdx.x < C) atomicAdd(class_mass + threadIdx.x, static_cast<float>(total[threadIdx.x]));
  // in place: hist[blk][c] -> sorted position of the first pixel of class c in chunk blk
  if (warp <= C) {
    uint32_t run = start[warp] + inc - mine;
    if (keep) {
#pragma unroll
      for (int i = 0; i < SCAN_KEEP; ++i)
        if (i < per && b0 + i < n_blk) {
          hist[static_cast<size_t>(b0 + i) * MAX_BINS + warp] = run;
          run += kept[i];
        }
    } else {
      for (int i = 0; i < per; ++i)
        if (b0 + i < n_blk) {
          const size_t idx = static_cast<size_t>(b0 + i) * MAX_BINS + warp;
          const uint32_t v = hist[idx];
          hist[idx] = run;
          run += v;
        }
    }
  }
  // tile class unions (also kept in shared memory for the numbering below when they fit: the plan bodies read every
  // tile's set twice)
  constexpr int SCAN_TOR_SM = 4096;
  __shared__ uint32_t tor_sm[SCAN_TOR_SM];
  const long long n_tiles = rows_pad / 128;
  const bool tor_in_sm = n_tiles <= SCAN_TOR_SM;
  for (long long tile = threadIdx.x; tile < n_tiles; tile += blockDim.x) {
    const long long r0 = tile * 128, r1 = r0 + 128;
    uint32_t u = 0;
    for (int c = 0; c < C; ++c)
      if (static_cast<long long>(start[c]) < r1 && static_cast<long long>(start[c + 1]) > r0) u |= 1u << c;
    if (r1 > static_cast<long long>(start[C])) u |= 0x80000000u;
    tor[tile] = u;
    if (tor_in_sm) tor_sm[tile] = u;
  }
  const uint32_t* tor_rd = tor_in_sm ? tor_sm : tor;
  // rows and columns are the same pixels (single GPU): number the live (row tile, column tile)
  // units here instead of in a launch of their own
  if (unit_off != nullptr) {
    __syncthreads();
    pair_plan_body<1>(tor_rd, static_cast<int>(n_tiles), tor_rd, static_cast<int>(n_tiles), unit_off);
  }
  if (unit_off_sym != nullptr) {   // and the units with column tile >= row tile for the symmetric row-sum pass
    __syncthreads();
    pair_plan_sym_body(tor_rd, static_cast<int>(n_tiles), unit_off_sym);
  }
}

// ------------------------------------------------------------------------------------------
struct PrepareParams {
  const float* x;
  const void* cls;
  int cls64;
  int layout;
  long long n;       // pixels
  long long ppi;     // pixels per image (NCHW)
  int C;
  float eps;
  const uint32_t* chunk_off;  // [n_blk][MAX_BINS]  scanned histogram (class_scan_kernel)
  __nv_bfloat16* ehat;        // [rows_pad][D]  sorted
  uint32_t* bits;             // [rows_pad]     sorted
  float* inv_norm;            // [rows_pad]     sorted
  int* perm;                  // [n]            source pixel -> sorted row
  float* class_sum;           // [C][D] zeroed, or null: prepare_nchw_kernel adds its block's per-class sums of the
                              // bf16-rounded rows (the job of class_sum_kernel, without the extra launch and pass over ehat)
};

// One block = 128 consecutive source pixels, one warp = 32 of them.  The sorted position of a
// pixel is  chunk_off[blk][class] + (pixels of that class earlier in the chunk), which keeps the
// source order inside every class (stable) and needs no atomics.
template <int D>
__global__ void __launch_bounds__(SORT_CHUNK) prepare_kernel(const PrepareParams p) {
  pdl_trigger();
  pdl_wait();
  constexpr int WPR = D / 64;          // 32-bit words (bf16 pairs) per lane per row
  constexpr int ROW_W = D / 2 + 1;     // smem row stride in words (odd: conflict-free columns)
  extern __shared__ uint32_t smem_u[];  // NCHW only: [SORT_WARPS][32][ROW_W]
  __shared__ uint32_t warp_hist[SORT_WARPS][MAX_BINS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < SORT_WARPS * MAX_BINS; i += SORT_CHUNK) (&warp_hist[0][0])[i] = 0u;
  __syncthreads();
  const long long n0 = static_cast<long long>(blockIdx.x) * SORT_CHUNK + warp * 32;
  const long long i = n0 + lane;
  const bool live = i < p.n;
  const int c = live ? load_class(p.cls, p.cls64, i, p.C) : (MAX_BINS + lane);  // dead lanes: unique keys
  const uint32_t peers = __match_any_sync(0xffffffffu, c);
  const int rank = __popc(peers & ((1u << lane) - 1u));
  if (live && rank == 0) warp_hist[warp][c] = __popc(peers);
  __syncthreads();
  int dest = 0;
  if (live) {
    uint32_t off = p.chunk_off[static_cast<size_t>(blockIdx.x) * MAX_BINS + c];
    for (int w = 0; w < warp; ++w) off += warp_hist[w][c];
    dest = static_cast<int>(off) + rank;
    p.perm[i] = dest;
    p.bits[dest] = (c < p.C) ? (1u << c) : 0u;
  }
  if (n0 >= p.n) return;

  if (p.layout == MUSCLE_LAYOUT_NCHW) {
    // lane = pixel: coalesced across the warp for every channel
    const long long b = live ? i / p.ppi : 0;
    const float* src = p.x + (b * D) * p.ppi + (live ? i - b * p.ppi : 0);
    float ss = 0.f;
    if (live) {
#pragma unroll 8
      for (int d = 0; d < D; ++d) {
        const float v = __ldg(src + static_cast<long long>(d) * p.ppi);
        ss = fmaf(v, v, ss);
      }
    }
    const float inv = 1.f / fmaxf(sqrtf(ss), p.eps);
    uint32_t* tile = smem_u + (warp * 32 + lane) * ROW_W;
    if (live) {
      p.inv_norm[dest] = inv;
#pragma unroll 8
      for (int d = 0; d < D; d += 2) {
        const float v0 = __ldg(src + static_cast<long long>(d) * p.ppi) * inv;
        const float v1 = __ldg(src + static_cast<long long>(d + 1) * p.ppi) * inv;
        tile[d >> 1] = pack_bf16x2(v0, v1);
      }
    }
    __syncwarp();
    const int n_live = static_cast<int>(min(32ll, p.n - n0));
    for (int r = 0; r < n_live; ++r) {
      const int dr = __shfl_sync(0xffffffffu, dest, r);
      const uint32_t* trow = smem_u + (warp * 32 + r) * ROW_W;
      uint32_t* out = reinterpret_cast<uint32_t*>(p.ehat + static_cast<size_t>(dr) * D);
#pragma unroll
      for (int k = 0; k < WPR; ++k) out[lane + 32 * k] = trow[lane + 32 * k];
    }
  } else {
    const int n_live = static_cast<int>(min(32ll, p.n - n0));
    for (int r = 0; r < n_live; ++r) {
      const int dr = __shfl_sync(0xffffffffu, dest, r);
      const float2* src = reinterpret_cast<const float2*>(p.x + (n0 + r) * D);
      float2 v[WPR];
      float ss = 0.f;
#pragma unroll
      for (int k = 0; k < WPR; ++k) {
        v[k] = __ldg(src + lane + 32 * k);
        ss += v[k].x * v[k].x + v[k].y * v[k].y;
      }
      ss = warp_sum(ss);
      const float inv = 1.f / fmaxf(sqrtf(ss), p.eps);
      uint32_t* out = reinterpret_cast<uint32_t*>(p.ehat + static_cast<size_t>(dr) * D);
#pragma unroll
      for (int k = 0; k < WPR; ++k) out[lane + 32 * k] = pack_bf16x2(v[k].x * inv, v[k].y * inv);
      if (lane == 0) p.inv_norm[dr] = inv;
    }
  }
}

// NCHW input: the same job with 16 warps per 128 pixels -- warp = (32 pixels, a quarter of the
// channels), lane = pixel, so every channel read is one coalesced 128-byte line and each value is
// read ONCE (kept in registers between the norm and the scale); 4x the loads in flight of the
// one-warp-per-32-pixels form above, which left this kernel latency bound (5 warps per SM).
constexpr int PREP_THREADS = 512;
template <int D>
__global__ void __launch_bounds__(PREP_THREADS) prepare_nchw_kernel(const PrepareParams p) {
  pdl_trigger();
  pdl_wait();
  constexpr int DQ = D / 4;            // channels per warp
  constexpr int WPR = D / 64;          // 32-bit words (bf16 pairs) per lane per row
  constexpr int ROW_W = D / 2 + 1;     // smem row stride in words (odd: conflict-free columns)
  extern __shared__ uint32_t smem_u[];  // [128][ROW_W]
  __shared__ uint32_t warp_hist[SORT_WARPS][MAX_BINS];
  __shared__ float ss_sm[4][SORT_CHUNK];
  __shared__ int dest_sm[SORT_CHUNK];
  __shared__ int cls_sm[SORT_CHUNK];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pg = warp & 3, cq = warp >> 2;
  for (int i = threadIdx.x; i < SORT_WARPS * MAX_BINS; i += PREP_THREADS) (&warp_hist[0][0])[i] = 0u;
  __syncthreads();
  const long long n0 = static_cast<long long>(blockIdx.x) * SORT_CHUNK + pg * 32;
  const long long i = n0 + lane;
  const bool live = i < p.n;
  // values first: the loads are in flight while the sort position is worked out
  float v[DQ];
  {
    const long long b = live ? i / p.ppi : 0;
    const float* src = p.x + (b * D + cq * DQ) * p.ppi + (live ? i - b * p.ppi : 0);
#pragma unroll
    for (int d = 0; d < DQ; ++d) v[d] = live ? __ldg(src + static_cast<long long>(d) * p.ppi) : 0.f;
  }
  int c = 0, rank = 0;
  if (cq == 0) {
    c = live ? load_class(p.cls, p.cls64, i, p.C) : (MAX_BINS + lane);  // dead lanes: unique keys
    const uint32_t peers = __match_any_sync(0xffffffffu, c);
    rank = __popc(peers & ((1u << lane) - 1u));
    if (live && rank == 0) warp_hist[pg][c] = __popc(peers);
  }
  float ss = 0.f;
#pragma unroll
  for (int d = 0; d < DQ; ++d) ss = fmaf(v[d], v[d], ss);
  ss_sm[cq][pg * 32 + lane] = ss;
  __syncthreads();
  if (cq == 0 && live) {
    uint32_t off = p.chunk_off[static_cast<size_t>(blockIdx.x) * MAX_BINS + c];
    for (int w = 0; w < pg; ++w) off += warp_hist[w][c];
    const int dest = static_cast<int>(off) + rank;
    dest_sm[pg * 32 + lane] = dest;
    p.perm[i] = dest;
    p.bits[dest] = (c < p.C) ? (1u << c) : 0u;
  }
  const int px = pg * 32 + lane;
  const float tot = (ss_sm[0][px] + ss_sm[1][px]) + (ss_sm[2][px] + ss_sm[3][px]);
  const float inv = 1.f / fmaxf(sqrtf(tot), p.eps);
  uint32_t* tile = smem_u + px * ROW_W + cq * (DQ / 2);
#pragma unroll
  for (int d = 0; d < DQ; d += 2) tile[d >> 1] = pack_bf16x2(v[d] * inv, v[d + 1] * inv);
  __syncthreads();
  if (cq == 0 && live) p.inv_norm[dest_sm[px]] = inv;
  // rows out: warp w writes rows w, w+16, ... (a full 2*D-byte row per step, coalesced)
  const int n_live = static_cast<int>(min(static_cast<long long>(SORT_CHUNK),
                                          p.n - static_cast<long long>(blockIdx.x) * SORT_CHUNK));
  for (int r = warp; r < n_live; r += PREP_THREADS / 32) {
    const uint32_t* trow = smem_u + r * ROW_W;
    uint32_t* out = reinterpret_cast<uint32_t*>(p.ehat + static_cast<size_t>(dest_sm[r]) * D);
#pragma unroll
    for (int k = 0; k < WPR; ++k) out[lane + 32 * k] = trow[lane + 32 * k];
  }
  // per-class column sums of this block's (bf16-rounded) rows: thread = (bf16-pair column, slice of the 128 rows); a
  // chunk of 128 consecutive pixels holds two or three classes on blob-like maps, so this is a few dozen shared loads
  if (p.class_sum != nullptr) {
    constexpr int NCOL = D / 2, NSL = PREP_THREADS / NCOL, RPS = SORT_CHUNK / NSL;
    __shared__ float2 csum_sm[PREP_THREADS];
    if (cq == 0) cls_sm[px] = live ? c : MAX_BINS;
    __syncthreads();
    const int col = threadIdx.x % NCOL, sl = threadIdx.x / NCOL;
    for (int cc = 0; cc < p.C; ++cc) {
      if (warp_hist[0][cc] + warp_hist[1][cc] + warp_hist[2][cc] + warp_hist[3][cc] == 0u) continue;   // block-uniform
      float a0 = 0.f, a1 = 0.f;
#pragma unroll 4
      for (int r = sl * RPS; r < (sl + 1) * RPS; ++r)
        if (cls_sm[r] == cc) {
          const uint32_t wv = smem_u[r * ROW_W + col];
          a0 += __uint_as_float(wv << 16);
          a1 += __uint_as_float(wv & 0xffff0000u);
        }
      csum_sm[threadIdx.x] = make_float2(a0, a1);
      __syncthreads();
      if (threadIdx.x < NCOL) {
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int k = 0; k < NSL; ++k) { const float2 t = csum_sm[k * NCOL + threadIdx.x]; s0 += t.x; s1 += t.y; }
        float* dst = p.class_sum + static_cast<size_t>(cc) * D + 2 * threadIdx.x;
        if (s0 != 0.f) atomicAdd(dst, s0);
        if (s1 != 0.f) atomicAdd(dst + 1, s1);
      }
      __syncthreads();
    }
  }
}

// ------------------------------------------------------------------------------------------
// Segmented column sums over the class-sorted rows.  One block per SEG_ROWS rows; a block is cut
// into at most a few single-class pieces at the segment starts, every piece is reduced in
// registers + shared memory and leaves the block as one atomicAdd per channel.
constexpr int SEG_THREADS = 256;
#ifndef MB_SEG_ROWS
#define MB_SEG_ROWS 64
#endif
constexpr int SEG_ROWS = MB_SEG_ROWS;   // sorted rows per block (small blocks: these kernels are latency bound;
                                        // measured 64 / 32 / 16: 0.2835 / 0.2870 / 0.2913 ms per config-2 step, profiles/r1l_variants_seg_rows.txt)
static_assert(128 % SEG_ROWS == 0 && SEG_ROWS >= 8, "a block covers a whole fraction of a 128-row tile");

template <int D>
__device__ __forceinline__ void load_row8(const __nv_bfloat16* row, int l, float (&v)[8]) {
  const uint4 t = __ldg(reinterpret_cast<const uint4*>(row) + l);
  const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
    v[2 * i] = __low2float(a);
    v[2 * i + 1] = __high2float(a);
  }
}

// red: shared [SEG_THREADS / (D/8)][D]; adds the block's column sums of `acc` to out[0..D)
template <int D>
__device__ __forceinline__ void block_colsum_flush(float (&acc)[8], float* red, float* out) {
  constexpr int TPR = D / 8;
  constexpr int RPP = SEG_THREADS / TPR;
  const int sub = threadIdx.x / TPR, l = threadIdx.x % TPR;
#pragma unroll
  for (int i = 0; i < 8; ++i) red[sub * D + l * 8 + i] = acc[i];
  __syncthreads();
  for (int d = threadIdx.x; d < D; d += SEG_THREADS) {
    float s = 0.f;
#pragma unroll 8
    for (int r = 0; r < RPP; ++r) s += red[r * D + d];
    if (s != 0.f) atomicAdd(out + d, s);
  }
  __syncthreads();
}

// Bit c set <=> class c has rows in [r0, r1): one parallel look at the segment table instead of a
// serial walk over all classes in every block (C <= 31).  Ends with a __syncthreads().
__device__ __forceinline__ uint32_t block_class_mask(const uint32_t* seg, int C, long long r0, long long r1,
                                                     uint32_t* sh) {
  if (threadIdx.x < 32) {
    const int c = threadIdx.x;
    bool ov = false;
    if (c < C) {
      const long long lo = max(r0, static_cast<long long>(seg[c]));
      const long long hi = min(r1, static_cast<long long>(seg[c + 1]));
      ov = lo < hi;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, ov);
    if (threadIdx.x == 0) *sh = m;
  }
  __syncthreads();
  return *sh;
}

// class_sum[c][:] += sum of the (bf16-rounded) rows of class c in this tile
template <int D>
__global__ void __launch_bounds__(SEG_THREADS) class_sum_kernel(const __nv_bfloat16* ehat,
                                                                const uint32_t* seg, int C,
                                                                float* class_sum) {
  pdl_trigger();
  pdl_wait();
  constexpr int TPR = D / 8;
  constexpr int RPP = SEG_THREADS / TPR;
  __shared__ float red[RPP * D];
  const int sub = threadIdx.x / TPR, l = threadIdx.x % TPR;
  __shared__ uint32_t cmask_sh;
  const long long r0 = static_cast<long long>(blockIdx.x) * SEG_ROWS, r1 = r0 + SEG_ROWS;
  for (uint32_t cm = block_class_mask(seg, C, r0, r1, &cmask_sh); cm != 0u; cm &= cm - 1u) {
    const int c = __ffs(cm) - 1;
    const long long lo = max(r0, static_cast<long long>(seg[c]));
    const long long hi = min(r1, static_cast<long long>(seg[c + 1]));
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (long long r = lo + sub; r < hi; r += RPP) {
      float v[8];
      load_row8<D>(ehat + r * D, l, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += v[i];
    }
    block_colsum_flush<D>(acc, red, class_sum + c * D);
  }
}

// class_sum[c][:] = proto[c][:] * class_mass[c]: prototypes handed in by the caller (e.g. CAM-weighted ones from
// muscle_pool_prototypes_fwd) take the place of the class means; everything downstream divides by the mass again.
__global__ void proto_to_sum_kernel(const float* proto, const float* class_mass, int C, int dim, float* class_sum) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < C * dim) class_sum[i] = proto[i] * class_mass[i / dim];
}

struct MidParams {
  const __nv_bfloat16* ehat;   // sorted rows
  const uint32_t* seg;         // [C+2]
  int C;
  const float* class_sum;      // [C][D]  (global sums when sharded)
  const float* class_mass;     // [C]
  float* rowsum;               // in: sum over negatives of exp((s-1)/tau); out: + the positive term
  float inv_tau;
  float* w;                    // [rows_pad] 1/rowsum, 0 for rows outside the loss
  float* gpos;                 // [rows_pad] d loss_q / d pos-logit_q (before the 1/n_valid factor)
  double* loss_acc;            // += sum_q (1/tau + ln rowsum_q - pos_q)
  float* dproto;               // [C][D] += sum_q gpos_q e_q
  float* loss_out;             // optional: the last block to finish writes loss_acc / n_valid here
  unsigned int* ticket;        // zeroed counter of finished blocks (with loss_out)
};

// Same tiling as class_sum_kernel.  Rows of the invalid segment and pad rows get w = gpos = 0.
template <int D>
__global__ void __launch_bounds__(SEG_THREADS) mid_kernel(const MidParams p) {
  pdl_trigger();
  pdl_wait();
  constexpr int TPR = D / 8;
  constexpr int RPP = SEG_THREADS / TPR;
  __shared__ float red[RPP * D];
  __shared__ double loss_sh[SEG_THREADS / 32];
  const int sub = threadIdx.x / TPR, l = threadIdx.x % TPR;
  const long long r0 = static_cast<long long>(blockIdx.x) * SEG_ROWS, r1 = r0 + SEG_ROWS;
  const float bt = p.inv_tau;  // the shift used by the pair passes: exp((s - 1)/tau)
  double loss = 0.0;
  __shared__ uint32_t cmask_sh;
  for (uint32_t cm = block_class_mask(p.seg, p.C, r0, r1, &cmask_sh); cm != 0u; cm &= cm - 1u) {
    const int c = __ffs(cm) - 1;
    const long long lo = max(r0, static_cast<long long>(p.seg[c]));
    const long long hi = min(r1, static_cast<long long>(p.seg[c + 1]));
    const float inv_mass = 1.f / __ldg(p.class_mass + c);
    float proto[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) proto[i] = __ldg(p.class_sum + c * D + l * 8 + i) * inv_mass;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // MID_BATCH passes of RPP rows at a time: all their rows and row sums are requested before the first dot product
    // (one pass per iteration left four dependent load -> shuffle -> exp chains in a row per 64-row block)
    constexpr int MID_BATCH = 4;
    for (long long rr = lo; rr < hi; rr += MID_BATCH * RPP) {   // uniform trip count: shuffles stay converged
      float v[MID_BATCH][8];
      float rs_in[MID_BATCH];
#pragma unroll
      for (int j = 0; j < MID_BATCH; ++j) {
        const long long r = rr + j * RPP + sub;
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j][i] = 0.f;
        rs_in[j] = 0.f;
        if (r < hi) {
          load_row8<D>(p.ehat + r * D, l, v[j]);
          rs_in[j] = p.rowsum[r];
        }
      }
#pragma unroll
      for (int j = 0; j < MID_BATCH; ++j) {
        if (rr + j * RPP >= hi) break;              // block-uniform
        const long long r = rr + j * RPP + sub;
        const bool on = r < hi;
        float dot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) dot = fmaf(v[j][i], proto[i], dot);
#pragma unroll
        for (int o = TPR / 2; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
        if (on) {
          const float pos = dot * p.inv_tau;
          const float ep = __expf(pos - bt);
          const float rs = rs_in[j] + ep;
          const float w = 1.f / rs;
          const float g = (ep * w - 1.f) * p.inv_tau;
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = fmaf(g, v[j][i], acc[i]);
          if (l == 0) {
            p.rowsum[r] = rs;
            p.w[r] = w;
            p.gpos[r] = g;
            loss += static_cast<double>(bt + logf(rs) - pos);
          }
        }
      }
    }
    block_colsum_flush<D>(acc, red, p.dproto + c * D);
  }
  // rows outside every class (invalid pixels, pad rows)
  {
    const long long lo = max(r0, static_cast<long long>(p.seg[p.C]));
    for (long long r = lo + threadIdx.x; r < r1; r += SEG_THREADS) {
      p.w[r] = 0.f;
      p.gpos[r] = 0.f;
    }
  }
  // block loss
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
  if ((threadIdx.x & 31) == 0) loss_sh[threadIdx.x >> 5] = loss;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < SEG_THREADS / 32; ++i) t += loss_sh[i];
    if (t != 0.0) atomicAdd(p.loss_acc, t);
    if (p.loss_out != nullptr) {   // single-GPU composite: the mean is written here, no extra launch
      __threadfence();
      if (atomicAdd(p.ticket, 1u) == gridDim.x - 1) {
        __threadfence();
        float nv = 0.f;
        for (int c = 0; c < p.C; ++c) nv += p.class_mass[c];
        const double tot = atomicAdd(p.loss_acc, 0.0);
        *p.loss_out = static_cast<float>(tot / static_cast<double>(nv));  // 0/0 -> NaN like the reference
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
struct FinalParams {
  const __nv_bfloat16* ehat;   // sorted
  const uint32_t* bits;        // sorted
  const float* inv_norm;       // sorted
  const int* perm;             // [n] source pixel -> sorted row
  const float* part;           // [PAIR_SLOTS][rows_pad][D]
  const uint32_t* unit_off;    // [n_row_blocks+1]  the pair passes' work partition (pair_plan_kernel)
  int n_row_blocks;
  int n_col_tiles;
  int n_ctas;                  // grid of the gradient pass
  long long rows_pad;
  long long n;
  long long ppi;
  int C;
  const float* class_sum;
  const float* class_mass;
  const float* gpos;           // sorted
  const float* dproto;
  const double* loss_acc;
  float inv_tau;
  int layout;
  float* dx;                   // source order, `layout`
  float* loss_out;
  const float* grad_scale;     // upstream gradient of the loss (device scalar), or null = 1
  int proto_external;          // the prototypes came from the caller: no gradient to the pixels through them here
  float* dproto_out;           // optional [C][D]: d loss / d prototype (times grad_scale)
};

// One block = FINAL_CHUNK consecutive source pixels, one warp per 8 pixels (two pixels in flight per
// warp, so the dependent gather -> reduce -> store chains of different pixels overlap).  Per pixel
// the warp gathers its sorted row (lane holds channels lane, lane+32, ...), forms
//   g = 1/n_valid (1/tau sum_slots part + gpos/mass class_sum_c + dproto_c/mass),
//   dx = inv_norm (g - e (g.e))
// and writes it back in the caller's layout (NCHW through a shared-memory transpose, so the global
// stores are coalesced along the pixel index).  64-pixel blocks (round 1: 128): 392 blocks instead of 196 at
// config 2 fill 148 SMs more evenly, and six of them fit an SM's shared memory instead of three.
#ifndef MB_FINAL_CHUNK
#define MB_FINAL_CHUNK 64
#endif
constexpr int FINAL_CHUNK = MB_FINAL_CHUNK;
constexpr int FINAL_ROWS_PER_WARP = 8;
constexpr int FINAL_THREADS = FINAL_CHUNK / FINAL_ROWS_PER_WARP * 32;
static_assert(FINAL_CHUNK == 64 || FINAL_CHUNK == 128, "final_kernel: 64 or 128 pixels per block");

template <int D>
__global__ void __launch_bounds__(FINAL_THREADS) final_kernel(const FinalParams p) {
  pdl_trigger();
  pdl_wait();
  constexpr int EPL = D / 32;
  extern __shared__ float smem_f[];  // NCHW only: [FINAL_CHUNK][D+1]
  __shared__ float n_valid_sh;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    float nv = 0.f;
    for (int c = 0; c < p.C; ++c) nv += p.class_mass[c];
    n_valid_sh = nv;
    if (blockIdx.x == 0 && p.loss_out != nullptr)
      *p.loss_out = static_cast<float>(*p.loss_acc / static_cast<double>(nv));  // 0/0 -> NaN like the reference
  }
  __syncthreads();
  if (p.dx == nullptr) return;
  const float inv_n = (n_valid_sh > 0.f ? 1.f / n_valid_sh : 0.f) * (p.grad_scale ? __ldg(p.grad_scale) : 1.f);
  if (p.dproto_out != nullptr && blockIdx.x == 0)
    for (int i = threadIdx.x; i < p.C * D; i += FINAL_THREADS) p.dproto_out[i] = inv_n * __ldg(p.dproto + i);
  constexpr int RB = 128;  // rows per row block of the pair passes (PAIR_RB)
  const uint32_t per = pair_units_per_cta(__ldg(p.unit_off + p.n_row_blocks), p.n_ctas, p.n_col_tiles);
  const long long blk0 = static_cast<long long>(blockIdx.x) * FINAL_CHUNK;

  // Per-row scalars of this warp's rows, fetched up front by lane (r mod 8 = row): the dependent
  // chain perm -> bits / unit_off / gpos / inv_norm is paid once per warp instead of once per row.
  static_assert(FINAL_ROWS_PER_WARP == 8, "lane & 7 = row of the warp");
  const long long my_i = blk0 + warp * FINAL_ROWS_PER_WARP + (lane & 7);
  long long q_l = 0;
  uint32_t bq_l = 0u;
  int ns_l = 0;
  float gq_l = 0.f, inv_l = 0.f, im_l = 0.f;
  if (my_i < p.n) {
    q_l = p.perm[my_i];
    bq_l = __ldg(p.bits + q_l);
    if (bq_l) {
      // the row block's live units were shared by CTAs U0/per .. (U1-1)/per, one slot each
      const int rbk = static_cast<int>(q_l / RB);
      const uint32_t U0 = __ldg(p.unit_off + rbk), U1 = __ldg(p.unit_off + rbk + 1);
      ns_l = (U1 > U0) ? static_cast<int>((U1 - 1) / per - U0 / per + 1) : 0;
      im_l = 1.f / __ldg(p.class_mass + class_of_bits(bq_l));
      gq_l = __ldg(p.gpos + q_l) * im_l;
      inv_l = __ldg(p.inv_norm + q_l);
    }
  }

  // Two phases per batch of FINAL_INFLIGHT rows, written out so that the loads of ALL rows of the batch are issued
  // before the first use: (1) the first two partial-sum slots (predicated; 1.75 slots per row on average at config 2)
  // and the bf16 row itself, (2) further slots (rare), prototype terms, the normalisation's backward.  With one lambda
  // per row (an early return for rows outside the loss, a slot loop of variable length) the compiler kept the rows of a
  // warp strictly one after the other, and neither smaller blocks nor more rows "in flight" changed the 23 us.
  // A row outside the loss (bq == 0: ignored pixel, or a pixel past n) has n_slots = 0 and inv = 0: it comes out as zeros.
#ifndef MB_FINAL_INFLIGHT
#define MB_FINAL_INFLIGHT 4
#endif
  constexpr int NF = MB_FINAL_INFLIGHT;
  static_assert(FINAL_ROWS_PER_WARP % NF == 0, "rows in flight divide the warp's rows");
#pragma unroll 1
  for (int r = 0; r < FINAL_ROWS_PER_WARP; r += NF) {
    float g[NF][EPL], e[NF][EPL];
    long long q[NF];
    int ns[NF];
    // ---- phase 1
#pragma unroll
    for (int j = 0; j < NF; ++j) {
      q[j] = __shfl_sync(0xffffffffu, q_l, r + j);
      ns[j] = __shfl_sync(0xffffffffu, ns_l, r + j);
      const float* s0 = p.part + q[j] * D + lane;
      const float* s1 = s0 + p.rows_pad * D;
#pragma unroll
      for (int k = 0; k < EPL; ++k) {
        const float a0 = ns[j] > 0 ? __ldg(s0 + 32 * k) : 0.f;
        const float a1 = ns[j] > 1 ? __ldg(s1 + 32 * k) : 0.f;
        g[j][k] = a0 + a1;
        e[j][k] = __bfloat162float(p.ehat[q[j] * D + lane + 32 * k]);
      }
    }
    // ---- phase 2
#pragma unroll
    for (int j = 0; j < NF; ++j) {
      const uint32_t bq = __shfl_sync(0xffffffffu, bq_l, r + j);
      const float gq = __shfl_sync(0xffffffffu, gq_l, r + j);
      const float inv = __shfl_sync(0xffffffffu, inv_l, r + j);
      const float inv_mass = __shfl_sync(0xffffffffu, im_l, r + j);
      const float key_mass = p.proto_external ? 0.f : inv_mass;   // pixel -> its class mean -> every query of the class
      const int c = bq ? class_of_bits(bq) : 0;
      for (int s = 2; s < ns[j]; ++s) {                            // warp-uniform trip count
        const float* src = p.part + (static_cast<long long>(s) * p.rows_pad + q[j]) * D + lane;
#pragma unroll
        for (int k = 0; k < EPL; ++k) g[j][k] += __ldg(src + 32 * k);
      }
      float dot = 0.f;
#pragma unroll
      for (int k = 0; k < EPL; ++k) {
        const int d = lane + 32 * k;
        g[j][k] = inv_n * (p.inv_tau * g[j][k] + gq * __ldg(p.class_sum + c * D + d) + key_mass * __ldg(p.dproto + c * D + d));
        dot = fmaf(g[j][k], e[j][k], dot);
      }
      dot = warp_sum(dot);
#pragma unroll
      for (int k = 0; k < EPL; ++k) g[j][k] = inv * (g[j][k] - e[j][k] * dot);
    }
    // ---- out
#pragma unroll
    for (int j = 0; j < NF; ++j) {
      const int l = warp * FINAL_ROWS_PER_WARP + r + j;      // local pixel index within the block
      const long long i = blk0 + l;
      if (i < p.n) {
        if (p.layout == MUSCLE_LAYOUT_NCHW) {
#pragma unroll
          for (int k = 0; k < EPL; ++k) smem_f[l * (D + 1) + lane + 32 * k] = g[j][k];
        } else {
#pragma unroll
          for (int k = 0; k < EPL; ++k) p.dx[i * D + lane + 32 * k] = g[j][k];
        }
      }
    }
  }
  if (p.layout == MUSCLE_LAYOUT_NCHW) {
    __syncthreads();
    // thread -> (pixel = tid % FINAL_CHUNK, channel group = tid / FINAL_CHUNK): consecutive threads write consecutive pixels
    const int px = threadIdx.x & (FINAL_CHUNK - 1);
    const long long i = blk0 + px;
    if (i < p.n) {
      const long long b = i / p.ppi;
      float* dst = p.dx + (b * D) * p.ppi + (i - b * p.ppi);
      for (int d = threadIdx.x / FINAL_CHUNK; d < D; d += FINAL_THREADS / FINAL_CHUNK)
        dst[static_cast<long long>(d) * p.ppi] = smem_f[px * (D + 1) + d];
    }
  }
}

}  // namespace mb200
