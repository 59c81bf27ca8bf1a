// C-ABI of the dense label-masked pixel InfoNCE (include/muscle_b200.h) and its launch logic.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>

#include "host_util.cuh"
#include "pair_kernel.cuh"
#include "pair_sym_kernel.cuh"
#include "row_kernels.cuh"

using namespace mb200;

// Scratch of muscle_infonce_plan_sym_ring_static (the allocation-free variant of muscle_infonce_plan_sym_ring): the
// per-row-tile live-unit counts.  Lives in the module image: every GPU that loads the library holds these 64 KB until
// exit.  In HOST code this name is a stand-in, not a device address -- static_plan_scratch() asks the runtime for it.
constexpr int kStaticPlanMaxTiles = 16384;       // row tiles of 128 pixels: rows_pad <= 2 097 152
static __device__ __align__(256) uint32_t g_static_plan_cnt[kStaticPlanMaxTiles + 1];

namespace {

constexpr float kLog2e = 1.4426950408889634f;

bool dim_ok(int dim) { return dim == 64 || dim == 128 || dim == 256; }

int check_device() {
  const DeviceInfo& di = device_info();
  if (!di.ok) return fail(MUSCLE_ERR_CUDA, "no CUDA device");
  if (di.cc_major != 10)
    return fail(MUSCLE_ERR_UNSUPPORTED, "muscle_b200 needs an sm_100 GPU (found cc %d.x)", di.cc_major);
  return MUSCLE_OK;
}

// grid of the pair passes: one persistent CTA per SM (MUSCLE_PAIR_GRID overrides it for experiments)
int pair_grid() {
  static const int g = getenv("MUSCLE_PAIR_GRID") ? atoi(getenv("MUSCLE_PAIR_GRID")) : 0;
  return g > 0 ? g : device_info().sm_count;
}

// Tensor maps depend only on (base, rows, D); a training loop presents the same few buffers again and
// again (the caching allocator hands the workspace back at the same address), and one
// cuTensorMapEncodeTiled costs a few microseconds of host time, four times per step.
int cached_panel_map(const void* base, int64_t rows, int dim, CUtensorMap* out) {
  struct MapEntry { const void* base; int64_t rows; int dim; CUtensorMap map; };
  static thread_local MapEntry cache[8] = {};
  static thread_local int next_slot = 0;
  for (const MapEntry& e : cache)
    if (e.base == base && e.rows == rows && e.dim == dim) { *out = e.map; return MUSCLE_OK; }
  MapEntry& e = cache[next_slot];
  next_slot = (next_slot + 1) % 8;
  e.base = nullptr;
  if (int rc2 = make_bf16_panel_map(&e.map, base, rows, dim)) return rc2;
  e.base = base; e.rows = rows; e.dim = dim;
  *out = e.map;
  return MUSCLE_OK;
}

#if MB_PAIR_TIMING
template <int D, int MODE>
void pair_timing_report(unsigned long long* dbg, int grid, cudaStream_t stream) {
    static int printed = 0;
    cudaStreamSynchronize(stream);
    unsigned long long h[148 * 16];
    cudaMemcpy(h, dbg, std::min(grid, 148) * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    if (printed++ < 6) {
      double s[16] = {0};
      int n = 0;
      for (int c = 0; c < std::min(grid, 148); ++c)
        if (h[c * 16 + 6]) { for (int k = 0; k < 15; ++k) s[k] += double(h[c * 16 + k]); ++n; }
      if (n) {
        const double tiles = s[6] / n;
        fprintf(stderr, "[pair D=%d mode=%d] active CTAs %d, units/CTA (group 0) %.1f | per own unit: q %.0f kfull %.0f sfull %.0f ld %.0f math %.0f tail %.0f cycles\n",
                D, MODE, n, tiles, s[0] / n / tiles, s[1] / n / tiles, s[2] / n / tiles, s[3] / n / tiles, s[4] / n / tiles, s[5] / n / tiles);
        {
          unsigned long long mn = ~0ull, mx = 0;
          for (int c = 0; c < std::min(grid, 148); ++c)
            if (h[c * 16 + 6]) { mn = std::min(mn, h[c * 16 + 12]); mx = std::max(mx, h[c * 16 + 12]); }
          fprintf(stderr, "    issuer lifetime per CTA: min %llu mean %.0f max %llu cycles\n", mn, s[12] / n, mx);
          unsigned long long t0 = ~0ull, t1 = 0, t0x = 0;
          double span = 0;
          for (int c = 0; c < std::min(grid, 148); ++c) {
            t0 = std::min(t0, h[c * 16 + 14]); t0x = std::max(t0x, h[c * 16 + 14]); t1 = std::max(t1, h[c * 16 + 15]);
            span += double(h[c * 16 + 15] - h[c * 16 + 14]);
          }
          fprintf(stderr, "    globaltimer: first CTA start -> last CTA end %.1f us, CTA starts spread over %.1f us, mean CTA span %.1f us"
                          " (mean SM clock while running %.0f MHz)\n",
                  (t1 - t0) / 1e3, (t0x - t0) / 1e3, span / std::min(grid, 148) / 1e3, s[12] / n / (span / std::min(grid, 148)) * 1e3);
          if (getenv("MUSCLE_PAIR_TIMING_ALL")) {
            fprintf(stderr, "    per CTA (issuer lifetime k cycles / group-0 units: all, select-masked, bias-masked / epilogue math k cycles):");
            for (int c = 0; c < std::min(grid, 148); ++c)
              fprintf(stderr, "%s%llu/%llu,%llu,%llu/%llu", c % 6 ? " " : "\n      ", h[c * 16 + 12] / 1000, h[c * 16 + 6], h[c * 16 + 7], h[c * 16 + 13], h[c * 16 + 4] / 1000);
            fprintf(stderr, "\n");
          }
        }
        fprintf(stderr, "    producer: wait k_empty %.0f of %.0f per unit | issuer: wait k_full %.0f, wait t_full %.0f of %.0f per unit\n",
                s[8] / n / tiles / 2, s[9] / n / tiles / 2, s[10] / n / tiles / 2, s[11] / n / tiles / 2, s[12] / n / tiles / 2);
      }
    }
}
#endif

template <int D, int MODE, int POLY>
int launch_pair(const void* ehat_rows, int64_t rows_pad, const void* ehat_cols, int64_t cols_pad,
                PairParams prm, cudaStream_t stream) {
  using C = PairCfg<D, MODE>;
  CUtensorMap tm_row, tm_col;
  int rc = cached_panel_map(ehat_rows, rows_pad, D, &tm_row);
  if (rc) return rc;
  rc = cached_panel_map(ehat_cols, cols_pad, D, &tm_col);
  if (rc) return rc;
  if (int rc2 = ensure_dynamic_smem<pair_kernel<D, MODE, POLY>>(C::SMEM_BYTES)) return rc2;
  const int grid = pair_grid();  // one persistent CTA per SM; shares come from unit_off
#if MB_PAIR_TIMING
  // measurement builds only (scripts/build_variants.sh with -DMB_PAIR_TIMING=1): per-phase cycle counters
  static const bool timing = getenv("MUSCLE_PAIR_TIMING") != nullptr;
  static unsigned long long* dbg = nullptr;
  static const int flags = getenv("MUSCLE_PAIR_FLAGS") ? atoi(getenv("MUSCLE_PAIR_FLAGS")) : 0;
  prm.dbg_flags = flags;
  if (timing) {
    if (!dbg) cudaMalloc(&dbg, grid * 16 * sizeof(unsigned long long));
    cudaMemsetAsync(dbg, 0, grid * 16 * sizeof(unsigned long long), stream);
    prm.dbg = dbg;
  }
#endif
  MB_CUDA(launch_pdl(pair_kernel<D, MODE, POLY>, dim3(grid), dim3(C::THREADS), C::SMEM_BYTES, stream, tm_row, tm_col, prm));
  MB_LAUNCH_CHECK();
#if MB_PAIR_TIMING
  if (timing) pair_timing_report<D, MODE>(dbg, grid, stream);
#endif
  return MUSCLE_OK;
}

// The FMA-pipe exponential is exact enough only while the exponent stays in range:
// x = (s - 1) log2(e)/T lies in [-2 log2(e)/T, ~0].
// Share on the FMA pipe: 1 of every 4 element pairs.  Measured on B200 (profiles/r1l_variants_poly.txt, three
// alternating runs): share 2/4 -> 1/4 takes the config-2 step from 0.277 to 0.268 ms and config 3 from 1.31
// to 1.22 ms (both passes are issue bound, and a polynomial pair costs ~10 issue slots against 2 for two
// MUFU.EX2); 0/4 is slower again in the row-sum pass (MUFU, 16/clk/SM, becomes the cap), 3/4 and 4/4 slower.
constexpr int kPoly = 1;
// (MASK_BIAS pushes masked columns another pair_mask_shift down: x in [-2 log2(e)/T - shift, ~0])
bool poly_ok(float temperature) { return pair_poly_ok(kLog2e / temperature); }
// exp((s - 1)/T) leaves the fp32 range for every s < 0 once log2(e)/T > 126: refuse instead of returning
// a silently wrong loss (the reference works in log space and has no such limit)
constexpr float kMinTemperature = MUSCLE_MIN_TEMPERATURE;

template <int MODE>
int launch_pair_dim(int dim, bool poly, const void* ehat_rows, int64_t rows_pad,
                    const void* ehat_cols, int64_t cols_pad, PairParams prm, cudaStream_t stream) {
#if MB_PAIR_TIMING
  static const int dbg = getenv("MUSCLE_DEBUG_PAIR") ? atoi(getenv("MUSCLE_DEBUG_PAIR")) : 0;  // measurement variants
#define MB_PAIR(DD)                                                                              \
  if (DD == 128 && dbg == 1) return launch_pair<128, MODE, -1>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream); \
  if (DD == 128 && dbg == 2) return launch_pair<128, MODE, 4>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream); \
  if (DD == 128 && dbg == 3) return launch_pair<128, MODE, 0>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream); \
  if (DD == 128 && dbg == 5) return launch_pair<128, MODE, 3>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream); \
  if (DD == 128 && dbg == 6) return launch_pair<128, MODE, 2>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream); \
  return poly ? launch_pair<DD, MODE, kPoly>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream) \
              : launch_pair<DD, MODE, 0>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream);
#else
#define MB_PAIR(DD)                                                                              \
  return poly ? launch_pair<DD, MODE, kPoly>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream) \
              : launch_pair<DD, MODE, 0>(ehat_rows, rows_pad, ehat_cols, cols_pad, prm, stream);
#endif
  switch (dim) {
    case 64: MB_PAIR(64)
    case 128: MB_PAIR(128)
    case 256: MB_PAIR(256)
  }
#undef MB_PAIR
  return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
}

// Row-sum pass over rows == columns (pair_sym_kernel.cuh): only the live units with c >= r.
template <int D, int POLY, int NG>
int launch_rowsum_sym_ng(const void* ehat, int64_t rows_pad, const void* ehat_cols, int64_t cols_pad, PairParams prm,
                         cudaStream_t stream) {
  using C = SymCfg<D, NG>;
  CUtensorMap tm, tm_col;
  if (int rc = cached_panel_map(ehat, rows_pad, D, &tm)) return rc;
  if (int rc = cached_panel_map(ehat_cols, cols_pad, D, &tm_col)) return rc;
  if (int rc = ensure_dynamic_smem<pair_sym_kernel<D, POLY, NG>>(C::SMEM_BYTES)) return rc;
#if MB_PAIR_TIMING
  static const bool timing = getenv("MUSCLE_PAIR_TIMING") != nullptr;
  static unsigned long long* dbg = nullptr;
  if (timing) {
    if (!dbg) cudaMalloc(&dbg, pair_grid() * 16 * sizeof(unsigned long long));
    cudaMemsetAsync(dbg, 0, pair_grid() * 16 * sizeof(unsigned long long), stream);
    prm.dbg = dbg;
  }
#endif
  MB_CUDA(launch_pdl(pair_sym_kernel<D, POLY, NG>, dim3(pair_grid()), dim3(C::THREADS), C::SMEM_BYTES, stream, tm, tm_col, prm));
  MB_LAUNCH_CHECK();
#if MB_PAIR_TIMING
  if (timing) {
    static int printed = 0;
    cudaStreamSynchronize(stream);
    const int grid = std::min(pair_grid(), 148);
    unsigned long long h[148 * 16];
    cudaMemcpy(h, dbg, grid * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    if (printed++ < 4) {
      double s[16] = {0};
      int n = 0;
      for (int c = 0; c < grid; ++c)
        if (h[c * 16 + 6]) { for (int k = 0; k < 16; ++k) s[k] += double(h[c * 16 + k]); ++n; }
      if (n) {
        const double own = s[6] / n;
        fprintf(stderr, "[pair_sym D=%d NG=%d] active CTAs %d, own units/CTA (group 0) %.1f | per own unit: item %.0f "
                        "ktile %.0f S %.0f drain %.0f ld %.0f math+store %.0f fence+publish %.0f cycles\n",
                D, NG, n, own, s[0] / n / own, s[1] / n / own, s[2] / n / own, s[3] / n / own, s[4] / n / own, s[5] / n / own, s[15] / n / own);
        fprintf(stderr, "    producer: wait k_empty %.0f of %.0f per unit (%.1f units/CTA) | S issuer: wait %.0f (column tile %.0f) of %.0f | column-sum issuer: wait %.0f of %.0f per unit\n",
                s[8] / s[10], s[9] / s[10], s[10] / n, s[11] / s[10], s[7] / s[10], s[12] / s[10], s[13] / s[10], s[14] / s[10]);
      }
    }
  }
#endif
  return MUSCLE_OK;
}
template <int D, int POLY>
int launch_rowsum_sym(const void* ehat, int64_t rows_pad, const void* ehat_cols, int64_t cols_pad, PairParams prm,
                      cudaStream_t stream) {
  return launch_rowsum_sym_ng<D, POLY, 3>(ehat, rows_pad, ehat_cols, cols_pad, prm, stream);
}

// MUSCLE_ROWSUM_SYM=0 keeps the ordered row-sum pass in the single-GPU composite (A/B measurements)
bool rowsum_sym_enabled(int dim) {
  static const bool off = getenv("MUSCLE_ROWSUM_SYM") != nullptr && atoi(getenv("MUSCLE_ROWSUM_SYM")) == 0;
  return !off && (dim == 64 || dim == 128);
}

// rows = this rank's pixels, cols = all ranks' pixels (one GPU: the same arrays, rank 0 of 1); rowsum_all [cols_pad]
// receives the row sums of the rank's own rows (at rank * rows_pad) AND its column-sum contributions to the other ranks'
int rowsum_sym(const void* ehat, const uint32_t* bits, const uint32_t* tor, int64_t rows_pad, const void* ehat_cols,
               const uint32_t* bits_cols, const uint32_t* tor_cols, int64_t cols_pad, const uint32_t* unit_off_sym,
               int rank, int world, int dim, float temperature, float* rowsum_all, cudaStream_t stream,
               int share = 0, int n_shares = 1) {
  PairParams prm{};
  prm.share_g = share; prm.share_G = n_shares;
  prm.bits_row = bits; prm.bits_col = bits_cols; prm.tor_row = tor; prm.tor_col = tor_cols;
  prm.rowsum = rowsum_all + static_cast<size_t>(rank) * rows_pad; prm.rowsum_cols = rowsum_all;
  prm.unit_off = unit_off_sym;
  prm.scale_log2 = kLog2e / temperature;
  const bool poly = poly_ok(temperature);
  prm.mask_shift = pair_mask_shift(prm.scale_log2, poly);
  prm.n_row_blocks = static_cast<int>(rows_pad / 128);
  prm.n_col_tiles = static_cast<int>(cols_pad / 128);
  prm.rows_pad = static_cast<int>(rows_pad);
  prm.rb_off = rank * prm.n_row_blocks; prm.own_g = rank; prm.own_G = world;
  if (dim == 64)
    return poly ? launch_rowsum_sym<64, kPoly>(ehat, rows_pad, ehat_cols, cols_pad, prm, stream)
                : launch_rowsum_sym<64, 0>(ehat, rows_pad, ehat_cols, cols_pad, prm, stream);
  return poly ? launch_rowsum_sym<128, kPoly>(ehat, rows_pad, ehat_cols, cols_pad, prm, stream)
              : launch_rowsum_sym<128, 0>(ehat, rows_pad, ehat_cols, cols_pad, prm, stream);
}

int rows_per_block(int) { return PAIR_RB; }

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" {

const char* muscle_last_error(void) { return err_buf(); }
int muscle_version(void) { return 200; }
int64_t muscle_pad_rows(int64_t n) { return round_up(std::max<int64_t>(n, 1), 256); }

size_t muscle_infonce_prepare_scratch_bytes(int64_t n_pixels) {
  const int64_t n_blk = (std::max<int64_t>(n_pixels, 1) + SORT_CHUNK - 1) / SORT_CHUNK;
  return align_up(static_cast<size_t>(n_blk) * MAX_BINS * sizeof(uint32_t), 256);
}

}  // extern "C"

namespace {
// unit_off != null: rows == columns, the work partition of the pair passes comes out of the scan kernel
int prepare_impl(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                 int dim, const void* cls, int cls_is_int64, int num_classes, float eps,
                 void* ehat, uint32_t* bits, float* inv_norm, int32_t* perm,
                 uint32_t* tor, uint32_t* seg, float* class_sum, float* class_mass,
                 void* scratch, size_t scratch_bytes, uint32_t* unit_off, uint32_t* unit_off_sym, const float* proto_in,
                 cudaStream_t stream) {
  if (int rc = check_device()) return rc;
  if (!x || !cls || !ehat || !bits || !inv_norm || !perm || !tor || !seg || !class_sum ||
      !class_mass || !scratch)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (num_classes < 1 || num_classes > MUSCLE_MAX_CLASSES)
    return fail(MUSCLE_ERR_BAD_ARG, "num_classes must be in [1,%d]", MUSCLE_MAX_CLASSES);
  if (layout != MUSCLE_LAYOUT_PIXEL_MAJOR && layout != MUSCLE_LAYOUT_NCHW)
    return fail(MUSCLE_ERR_BAD_ARG, "unknown layout %d", layout);
  if (batch <= 0 || pixels_per_image <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  const int64_t n = batch * pixels_per_image;
  if (n >= (1ll << 31) - 256) return fail(MUSCLE_ERR_BAD_ARG, "too many pixels for one rank");
  if (scratch_bytes < muscle_infonce_prepare_scratch_bytes(n))
    return fail(MUSCLE_ERR_WORKSPACE, "prepare scratch too small");
  const int64_t rows_pad = muscle_pad_rows(n);
  const int n_blk = static_cast<int>((n + SORT_CHUNK - 1) / SORT_CHUNK);
  uint32_t* hist = static_cast<uint32_t*>(scratch);
  // (pad rows -- zero vectors that belong to no class -- are cleared by the histogram kernel)
  MB_CUDA(launch_pdl(class_hist_kernel, dim3(n_blk), dim3(SORT_CHUNK), 0, stream, cls, cls_is_int64, static_cast<long long>(n), num_classes, hist,
                     static_cast<long long>(rows_pad), dim, static_cast<uint32_t*>(ehat), bits, inv_norm));
  MB_LAUNCH_CHECK();
  MB_CUDA(launch_pdl(class_scan_kernel, dim3(1), dim3(1024), 0, stream, hist, n_blk, num_classes, static_cast<long long>(n), static_cast<long long>(rows_pad), seg, class_mass, tor, unit_off, unit_off_sym));
  MB_LAUNCH_CHECK();
  PrepareParams p;
  p.x = x; p.cls = cls; p.cls64 = cls_is_int64; p.layout = layout; p.n = n;
  p.ppi = pixels_per_image; p.C = num_classes; p.eps = eps; p.chunk_off = hist;
  p.ehat = static_cast<__nv_bfloat16*>(ehat); p.bits = bits; p.inv_norm = inv_norm; p.perm = perm;
  // NCHW input: the class sums come out of prepare_nchw_kernel itself (one launch and one pass over ehat less)
  const bool fold_class_sum = layout == MUSCLE_LAYOUT_NCHW && proto_in == nullptr;
  p.class_sum = fold_class_sum ? class_sum : nullptr;
  const size_t smem = (layout == MUSCLE_LAYOUT_NCHW)
                          ? static_cast<size_t>(SORT_CHUNK) * (dim / 2 + 1) * sizeof(uint32_t) : 0;
  const int n_tiles = static_cast<int>(rows_pad / 128);
#define MB_PREP(DD)                                                                               \
  {                                                                                               \
    if (layout == MUSCLE_LAYOUT_NCHW) {                                                           \
      if (int rca = ensure_dynamic_smem<prepare_nchw_kernel<DD>>(static_cast<int>(SORT_CHUNK * (DD / 2 + 1) * sizeof(uint32_t)))) return rca; \
      MB_CUDA(launch_pdl(prepare_nchw_kernel<DD>, dim3(n_blk), dim3(PREP_THREADS), smem, stream, p)); \
    } else {                                                                                      \
      MB_CUDA(launch_pdl(prepare_kernel<DD>, dim3(n_blk), dim3(SORT_CHUNK), 0, stream, p));       \
    }                                                                                             \
    MB_LAUNCH_CHECK();                                                                            \
    if (proto_in != nullptr)                                                                      \
      MB_CUDA(launch_pdl(proto_to_sum_kernel, dim3((num_classes * DD + 255) / 256), dim3(256), 0, stream, proto_in, \
                         static_cast<const float*>(class_mass), num_classes, DD, class_sum));     \
    else if (!fold_class_sum)                                                                     \
      MB_CUDA(launch_pdl(class_sum_kernel<DD>, dim3(n_tiles * (128 / SEG_ROWS)), dim3(SEG_THREADS), 0, stream, \
                         static_cast<const __nv_bfloat16*>(p.ehat), static_cast<const uint32_t*>(seg), num_classes, class_sum)); \
  }
  if (dim == 64) MB_PREP(64) else if (dim == 128) MB_PREP(128) else MB_PREP(256)
#undef MB_PREP
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
}  // namespace

extern "C" {

int muscle_infonce_prepare(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                           int dim, const void* cls, int cls_is_int64, int num_classes, float eps,
                           void* ehat, uint32_t* bits, float* inv_norm, int32_t* perm,
                           uint32_t* tor, uint32_t* seg, float* class_sum, float* class_mass,
                           void* scratch, size_t scratch_bytes, void* stream_) {
  return prepare_impl(x, layout, batch, pixels_per_image, dim, cls, cls_is_int64, num_classes, eps, ehat,
                      bits, inv_norm, perm, tor, seg, class_sum, class_mass, scratch, scratch_bytes, nullptr, nullptr, nullptr,
                      static_cast<cudaStream_t>(stream_));
}

int muscle_infonce_rowsum(const void* ehat_rows, const uint32_t* bits_rows,
                          const uint32_t* tor_rows, int64_t rows_pad, const void* ehat_cols,
                          const uint32_t* bits_cols, const uint32_t* tor_cols, int64_t cols_pad,
                          const uint32_t* unit_off, int dim, float temperature, float* rowsum,
                          void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device()) return rc;
  if (!ehat_rows || !bits_rows || !tor_rows || !ehat_cols || !bits_cols || !tor_cols || !unit_off || !rowsum)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (rows_pad % 256 || cols_pad % 256 || rows_pad <= 0 || cols_pad <= 0)
    return fail(MUSCLE_ERR_BAD_ARG, "padded sizes must be positive multiples of 256");
  PairParams prm{};
  prm.bits_row = bits_rows; prm.bits_col = bits_cols; prm.tor_row = tor_rows; prm.tor_col = tor_cols;
  prm.rowsum = rowsum; prm.unit_off = unit_off;
  prm.scale_log2 = kLog2e / temperature;
  prm.mask_shift = pair_mask_shift(prm.scale_log2, poly_ok(temperature));
  prm.n_row_blocks = static_cast<int>(rows_pad / rows_per_block(dim));
  prm.n_col_tiles = static_cast<int>(cols_pad / 128);
  prm.rows_pad = static_cast<int>(rows_pad);
  return launch_pair_dim<MODE_ROWSUM>(dim, poly_ok(temperature), ehat_rows, rows_pad, ehat_cols,
                                      cols_pad, prm, stream);
}

}  // extern "C"

namespace {
int launch_mid(const void* ehat, const uint32_t* seg, int64_t rows_pad, int dim, int num_classes,
               const float* class_sum, const float* class_mass, float* rowsum, float temperature,
               float* w, float* gpos, double* loss_acc, float* dproto, float* loss_out,
               unsigned int* ticket, cudaStream_t stream) {
  if (int rc = check_device()) return rc;
  if (!ehat || !seg || !class_sum || !class_mass || !rowsum || !w || !gpos || !loss_acc || !dproto)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (num_classes < 1 || num_classes > MUSCLE_MAX_CLASSES)
    return fail(MUSCLE_ERR_BAD_ARG, "num_classes must be in [1,%d]", MUSCLE_MAX_CLASSES);
  if (rows_pad <= 0 || rows_pad % 128) return fail(MUSCLE_ERR_BAD_ARG, "rows_pad must be a positive multiple of 128");
  MidParams p;
  p.ehat = static_cast<const __nv_bfloat16*>(ehat); p.seg = seg; p.C = num_classes;
  p.class_sum = class_sum; p.class_mass = class_mass; p.rowsum = rowsum;
  p.inv_tau = 1.f / temperature; p.w = w; p.gpos = gpos; p.loss_acc = loss_acc; p.dproto = dproto;
  p.loss_out = loss_out; p.ticket = ticket;
  const int grid = static_cast<int>(rows_pad / SEG_ROWS);
  if (dim == 64) MB_CUDA(launch_pdl(mid_kernel<64>, dim3(grid), dim3(SEG_THREADS), 0, stream, p));
  else if (dim == 128) MB_CUDA(launch_pdl(mid_kernel<128>, dim3(grid), dim3(SEG_THREADS), 0, stream, p));
  else MB_CUDA(launch_pdl(mid_kernel<256>, dim3(grid), dim3(SEG_THREADS), 0, stream, p));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
}  // namespace

extern "C" {

int muscle_infonce_mid(const void* ehat, const uint32_t* seg, int64_t rows_pad, int dim,
                       int num_classes, const float* class_sum, const float* class_mass,
                       float* rowsum, float temperature, float* w, float* gpos, double* loss_acc,
                       float* dproto, void* stream_) {
  return launch_mid(ehat, seg, rows_pad, dim, num_classes, class_sum, class_mass, rowsum, temperature, w,
                    gpos, loss_acc, dproto, nullptr, nullptr, static_cast<cudaStream_t>(stream_));
}

int muscle_infonce_part_slots(void) { return PAIR_SLOTS; }

int muscle_infonce_plan(const uint32_t* tor_rows, int64_t rows_pad, const uint32_t* tor_cols,
                        int64_t cols_pad, int dim, uint32_t* unit_off, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device()) return rc;
  if (!tor_rows || !tor_cols || !unit_off) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (rows_pad % 256 || cols_pad % 256 || rows_pad <= 0 || cols_pad <= 0)
    return fail(MUSCLE_ERR_BAD_ARG, "padded sizes must be positive multiples of 256");
  const int n_row_blocks = static_cast<int>(rows_pad / rows_per_block(dim));
  const int n_col_tiles = static_cast<int>(cols_pad / 128);
  if (static_cast<double>(n_row_blocks) * n_col_tiles >= 4.0e9)
    return fail(MUSCLE_ERR_BAD_ARG, "too many tile units for one rank");
  MB_CUDA(launch_pdl(pair_plan_kernel<1>, dim3(1), dim3(1024), 0, stream, tor_rows, n_row_blocks, tor_cols, n_col_tiles, unit_off));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_infonce_plan_sym(const uint32_t* tor, int64_t rows_pad, uint32_t* unit_off_sym, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device()) return rc;
  if (!tor || !unit_off_sym) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (rows_pad % 256 || rows_pad <= 0) return fail(MUSCLE_ERR_BAD_ARG, "padded size must be a positive multiple of 256");
  const int n_tiles = static_cast<int>(rows_pad / 128);
  if (static_cast<double>(n_tiles) * n_tiles >= 8.0e9) return fail(MUSCLE_ERR_BAD_ARG, "too many tile units for one rank");
  MB_CUDA(launch_pdl(pair_plan_sym_kernel, dim3(1), dim3(1024), 0, stream, tor, n_tiles, unit_off_sym));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_infonce_rowsum_sym(const void* ehat, const uint32_t* bits, const uint32_t* tor, int64_t rows_pad,
                              const uint32_t* unit_off_sym, int dim, float temperature, float* rowsum,
                              void* stream_) {
  if (int rc = check_device()) return rc;
  if (!ehat || !bits || !tor || !unit_off_sym || !rowsum) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (dim != 64 && dim != 128) return fail(MUSCLE_ERR_BAD_ARG, "symmetric row-sum pass: dim must be 64 or 128 (got %d)", dim);
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (rows_pad % 256 || rows_pad <= 0) return fail(MUSCLE_ERR_BAD_ARG, "padded size must be a positive multiple of 256");
  return rowsum_sym(ehat, bits, tor, rows_pad, ehat, bits, tor, rows_pad, unit_off_sym, 0, 1, dim, temperature, rowsum,
                    static_cast<cudaStream_t>(stream_));
}

size_t muscle_infonce_plan_sym_ring_scratch_bytes(int64_t rows_pad) { return static_cast<size_t>(rows_pad / 128 + 1) * 4; }

int muscle_infonce_plan_sym_ring(const uint32_t* tor_rows, int64_t rows_pad, const uint32_t* tor_cols, int64_t cols_pad,
                                 int rank, int world, void* scratch, uint32_t* unit_off_sym, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device()) return rc;
  if (!tor_rows || !tor_cols || !scratch || !unit_off_sym) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (rows_pad % 256 || rows_pad <= 0 || world < 1 || rank < 0 || rank >= world || cols_pad != rows_pad * world)
    return fail(MUSCLE_ERR_BAD_ARG, "every rank holds rows_pad rows (a multiple of 256) and cols_pad = world * rows_pad");
  const int n_rb = static_cast<int>(rows_pad / 128), n_ct = static_cast<int>(cols_pad / 128);
  uint32_t* cnt = static_cast<uint32_t*>(scratch);
  MB_CUDA(launch_pdl(pair_plan_sym_count_kernel, dim3((n_rb + 7) / 8), dim3(256), 0, stream, tor_rows, n_rb, tor_cols, n_ct,
                     rank, world, cnt));
  MB_LAUNCH_CHECK();
  MB_CUDA(launch_pdl(pair_plan_scan_kernel, dim3(1), dim3(1024), 0, stream, static_cast<const uint32_t*>(cnt), n_rb, unit_off_sym));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_infonce_plan_sym_ring_static_max_rows(void) { return kStaticPlanMaxTiles * 128; }

int muscle_infonce_plan_sym_ring_static(const uint32_t* tor_rows, int64_t rows_pad, const uint32_t* tor_cols,
                                        int64_t cols_pad, int rank, int world, uint32_t* unit_off_sym, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device()) return rc;
  if (!tor_rows || !tor_cols || !unit_off_sym) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (rows_pad % 256 || rows_pad <= 0 || world < 1 || rank < 0 || rank >= world || cols_pad != rows_pad * world)
    return fail(MUSCLE_ERR_BAD_ARG, "every rank holds rows_pad rows (a multiple of 256) and cols_pad = world * rows_pad");
  if (rows_pad > static_cast<int64_t>(kStaticPlanMaxTiles) * 128)
    return fail(MUSCLE_ERR_BAD_ARG,
                "muscle_infonce_plan_sym_ring_static: rows_pad %lld exceeds the built-in scratch (at most %d rows); use "
                "muscle_infonce_plan_sym_ring with caller-owned scratch",
                static_cast<long long>(rows_pad), kStaticPlanMaxTiles * 128);
  // device address of the scratch array, looked up once per device with that device current
  constexpr int kMaxDevices = 64;
  static std::atomic<uint32_t*> addr[kMaxDevices] = {};
  int dev = -1;
  MB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(MUSCLE_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  uint32_t* cnt = addr[dev].load(std::memory_order_acquire);
  if (cnt == nullptr) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_plan_cnt);
    if (e != cudaSuccess || ptr == nullptr)
      return fail(MUSCLE_ERR_CUDA, "cudaGetSymbolAddress(g_static_plan_cnt) failed on device %d: %s", dev, cudaGetErrorString(e));
    cnt = static_cast<uint32_t*>(ptr);
    addr[dev].store(cnt, std::memory_order_release);
  }
  // cnt[0, n_rb) is written in full by the count kernel before the scan reads it: nothing has to start at zero, and what
  // the previous call left behind is never read
  const int n_rb = static_cast<int>(rows_pad / 128), n_ct = static_cast<int>(cols_pad / 128);
  MB_CUDA(launch_pdl(pair_plan_sym_count_kernel, dim3((n_rb + 7) / 8), dim3(256), 0, stream, tor_rows, n_rb, tor_cols, n_ct,
                     rank, world, cnt));
  MB_LAUNCH_CHECK();
  MB_CUDA(launch_pdl(pair_plan_scan_kernel, dim3(1), dim3(1024), 0, stream, static_cast<const uint32_t*>(cnt), n_rb, unit_off_sym));
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}

int muscle_infonce_rowsum_sym_ring(const void* ehat_rows, const uint32_t* bits_rows, const uint32_t* tor_rows,
                                   int64_t rows_pad, const void* ehat_cols, const uint32_t* bits_cols,
                                   const uint32_t* tor_cols, int64_t cols_pad, const uint32_t* unit_off_sym, int rank,
                                   int world, int dim, float temperature, float* rowsum_all, void* stream_) {
  if (int rc = check_device()) return rc;
  if (!ehat_rows || !bits_rows || !tor_rows || !ehat_cols || !bits_cols || !tor_cols || !unit_off_sym || !rowsum_all)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (dim != 64 && dim != 128) return fail(MUSCLE_ERR_BAD_ARG, "symmetric row-sum pass: dim must be 64 or 128 (got %d)", dim);
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (rows_pad % 256 || rows_pad <= 0 || world < 1 || rank < 0 || rank >= world || cols_pad != rows_pad * world)
    return fail(MUSCLE_ERR_BAD_ARG, "every rank holds rows_pad rows (a multiple of 256) and cols_pad = world * rows_pad");
  return rowsum_sym(ehat_rows, bits_rows, tor_rows, rows_pad, ehat_cols, bits_cols, tor_cols, cols_pad, unit_off_sym, rank,
                    world, dim, temperature, rowsum_all, static_cast<cudaStream_t>(stream_));
}

int muscle_infonce_rowsum_sym_share(const void* ehat_all, const uint32_t* bits_all, const uint32_t* tor_all,
                                    int64_t cols_pad, const uint32_t* unit_off_sym, int share, int n_shares, int dim,
                                    float temperature, float* rowsum_all, void* stream_) {
  if (int rc = check_device()) return rc;
  if (!ehat_all || !bits_all || !tor_all || !unit_off_sym || !rowsum_all) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (dim != 64 && dim != 128) return fail(MUSCLE_ERR_BAD_ARG, "symmetric row-sum pass: dim must be 64 or 128 (got %d)", dim);
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (cols_pad % 256 || cols_pad <= 0 || n_shares < 1 || share < 0 || share >= n_shares)
    return fail(MUSCLE_ERR_BAD_ARG, "cols_pad must be a positive multiple of 256 and 0 <= share < n_shares");
  return rowsum_sym(ehat_all, bits_all, tor_all, cols_pad, ehat_all, bits_all, tor_all, cols_pad, unit_off_sym, 0, 1, dim,
                    temperature, rowsum_all, static_cast<cudaStream_t>(stream_), share, n_shares);
}

int muscle_infonce_grad(const void* ehat_rows, const uint32_t* bits_rows,
                        const uint32_t* tor_rows, const float* w_rows, int64_t rows_pad,
                        const void* ehat_cols, const uint32_t* bits_cols,
                        const uint32_t* tor_cols, const float* w_cols, int64_t cols_pad,
                        const uint32_t* unit_off, int dim, float temperature, float* part,
                        void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_device()) return rc;
  if (!ehat_rows || !bits_rows || !tor_rows || !w_rows || !ehat_cols || !bits_cols || !tor_cols ||
      !w_cols || !unit_off || !part)
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (rows_pad % 256 || cols_pad % 256 || rows_pad <= 0 || cols_pad <= 0)
    return fail(MUSCLE_ERR_BAD_ARG, "padded sizes must be positive multiples of 256");
  PairParams prm{};
  prm.bits_row = bits_rows; prm.bits_col = bits_cols; prm.tor_row = tor_rows; prm.tor_col = tor_cols;
  prm.w_row = w_rows; prm.w_col = w_cols; prm.part = part; prm.unit_off = unit_off;
  prm.scale_log2 = kLog2e / temperature;
  prm.mask_shift = pair_mask_shift(prm.scale_log2, poly_ok(temperature));
  prm.n_row_blocks = static_cast<int>(rows_pad / rows_per_block(dim));
  prm.n_col_tiles = static_cast<int>(cols_pad / 128);
  prm.rows_pad = static_cast<int>(rows_pad);
  return launch_pair_dim<MODE_GRAD>(dim, poly_ok(temperature), ehat_rows, rows_pad, ehat_cols,
                                    cols_pad, prm, stream);
}

}  // extern "C"

namespace {
int final_impl(const void* ehat, const uint32_t* bits, const float* inv_norm,
               const int32_t* perm, const float* part, const uint32_t* unit_off,
               int64_t rows_pad, int64_t cols_pad, int64_t batch, int64_t pixels_per_image, int dim, int num_classes,
               const float* class_sum, const float* class_mass, const float* gpos,
               const float* dproto, const double* loss_acc, float temperature,
               int layout, float* dx, float* loss_out, const float* grad_scale, int proto_external,
               float* dproto_out, cudaStream_t stream) {
  if (int rc = check_device()) return rc;
  if (!class_mass || !loss_acc) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (dx && (!ehat || !bits || !inv_norm || !perm || !part || !unit_off || !class_sum || !gpos || !dproto))
    return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (layout != MUSCLE_LAYOUT_PIXEL_MAJOR && layout != MUSCLE_LAYOUT_NCHW)
    return fail(MUSCLE_ERR_BAD_ARG, "unknown layout %d", layout);
  FinalParams p;
  p.ehat = static_cast<const __nv_bfloat16*>(ehat); p.bits = bits; p.inv_norm = inv_norm;
  p.perm = perm; p.part = part; p.unit_off = unit_off; p.rows_pad = rows_pad;
  p.n_row_blocks = static_cast<int>(rows_pad / rows_per_block(dim));
  p.n_col_tiles = static_cast<int>(cols_pad / 128); p.n_ctas = pair_grid();
  p.n = batch * pixels_per_image; p.ppi = pixels_per_image; p.C = num_classes;
  p.class_sum = class_sum; p.class_mass = class_mass; p.gpos = gpos; p.dproto = dproto;
  p.loss_acc = loss_acc; p.inv_tau = 1.f / temperature; p.layout = layout; p.dx = dx;
  p.loss_out = loss_out; p.grad_scale = grad_scale; p.proto_external = proto_external; p.dproto_out = dproto_out;
  const size_t smem = (dx && layout == MUSCLE_LAYOUT_NCHW)
                          ? static_cast<size_t>(FINAL_CHUNK) * (dim + 1) * sizeof(float) : 0;
  const int grid = dx ? static_cast<int>((p.n + FINAL_CHUNK - 1) / FINAL_CHUNK) : 1;
#define MB_FINAL(DD)                                                                            \
  {                                                                                             \
    if (int rca = ensure_dynamic_smem<final_kernel<DD>>(static_cast<int>(FINAL_CHUNK * (DD + 1) * sizeof(float)))) return rca; \
    MB_CUDA(launch_pdl(final_kernel<DD>, dim3(grid), dim3(FINAL_THREADS), smem, stream, p));     \
  }
  if (dim == 64) MB_FINAL(64) else if (dim == 128) MB_FINAL(128) else MB_FINAL(256)
#undef MB_FINAL
  MB_LAUNCH_CHECK();
  return MUSCLE_OK;
}
}  // namespace

extern "C" {

int muscle_infonce_final(const void* ehat, const uint32_t* bits, const float* inv_norm,
                         const int32_t* perm, const float* part, const uint32_t* unit_off,
                         int64_t rows_pad, int64_t cols_pad, int64_t batch, int64_t pixels_per_image, int dim, int num_classes,
                         const float* class_sum, const float* class_mass, const float* gpos,
                         const float* dproto, const double* loss_acc, float temperature,
                         int layout, float* dx, float* loss_out, const float* grad_scale, void* stream_) {
  return final_impl(ehat, bits, inv_norm, perm, part, unit_off, rows_pad, cols_pad, batch, pixels_per_image, dim,
                    num_classes, class_sum, class_mass, gpos, dproto, loss_acc, temperature, layout, dx, loss_out,
                    grad_scale, 0, nullptr, static_cast<cudaStream_t>(stream_));
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Single-GPU composite: the whole op on one stream, no host synchronisation.
namespace {

struct Workspace {
  size_t ehat, bits, inv_norm, perm, tor, unit_off, unit_off_sym, rowsum, w, gpos, small, scratch, part, total;
  size_t seg, class_sum, class_mass, dproto, loss_acc, ticket, small_bytes;  // offsets inside `small`
  size_t zero_lo, zero_bytes;  // rowsum .. small are zeroed with one memset
  int64_t rows_pad;
};

Workspace plan_workspace(int64_t n, int dim, int C) {
  Workspace ws{};
  ws.rows_pad = muscle_pad_rows(n);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  const size_t rp = static_cast<size_t>(ws.rows_pad);
  ws.ehat = take(rp * dim * 2);
  ws.bits = take(rp * 4);
  ws.inv_norm = take(rp * 4);
  ws.perm = take(rp * 4);
  ws.tor = take(rp / 128 * 4);
  ws.unit_off = take((rp / 128 + 1) * 4);
  ws.unit_off_sym = take((rp / 128 + 1) * 4);
  ws.w = take(rp * 4);
  ws.gpos = take(rp * 4);
  ws.rowsum = take(rp * 4);
  size_t s = 0;
  auto take_small = [&](size_t bytes) { size_t o = s; s = align_up(s + bytes, 16); return o; };
  ws.class_sum = take_small(static_cast<size_t>(C) * dim * 4);
  ws.class_mass = take_small(static_cast<size_t>(C) * 4);
  ws.dproto = take_small(static_cast<size_t>(C) * dim * 4);
  ws.loss_acc = take_small(8);
  ws.ticket = take_small(4);
  ws.seg = take_small(static_cast<size_t>(C + 2) * 4);
  ws.small_bytes = s;
  ws.small = take(s);
  ws.zero_lo = ws.rowsum;
  ws.zero_bytes = off - ws.rowsum;
  ws.scratch = take(muscle_infonce_prepare_scratch_bytes(n));
  ws.part = take(static_cast<size_t>(PAIR_SLOTS) * rp * dim * 4);
  ws.total = off;
  return ws;
}

}  // namespace

extern "C" {

size_t muscle_dense_infonce_workspace_bytes(int64_t n_pixels, int dim, int num_classes) {
  if (!dim_ok(dim) || n_pixels < 0 || num_classes < 1 || num_classes > MUSCLE_MAX_CLASSES) return 0;
  return plan_workspace(n_pixels, dim, num_classes).total;
}

namespace {
// the composite's view of its workspace
struct Carved {
  void* ehat; uint32_t* bits; float* inv_norm; int32_t* perm; uint32_t* tor; uint32_t* unit_off; uint32_t* unit_off_sym;
  float *rowsum, *w, *gpos, *class_sum, *class_mass, *dproto, *part;
  double* loss_acc; uint32_t* seg; char* scratch; unsigned int* ticket;
};
Carved carve(void* workspace, const Workspace& ws) {
  char* base = static_cast<char*>(workspace);
  char* small = base + ws.small;
  Carved c;
  c.ehat = base + ws.ehat;
  c.bits = reinterpret_cast<uint32_t*>(base + ws.bits);
  c.inv_norm = reinterpret_cast<float*>(base + ws.inv_norm);
  c.perm = reinterpret_cast<int32_t*>(base + ws.perm);
  c.tor = reinterpret_cast<uint32_t*>(base + ws.tor);
  c.unit_off = reinterpret_cast<uint32_t*>(base + ws.unit_off);
  c.unit_off_sym = reinterpret_cast<uint32_t*>(base + ws.unit_off_sym);
  c.rowsum = reinterpret_cast<float*>(base + ws.rowsum);
  c.w = reinterpret_cast<float*>(base + ws.w);
  c.gpos = reinterpret_cast<float*>(base + ws.gpos);
  c.class_sum = reinterpret_cast<float*>(small + ws.class_sum);
  c.class_mass = reinterpret_cast<float*>(small + ws.class_mass);
  c.dproto = reinterpret_cast<float*>(small + ws.dproto);
  c.loss_acc = reinterpret_cast<double*>(small + ws.loss_acc);
  c.ticket = reinterpret_cast<unsigned int*>(small + ws.ticket);
  c.seg = reinterpret_cast<uint32_t*>(small + ws.seg);
  c.part = reinterpret_cast<float*>(base + ws.part);
  c.scratch = base + ws.scratch;
  return c;
}

int check_composite(int dim, int num_classes, float temperature, int64_t batch, int64_t pixels_per_image,
                    const void* workspace, size_t workspace_bytes, Workspace* ws) {
  if (int rc = check_device()) return rc;
  if (!workspace) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  if (!dim_ok(dim)) return fail(MUSCLE_ERR_BAD_ARG, "dim must be 64, 128 or 256 (got %d)", dim);
  if (num_classes < 1 || num_classes > MUSCLE_MAX_CLASSES)
    return fail(MUSCLE_ERR_BAD_ARG, "num_classes must be in [1,%d]", MUSCLE_MAX_CLASSES);
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (batch <= 0 || pixels_per_image <= 0) return fail(MUSCLE_ERR_BAD_ARG, "empty input");
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return fail(MUSCLE_ERR_BAD_ARG, "workspace must be 256-byte aligned");
  *ws = plan_workspace(batch * pixels_per_image, dim, num_classes);
  if (workspace_bytes < ws->total)
    return fail(MUSCLE_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes, ws->total);
  return MUSCLE_OK;
}

// MUSCLE_STAGE_TIMING=1: CUDA events around every stage of the composite (debugging aid; the event
// records break the programmatic-launch chain, so the sum is a little above the real step)
struct StageTimer {
  bool on;
  cudaStream_t stream;
  cudaEvent_t ev[16];
  const char* name[16];
  int n = 0;
  explicit StageTimer(cudaStream_t s) : stream(s) {
    static const bool env = getenv("MUSCLE_STAGE_TIMING") != nullptr;
    on = env;
  }
  void mark(const char* what) {
    if (!on || n >= 16) return;
    cudaEventCreate(&ev[n]);
    cudaEventRecord(ev[n], stream);
    name[n++] = what;
  }
  void report(const char* title) {
    if (!on || n < 2) return;
    cudaEventSynchronize(ev[n - 1]);
    fprintf(stderr, "[stages %s]", title);
    for (int i = 1; i < n; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[i - 1], ev[i]);
      fprintf(stderr, " %s %.1f us |", name[i], ms * 1e3f);
    }
    fprintf(stderr, "\n");
    for (int i = 0; i < n; ++i) cudaEventDestroy(ev[i]);
  }
};

// prepare -> plan -> row sums -> mid: everything the loss needs
int composite_forward(const float* x, int layout, int64_t batch, int64_t pixels_per_image, int dim,
                      const void* cls, int cls_is_int64, int num_classes, float temperature, float eps,
                      float* loss_out, void* workspace, const Workspace& ws, const Carved& c,
                      const float* proto_in, cudaStream_t stream) {
  const int64_t n = batch * pixels_per_image;
  StageTimer tm(stream);
  tm.mark("start");
  MB_CUDA(cudaMemsetAsync(static_cast<char*>(workspace) + ws.zero_lo, 0, ws.zero_bytes, stream));
  if (static_cast<double>(ws.rows_pad / 128) * static_cast<double>(ws.rows_pad / 128) >= 4.0e9)
    return fail(MUSCLE_ERR_BAD_ARG, "too many tile units for one rank");
  // (the work partition `unit_off` comes out of the scan kernel: no plan launch)
  int rc = prepare_impl(x, layout, batch, pixels_per_image, dim, cls, cls_is_int64, num_classes, eps,
                        c.ehat, c.bits, c.inv_norm, c.perm, c.tor, c.seg, c.class_sum, c.class_mass,
                        c.scratch, muscle_infonce_prepare_scratch_bytes(n), c.unit_off,
                        rowsum_sym_enabled(dim) ? c.unit_off_sym : nullptr, proto_in, stream);
  if (rc) return rc;
  tm.mark("memset+hist+scan(+plan)+prepare+class_sum");
  if (!(temperature >= kMinTemperature))
    return fail(MUSCLE_ERR_BAD_ARG, "temperature must be >= %g (exp((s-1)/T) underflows fp32 below it)", kMinTemperature);
  if (rowsum_sym_enabled(dim))   // rows == columns: the symmetric pass (half the products and exponentials)
    rc = rowsum_sym(c.ehat, c.bits, c.tor, ws.rows_pad, c.ehat, c.bits, c.tor, ws.rows_pad, c.unit_off_sym, 0, 1, dim,
                    temperature, c.rowsum, stream);
  else
    rc = muscle_infonce_rowsum(c.ehat, c.bits, c.tor, ws.rows_pad, c.ehat, c.bits, c.tor, ws.rows_pad,
                               c.unit_off, dim, temperature, c.rowsum, stream);
  if (rc) return rc;
  tm.mark("rowsum");
  rc = launch_mid(c.ehat, c.seg, ws.rows_pad, dim, num_classes, c.class_sum, c.class_mass, c.rowsum,
                  temperature, c.w, c.gpos, c.loss_acc, c.dproto, loss_out, c.ticket, stream);
  tm.mark("mid (+ loss)");
  tm.report("forward");
  return rc;
}

// gradient pass + final (dx may be null: loss only)
int composite_backward(int layout, int64_t batch, int64_t pixels_per_image, int dim, int num_classes,
                       float temperature, const float* grad_scale, float* dx_out, float* loss_out,
                       const Workspace& ws, const Carved& c, int proto_external, float* dproto_out,
                       cudaStream_t stream) {
  StageTimer tm(stream);
  tm.mark("start");
  if (dx_out) {
    int rc = muscle_infonce_grad(c.ehat, c.bits, c.tor, c.w, ws.rows_pad, c.ehat, c.bits, c.tor, c.w,
                                 ws.rows_pad, c.unit_off, dim, temperature, c.part, stream);
    if (rc) return rc;
    tm.mark("grad");
  }
  const int rcf = final_impl(c.ehat, c.bits, c.inv_norm, c.perm, c.part, c.unit_off, ws.rows_pad,
                             ws.rows_pad, batch, pixels_per_image, dim, num_classes, c.class_sum,
                             c.class_mass, c.gpos, c.dproto, c.loss_acc, temperature, layout, dx_out,
                             loss_out, grad_scale, proto_external, dproto_out, stream);
  tm.mark(dx_out ? "final" : "final (loss only)");
  tm.report(dx_out ? "backward" : "loss");
  return rcf;
}
}  // namespace

int muscle_dense_infonce_proto_fwd(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                                   int dim, const void* cls, int cls_is_int64, int num_classes,
                                   float temperature, float eps, const float* proto_in, float* loss_out,
                                   void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!x || !cls || !loss_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  Workspace ws;
  if (int rc = check_composite(dim, num_classes, temperature, batch, pixels_per_image, workspace, workspace_bytes, &ws))
    return rc;
  const Carved c = carve(workspace, ws);
  // the loss is written by the last block of mid_kernel: no separate launch for it
  return composite_forward(x, layout, batch, pixels_per_image, dim, cls, cls_is_int64, num_classes,
                           temperature, eps, loss_out, workspace, ws, c, proto_in, stream);
}

int muscle_dense_infonce_proto_bwd(int layout, int64_t batch, int64_t pixels_per_image, int dim,
                                   int num_classes, float temperature, const float* grad_scale,
                                   int proto_external, float* dx_out, float* dproto_out, void* workspace,
                                   size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!dx_out) return fail(MUSCLE_ERR_BAD_ARG, "null pointer argument");
  Workspace ws;
  if (int rc = check_composite(dim, num_classes, temperature, batch, pixels_per_image, workspace, workspace_bytes, &ws))
    return rc;
  const Carved c = carve(workspace, ws);
  return composite_backward(layout, batch, pixels_per_image, dim, num_classes, temperature, grad_scale, dx_out,
                            nullptr, ws, c, proto_external, dproto_out, stream);
}

int muscle_dense_infonce_fwd_bwd(const float* x, int layout, int64_t batch,
                                 int64_t pixels_per_image, int dim, const void* cls,
                                 int cls_is_int64, int num_classes, float temperature, float eps,
                                 float* loss_out, float* dx_out, void* workspace,
                                 size_t workspace_bytes, void* stream_) {
  if (int rc = muscle_dense_infonce_proto_fwd(x, layout, batch, pixels_per_image, dim, cls, cls_is_int64, num_classes,
                                              temperature, eps, nullptr, loss_out, workspace, workspace_bytes, stream_))
    return rc;
  if (!dx_out) return MUSCLE_OK;
  return muscle_dense_infonce_proto_bwd(layout, batch, pixels_per_image, dim, num_classes, temperature, nullptr, 0,
                                        dx_out, nullptr, workspace, workspace_bytes, stream_);
}

int muscle_dense_infonce_fwd(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                             int dim, const void* cls, int cls_is_int64, int num_classes,
                             float temperature, float eps, float* loss_out, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  return muscle_dense_infonce_proto_fwd(x, layout, batch, pixels_per_image, dim, cls, cls_is_int64, num_classes,
                                        temperature, eps, nullptr, loss_out, workspace, workspace_bytes, stream_);
}

int muscle_dense_infonce_bwd(int layout, int64_t batch, int64_t pixels_per_image, int dim,
                             int num_classes, float temperature, const float* grad_scale,
                             float* dx_out, void* workspace, size_t workspace_bytes, void* stream_) {
  return muscle_dense_infonce_proto_bwd(layout, batch, pixels_per_image, dim, num_classes, temperature, grad_scale, 0,
                                        dx_out, nullptr, workspace, workspace_bytes, stream_);
}

}  // extern "C"
