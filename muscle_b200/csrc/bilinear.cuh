// Bilinear re-sampling taps with align_corners=True, exactly as ATen's upsample_bilinear2d computes them
// (F.interpolate as called at src/MuSCLe.py:256-257 and :285): source index = o * (n_in - 1) / (n_out - 1).
#pragma once

namespace mb200 {

struct Tap { int i0, i1; float w0, w1; };
__device__ __forceinline__ Tap tap_of(int o, float scale, int n_in) {   // area_pixel_compute_source_index
  const float src = scale * static_cast<float>(o);
  int i0 = static_cast<int>(src);
  if (i0 > n_in - 1) i0 = n_in - 1;
  const int i1 = i0 + (i0 < n_in - 1 ? 1 : 0);
  const float l1 = src - static_cast<float>(i0);
  return Tap{i0, i1, 1.f - l1, l1};
}
inline float bilinear_scale(int n_in, int n_out) {
  return n_out > 1 ? static_cast<float>(n_in - 1) / static_cast<float>(n_out - 1) : 0.f;
}

}  // namespace mb200
