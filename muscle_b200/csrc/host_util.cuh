// Host-side helpers shared by the C-ABI translation units: error reporting, device queries and
// TMA tensor-map construction (driver entry point fetched at run time, so the library links
// against cudart only and still loads on a machine without libcuda for symbol checks).
#pragma once
#include <atomic>
#include <cstdlib>
#include <utility>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cuda.h>
#include <cuda_runtime.h>

#include "../../include/muscle_b200.h"

namespace mb200 {

inline char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define MB_CUDA(expr)                                                                        \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ::mb200::fail(MUSCLE_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                  \
                           cudaGetErrorString(_e), __FILE__, __LINE__);                      \
  } while (0)

#define MB_LAUNCH_CHECK()                                                                    \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess)                                                                   \
      return ::mb200::fail(MUSCLE_ERR_CUDA, "kernel launch failed: %s (%s:%d)",              \
                           cudaGetErrorString(_e), __FILE__, __LINE__);                      \
  } while (0)

struct DeviceInfo {
  int sm_count = 0;
  int cc_major = 0;
  bool ok = false;
};
inline const DeviceInfo& device_info() {
  static thread_local DeviceInfo info[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  DeviceInfo& d = info[dev];
  if (!d.ok) {
    cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    d.ok = d.sm_count > 0;
  }
  return d;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  // cuTensorMapEncodeTiled is a driver call and needs the primary context current on THIS thread;
  // a thread that has only been handed device pointers (autograd's backward thread) has none until
  // its first runtime call that touches the device (CUDA_ERROR_INVALID_CONTEXT otherwise).
  static thread_local bool bound = false;
  if (!bound) {
    cudaFree(nullptr);
    bound = true;
  }
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 [rows, cols] row-major -> boxes of [box_rows rows x 64 elements], 128-B swizzle.
inline int make_bf16_box_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(MUSCLE_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 127) != 0)
    return fail(MUSCLE_ERR_BAD_ARG, "bf16 operand must be 128-byte aligned");
  if (cols % 8 != 0) return fail(MUSCLE_ERR_BAD_ARG, "bf16 operand rows must be multiples of 16 bytes");
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(cols) * 2};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(MUSCLE_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return MUSCLE_OK;
}

// bf16 [rows, dim] row-major -> boxes of [128 rows x 64 elements], 128-B swizzle.
inline int make_bf16_tile_map(CUtensorMap* map, const void* base, int64_t rows, int dim) {
  return make_bf16_box_map(map, base, rows, dim, 128);
}

// bf16 [rows, dim] row-major seen as [dim/64 panels][rows][64]: ONE copy brings a whole 128-row tile as
// dim/64 consecutive [128 x 64] 128-B-swizzled panels (coordinates: 0, row, 0).
inline int make_bf16_panel_map(CUtensorMap* map, const void* base, int64_t rows, int dim, int box_rows = 128) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(MUSCLE_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 127) != 0)
    return fail(MUSCLE_ERR_BAD_ARG, "bf16 operand must be 128-byte aligned");
  if (dim % 64 != 0) return fail(MUSCLE_ERR_BAD_ARG, "bf16 operand rows must be multiples of 64 elements");
  cuuint64_t gdim[3] = {64, static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(dim / 64)};
  cuuint64_t gstride[2] = {static_cast<cuuint64_t>(dim) * 2, 128};
  cuuint32_t box[3] = {64, static_cast<cuuint32_t>(box_rows), static_cast<cuuint32_t>(dim / 64)};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstride,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(MUSCLE_ERR_CUDA, "cuTensorMapEncodeTiled (panel map) failed with CUresult %d", (int)r);
  return MUSCLE_OK;
}

// Launch with programmatic stream serialization (see ptx.cuh: pdl_wait / pdl_trigger).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  static const bool off = getenv("MUSCLE_NO_PDL") != nullptr;   // debugging: plain stream order
  cfg.attrs = attr;
  cfg.numAttrs = off ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device property of a kernel: set it once per
// (kernel instantiation, device) -- not once per thread, which left a thread's second GPU without it.
template <auto Kern>
inline int ensure_dynamic_smem(int bytes) {
  static std::atomic<unsigned long long> done{0ull};   // bit d: set on device d
  int dev = 0;
  cudaGetDevice(&dev);
  const bool tracked = dev >= 0 && dev < 64;
  if (!tracked || !((done.load(std::memory_order_acquire) >> dev) & 1ull)) {
    MB_CUDA(cudaFuncSetAttribute(Kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    if (tracked) done.fetch_or(1ull << dev, std::memory_order_release);
  }
  return MUSCLE_OK;
}

inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

}  // namespace mb200
