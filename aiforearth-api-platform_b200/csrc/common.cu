// Error plumbing and tensor-map encoders shared by every kernel file.
#include "common.cuh"

#include <cstdlib>
#include <cstring>
#include <mutex>

namespace ai4e {

static thread_local char g_err[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_last_error() { return g_err; }

bool pdl_enabled() {
  static const bool on = [] {
    const char* v = getenv("AI4E_B200_NO_PDL");
    return !(v && v[0] == '1');
  }();
  return on;
}

// ----------------------------------------------------------------------------------------
// cuTensorMapEncode* resolved through the runtime so the library has no link-time
// dependency on libcuda.so (it is absent on build machines without a driver).
// ----------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                   cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

static EncodeTiledFn g_encode_tiled = nullptr;
static EncodeIm2colFn g_encode_im2col = nullptr;
static int g_driver_version = 0;
static std::once_flag g_once;

static void resolve_driver_entry_points() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    g_encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
  cudaDriverGetVersion(&g_driver_version);
}

int tmap_encode_tiled(CUtensorMap* out, const void* gaddr, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box,
                      const uint32_t* elem_strides, CUtensorMapSwizzle swizzle) {
  return tmap_encode_tiled_as(out, gaddr, rank, dims, strides_bytes, box, elem_strides, swizzle, /*u32=*/false);
}

// same with 32-bit elements (raw bytes moved four at a time: uint8 image rows)
int tmap_encode_tiled_as(CUtensorMap* out, const void* gaddr, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box,
                         const uint32_t* elem_strides, CUtensorMapSwizzle swizzle, bool u32) {
  std::call_once(g_once, resolve_driver_entry_points);
  if (!g_encode_tiled) {
    set_last_error("cuTensorMapEncodeTiled is not available from this driver");
    return AI4E_ERR_UNSUPPORTED;
  }
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = elem_strides[i]; }
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
  CUresult r = g_encode_tiled(out, u32 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                              const_cast<void*>(gaddr), d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu)", (int)r,
                   rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 1));
    return AI4E_ERR_CUDA;
  }
  return AI4E_OK;
}

// general im2col-mode map (explicit bounding-box corners, traversal strides and swizzle)
int tmap_encode_im2col(CUtensorMap* out, const void* gaddr, int n, int h, int w, int c, const int lower[2],
                       const int upper[2], const int stride_wh[2], int channels_per_pixel,
                       int pixels_per_column, CUtensorMapSwizzle swizzle) {
  std::call_once(g_once, resolve_driver_entry_points);
  if (!g_encode_im2col) {
    set_last_error("cuTensorMapEncodeIm2col is not available from this driver");
    return AI4E_ERR_UNSUPPORTED;
  }
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)w * c * 2, (cuuint64_t)h * w * c * 2};
  int lo[2] = {lower[0], lower[1]}, up[2] = {upper[0], upper[1]};
  cuuint32_t es[4] = {1, (cuuint32_t)stride_wh[0], (cuuint32_t)stride_wh[1], 1};
  CUresult r = g_encode_im2col(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(gaddr), dims, strides,
                               lo, up, (cuuint32_t)channels_per_pixel, (cuuint32_t)pixels_per_column, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeIm2col failed with CUresult %d (N%d H%d W%d C%d, corners [%d,%d]/[%d,%d], "
                   "stride %dx%d)", (int)r, n, h, w, c, lo[0], lo[1], up[0], up[1], stride_wh[0], stride_wh[1]);
    return AI4E_ERR_CUDA;
  }
  // Drivers up to CUDA 13.1 set a descriptor bit that breaks im2col loads from tensors
  // smaller than 128 KiB; CUTLASS (cute/atom/copy_traits_sm90_im2col.hpp) clears bit 21 of
  // the second 64-bit word in that case, and so do we.
  if (g_driver_version <= 13010 && (uint64_t)n * h * w * c * 2 < 131072ull)
    reinterpret_cast<uint64_t*>(out)[1] &= ~(1ull << 21);
  return AI4E_OK;
}

// im2col-mode map over an NHWC bf16 activation tensor for a forward convolution.
//   dims (innermost first) {C, W, H, N}; bounding box of base pixels: lower corner = -pad,
//   upper corner = pad - (filter - 1); traversal stride = convolution stride.  A load at base
//   pixel {c, w, h, n} with offsets {s, r} then returns, for `pixels_per_column` consecutive
//   output pixels starting at ((w + pad)/stride, (h + pad)/stride, n) and wrapping over rows
//   and images, the `channels_per_pixel` channels of input pixel (w + s, h + r); out-of-image
//   taps read as zero.
int tmap_encode_im2col_nhwc(CUtensorMap* out, const void* gaddr, int n, int h, int w, int c,
                            int filt_r, int filt_s, int stride, int pad, int channels_per_pixel,
                            int pixels_per_column) {
  const int lower[2] = {-pad, -pad};
  const int upper[2] = {pad - (filt_s - 1), pad - (filt_r - 1)};
  const int st[2] = {stride, stride};
  return tmap_encode_im2col(out, gaddr, n, h, w, c, lower, upper, st, channels_per_pixel, pixels_per_column,
                            CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace ai4e
