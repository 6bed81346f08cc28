// extern "C" stateless entry points of include/ai4e_b200.h (the parity-test boundary).
#include "../../include/ai4e_b200.h"

#include "common.cuh"
#include "kernels.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace ai4e;

extern "C" {

int ai4e_version(void) { return 100; }  // 0.1.0

const char* ai4e_last_error(void) { return get_last_error(); }

int ai4e_device_count(int* out_count) {
  AI4E_REQUIRE(out_count, "device_count: null pointer");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    *out_count = 0;
    set_last_error("cudaGetDeviceCount failed: %s", cudaGetErrorString(e));
    return AI4E_ERR_CUDA;
  }
  *out_count = n;
  return AI4E_OK;
}

int ai4e_preprocess(const uint8_t* d_src, int batch, int src_h, int src_w, int channels,
                    int64_t src_image_stride, int src_row_stride, int resize_h, int resize_w,
                    int out_h, int out_w, const float* h_mean, const float* h_std, int layout,
                    int out_channels, void* d_dst_bf16, void* stream) {
  const int rc = launch_preprocess(d_src, batch, src_h, src_w, channels, src_image_stride,
                                   src_row_stride, resize_h, resize_w, out_h, out_w, h_mean, h_std,
                                   layout, out_channels, d_dst_bf16, (cudaStream_t)stream);
  return rc < 0 ? rc : AI4E_OK;
}

int ai4e_preprocess_tile(const uint8_t* d_tile, int tile_h, int tile_w, int channels, int tile_row_stride, int chip,
                         int first_chip, int n_chips, int resize_h, int resize_w, int out_h, int out_w,
                         const float* h_mean, const float* h_std, int layout, int out_channels, void* d_dst_bf16,
                         void* stream) {
  AI4E_REQUIRE(d_tile && chip > 0 && tile_h >= chip && tile_w >= chip, "preprocess_tile: tile %d x %d smaller than a %d chip",
               tile_h, tile_w, chip);
  AI4E_REQUIRE(channels == 3 || channels == 4, "preprocess_tile: channels must be 3 or 4");
  if (tile_row_stride <= 0) tile_row_stride = tile_w * channels;
  AI4E_REQUIRE(tile_row_stride >= tile_w * channels, "preprocess_tile: row stride too small");
  const int gh = tile_h / chip, gw = tile_w / chip;
  AI4E_REQUIRE(first_chip >= 0 && n_chips >= 1 && first_chip + n_chips <= gh * gw,
               "preprocess_tile: chips [%d, %d) outside the %d x %d grid", first_chip, first_chip + n_chips, gh, gw);
  ChipGrid g;
  g.grid_w = gw; g.first = first_chip;
  g.row_bytes = tile_row_stride;
  g.x_stride = (long long)chip * channels;
  g.y_stride = (long long)chip * tile_row_stride;
  const int rc = launch_preprocess(d_tile, n_chips, chip, chip, channels, 0, tile_row_stride, resize_h, resize_w, out_h, out_w,
                                   h_mean, h_std, layout, out_channels, d_dst_bf16, (cudaStream_t)stream, &g);
  return rc < 0 ? rc : AI4E_OK;
}

int ai4e_png_decode(const uint8_t* png, uint64_t len, uint8_t* h_rgb_hwc, uint64_t capacity_bytes, int32_t* out_height,
                    int32_t* out_width) {
  int h = 0, w = 0;
  const int rc = png_host_decode(png, len, h_rgb_hwc, capacity_bytes, &h, &w);
  if (out_height) *out_height = h;
  if (out_width) *out_width = w;
  return rc;
}

int ai4e_image_probe(const uint8_t* data, uint64_t len, int32_t* out_kind, int32_t* out_height, int32_t* out_width) {
  AI4E_REQUIRE(data && out_kind && out_height && out_width, "image_probe: null pointer");
  *out_kind = AI4E_IMAGE_UNKNOWN; *out_height = *out_width = 0;
  if (len >= 3 && data[0] == 0xFF && data[1] == 0xD8) {
    ai4e_jpeg_info info;
    const int rc = jpeg_host_decode(data, len, nullptr, 0, &info);
    if (rc) return rc;
    *out_kind = AI4E_IMAGE_JPEG; *out_height = info.height; *out_width = info.width;
    return AI4E_OK;
  }
  if (len >= 8 && data[0] == 0x89 && data[1] == 'P' && data[2] == 'N' && data[3] == 'G') {
    int h = 0, w = 0;
    const int rc = png_host_decode(data, len, nullptr, 0, &h, &w);
    if (rc) return rc;
    *out_kind = AI4E_IMAGE_PNG; *out_height = h; *out_width = w;
    return AI4E_OK;
  }
  set_last_error("image_probe: neither a JPEG nor a PNG stream (first bytes %02x %02x)", len > 0 ? data[0] : 0, len > 1 ? data[1] : 0);
  return AI4E_ERR_UNSUPPORTED;
}

int ai4e_conv_layer(const void* d_in_nhwc, int batch, int h, int w, int cin,
                    const void* d_weight_krsc, const float* d_bias, int cout, int kh, int kw,
                    int stride, int pad, const void* d_residual_or_null, int relu,
                    void* d_out_nhwc, int algo, void* stream) {
  AI4E_REQUIRE(d_in_nhwc && d_weight_krsc && d_bias && d_out_nhwc, "conv: null pointer");
  AI4E_REQUIRE(batch > 0 && h > 0 && w > 0 && cin > 0 && cout > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0,
               "conv: bad geometry");
  AI4E_REQUIRE(h + 2 * pad >= kh && w + 2 * pad >= kw, "conv: filter larger than padded input");
  AI4E_REQUIRE(algo == AI4E_CONV_AUTO || algo == AI4E_CONV_IGEMM || algo == AI4E_CONV_DIRECT,
               "conv: unknown algo %d", algo);
  const bool can = igemm_supported(cin, cout, kh, kw, stride, pad) != 0;
  if (algo != AI4E_CONV_DIRECT && !can && stem_supported(h, w, cin, cout, kh, kw, stride, pad)) {
    // tensor-core stem.  This stateless entry repacks the weights on every call (a test
    // convenience); the engine repacks once at creation.
    AI4E_REQUIRE(d_residual_or_null == nullptr, "conv(stem): no residual input");
    cudaStream_t s = (cudaStream_t)stream;
    std::vector<uint16_t> w_in(64 * 7 * 7 * 4), w_sp(64 * 7 * 4 * 8);
    AI4E_CUDA_TRY(cudaMemcpyAsync(w_in.data(), d_weight_krsc, w_in.size() * 2, cudaMemcpyDeviceToHost, s));
    AI4E_CUDA_TRY(cudaStreamSynchronize(s));
    stem_repack_weights(w_in.data(), w_sp.data());
    void* d_sp = nullptr;
    AI4E_CUDA_TRY(cudaMallocAsync(&d_sp, w_sp.size() * 2, s));
    AI4E_CUDA_TRY(cudaMemcpyAsync(d_sp, w_sp.data(), w_sp.size() * 2, cudaMemcpyHostToDevice, s));
    AI4E_CUDA_TRY(cudaStreamSynchronize(s));  // w_sp is a stack-lifetime host buffer
    IgemmPlan plan;
    int rc = igemm_make_plan_stem(&plan, d_in_nhwc, batch, h, w, d_sp, d_bias, relu, d_out_nhwc);
    if (rc == AI4E_OK) rc = igemm_launch(plan, batch, 0, s);
    cudaFreeAsync(d_sp, s);
    return rc;
  }
  if (algo == AI4E_CONV_DIRECT || (algo == AI4E_CONV_AUTO && !can))
    return launch_conv_direct(d_in_nhwc, batch, h, w, cin, d_weight_krsc, d_bias, cout, kh, kw, stride,
                              pad, d_residual_or_null, relu, d_out_nhwc, (cudaStream_t)stream);
  IgemmPlan plan;
  int rc = igemm_make_plan(&plan, d_in_nhwc, batch, h, w, cin, d_weight_krsc, d_bias, cout, kh, kw,
                           stride, pad, d_residual_or_null, relu, d_out_nhwc);
  if (rc) return rc;
  if (const char* trace = getenv("AI4E_B200_PIPE_TRACE")) {  // developer aid: dump CTA 0's pipeline timestamps
    long long* d = nullptr;
    AI4E_CUDA_TRY(cudaMalloc((void**)&d, 1024 * sizeof(long long)));
    AI4E_CUDA_TRY(cudaMemset(d, 0, 1024 * sizeof(long long)));
    plan.prm.dbg = d;
    rc = igemm_launch(plan, batch, 0, (cudaStream_t)stream);
    AI4E_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    std::vector<long long> h(1024);
    AI4E_CUDA_TRY(cudaMemcpy(h.data(), d, 1024 * sizeof(long long), cudaMemcpyDeviceToHost));
    cudaFree(d);
    if (FILE* f = fopen(trace, "w")) {  // columns: slot free, A issued (producer) | data seen, MMAs issued (MMA warp)
      for (int i = 0; i < 192; ++i) fprintf(f, "%d %lld %lld %lld %lld\n", i, h[i], h[192 + i], h[384 + i], h[576 + i]);
      for (int i = 0; i < 30; ++i) {  // epilogue warp 0, per 64-channel chunk: 8 stamps
        fprintf(f, "E %d", i);
        for (int j = 0; j < 8; ++j) fprintf(f, " %lld", h[768 + 8 * i + j]);
        fprintf(f, "\n");
      }
      // phases of CTA 0: entry, set-up done, previous kernel finished, MMAs issued, epilogue done, all warps done
      fprintf(f, "T %lld %lld %lld %lld %lld %lld\n", h[1008], h[1009], h[1010], h[1011], h[1012], h[1013]);
      fclose(f);
    }
    return rc;
  }
  return igemm_launch(plan, batch, 0, (cudaStream_t)stream);
}

static int stem_pool_any(const void* d_in, bool u8, const float* h_mean, const float* h_std, int batch, int h, int w,
                         const void* d_weight_krsc, const float* d_bias, int relu, void* d_out_nhwc, void* stream);

int ai4e_stem_pool(const void* d_in_nhwc4, int batch, int h, int w, const void* d_weight_krsc, const float* d_bias,
                   int relu, void* d_out_nhwc, void* stream) {
  return stem_pool_any(d_in_nhwc4, false, nullptr, nullptr, batch, h, w, d_weight_krsc, d_bias, relu, d_out_nhwc, stream);
}

int ai4e_preprocess_stem_pool(const uint8_t* d_images_hwc3, int batch, int h, int w, const float* h_mean, const float* h_std,
                              const void* d_weight_krsc, const float* d_bias, int relu, void* d_out_nhwc, void* stream) {
  AI4E_REQUIRE(h_mean && h_std, "preprocess_stem_pool: null mean/std");
  for (int c = 0; c < 3; ++c) AI4E_REQUIRE(h_std[c] != 0.0f, "preprocess_stem_pool: std[%d] is zero", c);
  return stem_pool_any(d_images_hwc3, true, h_mean, h_std, batch, h, w, d_weight_krsc, d_bias, relu, d_out_nhwc, stream);
}

static int stem_pool_any(const void* d_in_nhwc4, bool u8, const float* h_mean, const float* h_std, int batch, int h, int w,
                         const void* d_weight_krsc, const float* d_bias, int relu, void* d_out_nhwc, void* stream) {
  AI4E_REQUIRE(d_in_nhwc4 && d_weight_krsc && d_bias && d_out_nhwc && batch > 0, "stem_pool: bad arguments");
  if (!stem_pool_supported(h, w, 4, 64, 7, 7, 2, 3, 3, 2, 1)) {
    set_last_error("stem_pool: input %d x %d is not supported (width a multiple of 16 and at most 256, height a multiple of 4)", h, w);
    return AI4E_ERR_UNSUPPORTED;
  }
  // like the stem branch of ai4e_conv_layer this stateless entry repacks the weights on every call
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<uint16_t> w_in(64 * 7 * 7 * 4), w_sp((size_t)stem_pool_weight_elems());
  AI4E_CUDA_TRY(cudaMemcpyAsync(w_in.data(), d_weight_krsc, w_in.size() * 2, cudaMemcpyDeviceToHost, s));
  AI4E_CUDA_TRY(cudaStreamSynchronize(s));
  stem_pool_repack_weights(w_in.data(), w_sp.data());
  void* d_sp = nullptr;
  AI4E_CUDA_TRY(cudaMallocAsync(&d_sp, w_sp.size() * 2, s));
  AI4E_CUDA_TRY(cudaMemcpyAsync(d_sp, w_sp.data(), w_sp.size() * 2, cudaMemcpyHostToDevice, s));
  AI4E_CUDA_TRY(cudaStreamSynchronize(s));
  StemPoolPlan plan;
  // (u8 mode: the plan's bf16 input map is unused; any aligned pointer encodes)
  int rc = stem_pool_make_plan(&plan, u8 ? d_out_nhwc : d_in_nhwc4, batch, h, w, d_sp, d_bias, relu, d_out_nhwc);
  if (rc == AI4E_OK)
    rc = u8 ? stem_pool_launch_u8(plan, (const uint8_t*)d_in_nhwc4, batch, h_mean, h_std, s) : stem_pool_launch(plan, batch, s);
  cudaFreeAsync(d_sp, s);
  return rc;
}

int ai4e_jpeg_entropy_decode(const uint8_t* jpeg, uint64_t len, int16_t* h_record, uint64_t capacity_int16,
                             ai4e_jpeg_info* out) {
  return jpeg_host_decode(jpeg, len, h_record, capacity_int16, out);
}

int ai4e_jpeg_reconstruct(const int16_t* d_records, int batch, const ai4e_jpeg_info* geometry, uint8_t* d_scratch,
                          uint8_t* d_rgb_hwc, void* stream) {
  return jpeg_reconstruct(d_records, batch, geometry, d_scratch, d_rgb_hwc, (cudaStream_t)stream);
}

int ai4e_jpeg_decode(const uint8_t* jpeg, uint64_t len, uint8_t* d_rgb_hwc, uint64_t capacity_bytes, int32_t* out_height,
                     int32_t* out_width, void* stream) {
  AI4E_REQUIRE(jpeg && d_rgb_hwc, "jpeg_decode: null pointer");
  ai4e_jpeg_info info;
  int rc = jpeg_host_decode(jpeg, len, nullptr, 0, &info);
  if (rc) return rc;
  if (out_height) *out_height = info.height;
  if (out_width) *out_width = info.width;
  AI4E_REQUIRE(capacity_bytes >= (uint64_t)info.width * info.height * 3, "jpeg_decode: output needs %lld bytes",
               (long long)info.width * info.height * 3);
  std::vector<int16_t> rec((size_t)info.record_int16);
  rc = jpeg_host_decode(jpeg, len, rec.data(), rec.size(), &info);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  int16_t* d_rec = nullptr;
  uint8_t* d_scratch = nullptr;
  AI4E_CUDA_TRY(cudaMallocAsync((void**)&d_rec, rec.size() * sizeof(int16_t), s));
  AI4E_CUDA_TRY(cudaMallocAsync((void**)&d_scratch, (size_t)info.plane_bytes, s));
  AI4E_CUDA_TRY(cudaMemcpyAsync(d_rec, rec.data(), rec.size() * sizeof(int16_t), cudaMemcpyHostToDevice, s));
  rc = jpeg_reconstruct(d_rec, 1, &info, d_scratch, d_rgb_hwc, s);
  AI4E_CUDA_TRY(cudaStreamSynchronize(s));  // `rec` is a stack-lifetime host buffer
  cudaFreeAsync(d_rec, s);
  cudaFreeAsync(d_scratch, s);
  return rc;
}

int ai4e_bottleneck_tail(const void* d_in_nhwc, int batch, int h, int w, int cin, const void* d_w2_krsc,
                         const float* d_bias2, int cmid, int kh, int kw, int stride, int pad, const void* d_w3_kc,
                         const float* d_bias3, int cout, const void* d_residual, void* d_out_nhwc, void* stream) {
  AI4E_REQUIRE(d_in_nhwc && d_w2_krsc && d_bias2 && d_w3_kc && d_bias3 && d_residual && d_out_nhwc,
               "bottleneck tail: null pointer");
  AI4E_REQUIRE(batch > 0 && h > 0 && w > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0, "bottleneck tail: bad geometry");
  AI4E_REQUIRE(h + 2 * pad >= kh && w + 2 * pad >= kw, "bottleneck tail: filter larger than padded input");
  FusedPlan plan;
  int rc = fused_tail_make_plan(&plan, d_in_nhwc, batch, h, w, cin, d_w2_krsc, d_bias2, cmid, kh, kw, stride, pad, d_w3_kc,
                                d_bias3, cout, d_residual, d_out_nhwc);
  if (rc) return rc;
  if (const char* trace = getenv("AI4E_B200_PIPE_TRACE")) {  // developer aid (needs a -DAI4E_PIPE_TRACE build): raw stamps of CTA 0
    long long* d = nullptr;
    AI4E_CUDA_TRY(cudaMalloc((void**)&d, 2048 * sizeof(long long)));
    AI4E_CUDA_TRY(cudaMemset(d, 0, 2048 * sizeof(long long)));
    plan.prm.dbg = d;
    rc = fused_tail_launch(plan, batch, (cudaStream_t)stream);
    AI4E_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    std::vector<long long> h(2048);
    AI4E_CUDA_TRY(cudaMemcpy(h.data(), d, 2048 * sizeof(long long), cudaMemcpyDeviceToHost));
    cudaFree(d);
    if (FILE* f = fopen(trace, "w")) {
      for (int i = 0; i < 2048; ++i) fprintf(f, "%d %lld\n", i, h[i]);
      fclose(f);
    }
    return rc;
  }
  return fused_tail_launch(plan, batch, (cudaStream_t)stream);
}

int ai4e_maxpool(const void* d_in_nhwc, int batch, int h, int w, int c, int k, int stride, int pad,
                 void* d_out_nhwc, void* stream) {
  return launch_maxpool(d_in_nhwc, batch, h, w, c, k, stride, pad, d_out_nhwc, (cudaStream_t)stream);
}

int ai4e_head(const void* d_in_nhwc, int batch, int hw, int c, const void* d_fc_w,
              const float* d_fc_b, int classes, float* d_logits, int32_t* d_top_idx,
              float* d_top_prob, int topk, void* stream) {
  AI4E_REQUIRE(batch > 0 && c > 0, "head: bad geometry");
  float* scratch = nullptr;
  cudaStream_t s = (cudaStream_t)stream;
  AI4E_CUDA_TRY(cudaMallocAsync(&scratch, (size_t)batch * c * sizeof(float), s));
  const int rc = launch_head(d_in_nhwc, batch, hw, c, d_fc_w, d_fc_b, classes, scratch, d_logits,
                             d_top_idx, d_top_prob, topk, s);
  cudaFreeAsync(scratch, s);
  return rc;
}

}  // extern "C"
