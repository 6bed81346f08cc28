// K1 — preprocessing: uint8 HWC -> normalised bf16 (NCHW planes, or NHWC with the channel
// count padded for the stem), optional bilinear resize + centre crop fused in.
//
// Replaces what a PyTorch CPU container does inside the user callback the reference calls at
// APIs/1.0/base-py/ai4e_service.py:203 (ToTensor -> Normalize, optionally Resize ->
// CenterCrop first); restated on the CPU in oracle/preprocess_oracle.py.
//
// Two kernels:
//   * preprocess_lut_kernel     no resize, dense rows, H*W % 16 == 0.  HBM-bound streaming
//     kernel: each CTA stages 4096-pixel chunks of the interleaved source into shared memory
//     with 1-D bulk async copies (TMA engine, mbarrier completion, two stages in flight),
//     every thread converts 16 pixels through a 256-entry-per-channel bf16 look-up table
//     held in shared memory and writes 16-byte vectors.  The table is built in the prologue
//     with IEEE-exact, separately rounded fp32 operations ((v/255 - mean)/std), so the
//     output is bit-identical to the oracle by construction: a byte has only 256 values.
//   * preprocess_generic_kernel everything else (resize/crop, row strides, ragged sizes):
//     one thread per output pixel, same fp32 operation order as the oracle, no FMA
//     contraction.
#include "common.cuh"
#include "kernels.h"
#include <cstdlib>

namespace ai4e {

struct PreprocParams {
  const uint8_t* src;
  uint16_t* dst;  // bf16 bits
  int batch, src_h, src_w, channels;
  long long src_image_stride;  // bytes
  int src_row_stride;          // bytes
  int resize_h, resize_w;      // 0 => no resize
  int crop_top, crop_left;     // offsets into the (resized) image
  int out_h, out_w;
  int layout;        // 0 = NCHW, 1 = NHWC
  int out_channels;  // NHWC: padded channel count (>= channels); NCHW: == channels
  float mean[4], stdv[4];
  // K6 chip grid (grid_w > 0): image i starts at ((first + i) / grid_w) * y_stride + ((first + i) % grid_w) * x_stride
  int grid_w, grid_first;
  long long grid_x_stride, grid_y_stride;
  int rows_per_chunk;  // streaming kernel with strided rows: a chunk is this many whole rows (0 = dense images)
};

__device__ __forceinline__ const uint8_t* image_base(const PreprocParams& p, int img) {
  if (p.grid_w > 0) {
    const int g = p.grid_first + img;
    const int gy = g / p.grid_w, gx = g - gy * p.grid_w;
    return p.src + (long long)gy * p.grid_y_stride + (long long)gx * p.grid_x_stride;
  }
  return p.src + (long long)img * p.src_image_stride;
}

constexpr int kPixPerThread = 16;
constexpr int kLutThreads = 256;
constexpr int kChunkPix = kPixPerThread * kLutThreads;  // 4096


// Shared-memory budget of the streaming kernel: the look-up table is replicated per lane
// (entry (c, v) for lane l lives in word (c * 256 + v) * 32 + l, i.e. always in bank l), so the
// random byte-indexed look-ups never conflict.  (With one 512-byte table per channel ncu showed the
// kernel bound by the shared-memory queue: mio_throttle 19 of 37 stall cycles per issue, 537 k bank
// conflicts, 15 % issue utilisation, 19.9 us for 35 MB.)  One CTA per SM; the source chunks of a CTA
// are all requested up front, kLutStages deep, because one chunk converts in a fraction of the
// ~1.5 us its load takes.
template <int C>
struct LutCfg {
  static constexpr int kStages = C == 3 ? 6 : 5;
  static constexpr int kLutBytes = C * 256 * 32 * 4;      // 96 / 128 KiB
  static constexpr int kStageBytes = kChunkPix * C;       // 12 / 16 KiB
  static constexpr int kSmem = kLutBytes + kStages * kStageBytes + kStages * 8;
};

template <int C>
__global__ void __launch_bounds__(kLutThreads, 1)
preprocess_lut_kernel(const PreprocParams p, int chunks_per_image, int total_chunks) {
  using Cfg = LutCfg<C>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  // layout: lut[C][256][32] u32 | [kStages][kChunkPix*C bytes] | mbarriers
  uint32_t* lut = reinterpret_cast<uint32_t*>(smem_raw);
  uint8_t* stage_buf = smem_raw + Cfg::kLutBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + Cfg::kLutBytes + kStages * Cfg::kStageBytes);

  const int tid = threadIdx.x, lane = tid & 31;
  pdl_launch_dependents();  // the stem may start its prologue while this kernel streams
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(smem_u32(&bars[s]), 1);
    fence_mbar_init();
  }
  __syncthreads();

  const int hw = p.out_h * p.out_w;
  auto issue = [&](int chunk, int stage) {
    const int img = chunk / chunks_per_image;
    const int cj = chunk - img * chunks_per_image;
    const int px0 = cj * kChunkPix;
    const int npx = min(kChunkPix, hw - px0);
    const uint32_t bytes = (uint32_t)npx * C;
    const uint8_t* base = image_base(p, img);
    const uint32_t bar = smem_u32(&bars[stage]);
    mbar_expect_tx(bar, bytes);
    if (p.rows_per_chunk == 0) {
      bulk_load_1d(smem_u32(stage_buf + stage * Cfg::kStageBytes), base + (long long)px0 * C, bytes, bar);
    } else {
      // strided rows (a chip inside a tile): the chunk is whole rows, one bulk copy per row, one barrier
      const int row0 = px0 / p.src_w, rows = npx / p.src_w;
      const uint32_t rb = (uint32_t)p.src_w * C;
      for (int r = 0; r < rows; ++r)
        bulk_load_1d(smem_u32(stage_buf + stage * Cfg::kStageBytes + r * rb),
                     base + (long long)(row0 + r) * p.src_row_stride, rb, bar);
    }
  };
  if (tid == 0) {  // the source is the caller's buffer, not a predecessor kernel's output
    for (int s = 0; s < kStages; ++s) {
      const int chunk = blockIdx.x + s * gridDim.x;
      if (chunk < total_chunks) issue(chunk, s);
    }
  }
  // table: thread v computes entry v of each channel once and writes its 32 copies, lane l starting
  // at copy l so that the 32 lanes of a warp hit 32 different banks
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const uint32_t e = norm_to_bf16((float)tid, p.mean[c], p.stdv[c]);
    uint32_t* row = lut + (c * 256 + tid) * 32;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) row[(k + lane) & 31] = e;
  }
  __syncthreads();
  const uint32_t* lut_lane = lut + lane;

  uint32_t it = 0;
  for (int chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x, ++it) {
    const int stage_i = it % kStages;
    mbar_wait(smem_u32(&bars[stage_i]), (it / kStages) & 1);

    const int img = chunk / chunks_per_image;
    const int cj = chunk - img * chunks_per_image;
    auto look = [&](int c, uint32_t v) -> uint32_t { return lut_lane[(c * 256 + v) * 32]; };
    const uint8_t* stage = stage_buf + stage_i * Cfg::kStageBytes;
    if (p.layout == 0) {
      // NCHW planes: a thread converts 16 consecutive pixels (C 16-byte vectors; the 16*C-byte stride
      // between lanes is bank-conflict free for C = 3 and C = 4) and writes 32 contiguous bytes per plane
      const int px = cj * kChunkPix + tid * kPixPerThread;
      if (px < hw) {
        const uint4* s4 = reinterpret_cast<const uint4*>(stage + tid * kPixPerThread * C);
        uint32_t w[4 * C];
#pragma unroll
        for (int i = 0; i < C; ++i) {
          uint4 v = s4[i];
          w[4 * i + 0] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
        }
        auto byte_at = [&](int k) -> uint32_t { return (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu; };
#pragma unroll
        for (int c = 0; c < C; ++c) {
          uint32_t o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t lo = look(c, byte_at((2 * j) * C + c));
            const uint32_t hi = look(c, byte_at((2 * j + 1) * C + c));
            o[j] = lo | (hi << 16);
          }
          uint4* d = reinterpret_cast<uint4*>(p.dst + ((long long)(img * C + c) * hw + px));
          d[0] = make_uint4(o[0], o[1], o[2], o[3]);
          d[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
      }
    } else {
      // NHWC with channel padding to 4 (C == 3: pad lane is +0.0; C == 4: all real).  Two pixels
      // make one 16-byte store; in each of 8 rounds a warp converts 64 consecutive pixels, lane l the
      // pair l, so the warp writes 512 contiguous bytes.  (16 consecutive pixels per thread made every
      // store instruction touch 32 different 128-byte lines.)
      const int npx = min(kChunkPix, hw - cj * kChunkPix);
      uint4* d = reinterpret_cast<uint4*>(p.dst + ((long long)img * hw + cj * kChunkPix) * 4);
      const int warp_px0 = (tid >> 5) * 32 * kPixPerThread;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int pix = warp_px0 + j * 64 + lane * 2;
        if (pix < npx) {
          uint32_t b[2 * C];  // the bytes of the two pixels
          if constexpr (C == 3) {
            const uint16_t* sp = reinterpret_cast<const uint16_t*>(stage + pix * 3);
            const uint32_t h0 = sp[0], h1 = sp[1], h2 = sp[2];
            b[0] = h0 & 0xFF; b[1] = h0 >> 8; b[2] = h1 & 0xFF; b[3] = h1 >> 8; b[4] = h2 & 0xFF; b[5] = h2 >> 8;
          } else {
            const uint2 v = *reinterpret_cast<const uint2*>(stage + pix * 4);
            b[0] = v.x & 0xFF; b[1] = (v.x >> 8) & 0xFF; b[2] = (v.x >> 16) & 0xFF; b[3] = v.x >> 24;
            b[4] = v.y & 0xFF; b[5] = (v.y >> 8) & 0xFF; b[6] = (v.y >> 16) & 0xFF; b[7] = v.y >> 24;
          }
          uint32_t o[4];
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const uint32_t c0 = look(0, b[q * C + 0]);
            const uint32_t c1 = look(1, b[q * C + 1]);
            const uint32_t c2 = look(2, b[q * C + 2]);
            const uint32_t c3 = (C == 4) ? look(C - 1, b[q * C + (C - 1)]) : 0u;
            o[2 * q] = c0 | (c1 << 16);
            o[2 * q + 1] = c2 | (c3 << 16);
          }
          d[pix >> 1] = make_uint4(o[0], o[1], o[2], o[3]);
        }
      }
    }
    __syncthreads();  // every thread has read this stage: refill it with the chunk kStages ahead
    const int next = chunk + kStages * gridDim.x;
    if (tid == 0 && next < total_chunks) issue(next, stage_i);
  }
}

__device__ __forceinline__ void bilinear_axis(int dst, int in_size, int out_size, int& i0, int& i1,
                                              float& l0, float& l1) {
  const float scale = __fdiv_rn((float)in_size, (float)out_size);
  float src = __fsub_rn(__fmul_rn(__fadd_rn((float)dst, 0.5f), scale), 0.5f);
  src = fmaxf(src, 0.0f);
  i0 = min((int)floorf(src), in_size - 1);
  i1 = min(i0 + 1, in_size - 1);
  l1 = __fsub_rn(src, (float)i0);
  l0 = __fsub_rn(1.0f, l1);
}

__global__ void __launch_bounds__(256)
preprocess_generic_kernel(const PreprocParams p) {
  const long long total = (long long)p.batch * p.out_h * p.out_w;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(idx % p.out_w);
    const int y = (int)((idx / p.out_w) % p.out_h);
    const int n = (int)(idx / ((long long)p.out_w * p.out_h));
    const uint8_t* img = image_base(p, n);
    const int C = p.channels;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (p.resize_h > 0) {
      int y0, y1, x0, x1;
      float ly0, ly1, lx0, lx1;
      bilinear_axis(y + p.crop_top, p.src_h, p.resize_h, y0, y1, ly0, ly1);
      bilinear_axis(x + p.crop_left, p.src_w, p.resize_w, x0, x1, lx0, lx1);
      const uint8_t* r0 = img + (long long)y0 * p.src_row_stride;
      const uint8_t* r1 = img + (long long)y1 * p.src_row_stride;
      for (int c = 0; c < C; ++c) {
        const float p00 = r0[x0 * C + c], p01 = r0[x1 * C + c];
        const float p10 = r1[x0 * C + c], p11 = r1[x1 * C + c];
        const float top = __fadd_rn(__fmul_rn(lx0, p00), __fmul_rn(lx1, p01));
        const float bot = __fadd_rn(__fmul_rn(lx0, p10), __fmul_rn(lx1, p11));
        v[c] = __fadd_rn(__fmul_rn(ly0, top), __fmul_rn(ly1, bot));
      }
    } else {
      const uint8_t* px = img + (long long)(y + p.crop_top) * p.src_row_stride +
                          (long long)(x + p.crop_left) * C;
      for (int c = 0; c < C; ++c) v[c] = (float)px[c];
    }
    if (p.layout == 0) {
      const long long plane = (long long)p.out_h * p.out_w;
      for (int c = 0; c < C; ++c)
        p.dst[((long long)n * C + c) * plane + (long long)y * p.out_w + x] =
            norm_to_bf16(v[c], p.mean[c], p.stdv[c]);
    } else {
      uint16_t* d = p.dst + idx * p.out_channels;
      for (int c = 0; c < p.out_channels; ++c)
        d[c] = (c < C) ? norm_to_bf16(v[c], p.mean[c], p.stdv[c]) : (uint16_t)0;
    }
  }
}

static std::atomic<int> g_sm_count{0};
int sm_count() {  // of the first device asked about; a box holds one kind of GPU
  int n = g_sm_count.load(std::memory_order_relaxed);
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    g_sm_count.store(n, std::memory_order_relaxed);
  }
  return n;
}

// returns 1 if the streaming (LUT) kernel was used, 0 for the generic kernel, <0 on error
int launch_preprocess(const uint8_t* d_src, int batch, int src_h, int src_w, int channels,
                      long long src_image_stride, int src_row_stride, int resize_h, int resize_w,
                      int out_h, int out_w, const float* mean, const float* stdv, int layout,
                      int out_channels, void* d_dst, cudaStream_t stream, const ChipGrid* chips) {
  AI4E_REQUIRE(d_src && d_dst && mean && stdv, "preprocess: null pointer");
  if (chips && chips->grid_w <= 0) chips = nullptr;
  AI4E_REQUIRE(batch > 0 && src_h > 0 && src_w > 0, "preprocess: empty input (%d x %d x %d)", batch,
               src_h, src_w);
  AI4E_REQUIRE(channels == 3 || channels == 4, "preprocess: channels must be 3 or 4, got %d",
               channels);
  AI4E_REQUIRE(layout == 0 || layout == 1, "preprocess: layout must be 0 (NCHW) or 1 (NHWC)");
  if (layout == 0) AI4E_REQUIRE(out_channels == channels, "preprocess: NCHW keeps the channel count");
  if (layout == 1)
    AI4E_REQUIRE(out_channels >= channels && out_channels <= 8,
                 "preprocess: NHWC out_channels %d outside [%d, 8]", out_channels, channels);
  const bool resize = resize_h > 0 || resize_w > 0;
  if (resize) AI4E_REQUIRE(resize_h > 0 && resize_w > 0, "preprocess: resize needs both sizes");
  const int full_h = resize ? resize_h : src_h, full_w = resize ? resize_w : src_w;
  AI4E_REQUIRE(out_h > 0 && out_w > 0 && out_h <= full_h && out_w <= full_w,
               "preprocess: crop %dx%d does not fit %dx%d", out_h, out_w, full_h, full_w);
  if (src_row_stride <= 0) src_row_stride = src_w * channels;
  if (src_image_stride <= 0) src_image_stride = (long long)src_row_stride * src_h;
  AI4E_REQUIRE(src_row_stride >= src_w * channels, "preprocess: row stride too small");
  for (int c = 0; c < channels; ++c)
    AI4E_REQUIRE(stdv[c] != 0.0f, "preprocess: std[%d] is zero", c);

  PreprocParams p{};
  p.src = d_src; p.dst = reinterpret_cast<uint16_t*>(d_dst);
  p.batch = batch; p.src_h = src_h; p.src_w = src_w; p.channels = channels;
  p.src_image_stride = src_image_stride; p.src_row_stride = src_row_stride;
  p.resize_h = resize ? resize_h : 0; p.resize_w = resize ? resize_w : 0;
  // torchvision CenterCrop: int(round((size - crop) / 2.0)), Python round-half-to-even
  auto centre = [](int full, int crop) {
    const int d = full - crop;
    int h = d / 2;
    if ((d & 1) && (h & 1)) h += 1;  // d/2 = h + 0.5 rounds to the even neighbour
    return h;
  };
  p.crop_top = centre(full_h, out_h); p.crop_left = centre(full_w, out_w);
  p.out_h = out_h; p.out_w = out_w; p.layout = layout; p.out_channels = out_channels;
  for (int c = 0; c < 4; ++c) { p.mean[c] = c < channels ? mean[c] : 0.f; p.stdv[c] = c < channels ? stdv[c] : 1.f; }
  if (chips) {
    AI4E_REQUIRE(chips->first >= 0 && chips->x_stride >= (long long)src_w * channels &&
                 chips->y_stride >= (long long)src_row_stride * src_h,
                 "preprocess: chip grid strides smaller than a chip");
    p.grid_w = chips->grid_w; p.grid_first = chips->first;
    p.grid_x_stride = chips->x_stride; p.grid_y_stride = chips->y_stride;
  }

  const long long hw = (long long)out_h * out_w;
  const bool same_size = !resize && out_h == src_h && out_w == src_w;
  const bool dense = same_size && src_row_stride == src_w * channels;
  // strided rows stream too when a 4096-pixel chunk is whole rows and every row is a 16-byte multiple
  // at a 16-byte aligned address (256 x 256 x 4 chips of a 4096-wide tile: 16 rows of 1 KB per chunk)
  const bool strided_ok = same_size && !dense && (kChunkPix % src_w == 0) && ((src_w * channels) % 16 == 0) &&
                          (src_row_stride % 16 == 0);
  const bool bases_ok = chips ? (chips->x_stride % 16 == 0 && chips->y_stride % 16 == 0) : (src_image_stride % 16 == 0);
  const bool fast = (dense || strided_ok) && (hw % kPixPerThread == 0) && bases_ok &&
                    ((reinterpret_cast<uintptr_t>(d_src) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(d_dst) & 15) == 0) &&
                    (layout == 0 || out_channels == 4);
  p.rows_per_chunk = (fast && !dense) ? kChunkPix / src_w : 0;
  if (fast) {
    const int chunks_per_image = (int)((hw + kChunkPix - 1) / kChunkPix);
    const long long total_chunks = (long long)chunks_per_image * batch;
    AI4E_REQUIRE(total_chunks < (1ll << 31), "preprocess: batch too large");
    const int grid = (int)std::min<long long>(total_chunks, (long long)sm_count());  // one CTA per SM (shared memory)
    if (channels == 3) {
      static std::atomic<unsigned long long> attr3{0};
      if (int rc = ensure_dyn_smem(preprocess_lut_kernel<3>, LutCfg<3>::kSmem, attr3)) return rc;
      preprocess_lut_kernel<3><<<grid, kLutThreads, LutCfg<3>::kSmem, stream>>>(p, chunks_per_image, (int)total_chunks);
    } else {
      static std::atomic<unsigned long long> attr4{0};
      if (int rc = ensure_dyn_smem(preprocess_lut_kernel<4>, LutCfg<4>::kSmem, attr4)) return rc;
      preprocess_lut_kernel<4><<<grid, kLutThreads, LutCfg<4>::kSmem, stream>>>(p, chunks_per_image, (int)total_chunks);
    }
    AI4E_CUDA_TRY(cudaGetLastError());
    return 1;
  }
  const long long total = (long long)batch * hw;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)sm_count() * 8);
  preprocess_generic_kernel<<<grid, 256, 0, stream>>>(p);
  AI4E_CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace ai4e
