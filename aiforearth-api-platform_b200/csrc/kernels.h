// Host-side launch interface of the kernels in this directory (internal to the library;
// the public boundary is include/ai4e_b200.h).
#pragma once

#include "common.cuh"

struct ai4e_jpeg_info;  // include/ai4e_b200.h

namespace ai4e {

int sm_count();

// K6: the images of a batch are chips of one larger tile.  Image i sits at
//   d_src + ((first + i) / grid_w) * y_stride + ((first + i) % grid_w) * x_stride        (bytes)
// and its rows are `row_bytes` apart (the tile's row pitch): the gather is address arithmetic inside
// K1, no copy.  grid_w == 0 (or a null pointer) means "images src_image_stride apart".
struct ChipGrid {
  int grid_w = 0, first = 0;
  long long x_stride = 0, y_stride = 0, row_bytes = 0;
};

// ---- K1 preprocess (preprocess.cu); returns 1 if the streaming LUT kernel ran, 0 generic
int launch_preprocess(const uint8_t* d_src, int batch, int src_h, int src_w, int channels,
                      long long src_image_stride, int src_row_stride, int resize_h, int resize_w,
                      int out_h, int out_w, const float* mean, const float* stdv, int layout,
                      int out_channels, void* d_dst, cudaStream_t stream, const ChipGrid* chips = nullptr);

// ---- K0 JPEG, baseline and progressive (jpeg.cu): host headers + entropy decode, device IDCT / upsampling / colour
int jpeg_host_decode(const uint8_t* data, uint64_t len, int16_t* record, uint64_t capacity_int16, ai4e_jpeg_info* out);
int jpeg_reconstruct(const int16_t* d_records, int batch, const ai4e_jpeg_info* geometry, uint8_t* d_scratch,
                     uint8_t* d_out, cudaStream_t stream, const float* norm_mean_std = nullptr);

// ---- PNG (png.cu): host only; rgb == nullptr sizes the output
int png_host_decode(const uint8_t* data, uint64_t len, uint8_t* rgb, uint64_t capacity, int* out_h, int* out_w);

// ---- direct convolution on CUDA cores (conv_direct.cu)
int launch_conv_direct(const void* in, int N, int H, int W, int Cin, const void* weight,
                       const float* bias, int Cout, int R, int S, int stride, int pad,
                       const void* residual, int relu, void* out, cudaStream_t stream);

// ---- K4 tcgen05 implicit-GEMM convolution (conv_igemm.cu)
struct IgemmParams {
  int M, P, Q, PQ;
  int Cout;
  int S;           // filter width (taps per filter row)
  int stride, pad;
  int cin_blocks;  // Cin / 64
  int num_k_blocks;
  int tiles_m, tiles_n, num_tiles;
  int relu;
  int reverse;     // walk the row tiles from the last to the first (set per layer by the engine)
  int stem_rb;     // row-mode stem: bytes per input-row buffer in shared memory
  long long* dbg;  // developer aid: clock64 stamps of CTA 0's pipeline (768 slots), null in production
  const float* bias;
  const bf16* residual;
  bf16* out;
};

struct IgemmPlan {
  CUtensorMap tmap_a;     // im2col-mode map over the input activations (N = max_batch)
  CUtensorMap tmap_b[3];  // weight tiles for BLOCK_N = 64 / 128 / 256
  CUtensorMap tmap_out;   // output tiles (TMA store), 128 rows x 64 channels
  CUtensorMap tmap_res;   // residual tiles (TMA load), same geometry
  bool has_bn[3];
  IgemmParams prm;        // everything but the batch-dependent fields
  int max_batch;
  int stem_mode;          // 0 generic conv; 1 stem via im2col super-pixels; 2 stem via row buffers
  int tuned_tiling;       // igemm_autotune's choice for launches with tuned_batch images (0 = none)
  int tuned_batch;
  int grid_sms;           // persistent grid sized for this many SMs (0 = all)
};

// K4b bottleneck tail (conv RxS + relu, then conv 1x1 + residual + relu, one kernel): conv_fused.cu
struct FusedParams {
  int M, P, Q, PQ;
  int Cmid, Cout;
  int S, stride, pad;
  int cin_blocks, num_k_blocks;
  int num_tiles;
  int reverse;  // walk the row tiles from the last to the first
  int interleave;  // spread the second GEMM's passes of tile t-1 among the K blocks of tile t (see conv_fused.cu)
  long long* dbg;  // developer aid (-DAI4E_PIPE_TRACE): clock64 stamps of CTA 0's pipeline (2048 slots), null in production
  const float* bias2;
  const float* bias3;
};
struct FusedPlan {
  CUtensorMap tmap_a;    // im2col-mode map over the input activations
  CUtensorMap tmap_w2;   // first convolution's weights, [Cmid][R*S*Cin], box 64 x Cmid
  CUtensorMap tmap_w3;   // second (1x1) convolution's weights, [Cout][Cmid], box 64 x 128
  CUtensorMap tmap_out;  // z, 32 rows x 64 channels per store
  CUtensorMap tmap_res;  // residual, 128 rows x 64 channels per load
  FusedParams prm;
  int max_batch;
  int grid_sms;          // persistent grid sized for this many SMs (0 = all)
};
int fused_tail_supported(int Cin, int Cmid, int Cout, int R, int S, int stride, int pad);
int fused_tail_make_plan(FusedPlan* plan, const void* in, int N, int H, int W, int Cin, const void* w2,
                         const float* bias2, int Cmid, int R, int S, int stride, int pad, const void* w3,
                         const float* bias3, int Cout, const void* residual, void* out);
int fused_tail_launch(const FusedPlan& plan, int batch, cudaStream_t stream);

int igemm_supported(int Cin, int Cout, int R, int S, int stride, int pad);
constexpr int kTilingPair = 1000;  // tiling code = BLOCK_N (+ kTilingPair: CTA-pair kernel, tcgen05 cta_group::2)
int igemm_pick_tiling(int Cout, long long M, int num_k_blocks, bool has_res, int sms = 0);
int igemm_make_plan(IgemmPlan* plan, const void* in, int N, int H, int W, int Cin,
                    const void* weight, const float* bias, int Cout, int R, int S, int stride,
                    int pad, const void* residual, int relu, void* out);
int igemm_launch(const IgemmPlan& plan, int batch, int tiling, cudaStream_t stream);
int igemm_autotune(IgemmPlan* plan, int batch, cudaStream_t stream, float* best_us_out);
// tensor-core stem: 7x7 stride 2 pad 3, NHWC4 input, 64 outputs
int stem_supported(int H, int W, int Cin, int Cout, int R, int S, int stride, int pad);
void stem_repack_weights(const uint16_t* w_krsc, uint16_t* w_sp);  // [64][7][7][4] -> [64][7][4][8]
int igemm_make_plan_stem(IgemmPlan* plan, const void* in, int N, int H, int W, const void* weight_sp,
                         const float* bias, int relu, void* out);

// ---- K7 tile summary (tile_summary.cu): label map, class histogram and 32.32 fixed-point probability mass of a tile's chips
int tile_summary_launch(const int32_t* d_idx5, const float* d_prob5, int count, int first, int num_classes,
                        int32_t* d_labels, float* d_label_prob, uint32_t* d_hist, unsigned long long* d_mass_q32,
                        cudaStream_t stream);

// ---- K2 + K3 in one kernel (stem_pool.cu): 7x7/2 stem + bias + ReLU + 3x3/2 max-pool, NHWC4 in, pooled NHWC out
struct StemPoolPlan {
  CUtensorMap tmap_a;  // input rows, [N][H][W/16][64 elements], nine rows per load
  CUtensorMap tmap_b;  // (unused)
  const void* weights;  // repacked weights, slab-major [36][128][8]: one bulk copy
  int H, P, Q, Pp, Qp, stem_rb, relu;
  const float* bias;
  void* out;
  int max_batch;
  int grid_sms;
};
int stem_pool_supported(int H, int W, int Cin, int Cout, int R, int S, int stride, int pad, int pool_k, int pool_stride,
                        int pool_pad);
int stem_pool_weight_elems();
void stem_pool_repack_weights(const uint16_t* w_krsc, uint16_t* w_sp);  // [64][7][7][4] -> [9][4][128][8]
int stem_pool_make_plan(StemPoolPlan* plan, const void* in, int N, int H, int W, const void* weight_sp,
                        const float* bias, int relu, void* out);
int stem_pool_launch(const StemPoolPlan& plan, int batch, cudaStream_t stream);
// K1 fused in as well: the uint8 HWC batch (3 channels, no resize) goes straight into the stem
int stem_pool_u8_supported(int H, int W, int channels);
int stem_pool_launch_u8(const StemPoolPlan& plan, const uint8_t* d_images, int batch, const float* mean, const float* stdv,
                        cudaStream_t stream);

// ---- K3 / K5 (pool_head.cu)
int launch_maxpool(const void* in, int N, int H, int W, int C, int k, int stride, int pad, void* out,
                   cudaStream_t stream, int reverse = 0);
int launch_head(const void* in, int N, int HW, int C, const void* fc_w, const float* fc_b,
                int classes, float* pooled_scratch, float* logits, int32_t* top_idx, float* top_prob,
                int topk, cudaStream_t stream);

}  // namespace ai4e
