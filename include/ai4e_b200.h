/* libai4e_b200 — C ABI of the B200-native classify engine that sits under the AI for Earth
 * API Platform's container host shell.
 *
 * The reference has NO foreign-function or plugin ABI on this path: its extension point is
 * the Python callback `r = func(*args, **kwargs)` at
 * /root/reference/APIs/1.0/base-py/ai4e_service.py:203, invoked from the views that
 * `APIService.api_sync_func` / `api_async_func` create (ai4e_service.py:72-109).  Everything
 * a PyTorch CPU container does inside that callback (decode -> preprocess -> model forward
 * -> softmax/top-5) is what this library replaces; each entry point below names the piece
 * of that callback it stands in for.  INTEGRATION.md shows the ctypes stub a maintainer adds
 * on the reference side.
 *
 * Conventions
 *   - every function returns 0 on success or a negative AI4E_ERR_* code; the message for
 *     the calling thread is `ai4e_last_error()`;
 *   - plain pointers and sizes only; outputs are caller-allocated; no exceptions cross;
 *   - `d_` pointers are device memory on the current CUDA device, `h_` pointers are host
 *     memory; `stream` is a `cudaStream_t` passed as `void*` (NULL = default stream);
 *   - activations are NHWC bfloat16; convolution weights are [Cout][kh][kw][Cin] bfloat16
 *     with BatchNorm already folded in, biases fp32;
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef AI4E_B200_H_
#define AI4E_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define AI4E_API __attribute__((visibility("default")))
#else
#define AI4E_API
#endif

#define AI4E_OK 0
#define AI4E_ERR_INVALID (-1)     /* bad argument / unsupported shape */
#define AI4E_ERR_CUDA (-2)        /* a CUDA runtime or driver call failed */
#define AI4E_ERR_UNSUPPORTED (-3) /* feature missing on this driver / device */
#define AI4E_ERR_NOMEM (-4)
#define AI4E_ERR_STATE (-5)       /* engine closed, unknown ticket, ... */
#define AI4E_ERR_TIMEOUT (-6)

#define AI4E_LAYOUT_NCHW 0
#define AI4E_LAYOUT_NHWC 1

#define AI4E_CONV_AUTO 0   /* tcgen05 implicit GEMM when the shape allows, else direct */
#define AI4E_CONV_IGEMM 1  /* tcgen05 implicit GEMM, error if the shape does not allow it */
#define AI4E_CONV_DIRECT 2 /* CUDA-core direct convolution (stem / cross-check) */

#define AI4E_OP_CONV 0
#define AI4E_OP_MAXPOOL 1
#define AI4E_OP_HEAD 2 /* global average pool -> FC -> softmax -> top-k */

#define AI4E_TOPK 5

/* Library version (major*10000 + minor*100 + patch).  No reference counterpart. */
AI4E_API int ai4e_version(void);

/* Message of the last error on the calling thread.  The reference reports errors by raising
 * inside the callback (ai4e_service.py:205-211); this is the C-ABI equivalent. */
AI4E_API const char* ai4e_last_error(void);

/* Number of CUDA devices (the reference scales by Kubernetes replicas, one GPU per pod:
 * APIs/Charts/camera-trap/detection-async/prod-values.yaml:18). */
AI4E_API int ai4e_device_count(int* out_count);

/* ---------------------------------------------------------------------------------------
 * Stateless kernels (device pointers).  These are the parity-test entry points.
 * ------------------------------------------------------------------------------------- */

/* K1.  uint8 HWC -> normalised bf16: [optional bilinear resize (align_corners = false, no
 * antialias) -> centre crop ->] (x/255 - mean)/std.  Stands in for torchvision
 * Resize/CenterCrop/ToTensor/Normalize inside the callback at ai4e_service.py:203.
 *   d_src            batch images, `src_image_stride` bytes apart (0 = dense), rows
 *                    `src_row_stride` bytes apart (0 = src_w*channels)
 *   channels         3 or 4
 *   resize_h/w       0,0 = no resize
 *   out_h/out_w      output size = centre crop of the (resized) image
 *   h_mean/h_std     host arrays, `channels` entries
 *   layout           AI4E_LAYOUT_NCHW ([B][C][out_h][out_w], out_channels == channels) or
 *                    AI4E_LAYOUT_NHWC ([B][out_h][out_w][out_channels], extra channels = 0)
 * Bit-exact against oracle/preprocess_oracle.py. */
AI4E_API int ai4e_preprocess(const uint8_t* d_src, int batch, int src_h, int src_w, int channels,
                    int64_t src_image_stride, int src_row_stride, int resize_h, int resize_w,
                    int out_h, int out_w, const float* h_mean, const float* h_std, int layout,
                    int out_channels, void* d_dst_bf16, void* stream);

/* K6 fused into K1: the images are chips of ONE larger tile already resident in device memory
 * (land-cover case, BASELINE.json configs[3]: a 4096 x 4096 x 4 tile -> 256 chips of 256 x 256).
 * Chips are numbered row-major over the (tile_h / chip) x (tile_w / chip) grid; chips
 * [first_chip, first_chip + n_chips) are preprocessed exactly as ai4e_preprocess would preprocess
 * their copies (same resize / crop / layout arguments, src size = chip x chip), but the gather is
 * address arithmetic inside the kernel: no chip is ever copied.  The reference has no such stage; a
 * PyTorch container slices the tile on the host before the model call at ai4e_service.py:203.
 * Bit-exact against oracle/preprocess_oracle.py (chips_from_tile + preprocess). */
AI4E_API int ai4e_preprocess_tile(const uint8_t* d_tile, int tile_h, int tile_w, int channels, int tile_row_stride,
                         int chip, int first_chip, int n_chips, int resize_h, int resize_w, int out_h, int out_w,
                         const float* h_mean, const float* h_std, int layout, int out_channels,
                         void* d_dst_bf16, void* stream);

/* K2/K4.  One convolution layer with folded BatchNorm, optional residual add and ReLU fused:
 *   out = act(conv(in, w) + bias [+ residual]).  Stands in for Conv2d + BatchNorm2d (+ add)
 * (+ ReLU) of torchvision's ResNet blocks.  `algo` is one of AI4E_CONV_*.  The tcgen05 path
 * needs cin % 64 == 0 and cout % 64 == 0. */
AI4E_API int ai4e_conv_layer(const void* d_in_nhwc, int batch, int h, int w, int cin,
                    const void* d_weight_krsc, const float* d_bias, int cout, int kh, int kw,
                    int stride, int pad, const void* d_residual_or_null, int relu,
                    void* d_out_nhwc, int algo, void* stream);

/* K0.  JPEG decode, baseline and progressive (requests carry encoded images; the reference
 * gates on image MIME types at APIs/1.0/base-py/ai4e_service.py:72,127-129 and leaves decoding to
 * the callback it calls at :203 - Pillow / libjpeg-turbo in a PyTorch container, which this
 * matches bit for bit: accurate integer IDCT, fancy chroma upsampling, fixed-point YCbCr->RGB).
 * Split like a hybrid decoder: the entropy (Huffman) decode is sequential and runs on the
 * submitting thread's host core; dequantisation, IDCT, upsampling and colour conversion run on the
 * GPU.  8-bit baseline / extended sequential and progressive (SOF0, SOF1, SOF2) Huffman, 1 or 3
 * components, 4:4:4 / 4:2:2 / 4:2:0 / 4:4:0 / 4:1:1 / 4:1:0 (chroma ratios 1, 2, 4 horizontally and 1, 2 vertically,
 * each component its own), restart intervals; a progressive file's scans accumulate into the
 * same record on the host.  Anything else (arithmetic coding, lossless, CMYK, 12-bit)
 * -> AI4E_ERR_UNSUPPORTED.
 *
 * A decoded image travels as a RECORD of int16: 3 x 64 quantisation values (natural order, per
 * component), then the quantised coefficients, component after component, block rows padded to
 * whole MCUs, 64 per block in natural order.  Images with equal geometry (ai4e_jpeg_info up to
 * `record_int16`) share one launch. */
typedef struct ai4e_jpeg_info {
  int32_t width, height, components;
  int32_t hmax, vmax;
  int32_t h[3], v[3];                 /* sampling factors */
  int32_t blocks_w[3], blocks_h[3];   /* 8x8 blocks per component, padded to whole MCUs */
  int32_t comp_w[3], comp_h[3];       /* real size of each component in samples */
  int32_t restart_interval;
  int64_t coef_off[3];                /* int16 offset of each component's coefficients in the record */
  int64_t record_int16;               /* int16 per record (multiple of 64) */
  int64_t plane_bytes;                /* scratch bytes per image for ai4e_jpeg_reconstruct */
} ai4e_jpeg_info;

/* Host only: parses the headers; with h_record != NULL also entropy-decodes into it
 * (capacity in int16, must be >= info.record_int16: call once with NULL to size it). */
AI4E_API int ai4e_jpeg_entropy_decode(const uint8_t* jpeg, uint64_t len, int16_t* h_record,
                                      uint64_t capacity_int16, ai4e_jpeg_info* out);

/* GPU: `batch` records of one geometry (device memory, record_int16 apart) -> dense uint8
 * [batch][height][width][3] RGB.  d_scratch: batch * plane_bytes. */
AI4E_API int ai4e_jpeg_reconstruct(const int16_t* d_records, int batch, const ai4e_jpeg_info* geometry,
                                   uint8_t* d_scratch, uint8_t* d_rgb_hwc, void* stream);

/* Convenience (tests, single images): entropy decode on the calling thread, H2D, reconstruct.
 * Scratch comes from the stream's memory pool. */
AI4E_API int ai4e_jpeg_decode(const uint8_t* jpeg, uint64_t len, uint8_t* d_rgb_hwc, uint64_t capacity_bytes,
                              int32_t* out_height, int32_t* out_width, void* stream);

/* PNG payloads.  PNG is lossless: chunk walk, zlib inflate and the scan-line filters
 * run on the calling thread's host core and there is no arithmetic for the GPU; the pixels then take
 * the ordinary submit path.  Output = `Image.open(BytesIO(data)).convert("RGB")` of Pillow, bit for
 * bit (the decode inside the callback at APIs/1.0/base-py/ai4e_service.py:203).  Every image type of the PNG
 * specification: grey (1 - 16 bits), RGB, palette (1 - 8 bits), with or without alpha, 8 or 16 bits per sample,
 * plain or Adam7-interlaced; 16-bit samples are reduced the way Pillow reduces them.  Anything else ->
 * AI4E_ERR_UNSUPPORTED.
 * h_rgb_hwc == NULL only reports the size. */
AI4E_API int ai4e_png_decode(const uint8_t* png, uint64_t len, uint8_t* h_rgb_hwc, uint64_t capacity_bytes,
                             int32_t* out_height, int32_t* out_width);

/* Admission check for an encoded payload, host only, headers only: AI4E_OK with the format and size
 * when the engine can decode it, AI4E_ERR_UNSUPPORTED otherwise (arithmetic-coded / CMYK JPEG, a file that is
 * neither JPEG nor PNG).  A request_processing_function calls it while the request is
 * still open and answers 415 instead of accepting a task that could only fail - the content gate
 * of APIs/1.0/base-py/ai4e_service.py:127-129 only looks at the MIME type. */
#define AI4E_IMAGE_UNKNOWN 0
#define AI4E_IMAGE_JPEG 1
#define AI4E_IMAGE_PNG 2
AI4E_API int ai4e_image_probe(const uint8_t* data, uint64_t len, int32_t* out_kind, int32_t* out_height,
                              int32_t* out_width);

/* K4b.  The tail of a ResNet bottleneck block in one kernel:
 *   y = relu(conv_khxkw(x; w2) + bias2)       [batch][p][q][cmid]   (conv2 + bn2 + relu, BN folded)
 *   z = relu(y * w3^T + bias3 + residual)     [batch][p][q][cout]   (conv3 1x1 + bn3 + add + relu)
 * y is rounded to bf16 exactly as if the two layers ran separately through ai4e_conv_layer, but
 * never leaves the SM.  cin % 64 == 0, cmid is 64 or 128, cout % 128 == 0; d_w2 is [cmid][kh][kw][cin],
 * d_w3 is [cout][cmid], residual and z are [batch][p][q][cout] bf16.  Stands in for the same
 * torchvision modules as ai4e_conv_layer (Bottleneck.conv2..relu, torchvision resnet.py), i.e. the
 * model the callback at APIs/1.0/base-py/ai4e_service.py:203 runs. */
AI4E_API int ai4e_bottleneck_tail(const void* d_in_nhwc, int batch, int h, int w, int cin, const void* d_w2_krsc,
                         const float* d_bias2, int cmid, int kh, int kw, int stride, int pad,
                         const void* d_w3_kc, const float* d_bias3, int cout, const void* d_residual,
                         void* d_out_nhwc, void* stream);

/* K2 + K3 in one kernel: the 7x7 stride-2 pad-3 stem convolution (4 input channels, 64 outputs, folded
 * BatchNorm, bias, optional ReLU) and the 3x3 stride-2 pad-1 max-pool behind it.  The stem's output -
 * the largest tensor of a ResNet - never reaches memory.  d_in is [batch][h][w][4] bf16 (the NHWC4
 * output of ai4e_preprocess), d_weight_krsc [64][7][7][4] bf16, d_out [batch][h/4][w/4][64] bf16;
 * w % 16 == 0, w <= 256, h % 4 == 0, else AI4E_ERR_UNSUPPORTED.  Bit-identical to ai4e_conv_layer
 * followed by ai4e_maxpool.  Stands in for conv1 + bn1 + relu + maxpool of torchvision's ResNet
 * inside the callback at APIs/1.0/base-py/ai4e_service.py:203; the engine uses it automatically. */
AI4E_API int ai4e_stem_pool(const void* d_in_nhwc4, int batch, int h, int w, const void* d_weight_krsc,
                   const float* d_bias, int relu, void* d_out_nhwc, void* stream);

/* K1 + K2 + K3 in one kernel: the same with the uint8 image as input.  d_images is the dense
 * [batch][h][w][3] uint8 batch; two warps of every CTA normalise the rows it needs through K1's
 * look-up table ((x/255 - mean)/std, bit-identical to ai4e_preprocess) while the tensor core works on
 * the previous ones, so the normalised bf16 image (401 KB per 224 x 224 image) is never written or read.
 * No resize: h, w are the model's input size.  Bit-identical to ai4e_preprocess (NHWC, 4 channels)
 * followed by ai4e_stem_pool.  Stands in for ToTensor + Normalize + conv1 + bn1 + relu + maxpool of the
 * callback at APIs/1.0/base-py/ai4e_service.py:203.  Measured slower than the two kernels at
 * micro-batch 64 (its two converter warps cannot keep up with the tensor core), so the engine uses it
 * only with AI4E_B200_U8_STEM=1; it saves a launch where latency matters more than throughput. */
AI4E_API int ai4e_preprocess_stem_pool(const uint8_t* d_images_hwc3, int batch, int h, int w, const float* h_mean,
                              const float* h_std, const void* d_weight_krsc, const float* d_bias, int relu,
                              void* d_out_nhwc, void* stream);

/* K3.  Max pooling (padding = -inf).  Stands in for torchvision's `maxpool`. */
AI4E_API int ai4e_maxpool(const void* d_in_nhwc, int batch, int h, int w, int c, int k, int stride, int pad,
                 void* d_out_nhwc, void* stream);

/* K5.  Global average pool -> FC -> softmax -> top-k.  Stands in for `avgpool`, `fc` and the
 * `softmax(...).topk(5)` of the callback.  d_in is [batch][hw][c] bf16, d_fc_w is
 * [classes][c] bf16.  d_logits ([batch][classes] fp32) is always written; top-k outputs
 * ([batch][topk]) may be NULL when topk == 0.  Allocates batch*c*4 bytes of scratch from the
 * stream's memory pool. */
AI4E_API int ai4e_head(const void* d_in_nhwc, int batch, int hw, int c, const void* d_fc_w,
              const float* d_fc_b, int classes, float* d_logits, int32_t* d_top_idx,
              float* d_top_prob, int topk, void* stream);

/* K5 for callers that must not allocate device memory: a request thread whose framework has
 * already given the whole GPU to its own pool allocator, or a latency-critical path that cannot
 * take the device-wide synchronisation of cudaMalloc / cudaFree.  Same arguments and results as
 * ai4e_head, but the batch*c floats of pooled-feature scratch live in a fixed array inside the
 * library (AI4E_HEAD_NOALLOC_MAX_FLOATS floats = 512 KiB, held by every GPU that loads the
 * library until exit).  batch*c above the limit -> AI4E_ERR_INVALID.  All calls on a device
 * share the array: calls that overlap on the GPU race, so keep them on one stream or order
 * them with an event.  The first call on each device resolves the array's address; make it
 * before any stream capture on that device.  This entry point has no counterpart in the
 * reference. */
#define AI4E_HEAD_NOALLOC_MAX_FLOATS (64 * 2048)
AI4E_API int ai4e_head_noalloc(const void* d_in_nhwc, int batch, int hw, int c, const void* d_fc_w,
                               const float* d_fc_b, int classes, float* d_logits, int32_t* d_top_idx,
                               float* d_top_prob, int topk, void* stream);

/* ---------------------------------------------------------------------------------------
 * Engine: one model replica on one GPU (the reference runs one replica per pod).
 * ------------------------------------------------------------------------------------- */
typedef struct ai4e_engine ai4e_engine;

typedef struct ai4e_layer_desc {
  int32_t op;      /* AI4E_OP_* */
  int32_t in_buf;  /* activation slot read  (slot 0 = preprocessed input) */
  int32_t out_buf; /* activation slot written (ignored for AI4E_OP_HEAD) */
  int32_t res_buf; /* slot added before the activation, or -1 */
  int32_t cin, cout;
  int32_t kh, kw, stride, pad;
  int32_t relu;
  int32_t reserved;
  int64_t w_off; /* byte offset of the bf16 weights in the blob (256-byte aligned) */
  int64_t b_off; /* byte offset of the fp32 bias in the blob (256-byte aligned) */
} ai4e_layer_desc;

typedef struct ai4e_model_desc {
  int32_t n_layers;
  int32_t n_bufs;       /* number of activation slots the layer list refers to */
  int32_t in_h, in_w;   /* network input size (after resize/crop) */
  int32_t in_c;         /* image channels (3 or 4) */
  int32_t in_c_pad;     /* channels of slot 0 (stem Cin; 4) */
  int32_t num_classes;
  int32_t reserved;     /* bit 0: keep every layer's output in memory (stage read-back): no cross-layer fusion */
  float mean[4], std[4];
  const ai4e_layer_desc* layers;
} ai4e_model_desc;

typedef struct ai4e_engine_stats {
  uint64_t images, batches;           /* completed through submit/wait or forward */
  uint64_t batch_hist[11];            /* batches of size <=1,2,4,...,512,>512 */
  uint64_t queue_depth;               /* submitted, not yet completed */
  double gpu_ms_total;                /* device time of all forwards (CUDA events) */
  double lat_ms_p50, lat_ms_p99;      /* submit -> result ready, last 4096 tasks */
  int32_t healthy;                    /* 0 after a CUDA error on this replica */
  int32_t launches_per_forward;       /* kernels one forward launches */
} ai4e_engine_stats;

/* Builds a replica: copies the weight blob to the device, sizes activation slots for
 * `max_batch`, starts the micro-batching thread.  Stands in for model construction +
 * `model.eval()` at container start-up. */
AI4E_API int ai4e_engine_create(int device, const ai4e_model_desc* desc, const void* h_weights_blob,
                       uint64_t blob_bytes, int max_batch, int max_wait_us, ai4e_engine** out);
AI4E_API int ai4e_engine_destroy(ai4e_engine* e);

/* One synchronous batch from HOST memory: H2D copy, preprocess (resize to the model input if
 * src size differs), forward, top-5, D2H.  h_images is dense [batch][src_h][src_w][in_c].
 * h_logits_or_null: [batch][num_classes].  This is `model(batch)` of the callback. */
AI4E_API int ai4e_engine_forward_host(ai4e_engine* e, const uint8_t* h_images, int batch, int src_h,
                             int src_w, int32_t* h_top_idx, float* h_top_prob,
                             float* h_logits_or_null);

/* Same with the uint8 batch already resident on the engine's device and results left on the
 * device (no host copies); returns after the work is enqueued on `stream`.  The engine owns
 * AI4E_B200_LANES (default 2) sets of activation buffers and hands them out round-robin: calls
 * on DIFFERENT streams overlap on the GPU (the second micro-batch fills the SMs the first leaves
 * idle), calls on one stream run in order.  Inputs and outputs of calls that may overlap must be
 * distinct buffers. */
AI4E_API int ai4e_engine_forward_device(ai4e_engine* e, const uint8_t* d_images, int batch, int src_h,
                               int src_w, int32_t* d_top_idx, float* d_top_prob,
                               float* d_logits_or_null, void* stream);

/* Micro-batched path.  `submit` copies one image into a pinned staging slot and returns a
 * ticket at once (the caller's buffer is free on return); the engine's batcher closes a batch
 * at `max_batch` images or `max_wait_us` after the first one.  Thread-safe; call with the GIL
 * released.  `wait` blocks until the ticket's batch has finished.  Together they replace the
 * per-task thread + per-image forward of ai4e_service.py:180-183,203. */
AI4E_API int ai4e_engine_submit(ai4e_engine* e, const uint8_t* h_hwc, int h, int w, int c, int row_stride,
                       uint64_t tag, uint64_t* out_ticket);
/* `n` images of one size in ONE call (one lock acquisition per micro-batch instead of per image, one
 * trip through the caller's FFI).  h_images[i] points at image i (rows `row_stride` bytes apart,
 * 0 = dense); tags_or_null[i] is an opaque caller value (a task index) handed back by
 * ai4e_engine_poll.  With notify != 0 every ticket is ANNOUNCED when its batch finishes: one
 * reaper thread drains ai4e_engine_poll instead of one blocked thread per task in `wait` - the
 * thread-per-task model of ai4e_service.py:180-183 turned into a queue.  row_stride must not be
 * negative (copy a flipped view first). */
AI4E_API int ai4e_engine_submit_many(ai4e_engine* e, const uint8_t* const* h_images, int n, int h, int w, int c,
                            int row_stride, const uint64_t* tags_or_null, int notify, uint64_t* out_tickets);
/* Up to `capacity` announced completions, oldest first; blocks up to timeout_us for the first one
 * (< 0: for ever, 0: never).  Any output array may be NULL.  status[i] is AI4E_OK or the error that
 * failed the ticket's batch; top_idx5 / top_prob5 are [capacity][5].  *out_n may be 0 (timeout).
 * Records stay queued until somebody polls them (40 bytes each): submit with notify only what a reaper
 * will collect.  The replicas of a pool announce into the POOL's queue: poll the pool, not its engines. */
AI4E_API int ai4e_engine_poll(ai4e_engine* e, int capacity, int64_t timeout_us, uint64_t* tickets, uint64_t* tags,
                     int32_t* status, int32_t* top_idx5, float* top_prob5, int* out_n);
/* A caller-made batch (e.g. the 256 chips of a land-cover tile, or a dispatcher that already
 * holds a full micro-batch): `count` dense images become one batch of their own, tickets
 * first..first+count-1.  If h_images is page-locked (cudaHostAlloc / cudaHostRegister /
 * torch pin_memory) the H2D copy reads it in place - keep it unchanged until the last ticket
 * has been waited for; pageable memory is copied to the staging slot before returning. */
AI4E_API int ai4e_engine_submit_batch(ai4e_engine* e, const uint8_t* h_images, int count, int h, int w,
                                      int c, uint64_t* out_first_ticket);
/* Same for one JPEG, baseline or progressive (K0): headers and the entropy decode run on the calling thread,
 * the record joins the open micro-batch (a batch holds images of one size and sampling) and the
 * GPU reconstructs the pixels in front of the preprocessing kernel.  The engine's model must
 * take 3 channels; unsupported streams (arithmetic coding, CMYK, ...) -> AI4E_ERR_UNSUPPORTED and
 * nothing is queued.  This is `Image.open(BytesIO(request.data)).convert("RGB")` + the rest of
 * the callback at APIs/1.0/base-py/ai4e_service.py:203. */
AI4E_API int ai4e_engine_submit_jpeg(ai4e_engine* e, const uint8_t* jpeg, uint64_t len, uint64_t tag,
                                     uint64_t* out_ticket);
/* An encoded payload of either supported format (sniffed from its first bytes): JPEG as
 * ai4e_engine_submit_jpeg, PNG decoded on the calling thread and submitted as pixels. */
AI4E_API int ai4e_engine_submit_image(ai4e_engine* e, const uint8_t* data, uint64_t len, uint64_t tag,
                                      uint64_t* out_ticket);
/* A tile cut into chip x chip chips (row-major over the chip grid; the remainder at the right /
 * bottom is ignored), in one call: tile -> chips -> classify of the land-cover case (BASELINE.json
 * configs[3]; the reference chains such stages through AddPipelineTask,
 * Common/task_management/distributed_api_task.py:67-100, one HTTP hop per stage).  The tile crosses
 * PCIe once (one 2-D H2D copy into a device tile buffer) and K1 reads the chips in place
 * (ai4e_preprocess_tile): no per-chip copy on the host or the device.  A page-locked tile is copied
 * asynchronously - keep it unchanged until the last ticket is done; a pageable one is free on
 * return.  out_tickets: one per chip, `capacity` >= (tile_h/chip)*(tile_w/chip). */
AI4E_API int ai4e_engine_submit_tile(ai4e_engine* e, const uint8_t* h_tile, int tile_h, int tile_w, int c,
                                     int row_stride, int chip, uint64_t* out_tickets, int capacity,
                                     int* out_count);
/* K7.  tile -> chips -> classify -> AGGREGATE, all on the device (the reference would chain an
 * aggregation stage as one more task through AddPipelineTask, Common/task_management/distributed_api_task.py:67-100).
 * Like ai4e_engine_submit_tile, but behind every batch's head one small launch folds the chips' top-5 into the
 * tile's summary in device memory, and ONE ticket stands for the whole tile:
 *   labels[i], label_prob[i]  top-1 class / probability of chip i (row-major chip grid)  - the tile's label map
 *   hist[k]                   chips whose top-1 class is k
 *   mass_q32[k]               sum over chips and their top-5 entries of class k of round(p * 2^32): the probability
 *                             mass in 32.32 fixed point (integer sums: independent of batch order and replica count)
 * wait_tile_summary blocks like ai4e_engine_wait (AI4E_B200_WAIT_TIMEOUT_MS); any output may be NULL; labels /
 * label_prob need capacity_chips >= the tile's chips, hist / mass_q32 num_classes entries.  The last kSumRing = 64
 * summaries are kept.  No per-chip record crosses PCIe for such a tile except the 40-byte top-5 the head always
 * returns. */
AI4E_API int ai4e_engine_submit_tile_summary(ai4e_engine* e, const uint8_t* h_tile, int tile_h, int tile_w, int c,
                                             int row_stride, int chip, uint64_t* out_tile_ticket, int* out_chips);
AI4E_API int ai4e_engine_wait_tile_summary(ai4e_engine* e, uint64_t tile_ticket, int32_t* labels, float* label_prob,
                                           int capacity_chips, uint32_t* hist, uint64_t* mass_q32, int* out_chips);
/* The aggregate as a stand-alone device op: folds the top-5 of `count` chips (chips [first, first + count) of the
 * summary) into labels / label_prob / hist / mass_q32, all device pointers; hist and mass_q32 are added to (zero
 * them before the first call of a tile).  No reference counterpart: the reference has no arithmetic. */
AI4E_API int ai4e_tile_summary(const int32_t* d_top_idx5, const float* d_top_prob5, int count, int first, int num_classes,
                               int32_t* d_labels, float* d_label_prob, uint32_t* d_hist, uint64_t* d_mass_q32,
                               void* stream);
/* ai4e_engine_wait for n tickets: top_idx5 / top_prob5 are [n][5]. */
AI4E_API int ai4e_engine_wait_many(ai4e_engine* e, const uint64_t* tickets, int n, int32_t* top_idx5,
                                   float* top_prob5);
/* Blocks until the ticket's batch has finished, at most AI4E_B200_WAIT_TIMEOUT_MS (default 60 s):
 * then AI4E_ERR_TIMEOUT, and the ticket may be waited for again. */
AI4E_API int ai4e_engine_wait(ai4e_engine* e, uint64_t ticket, int32_t* top_idx5, float* top_prob5,
                     float* logits_or_null);
/* Same with an explicit limit: timeout_us < 0 waits for ever, 0 only reports a finished ticket. */
AI4E_API int ai4e_engine_wait_for(ai4e_engine* e, uint64_t ticket, int64_t timeout_us, int32_t* top_idx5,
                         float* top_prob5, float* logits_or_null);
AI4E_API int ai4e_engine_flush(ai4e_engine* e);
AI4E_API int ai4e_engine_stats_get(ai4e_engine* e, ai4e_engine_stats* out);

/* Measurement: runs `reps` forwards on a device-resident batch with a CUDA event between
 * kernels and returns the mean device time (ms) of the preprocess kernel (entry 0) and of
 * every layer of the program (entries 1..n_layers).  bench.py derives the per-kernel roofline
 * from it.  No reference counterpart (the reference has no profile). */
AI4E_API int ai4e_engine_profile(ai4e_engine* e, const uint8_t* d_images, int batch, int src_h, int src_w,
                                 int reps, float* h_ms_out, int n_out);

/* Fault drill: the next `n_batches` micro-batches fail as if their launch had hit a sticky CUDA
 * error - their tickets fail with AI4E_ERR_CUDA, the replica turns unhealthy and a pool routes
 * around it (SURVEY 5's failure model; the shell turns the error into
 * FailTask(taskId, 'Task failed - try again'), ai4e_service.py:185-195).  No reference counterpart. */
AI4E_API int ai4e_engine_debug_inject_fault(ai4e_engine* e, int n_batches);

/* Debug/parity: copy activation slot `buf` of the last forward to the host (bf16 NHWC). */
AI4E_API int ai4e_engine_read_activation(ai4e_engine* e, int buf, void* h_dst, uint64_t bytes);

/* ---------------------------------------------------------------------------------------
 * Pool: N independent replicas, one per GPU, fed by one dispatcher (least-loaded routing).
 * The reference's equivalent is Istio ROUND_ROBIN over pods (APIs/Charts/templates/routing.yml:20-28).
 * No collective is involved.
 * ------------------------------------------------------------------------------------- */
typedef struct ai4e_pool ai4e_pool;
AI4E_API int ai4e_pool_create(const int* devices, int n_devices, const ai4e_model_desc* desc,
                     const void* h_weights_blob, uint64_t blob_bytes, int max_batch, int max_wait_us,
                     ai4e_pool** out);
AI4E_API int ai4e_pool_destroy(ai4e_pool* p);
AI4E_API int ai4e_pool_submit(ai4e_pool* p, const uint8_t* h_hwc, int h, int w, int c, int row_stride,
                     uint64_t tag, uint64_t* out_ticket);
AI4E_API int ai4e_pool_submit_jpeg(ai4e_pool* p, const uint8_t* jpeg, uint64_t len, uint64_t tag,
                                   uint64_t* out_ticket);
AI4E_API int ai4e_pool_submit_image(ai4e_pool* p, const uint8_t* data, uint64_t len, uint64_t tag,
                                    uint64_t* out_ticket);
/* The ai4e_engine_* calls of the same name behind the dispatcher.  submit_many deals the images to
 * the replicas in runs, submit_batch gives the whole batch to one replica (tickets consecutive),
 * submit_tile deals contiguous chip ranges, one per healthy replica (256 chips on 8 GPUs: 32 each)
 * and sends each replica only the band of chip rows it needs; poll drains ONE queue that all
 * replicas announce into. */
AI4E_API int ai4e_pool_submit_many(ai4e_pool* p, const uint8_t* const* h_images, int n, int h, int w, int c,
                          int row_stride, const uint64_t* tags_or_null, int notify, uint64_t* out_tickets);
AI4E_API int ai4e_pool_submit_batch(ai4e_pool* p, const uint8_t* h_images, int count, int h, int w, int c,
                           uint64_t* out_first_ticket);
AI4E_API int ai4e_pool_submit_tile(ai4e_pool* p, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride,
                          int chip, uint64_t* out_tickets, int capacity, int* out_count);
/* K7 behind the dispatcher: every replica summarises its own chip range on its GPU; wait puts the ranges together
 * (labels side by side, hist and mass_q32 added - integer sums, the same result for any number of replicas). */
AI4E_API int ai4e_pool_submit_tile_summary(ai4e_pool* p, const uint8_t* h_tile, int tile_h, int tile_w, int c,
                                           int row_stride, int chip, uint64_t* out_tile_ticket, int* out_chips);
AI4E_API int ai4e_pool_wait_tile_summary(ai4e_pool* p, uint64_t tile_ticket, int32_t* labels, float* label_prob,
                                         int capacity_chips, uint32_t* hist, uint64_t* mass_q32, int* out_chips);
/* Measurement aid, no reference counterpart (BASELINE.json configs[4]: Poisson arrivals at a fraction of peak): `feeders`
 * native threads offer `rate_per_s` arrivals for `seconds` through ai4e_pool_submit_many (every arrival of a feeder that
 * is due per call, notify) and one thread drains ai4e_pool_poll - nobody else may poll this pool meanwhile.  h_images:
 * n_images dense h x w x c images, cycled.  out_latency_ms[i] = announced - SCHEDULED time of arrival i in ms (-1: not
 * completed or failed), i < *out_offered <= capacity; *out_span_s = when the last result arrived. */
AI4E_API int ai4e_pool_openloop(ai4e_pool* p, const uint8_t* h_images, int n_images, int h, int w, int c, double rate_per_s,
                                double seconds, int feeders, uint64_t seed, float* out_latency_ms, int64_t capacity,
                                int64_t* out_offered, int64_t* out_completed, int64_t* out_failed, double* out_span_s,
                                double* out_feeder_lateness_ms);
AI4E_API int ai4e_pool_wait_many(ai4e_pool* p, const uint64_t* tickets, int n, int32_t* top_idx5, float* top_prob5);
AI4E_API int ai4e_pool_poll(ai4e_pool* p, int capacity, int64_t timeout_us, uint64_t* tickets, uint64_t* tags,
                   int32_t* status, int32_t* top_idx5, float* top_prob5, int* out_n);
AI4E_API int ai4e_pool_wait(ai4e_pool* p, uint64_t ticket, int32_t* top_idx5, float* top_prob5,
                   float* logits_or_null);
AI4E_API int ai4e_pool_wait_for(ai4e_pool* p, uint64_t ticket, int64_t timeout_us, int32_t* top_idx5,
                       float* top_prob5, float* logits_or_null);
AI4E_API int ai4e_pool_flush(ai4e_pool* p);
AI4E_API int ai4e_pool_size(ai4e_pool* p);
AI4E_API int ai4e_pool_engine(ai4e_pool* p, int i, ai4e_engine** out);

#ifdef __cplusplus
}
#endif
#endif /* AI4E_B200_H_ */
