// Engine: one model replica on one GPU, plus the micro-batching request queue and the
// multi-GPU pool (extern "C" entry points `ai4e_engine_*` / `ai4e_pool_*`).
//
// What it replaces in the reference: the per-task thread + per-image model call of
// APIs/1.0/base-py/ai4e_service.py:180-183,203 (one `threading.Thread` per task, no batching;
// upstream pushes one message at a time, ProcessManager/BackendQueueProcessor/host.json:5-9).
// Here request threads copy their payload straight into a pinned staging slot and get a
// ticket; a batcher thread closes the batch on `max_batch` or `max_wait_us`, issues
// H2D -> preprocess -> forward -> top-5 -> D2H on the batch's own stream, and a completer
// thread publishes results.  Three batch contexts are in flight so the H2D copy of batch k+1
// overlaps the forward of batch k (forwards themselves are chained by an event because they
// share the activation slots).
#include "../../include/ai4e_b200.h"

#include "common.cuh"
#include "kernels.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

using namespace ai4e;
using Clock = std::chrono::steady_clock;

namespace {

constexpr int kNumCtx = 4;
constexpr int kNumTileBufs = 3;
constexpr int kKindTile = -1;  // BatchCtx::kind of a batch whose images are chips of a device-resident tile
constexpr uint64_t kResultRing = 1ull << 16;
constexpr uint64_t kLogitRing = 1ull << 12;
constexpr int kLatRing = 4096;

struct LayerRt {
  ai4e_layer_desc d;
  int in_h, in_w, out_h, out_w;
  bool igemm;
  IgemmPlan plan;
  bool fused_first = false;  // this layer and the next run as one bottleneck_tail_kernel (fplan)
  bool fused_second = false; // computed by the previous layer's launch
  FusedPlan fplan;
  bool stem_pool = false;    // this stem convolution and the max-pool behind it run as one stem_pool_kernel (splan)
  StemPoolPlan splan;
  // side branch (a block's downsample convolution): runs on the engine's second stream next to the
  // layers that follow it; `wait_ev` makes its first consumer wait for it
  bool side = false;
  bool reverse = false;      // walks its row tiles last-to-first (opposite to the producer of its input)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, wait_ev = nullptr;
};

// A tile (or the band of chip rows one replica needs of it) resident in device memory: it crosses
// PCIe once and K1 reads the chips in place through the chip-grid strides (K6).
struct TileBuf {
  uint8_t* d = nullptr;
  size_t bytes = 0;
  cudaEvent_t ev_ready = nullptr;  // the H2D copy of the band has finished
  int users = 0;                   // batches that still read it (guarded by ai4e_engine::mu)
  bool busy = false;               // being filled by a submitter
  // K7, summary submits only: the tile's aggregate is accumulated in d_sum by one small launch behind every batch's
  // head and read back once, by the completer, when the tile's last batch has finished.
  // Layout of d_sum / h_sum: labels int32[n] | label_prob float[n] | hist uint32[K] | (8-byte aligned) mass u64[K]
  uint8_t* d_sum = nullptr;
  uint8_t* h_sum = nullptr;        // page-locked
  size_t sum_cap = 0;
  bool summary = false;
  int sum_chips = 0, sum_left = 0, sum_rc = 0;  // chips of the summary, batches not yet completed, first error
  uint64_t sum_ticket = 0;
};

struct SumLayout {
  size_t labels, prob, hist, mass, bytes;
  SumLayout(int n, int K) {
    labels = 0; prob = (size_t)n * 4; hist = prob + (size_t)n * 4;
    mass = (hist + (size_t)K * 4 + 7) & ~(size_t)7;
    bytes = mass + (size_t)K * 8;
  }
};

// A finished (or pending) tile summary, kept until kSumRing newer ones have been submitted.
constexpr int kSumRing = 64;
struct TileSumRec {
  uint64_t id = ~0ull;
  int status = 1;  // 1 = pending, AI4E_OK, or the error that failed one of its batches
  int n = 0;
  std::vector<uint8_t> data;  // SumLayout(n, num_classes)
};

// One finished ticket that asked to be announced (submit_many with notify): what ai4e_*_poll returns.
struct DoneRec {
  uint64_t ticket, tag;
  int32_t status;
  int32_t idx[AI4E_TOPK];
  float prob[AI4E_TOPK];
};
struct CompletionQueue {
  std::mutex mu;
  std::condition_variable cv;
  std::deque<DoneRec> q;
};

struct BatchCtx {
  uint8_t* h_in = nullptr;  // pinned, max_batch * slot_bytes
  uint8_t* d_in = nullptr;
  int32_t *h_idx = nullptr, *d_idx = nullptr;
  float *h_prob = nullptr, *d_prob = nullptr;
  float *h_logits = nullptr, *d_logits = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_start = nullptr, ev_fwd = nullptr, ev_done = nullptr;
  int count = 0, src_h = 0, src_w = 0;
  uint64_t first_ticket = 0;
  std::atomic<int> copied{0};
  Clock::time_point t_first;
  int launch_rc = 0;
  const uint8_t* ext_src = nullptr;  // zero-copy source (page-locked caller memory), else h_in
  // JPEG batches: h_in holds entropy-decoded records of one geometry instead of pixels
  int kind = 0;                      // 0 = raw pixels; else a key of the JPEG geometry (see jpeg_kind)
  ai4e_jpeg_info jinfo{};
  int16_t* d_rec = nullptr;          // device copy of the records (3-channel models only)
  uint8_t* d_planes = nullptr;       // component planes between the IDCT and the colour kernel
  // tile batches (kind == kKindTile): chips [tile_first, tile_first + count) of the band in `tile`
  TileBuf* tile = nullptr;
  int tile_first = 0, tile_grid_w = 0, tile_chip = 0;
  long long tile_row_bytes = 0;
  int sum_first = 0;                 // summary submits: where this batch's chips sit in the summary
};

}  // namespace

// A lane = one set of activation slots with the tensor maps that point into them.  Forwards that
// use different lanes run concurrently (each on its caller's / batch context's stream): at
// micro-batch 64 many layers leave SMs idle (25-98 M tiles, launch fill and epilogue tails), and a
// second micro-batch in flight fills them.  Forwards on the same lane are chained by an event.
struct Lane {
  std::vector<LayerRt> layers;
  std::vector<void*> bufs;
  float* d_pooled = nullptr;
  float* d_logits = nullptr;  // where the logits go when the caller of forward_device passes no buffer
  cudaEvent_t ev_chain = nullptr;
  bool chain_valid = false;
  bool u8_stem = false;       // layer 0 is the fused stem + max-pool and may take the uint8 batch itself (K1 fused in)
  cudaStream_t side_stream = nullptr;
};

struct ai4e_engine {
  int device = 0, max_batch = 0, max_wait_us = 0;
  ai4e_model_desc desc{};
  std::vector<LayerRt> layers;  // validated program (shapes, igemm flag); each lane has its own copy with plans
  std::vector<Lane> lanes;
  std::vector<size_t> buf_elems_per_image;
  uint8_t* d_blob = nullptr;
  uint16_t* d_stem_w = nullptr;  // stem weights repacked for the tensor-core stem
  uint16_t* d_stem_pool_w = nullptr;  // ... and for the fused stem + max-pool kernel
  size_t slot_bytes = 0;  // staging bytes per image
  int launches = 0;       // kernels one forward launches
  int launches_u8 = 0;    // ... when the batch is raw pixels of the model's input size (K1 runs inside the stem); 0 = never

  // lanes are handed out round-robin under launch_mu
  std::mutex launch_mu;
  unsigned next_lane = 0;
  int last_lane = 0;  // lane of the most recent forward (ai4e_engine_read_activation reads it)

  // synchronous path
  std::mutex sync_mu;
  BatchCtx sync_ctx;

  // micro-batcher
  BatchCtx ctx[kNumCtx];
  std::mutex mu;
  std::condition_variable cv_work, cv_free, cv_done, cv_inflight;
  std::deque<BatchCtx*> free_q, closed_q, inflight_q;
  BatchCtx* open_ctx = nullptr;
  uint64_t next_ticket = 0, done_upto = 0;
  bool stopping = false, flush_req = false, batcher_done = false, completer_done = false;
  std::atomic<int> healthy{1};
  std::atomic<int> inject_faults{0};          // debug hook: fail this many upcoming batches like a CUDA fault
  std::atomic<uint64_t> n_submitted{0}, n_completed{0};  // queue depth without taking `mu` (pool routing)
  std::thread batcher, completer;
  int wait_timeout_ms = 60000;                // ai4e_engine_wait gives up after this (AI4E_B200_WAIT_TIMEOUT_MS)

  // device-resident tiles (K6)
  TileBuf tiles[kNumTileBufs];
  cudaStream_t copy_stream = nullptr;
  cudaStream_t sum_stream = nullptr;          // the completer's read-back of a finished tile summary
  TileSumRec sums[kSumRing];
  uint64_t next_sum_ticket = 0;

  // announced completions (ai4e_engine_poll); a pool points all its replicas at one shared queue
  CompletionQueue own_cq;
  CompletionQueue* cq = &own_cq;
  uint64_t ticket_prefix = 0;                 // replica index << 56 inside a pool

  // results
  std::vector<int32_t> r_idx;
  std::vector<float> r_prob, r_logits;
  std::vector<int8_t> r_status;
  std::vector<uint8_t> r_notify;
  std::vector<uint64_t> r_tag;
  std::vector<Clock::time_point> r_submit;
  std::vector<float> lat_ms;
  uint64_t lat_n = 0;
  ai4e_engine_stats st{};
};

namespace {

// two JPEGs may share a launch only if their records have the same layout, component by component
// (equal totals can hide different per-component sampling)
bool jpeg_same_geometry(const ai4e_jpeg_info& a, const ai4e_jpeg_info& b) {
  if (a.width != b.width || a.height != b.height || a.components != b.components || a.hmax != b.hmax ||
      a.vmax != b.vmax || a.record_int16 != b.record_int16 || a.plane_bytes != b.plane_bytes)
    return false;
  for (int c = 0; c < 3; ++c)
    if (a.h[c] != b.h[c] || a.v[c] != b.v[c] || a.blocks_w[c] != b.blocks_w[c] || a.blocks_h[c] != b.blocks_h[c] ||
        a.comp_w[c] != b.comp_w[c] || a.comp_h[c] != b.comp_h[c] || a.coef_off[c] != b.coef_off[c])
      return false;
  return true;
}

int alloc_ctx(ai4e_engine* e, BatchCtx& c) {
  const size_t B = e->max_batch;
  const int K = e->desc.num_classes;
  AI4E_CUDA_TRY(cudaHostAlloc((void**)&c.h_in, B * e->slot_bytes, cudaHostAllocDefault));
  AI4E_CUDA_TRY(cudaMalloc((void**)&c.d_in, B * e->slot_bytes));
  if (e->desc.in_c == 3) {
    // JPEG batches: records and component planes are bounded by the staging slot (checked at submit),
    // so they are sized once here - no cudaMalloc / cudaFree (device-wide syncs) on the batcher thread
    AI4E_CUDA_TRY(cudaMalloc((void**)&c.d_rec, B * e->slot_bytes));
    AI4E_CUDA_TRY(cudaMalloc((void**)&c.d_planes, B * e->slot_bytes));
  }
  AI4E_CUDA_TRY(cudaHostAlloc((void**)&c.h_idx, B * AI4E_TOPK * sizeof(int32_t), cudaHostAllocDefault));
  AI4E_CUDA_TRY(cudaHostAlloc((void**)&c.h_prob, B * AI4E_TOPK * sizeof(float), cudaHostAllocDefault));
  AI4E_CUDA_TRY(cudaHostAlloc((void**)&c.h_logits, B * K * sizeof(float), cudaHostAllocDefault));
  AI4E_CUDA_TRY(cudaMalloc((void**)&c.d_idx, B * AI4E_TOPK * sizeof(int32_t)));
  AI4E_CUDA_TRY(cudaMalloc((void**)&c.d_prob, B * AI4E_TOPK * sizeof(float)));
  AI4E_CUDA_TRY(cudaMalloc((void**)&c.d_logits, B * K * sizeof(float)));
  AI4E_CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  AI4E_CUDA_TRY(cudaEventCreate(&c.ev_start));
  AI4E_CUDA_TRY(cudaEventCreate(&c.ev_fwd));
  AI4E_CUDA_TRY(cudaEventCreateWithFlags(&c.ev_done, cudaEventDisableTiming));
  return AI4E_OK;
}

void free_ctx(BatchCtx& c) {
  if (c.h_in) cudaFreeHost(c.h_in);
  if (c.d_in) cudaFree(c.d_in);
  if (c.d_rec) cudaFree(c.d_rec);
  if (c.d_planes) cudaFree(c.d_planes);
  c.d_rec = nullptr; c.d_planes = nullptr;
  if (c.h_idx) cudaFreeHost(c.h_idx);
  if (c.h_prob) cudaFreeHost(c.h_prob);
  if (c.h_logits) cudaFreeHost(c.h_logits);
  if (c.d_idx) cudaFree(c.d_idx);
  if (c.d_prob) cudaFree(c.d_prob);
  if (c.d_logits) cudaFree(c.d_logits);
  if (c.stream) cudaStreamDestroy(c.stream);
  if (c.ev_start) cudaEventDestroy(c.ev_start);
  if (c.ev_fwd) cudaEventDestroy(c.ev_fwd);
  if (c.ev_done) cudaEventDestroy(c.ev_done);
  c.h_in = c.d_in = nullptr;
  c.h_idx = c.d_idx = nullptr;
  c.h_prob = c.d_prob = c.h_logits = c.d_logits = nullptr;
  c.stream = nullptr;
  c.ev_start = c.ev_fwd = c.ev_done = nullptr;
}

// preprocess + every layer, enqueued on `stream`.  Caller holds e->launch_mu.
int forward_enqueue(ai4e_engine* eng, Lane* e, const uint8_t* d_images, int batch, int src_h, int src_w,
                    int32_t* d_idx, float* d_prob, float* d_logits, cudaStream_t stream,
                    cudaEvent_t* marks = nullptr, const BatchCtx* jpeg = nullptr, const ChipGrid* chips = nullptr) {
  const ai4e_model_desc& m = eng->desc;
  int mark = 0;
  if (marks) AI4E_CUDA_TRY(cudaEventRecord(marks[mark++], stream));
  const bool resize = src_h != m.in_h || src_w != m.in_w;
  int rc = AI4E_OK;
  // K1 inside the stem: a raw 3-channel batch of the model's input size needs no preprocessing launch at
  // all when the first layer is the fused stem + max-pool kernel (its converter warps apply K1's table)
  const bool u8_stem = e->u8_stem && !jpeg && !chips && !resize;
  if (u8_stem) {
    // nothing: the stem reads d_images
  } else if (jpeg && !resize && m.in_c == 3 && m.in_c_pad == 4) {
    // K0 fused into K1: the JPEG colour kernel writes the normalised bf16 NHWC4 pixels itself
    const float ms[6] = {m.mean[0], m.mean[1], m.mean[2], m.std[0], m.std[1], m.std[2]};
    rc = jpeg_reconstruct(jpeg->d_rec, batch, &jpeg->jinfo, jpeg->d_planes, (uint8_t*)e->bufs[0], stream, ms);
  } else {
    if (jpeg) {  // other sizes: pixels first, then K1's resize + centre crop
      rc = jpeg_reconstruct(jpeg->d_rec, batch, &jpeg->jinfo, jpeg->d_planes, const_cast<uint8_t*>(d_images), stream);
      if (rc) return rc;
    }
    rc = launch_preprocess(d_images, batch, src_h, src_w, m.in_c, 0, chips ? (int)chips->row_bytes : 0, resize ? m.in_h : 0,
                           resize ? m.in_w : 0, m.in_h, m.in_w, m.mean, m.std, AI4E_LAYOUT_NHWC,
                           m.in_c_pad, e->bufs[0], stream, chips);
  }
  if (rc < 0) return rc;
  if (marks) AI4E_CUDA_TRY(cudaEventRecord(marks[mark++], stream));
  for (LayerRt& L : e->layers) {
    const ai4e_layer_desc& d = L.d;
    const bool branch = !marks && e->side_stream;  // per-layer timing runs everything in order on one stream
    if (branch && L.wait_ev) AI4E_CUDA_TRY(cudaStreamWaitEvent(stream, L.wait_ev, 0));
    if (branch && L.side) {
      AI4E_CUDA_TRY(cudaEventRecord(L.ev_fork, stream));
      AI4E_CUDA_TRY(cudaStreamWaitEvent(e->side_stream, L.ev_fork, 0));
      rc = igemm_launch(L.plan, batch, 0, e->side_stream);
      if (rc) return rc;
      AI4E_CUDA_TRY(cudaEventRecord(L.ev_join, e->side_stream));
      continue;
    }
    if (d.op == AI4E_OP_CONV) {
      if (L.fused_second) {
        rc = AI4E_OK;
      } else if (L.stem_pool) {
        rc = (u8_stem && &L == &e->layers.front()) ? stem_pool_launch_u8(L.splan, d_images, batch, m.mean, m.std, stream)
                                                   : stem_pool_launch(L.splan, batch, stream);
      } else if (L.fused_first) {
        rc = fused_tail_launch(L.fplan, batch, stream);
      } else if (L.igemm) {
        rc = igemm_launch(L.plan, batch, 0, stream);
      } else {
        rc = launch_conv_direct(e->bufs[d.in_buf], batch, L.in_h, L.in_w, d.cin, eng->d_blob + d.w_off,
                                (const float*)(eng->d_blob + d.b_off), d.cout, d.kh, d.kw, d.stride, d.pad,
                                d.res_buf >= 0 ? e->bufs[d.res_buf] : nullptr, d.relu, e->bufs[d.out_buf],
                                stream);
      }
    } else if (d.op == AI4E_OP_MAXPOOL) {
      rc = L.fused_second ? AI4E_OK
                          : launch_maxpool(e->bufs[d.in_buf], batch, L.in_h, L.in_w, d.cin, d.kh, d.stride, d.pad,
                                           e->bufs[d.out_buf], stream, L.reverse ? 1 : 0);
    } else {
      rc = launch_head(e->bufs[d.in_buf], batch, L.in_h * L.in_w, d.cin, eng->d_blob + d.w_off,
                       (const float*)(eng->d_blob + d.b_off), d.cout, e->d_pooled, d_logits, d_idx, d_prob,
                       AI4E_TOPK, stream);
    }
    if (rc) return rc;
    if (marks) AI4E_CUDA_TRY(cudaEventRecord(marks[mark++], stream));
  }
  return AI4E_OK;
}

// Next lane (round-robin) with `stream` made to wait for that lane's previous forward.  Caller
// holds e->launch_mu and calls lane_done() after enqueueing.
int lane_begin(ai4e_engine* e, cudaStream_t stream, Lane** out, int force_lane = -1) {
  const int li = force_lane >= 0 ? force_lane : (int)(e->next_lane++ % e->lanes.size());
  Lane& ln = e->lanes[li];
  if (ln.chain_valid) AI4E_CUDA_TRY(cudaStreamWaitEvent(stream, ln.ev_chain, 0));
  e->last_lane = li;
  *out = &ln;
  return AI4E_OK;
}
int lane_done(Lane* ln, cudaStream_t stream) {
  AI4E_CUDA_TRY(cudaEventRecord(ln->ev_chain, stream));
  ln->chain_valid = true;
  return AI4E_OK;
}

// H2D (already staged in c.h_in, dense) -> forward -> D2H on c.stream
int launch_batch(ai4e_engine* e, BatchCtx& c, bool want_logits) {
  const int K = e->desc.num_classes;
  const size_t img_bytes = (size_t)c.src_h * c.src_w * e->desc.in_c;
  const uint8_t* h_src = c.ext_src ? c.ext_src : c.h_in;  // ext_src: caller's page-locked buffer
  ChipGrid grid{};
  if (c.kind == kKindTile) {
    grid.grid_w = c.tile_grid_w; grid.first = c.tile_first;
    grid.row_bytes = c.tile_row_bytes;
    grid.x_stride = (long long)c.tile_chip * e->desc.in_c;
    grid.y_stride = (long long)c.tile_chip * c.tile_row_bytes;
  }
  if (c.kind > 0) {
    // K0: the slots hold entropy-decoded JPEG records (one geometry per batch): copy them, then
    // dequantise + IDCT + upsample + colour-convert straight into the pixel slots K1 reads
    const size_t rec_bytes = (size_t)c.jinfo.record_int16 * sizeof(int16_t);
    AI4E_REQUIRE(c.d_rec && rec_bytes <= e->slot_bytes && (size_t)c.jinfo.plane_bytes <= e->slot_bytes,
                 "jpeg batch exceeds the staging slot");
    AI4E_CUDA_TRY(cudaMemcpyAsync(c.d_rec, c.h_in, rec_bytes * c.count, cudaMemcpyHostToDevice, c.stream));
  } else if (c.kind == kKindTile) {
    // K6: the band is already on its way (one H2D on the copy stream); nothing is copied per chip
    AI4E_CUDA_TRY(cudaStreamWaitEvent(c.stream, c.tile->ev_ready, 0));
  } else {
    AI4E_CUDA_TRY(cudaMemcpyAsync(c.d_in, h_src, img_bytes * c.count, cudaMemcpyHostToDevice, c.stream));
  }
  {
    std::lock_guard<std::mutex> lk(e->launch_mu);
    Lane* ln = nullptr;
    int rc = lane_begin(e, c.stream, &ln);
    if (rc) return rc;
    AI4E_CUDA_TRY(cudaEventRecord(c.ev_start, c.stream));
    rc = forward_enqueue(e, ln, c.kind == kKindTile ? c.tile->d : c.d_in, c.count, c.src_h, c.src_w, c.d_idx, c.d_prob,
                         c.d_logits, c.stream, nullptr, c.kind > 0 ? &c : nullptr, c.kind == kKindTile ? &grid : nullptr);
    if (rc) return rc;
    AI4E_CUDA_TRY(cudaEventRecord(c.ev_fwd, c.stream));
    rc = lane_done(ln, c.stream);
    if (rc) return rc;
  }
  if (c.kind == kKindTile && c.tile->summary) {  // K7: this batch's chips into the tile's aggregate, still on the device
    const SumLayout sl(c.tile->sum_chips, K);
    uint8_t* d = c.tile->d_sum;
    const int rc = tile_summary_launch(c.d_idx, c.d_prob, c.count, c.sum_first, K, (int32_t*)(d + sl.labels), (float*)(d + sl.prob),
                                       (uint32_t*)(d + sl.hist), (unsigned long long*)(d + sl.mass), c.stream);
    if (rc) return rc;
  }
  AI4E_CUDA_TRY(cudaMemcpyAsync(c.h_idx, c.d_idx, (size_t)c.count * AI4E_TOPK * sizeof(int32_t),
                                cudaMemcpyDeviceToHost, c.stream));
  AI4E_CUDA_TRY(cudaMemcpyAsync(c.h_prob, c.d_prob, (size_t)c.count * AI4E_TOPK * sizeof(float),
                                cudaMemcpyDeviceToHost, c.stream));
  if (want_logits)
    AI4E_CUDA_TRY(cudaMemcpyAsync(c.h_logits, c.d_logits, (size_t)c.count * K * sizeof(float),
                                  cudaMemcpyDeviceToHost, c.stream));
  AI4E_CUDA_TRY(cudaEventRecord(c.ev_done, c.stream));
  return AI4E_OK;
}

bool is_page_locked(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

void account_batch(ai4e_engine* e, BatchCtx& c) {  // caller holds e->mu
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, c.ev_start, c.ev_fwd) == cudaSuccess) e->st.gpu_ms_total += ms;
  e->st.images += c.count;
  e->st.batches += 1;
  int b = 0;
  while ((1 << b) < c.count && b < 10) ++b;
  e->st.batch_hist[b] += 1;
}

void batcher_main(ai4e_engine* e) {
  cudaSetDevice(e->device);
  std::unique_lock<std::mutex> lk(e->mu);
  for (;;) {
    for (;;) {
      if (!e->closed_q.empty() || e->stopping) break;
      BatchCtx* o = e->open_ctx;
      if (o && o->count > 0) {
        const auto deadline = o->t_first + std::chrono::microseconds(e->max_wait_us);
        if (e->flush_req || Clock::now() >= deadline) {
          e->closed_q.push_back(o);
          e->open_ctx = nullptr;
          break;
        }
        e->cv_work.wait_until(lk, deadline);
      } else {
        e->flush_req = false;
        e->cv_work.wait(lk);
      }
    }
    if (e->closed_q.empty()) {
      if (e->stopping) {
        if (e->open_ctx && e->open_ctx->count > 0) {  // drain what was submitted
          e->closed_q.push_back(e->open_ctx);
          e->open_ctx = nullptr;
          continue;
        }
        break;
      }
      continue;
    }
    BatchCtx* c = e->closed_q.front();
    e->closed_q.pop_front();
    lk.unlock();
    // every submitter of this batch has reserved its slot; wait for the copies to land - bounded, so a
    // submitter that never finishes (a killed thread) fails this batch instead of hanging the queue
    const auto copy_deadline = Clock::now() + std::chrono::seconds(10);
    bool copies_ok = true;
    while (c->copied.load(std::memory_order_acquire) < c->count) {
      if (Clock::now() > copy_deadline) { copies_ok = false; break; }
      std::this_thread::yield();
    }
    if (!copies_ok) {
      set_last_error("batch of %d: only %d image copies landed within 10 s", c->count, c->copied.load());
      c->launch_rc = AI4E_ERR_TIMEOUT;
    } else if (!e->healthy.load()) {
      c->launch_rc = AI4E_ERR_STATE;
    } else if (e->inject_faults.load() > 0 && e->inject_faults.fetch_sub(1) > 0) {
      c->launch_rc = AI4E_ERR_CUDA;  // ai4e_engine_debug_inject_fault: behave as if the launch hit a sticky CUDA error
    } else {
      c->launch_rc = launch_batch(e, *c, true);
    }
    lk.lock();
    e->inflight_q.push_back(c);
    e->cv_inflight.notify_one();
  }
  e->batcher_done = true;
  e->cv_inflight.notify_all();
}

void completer_main(ai4e_engine* e) {
  cudaSetDevice(e->device);
  std::unique_lock<std::mutex> lk(e->mu);
  for (;;) {
    e->cv_inflight.wait(lk, [&] { return !e->inflight_q.empty() || e->batcher_done; });
    if (e->inflight_q.empty()) break;
    BatchCtx* c = e->inflight_q.front();
    e->inflight_q.pop_front();
    lk.unlock();
    int rc = c->launch_rc;
    if (rc == AI4E_OK) {
      cudaError_t ce = cudaEventSynchronize(c->ev_done);
      if (ce != cudaSuccess) {
        set_last_error("batch failed on device %d: %s", e->device, cudaGetErrorString(ce));
        rc = AI4E_ERR_CUDA;
      }
    }
    // a CUDA fault is sticky (the context is gone): the replica stops taking work.  Anything else
    // (bad geometry, a copy that never landed, no memory) fails this batch's tickets only.
    if (rc == AI4E_ERR_CUDA) e->healthy.store(0);
    // K7: the last batch of a summarised tile brings the aggregate back (every earlier batch of the tile has been
    // synchronised by this thread already; sum_left was set before the first of them was queued)
    TileBuf* const stb = (c->tile && c->tile->summary) ? c->tile : nullptr;
    bool sum_final = false;
    if (stb) {
      {
        std::lock_guard<std::mutex> g(e->mu);  // a submitter that gave up half-way adjusts these under the lock
        if (rc && !stb->sum_rc) stb->sum_rc = rc;
        sum_final = stb->sum_left == 1;
      }
      if (sum_final && !stb->sum_rc) {
        const SumLayout sl(stb->sum_chips, e->desc.num_classes);
        cudaError_t ce = cudaMemcpyAsync(stb->h_sum, stb->d_sum, sl.bytes, cudaMemcpyDeviceToHost, e->sum_stream);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->sum_stream);
        if (ce != cudaSuccess) {
          set_last_error("tile summary read-back failed on device %d: %s", e->device, cudaGetErrorString(ce));
          stb->sum_rc = AI4E_ERR_CUDA;
          e->healthy.store(0);
        }
      }
    }
    const auto now = Clock::now();
    std::vector<DoneRec> announce;
    lk.lock();
    const int K = e->desc.num_classes;
    for (int i = 0; i < c->count; ++i) {
      const uint64_t t = c->first_ticket + i;
      const uint64_t r = t % kResultRing;
      e->r_status[r] = (int8_t)rc;
      if (e->r_notify[r]) {
        DoneRec d{};
        d.ticket = e->ticket_prefix | t; d.tag = e->r_tag[r]; d.status = rc;
        if (rc == AI4E_OK) {
          memcpy(d.idx, c->h_idx + (size_t)i * AI4E_TOPK, sizeof(d.idx));
          memcpy(d.prob, c->h_prob + (size_t)i * AI4E_TOPK, sizeof(d.prob));
        }
        announce.push_back(d);
        e->r_notify[r] = 0;
      }
      if (rc == AI4E_OK) {
        memcpy(&e->r_idx[r * AI4E_TOPK], c->h_idx + (size_t)i * AI4E_TOPK, AI4E_TOPK * sizeof(int32_t));
        memcpy(&e->r_prob[r * AI4E_TOPK], c->h_prob + (size_t)i * AI4E_TOPK, AI4E_TOPK * sizeof(float));
        memcpy(&e->r_logits[(t % kLogitRing) * K], c->h_logits + (size_t)i * K, K * sizeof(float));
        e->lat_ms[e->lat_n++ % kLatRing] =
            std::chrono::duration<float, std::milli>(now - e->r_submit[r]).count();
      }
    }
    if (rc == AI4E_OK) account_batch(e, *c);
    e->done_upto = c->first_ticket + c->count;
    e->n_completed.fetch_add((uint64_t)c->count);
    if (stb) {
      stb->sum_left -= 1;
      if (sum_final) {
        TileSumRec& sr = e->sums[stb->sum_ticket % kSumRing];
        if (sr.id == stb->sum_ticket) {
          if (!stb->sum_rc) sr.data.assign(stb->h_sum, stb->h_sum + SumLayout(stb->sum_chips, K).bytes);
          sr.status = stb->sum_rc;
        }
        stb->summary = false;
      }
    }
    if (c->tile) { c->tile->users -= 1; c->tile = nullptr; }
    c->count = 0;
    c->kind = 0;
    c->ext_src = nullptr;
    c->copied.store(0);
    e->free_q.push_back(c);
    e->cv_free.notify_all();
    e->cv_done.notify_all();
    if (!announce.empty()) {
      lk.unlock();
      {
        std::lock_guard<std::mutex> ql(e->cq->mu);
        for (const DoneRec& d : announce) e->cq->q.push_back(d);
      }
      e->cq->cv.notify_all();
      lk.lock();
    }
  }
  e->completer_done = true;
  e->cv_done.notify_all();
}

int engine_create(int device, const ai4e_model_desc* desc, const void* blob, uint64_t blob_bytes, int max_batch,
                  int max_wait_us, ai4e_engine** out) {
  AI4E_REQUIRE(desc && blob && out, "engine_create: null pointer");
  AI4E_REQUIRE(desc->n_layers > 0 && desc->layers, "engine_create: empty layer list");
  AI4E_REQUIRE(max_batch >= 1 && max_batch <= 4096, "engine_create: max_batch %d outside [1, 4096]", max_batch);
  AI4E_REQUIRE(desc->in_c == 3 || desc->in_c == 4, "engine_create: in_c must be 3 or 4");
  AI4E_REQUIRE(desc->in_c_pad == 4, "engine_create: in_c_pad must be 4");
  AI4E_REQUIRE(desc->n_bufs >= 2 && desc->n_bufs <= 64, "engine_create: n_bufs outside [2, 64]");
  AI4E_REQUIRE(desc->num_classes >= AI4E_TOPK, "engine_create: fewer classes than top-k");
  int ndev = 0;
  AI4E_CUDA_TRY(cudaGetDeviceCount(&ndev));
  AI4E_REQUIRE(device >= 0 && device < ndev, "engine_create: device %d of %d", device, ndev);
  AI4E_CUDA_TRY(cudaSetDevice(device));

  ai4e_engine* e = new ai4e_engine();
  e->device = device; e->max_batch = max_batch; e->max_wait_us = std::max(0, max_wait_us);
  e->desc = *desc;
  e->desc.layers = nullptr;
  auto fail = [&](int rc) { ai4e_engine_destroy(e); return rc; };

  // ---- shape inference over the slot graph
  struct Shape { int h = 0, w = 0, c = 0; bool set = false; };
  std::vector<Shape> shp(desc->n_bufs);
  shp[0] = {desc->in_h, desc->in_w, desc->in_c_pad, true};
  e->buf_elems_per_image.assign(desc->n_bufs, 0);
  e->buf_elems_per_image[0] = (size_t)desc->in_h * desc->in_w * desc->in_c_pad;
  bool seen_head = false;
  for (int i = 0; i < desc->n_layers; ++i) {
    LayerRt L{};
    L.d = desc->layers[i];
    const ai4e_layer_desc& d = L.d;
    if (!(d.in_buf >= 0 && d.in_buf < desc->n_bufs && shp[d.in_buf].set)) {
      set_last_error("layer %d reads slot %d before it is written", i, d.in_buf);
      return fail(AI4E_ERR_INVALID);
    }
    const Shape in = shp[d.in_buf];
    L.in_h = in.h; L.in_w = in.w;
    if (d.cin != in.c) {
      set_last_error("layer %d: cin %d but slot %d holds %d channels", i, d.cin, d.in_buf, in.c);
      return fail(AI4E_ERR_INVALID);
    }
    if (d.op == AI4E_OP_HEAD) {
      if (d.w_off < 0 || d.b_off < 0 || (uint64_t)d.w_off + (uint64_t)d.cout * d.cin * 2 > blob_bytes ||
          (uint64_t)d.b_off + (uint64_t)d.cout * 4 > blob_bytes || d.cout != desc->num_classes || d.cin % 8) {
        set_last_error("layer %d (head): bad weights/geometry", i);
        return fail(AI4E_ERR_INVALID);
      }
      seen_head = true;
      e->layers.push_back(L);
      continue;
    }
    if (!(d.out_buf > 0 && d.out_buf < desc->n_bufs && d.out_buf != d.in_buf && d.out_buf != d.res_buf) ||
        d.kh <= 0 || d.kw <= 0 || d.stride <= 0 || d.pad < 0 || in.h + 2 * d.pad < d.kh || in.w + 2 * d.pad < d.kw) {
      set_last_error("layer %d: bad slots or geometry", i);
      return fail(AI4E_ERR_INVALID);
    }
    L.out_h = (in.h + 2 * d.pad - d.kh) / d.stride + 1;
    L.out_w = (in.w + 2 * d.pad - d.kw) / d.stride + 1;
    int cout = d.cout;
    if (d.op == AI4E_OP_MAXPOOL) {
      cout = d.cin;
      if (d.kh != d.kw || d.cin % 8) { set_last_error("layer %d (maxpool): bad geometry", i); return fail(AI4E_ERR_INVALID); }
    } else if (d.op == AI4E_OP_CONV) {
      const uint64_t wbytes = (uint64_t)d.cout * d.kh * d.kw * d.cin * 2;
      if (d.w_off < 0 || d.b_off < 0 || (d.w_off & 255) || (d.b_off & 255) || (uint64_t)d.w_off + wbytes > blob_bytes ||
          (uint64_t)d.b_off + (uint64_t)d.cout * 4 > blob_bytes || d.cout % 8) {
        set_last_error("layer %d (conv): weights outside the blob or misaligned", i);
        return fail(AI4E_ERR_INVALID);
      }
      if (d.res_buf >= 0) {
        const bool ok = d.res_buf < desc->n_bufs && shp[d.res_buf].set && shp[d.res_buf].h == L.out_h &&
                        shp[d.res_buf].w == L.out_w && shp[d.res_buf].c == d.cout;
        if (!ok) { set_last_error("layer %d: residual slot %d does not match the output", i, d.res_buf); return fail(AI4E_ERR_INVALID); }
      }
      L.igemm = igemm_supported(d.cin, d.cout, d.kh, d.kw, d.stride, d.pad) != 0;
    } else {
      set_last_error("layer %d: unknown op %d", i, d.op);
      return fail(AI4E_ERR_INVALID);
    }
    shp[d.out_buf] = {L.out_h, L.out_w, cout, true};
    e->buf_elems_per_image[d.out_buf] =
        std::max(e->buf_elems_per_image[d.out_buf], (size_t)L.out_h * L.out_w * cout);
    e->layers.push_back(L);
  }
  if (!seen_head || e->layers.back().d.op != AI4E_OP_HEAD) {
    set_last_error("engine_create: the last layer must be the head");
    return fail(AI4E_ERR_INVALID);
  }

  // ---- device memory: weights, activation slots, scratch
#define TRY_OR_FAIL(expr)                                                                    \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      set_last_error("%s failed: %s", #expr, cudaGetErrorString(_e));                        \
      return fail(_e == cudaErrorMemoryAllocation ? AI4E_ERR_NOMEM : AI4E_ERR_CUDA);         \
    }                                                                                        \
  } while (0)
  TRY_OR_FAIL(cudaMalloc((void**)&e->d_blob, blob_bytes));
  TRY_OR_FAIL(cudaMemcpy(e->d_blob, blob, blob_bytes, cudaMemcpyHostToDevice));
  int n_lanes = 2;
  if (const char* v = getenv("AI4E_B200_LANES")) n_lanes = std::max(1, std::min(4, atoi(v)));
  e->lanes.resize(n_lanes);
  // With two lanes a kernel that occupies every SM keeps the other lane's next kernel from even
  // starting its prologue and weight prefetch; grids sized for 16 fewer SMs measured +2 % (sweep in
  // DESIGN.md).  AI4E_B200_GRID_SMS overrides.
  int grid_sms = n_lanes > 1 ? std::max(2, sm_count() - 16) : 0;
  if (const char* v = getenv("AI4E_B200_GRID_SMS")) grid_sms = atoi(v);
  for (int li = 0; li < n_lanes; ++li) {
  Lane& ln = e->lanes[li];
  ln.layers = e->layers;
  ln.bufs.assign(desc->n_bufs, nullptr);
  for (int b = 0; b < desc->n_bufs; ++b) {
    if (!e->buf_elems_per_image[b]) continue;
    const size_t bytes = e->buf_elems_per_image[b] * 2 * (size_t)max_batch;
    TRY_OR_FAIL(cudaMalloc(&ln.bufs[b], bytes));
    TRY_OR_FAIL(cudaMemset(ln.bufs[b], 0, bytes));
  }
  int head_c = ln.layers.back().d.cin;
  TRY_OR_FAIL(cudaMalloc((void**)&ln.d_pooled, (size_t)max_batch * head_c * sizeof(float)));
  TRY_OR_FAIL(cudaMalloc((void**)&ln.d_logits, (size_t)max_batch * desc->num_classes * sizeof(float)));
  TRY_OR_FAIL(cudaEventCreateWithFlags(&ln.ev_chain, cudaEventDisableTiming));

  // ---- tensor maps for the tcgen05 layers (batch-independent, see igemm_make_plan)
  if (li == 0) e->launches = 1;  // preprocess
  for (LayerRt& L : ln.layers) {
    const ai4e_layer_desc& d = L.d;
    if (li == 0) e->launches += d.op == AI4E_OP_HEAD ? 3 : 1;
    if (d.op != AI4E_OP_CONV) continue;
    if (!L.igemm && d.res_buf < 0 && stem_supported(L.in_h, L.in_w, d.cin, d.cout, d.kh, d.kw, d.stride, d.pad) &&
        !getenv("AI4E_B200_DIRECT_STEM")) {
      std::vector<uint16_t> w_sp(64 * 7 * 4 * 8);
      stem_repack_weights(reinterpret_cast<const uint16_t*>(static_cast<const uint8_t*>(blob) + d.w_off), w_sp.data());
      if (!e->d_stem_w) TRY_OR_FAIL(cudaMalloc((void**)&e->d_stem_w, w_sp.size() * 2));
      TRY_OR_FAIL(cudaMemcpy(e->d_stem_w, w_sp.data(), w_sp.size() * 2, cudaMemcpyHostToDevice));
      int rc = igemm_make_plan_stem(&L.plan, ln.bufs[d.in_buf], max_batch, L.in_h, L.in_w, e->d_stem_w,
                                    (const float*)(e->d_blob + d.b_off), d.relu, ln.bufs[d.out_buf]);
      if (rc) return fail(rc);
      L.plan.grid_sms = grid_sms;
      L.igemm = true;
      continue;
    }
    if (!L.igemm) continue;
    int rc = igemm_make_plan(&L.plan, ln.bufs[d.in_buf], max_batch, L.in_h, L.in_w, d.cin, e->d_blob + d.w_off,
                             (const float*)(e->d_blob + d.b_off), d.cout, d.kh, d.kw, d.stride, d.pad,
                             d.res_buf >= 0 ? ln.bufs[d.res_buf] : nullptr, d.relu, ln.bufs[d.out_buf]);
    if (rc) return fail(rc);
    L.plan.grid_sms = grid_sms;
  }

  // ---- stem + max-pool: a row-mode stem whose only consumer is a 3x3/2 max-pool runs with it in one
  // kernel (stem_pool.cu); the stem's output - the largest tensor of the network - is never written.
  // Not when the program keeps every stage in memory (desc.reserved bit 0: stage read-back).
  if (!getenv("AI4E_B200_NO_STEM_POOL") && !(desc->reserved & 1)) {
    for (size_t i = 0; i + 1 < ln.layers.size(); ++i) {
      LayerRt& A = ln.layers[i];
      LayerRt& B = ln.layers[i + 1];
      const ai4e_layer_desc& a = A.d;
      const ai4e_layer_desc& b = B.d;
      if (a.op != AI4E_OP_CONV || !A.igemm || A.plan.stem_mode == 0 || b.op != AI4E_OP_MAXPOOL || b.in_buf != a.out_buf ||
          b.out_buf == a.in_buf || a.res_buf >= 0)
        continue;
      if (!stem_pool_supported(A.in_h, A.in_w, a.cin, a.cout, a.kh, a.kw, a.stride, a.pad, b.kh, b.stride, b.pad)) continue;
      bool sole_consumer = true;
      for (size_t k = i + 2; k < ln.layers.size() && sole_consumer; ++k) {
        const ai4e_layer_desc& c = ln.layers[k].d;
        if (c.in_buf == a.out_buf || c.res_buf == a.out_buf) sole_consumer = false;
        if (c.out_buf == a.out_buf) break;
      }
      if (!sole_consumer) continue;
      if (!e->d_stem_pool_w) {
        std::vector<uint16_t> w_sp((size_t)stem_pool_weight_elems());
        stem_pool_repack_weights(reinterpret_cast<const uint16_t*>(static_cast<const uint8_t*>(blob) + a.w_off), w_sp.data());
        TRY_OR_FAIL(cudaMalloc((void**)&e->d_stem_pool_w, w_sp.size() * 2));
        TRY_OR_FAIL(cudaMemcpy(e->d_stem_pool_w, w_sp.data(), w_sp.size() * 2, cudaMemcpyHostToDevice));
      }
      int rc = stem_pool_make_plan(&A.splan, ln.bufs[a.in_buf], max_batch, A.in_h, A.in_w, e->d_stem_pool_w,
                                   (const float*)(e->d_blob + a.b_off), a.relu, ln.bufs[b.out_buf]);
      if (rc) continue;  // (a driver that refuses the tensor maps: the two kernels run separately)
      A.splan.grid_sms = grid_sms;
      A.stem_pool = true;
      B.fused_second = true;
      if (li == 0) e->launches -= 1;
      // K1 inside the stem as well is built and bit-exact but OFF by default: two converter warps per CTA
      // cannot keep up with the tensor core (92 us against 51 + 20 us for the two kernels at micro-batch 64,
      // profiles/r02_ncu_full.md), so it only pays where a launch matters more than throughput.
      // AI4E_B200_U8_STEM=1 turns it on.
      if (i == 0 && a.in_buf == 0 && desc->in_c == 3 && stem_pool_u8_supported(A.in_h, A.in_w, desc->in_c) &&
          getenv("AI4E_B200_U8_STEM")) {
        ln.u8_stem = true;
        if (li == 0) e->launches_u8 = -1;  // ... and no preprocessing launch for raw batches of the input size (counted below)
      }
    }
  }

  // ---- bottleneck tails: a convolution + ReLU whose only consumer is a 1x1 convolution with a
  // residual runs with it in one kernel (conv_fused.cu); the intermediate buffer is never written.
  if (!getenv("AI4E_B200_NO_FUSE")) {
    for (size_t i = 0; i + 1 < ln.layers.size(); ++i) {
      LayerRt& A = ln.layers[i];
      LayerRt& B = ln.layers[i + 1];
      const ai4e_layer_desc& a = A.d;
      const ai4e_layer_desc& b = B.d;
      if (a.op != AI4E_OP_CONV || b.op != AI4E_OP_CONV || !A.igemm || A.plan.stem_mode != 0 || A.fused_second) continue;
      if (a.res_buf >= 0 || !a.relu || b.in_buf != a.out_buf || b.kh != 1 || b.kw != 1 || b.stride != 1 || b.pad != 0 ||
          b.res_buf < 0 || !b.relu || b.res_buf == a.out_buf || b.out_buf == a.out_buf || b.out_buf == a.in_buf)
        continue;
      if (!fused_tail_supported(a.cin, a.cout, b.cout, a.kh, a.kw, a.stride, a.pad)) continue;
      bool sole_consumer = true;  // nobody else reads the intermediate before it is overwritten
      for (size_t k = i + 2; k < ln.layers.size() && sole_consumer; ++k) {
        const ai4e_layer_desc& c = ln.layers[k].d;
        if (c.in_buf == a.out_buf || c.res_buf == a.out_buf) sole_consumer = false;
        if (c.out_buf == a.out_buf) break;
      }
      if (!sole_consumer) continue;
      int rc = fused_tail_make_plan(&A.fplan, ln.bufs[a.in_buf], max_batch, A.in_h, A.in_w, a.cin, e->d_blob + a.w_off,
                                    (const float*)(e->d_blob + a.b_off), a.cout, a.kh, a.kw, a.stride, a.pad,
                                    e->d_blob + b.w_off, (const float*)(e->d_blob + b.b_off), b.cout, ln.bufs[b.res_buf],
                                    ln.bufs[b.out_buf]);
      if (rc) return fail(rc);
      A.fplan.grid_sms = grid_sms;
      A.fused_first = true;
      B.fused_second = true;
      if (li == 0) e->launches -= 1;
    }
  }

  // ---- row-tile direction: opposite to the layer that wrote the input, so that a layer starts with
  // the rows still in L2.  Slot 0 (the preprocessed image) is written first-to-last.
  if (!getenv("AI4E_B200_NO_REVERSE")) {
    std::vector<char> dir(desc->n_bufs, 0);  // direction in which each slot was last written
    for (size_t i = 0; i < ln.layers.size(); ++i) {
      LayerRt& L = ln.layers[i];
      const ai4e_layer_desc& d = L.d;
      if (d.op == AI4E_OP_HEAD) continue;
      if (L.fused_second) {  // written by the previous layer's launch, in that launch's order
        L.reverse = ln.layers[i - 1].reverse;
      } else {
        const bool stem = d.op == AI4E_OP_CONV && L.igemm && L.plan.stem_mode != 0;
        const bool can = d.op == AI4E_OP_MAXPOOL || (d.op == AI4E_OP_CONV && L.igemm && !stem);
        L.reverse = can && !dir[d.in_buf];
        if (d.op == AI4E_OP_CONV && L.igemm) L.plan.prm.reverse = L.reverse;
        if (L.fused_first) L.fplan.prm.reverse = L.reverse;
      }
      dir[d.out_buf] = L.reverse;
    }
  }

  // ---- side branches: a convolution whose output the next layer does not read, while that next
  // layer reads the same input (a block's downsample path), goes to a second stream and overlaps
  // the layers up to its first consumer.  Those layers rarely fill all SMs at micro-batch 64.
  if (!getenv("AI4E_B200_NO_SIDE_STREAM")) {
    const int n = (int)ln.layers.size();
    for (int i = 0; i + 1 < n; ++i) {
      LayerRt& A = ln.layers[i];
      const ai4e_layer_desc& a = A.d;
      const ai4e_layer_desc& b = ln.layers[i + 1].d;
      if (a.op != AI4E_OP_CONV || !A.igemm || A.plan.stem_mode != 0 || A.fused_first || A.fused_second) continue;
      if (b.in_buf != a.in_buf || b.in_buf == a.out_buf || b.res_buf == a.out_buf || b.out_buf == a.out_buf) continue;
      int consumer = -1;
      bool ok = true;
      for (int k = i + 1; k < n && ok; ++k) {
        const ai4e_layer_desc& c = ln.layers[k].d;
        if (c.in_buf == a.out_buf || c.res_buf == a.out_buf) { consumer = k; break; }
        if (c.out_buf == a.in_buf || c.out_buf == a.out_buf) ok = false;  // someone overwrites a buffer the branch uses
        if (a.res_buf >= 0 && c.out_buf == a.res_buf) ok = false;
      }
      if (!ok || consumer < 0) continue;
      if (ln.layers[consumer].fused_second) --consumer;  // that layer runs inside the previous launch
      if (consumer <= i) continue;
      if (!ln.side_stream) TRY_OR_FAIL(cudaStreamCreateWithFlags(&ln.side_stream, cudaStreamNonBlocking));
      TRY_OR_FAIL(cudaEventCreateWithFlags(&A.ev_fork, cudaEventDisableTiming));
      TRY_OR_FAIL(cudaEventCreateWithFlags(&A.ev_join, cudaEventDisableTiming));
      A.side = true;
      ln.layers[consumer].wait_ev = A.ev_join;
    }
  }

  // ---- per-layer tiling: time the candidates once, at the full micro-batch (smaller batches fall
  // back to the cost model).  ~50 ms for ResNet-50.  AI4E_B200_NO_AUTOTUNE=1 skips it.
  if (li > 0) {  // same shapes, same tilings as lane 0 measured
    for (size_t i = 0; i < ln.layers.size(); ++i) {
      ln.layers[i].plan.tuned_tiling = e->lanes[0].layers[i].plan.tuned_tiling;
      ln.layers[i].plan.tuned_batch = e->lanes[0].layers[i].plan.tuned_batch;
    }
  } else if (!getenv("AI4E_B200_NO_AUTOTUNE")) {
    const bool log = getenv("AI4E_B200_TUNE_LOG") != nullptr;
    for (LayerRt& L : ln.layers) {
      if (L.d.op != AI4E_OP_CONV || !L.igemm || L.fused_first || L.fused_second) continue;
      float us = 0.f;
      int rc = igemm_autotune(&L.plan, max_batch, nullptr, &us);
      if (rc) return fail(rc);
      if (log && L.plan.tuned_tiling)
        fprintf(stderr, "ai4e_b200: %dx%d %d->%d k%d s%d%s: tiling %d%s (%.1f us)\n", L.in_h, L.in_w, L.d.cin, L.d.cout,
                L.d.kh, L.d.stride, L.d.res_buf >= 0 ? " +res" : "", L.plan.tuned_tiling % kTilingPair,
                L.plan.tuned_tiling >= kTilingPair ? " pair" : "", us);
    }
    // the timed launches wrote garbage into the activation buffers; nothing reads them before
    // the next forward overwrites them
  }

  }  // lanes
  if (e->launches_u8 < 0) e->launches_u8 = e->launches - 1;

  // ---- staging: images up to 2x the model input's bytes: covers a 256 -> 224 resize+crop (1.31x)
  // and the int16 coefficient record of a 4:4:4 JPEG of the input size (2x + tables)
  const size_t in_bytes = (size_t)desc->in_h * desc->in_w * desc->in_c;
  size_t slot = 2 * in_bytes + 4096;
  if (const char* s = getenv("AI4E_B200_MAX_SRC_BYTES")) slot = std::max(in_bytes, (size_t)strtoull(s, nullptr, 10));
  e->slot_bytes = (slot + 255) & ~(size_t)255;
  int rc = alloc_ctx(e, e->sync_ctx);
  if (rc) return fail(rc);
  for (int i = 0; i < kNumCtx; ++i) {
    rc = alloc_ctx(e, e->ctx[i]);
    if (rc) return fail(rc);
    e->free_q.push_back(&e->ctx[i]);
  }
  const int K = desc->num_classes;
  e->r_idx.assign(kResultRing * AI4E_TOPK, 0);
  e->r_prob.assign(kResultRing * AI4E_TOPK, 0.f);
  e->r_logits.assign(kLogitRing * K, 0.f);
  e->r_status.assign(kResultRing, 0);
  e->r_notify.assign(kResultRing, 0);
  e->r_tag.assign(kResultRing, 0);
  e->r_submit.assign(kResultRing, Clock::now());
  e->lat_ms.assign(kLatRing, 0.f);
  if (const char* v = getenv("AI4E_B200_WAIT_TIMEOUT_MS")) e->wait_timeout_ms = std::max(1, atoi(v));
  TRY_OR_FAIL(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
  TRY_OR_FAIL(cudaStreamCreateWithFlags(&e->sum_stream, cudaStreamNonBlocking));
  for (TileBuf& tb : e->tiles) TRY_OR_FAIL(cudaEventCreateWithFlags(&tb.ev_ready, cudaEventDisableTiming));
  e->batcher = std::thread(batcher_main, e);
  e->completer = std::thread(completer_main, e);
  *out = e;
  return AI4E_OK;
}

}  // namespace

extern "C" {

int ai4e_engine_create(int device, const ai4e_model_desc* desc, const void* h_weights_blob, uint64_t blob_bytes,
                       int max_batch, int max_wait_us, ai4e_engine** out) {
  return engine_create(device, desc, h_weights_blob, blob_bytes, max_batch, max_wait_us, out);
}

int ai4e_engine_destroy(ai4e_engine* e) {
  if (!e) return AI4E_OK;
  cudaSetDevice(e->device);
  {
    std::lock_guard<std::mutex> lk(e->mu);
    e->stopping = true;
    e->cv_work.notify_all();
    e->cv_free.notify_all();
  }
  if (e->batcher.joinable()) e->batcher.join();
  {
    std::lock_guard<std::mutex> lk(e->mu);
    e->batcher_done = true;
    e->cv_inflight.notify_all();
  }
  if (e->completer.joinable()) e->completer.join();
  {
    std::lock_guard<std::mutex> lk(e->mu);
    e->cv_done.notify_all();
  }
  cudaDeviceSynchronize();
  free_ctx(e->sync_ctx);
  for (int i = 0; i < kNumCtx; ++i) free_ctx(e->ctx[i]);
  for (Lane& ln : e->lanes) {
    for (void* b : ln.bufs)
      if (b) cudaFree(b);
    if (ln.d_pooled) cudaFree(ln.d_pooled);
    if (ln.d_logits) cudaFree(ln.d_logits);
    if (ln.ev_chain) cudaEventDestroy(ln.ev_chain);
    for (LayerRt& L : ln.layers) {
      if (L.ev_fork) cudaEventDestroy(L.ev_fork);
      if (L.ev_join) cudaEventDestroy(L.ev_join);
    }
    if (ln.side_stream) cudaStreamDestroy(ln.side_stream);
  }
  for (TileBuf& tb : e->tiles) {
    if (tb.d) cudaFree(tb.d);
    if (tb.d_sum) cudaFree(tb.d_sum);
    if (tb.h_sum) cudaFreeHost(tb.h_sum);
    if (tb.ev_ready) cudaEventDestroy(tb.ev_ready);
  }
  if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
  if (e->sum_stream) cudaStreamDestroy(e->sum_stream);
  if (e->d_blob) cudaFree(e->d_blob);
  if (e->d_stem_w) cudaFree(e->d_stem_w);
  if (e->d_stem_pool_w) cudaFree(e->d_stem_pool_w);
  delete e;
  return AI4E_OK;
}

int ai4e_engine_forward_host(ai4e_engine* e, const uint8_t* h_images, int batch, int src_h, int src_w,
                             int32_t* h_top_idx, float* h_top_prob, float* h_logits_or_null) {
  AI4E_REQUIRE(e && h_images && h_top_idx && h_top_prob, "forward_host: null pointer");
  AI4E_REQUIRE(batch >= 1 && batch <= e->max_batch, "forward_host: batch %d outside [1, %d]", batch, e->max_batch);
  AI4E_REQUIRE(src_h > 0 && src_w > 0, "forward_host: empty image");
  const size_t img_bytes = (size_t)src_h * src_w * e->desc.in_c;
  AI4E_REQUIRE(img_bytes <= e->slot_bytes, "forward_host: image of %zu bytes exceeds the staging slot (%zu)",
               img_bytes, e->slot_bytes);
  if (!e->healthy.load()) { set_last_error("engine on device %d is unhealthy after a CUDA error", e->device); return AI4E_ERR_STATE; }
  AI4E_CUDA_TRY(cudaSetDevice(e->device));
  std::lock_guard<std::mutex> lk(e->sync_mu);
  BatchCtx& c = e->sync_ctx;
  c.count = batch; c.src_h = src_h; c.src_w = src_w; c.kind = 0;
  c.ext_src = is_page_locked(h_images) ? h_images : nullptr;  // page-locked input: H2D straight from it
  if (!c.ext_src) memcpy(c.h_in, h_images, img_bytes * batch);
  int rc = launch_batch(e, c, h_logits_or_null != nullptr);
  if (rc) return rc;
  cudaError_t ce = cudaEventSynchronize(c.ev_done);
  if (ce != cudaSuccess) {
    e->healthy.store(0);
    set_last_error("forward failed on device %d: %s", e->device, cudaGetErrorString(ce));
    return AI4E_ERR_CUDA;
  }
  memcpy(h_top_idx, c.h_idx, (size_t)batch * AI4E_TOPK * sizeof(int32_t));
  memcpy(h_top_prob, c.h_prob, (size_t)batch * AI4E_TOPK * sizeof(float));
  if (h_logits_or_null) memcpy(h_logits_or_null, c.h_logits, (size_t)batch * e->desc.num_classes * sizeof(float));
  {
    std::lock_guard<std::mutex> lk2(e->mu);
    account_batch(e, c);
  }
  return AI4E_OK;
}

int ai4e_engine_forward_device(ai4e_engine* e, const uint8_t* d_images, int batch, int src_h, int src_w,
                               int32_t* d_top_idx, float* d_top_prob, float* d_logits_or_null, void* stream) {
  AI4E_REQUIRE(e && d_images && d_top_idx && d_top_prob, "forward_device: null pointer");
  AI4E_REQUIRE(batch >= 1 && batch <= e->max_batch, "forward_device: batch %d outside [1, %d]", batch, e->max_batch);
  if (!e->healthy.load()) { set_last_error("engine on device %d is unhealthy after a CUDA error", e->device); return AI4E_ERR_STATE; }
  AI4E_CUDA_TRY(cudaSetDevice(e->device));
  cudaStream_t s = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(e->launch_mu);
  Lane* ln = nullptr;
  int rc = lane_begin(e, s, &ln);
  if (rc) return rc;
  // without a caller buffer the logits go to the lane's own scratch (never read back)
  rc = forward_enqueue(e, ln, d_images, batch, src_h, src_w, d_top_idx, d_top_prob,
                       d_logits_or_null ? d_logits_or_null : ln->d_logits, s);
  if (rc) return rc;
  return lane_done(ln, s);
}

// Reserves up to `want` consecutive slots of the open batch for images of one size and kind (0 = raw
// pixels, > 0 = a JPEG geometry key) and hands out their tickets; closes the batch when it is full.
// The caller copies its payload into the slots and then adds the count to ctx->copied.
static int reserve_slots(ai4e_engine* e, const char* who, int h, int w, int kind, const ai4e_jpeg_info* jinfo, int want,
                         const uint64_t* tags, uint64_t tag, int notify, BatchCtx** out_ctx, int* out_slot0, int* out_got,
                         uint64_t* out_tickets) {
  std::unique_lock<std::mutex> lk(e->mu);
  BatchCtx* ctx = nullptr;
  for (;;) {
    if (e->stopping) { set_last_error("%s: engine is shutting down", who); return AI4E_ERR_STATE; }
    if (!e->healthy.load()) { set_last_error("%s: engine on device %d is unhealthy", who, e->device); return AI4E_ERR_STATE; }
    if (!e->open_ctx) {
      if (e->free_q.empty()) { e->cv_free.wait(lk); continue; }
      e->open_ctx = e->free_q.front();
      e->free_q.pop_front();
      e->open_ctx->count = 0;
      e->open_ctx->copied.store(0);
    }
    ctx = e->open_ctx;
    // a batch holds one image size and kind; JPEG batches additionally one record layout
    const bool other = ctx->count > 0 && (ctx->src_h != h || ctx->src_w != w || ctx->kind != kind ||
                                          (kind > 0 && !jpeg_same_geometry(ctx->jinfo, *jinfo)));
    if (other) {
      e->closed_q.push_back(ctx);
      e->open_ctx = nullptr;
      e->cv_work.notify_one();
      continue;
    }
    break;
  }
  const auto now = Clock::now();
  if (ctx->count == 0) {
    ctx->src_h = h; ctx->src_w = w;
    ctx->kind = kind;
    if (jinfo) ctx->jinfo = *jinfo;
    ctx->tile = nullptr;
    ctx->t_first = now;
    ctx->first_ticket = e->next_ticket;
  }
  const int got = std::min(want, e->max_batch - ctx->count);
  *out_ctx = ctx; *out_slot0 = ctx->count; *out_got = got;
  for (int i = 0; i < got; ++i) {
    const uint64_t t = e->next_ticket++;
    const uint64_t r = t % kResultRing;
    out_tickets[i] = t;
    e->r_submit[r] = now;
    e->r_status[r] = 1;  // pending
    e->r_notify[r] = notify ? 1 : 0;
    e->r_tag[r] = tags ? tags[i] : tag;
  }
  ctx->count += got;
  e->n_submitted.fetch_add((uint64_t)got);
  if (ctx->count == e->max_batch) {
    e->closed_q.push_back(ctx);
    e->open_ctx = nullptr;
  }
  e->cv_work.notify_one();
  return AI4E_OK;
}

static inline void copy_image(uint8_t* dst, const uint8_t* src, int h, int w, int c, int row_stride) {
  if (row_stride == w * c) {
    memcpy(dst, src, (size_t)h * w * c);
  } else {
    for (int y = 0; y < h; ++y) memcpy(dst + (size_t)y * w * c, src + (size_t)y * row_stride, (size_t)w * c);
  }
}

int ai4e_engine_submit_many(ai4e_engine* e, const uint8_t* const* h_images, int n, int h, int w, int c, int row_stride,
                            const uint64_t* tags_or_null, int notify, uint64_t* out_tickets) {
  AI4E_REQUIRE(e && h_images && out_tickets, "submit_many: null pointer");
  AI4E_REQUIRE(n >= 0, "submit_many: negative count");
  AI4E_REQUIRE(c == e->desc.in_c, "submit_many: images have %d channels, model takes %d", c, e->desc.in_c);
  AI4E_REQUIRE(h > 0 && w > 0, "submit_many: empty image");
  AI4E_REQUIRE(row_stride >= 0, "submit_many: negative row stride (pass a copy of a flipped view)");
  if (row_stride == 0) row_stride = w * c;
  AI4E_REQUIRE(row_stride >= w * c, "submit_many: row stride too small");
  const size_t img_bytes = (size_t)h * w * c;
  AI4E_REQUIRE(img_bytes <= e->slot_bytes, "submit_many: image of %zu bytes exceeds the staging slot (%zu)", img_bytes,
               e->slot_bytes);
  for (int i = 0; i < n; ++i) AI4E_REQUIRE(h_images[i], "submit_many: image %d is null", i);
  int done = 0;
  while (done < n) {
    BatchCtx* ctx = nullptr;
    int slot0 = 0, got = 0;
    int rc = reserve_slots(e, "submit", h, w, 0, nullptr, n - done, tags_or_null ? tags_or_null + done : nullptr, 0, notify,
                           &ctx, &slot0, &got, out_tickets + done);
    if (rc) return rc;  // tickets handed out so far stay valid
    for (int i = 0; i < got; ++i)
      copy_image(ctx->h_in + (size_t)(slot0 + i) * img_bytes, h_images[done + i], h, w, c, row_stride);
    ctx->copied.fetch_add(got, std::memory_order_release);
    done += got;
  }
  return AI4E_OK;
}

int ai4e_engine_submit(ai4e_engine* e, const uint8_t* h_hwc, int h, int w, int c, int row_stride, uint64_t tag,
                       uint64_t* out_ticket) {
  AI4E_REQUIRE(e && h_hwc && out_ticket, "submit: null pointer");
  AI4E_REQUIRE(c == e->desc.in_c, "submit: image has %d channels, model takes %d", c, e->desc.in_c);
  AI4E_REQUIRE(h > 0 && w > 0, "submit: empty image");
  AI4E_REQUIRE(row_stride >= 0, "submit: negative row stride (pass a copy of a flipped view)");
  if (row_stride == 0) row_stride = w * c;
  AI4E_REQUIRE(row_stride >= w * c, "submit: row stride too small");
  const size_t img_bytes = (size_t)h * w * c;
  AI4E_REQUIRE(img_bytes <= e->slot_bytes, "submit: image of %zu bytes exceeds the staging slot (%zu)", img_bytes,
               e->slot_bytes);
  BatchCtx* ctx = nullptr;
  int slot = 0, got = 0;
  int rc = reserve_slots(e, "submit", h, w, 0, nullptr, 1, nullptr, tag, 0, &ctx, &slot, &got, out_ticket);
  if (rc) return rc;
  copy_image(ctx->h_in + (size_t)slot * img_bytes, h_hwc, h, w, c, row_stride);
  ctx->copied.fetch_add(1, std::memory_order_release);
  return AI4E_OK;
}

// geometry key of a JPEG batch (reserve_slots compares the complete record layout on top of it)
static int jpeg_kind(const ai4e_jpeg_info& j) { return 1 + j.components * 64 + j.hmax * 16 + j.vmax * 4 + (j.h[0] * 2 + j.v[0]) % 4; }

int ai4e_engine_submit_jpeg(ai4e_engine* e, const uint8_t* jpeg, uint64_t len, uint64_t tag, uint64_t* out_ticket) {
  AI4E_REQUIRE(e && jpeg && out_ticket, "submit_jpeg: null pointer");
  AI4E_REQUIRE(e->desc.in_c == 3, "submit_jpeg: the model takes %d channels, a JPEG decodes to 3", e->desc.in_c);
  // entropy decode on the caller's thread, into a thread-local record: nothing is reserved if it fails
  thread_local std::vector<int16_t> rec;
  ai4e_jpeg_info info;
  int rc = jpeg_host_decode(jpeg, len, nullptr, 0, &info);
  if (rc) return rc;
  const size_t rec_bytes = (size_t)info.record_int16 * sizeof(int16_t);
  const size_t rgb_bytes = (size_t)info.width * info.height * 3;
  AI4E_REQUIRE(rec_bytes <= e->slot_bytes && rgb_bytes <= e->slot_bytes && (size_t)info.plane_bytes <= e->slot_bytes,
               "submit_jpeg: a %d x %d image needs %zu bytes of staging, the slot holds %zu (AI4E_B200_MAX_SRC_BYTES)",
               info.width, info.height, std::max(rec_bytes, rgb_bytes), e->slot_bytes);
  if (rec.size() < (size_t)info.record_int16) rec.resize((size_t)info.record_int16);
  rc = jpeg_host_decode(jpeg, len, rec.data(), rec.size(), &info);
  if (rc) return rc;
  BatchCtx* ctx = nullptr;
  int slot = 0, got = 0;
  rc = reserve_slots(e, "submit_jpeg", info.height, info.width, jpeg_kind(info), &info, 1, nullptr, tag, 0, &ctx, &slot, &got,
                     out_ticket);
  if (rc) return rc;
  memcpy(ctx->h_in + (size_t)slot * rec_bytes, rec.data(), rec_bytes);
  ctx->copied.fetch_add(1, std::memory_order_release);
  return AI4E_OK;
}

// An encoded payload of either supported format: a JPEG goes the hybrid K0 way, a PNG is inflated and
// un-filtered on this thread (lossless: nothing for the GPU to do) and its pixels take the plain path.
int ai4e_engine_submit_image(ai4e_engine* e, const uint8_t* data, uint64_t len, uint64_t tag, uint64_t* out_ticket) {
  AI4E_REQUIRE(e && data && out_ticket, "submit_image: null pointer");
  if (len >= 3 && data[0] == 0xFF && data[1] == 0xD8) return ai4e_engine_submit_jpeg(e, data, len, tag, out_ticket);
  if (len >= 8 && data[0] == 0x89 && data[1] == 'P' && data[2] == 'N' && data[3] == 'G') {
    AI4E_REQUIRE(e->desc.in_c == 3, "submit_image: the model takes %d channels, a PNG decodes to 3", e->desc.in_c);
    int h = 0, w = 0;
    int rc = png_host_decode(data, len, nullptr, 0, &h, &w);
    if (rc) return rc;
    const size_t bytes = (size_t)h * w * 3;
    AI4E_REQUIRE(bytes <= e->slot_bytes, "submit_image: a %d x %d image needs %zu bytes of staging, the slot holds %zu (AI4E_B200_MAX_SRC_BYTES)",
                 w, h, bytes, e->slot_bytes);
    thread_local std::vector<uint8_t> px;
    if (px.size() < bytes) px.resize(bytes);
    rc = png_host_decode(data, len, px.data(), px.size(), &h, &w);
    if (rc) return rc;
    return ai4e_engine_submit(e, px.data(), h, w, 3, 0, tag, out_ticket);
  }
  set_last_error("submit_image: neither a JPEG nor a PNG stream");
  return AI4E_ERR_UNSUPPORTED;
}

int ai4e_engine_submit_batch(ai4e_engine* e, const uint8_t* h_images, int count, int h, int w, int c,
                             uint64_t* out_first_ticket) {
  AI4E_REQUIRE(e && h_images && out_first_ticket, "submit_batch: null pointer");
  AI4E_REQUIRE(c == e->desc.in_c, "submit_batch: images have %d channels, model takes %d", c, e->desc.in_c);
  AI4E_REQUIRE(count >= 1 && count <= e->max_batch, "submit_batch: count %d outside [1, %d]", count, e->max_batch);
  AI4E_REQUIRE(h > 0 && w > 0, "submit_batch: empty image");
  const size_t img_bytes = (size_t)h * w * c;
  AI4E_REQUIRE(img_bytes <= e->slot_bytes, "submit_batch: image of %zu bytes exceeds the staging slot (%zu)",
               img_bytes, e->slot_bytes);
  const bool pinned = is_page_locked(h_images);
  BatchCtx* ctx = nullptr;
  {
    std::unique_lock<std::mutex> lk(e->mu);
    for (;;) {
      if (e->stopping) { set_last_error("submit_batch: engine is shutting down"); return AI4E_ERR_STATE; }
      if (!e->healthy.load()) { set_last_error("submit_batch: engine on device %d is unhealthy", e->device); return AI4E_ERR_STATE; }
      if (e->open_ctx && e->open_ctx->count > 0) {  // a caller-made batch is never mixed with singles
        e->closed_q.push_back(e->open_ctx);
        e->open_ctx = nullptr;
        e->cv_work.notify_one();
      }
      if (!e->open_ctx) {
        if (e->free_q.empty()) { e->cv_free.wait(lk); continue; }
        e->open_ctx = e->free_q.front();
        e->free_q.pop_front();
      }
      break;
    }
    ctx = e->open_ctx;
    e->open_ctx = nullptr;
    ctx->count = count; ctx->src_h = h; ctx->src_w = w; ctx->kind = 0;
    ctx->tile = nullptr;
    ctx->t_first = Clock::now();
    ctx->first_ticket = e->next_ticket;
    ctx->ext_src = pinned ? h_images : nullptr;
    ctx->copied.store(pinned ? count : 0);
    *out_first_ticket = e->next_ticket;
    for (int i = 0; i < count; ++i) {
      const uint64_t r = (e->next_ticket + i) % kResultRing;
      e->r_submit[r] = ctx->t_first;
      e->r_status[r] = 1;
      e->r_notify[r] = 0;
      e->r_tag[r] = 0;
    }
    e->next_ticket += count;
    e->n_submitted.fetch_add((uint64_t)count);
    e->closed_q.push_back(ctx);
    e->cv_work.notify_one();
  }
  if (!pinned) {
    memcpy(ctx->h_in, h_images, img_bytes * count);
    ctx->copied.store(count, std::memory_order_release);
  }
  return AI4E_OK;
}

int ai4e_engine_profile(ai4e_engine* e, const uint8_t* d_images, int batch, int src_h, int src_w, int reps,
                        float* h_ms_out, int n_out) {
  AI4E_REQUIRE(e && d_images && h_ms_out, "profile: null pointer");
  AI4E_REQUIRE(batch >= 1 && batch <= e->max_batch && reps >= 1, "profile: bad batch/reps");
  const int n = (int)e->layers.size() + 1;  // preprocess + every layer
  AI4E_REQUIRE(n_out >= n, "profile: need room for %d timings", n);
  AI4E_CUDA_TRY(cudaSetDevice(e->device));
  std::vector<cudaEvent_t> ev(n + 1);
  for (auto& x : ev) AI4E_CUDA_TRY(cudaEventCreate(&x));
  std::vector<double> acc(n, 0.0);
  std::lock_guard<std::mutex> lk0(e->sync_mu);
  BatchCtx& c = e->sync_ctx;
  int rc = AI4E_OK;
  for (int r = 0; r < reps && rc == AI4E_OK; ++r) {
    {
      std::lock_guard<std::mutex> lk(e->launch_mu);
      Lane* ln = nullptr;
      rc = lane_begin(e, c.stream, &ln, 0);  // lane 0 holds the autotuned plans' buffers; any lane would do
      if (rc == AI4E_OK) rc = forward_enqueue(e, ln, d_images, batch, src_h, src_w, c.d_idx, c.d_prob, c.d_logits, c.stream, ev.data());
      if (rc == AI4E_OK) rc = lane_done(ln, c.stream);
    }
    if (rc) break;
    cudaError_t ce = cudaStreamSynchronize(c.stream);
    if (ce != cudaSuccess) { set_last_error("profile: %s", cudaGetErrorString(ce)); rc = AI4E_ERR_CUDA; break; }
    for (int i = 0; i < n; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
      acc[i] += ms;
    }
  }
  for (auto& x : ev) cudaEventDestroy(x);
  if (rc) return rc;
  for (int i = 0; i < n; ++i) h_ms_out[i] = (float)(acc[i] / reps);
  return AI4E_OK;
}

// timeout_us < 0: wait for ever; 0: only report what is there
int ai4e_engine_wait_for(ai4e_engine* e, uint64_t ticket, int64_t timeout_us, int32_t* top_idx5, float* top_prob5,
                         float* logits_or_null) {
  AI4E_REQUIRE(e && top_idx5 && top_prob5, "wait: null pointer");
  std::unique_lock<std::mutex> lk(e->mu);
  AI4E_REQUIRE(ticket < e->next_ticket, "wait: unknown ticket %llu", (unsigned long long)ticket);
  auto ready = [&] { return e->done_upto > ticket || e->completer_done; };
  if (timeout_us < 0) {
    e->cv_done.wait(lk, ready);
  } else if (!e->cv_done.wait_for(lk, std::chrono::microseconds(timeout_us), ready)) {
    set_last_error("wait: ticket %llu not finished after %lld us (%llu tickets in the queue ahead of it)",
                   (unsigned long long)ticket, (long long)timeout_us, (unsigned long long)(ticket - e->done_upto));
    return AI4E_ERR_TIMEOUT;
  }
  if (e->done_upto <= ticket) { set_last_error("wait: engine shut down before ticket completed"); return AI4E_ERR_STATE; }
  AI4E_REQUIRE(e->next_ticket - ticket <= kResultRing, "wait: result of ticket %llu has been overwritten",
               (unsigned long long)ticket);
  const uint64_t r = ticket % kResultRing;
  const int rc = e->r_status[r];
  if (rc != AI4E_OK) { set_last_error("wait: the batch of ticket %llu failed (%d)", (unsigned long long)ticket, rc); return rc < 0 ? rc : AI4E_ERR_STATE; }
  memcpy(top_idx5, &e->r_idx[r * AI4E_TOPK], AI4E_TOPK * sizeof(int32_t));
  memcpy(top_prob5, &e->r_prob[r * AI4E_TOPK], AI4E_TOPK * sizeof(float));
  if (logits_or_null) {
    AI4E_REQUIRE(e->next_ticket - ticket <= kLogitRing, "wait: logits of ticket %llu have been overwritten",
                 (unsigned long long)ticket);
    const int K = e->desc.num_classes;
    memcpy(logits_or_null, &e->r_logits[(ticket % kLogitRing) * K], K * sizeof(float));
  }
  return AI4E_OK;
}

int ai4e_engine_wait(ai4e_engine* e, uint64_t ticket, int32_t* top_idx5, float* top_prob5, float* logits_or_null) {
  AI4E_REQUIRE(e, "wait: null engine");
  return ai4e_engine_wait_for(e, ticket, (int64_t)e->wait_timeout_ms * 1000, top_idx5, top_prob5, logits_or_null);
}

// Drains up to `capacity` announced completions from `cq`, blocking up to timeout_us for the first one.
static int poll_queue(CompletionQueue* cq, int capacity, int64_t timeout_us, uint64_t* tickets, uint64_t* tags,
                      int32_t* status, int32_t* top_idx5, float* top_prob5, int* out_n) {
  std::unique_lock<std::mutex> lk(cq->mu);
  if (cq->q.empty() && timeout_us != 0) {
    if (timeout_us < 0) cq->cv.wait(lk, [&] { return !cq->q.empty(); });
    else cq->cv.wait_for(lk, std::chrono::microseconds(timeout_us), [&] { return !cq->q.empty(); });
  }
  int n = 0;
  while (n < capacity && !cq->q.empty()) {
    const DoneRec& d = cq->q.front();
    if (tickets) tickets[n] = d.ticket;
    if (tags) tags[n] = d.tag;
    if (status) status[n] = d.status;
    if (top_idx5) memcpy(top_idx5 + (size_t)n * AI4E_TOPK, d.idx, sizeof(d.idx));
    if (top_prob5) memcpy(top_prob5 + (size_t)n * AI4E_TOPK, d.prob, sizeof(d.prob));
    cq->q.pop_front();
    ++n;
  }
  *out_n = n;
  return AI4E_OK;
}

int ai4e_engine_poll(ai4e_engine* e, int capacity, int64_t timeout_us, uint64_t* tickets, uint64_t* tags, int32_t* status,
                     int32_t* top_idx5, float* top_prob5, int* out_n) {
  AI4E_REQUIRE(e && out_n && capacity >= 0, "poll: bad arguments");
  return poll_queue(e->cq, capacity, timeout_us, tickets, tags, status, top_idx5, top_prob5, out_n);
}

int ai4e_engine_debug_inject_fault(ai4e_engine* e, int n_batches) {
  AI4E_REQUIRE(e && n_batches >= 0, "inject_fault: bad arguments");
  e->inject_faults.store(n_batches);
  return AI4E_OK;
}

int ai4e_engine_flush(ai4e_engine* e) {
  AI4E_REQUIRE(e, "flush: null engine");
  std::unique_lock<std::mutex> lk(e->mu);
  const uint64_t target = e->next_ticket;
  e->flush_req = true;
  e->cv_work.notify_all();
  e->cv_done.wait(lk, [&] { return e->done_upto >= target || e->completer_done; });
  return AI4E_OK;
}

int ai4e_engine_stats_get(ai4e_engine* e, ai4e_engine_stats* out) {
  AI4E_REQUIRE(e && out, "stats: null pointer");
  std::lock_guard<std::mutex> lk(e->mu);
  *out = e->st;
  out->queue_depth = e->next_ticket - e->done_upto;  // == n_submitted - n_completed
  out->healthy = e->healthy.load();
  out->launches_per_forward = e->launches_u8 > 0 ? e->launches_u8 : e->launches;  // the serving path's usual case
  const uint64_t n = std::min<uint64_t>(e->lat_n, kLatRing);
  if (n) {
    std::vector<float> v(e->lat_ms.begin(), e->lat_ms.begin() + n);
    std::sort(v.begin(), v.end());
    out->lat_ms_p50 = v[(size_t)((n - 1) * 0.50)];
    out->lat_ms_p99 = v[(size_t)((n - 1) * 0.99)];
  }
  return AI4E_OK;
}

int ai4e_engine_read_activation(ai4e_engine* e, int buf, void* h_dst, uint64_t bytes) {
  AI4E_REQUIRE(e && h_dst, "read_activation: null pointer");
  std::lock_guard<std::mutex> lk_lane(e->launch_mu);
  Lane& lane_r = e->lanes[e->last_lane];  // the lane the most recent forward ran on
  AI4E_REQUIRE(buf >= 0 && buf < (int)lane_r.bufs.size() && lane_r.bufs[buf], "read_activation: no slot %d", buf);
  AI4E_REQUIRE(bytes <= e->buf_elems_per_image[buf] * 2 * (uint64_t)e->max_batch, "read_activation: too many bytes");
  AI4E_CUDA_TRY(cudaSetDevice(e->device));
  AI4E_CUDA_TRY(cudaDeviceSynchronize());
  AI4E_CUDA_TRY(cudaMemcpy(h_dst, lane_r.bufs[buf], bytes, cudaMemcpyDeviceToHost));
  return AI4E_OK;
}

// ---------------------------------------------------------------------------------------- tiles (K6)
// Chips [first_chip, first_chip + n_chips) of a tile cut into chip x chip squares (row-major over the
// chip grid).  The band of chip rows that holds them crosses PCIe ONCE (one 2-D copy on the copy
// stream into a device tile buffer); the batches made from it carry only (first chip, count) and K1
// reads every chip in place through the chip-grid strides - no per-chip host copy, no gather kernel.
// With out_sum_ticket the chips' results are also aggregated on the device (K7) into one tile summary, waited for
// with ai4e_engine_wait_tile_summary; out_tickets may then be null (nobody wants the per-chip records).
static int submit_tile_range(ai4e_engine* e, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride,
                             int chip, int first_chip, int n_chips, uint64_t* out_tickets, uint64_t* out_sum_ticket = nullptr) {
  AI4E_REQUIRE(e && h_tile && (out_tickets || out_sum_ticket), "submit_tile: null pointer");
  AI4E_REQUIRE(c == e->desc.in_c, "submit_tile: tile has %d channels, model takes %d", c, e->desc.in_c);
  AI4E_REQUIRE(chip > 0 && tile_h >= chip && tile_w >= chip, "submit_tile: tile %d x %d smaller than a %d chip", tile_h, tile_w, chip);
  AI4E_REQUIRE(row_stride >= 0, "submit_tile: negative row stride");
  if (row_stride == 0) row_stride = tile_w * c;
  AI4E_REQUIRE(row_stride >= tile_w * c, "submit_tile: row stride too small");
  const int gh = tile_h / chip, gw = tile_w / chip;
  AI4E_REQUIRE(first_chip >= 0 && n_chips >= 1 && first_chip + n_chips <= gh * gw, "submit_tile: chips [%d, %d) outside the %d x %d grid",
               first_chip, first_chip + n_chips, gh, gw);
  const int r0 = first_chip / gw, r1 = (first_chip + n_chips - 1) / gw;
  const size_t band_rows = (size_t)(r1 - r0 + 1) * chip;
  const size_t row_bytes = ((size_t)gw * chip * c + 15) & ~(size_t)15;  // device rows: only the columns chips cover, 16-byte aligned
  const size_t band_bytes = band_rows * row_bytes;
  AI4E_CUDA_TRY(cudaSetDevice(e->device));

  // ---- a free tile buffer
  TileBuf* tb = nullptr;
  {
    std::unique_lock<std::mutex> lk(e->mu);
    for (;;) {
      if (e->stopping) { set_last_error("submit_tile: engine is shutting down"); return AI4E_ERR_STATE; }
      if (!e->healthy.load()) { set_last_error("submit_tile: engine on device %d is unhealthy", e->device); return AI4E_ERR_STATE; }
      for (TileBuf& t : e->tiles)
        if (!t.busy && t.users == 0) { tb = &t; break; }
      if (tb) break;
      e->cv_free.wait(lk);
    }
    tb->busy = true;
  }
  auto release = [&](int rc) {
    std::lock_guard<std::mutex> lk(e->mu);
    tb->busy = false;
    e->cv_free.notify_all();
    return rc;
  };
  if (tb->bytes < band_bytes) {  // first tile of this size on this buffer
    if (tb->d) cudaFree(tb->d);
    tb->d = nullptr; tb->bytes = 0;
    cudaError_t ce = cudaMalloc((void**)&tb->d, band_bytes);
    if (ce != cudaSuccess) { set_last_error("submit_tile: %zu bytes of tile buffer: %s", band_bytes, cudaGetErrorString(ce)); cudaGetLastError(); return release(AI4E_ERR_NOMEM); }
    tb->bytes = band_bytes;
  }
  if (out_sum_ticket) {  // K7: a zeroed aggregate next to the band, ready (like the band) when ev_ready fires
    const SumLayout sl(n_chips, e->desc.num_classes);
    if (tb->sum_cap < sl.bytes) {
      if (tb->d_sum) cudaFree(tb->d_sum);
      if (tb->h_sum) cudaFreeHost(tb->h_sum);
      tb->d_sum = tb->h_sum = nullptr; tb->sum_cap = 0;
      cudaError_t ce = cudaMalloc((void**)&tb->d_sum, sl.bytes);
      if (ce == cudaSuccess) ce = cudaHostAlloc((void**)&tb->h_sum, sl.bytes, cudaHostAllocDefault);
      if (ce != cudaSuccess) { set_last_error("submit_tile: %zu bytes of summary buffer: %s", sl.bytes, cudaGetErrorString(ce)); cudaGetLastError(); return release(AI4E_ERR_NOMEM); }
      tb->sum_cap = sl.bytes;
    }
    cudaError_t ce = cudaMemsetAsync(tb->d_sum, 0, sl.bytes, e->copy_stream);
    if (ce != cudaSuccess) { set_last_error("submit_tile: summary memset failed: %s", cudaGetErrorString(ce)); return release(AI4E_ERR_CUDA); }
    std::lock_guard<std::mutex> lk(e->mu);
    tb->summary = true; tb->sum_chips = n_chips; tb->sum_rc = 0;
    tb->sum_left = (n_chips + e->max_batch - 1) / e->max_batch;
    tb->sum_ticket = e->next_sum_ticket++;
    TileSumRec& sr = e->sums[tb->sum_ticket % kSumRing];
    sr.id = tb->sum_ticket; sr.status = 1; sr.n = n_chips; sr.data.clear();
    *out_sum_ticket = tb->sum_ticket;
  }
  // page-locked source: asynchronous, keep the tile unchanged until its tickets are done; pageable
  // source: the runtime stages it and the tile is free on return
  cudaError_t ce = cudaMemcpy2DAsync(tb->d, row_bytes, h_tile + (size_t)r0 * chip * row_stride, (size_t)row_stride,
                                     (size_t)gw * chip * c, band_rows, cudaMemcpyHostToDevice, e->copy_stream);
  if (ce == cudaSuccess) ce = cudaEventRecord(tb->ev_ready, e->copy_stream);
  if (ce != cudaSuccess) {
    set_last_error("submit_tile: H2D copy failed: %s", cudaGetErrorString(ce));
    if (out_sum_ticket) { std::lock_guard<std::mutex> lk(e->mu); tb->summary = false; e->sums[tb->sum_ticket % kSumRing].status = AI4E_ERR_CUDA; }
    return release(AI4E_ERR_CUDA);
  }

  // ---- batches over the band
  int done = 0;
  while (done < n_chips) {
    const int count = std::min(n_chips - done, e->max_batch);
    std::unique_lock<std::mutex> lk(e->mu);
    for (;;) {
      if (e->stopping || !e->healthy.load()) {
        tb->busy = false;
        if (out_sum_ticket) {  // the batches that will never be queued: the summary fails instead of waiting for them
          TileSumRec& sr = e->sums[tb->sum_ticket % kSumRing];
          if (sr.id == tb->sum_ticket) sr.status = AI4E_ERR_STATE;
          if (tb->users == 0) tb->summary = false;
          else { tb->sum_rc = AI4E_ERR_STATE; tb->sum_left -= (n_chips - done + e->max_batch - 1) / e->max_batch; }
        }
        e->cv_free.notify_all();
        e->cv_done.notify_all();
        set_last_error("submit_tile: engine stopped or unhealthy");
        return AI4E_ERR_STATE;
      }
      if (e->open_ctx && e->open_ctx->count > 0) {  // never mixed with single submits
        e->closed_q.push_back(e->open_ctx);
        e->open_ctx = nullptr;
        e->cv_work.notify_one();
      }
      if (!e->open_ctx) {
        if (e->free_q.empty()) { e->cv_free.wait(lk); continue; }
        e->open_ctx = e->free_q.front();
        e->free_q.pop_front();
      }
      break;
    }
    BatchCtx* ctx = e->open_ctx;
    e->open_ctx = nullptr;
    ctx->count = count; ctx->src_h = chip; ctx->src_w = chip; ctx->kind = kKindTile;
    ctx->tile = tb; ctx->tile_first = first_chip - r0 * gw + done; ctx->tile_grid_w = gw; ctx->tile_chip = chip;
    ctx->tile_row_bytes = (long long)row_bytes;
    ctx->sum_first = done;
    ctx->ext_src = nullptr;
    ctx->t_first = Clock::now();
    ctx->first_ticket = e->next_ticket;
    ctx->copied.store(count);
    tb->users += 1;
    for (int i = 0; i < count; ++i) {
      const uint64_t t = e->next_ticket++;
      const uint64_t r = t % kResultRing;
      if (out_tickets) out_tickets[done + i] = t;
      e->r_submit[r] = ctx->t_first;
      e->r_status[r] = 1;
      e->r_notify[r] = 0;
      e->r_tag[r] = 0;
    }
    e->n_submitted.fetch_add((uint64_t)count);
    e->closed_q.push_back(ctx);
    e->cv_work.notify_one();
    done += count;
  }
  return release(AI4E_OK);
}

int ai4e_engine_submit_tile(ai4e_engine* e, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride,
                            int chip, uint64_t* out_tickets, int capacity, int* out_count) {
  AI4E_REQUIRE(e && h_tile && out_tickets && out_count, "submit_tile: null pointer");
  AI4E_REQUIRE(chip > 0 && tile_h >= chip && tile_w >= chip, "submit_tile: tile %d x %d smaller than a %d chip", tile_h, tile_w, chip);
  const int total = (tile_h / chip) * (tile_w / chip);
  AI4E_REQUIRE(capacity >= total, "submit_tile: %d tickets needed, room for %d", total, capacity);
  *out_count = 0;
  int rc = submit_tile_range(e, h_tile, tile_h, tile_w, c, row_stride, chip, 0, total, out_tickets);
  if (rc) return rc;
  *out_count = total;
  return AI4E_OK;
}

int ai4e_engine_submit_tile_summary(ai4e_engine* e, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride,
                                    int chip, uint64_t* out_tile_ticket, int* out_chips) {
  AI4E_REQUIRE(e && h_tile && out_tile_ticket, "submit_tile_summary: null pointer");
  AI4E_REQUIRE(chip > 0 && tile_h >= chip && tile_w >= chip, "submit_tile_summary: tile %d x %d smaller than a %d chip", tile_h, tile_w, chip);
  const int total = (tile_h / chip) * (tile_w / chip);
  int rc = submit_tile_range(e, h_tile, tile_h, tile_w, c, row_stride, chip, 0, total, nullptr, out_tile_ticket);
  if (rc) return rc;
  if (out_chips) *out_chips = total;
  return AI4E_OK;
}

// The summary of one tile (or of one replica's chip range): any output may be null.  labels / label_prob hold
// `capacity_chips` entries at least; hist / mass_q32 hold num_classes.
static int wait_tile_summary(ai4e_engine* e, uint64_t ticket, int64_t timeout_us, int32_t* labels, float* label_prob,
                             int capacity_chips, uint32_t* hist, uint64_t* mass_q32, bool accumulate, int* out_chips) {
  AI4E_REQUIRE(e, "wait_tile_summary: null engine");
  std::unique_lock<std::mutex> lk(e->mu);
  AI4E_REQUIRE(ticket < e->next_sum_ticket, "wait_tile_summary: unknown tile ticket %llu", (unsigned long long)ticket);
  TileSumRec& sr = e->sums[ticket % kSumRing];
  AI4E_REQUIRE(sr.id == ticket, "wait_tile_summary: the summary of tile ticket %llu has been overwritten (%d are kept)",
               (unsigned long long)ticket, kSumRing);
  auto ready = [&] { return sr.id != ticket || sr.status != 1 || e->completer_done; };
  if (timeout_us < 0) {
    e->cv_done.wait(lk, ready);
  } else if (!e->cv_done.wait_for(lk, std::chrono::microseconds(timeout_us), ready)) {
    set_last_error("wait_tile_summary: tile ticket %llu not finished after %lld us", (unsigned long long)ticket, (long long)timeout_us);
    return AI4E_ERR_TIMEOUT;
  }
  AI4E_REQUIRE(sr.id == ticket, "wait_tile_summary: the summary of tile ticket %llu has been overwritten", (unsigned long long)ticket);
  if (sr.status == 1) { set_last_error("wait_tile_summary: engine shut down before the tile completed"); return AI4E_ERR_STATE; }
  if (sr.status != AI4E_OK) {
    set_last_error("wait_tile_summary: a batch of tile ticket %llu failed (%d)", (unsigned long long)ticket, sr.status);
    return sr.status < 0 ? sr.status : AI4E_ERR_STATE;
  }
  const int K = e->desc.num_classes;
  const SumLayout sl(sr.n, K);
  AI4E_REQUIRE(sr.data.size() == sl.bytes, "wait_tile_summary: internal size mismatch");
  if (out_chips) *out_chips = sr.n;
  if (labels || label_prob) AI4E_REQUIRE(capacity_chips >= sr.n, "wait_tile_summary: %d chips, room for %d", sr.n, capacity_chips);
  if (labels) memcpy(labels, sr.data.data() + sl.labels, (size_t)sr.n * 4);
  if (label_prob) memcpy(label_prob, sr.data.data() + sl.prob, (size_t)sr.n * 4);
  const uint32_t* h = (const uint32_t*)(sr.data.data() + sl.hist);
  const uint64_t* m = (const uint64_t*)(sr.data.data() + sl.mass);
  if (hist) for (int k = 0; k < K; ++k) hist[k] = (accumulate ? hist[k] : 0u) + h[k];
  if (mass_q32) for (int k = 0; k < K; ++k) mass_q32[k] = (accumulate ? mass_q32[k] : 0ull) + m[k];
  return AI4E_OK;
}

int ai4e_engine_wait_tile_summary(ai4e_engine* e, uint64_t tile_ticket, int32_t* labels, float* label_prob, int capacity_chips,
                                  uint32_t* hist, uint64_t* mass_q32, int* out_chips) {
  AI4E_REQUIRE(e, "wait_tile_summary: null engine");
  return wait_tile_summary(e, tile_ticket, (int64_t)e->wait_timeout_ms * 1000, labels, label_prob, capacity_chips, hist,
                           mass_q32, false, out_chips);
}

// results of `n` tickets in order: idx [n][5], prob [n][5]
int ai4e_engine_wait_many(ai4e_engine* e, const uint64_t* tickets, int n, int32_t* top_idx5, float* top_prob5) {
  AI4E_REQUIRE(e && tickets && top_idx5 && top_prob5 && n >= 0, "wait_many: bad arguments");
  for (int i = 0; i < n; ++i) {
    int rc = ai4e_engine_wait(e, tickets[i], top_idx5 + (size_t)i * AI4E_TOPK, top_prob5 + (size_t)i * AI4E_TOPK, nullptr);
    if (rc) return rc;
  }
  return AI4E_OK;
}

// ---------------------------------------------------------------------------------------- pool
// ticket = replica index in the top byte | replica-local ticket.  Routing: the healthy replica
// with the shortest queue, round-robin among equals (the reference: Istio ROUND_ROBIN).
// A tile summary dealt over the replicas: which replica summarises which chip range under which of its tickets.
struct PoolTileRec {
  uint64_t id = ~0ull;
  int total = 0;
  struct Part { int engine, first, n; uint64_t ticket; };
  std::vector<Part> parts;
};

struct ai4e_pool {
  std::vector<ai4e_engine*> engines;
  std::atomic<uint64_t> rr{0};
  CompletionQueue cq;  // shared by the replicas: one reaper thread serves all GPUs
  std::mutex sum_mu;
  PoolTileRec sums[kSumRing];
  uint64_t next_sum_ticket = 0;
};

static constexpr uint64_t kLocalMask = (1ull << 56) - 1;

int ai4e_pool_create(const int* devices, int n_devices, const ai4e_model_desc* desc, const void* h_weights_blob,
                     uint64_t blob_bytes, int max_batch, int max_wait_us, ai4e_pool** out) {
  AI4E_REQUIRE(devices && n_devices >= 1 && n_devices <= 64 && out, "pool_create: bad arguments");
  ai4e_pool* p = new ai4e_pool();
  for (int i = 0; i < n_devices; ++i) {
    ai4e_engine* e = nullptr;
    int rc = ai4e_engine_create(devices[i], desc, h_weights_blob, blob_bytes, max_batch, max_wait_us, &e);
    if (rc) { ai4e_pool_destroy(p); return rc; }
    e->cq = &p->cq;
    e->ticket_prefix = (uint64_t)i << 56;
    p->engines.push_back(e);
  }
  *out = p;
  return AI4E_OK;
}

int ai4e_pool_destroy(ai4e_pool* p) {
  if (!p) return AI4E_OK;
  for (ai4e_engine* e : p->engines) ai4e_engine_destroy(e);
  delete p;
  return AI4E_OK;
}

// healthy replica with the shortest queue, round-robin among equals (-1: none).  Reads two atomics
// per replica: no engine lock is taken on the routing path.
static int pool_pick(ai4e_pool* p) {
  const size_t n = p->engines.size();
  const uint64_t start = p->rr.fetch_add(1);
  int best = -1;
  uint64_t best_depth = ~0ull;
  for (size_t k = 0; k < n; ++k) {
    const size_t i = (start + k) % n;
    ai4e_engine* e = p->engines[i];
    if (!e->healthy.load()) continue;
    const uint64_t done = e->n_completed.load(), sub = e->n_submitted.load();
    const uint64_t depth = sub > done ? sub - done : 0;
    if (depth < best_depth) { best_depth = depth; best = (int)i; }
  }
  return best;
}

#define POOL_PICK_OR_FAIL(who)                                                       \
  const int best = pool_pick(p);                                                     \
  if (best < 0) { set_last_error(who ": no healthy replica"); return AI4E_ERR_STATE; }

int ai4e_pool_submit(ai4e_pool* p, const uint8_t* h_hwc, int h, int w, int c, int row_stride, uint64_t tag,
                     uint64_t* out_ticket) {
  AI4E_REQUIRE(p && out_ticket, "pool_submit: null pointer");
  POOL_PICK_OR_FAIL("pool_submit");
  uint64_t t = 0;
  int rc = ai4e_engine_submit(p->engines[best], h_hwc, h, w, c, row_stride, tag, &t);
  if (rc) return rc;
  *out_ticket = ((uint64_t)best << 56) | (t & kLocalMask);
  return AI4E_OK;
}

int ai4e_pool_submit_jpeg(ai4e_pool* p, const uint8_t* jpeg, uint64_t len, uint64_t tag, uint64_t* out_ticket) {
  AI4E_REQUIRE(p && out_ticket, "pool_submit_jpeg: null pointer");
  POOL_PICK_OR_FAIL("pool_submit_jpeg");
  uint64_t t = 0;
  int rc = ai4e_engine_submit_jpeg(p->engines[best], jpeg, len, tag, &t);
  if (rc) return rc;
  *out_ticket = ((uint64_t)best << 56) | (t & kLocalMask);
  return AI4E_OK;
}

int ai4e_pool_submit_image(ai4e_pool* p, const uint8_t* data, uint64_t len, uint64_t tag, uint64_t* out_ticket) {
  AI4E_REQUIRE(p && out_ticket, "pool_submit_image: null pointer");
  POOL_PICK_OR_FAIL("pool_submit_image");
  uint64_t t = 0;
  int rc = ai4e_engine_submit_image(p->engines[best], data, len, tag, &t);
  if (rc) return rc;
  *out_ticket = ((uint64_t)best << 56) | (t & kLocalMask);
  return AI4E_OK;
}

// n images in one call: dealt to the replicas in runs (n / replicas, at most a micro-batch), each run
// to the replica with the shortest queue at that moment.
int ai4e_pool_submit_many(ai4e_pool* p, const uint8_t* const* h_images, int n, int h, int w, int c, int row_stride,
                          const uint64_t* tags_or_null, int notify, uint64_t* out_tickets) {
  AI4E_REQUIRE(p && h_images && out_tickets && n >= 0, "pool_submit_many: bad arguments");
  const int ne = (int)p->engines.size();
  const int run = std::max(1, std::min(p->engines[0]->max_batch, (n + ne - 1) / ne));
  for (int i = 0; i < n; i += run) {
    const int k = std::min(run, n - i);
    POOL_PICK_OR_FAIL("pool_submit_many");
    int rc = ai4e_engine_submit_many(p->engines[best], h_images + i, k, h, w, c, row_stride,
                                     tags_or_null ? tags_or_null + i : nullptr, notify, out_tickets + i);
    if (rc) return rc;
    for (int j = 0; j < k; ++j) out_tickets[i + j] = ((uint64_t)best << 56) | (out_tickets[i + j] & kLocalMask);
  }
  return AI4E_OK;
}

int ai4e_pool_submit_batch(ai4e_pool* p, const uint8_t* h_images, int count, int h, int w, int c,
                           uint64_t* out_first_ticket) {
  AI4E_REQUIRE(p && out_first_ticket, "pool_submit_batch: null pointer");
  POOL_PICK_OR_FAIL("pool_submit_batch");
  uint64_t t = 0;
  int rc = ai4e_engine_submit_batch(p->engines[best], h_images, count, h, w, c, &t);
  if (rc) return rc;
  *out_first_ticket = ((uint64_t)best << 56) | (t & kLocalMask);  // the batch's tickets are consecutive
  return AI4E_OK;
}

// The chips of a tile in contiguous ranges, one range per healthy replica (256 chips on 8 GPUs: 32
// each); every replica receives only the band of chip rows it needs.
static int pool_submit_tile_any(ai4e_pool* p, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride, int chip,
                                uint64_t* out_tickets, int capacity, int* out_count, PoolTileRec* rec);

int ai4e_pool_submit_tile(ai4e_pool* p, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride, int chip,
                          uint64_t* out_tickets, int capacity, int* out_count) {
  AI4E_REQUIRE(p && h_tile && out_tickets && out_count, "pool_submit_tile: null pointer");
  return pool_submit_tile_any(p, h_tile, tile_h, tile_w, c, row_stride, chip, out_tickets, capacity, out_count, nullptr);
}

// The same deal with the K7 aggregate: every replica summarises its own chip range on its GPU; the ranges are
// put together on the host when the summary is waited for (labels side by side, histogram and mass added - integer
// sums, so the result does not depend on how many replicas shared the tile).
int ai4e_pool_submit_tile_summary(ai4e_pool* p, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride, int chip,
                                  uint64_t* out_tile_ticket, int* out_chips) {
  AI4E_REQUIRE(p && h_tile && out_tile_ticket, "pool_submit_tile_summary: null pointer");
  PoolTileRec rec;
  int count = 0;
  const int rc = pool_submit_tile_any(p, h_tile, tile_h, tile_w, c, row_stride, chip, nullptr, 0, &count, &rec);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(p->sum_mu);
  rec.id = p->next_sum_ticket++;
  rec.total = count;
  *out_tile_ticket = rec.id;
  if (out_chips) *out_chips = count;
  p->sums[rec.id % kSumRing] = std::move(rec);
  return AI4E_OK;
}

int ai4e_pool_wait_tile_summary(ai4e_pool* p, uint64_t tile_ticket, int32_t* labels, float* label_prob, int capacity_chips,
                                uint32_t* hist, uint64_t* mass_q32, int* out_chips) {
  AI4E_REQUIRE(p, "pool_wait_tile_summary: null pool");
  PoolTileRec rec;
  {
    std::lock_guard<std::mutex> lk(p->sum_mu);
    AI4E_REQUIRE(tile_ticket < p->next_sum_ticket, "pool_wait_tile_summary: unknown tile ticket %llu", (unsigned long long)tile_ticket);
    rec = p->sums[tile_ticket % kSumRing];
  }
  AI4E_REQUIRE(rec.id == tile_ticket, "pool_wait_tile_summary: the summary of tile ticket %llu has been overwritten (%d are kept)",
               (unsigned long long)tile_ticket, kSumRing);
  if (labels || label_prob) AI4E_REQUIRE(capacity_chips >= rec.total, "pool_wait_tile_summary: %d chips, room for %d", rec.total, capacity_chips);
  const int K = p->engines[0]->desc.num_classes;
  if (hist) memset(hist, 0, (size_t)K * sizeof(uint32_t));
  if (mass_q32) memset(mass_q32, 0, (size_t)K * sizeof(uint64_t));
  for (const PoolTileRec::Part& part : rec.parts) {
    ai4e_engine* e = p->engines[part.engine];
    const int rc = wait_tile_summary(e, part.ticket, (int64_t)e->wait_timeout_ms * 1000, labels ? labels + part.first : nullptr,
                                     label_prob ? label_prob + part.first : nullptr, part.n, hist, mass_q32, true, nullptr);
    if (rc) return rc;
  }
  if (out_chips) *out_chips = rec.total;
  return AI4E_OK;
}

static int pool_submit_tile_any(ai4e_pool* p, const uint8_t* h_tile, int tile_h, int tile_w, int c, int row_stride, int chip,
                                uint64_t* out_tickets, int capacity, int* out_count, PoolTileRec* rec) {
  AI4E_REQUIRE(chip > 0 && tile_h >= chip && tile_w >= chip, "pool_submit_tile: tile %d x %d smaller than a %d chip", tile_h, tile_w, chip);
  const int total = (tile_h / chip) * (tile_w / chip);
  AI4E_REQUIRE(rec || capacity >= total, "pool_submit_tile: %d tickets needed, room for %d", total, capacity);
  *out_count = 0;
  std::vector<int> alive;
  const uint64_t start = p->rr.fetch_add(1);
  for (size_t k = 0; k < p->engines.size(); ++k) {
    const size_t i = (start + k) % p->engines.size();
    if (p->engines[i]->healthy.load()) alive.push_back((int)i);
  }
  if (alive.empty()) { set_last_error("pool_submit_tile: no healthy replica"); return AI4E_ERR_STATE; }
  const int parts = (int)std::min<size_t>(alive.size(), (size_t)total);
  const int per = (total + parts - 1) / parts;
  std::vector<int> rcs(parts, AI4E_OK);
  std::vector<std::string> msgs(parts);
  std::vector<uint64_t> sum_tickets(parts, 0);
  auto work = [&](int k) {  // one submitter per replica: the H2D copies of the bands run side by side
    const int first = k * per, n = std::min(per, total - first);
    if (n <= 0) return;
    const int i = alive[k];
    const int rc = submit_tile_range(p->engines[i], h_tile, tile_h, tile_w, c, row_stride, chip, first, n,
                                     out_tickets ? out_tickets + first : nullptr, rec ? &sum_tickets[k] : nullptr);
    if (rc) { rcs[k] = rc; msgs[k] = ai4e_last_error(); return; }
    if (out_tickets)
      for (int j = 0; j < n; ++j) out_tickets[first + j] = ((uint64_t)i << 56) | (out_tickets[first + j] & kLocalMask);
  };
  std::vector<std::thread> th;
  for (int k = 1; k < parts; ++k) th.emplace_back(work, k);
  work(0);
  for (auto& t : th) t.join();
  for (int k = 0; k < parts; ++k)
    if (rcs[k]) { set_last_error("%s", msgs[k].c_str()); return rcs[k]; }
  if (rec)
    for (int k = 0; k < parts; ++k) {
      const int first = k * per, n = std::min(per, total - first);
      if (n > 0) rec->parts.push_back({alive[k], first, n, sum_tickets[k]});
    }
  *out_count = total;
  return AI4E_OK;
}

int ai4e_pool_wait_for(ai4e_pool* p, uint64_t ticket, int64_t timeout_us, int32_t* top_idx5, float* top_prob5,
                       float* logits_or_null) {
  AI4E_REQUIRE(p, "pool_wait: null pool");
  const size_t i = ticket >> 56;
  AI4E_REQUIRE(i < p->engines.size(), "pool_wait: bad ticket");
  return ai4e_engine_wait_for(p->engines[i], ticket & kLocalMask, timeout_us, top_idx5, top_prob5, logits_or_null);
}

int ai4e_pool_wait(ai4e_pool* p, uint64_t ticket, int32_t* top_idx5, float* top_prob5, float* logits_or_null) {
  AI4E_REQUIRE(p, "pool_wait: null pool");
  const size_t i = ticket >> 56;
  AI4E_REQUIRE(i < p->engines.size(), "pool_wait: bad ticket");
  return ai4e_engine_wait(p->engines[i], ticket & kLocalMask, top_idx5, top_prob5, logits_or_null);
}

int ai4e_pool_wait_many(ai4e_pool* p, const uint64_t* tickets, int n, int32_t* top_idx5, float* top_prob5) {
  AI4E_REQUIRE(p && tickets && top_idx5 && top_prob5 && n >= 0, "pool_wait_many: bad arguments");
  for (int i = 0; i < n; ++i) {
    int rc = ai4e_pool_wait(p, tickets[i], top_idx5 + (size_t)i * AI4E_TOPK, top_prob5 + (size_t)i * AI4E_TOPK, nullptr);
    if (rc) return rc;
  }
  return AI4E_OK;
}

int ai4e_pool_poll(ai4e_pool* p, int capacity, int64_t timeout_us, uint64_t* tickets, uint64_t* tags, int32_t* status,
                   int32_t* top_idx5, float* top_prob5, int* out_n) {
  AI4E_REQUIRE(p && out_n && capacity >= 0, "pool_poll: bad arguments");
  return poll_queue(&p->cq, capacity, timeout_us, tickets, tags, status, top_idx5, top_prob5, out_n);
}

int ai4e_pool_flush(ai4e_pool* p) {
  AI4E_REQUIRE(p, "pool_flush: null pool");
  for (ai4e_engine* e : p->engines) ai4e_engine_flush(e);
  return AI4E_OK;
}

int ai4e_pool_size(ai4e_pool* p) { return p ? (int)p->engines.size() : 0; }

int ai4e_pool_engine(ai4e_pool* p, int i, ai4e_engine** out) {
  AI4E_REQUIRE(p && out && i >= 0 && i < (int)p->engines.size(), "pool_engine: bad index");
  *out = p->engines[i];
  return AI4E_OK;
}

}  // extern "C"
