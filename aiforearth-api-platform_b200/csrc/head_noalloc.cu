// ai4e_head_noalloc - the classifier head (global average pool -> FC -> softmax -> top-k) for
// callers that are not allowed to allocate device memory.
//
// Who that is: a request thread in a serving container whose framework has already handed the
// whole GPU to its own pool allocator (a late cudaMalloc then fails or fights the pool), and any
// latency-critical path that cannot afford cudaMalloc / cudaFree because they synchronise the
// device.  `ai4e_head` needs batch*c floats of scratch for the pooled features and takes them
// from the stream's memory pool; this variant keeps that scratch in a fixed array that lives in
// the library's device image instead, so the call touches no allocator at all.
//
// Cost and limits, stated plainly:
//   * the array is AI4E_HEAD_NOALLOC_MAX_FLOATS floats (512 KiB); every GPU that loads the
//     library holds it until the process exits, used or not;
//   * batch*c above that limit is refused with AI4E_ERR_INVALID - no silent fallback;
//   * all calls on one device share the one array.  Nothing in it has to start at zero (the
//     pooling kernel overwrites batch*c floats before the FC reads them), but it keeps the last
//     call's contents, and two calls that overlap on the GPU (different streams, or two host
//     threads) race on it.  Serialise such calls - same stream, or an event between them;
//   * the first call on each device looks up the array's device address (below).  Make that
//     first call before any stream capture begins on that device.
#include "../../include/ai4e_b200.h"

#include "common.cuh"
#include "kernels.h"

#include <mutex>

using namespace ai4e;

namespace {

// In host code this name is only a handle for the runtime: it is NOT a device address, and a
// kernel handed `g_head_scratch` directly would fault (or, where the GPU can reach host memory,
// quietly use the wrong bytes).  Host code must translate it with cudaGetSymbolAddress.
static __device__ __align__(256) float g_head_scratch[AI4E_HEAD_NOALLOC_MAX_FLOATS];

constexpr int kMaxDevices = 64;
float* g_scratch_dev_ptr[kMaxDevices] = {};
std::mutex g_setup_mu;

// One-time set-up per device, with that device current: resolve the device address of the array.
int scratch_for_current_device(float** out) {
  int dev = -1;
  AI4E_CUDA_TRY(cudaGetDevice(&dev));
  AI4E_REQUIRE(dev >= 0 && dev < kMaxDevices, "head_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (!g_scratch_dev_ptr[dev]) {
    float* p = nullptr;
    const cudaError_t rc = cudaGetSymbolAddress((void**)&p, g_head_scratch);
    if (rc != cudaSuccess || !p) {
      set_last_error("head_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(rc));
      return AI4E_ERR_CUDA;
    }
    g_scratch_dev_ptr[dev] = p;
  }
  *out = g_scratch_dev_ptr[dev];
  return AI4E_OK;
}

}  // namespace

extern "C" int ai4e_head_noalloc(const void* d_in_nhwc, int batch, int hw, int c, const void* d_fc_w,
                                 const float* d_fc_b, int classes, float* d_logits, int32_t* d_top_idx,
                                 float* d_top_prob, int topk, void* stream) {
  AI4E_REQUIRE(batch > 0 && c > 0, "head_noalloc: bad geometry");
  AI4E_REQUIRE((long long)batch * c <= (long long)AI4E_HEAD_NOALLOC_MAX_FLOATS,
               "head_noalloc: batch*c = %lld floats of scratch needed, the built-in array holds %d "
               "(use ai4e_head, or split the batch)",
               (long long)batch * c, AI4E_HEAD_NOALLOC_MAX_FLOATS);
  float* scratch = nullptr;
  const int rc = scratch_for_current_device(&scratch);
  if (rc) return rc;
  return launch_head(d_in_nhwc, batch, hw, c, d_fc_w, d_fc_b, classes, scratch, d_logits, d_top_idx, d_top_prob,
                     topk, (cudaStream_t)stream);
}
