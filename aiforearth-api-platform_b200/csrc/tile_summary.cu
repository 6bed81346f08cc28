// K7 - on-device aggregate of a tile's per-chip results (tile -> chips -> classify ->
// aggregate without returning to the host between stages).  The reference chains such stages as separate
// tasks, one HTTP hop and one task-store round trip per stage
// (APIs/1.0/Common/task_management/distributed_api_task.py:67-100, AddPipelineTask); here the aggregate is
// one more launch on the stream that classified the chips, and what leaves the GPU is the tile's summary.
//
// From the head's top-5 of `count` chips (chips [first, first + count) of the tile):
//   labels[first + i]     = top-1 class of chip i                      (the tile's label map, row-major chip grid)
//   label_prob[first + i] = its probability
//   hist[k]              += 1 for every chip whose top-1 class is k    (class histogram)
//   mass_q32[k]          += round(p * 2^32) for every (chip, top-5 entry) of class k with probability p
// The probability mass is accumulated in 32.32 fixed point with integer atomics, so the sum does not depend on
// the order in which the batches of a tile (on different streams, different replicas) arrive: p * 2^32 is
// exact in fp32 and converts exactly (p <= 1 has 24 significant bits).  Oracle: oracle/tile_summary_oracle.py.
//
// HBM-bound and tiny (40 bytes in, 8 bytes + <= 6 atomics out per chip): one CTA per 256 chips.
#include "../../include/ai4e_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace ai4e {

namespace {

__global__ void tile_summary_kernel(const int32_t* __restrict__ idx5, const float* __restrict__ prob5, int count, int first,
                                    int num_classes, int32_t* __restrict__ labels, float* __restrict__ label_prob,
                                    uint32_t* __restrict__ hist, unsigned long long* __restrict__ mass_q32) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  int32_t id[AI4E_TOPK];
  float pr[AI4E_TOPK];
#pragma unroll
  for (int k = 0; k < AI4E_TOPK; ++k) {
    id[k] = idx5[i * AI4E_TOPK + k];
    pr[k] = prob5[i * AI4E_TOPK + k];
  }
  labels[first + i] = id[0];
  label_prob[first + i] = pr[0];
  if ((unsigned)id[0] < (unsigned)num_classes) atomicAdd(&hist[id[0]], 1u);
#pragma unroll
  for (int k = 0; k < AI4E_TOPK; ++k) {
    if ((unsigned)id[k] >= (unsigned)num_classes) continue;
    const float p = fminf(fmaxf(pr[k], 0.0f), 1.0f);  // a NaN probability adds nothing
    atomicAdd(&mass_q32[id[k]], __float2ull_rn(p * 4294967296.0f));
  }
}

}  // namespace

int tile_summary_launch(const int32_t* d_idx5, const float* d_prob5, int count, int first, int num_classes,
                        int32_t* d_labels, float* d_label_prob, uint32_t* d_hist, unsigned long long* d_mass_q32,
                        cudaStream_t stream) {
  if (count <= 0) return AI4E_OK;
  tile_summary_kernel<<<(count + 255) / 256, 256, 0, stream>>>(d_idx5, d_prob5, count, first, num_classes, d_labels,
                                                                d_label_prob, d_hist, d_mass_q32);
  AI4E_CUDA_TRY(cudaGetLastError());
  return AI4E_OK;
}

}  // namespace ai4e

extern "C" int ai4e_tile_summary(const int32_t* d_top_idx5, const float* d_top_prob5, int count, int first, int num_classes,
                                 int32_t* d_labels, float* d_label_prob, uint32_t* d_hist, uint64_t* d_mass_q32,
                                 void* stream) {
  AI4E_REQUIRE(d_top_idx5 && d_top_prob5 && d_labels && d_label_prob && d_hist && d_mass_q32, "tile_summary: null pointer");
  AI4E_REQUIRE(count >= 0 && first >= 0 && num_classes >= 1, "tile_summary: bad sizes");
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "64-bit atomics");
  return ai4e::tile_summary_launch(d_top_idx5, d_top_prob5, count, first, num_classes, d_labels, d_label_prob, d_hist,
                             (unsigned long long*)d_mass_q32, (cudaStream_t)stream);
}
