// K3 max-pool and K5 classifier head (global average pool -> FC -> softmax -> top-k),
// NHWC bf16 activations.  All HBM/latency-bound: 16-byte vector access, warp-shuffle
// reductions, fp32 arithmetic.  Mirrors torchvision's `maxpool`, `avgpool`, `fc` and the
// `softmax(...).topk(5)` a classification callback applies (reference call site
// APIs/1.0/base-py/ai4e_service.py:203; oracle/resnet_oracle.py restates it).
#include "common.cuh"

namespace ai4e {

int sm_count();

// ---------------------------------------------------------------------------------------
// max-pool k x k / stride / pad (padding contributes -inf, as torch)
// one thread = one output pixel x 8 channels (16 bytes)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
maxpool_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int N, int H, int W, int C8,
               int k, int stride, int pad, int P, int Q) {
  const long long total = (long long)N * P * Q * C8;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % C8);
    long long t = idx / C8;
    const int q = (int)(t % Q); t /= Q;
    const int pp = (int)(t % P);
    const int n = (int)(t / P);
    float m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
    for (int r = 0; r < k; ++r) {
      const int h = pp * stride - pad + r;
      if (h < 0 || h >= H) continue;
      for (int s = 0; s < k; ++s) {
        const int w = q * stride - pad + s;
        if (w < 0 || w >= W) continue;
        const uint4 v = __ldg(in + (((long long)n * H + h) * W + w) * C8 + c8);
        const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          m[2 * j] = fmaxf(m[2 * j], bf16_lo(u[j]));
          m[2 * j + 1] = fmaxf(m[2 * j + 1], bf16_hi(u[j]));
        }
      }
    }
    out[idx] = make_uint4(pack_bf16x2(m[0], m[1]), pack_bf16x2(m[2], m[3]),
                          pack_bf16x2(m[4], m[5]), pack_bf16x2(m[6], m[7]));
  }
}

int launch_maxpool(const void* in, int N, int H, int W, int C, int k, int stride, int pad, void* out,
                   cudaStream_t stream) {
  AI4E_REQUIRE(in && out, "maxpool: null pointer");
  AI4E_REQUIRE(C % 8 == 0, "maxpool: C %d not a multiple of 8", C);
  AI4E_REQUIRE(N > 0 && H > 0 && W > 0 && k > 0 && stride > 0 && pad >= 0 && 2 * pad <= k,
               "maxpool: bad geometry");
  const int P = (H + 2 * pad - k) / stride + 1, Q = (W + 2 * pad - k) / stride + 1;
  const long long total = (long long)N * P * Q * (C / 8);
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)sm_count() * 16);
  maxpool_kernel<<<grid, 256, 0, stream>>>((const uint4*)in, (uint4*)out, N, H, W, C / 8, k, stride,
                                           pad, P, Q);
  AI4E_CUDA_TRY(cudaGetLastError());
  return AI4E_OK;
}

// ---------------------------------------------------------------------------------------
// global average pool: [N, HW, C] bf16 -> [N, C] fp32.  thread = (n, 8 channels)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
avgpool_kernel(const uint4* __restrict__ in, float* __restrict__ out, int N, int HW, int C8) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * C8) return;
  const int n = idx / C8, c8 = idx % C8;
  float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const uint4* p = in + (long long)n * HW * C8 + c8;
  for (int i = 0; i < HW; ++i) {
    const uint4 v = __ldg(p + (long long)i * C8);
    const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { s[2 * j] += bf16_lo(u[j]); s[2 * j + 1] += bf16_hi(u[j]); }
  }
  const float inv = 1.0f / (float)HW;
  float4* o = reinterpret_cast<float4*>(out + (long long)n * C8 * 8 + c8 * 8);
  o[0] = make_float4(s[0] * inv, s[1] * inv, s[2] * inv, s[3] * inv);
  o[1] = make_float4(s[4] * inv, s[5] * inv, s[6] * inv, s[7] * inv);
}

// ---------------------------------------------------------------------------------------
// FC: logits[n][k] = bias[k] + sum_c pooled[n][c] * W[k][c]   (W bf16 [classes][C])
// CTA = 8 warps = 8 classes x up to 8 images whose pooled vectors sit in shared memory.
// ---------------------------------------------------------------------------------------
constexpr int kFcImgs = 8;
__global__ void __launch_bounds__(256)
fc_kernel(const float* __restrict__ pooled, const uint4* __restrict__ w, const float* __restrict__ bias,
          float* __restrict__ logits, int N, int C, int classes) {
  extern __shared__ float x_s[];  // [kFcImgs][C]
  const int n0 = blockIdx.y * kFcImgs;
  const int nimg = min(kFcImgs, N - n0);
  for (int i = threadIdx.x; i < nimg * C; i += blockDim.x) x_s[i] = pooled[(long long)n0 * C + i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cls = blockIdx.x * 8 + warp;
  if (cls >= classes) return;
  float acc[kFcImgs];
#pragma unroll
  for (int i = 0; i < kFcImgs; ++i) acc[i] = 0.f;
  const uint4* wr = w + (long long)cls * (C / 8);
  for (int c8 = lane; c8 < C / 8; c8 += 32) {
    const uint4 v = __ldg(wr + c8);
    const uint32_t u[4] = {v.x, v.y, v.z, v.w};
    float wf[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) { wf[2 * j] = bf16_lo(u[j]); wf[2 * j + 1] = bf16_hi(u[j]); }
#pragma unroll
    for (int i = 0; i < kFcImgs; ++i) {
      if (i < nimg) {
        const float* xp = x_s + i * C + c8 * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i] = fmaf(wf[j], xp[j], acc[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kFcImgs; ++i) {
    float v = acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0 && i < nimg) logits[(long long)(n0 + i) * classes + cls] = v + bias[cls];
  }
}

// ---------------------------------------------------------------------------------------
// softmax + top-k (k <= 8): one warp per image.  Ties resolve to the lower class index,
// matching a stable descending sort.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
softmax_topk_kernel(const float* __restrict__ logits, int N, int classes, int topk,
                    int32_t* __restrict__ top_idx, float* __restrict__ top_prob) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= N) return;
  const float* z = logits + (long long)warp * classes;
  float mx = -INFINITY;
  for (int i = lane; i < classes; i += 32) mx = fmaxf(mx, z[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
  for (int i = lane; i < classes; i += 32) sum += expf(z[i] - mx);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  float prev_v = INFINITY;
  int prev_i = -1;
  for (int t = 0; t < topk; ++t) {
    // best (value, index) strictly after (prev_v, prev_i) in (value desc, index asc) order
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = lane; i < classes; i += 32) {
      const float v = z[i];
      const bool after = (v < prev_v) || (v == prev_v && i > prev_i);
      if (after && (v > bv || (v == bv && i < bi))) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) {
      top_idx[(long long)warp * topk + t] = bi;
      top_prob[(long long)warp * topk + t] = expf(bv - mx) / sum;
    }
    prev_v = bv; prev_i = bi;
  }
}

int launch_head(const void* in, int N, int HW, int C, const void* fc_w, const float* fc_b,
                int classes, float* pooled_scratch, float* logits, int32_t* top_idx, float* top_prob,
                int topk, cudaStream_t stream) {
  AI4E_REQUIRE(in && fc_w && fc_b && pooled_scratch && logits, "head: null pointer");
  AI4E_REQUIRE(N > 0 && HW > 0 && C % 8 == 0 && classes > 0, "head: bad geometry");
  AI4E_REQUIRE(topk >= 0 && topk <= 8 && topk <= classes, "head: topk %d unsupported", topk);
  avgpool_kernel<<<(N * (C / 8) + 255) / 256, 256, 0, stream>>>((const uint4*)in, pooled_scratch, N,
                                                                HW, C / 8);
  AI4E_CUDA_TRY(cudaGetLastError());
  const size_t smem = (size_t)kFcImgs * C * sizeof(float);
  AI4E_REQUIRE(smem <= 96 * 1024, "head: C %d too large", C);
  AI4E_CUDA_TRY(cudaFuncSetAttribute(fc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((classes + 7) / 8, (N + kFcImgs - 1) / kFcImgs);
  fc_kernel<<<grid, 256, smem, stream>>>(pooled_scratch, (const uint4*)fc_w, fc_b, logits, N, C, classes);
  AI4E_CUDA_TRY(cudaGetLastError());
  if (topk > 0) {
    AI4E_REQUIRE(top_idx && top_prob, "head: null top-k outputs");
    softmax_topk_kernel<<<(N * 32 + 127) / 128, 128, 0, stream>>>(logits, N, classes, topk, top_idx, top_prob);
    AI4E_CUDA_TRY(cudaGetLastError());
  }
  return AI4E_OK;
}

}  // namespace ai4e
