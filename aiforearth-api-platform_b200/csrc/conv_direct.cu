// Direct convolution on CUDA cores, NHWC bf16 in/out, KRSC bf16 weights, fp32 bias and
// accumulation, fused bias + residual + ReLU.
//
// Role: (1) the stem (7x7 s2, Cin = 3/4) until the tensor-core stem replaces it — Cin < 64
// cannot feed the 64-wide K blocks of the implicit-GEMM kernel; (2) an on-device
// cross-check for the tcgen05 path in the GPU tests (`algo = AI4E_CONV_DIRECT`).  It is not
// the product path for the 52 remaining convolutions.
#include "common.cuh"

namespace ai4e {

struct DirectParams {
  const bf16* in; const bf16* w; const float* bias; const bf16* res; bf16* out;
  int N, H, W, Cin, Cout, R, S, stride, pad, P, Q, relu;
};

// One thread = one output pixel x 8 consecutive output channels.  Weights for the CTA's
// channel group live in shared memory as fp32 [R*S*Cin][8].
template <int CO_PER_THREAD>
__global__ void __launch_bounds__(128)
conv_direct_kernel(const DirectParams p) {
  extern __shared__ float w_s[];  // [K][CO_PER_THREAD]
  const int K = p.R * p.S * p.Cin;
  const int co0 = blockIdx.y * CO_PER_THREAD;
  for (int i = threadIdx.x; i < K * CO_PER_THREAD; i += blockDim.x) {
    const int k = i / CO_PER_THREAD, j = i % CO_PER_THREAD;
    w_s[i] = __bfloat162float(p.w[(long long)(co0 + j) * K + k]);
  }
  __syncthreads();
  const long long M = (long long)p.N * p.P * p.Q;
  const long long m = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (m >= M) return;
  const int q = (int)(m % p.Q);
  const int pp = (int)((m / p.Q) % p.P);
  const int n = (int)(m / ((long long)p.P * p.Q));
  float acc[CO_PER_THREAD];
#pragma unroll
  for (int j = 0; j < CO_PER_THREAD; ++j) acc[j] = 0.f;
  for (int r = 0; r < p.R; ++r) {
    const int h = pp * p.stride - p.pad + r;
    if (h < 0 || h >= p.H) continue;
    for (int s = 0; s < p.S; ++s) {
      const int w = q * p.stride - p.pad + s;
      if (w < 0 || w >= p.W) continue;
      const bf16* ip = p.in + (((long long)n * p.H + h) * p.W + w) * p.Cin;
      const float* wp = w_s + (r * p.S + s) * p.Cin * CO_PER_THREAD;
      for (int c = 0; c < p.Cin; ++c) {
        const float x = __bfloat162float(ip[c]);
#pragma unroll
        for (int j = 0; j < CO_PER_THREAD; ++j) acc[j] = fmaf(x, wp[c * CO_PER_THREAD + j], acc[j]);
      }
    }
  }
  bf16* op = p.out + m * p.Cout + co0;
  const bf16* rp = p.res ? p.res + m * p.Cout + co0 : nullptr;
#pragma unroll
  for (int j = 0; j < CO_PER_THREAD; ++j) {
    float v = acc[j] + p.bias[co0 + j];
    if (rp) v += __bfloat162float(rp[j]);
    if (p.relu) v = fmaxf(v, 0.f);
    op[j] = __float2bfloat16_rn(v);
  }
}

int launch_conv_direct(const void* in, int N, int H, int W, int Cin, const void* weight,
                       const float* bias, int Cout, int R, int S, int stride, int pad,
                       const void* residual, int relu, void* out, cudaStream_t stream) {
  constexpr int CO = 8;
  AI4E_REQUIRE(Cout % CO == 0, "conv(direct): Cout %d not a multiple of %d", Cout, CO);
  DirectParams p{};
  p.in = (const bf16*)in; p.w = (const bf16*)weight; p.bias = bias; p.res = (const bf16*)residual;
  p.out = (bf16*)out; p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout; p.R = R; p.S = S;
  p.stride = stride; p.pad = pad; p.relu = relu;
  p.P = (H + 2 * pad - R) / stride + 1;
  p.Q = (W + 2 * pad - S) / stride + 1;
  const long long M = (long long)N * p.P * p.Q;
  const size_t smem = (size_t)R * S * Cin * CO * sizeof(float);
  AI4E_REQUIRE(smem <= 200 * 1024, "conv(direct): filter too large for shared memory (%zu B)", smem);
  AI4E_CUDA_TRY(cudaFuncSetAttribute(conv_direct_kernel<CO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)((M + 127) / 128), (unsigned)(Cout / CO));
  conv_direct_kernel<CO><<<grid, 128, smem, stream>>>(p);
  AI4E_CUDA_TRY(cudaGetLastError());
  return AI4E_OK;
}

}  // namespace ai4e
