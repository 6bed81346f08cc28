// Micro-probe of the MMA warp's per-K-block loop (companion of mma_probe.cu): what one iteration of
//   [wait for a full barrier] [tcgen05 fence] [elect] [4 x tcgen05.mma, M = 128, K = 16] [commit] [__syncwarp]
// costs against the 4 x N / 2 cycles its MMAs need, piece by piece, alone on the SM and next to ALU-busy warps on the
// same scheduler (what the epilogue warps are to the MMA warp in conv_igemm.cu / conv_fused.cu).
//   nvcc -gencode arch=compute_100a,code=sm_100a tools/mma_loop_probe.cu -o mma_loop_probe
#include "../aiforearth-api-platform_b200/csrc/common.cuh"
#include <cstdio>
using namespace ai4e;

// variant bits: 1 = wait on an (already complete) mbarrier per unit, 2 = tcgen05.fence::after_thread_sync per unit,
// 4 = commit per unit (else one at the end), 8 = two ALU-busy warps share the MMA warp's scheduler,
// 16 = the wait is taken by the elected thread between its third and fourth MMA (for the next unit)
template <int N, int kVariant>
__global__ void __launch_bounds__(384, 1) probe(int units, long long* out, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar_done, bar_ready, bar_unit;
  __shared__ uint32_t tmem_ptr;
  __shared__ int stop;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar_done), 1); mbar_init(smem_u32(&bar_ready), 1); mbar_init(smem_u32(&bar_unit), 1);
    fence_mbar_init();
    stop = 0;
  }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_ptr), 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) mbar_arrive(smem_u32(&bar_ready));  // phase 0 of bar_ready is complete from here on
  __syncthreads();
  if (warp == 1) {  // like the kernels: the MMA warp is warp 1 (scheduler 1)
    const uint32_t a_lo = umma_desc_lo(base), b_lo = umma_desc_lo(base + 16384);
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, N);
    const long long t0 = clock64();
    for (int u = 0; u < units; ++u) {
      if ((kVariant & 1) && !(kVariant & 16)) mbar_wait(smem_u32(&bar_ready), 0);
      if (kVariant & 2) tc_fence_after();
      const uint32_t d = tm + (u & 1) * N;
      if (elect_one()) {
        if (kVariant & 16) {  // the elected thread takes the NEXT unit's wait between its third and fourth MMA
#pragma unroll
          for (int k = 0; k < 3; ++k) umma_bf16_ss(d, umma_desc_join(a_lo + 2u * k), umma_desc_join(b_lo + 2u * k), idesc, k ? 1u : 0u);
          mbar_wait(smem_u32(&bar_ready), 0);
          umma_bf16_ss(d, umma_desc_join(a_lo + 6u), umma_desc_join(b_lo + 6u), idesc, 1u);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(d, umma_desc_join(a_lo + 2u * k), umma_desc_join(b_lo + 2u * k), idesc, k ? 1u : 0u);
        }
        if (kVariant & 4) umma_commit(smem_u32(&bar_unit));
      }
      __syncwarp();
    }
    const long long t1 = clock64();
    if (elect_one()) umma_commit(smem_u32(&bar_done));
    __syncwarp();
    mbar_wait(smem_u32(&bar_done), 0);
    const long long t2 = clock64();
    if (lane == 0) { stop = 1; if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; } }
  } else if ((kVariant & 8) && (warp == 5 || warp == 9)) {  // same scheduler as warp 1: dependent FMAs and shared loads
    float a = (float)lane, b = 1.0001f;
    volatile int* vs = &stop;
    while (!*vs) {
#pragma unroll
      for (int i = 0; i < 64; ++i) a = fmaf(a, b, 0.5f);
    }
    if (a == 123.456f) sink[0] = a;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

template <int N, int V>
void run(const char* what) {
  long long* d; cudaMalloc(&d, 16);
  float* sink; cudaMalloc(&sink, 4);
  cudaFuncSetAttribute(probe<N, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
  const int units = 256;
  for (int rep = 0; rep < 2; ++rep) probe<N, V><<<148, 384, 60000>>>(units, d, sink);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("N=%3d %-58s loop %6.1f cyc/unit, to completion %6.1f cyc/unit (MMAs alone %d)  %s\n", N, what, (double)h[0] / units,
         (double)h[1] / units, 4 * N / 2, cudaGetErrorString(e));
  cudaFree(d); cudaFree(sink);
}
template <int N>
void all() {
  run<N, 0>("elect + 4 MMA + syncwarp");
  run<N, 4>("... + commit per unit");
  run<N, 5>("... + commit + barrier wait (complete)");
  run<N, 7>("... + commit + wait + tcgen05 fence  (= the kernels' loop)");
  run<N, 15>("the kernels' loop next to two ALU-busy warps");
  run<N, 23>("next unit's wait by the elected thread before its last MMA");
  run<N, 31>("... next to two ALU-busy warps");
}
int main() {
  all<64>(); all<128>(); all<256>();
  return 0;
}
