// Micro-probe: cost of back-to-back tcgen05.mma (M=128, K=16, bf16) for several N and for
// dependent vs independent accumulators.  nvcc -arch sm_100a tools/mma_probe.cu -o mma_probe
#include "../aiforearth-api-platform_b200/csrc/common.cuh"
#include <cstdio>
using namespace ai4e;

template <int N>
__global__ void probe(int reps, int n_acc, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_ptr), 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = tmem_ptr;
  if (threadIdx.x < 32) {
    const uint64_t adesc = umma_smem_desc_kmajor(base, 128, 2);
    const uint64_t bdesc = umma_smem_desc_kmajor(base + 16384, 128, 2);
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, N);
    long long t0 = clock64();
    if (elect_one()) {
      for (int r = 0; r < reps; ++r) {
        const uint32_t d = tm + (r % n_acc) * N;
        umma_bf16_ss(d, adesc + 2u * (r & 3), bdesc + 2u * (r & 3), idesc, r >= n_acc ? 1u : 0u);
      }
      umma_commit(smem_u32(&bar));
    }
    __syncwarp();
    long long t1 = clock64();
    mbar_wait(smem_u32(&bar), 0);
    long long t2 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

template <int N>
void run(int grid) {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(probe<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
  for (int n_acc : {1, 2, 4}) {
    if (n_acc * N > 512) continue;
    const int reps = 512;
    probe<N><<<grid, 128, 60000>>>(reps, n_acc, d);
    cudaDeviceSynchronize();
    probe<N><<<grid, 128, 60000>>>(reps, n_acc, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("grid %3d N=%3d accumulators=%d: issue %.1f cyc/MMA, complete %.1f cyc/MMA (ideal %d)  %s\n", grid, N, n_acc,
           (double)h[0] / reps, (double)h[1] / reps, N / 2, cudaGetErrorString(e));
  }
  cudaFree(d);
}
int main() {
  for (int grid : {1, 148}) { run<64>(grid); run<128>(grid); run<256>(grid); }
  return 0;
}
