// Micro-probe: cost of the synchronisation primitives in the MMA warp's loop.
#include "../aiforearth-api-platform_b200/csrc/common.cuh"
#include <cstdio>
using namespace ai4e;
__global__ void probe(long long* out) {
  __shared__ uint64_t bar;
  __shared__ volatile uint32_t word;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); word = 0; }
  __syncthreads();
  if (threadIdx.x == 0) mbar_arrive(smem_u32(&bar));  // phase 0 completes
  __syncthreads();
  if (threadIdx.x < 32) {
    const int R = 256;
    long long t[8];
    t[0] = clock64();
    for (int i = 0; i < R; ++i) { if (!mbar_try_wait(smem_u32(&bar), 0)) word = 1; }
    t[1] = clock64();
    for (int i = 0; i < R; ++i) { if (!mbar_test_wait(smem_u32(&bar), 0)) word = 1; }
    t[2] = clock64();
    for (int i = 0; i < R; ++i) tc_fence_after();
    t[3] = clock64();
    for (int i = 0; i < R; ++i) tc_fence_before();
    t[4] = clock64();
    uint32_t acc = 0;
    for (int i = 0; i < R; ++i) acc += elect_one();
    t[5] = clock64();
    for (int i = 0; i < R; ++i) __syncwarp();
    t[6] = clock64();
    for (int i = 0; i < R; ++i) acc += word;
    t[7] = clock64();
    if (threadIdx.x == 0) { for (int i = 0; i < 7; ++i) out[i] = (t[i + 1] - t[i]) / R; out[7] = acc; }
  }
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  probe<<<1, 64>>>(d); cudaDeviceSynchronize();
  probe<<<1, 64>>>(d); printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  long long h[8]; cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
  const char* names[7] = {"try_wait (complete)", "test_wait (complete)", "tcgen05.fence::after", "tcgen05.fence::before", "elect.sync", "__syncwarp", "volatile LDS"};
  for (int i = 0; i < 7; ++i) printf("%-24s %lld cycles\n", names[i], h[i]);
  return 0;
}
