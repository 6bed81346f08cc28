// L2 -> shared-memory ingest probe: every CTA streams 32 KB bulk copies (cp.async.bulk, the 1-D TMA
// path) from an L2-resident region into a ring of shared-memory buffers.  Prints bytes/clk/SM and
// aggregate TB/s for several grid sizes.   nvcc -arch=sm_100a -o l2_probe l2_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int kStages, int kBytes>
__global__ void __launch_bounds__(128, 1) probe(const uint8_t* src, size_t region, size_t cta_stride, int iters,
                                                unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[kStages];
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t* base = src + (blockIdx.x * cta_stride) % region;
    size_t off = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters + kStages; ++i) {
      const int s = i % kStages;
      if (i >= kStages) {  // wait for the copy issued kStages ago into this slot
        const uint32_t parity = ((i / kStages) - 1) & 1;
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p; }"
                       : "=r"(ok) : "r"(smem_u32(&bars[s])), "r"(parity) : "memory");
      }
      if (i < iters) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[s])), "r"(kBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + s * kBytes)), "l"(base + off), "r"(kBytes), "r"(smem_u32(&bars[s])) : "memory");
        off += kBytes;
        if (off + kBytes > cta_stride && cta_stride >= (size_t)kBytes) off = 0;
        if ((blockIdx.x * cta_stride) % region + off + kBytes > region) off = 0;
      }
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const size_t region = 64ull << 20;  // 64 MB: L2-resident (126 MB L2)
  uint8_t* d; cudaMalloc(&d, region); cudaMemset(d, 1, region);
  unsigned long long* out; cudaMalloc(&out, 148 * 8);
  constexpr int kStages = 6, kBytes = 32768;
  cudaFuncSetAttribute(probe<kStages, kBytes>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStages * kBytes);
  int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const int iters = 4000;
  for (int shared_addr = 0; shared_addr < 2; ++shared_addr) {
    for (int grid : {1, 8, 37, 74, 148}) {
      const size_t stride = shared_addr ? 0 : (region / 148) & ~(size_t)32767;
      for (int rep = 0; rep < 2; ++rep) {
        probe<kStages, kBytes><<<grid, 128, kStages * kBytes>>>(d, region, shared_addr ? (size_t)kBytes * 8 : stride, iters, out);
        if (shared_addr) {}
      }
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      unsigned long long h[148]; cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
      double mx = 0, sum = 0;
      for (int i = 0; i < grid; ++i) { mx = h[i] > mx ? h[i] : mx; sum += h[i]; }
      const double bpc = (double)iters * kBytes / (sum / grid);
      printf("%s grid %3d: %.1f B/clk/SM (avg), slowest CTA %.1f B/clk; aggregate at %.2f GHz = %.2f TB/s\n",
             shared_addr ? "same 256 KB for all CTAs" : "distinct 442 KB per CTA   ", grid, bpc, (double)iters * kBytes / mx,
             clk_khz / 1e6, bpc * grid * clk_khz * 1e3 / 1e12);
    }
  }
  return 0;
}
