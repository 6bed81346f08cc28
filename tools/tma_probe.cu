// L2 -> shared-memory ingest by TMA mode (companion of l2_probe.cu, which streams 1-D bulk copies): every CTA
// streams 16 KB tiles from an L2-resident region into a six-slot shared-memory ring, one issuing thread, and
// the probe prints bytes / clk / SM and the aggregate for
//   bulk     cp.async.bulk, 16 KB contiguous
//   tiled    cp.async.bulk.tensor.2d, box 64 x 128 bf16 (128 rows of 128 bytes), SWIZZLE_128B, row pitch
//            128 B (a C = 64 activation: the rows are contiguous), 512 B (C = 256) and 2304 B (a w2 weight matrix)
//   im2col   cp.async.bulk.tensor.4d...im2col over NHWC [N, 28, 28, 128], 3 x 3 / pad 1, 64 channels x 128 pixels
// i.e. the operand loads of conv_igemm.cu / conv_fused.cu.  What the convolution K loops can ingest is bounded by
// the tiled / im2col rows, not by the bulk row.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tma_probe.cu -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int kStages = 6, kBytes = 16384;
enum Mode { kBulk = 0, kTiled = 1, kIm2col = 2 };

struct Args {
  const uint8_t* src;        // bulk
  unsigned long long region;
  int rows_per_cta;          // tiled: this CTA walks rows [cta * rows_per_cta, +rows_per_cta) in steps of 128
  int kcols;                 // tiled: columns (elements) of the matrix; the walk also moves along K in steps of 64
  int images_per_cta;        // im2col
  int iters;
};

template <int kMode>
__global__ void __launch_bounds__(128, 1) probe(const __grid_constant__ CUtensorMap tmap, const Args a,
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
    const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
    int row = 0, kc = 0, img = 0, pix = 0, tap = 0;
    unsigned long long off = 0;
    const unsigned long long cta_bytes = a.region / gridDim.x & ~(unsigned long long)(kBytes - 1);
    const long long t0 = clock64();
    for (int i = 0; i < a.iters + kStages; ++i) {
      const int s = i % kStages;
      if (i >= kStages) {
        const uint32_t parity = ((i / kStages) - 1) & 1;
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p; }"
                       : "=r"(ok) : "r"(smem_u32(&bars[s])), "r"(parity) : "memory");
      }
      if (i < a.iters) {
        const uint32_t bar = smem_u32(&bars[s]), dst = base + s * kBytes;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kBytes) : "memory");
        if (kMode == kBulk) {
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(dst), "l"(a.src + blockIdx.x * cta_bytes + off), "r"(kBytes), "r"(bar) : "memory");
          off += kBytes;
          if (off + kBytes > cta_bytes) off = 0;
        } else if (kMode == kTiled) {
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(dst), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(bar), "r"(kc), "r"((int)blockIdx.x * a.rows_per_cta + row)
                       : "memory");
          kc += 64;  // K-major walk like a GEMM K loop, then the next 128 rows
          if (kc >= a.kcols) { kc = 0; row += 128; if (row + 128 > a.rows_per_cta) row = 0; }
        } else {
          const int n = (int)blockIdx.x * a.images_per_cta + img;
          const int p0 = pix / 28, q0 = pix - p0 * 28;
          const uint16_t s_ = (uint16_t)(tap % 3), r_ = (uint16_t)(tap / 3);
          asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
                       " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
                       ::"r"(dst), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(bar), "r"((tap & 1) * 64), "r"(q0 - 1), "r"(p0 - 1), "r"(n),
                       "h"(s_), "h"(r_) : "memory");
          if (++tap == 9) { tap = 0; pix += 128; if (pix >= 784) { pix -= 784; if (++img >= a.images_per_cta) img = 0; } }
        }
      }
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int kMode>
static void run(const char* what, const CUtensorMap& tm, Args a, int clk_khz, unsigned long long* out) {
  cudaFuncSetAttribute(probe<kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStages * kBytes + 1024);
  for (int grid : {1, 37, 148}) {
    for (int rep = 0; rep < 2; ++rep) probe<kMode><<<grid, 128, kStages * kBytes + 1024>>>(tm, a, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: error %s\n", what, cudaGetErrorString(e)); exit(1); }
    unsigned long long h[148];
    cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
    double mx = 0, sum = 0;
    for (int i = 0; i < grid; ++i) { mx = h[i] > mx ? h[i] : mx; sum += h[i]; }
    const double bpc = (double)a.iters * kBytes / (sum / grid);
    printf("%-44s grid %3d: %6.1f B/clk/SM (slowest CTA %6.1f), aggregate at %.2f GHz %6.2f TB/s\n", what, grid, bpc,
           (double)a.iters * kBytes / mx, clk_khz / 1e6, bpc * grid * clk_khz * 1e3 / 1e12);
  }
}

int main() {
  const size_t region = 64ull << 20;  // L2-resident (126 MB L2)
  uint8_t* d;
  cudaMalloc(&d, region);
  cudaMemset(d, 1, region);
  unsigned long long* out;
  cudaMalloc(&out, 148 * 8);
  int clk_khz;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  EncodeTiled enc_tiled = (EncodeTiled)fn;
  cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q);
  EncodeIm2col enc_im2col = (EncodeIm2col)fn;
  if (!enc_tiled || !enc_im2col) { printf("no tensor-map encoders\n"); return 1; }
  Args a{};
  a.src = d; a.region = region; a.iters = 4000;
  CUtensorMap none{};
  run<kBulk>("bulk 1-D, 16 KB contiguous", none, a, clk_khz, out);
  for (int kcols : {64, 256, 1152}) {
    const size_t pitch = (size_t)kcols * 2;
    const size_t rows = region / pitch;
    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t)kcols, (cuuint64_t)rows};
    const cuuint64_t str[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {64, 128}, es[2] = {1, 1};
    CUresult r = enc_tiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode tiled failed %d\n", (int)r); return 1; }
    a.kcols = kcols;
    a.rows_per_cta = (int)((rows / 148) & ~(size_t)127);
    char what[96];
    snprintf(what, sizeof what, "tiled 64 x 128 bf16, 128B swizzle, pitch %zu B", pitch);
    run<kTiled>(what, tm, a, clk_khz, out);
  }
  {
    const int C = 128, H = 28, W = 28;
    const int N = (int)(region / ((size_t)C * H * W * 2));
    CUtensorMap tm;
    const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    const cuuint64_t str[3] = {(cuuint64_t)C * 2, (cuuint64_t)C * W * 2, (cuuint64_t)C * W * H * 2};
    const int lower[2] = {-1, -1}, upper[2] = {-1, -1};  // 3 x 3, pad 1, stride 1: corners -pad and pad - (k - 1)
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc_im2col(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, d, dims, str, lower, upper, 64, 128, es,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode im2col failed %d\n", (int)r); return 1; }
    a.images_per_cta = N / 148;
    run<kIm2col>("im2col 3x3 over [N,28,28,128], 64 ch x 128 px", tm, a, clk_khz, out);
  }
  return 0;
}
