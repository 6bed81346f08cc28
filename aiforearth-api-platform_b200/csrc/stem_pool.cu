// K2+K3 in one kernel: the 7x7/2 stem convolution (+ folded BatchNorm, bias, ReLU) and the 3x3/2
// max-pool behind it.  The stem's output (103 MB of bf16 at micro-batch 64, the largest tensor of the
// network) is then never written to HBM and never read back: the kernel reads the preprocessed image
// (25.7 MB) and writes the pooled tensor (25.7 MB).
//
// It builds on the row mode of the tensor-core stem (conv_igemm.cu, kModeStemRows): input rows sit in
// plain shared-memory row buffers and the A operand of a filter row is that buffer seen through an
// un-swizzled K-major descriptor whose core matrices overlap.  New here:
//   * one tile = one POOLED output row j: stem rows 2j and 2j+1 come out of ONE accumulator of 128
//     columns.  Input row i of the tile (image row 4j-3+i, i = 0..8) meets filter row i of stem row 2j
//     and filter row i-2 of stem row 2j+1, so the weights are repacked (host, once) into nine
//     [128][32] blocks [W_i ; W_(i-2)] (zero where a filter row does not exist) and every MMA has
//     N = 128 - which costs the tensor core the same 67 cycles as the N = 64 of the plain stem:
//     18 MMAs per two stem rows instead of 28.
//   * a CTA walks RUNS of consecutive pooled rows of one image, so that stem row 2j-1, the third row
//     of the pooling window, is simply what the previous tile left in registers; a run that does not
//     start at the top of an image first computes the tile above it and stores nothing.
//   * epilogue: eight warps, (TMEM lane quadrant) x (channel half).  A thread owns one stem column:
//     bias + ReLU + rounding to bf16 as the plain stem does, vertical max over the three rows in
//     packed bf16 (max commutes with the monotonic bias/ReLU/rounding, so the result is bit-identical
//     to pooling the stored stem output), horizontal max through warp shuffles - the one column that
//     lies in the neighbouring quadrant crosses through 64 bytes of shared memory and a 128-thread
//     named barrier - and 64-byte global stores of the pooled pixels from the even lanes.
//
// Stands in for conv1 + bn1 + relu + maxpool of torchvision's ResNet, i.e. the first four modules of
// the model the callback at /root/reference/APIs/1.0/base-py/ai4e_service.py:203 runs.  Oracle:
// oracle/resnet_oracle.py forward_bf16_emul stages "stem" -> "pool".
#include "../../include/ai4e_b200.h"
#include "common.cuh"
#include "kernels.h"
#include "igemm_tiles.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

namespace ai4e {

namespace {

constexpr int kSpRows = 9;                    // input rows per tile
constexpr int kSpTaps = 4;                    // super-pixel taps per filter row (see conv_igemm.cu)
constexpr int kSpN = 128;                     // accumulator columns: stem row 2j | stem row 2j+1
constexpr int kSpMaxGroups = 18;              // 128-byte groups per row buffer (output rows up to 128 wide)
constexpr int kSpStage = kSpRows * kSpMaxGroups * 128 + 256;   // 20,992 B incl. over-read slack
constexpr int kSpSlab = kSpN * 16;            // one tap of one input row: [128 channels][16 B]
constexpr int kSpResidentB = kSpRows * kSpTaps * kSpSlab;      // 73,728 B of weights, resident
constexpr int kSpParts = 4;                   // epilogue: channel quarters (16 channels per warp)
constexpr int kSpEpiWarps = 4 * kSpParts;     // x four TMEM lane quadrants
constexpr int kSpThreads = (2 + kSpEpiWarps + 2) * 32;  // producer, MMA, 16 epilogue warps, 2 converter warps (u8 mode)
constexpr int kSpXchg = 2 * kSpParts * 4 * 32;  // [tile parity][channel quarter][quadrant][16 channels bf16]
constexpr int kSpMaxStages = 8;
constexpr int kSpBarBytes = (2 * kSpMaxStages + 4 + 2) * 8 + 16;
constexpr int kSpSmemLimit = 227 * 1024;
constexpr int kSpStages = (kSpSmemLimit - kSpResidentB - kSpXchg - 256 - kSpBarBytes) / kSpStage > kSpMaxStages
                              ? kSpMaxStages
                              : (kSpSmemLimit - kSpResidentB - kSpXchg - 256 - kSpBarBytes) / kSpStage;
constexpr int kSpSmem = kSpStages * kSpStage + kSpResidentB + kSpXchg + 256 + kSpBarBytes;
// u8 mode (K1 fused in): the image arrives as uint8 HWC rows, two converter warps turn them into the bf16
// row buffers through K1's look-up table.  Its ring pairs every A stage with a stage of raw rows.
constexpr int kSpU8Stage = 7168;              // 9 rows x 768 B (W <= 256, 3 channels), rounded up
constexpr int kSpLutBytes = 3 * 256 * 2;
constexpr int kSpStagesU8 = (kSpSmemLimit - kSpResidentB - kSpXchg - 256 - kSpLutBytes - 2 * kSpBarBytes) / (kSpStage + kSpU8Stage);
constexpr int kSpSmemU8 = kSpStagesU8 * (kSpStage + kSpU8Stage) + kSpResidentB + kSpXchg + 256 + kSpLutBytes + 2 * kSpBarBytes;
static_assert(kSpStagesU8 >= 3 && kSpStagesU8 <= kSpMaxStages, "u8 ring");
constexpr uint32_t kNegInf2 = 0xFF80FF80u;    // two bf16 -inf: max-pool padding

__device__ __forceinline__ uint32_t hmax2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

}  // namespace

struct StemPoolParams {
  int W;            // input columns (u8 mode)
  float mean[3], stdv[3];  // u8 mode: K1's normalisation
  int H;            // input rows
  int P, Q;         // stem output rows / columns
  int Pp, Qp;       // pooled rows / columns
  int run_len;      // pooled rows per run
  int runs_per_img, num_runs;
  int stem_rb;      // bytes per input-row buffer
  int relu;
  const float* bias;
  const void* weights;  // repacked, slab-major
  bf16* out;        // [N][Pp][Qp][64]
};

template <bool kU8>
__global__ void __launch_bounds__(kSpThreads, 1)
stem_pool_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const StemPoolParams p) {
  constexpr int kStages = kU8 ? kSpStagesU8 : kSpStages;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  if (smem_base & 127u) __trap();
  const uint32_t u8_base = smem_base + kStages * kSpStage;                       // u8 mode: raw rows, one stage per A stage
  const uint32_t bres_base = u8_base + (kU8 ? kStages * kSpU8Stage : 0);
  const uint32_t xchg_base = bres_base + kSpResidentB;
  const uint32_t bias_base = xchg_base + kSpXchg;
  const uint32_t lut_base = bias_base + 256;                                     // u8 mode: [3][256] bf16
  const uint32_t bar_base = lut_base + (kU8 ? kSpLutBytes : 0);
  auto u8_full_bar = [&](int s) { return bar_base + kSpBarBytes + 8u * s; };     // (second barrier block, u8 mode only)
  auto u8_empty_bar = [&](int s) { return bar_base + kSpBarBytes + 8u * (kSpMaxStages + s); };
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kSpMaxStages + s); };
  auto tmem_full_bar = [&](int s) { return bar_base + 8u * (2 * kSpMaxStages + s); };
  auto tmem_empty_bar = [&](int s) { return bar_base + 8u * (2 * kSpMaxStages + 2 + s); };
  const uint32_t b_full_bar = bar_base + 8u * (2 * kSpMaxStages + 4);
  const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * kSpMaxStages + 5);
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_addr - smem_base));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), kU8 ? 2 : 1);  // u8 mode: the two converter warps; else the TMA load
      mbar_init(empty_bar(s), 1);
      if constexpr (kU8) {
        mbar_init(u8_full_bar(s), 1);
        mbar_init(u8_empty_bar(s), 2);
      }
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tmem_full_bar(s), 1);
      mbar_init(tmem_empty_bar(s), kSpEpiWarps);
    }
    mbar_init(b_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_addr, 2 * kSpN);
    tmem_relinquish();
  }
  if (warp == 2 && lane < 16) {
    // bias: 64 floats, read by every epilogue warp
    reinterpret_cast<float4*>(smem_raw + (bias_base - smem_base))[lane] = __ldg(reinterpret_cast<const float4*>(p.bias) + lane);
  }
  if constexpr (kU8) {
    // the table of K1 (three separately rounded IEEE operations per entry: bit-identical to preprocess.cu)
    uint16_t* lut = reinterpret_cast<uint16_t*>(smem_raw + (lut_base - smem_base));
    for (int e = threadIdx.x; e < 768; e += blockDim.x) lut[e] = norm_to_bf16((float)(e & 255), p.mean[e >> 8], p.stdv[e >> 8]);
    // the A stages start as zeros and the converters only ever write pixel units: the padding units of
    // every row buffer (left of the image, right of it, the over-read slack) stay zero for good
    uint4* z = reinterpret_cast<uint4*>(smem_raw);
    for (int e = threadIdx.x; e < kStages * kSpStage / 16; e += blockDim.x) z[e] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();  // the tensor core reads them through the async proxy
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;
  pdl_launch_dependents();

  // Every role walks the same tile sequence: runs blockIdx.x, + gridDim.x, ...; inside a run the
  // pooled rows j0 - 1 (only to seed the window, skipped at the top of an image) .. j0 + run_len - 1.
  auto run_geom = [&](int run, int& n_img, int& j_first, int& j_end, int& j_store) {
    n_img = run / p.runs_per_img;
    const int j0 = (run - n_img * p.runs_per_img) * p.run_len;
    j_store = j0;
    j_first = j0 > 0 ? j0 - 1 : 0;
    j_end = min(j0 + p.run_len, p.Pp);
  };

  if (warp == 0) {
    // ------------------------------- TMA producer -------------------------------
    if (elect_one()) {
      // all 36 weight slabs in ONE bulk copy: the host repacks them slab-major ([slab][channel][16 B], the
      // resident layout itself).  (36 separate tensor loads took the issuing thread ~10 k cycles before the
      // first MMA could start.)
      mbar_expect_tx(b_full_bar, kSpResidentB);
      bulk_load_1d(bres_base, p.weights, kSpResidentB, b_full_bar);
    }
    __syncwarp();
    pdl_wait();  // the image is the preprocessing kernel's output
    int stage = 0;
    uint32_t phase = 0;
    for (int run = blockIdx.x; run < p.num_runs; run += gridDim.x) {
      int n_img, j_first, j_end, j_store;
      run_geom(run, n_img, j_first, j_end, j_store);
      for (int j = j_first; j < j_end; ++j) {
        if constexpr (kU8) {
          // raw uint8 rows: box = (W * 3 / 4 words) x 9 rows x 1 image, rows outside the image arrive as
          // zero BYTES (the converters turn them into zero VALUES, not into lut[0])
          mbar_wait(u8_empty_bar(stage), phase ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(u8_full_bar(stage), (uint32_t)kSpRows * p.W * 3);
            tma_load_3d(u8_base + stage * kSpU8Stage, &tmap_a, u8_full_bar(stage), 0, 4 * j - 3, n_img);
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
          continue;
        }
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t a_dst = smem_base + stage * kSpStage;
        if (elect_one()) {
          // ONE load for the tile's nine rows (box = 64 x groups x 9 x 1; row i lands at i * stem_rb): a
          // thread needs ~200-350 cycles per TMA instruction, and nine of them per tile - not the MMAs -
          // set the pace of the first version (49 us).  Group -1 = super-pixels -8..-1 and rows outside
          // the image are filled with zeros: the convolution's padding.
          mbar_expect_tx(full_bar(stage), (uint32_t)kSpRows * p.stem_rb);
          tma_load_4d(a_dst, &tmap_a, full_bar(stage), 0, -1, 4 * j - 3, n_img);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (kU8 && warp >= 2 + kSpEpiWarps) {
    // ------------------------------- converters (u8 mode): K1 inside the stem ------
    // Two warps turn a stage of raw rows into the stage's bf16 row buffers: a lane takes four pixels
    // (12 bytes, three aligned words), looks every byte up in its channel's table and writes
    // (c0, c1, c2, 0) x 4 = 32 bytes.  Pixel x of row i lives in unit 8 + x / 2 of row buffer i.
    const int cw = warp - (2 + kSpEpiWarps);
    const uint16_t* lut = reinterpret_cast<const uint16_t*>(smem_raw + (lut_base - smem_base));
    const int quads_per_row = p.W >> 2, row_bytes = p.W * 3;
    const int slots = kSpRows * quads_per_row;
    pdl_wait();
    int stage = 0;
    uint32_t phase = 0;
    for (int run = blockIdx.x; run < p.num_runs; run += gridDim.x) {
      int n_img, j_first, j_end, j_store;
      run_geom(run, n_img, j_first, j_end, j_store);
      for (int j = j_first; j < j_end; ++j) {
        mbar_wait(u8_full_bar(stage), phase);
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint8_t* src = smem_raw + (u8_base - smem_base) + stage * kSpU8Stage;
        uint8_t* dst = smem_raw + stage * kSpStage;
        for (int slot = cw * 32 + lane; slot < slots; slot += 64) {
          const int i = slot / quads_per_row, xg = slot - i * quads_per_row;
          const int y = 4 * j - 3 + i;
          uint4 o0 = make_uint4(0u, 0u, 0u, 0u), o1 = o0;
          if (y >= 0 && y < p.H) {
            const uint32_t* w = reinterpret_cast<const uint32_t*>(src + i * row_bytes + xg * 12);
            const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
            auto px = [&](uint32_t r, uint32_t g, uint32_t b, uint32_t& lo, uint32_t& hi) {
              lo = (uint32_t)lut[r] | ((uint32_t)lut[256 + g] << 16);
              hi = (uint32_t)lut[512 + b];
            };
            px(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, o0.x, o0.y);
            px(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u, o0.z, o0.w);
            px((w1 >> 16) & 255u, w1 >> 24, w2 & 255u, o1.x, o1.y);
            px((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24, o1.z, o1.w);
          }
          uint4* d = reinterpret_cast<uint4*>(dst + i * p.stem_rb + 128 + xg * 32);
          d[0] = o0;
          d[1] = o1;
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(full_bar(stage));
          mbar_arrive(u8_empty_bar(stage));
        }
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer ---------------------------------
    constexpr uint32_t idesc = umma_idesc_bf16_f32(kBlockM, kSpN);
    int stage = 0;
    uint32_t phase = 0, it = 0;
    mbar_wait(b_full_bar, 0);
    for (int run = blockIdx.x; run < p.num_runs; run += gridDim.x) {
      int n_img, j_first, j_end, j_store;
      run_geom(run, n_img, j_first, j_end, j_store);
      for (int j = j_first; j < j_end; ++j, ++it) {
        const uint32_t acc = it & 1u;
        mbar_wait(tmem_empty_bar(acc), ((it >> 1) & 1u) ^ 1u);
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kSpN;
        const uint32_t a_addr = smem_base + stage * kSpStage;
        if (elect_one()) {
#pragma unroll
          for (int i = 0; i < kSpRows; ++i) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {  // taps (2h, 2h+1) = one K = 16 step
              const uint64_t adesc = umma_smem_desc(a_addr + i * p.stem_rb + (6 + 2 * h) * 16, 16, 128, 0);
              const uint64_t bdesc = umma_smem_desc(bres_base + (i * kSpTaps + 2 * h) * kSpSlab, kSpSlab, 128, 0);
              umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (i | h) != 0 ? 1u : 0u);
            }
          }
          umma_commit(empty_bar(stage));
          umma_commit(tmem_full_bar(acc));
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp < 2 + kSpEpiWarps) {
    // ------------------------------- epilogue -----------------------------------
    // Sixteen warps: (TMEM lane quadrant) x (channel quarter).  With eight warps of 32 channels the kernel
    // ran at the pace of one warp's dependent instruction chain per tile (~2,150 cycles against 1,150 of
    // MMAs, ncu: profiles/r02_ncu_full.md); 16 channels per warp halve the chain and the registers.
    const int part = (warp - 2) >> 2;   // channel quarter: channels [16 * part, 16 * part + 16)
    const int quad = warp & 3;          // TMEM lane quadrant = stem columns [32 * quad, 32 * quad + 32)
    const float4* const b4 = reinterpret_cast<const float4*>(smem_raw + (bias_base - smem_base)) + part * 4;
    const uint32_t lane_tmem = (uint32_t)(quad * 32) << 16;
    const int q = quad * 32 + lane;     // this thread's stem column
    uint4* const xch = reinterpret_cast<uint4*>(smem_raw + (xchg_base - smem_base));  // [parity][part][quad][2 x uint4]
    const bool relu = p.relu != 0;
    pdl_wait();  // the pooled tensor's buffer may still be read by an earlier kernel
    uint32_t it = 0;      // tiles computed (accumulator stage and phase)
    uint32_t stored = 0;  // tiles stored (exchange-buffer parity: consecutive barriers must alternate buffers)
    for (int run = blockIdx.x; run < p.num_runs; run += gridDim.x) {
      int n_img, j_first, j_end, j_store;
      run_geom(run, n_img, j_first, j_end, j_store);
      uint32_t prev[8];  // stem row 2j - 1 of this thread's column: bf16 pairs, after bias / ReLU / rounding
#pragma unroll
      for (int k = 0; k < 8; ++k) prev[k] = kNegInf2;
      for (int j = j_first; j < j_end; ++j, ++it) {
        const uint32_t acc = it & 1u;
        mbar_wait(tmem_full_bar(acc), (it >> 1) & 1u);
        tc_fence_after();
        uint32_t ra[16], rb[16];
        tmem_ld_32x32b_x16(tmem_base + acc * kSpN + part * 16 + lane_tmem, ra);        // stem row 2j
        tmem_ld_32x32b_x16(tmem_base + acc * kSpN + 64 + part * 16 + lane_tmem, rb);   // stem row 2j + 1
        tmem_wait_ld_pin16(ra);
        tmem_wait_ld_pin16(rb);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tmem_empty_bar(acc));
        uint32_t vm[8];  // max over the window's three rows
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 b = b4[k];
          uint32_t a0 = pack_bf16x2(__uint_as_float(ra[4 * k]) + b.x, __uint_as_float(ra[4 * k + 1]) + b.y);
          uint32_t a1 = pack_bf16x2(__uint_as_float(ra[4 * k + 2]) + b.z, __uint_as_float(ra[4 * k + 3]) + b.w);
          uint32_t c0 = pack_bf16x2(__uint_as_float(rb[4 * k]) + b.x, __uint_as_float(rb[4 * k + 1]) + b.y);
          uint32_t c1 = pack_bf16x2(__uint_as_float(rb[4 * k + 2]) + b.z, __uint_as_float(rb[4 * k + 3]) + b.w);
          if (relu) { a0 = relu_bf16x2(a0); a1 = relu_bf16x2(a1); c0 = relu_bf16x2(c0); c1 = relu_bf16x2(c1); }
          vm[2 * k] = hmax2(hmax2(prev[2 * k], a0), c0);
          vm[2 * k + 1] = hmax2(hmax2(prev[2 * k + 1], a1), c1);
          prev[2 * k] = c0;
          prev[2 * k + 1] = c1;
        }
        if (j < j_store) continue;  // the tile above the run: it only seeded `prev`
        // horizontal max over columns q - 1, q, q + 1 (needed on the even columns).  Column q - 1 of a
        // quadrant's first lane belongs to the neighbouring warp: it crosses through shared memory.
        const uint32_t par = stored++ & 1u;
        if (lane == 31) {
          uint4* dst = xch + ((par * kSpParts + part) * 4 + quad) * 2;
          dst[0] = make_uint4(vm[0], vm[1], vm[2], vm[3]);
          dst[1] = make_uint4(vm[4], vm[5], vm[6], vm[7]);
        }
        named_bar_sync(1 + part, 128);  // the four quadrant warps of this channel quarter
        uint32_t o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = __shfl_up_sync(0xffffffffu, vm[k], 1);
        if (lane == 0) {
          if (quad > 0) {
            const uint4* src = xch + ((par * kSpParts + part) * 4 + quad - 1) * 2;
            const uint4 t0 = src[0], t1 = src[1];
            o[0] = t0.x; o[1] = t0.y; o[2] = t0.z; o[3] = t0.w;
            o[4] = t1.x; o[5] = t1.y; o[6] = t1.z; o[7] = t1.w;
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = kNegInf2;  // column -1: padding
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = hmax2(hmax2(o[k], vm[k]), __shfl_down_sync(0xffffffffu, vm[k], 1));
        const int i = q >> 1;  // pooled column
        if ((lane & 1) == 0 && i < p.Qp) {
          uint4* dst = reinterpret_cast<uint4*>(p.out + (((long long)n_img * p.Pp + j) * p.Qp + i) * 64 + part * 16);
          dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
          dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * kSpN);
  }
}

// ----------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------
int stem_pool_supported(int H, int W, int Cin, int Cout, int R, int S, int stride, int pad, int pool_k, int pool_stride,
                        int pool_pad) {
  if (!stem_supported(H, W, Cin, Cout, R, S, stride, pad)) return 0;
  const int P = (H + 6 - 7) / 2 + 1, Q = (W + 6 - 7) / 2 + 1;
  return pool_k == 3 && pool_stride == 2 && pool_pad == 1 && W % 16 == 0 && H % 4 == 0 && H >= 12 && Q <= kBlockM && P % 2 == 0 &&
         Q % 2 == 0 && (Q + 10 + 7) / 8 <= kSpMaxGroups;
}

// [64][7][7][4] KRSC bf16 weights -> [9 input rows][4 taps][128][8] (slab-major: exactly the resident
// shared-memory layout): row n < 64 of slab (i, t) is output channel n of stem row 2j (filter row i), row
// 64 + n the same channel of stem row 2j + 1 (filter row i - 2); taps as in stem_repack_weights (tap t:
// channels 0-3 <- filter column 2t - 1, channels 4-7 <- column 2t).
void stem_pool_repack_weights(const uint16_t* w_krsc, uint16_t* w_sp) {
  memset(w_sp, 0, (size_t)kSpN * kSpRows * kSpTaps * 8 * sizeof(uint16_t));
  for (int n = 0; n < kSpN; ++n) {
    const int co = n & 63, shift = n < 64 ? 0 : 2;
    for (int i = 0; i < kSpRows; ++i) {
      const int r = i - shift;
      if (r < 0 || r >= 7) continue;
      for (int t = 0; t < kSpTaps; ++t)
        for (int c = 0; c < 8; ++c) {
          const int s = (c < 4) ? 2 * t - 1 : 2 * t;
          if (s < 0) continue;
          w_sp[(((i * kSpTaps + t) * kSpN) + n) * 8 + c] = w_krsc[((co * 7 + r) * 7 + s) * 4 + (c & 3)];
        }
    }
  }
}

int stem_pool_weight_elems() { return kSpN * kSpRows * kSpTaps * 8; }

int stem_pool_make_plan(StemPoolPlan* plan, const void* in, int N, int H, int W, const void* weight_sp,
                        const float* bias, int relu, void* out) {
  AI4E_REQUIRE(stem_pool_supported(H, W, 4, 64, 7, 7, 2, 3, 3, 2, 1), "stem+pool: unsupported input %dx%d", H, W);
  AI4E_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(weight_sp) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(out) & 15) == 0 && (reinterpret_cast<uintptr_t>(bias) & 15) == 0,
               "stem+pool: pointers must be 16-byte aligned");
  const int P = (H + 6 - 7) / 2 + 1, Q = (W + 6 - 7) / 2 + 1;
  plan->H = H; plan->P = P; plan->Q = Q; plan->Pp = P / 2; plan->Qp = Q / 2;
  const int groups = (Q + 10 + 7) / 8;
  plan->stem_rb = groups * 128;
  plan->relu = relu; plan->bias = bias; plan->out = out;
  plan->max_batch = N;
  plan->grid_sms = 0;
  const uint32_t es4[4] = {1, 1, 1, 1};
  const uint64_t adims[4] = {64, (uint64_t)W / 16, (uint64_t)H, (uint64_t)N};
  const uint64_t astr[3] = {128, (uint64_t)W * 8, (uint64_t)H * W * 8};
  const uint32_t abox[4] = {64, (uint32_t)groups, (uint32_t)kSpRows, 1};
  int rc = tmap_encode_tiled(&plan->tmap_a, in, 4, adims, astr, abox, es4, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc) return rc;
  plan->weights = weight_sp;  // [36 slabs][128 channels][8], fetched by one bulk copy
  plan->tmap_b = plan->tmap_a;  // (unused)
  return AI4E_OK;
}

// Runs are the unit of work of a CTA: `run_len` pooled rows of one image plus, below the top of the image,
// one seed tile.  Short runs balance better, long runs recompute fewer seed tiles; the pick minimises the
// tiles of the busiest CTA (micro-batch 64, 56 pooled rows, 132 CTAs: 128 runs of 28 rows -> 29 tiles per
// CTA, where runs of 8 gave 36).
static void stem_pool_pick_runs(int Pp, int batch, int ctas, StemPoolParams* prm) {
  long long best_cost = -1;
  for (int len = 2; len <= Pp; ++len) {
    const int rpi = (Pp + len - 1) / len;
    const long long runs = (long long)rpi * batch;
    const long long rounds = (runs + ctas - 1) / ctas;
    const long long cost = rounds * (len + 1) * 1000 + rpi;  // ties: fewer seed tiles
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      prm->run_len = len; prm->runs_per_img = rpi; prm->num_runs = (int)runs;
    }
  }
  if (best_cost < 0) { prm->run_len = Pp; prm->runs_per_img = 1; prm->num_runs = batch; }
}

int stem_pool_launch(const StemPoolPlan& plan, int batch, cudaStream_t stream) {
  AI4E_REQUIRE(batch > 0 && batch <= plan.max_batch, "stem+pool: batch %d outside plan (max %d)", batch, plan.max_batch);
  StemPoolParams prm{};
  prm.H = plan.H; prm.P = plan.P; prm.Q = plan.Q; prm.Pp = plan.Pp; prm.Qp = plan.Qp;
  prm.stem_rb = plan.stem_rb; prm.relu = plan.relu; prm.bias = plan.bias; prm.out = (bf16*)plan.out; prm.weights = plan.weights;
  const int sms = plan.grid_sms > 0 ? std::min(plan.grid_sms, sm_count()) : sm_count();
  stem_pool_pick_runs(plan.Pp, batch, sms, &prm);
  static std::atomic<unsigned long long> attr_done{0};
  if (int rc = ensure_dyn_smem(stem_pool_kernel<false>, kSpSmem, attr_done)) return rc;
  AI4E_CUDA_TRY(launch_pdl(stem_pool_kernel<false>, dim3(std::min(prm.num_runs, sms)), dim3(kSpThreads), (size_t)kSpSmem,
                           stream, plan.tmap_a, plan.tmap_b, prm));
  return AI4E_OK;
}

// u8 mode: K1 (no resize, 3 channels) inside the kernel.  `d_images` is the dense uint8 [batch][H][W][3]
// batch; its tensor map is encoded per call (a host-side computation) because the batch lives wherever the
// caller put it.
int stem_pool_u8_supported(int H, int W, int channels) {
  return channels == 3 && stem_pool_supported(H, W, 4, 64, 7, 7, 2, 3, 3, 2, 1) && W * 3 / 4 <= 256;
}

int stem_pool_launch_u8(const StemPoolPlan& plan, const uint8_t* d_images, int batch, const float* mean, const float* stdv,
                        cudaStream_t stream) {
  AI4E_REQUIRE(batch > 0 && batch <= plan.max_batch, "stem+pool: batch %d outside plan (max %d)", batch, plan.max_batch);
  const int W = plan.Q * 2;
  AI4E_REQUIRE(stem_pool_u8_supported(plan.H, W, 3), "stem+pool(u8): unsupported input %d x %d", plan.H, W);
  AI4E_REQUIRE((reinterpret_cast<uintptr_t>(d_images) & 15) == 0, "stem+pool(u8): the image batch must be 16-byte aligned");
  CUtensorMap tmap_u8;
  const uint32_t es3[3] = {1, 1, 1};
  const uint64_t dims[3] = {(uint64_t)W * 3 / 4, (uint64_t)plan.H, (uint64_t)batch};
  const uint64_t str[2] = {(uint64_t)W * 3, (uint64_t)plan.H * W * 3};
  const uint32_t box[3] = {(uint32_t)(W * 3 / 4), (uint32_t)kSpRows, 1};
  int rc = tmap_encode_tiled_as(&tmap_u8, d_images, 3, dims, str, box, es3, CU_TENSOR_MAP_SWIZZLE_NONE, /*u32=*/true);
  if (rc) return rc;
  StemPoolParams prm{};
  prm.W = W;
  for (int c = 0; c < 3; ++c) { prm.mean[c] = mean[c]; prm.stdv[c] = stdv[c]; }
  prm.H = plan.H; prm.P = plan.P; prm.Q = plan.Q; prm.Pp = plan.Pp; prm.Qp = plan.Qp;
  prm.stem_rb = plan.stem_rb; prm.relu = plan.relu; prm.bias = plan.bias; prm.out = (bf16*)plan.out; prm.weights = plan.weights;
  const int sms = plan.grid_sms > 0 ? std::min(plan.grid_sms, sm_count()) : sm_count();
  stem_pool_pick_runs(plan.Pp, batch, sms, &prm);
  static std::atomic<unsigned long long> attr_done{0};
  if (int rc2 = ensure_dyn_smem(stem_pool_kernel<true>, kSpSmemU8, attr_done)) return rc2;
  AI4E_CUDA_TRY(launch_pdl(stem_pool_kernel<true>, dim3(std::min(prm.num_runs, sms)), dim3(kSpThreads), (size_t)kSpSmemU8,
                           stream, tmap_u8, plan.tmap_b, prm));
  return AI4E_OK;
}

}  // namespace ai4e
