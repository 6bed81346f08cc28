// K4 — implicit-GEMM convolution on the 5th-generation tensor cores (sm_100a).
//
//   out[m, co] = act( sum_{r,s,c} in[n, p*stride-pad+r, q*stride-pad+s, c] * w[co, r, s, c]
//                     + bias[co] (+ residual[m, co]) ),     m = (n*P + p)*Q + q
//
// NHWC bf16 activations, KRSC bf16 weights (BatchNorm folded), fp32 bias, fp32 accumulation.
// GEMM view: D[M = N*P*Q, Cout] = A[M, K = R*S*Cin] * B[Cout, K]^T, both operands K-major.
//
// This is the B200 replacement for the cuDNN/oneDNN convolution a PyTorch container runs
// inside the callback the reference invokes at APIs/1.0/base-py/ai4e_service.py:203;
// oracle: torch.nn.functional.conv2d in oracle/resnet_oracle.py (_conv_bf16).
//
// Structure (one persistent CTA per SM, 192 threads, warp-specialised):
//   warp 0   TMA producer.  A tiles come straight from the NHWC activation tensor through an
//            IM2COL-mode tensor map: one `cp.async.bulk.tensor.4d...im2col` per K block
//            fetches 128 consecutive output pixels x 64 channels for filter tap (r, s), with
//            stride, padding (zero fill) and the wrap across rows / images done by the TMA
//            unit.  B tiles are plain 2-D tiles of the weight matrix.  Both land in
//            128-byte-swizzled shared memory, STAGES deep, mbarrier `full`/`empty` rings.
//   warp 1   allocates TMEM and issues `tcgen05.mma.cta_group::1.kind::f16` (M = 128,
//            N = BLOCK_N, K = 16) from one elected lane; `tcgen05.commit` releases smem
//            stages and signals the epilogue.  Two accumulator stages in TMEM
//            (2 x BLOCK_N fp32 columns) let the epilogue of tile i overlap the MMAs of
//            tile i+1.
//   warps 2-5  epilogue: `tcgen05.ld` 32 lanes x 32 columns at a time, + bias, + residual,
//            ReLU, pack to bf16, 16-byte stores.  A warp may only touch the TMEM lane
//            quadrant `warp_id % 4`, hence four consecutive warps.
#include "common.cuh"
#include "kernels.h"

#include <algorithm>

namespace ai4e {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int kUmmaK = 16;
constexpr int kIgemmThreads = 192;
constexpr int kABytes = kBlockM * kBlockK * 2;  // 16 KiB

template <int BLOCK_N>
struct IgemmCfg {
  static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BLOCK_N == 256) ? 4 : (BLOCK_N == 128 ? 6 : 8);
  static constexpr int kTmemCols = 2 * BLOCK_N;  // 128 / 256 / 512: power of two >= 32
  // 1024-byte alignment slack + stages + barriers (full, empty, tmem_full[2], tmem_empty[2]) + tmem ptr
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + (2 * kStages + 4) * 8 + 16;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kIgemmThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a,
                  const __grid_constant__ CUtensorMap tmap_b, const IgemmParams p) {
  using Cfg = IgemmCfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + kStages * Cfg::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tmem_full_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
  auto tmem_empty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 2 + s); };
  const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * kStages + 4);
  volatile uint32_t* tmem_ptr_gen =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_addr - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tmem_full_bar(s), 1);
      mbar_init(tmem_empty_bar(s), 128);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_addr, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ------------------------------- TMA producer -------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int tile_n = tile % p.tiles_n;
        const int tile_m = tile / p.tiles_n;
        const int m0 = tile_m * kBlockM;
        const int n_img = m0 / p.PQ;
        const int rem = m0 - n_img * p.PQ;
        const int p0 = rem / p.Q;
        const int q0 = rem - p0 * p.Q;
        const int w_base = q0 * p.stride - p.pad;
        const int h_base = p0 * p.stride - p.pad;
        int r = 0, s = 0, cb = 0;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t a_dst = smem_base + stage * Cfg::kStageBytes;
          const uint32_t b_dst = a_dst + kABytes;
          mbar_expect_tx(full_bar(stage), Cfg::kStageBytes);
          tma_load_im2col_4d(a_dst, &tmap_a, full_bar(stage), cb * kBlockK, w_base, h_base, n_img,
                             (uint16_t)s, (uint16_t)r);
          tma_load_2d(b_dst, &tmap_b, full_bar(stage), kb * kBlockK, tile_n * BLOCK_N);
          if (++cb == p.cin_blocks) { cb = 0; if (++s == p.S) { s = 0; ++r; } }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer ---------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(kBlockM, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
        const uint32_t acc = it & 1u;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(tmem_empty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * Cfg::kStageBytes;
          const uint64_t adesc = umma_smem_desc_kmajor(a_addr, 128, 2);
          const uint64_t bdesc = umma_smem_desc_kmajor(a_addr + kABytes, 128, 2);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // advancing 16 bf16 = 32 bytes along K inside the 128-byte swizzle row:
            // +2 in the (addr >> 4) start-address field
            umma_bf16_ss(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));  // smem stage reusable once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tmem_full_bar(acc));  // accumulator complete -> epilogue
      }
    }
  } else {
    // ------------------------------- epilogue -----------------------------------
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const uint32_t acc = it & 1u;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const int tile_n = tile % p.tiles_n;
      const int tile_m = tile / p.tiles_n;
      const long long m = (long long)tile_m * kBlockM + row;
      const bool valid = m < p.M;
      const int n0 = tile_n * BLOCK_N;
      mbar_wait(tmem_full_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + acc * BLOCK_N + ((uint32_t)(quad * 32) << 16);
      bf16* out_row = p.out + m * p.Cout + n0;
      const bf16* res_row = p.residual ? p.residual + m * p.Cout + n0 : nullptr;
#pragma unroll 1
      for (int ch = 0; ch < BLOCK_N / 32; ++ch) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(t_addr + ch * 32, v);
        tmem_wait_ld();
        if (valid) {
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + n0 + ch * 32);
          uint32_t o[16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = __ldg(b4 + j);
            float x0 = __uint_as_float(v[4 * j + 0]) + b.x;
            float x1 = __uint_as_float(v[4 * j + 1]) + b.y;
            float x2 = __uint_as_float(v[4 * j + 2]) + b.z;
            float x3 = __uint_as_float(v[4 * j + 3]) + b.w;
            if (res_row) {
              const uint2 rr = __ldg(reinterpret_cast<const uint2*>(res_row + ch * 32 + 4 * j));
              x0 += bf16_lo(rr.x); x1 += bf16_hi(rr.x); x2 += bf16_lo(rr.y); x3 += bf16_hi(rr.y);
            }
            if (p.relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f); }
            o[2 * j] = pack_bf16x2(x0, x1);
            o[2 * j + 1] = pack_bf16x2(x2, x3);
          }
          uint4* dst = reinterpret_cast<uint4*>(out_row + ch * 32);
#pragma unroll
          for (int j = 0; j < 4; ++j) dst[j] = make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        }
      }
      tc_fence_before();
      mbar_arrive(tmem_empty_bar(acc));  // 128 arrivals free this accumulator stage
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ----------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------
int igemm_pick_block_n(int Cout, long long M) {
  const int sms = sm_count();
  const long long tiles_m = (M + kBlockM - 1) / kBlockM;
  // the widest tile that still leaves at least one tile per SM; wider tiles re-read A less
  for (int bn : {256, 128, 64}) {
    if (Cout % bn) continue;
    if (tiles_m * (Cout / bn) >= sms || bn == 64) return bn;
  }
  for (int bn : {64, 128, 256})
    if (Cout % bn == 0) return bn;
  return 0;
}

int igemm_supported(int Cin, int Cout, int R, int S, int stride, int pad) {
  // im2col-mode TMA corner offsets are 8-bit signed, filter offsets 8-bit unsigned (2 spatial dims)
  return Cin % kBlockK == 0 && Cout % 64 == 0 && R >= 1 && S >= 1 && R <= 16 && S <= 16 &&
         pad >= 0 && pad < 128 && stride >= 1 && stride <= 8;
}

// Builds the tensor maps for a layer whose activation tensors hold up to `N` images.  The
// same plan then serves any batch <= N: only M (hence the tile count) changes per launch;
// the last tile of a smaller batch reads rows of images >= batch (stale but in-bounds) and
// the epilogue masks them.
int igemm_make_plan(IgemmPlan* plan, const void* in, int N, int H, int W, int Cin,
                    const void* weight, const float* bias, int Cout, int R, int S, int stride,
                    int pad, const void* residual, int relu, void* out) {
  AI4E_REQUIRE(igemm_supported(Cin, Cout, R, S, stride, pad),
               "conv(igemm): unsupported shape Cin=%d Cout=%d %dx%d s%d p%d (needs Cin%%64==0, Cout%%64==0)",
               Cin, Cout, R, S, stride, pad);
  AI4E_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(weight) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(out) & 15) == 0 && (reinterpret_cast<uintptr_t>(bias) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(residual) & 15) == 0,
               "conv(igemm): pointers must be 16-byte aligned");
  const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
  AI4E_REQUIRE(P > 0 && Q > 0, "conv(igemm): empty output");
  AI4E_REQUIRE((long long)N * P * Q < (1ll << 31) - kBlockM, "conv(igemm): M too large");
  IgemmParams& q = plan->prm;
  q.M = 0; q.P = P; q.Q = Q; q.PQ = P * Q; q.Cout = Cout; q.S = S; q.stride = stride; q.pad = pad;
  q.cin_blocks = Cin / kBlockK; q.num_k_blocks = R * S * q.cin_blocks;
  q.tiles_m = q.tiles_n = q.num_tiles = 0; q.relu = relu; q.bias = bias;
  q.residual = (const bf16*)residual; q.out = (bf16*)out;
  plan->max_batch = N;
  int rc = tmap_encode_im2col_nhwc(&plan->tmap_a, in, N, H, W, Cin, R, S, stride, pad, kBlockK, kBlockM);
  if (rc) return rc;
  const uint64_t ktot = (uint64_t)R * S * Cin;
  const uint64_t dims[2] = {ktot, (uint64_t)Cout};
  const uint64_t strides[1] = {ktot * 2};
  const uint32_t es[2] = {1, 1};
  const int bns[3] = {64, 128, 256};
  for (int i = 0; i < 3; ++i) {
    plan->has_bn[i] = Cout % bns[i] == 0;
    if (!plan->has_bn[i]) continue;
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)bns[i]};
    rc = tmap_encode_tiled(&plan->tmap_b[i], weight, 2, dims, strides, box, es, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  return AI4E_OK;
}

template <int BN>
static int igemm_launch_bn(const CUtensorMap& ta, const CUtensorMap& tb, const IgemmParams& prm, int grid,
                           cudaStream_t stream) {
  static unsigned long long attr_done = 0;  // bit per device
  int dev = 0;
  AI4E_CUDA_TRY(cudaGetDevice(&dev));
  if (!(attr_done >> dev & 1ull)) {
    AI4E_CUDA_TRY(cudaFuncSetAttribute(conv_igemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       IgemmCfg<BN>::kSmemBytes));
    attr_done |= 1ull << dev;
  }
  conv_igemm_kernel<BN><<<grid, kIgemmThreads, IgemmCfg<BN>::kSmemBytes, stream>>>(ta, tb, prm);
  AI4E_CUDA_TRY(cudaGetLastError());
  return AI4E_OK;
}

int igemm_launch(const IgemmPlan& plan, int batch, int force_block_n, cudaStream_t stream) {
  AI4E_REQUIRE(batch > 0 && batch <= plan.max_batch, "conv(igemm): batch %d outside plan (max %d)", batch,
               plan.max_batch);
  IgemmParams prm = plan.prm;
  const long long M = (long long)batch * prm.PQ;
  const int bn = force_block_n ? force_block_n : igemm_pick_block_n(prm.Cout, M);
  const int bi = bn == 64 ? 0 : bn == 128 ? 1 : bn == 256 ? 2 : -1;
  AI4E_REQUIRE(bi >= 0 && plan.has_bn[bi], "conv(igemm): BLOCK_N %d not available for Cout %d", bn, prm.Cout);
  prm.M = (int)M;
  prm.tiles_m = (int)((M + kBlockM - 1) / kBlockM);
  prm.tiles_n = prm.Cout / bn;
  prm.num_tiles = prm.tiles_m * prm.tiles_n;
  const int grid = std::min(prm.num_tiles, sm_count());
  switch (bn) {
    case 64: return igemm_launch_bn<64>(plan.tmap_a, plan.tmap_b[bi], prm, grid, stream);
    case 128: return igemm_launch_bn<128>(plan.tmap_a, plan.tmap_b[bi], prm, grid, stream);
    default: return igemm_launch_bn<256>(plan.tmap_a, plan.tmap_b[bi], prm, grid, stream);
  }
}

}  // namespace ai4e
