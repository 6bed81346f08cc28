// K4b - the tail of a ResNet bottleneck in one kernel:
//
//   y = relu(conv_RxS(x; w2) + b2)            [M, Cmid]     (torchvision: conv2 + bn2 + relu)
//   z = relu(y * w3^T + b3 + residual)        [M, Cout]     (conv3 1x1 + bn3 + add + relu)
//
// y never goes to HBM.  Separately these are the 3x3 convolution (K4) followed by the most
// memory-bound layer of the network - the 1x1 expansion reads y and the residual and writes z,
// 16 FLOP per byte - which ran at the HBM bound while the tensor cores idled, and paid a launch, a
// pipeline fill and an epilogue tail of its own.  Fused, z's GEMM (K = Cmid) runs in the shadow of
// the first one and HBM sees x, the residual and z only.
//
// One persistent CTA per SM, 384 threads, roles as in conv_igemm.cu.  Per 128-pixel tile:
//   1. K loop of the first GEMM (im2col-mode TMA for A, w2 tiles for B) -> acc1 in TMEM
//      (2 x Cmid columns: the next tile's K loop overlaps this tile's second GEMM).
//   2. The epilogue warps turn acc1 into relu(. + b2) in bf16 and write it to the `mid` tile in
//      shared memory, in exactly the layout the tensor core wants for a K-major A operand
//      (Cmid / 64 blocks of [128 rows][128 bytes], 128-byte swizzle) - the layout the other
//      kernel's store staging already uses.
//   3. Second GEMM, 128 output channels at a time: A = mid, B = a [128][64] tile of w3 that
//      travels through the A slot of the same stage ring -> acc2 (2 x 128 columns, so the MMAs
//      of one 128-channel pass overlap the epilogue of the previous one).
//   4. The epilogue warps add b3 and the residual (TMA-loaded by the residual warp), apply ReLU
//      and store through TMA.  The residual buffer doubles as the store staging: a thread
//      overwrites the 128 bytes it has just read, so no separate staging tiles are needed and the
//      shared memory goes to pipeline stages instead.
// The MMA warp issues  KL1(0) KL1(1) P2(0) KL1(2) P2(1) ...  and both producer warps fill the
// stage ring in that order.
//
// Arithmetic is identical to running the two layers separately (y is rounded to bf16 either way).
// Oracle: the same two conv calls in oracle/resnet_oracle.py.  Reference call site:
// APIs/1.0/base-py/ai4e_service.py:203 (the user callback that runs the PyTorch model).
#include "common.cuh"
#include "kernels.h"
#include "igemm_tiles.cuh"

#include <algorithm>
#include <cstdlib>

// Developer aid, as in conv_igemm.cu: -DAI4E_PIPE_TRACE makes CTA 0 stamp its pipeline into FusedParams::dbg
// (dumped by ai4e_bottleneck_tail when AI4E_B200_PIPE_TRACE=<file> is set; tools/trace_decode_fused.py reads it).
// [0,256) A producer: slot free for ring unit g | [256,512) MMA warp starts waiting for unit g | [512,768) data seen |
// [768,1024) MMAs issued | 1024 + 16 * tile + j: MMA warp (0/1 acc1_empty wait, 2/3 mid_full wait, 4..7 acc2_empty of pass
// 0..3 granted), epilogue warp 0 (8 acc1_full seen, 9 mid_empty seen, 10 step 2 done, 11..14 acc2_full of pass 0..3 seen,
// 15 last store issued) | 1600.. kernel phases
// (-DAI4E_PIPE_TRACE=1: the per-tile and phase stamps only; =2: also the per-unit stamps, which cost the MMA warp ~200
// cycles per unit and so overstate the K loop.)
#ifdef AI4E_PIPE_TRACE
#define FT(cond, slot) do { if (p.dbg && blockIdx.x == 0 && ((slot) >= 1024 || AI4E_PIPE_TRACE >= 2) && (cond) && (slot) < 2048) p.dbg[(slot)] = clock64(); } while (0)
#else
#define FT(cond, slot) do { } while (0)
#endif

namespace ai4e {

template <int N1>  // Cmid: 64 or 128
struct FusedCfg {
  static_assert(N1 == 64 || N1 == 128, "Cmid 64 or 128");
  static constexpr int kKB2 = N1 / kBlockK;               // K blocks of the second GEMM
  static constexpr int kBBytes = N1 * kBlockK * 2;        // w2 tile
  static constexpr int kStageBytes = kABytes + kBBytes;   // 24 / 32 KiB
  static constexpr int kMidBytes = kKB2 * kChunkBytes;    // 16 / 32 KiB
  static constexpr int kBiasBytes = kEpiWarps * kChunkN * 4;
  static constexpr int kResBytes = 4 * kChunkBytes;       // two residual/staging buffers per epilogue side
#ifndef AI4E_FUSED_MAX_STAGES
#define AI4E_FUSED_MAX_STAGES 8
#endif
  static constexpr int kMaxStages = AI4E_FUSED_MAX_STAGES;  // (-DAI4E_FUSED_MAX_STAGES=n: stage-count experiments)
  // full[], empty[], acc1_full/empty[2], mid_full/empty, acc2_full/empty[2], res_full/empty[4], tmem ptr
  static constexpr int kBarBytes = (2 * kMaxStages + 4 + 2 + 4 + 8 + 1) * 8 + 16;
  static constexpr int kSmemLimit = 227 * 1024;
  static constexpr int kStages =
      std::min(kMaxStages, (kSmemLimit - kMidBytes - kBiasBytes - kResBytes - kBarBytes) / kStageBytes);
  static constexpr int kSmem = kStages * kStageBytes + kMidBytes + kBiasBytes + kResBytes + kBarBytes;
  static constexpr int kAcc2Col = 2 * N1;                 // acc1: [0, 2 N1), acc2: two 128-column buffers after it
  static constexpr int kTmemCols = 512;
  static constexpr int kE1Warps = 4 * (kKB2 < 2 ? kKB2 : 2);  // epilogue warps that take part in step 2
  // Cmid = 128: the two K blocks of a w3 pass travel in ONE stage (A slot + B slot, 16 KiB each).  The
  // ring is latency-bound (a slot takes ~3100 cycles to go round, 4 stages fit), so what counts is the
  // number of stage fills per pixel tile: 18 + 4 instead of 18 + 8.
  static constexpr bool kPackW3 = kKB2 == 2 && kBBytes == kABytes;
  static constexpr int kUnitsPerPass = kPackW3 ? 1 : kKB2;
};

// Order of the ring's units inside one iteration `st` of the persistent loop: the nk K blocks of tile st's first
// GEMM and the nu units of tile st-1's second GEMM.  Plain order: all K blocks, then all units - the MMA warp
// then sits in front of acc2 buffers the epilogue has not drained yet (a pass takes the epilogue ~2,400 cycles,
// the tensor core ~540) with the ring full and no load in flight behind it.  Interleaved: unit u goes after K
// block fused_unit_slot(u), evenly spread, so every pass finds its accumulator free and the ring keeps moving.
// Both producers and the MMA warp walk the same sequence.  (slot <= 0 or no K blocks: the units follow at the end.)
__device__ __forceinline__ int fused_unit_slot(int u, int nk, int nu, int interleave) {
  if (interleave == 2) return ((2 * u + 3) * nk) / (2 * (nu + 1));  // half a step later: more room for the first epilogue step
  return interleave ? ((u + 1) * nk) / (nu + 1) : 0;
}

template <int N1>
__global__ void __launch_bounds__(kIgemmThreads, 1)
bottleneck_tail_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w2,
                       const __grid_constant__ CUtensorMap tmap_w3, const __grid_constant__ CUtensorMap tmap_out,
                       const __grid_constant__ CUtensorMap tmap_res, const FusedParams p) {
  using Cfg = FusedCfg<N1>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kKB2 = Cfg::kKB2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  if (smem_base & 1023u) __trap();
  const uint32_t mid_base = smem_base + kStages * Cfg::kStageBytes;    // [kKB2][128 rows][128 B]
  const uint32_t res_base = mid_base + Cfg::kMidBytes;                 // [side][2][128 rows][128 B]
  const uint32_t bias_base = res_base + Cfg::kResBytes;                // [8 warps][64 floats]
  const uint32_t bar_base = bias_base + Cfg::kBiasBytes;
  constexpr int kMS = Cfg::kMaxStages;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMS + s); };
  auto acc1_full = [&](int s) { return bar_base + 8u * (2 * kMS + s); };
  auto acc1_empty = [&](int s) { return bar_base + 8u * (2 * kMS + 2 + s); };
  const uint32_t mid_full = bar_base + 8u * (2 * kMS + 4);
  const uint32_t mid_empty = bar_base + 8u * (2 * kMS + 5);
  auto acc2_full = [&](int s) { return bar_base + 8u * (2 * kMS + 6 + s); };
  auto acc2_empty = [&](int s) { return bar_base + 8u * (2 * kMS + 8 + s); };
  auto res_full = [&](int s) { return bar_base + 8u * (2 * kMS + 10 + s); };
  auto res_empty = [&](int s) { return bar_base + 8u * (2 * kMS + 14 + s); };
  const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * kMS + 18);
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_addr - smem_base));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  FT(threadIdx.x == 0, 1600);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w2);
    tma_prefetch_desc(&tmap_w3);
    tma_prefetch_desc(&tmap_out);
    tma_prefetch_desc(&tmap_res);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 2);  // one arrival from each producer warp
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(acc1_full(s), 1);
      mbar_init(acc1_empty(s), Cfg::kE1Warps);
      mbar_init(acc2_full(s), 1);
      mbar_init(acc2_empty(s), kEpiWarps);
    }
    mbar_init(mid_full, Cfg::kE1Warps);
    mbar_init(mid_empty, 1);
    for (int s = 0; s < 4; ++s) {
      mbar_init(res_full(s), 1);
      mbar_init(res_empty(s), 4);
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
  pdl_launch_dependents();
  FT(threadIdx.x == 0, 1601);

  // this CTA's tiles: blockIdx.x, + gridDim.x, ...
  const int my_tiles = p.num_tiles > (int)blockIdx.x ? (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int passes = p.Cout / 128;
  // first pixel row of this CTA's i-th tile; the engine alternates the direction from layer to layer
  // so that a consumer starts with what its producer wrote last (still in L2)
  auto tile_row = [&](int i) {
    const int t = (int)blockIdx.x + i * (int)gridDim.x;
    return (p.reverse ? p.num_tiles - 1 - t : t) * kBlockM;
  };
  const int units2 = passes * Cfg::kUnitsPerPass;  // ring stages the second GEMM of a pixel tile takes

  if (warp == 0) {
    // ------------------- producer: A tiles (im2col) of the first GEMM, w3 tiles of the second ----
    int stage = 0;
    uint32_t phase = 0;
    bool waited = false;
    [[maybe_unused]] int g = 0;  // ring units so far (trace)
    for (int st = 0; st <= my_tiles; ++st) {
      const int nk = st < my_tiles ? p.num_k_blocks : 0;
      const int nu = st >= 1 ? units2 : 0;
      int w_base = 0, h_base = 0, n_img = 0;
      if (nk) {
        const int m0 = tile_row(st);
        n_img = m0 / p.PQ;
        const int rem = m0 - n_img * p.PQ;
        const int p0 = rem / p.Q;
        const int q0 = rem - p0 * p.Q;
        w_base = q0 * p.stride - p.pad; h_base = p0 * p.stride - p.pad;
        if (!waited) {  // x is the previous layer's output
          pdl_wait();
          waited = true;
          FT(lane == 0, 1602);
        }
      }
      auto unit = [&](int u) {  // pass-major: w3 rows [128 * pass, +128), K block u % kKB2
        mbar_wait(empty_bar(stage), phase ^ 1u);
        FT(lane == 0 && g < 256, g); ++g;
        if (elect_one()) {
          mbar_expect_tx(full_bar(stage), kABytes);
          if constexpr (Cfg::kPackW3)  // K block 0 of pass u here, K block 1 in the B slot (the other producer)
            tma_load_2d(smem_base + stage * Cfg::kStageBytes, &tmap_w3, full_bar(stage), 0, u * 128);
          else
            tma_load_2d(smem_base + stage * Cfg::kStageBytes, &tmap_w3, full_bar(stage), (u % kKB2) * kBlockK,
                        (u / kKB2) * 128);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      };
      int u = 0;
      int next_at = fused_unit_slot(0, nk, nu, p.interleave);  // the K block after which unit u goes (0: at the end)
      int r = 0, s = 0, cb = 0;
      for (int kb = 0; kb < nk; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        FT(lane == 0 && g < 256, g); ++g;
        if (elect_one()) {
          mbar_expect_tx(full_bar(stage), kABytes);
          tma_load_im2col_4d(smem_base + stage * Cfg::kStageBytes, &tmap_a, full_bar(stage), cb * kBlockK, w_base,
                             h_base, n_img, (uint16_t)s, (uint16_t)r);
        }
        __syncwarp();
        if (++cb == p.cin_blocks) { cb = 0; if (++s == p.S) { s = 0; ++r; } }
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
        while (u < nu && next_at == kb + 1) { unit(u++); next_at = fused_unit_slot(u, nk, nu, p.interleave); }
      }
      while (u < nu) unit(u++);
    }
  } else if (warp == 2 + kEpiWarps) {
    // ------------------- producer: w2 tiles; the w3 stages carry nothing in their B slot -------
    int stage = 0;
    uint32_t phase = 0;
    for (int st = 0; st <= my_tiles; ++st) {
      const int nk = st < my_tiles ? p.num_k_blocks : 0;
      const int nu = st >= 1 ? units2 : 0;
      auto unit = [&](int u) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        if (elect_one()) {
          if constexpr (Cfg::kPackW3) {
            mbar_expect_tx(full_bar(stage), Cfg::kBBytes);
            tma_load_2d(smem_base + stage * Cfg::kStageBytes + kABytes, &tmap_w3, full_bar(stage), kBlockK, u * 128);
          } else {
            mbar_arrive(full_bar(stage));
          }
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      };
      int u = 0;
      int next_at = fused_unit_slot(0, nk, nu, p.interleave);  // the K block after which unit u goes (0: at the end)
      for (int kb = 0; kb < nk; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(full_bar(stage), Cfg::kBBytes);
          tma_load_2d(smem_base + stage * Cfg::kStageBytes + kABytes, &tmap_w2, full_bar(stage), kb * kBlockK, 0);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
        while (u < nu && next_at == kb + 1) { unit(u++); next_at = fused_unit_slot(u, nk, nu, p.interleave); }
      }
      while (u < nu) unit(u++);
    }
  } else if (warp == 1) {
    // ------------------- MMA issuer (convergent warp, one elected lane) -----------------------
    constexpr uint32_t idesc1 = umma_idesc_bf16_f32(kBlockM, N1);
    constexpr uint32_t idesc2 = umma_idesc_bf16_f32(kBlockM, 128);
    int stage = 0;
    uint32_t phase = 0;
    // descriptor low word and barrier address of the current stage, advanced with it (see umma_desc_join)
    const uint32_t ring_lo0 = umma_desc_lo(smem_base), mid_lo = umma_desc_lo(mid_base);
    uint32_t ring_lo = ring_lo0, fbar = full_bar(0);
    constexpr uint32_t kALo = kABytes >> 4, kStageLo = Cfg::kStageBytes >> 4, kChunkLo = kChunkBytes >> 4, kEmptyOff = 8u * kMS;
    auto advance = [&] {
      ring_lo += kStageLo; fbar += 8u;
      if (++stage == kStages) { stage = 0; phase ^= 1u; ring_lo = ring_lo0; fbar = full_bar(0); }
    };
    uint32_t gp = 0;  // 128-channel passes issued so far: selects the acc2 buffer and its phase
    [[maybe_unused]] int g = 0;
    for (int st = 0; st <= my_tiles; ++st) {
      const int nk = st < my_tiles ? p.num_k_blocks : 0;
      const int nu = st >= 1 ? units2 : 0;
      // one unit of the second GEMM of tile st - 1 (its `mid` tile must be written before the first one)
      auto unit = [&](int u) {
        const int kb = u % Cfg::kUnitsPerPass;
        const uint32_t b = gp & 1u;
        if (u == 0) {
          FT(lane == 0 && st - 1 < 32, 1024 + 16 * (st - 1) + 2);
          mbar_wait(mid_full, (uint32_t)(st - 1) & 1u);
          tc_fence_after();
          FT(lane == 0 && st - 1 < 32, 1024 + 16 * (st - 1) + 3);
        }
        if (kb == 0) {
          mbar_wait(acc2_empty(b), ((gp >> 1) & 1u) ^ 1u);
          tc_fence_after();
          FT(lane == 0 && st - 1 < 32 && u / Cfg::kUnitsPerPass < 4, 1024 + 16 * (st - 1) + 4 + u / Cfg::kUnitsPerPass);
        }
        const uint32_t d_tmem = tmem_base + Cfg::kAcc2Col + b * 128;
        FT(lane == 0 && g < 256, 256 + g);
        mbar_wait(fbar, phase);
        tc_fence_after();
        FT(lane == 0 && g < 256, 512 + g);
        if (elect_one()) {
          if constexpr (Cfg::kPackW3) {
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {  // both K blocks of the pass sit in this stage
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_bf16_ss(d_tmem, umma_desc_join(mid_lo + kk * kChunkLo + 2u * k), umma_desc_join(ring_lo + kk * kALo + 2u * k),
                             idesc2, (kk | k) != 0 ? 1u : 0u);
            }
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_ss(d_tmem, umma_desc_join(mid_lo + kb * kChunkLo + 2u * k), umma_desc_join(ring_lo + 2u * k), idesc2,
                           (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(fbar + kEmptyOff);
          if (kb == Cfg::kUnitsPerPass - 1) umma_commit(acc2_full(b));
          if (u == nu - 1) umma_commit(mid_empty);  // `mid` may be overwritten once these MMAs retire
        }
        __syncwarp();
        FT(lane == 0 && g < 256, 768 + g); ++g;
        advance();
        if (kb == Cfg::kUnitsPerPass - 1) ++gp;
      };
      int u = 0;
      int next_at = fused_unit_slot(0, nk, nu, p.interleave);  // the K block after which unit u goes (0: at the end)
      const uint32_t a = st & 1u;
      const uint32_t d_tmem1 = tmem_base + a * N1;
      if (nk) {  // first GEMM of tile st
        FT(lane == 0 && st < 32, 1024 + 16 * st + 0);
        mbar_wait(acc1_empty(a), ((st >> 1) & 1u) ^ 1u);
        tc_fence_after();
        FT(lane == 0 && st < 32, 1024 + 16 * st + 1);
      }
      for (int kb = 0; kb < nk; ++kb) {
        FT(lane == 0 && g < 256, 256 + g);
        mbar_wait(fbar, phase);
        tc_fence_after();
        FT(lane == 0 && g < 256, 512 + g);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16_ss(d_tmem1, umma_desc_join(ring_lo + 2u * k), umma_desc_join(ring_lo + kALo + 2u * k), idesc1,
                         (kb | k) != 0 ? 1u : 0u);
          umma_commit(fbar + kEmptyOff);
          if (kb == nk - 1) umma_commit(acc1_full(a));
        }
        __syncwarp();
        FT(lane == 0 && g < 256, 768 + g); ++g;
        advance();
        while (u < nu && next_at == kb + 1) { unit(u++); next_at = fused_unit_slot(u, nk, nu, p.interleave); }
      }
      while (u < nu) unit(u++);
    }
  } else if (warp < 2 + kEpiWarps) {
    // ------------------- epilogue warps ------------------------------------------------------
    const int ew = warp - 2;
    const int side = ew >> 2;
    const int quad = warp & 3;
    float* const bias_gen = reinterpret_cast<float*>(smem_raw + (bias_base - smem_base)) + ew * kChunkN;
    const float4* const b4 = reinterpret_cast<const float4*>(bias_gen);
    const uint32_t sw = (uint32_t)(lane & 7);
    const uint32_t row_off = (uint32_t)(quad * 32 + lane) * 128u;
    const uint32_t lane_tmem = (uint32_t)(quad * 32) << 16;
    uint32_t v[64];
    uint32_t (&vlo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[0]);
    uint32_t (&vhi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[32]);
    uint4 o[8];
    pdl_wait();  // z may still be read by the previous layer (it can be that layer's input buffer)
    uint32_t gs = 0;          // second-step chunks this warp has processed (= passes so far)
    int pending_buf = -1;     // residual/staging buffer whose TMA store may still be reading it
    for (int i = 0; i < my_tiles; ++i) {
      const int m0 = tile_row(i);
      // ---- step 2: acc1 -> relu(. + b2) -> bf16 `mid` tile (chunks c = side, side + 2, ...)
      if (side < kKB2) {
        const uint32_t a = i & 1u;
        mbar_wait(acc1_full(a), (i >> 1) & 1u);
        tc_fence_after();
        FT(ew == 0 && lane == 0 && i < 32, 1024 + 16 * i + 8);
        if (i >= 1) mbar_wait(mid_empty, (i - 1) & 1u);  // the previous tile's second GEMM has read `mid`
        FT(ew == 0 && lane == 0 && i < 32, 1024 + 16 * i + 9);
        for (int c = side; c < kKB2; c += 2) {
          const uint32_t t_addr = tmem_base + a * N1 + c * kChunkN + lane_tmem;
          tmem_ld_32x32b_x32(t_addr, vlo);
          tmem_ld_32x32b_x32(t_addr + 32, vhi);
          __syncwarp();  // the previous chunk's bias readers are done
          reinterpret_cast<float2*>(bias_gen)[lane] = __ldg(reinterpret_cast<const float2*>(p.bias2 + c * kChunkN) + lane);
          __syncwarp();
          tmem_wait_ld_pin(vlo);
          tmem_wait_ld_pin(vhi);
          epi_row<false, true>(v, b4, nullptr, sw, o);
          uint8_t* dst = smem_raw + (mid_base - smem_base) + c * kChunkBytes + row_off;
#pragma unroll
          for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(dst + ((((uint32_t)j) ^ sw) << 4)) = o[j];
        }
        tc_fence_before();
        fence_proxy_async_smem();  // the tensor core reads `mid` through the async proxy
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(acc1_empty(a));
          mbar_arrive(mid_full);
        }
        FT(ew == 0 && lane == 0 && i < 32, 1024 + 16 * i + 10);
      }
      // ---- step 4: acc2 (+ b3 + residual, relu) -> z, one 64-channel chunk per pass and side
      for (int ps = 0; ps < passes; ++ps, ++gs) {
        const uint32_t b = gs & 1u;  // every pass is taken by both sides: gs counts passes, like the MMA warp's gp
        const int n0c = ps * 128 + side * kChunkN;
        const float2 bias2 = __ldg(reinterpret_cast<const float2*>(p.bias3 + n0c) + lane);
        mbar_wait(acc2_full(b), (gs >> 1) & 1u);
        tc_fence_after();
        FT(ew == 0 && lane == 0 && i < 32 && ps < 4, 1024 + 16 * i + 11 + ps);
        const uint32_t t_addr = tmem_base + Cfg::kAcc2Col + b * 128 + side * kChunkN + lane_tmem;
        tmem_ld_32x32b_x32(t_addr, vlo);
        tmem_ld_32x32b_x32(t_addr + 32, vhi);
        __syncwarp();
        reinterpret_cast<float2*>(bias_gen)[lane] = bias2;
        const int rb = 2 * side + (int)(gs & 1u);  // residual buffer = store staging of this chunk
        uint8_t* const res_row = smem_raw + (res_base - smem_base) + rb * kChunkBytes + row_off;
        mbar_wait(res_full(rb), (gs >> 1) & 1u);
        __syncwarp();
        tmem_wait_ld_pin(vlo);
        tmem_wait_ld_pin(vhi);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc2_empty(b));
        epi_row<true, true>(v, b4, res_row, sw, o);
        // the store of the previous chunk has read its buffer by now: hand that buffer back
        if (lane == 0 && pending_buf >= 0) {
          bulk_wait_group_read<0>();
          mbar_arrive(res_empty(pending_buf));
        }
        // each thread overwrites the 128 bytes of residual it has just consumed
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(res_row + ((((uint32_t)j) ^ sw) << 4)) = o[j];
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (m0 + quad * 32 < p.M)
            tma_store_2d(&tmap_out, res_base + rb * kChunkBytes + quad * kSliceBytes, n0c, m0 + quad * 32);
          bulk_commit_group();
        }
        FT(ew == 0 && lane == 0 && i < 32 && ps == passes - 1, 1024 + 16 * i + 15);
        pending_buf = rb;
      }
    }
    if (lane == 0 && pending_buf >= 0) {
      bulk_wait_group_read<0>();
      mbar_arrive(res_empty(pending_buf));
    }
  } else if (warp == 3 + kEpiWarps) {
    // ------------------- residual warp --------------------------------------------------------
    if (lane == 0) {
      pdl_wait();
      const int total = my_tiles * passes;  // chunks per side
      auto coords = [&](int n, int sd, int& m0, int& n0c) {
        const int i = n / passes;
        m0 = tile_row(i);
        n0c = (n - i * passes) * 128 + sd * kChunkN;
      };
      int m0, n0c;
      for (int n = 0; n < total && n < 2; ++n)
        for (int sd = 0; sd < 2; ++sd) { coords(n, sd, m0, n0c); tma_prefetch_l2_2d(&tmap_res, n0c, m0); }
      for (int n = 0; n < total; ++n) {
        for (int sd = 0; sd < 2; ++sd) {
          const int rb = 2 * sd + (n & 1);
          // all four warps of the side hand the buffer back after their stores have read it
          if (n >= 2) mbar_wait(res_empty(rb), (uint32_t)((n >> 1) - 1) & 1u);
          coords(n, sd, m0, n0c);
          mbar_expect_tx(res_full(rb), kChunkBytes);
          tma_load_2d(res_base + rb * kChunkBytes, &tmap_res, res_full(rb), n0c, m0);
          if (n + 2 < total) { coords(n + 2, sd, m0, n0c); tma_prefetch_l2_2d(&tmap_res, n0c, m0); }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  FT(threadIdx.x == 0, 1603);
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ----------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------
int fused_tail_supported(int Cin, int Cmid, int Cout, int R, int S, int stride, int pad) {
  return igemm_supported(Cin, Cmid, R, S, stride, pad) && (Cmid == 64 || Cmid == 128) && Cout % 128 == 0 && Cout >= 128;
}

int fused_tail_make_plan(FusedPlan* plan, const void* in, int N, int H, int W, int Cin, const void* w2,
                         const float* bias2, int Cmid, int R, int S, int stride, int pad, const void* w3,
                         const float* bias3, int Cout, const void* residual, void* out) {
  AI4E_REQUIRE(fused_tail_supported(Cin, Cmid, Cout, R, S, stride, pad),
               "bottleneck tail: unsupported shape Cin=%d Cmid=%d Cout=%d %dx%d s%d p%d", Cin, Cmid, Cout, R, S, stride, pad);
  AI4E_REQUIRE(in && w2 && bias2 && w3 && bias3 && residual && out, "bottleneck tail: null pointer");
  const uintptr_t all = reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(w2) | reinterpret_cast<uintptr_t>(w3) |
                        reinterpret_cast<uintptr_t>(bias2) | reinterpret_cast<uintptr_t>(bias3) |
                        reinterpret_cast<uintptr_t>(residual) | reinterpret_cast<uintptr_t>(out);
  AI4E_REQUIRE((all & 15) == 0, "bottleneck tail: pointers must be 16-byte aligned");
  const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
  AI4E_REQUIRE(P > 0 && Q > 0, "bottleneck tail: empty output");
  AI4E_REQUIRE((long long)N * P * Q < (1ll << 31) - kBlockM, "bottleneck tail: M too large");
  FusedParams& q = plan->prm;
  q.M = 0; q.P = P; q.Q = Q; q.PQ = P * Q; q.Cmid = Cmid; q.Cout = Cout; q.S = S; q.stride = stride; q.pad = pad;
  q.cin_blocks = Cin / kBlockK; q.num_k_blocks = R * S * q.cin_blocks; q.num_tiles = 0; q.reverse = 0;
  // measured (B200, micro-batch 64, launch by launch, profiles/r02_k_mma_loop_ab.json): the four Cmid 128 tails of layer2
  // (four passes per tile) 184 -> 174 us with the passes interleaved, the three Cmid 64 tails of layer1 (two passes, six
  // stages) 178 -> 190 us: interleaved only where it pays
  q.interleave = Cmid == 128 ? 1 : 0;
  q.dbg = nullptr;
  if (const char* v = getenv("AI4E_B200_FUSED_INTERLEAVE")) q.interleave = atoi(v);  // 0 plain order, 1 / 2 interleaved (see fused_unit_slot)
  q.bias2 = bias2; q.bias3 = bias3;
  plan->max_batch = N;
  plan->grid_sms = 0;
  int rc = tmap_encode_im2col_nhwc(&plan->tmap_a, in, N, H, W, Cin, R, S, stride, pad, kBlockK, kBlockM);
  if (rc) return rc;
  const uint32_t es[2] = {1, 1};
  {
    const uint64_t k2 = (uint64_t)R * S * Cin;
    const uint64_t dims[2] = {k2, (uint64_t)Cmid};
    const uint64_t str[1] = {k2 * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)Cmid};
    rc = tmap_encode_tiled(&plan->tmap_w2, w2, 2, dims, str, box, es, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)Cmid, (uint64_t)Cout};
    const uint64_t str[1] = {(uint64_t)Cmid * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, 128};
    rc = tmap_encode_tiled(&plan->tmap_w3, w3, 2, dims, str, box, es, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    const uint64_t odims[2] = {(uint64_t)Cout, (uint64_t)N * P * Q};
    const uint64_t ostr[1] = {(uint64_t)Cout * 2};
    const uint32_t obox[2] = {(uint32_t)kChunkN, 32};
    rc = tmap_encode_tiled(&plan->tmap_out, out, 2, odims, ostr, obox, es, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    const uint32_t rbox[2] = {(uint32_t)kChunkN, (uint32_t)kBlockM};
    rc = tmap_encode_tiled(&plan->tmap_res, residual, 2, odims, ostr, rbox, es, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  return AI4E_OK;
}

template <int N1>
static int fused_launch_n1(const FusedPlan& plan, const FusedParams& prm, int grid, cudaStream_t stream) {
  static std::atomic<unsigned long long> attr_done{0};  // bit per device
  if (int rc = ensure_dyn_smem(bottleneck_tail_kernel<N1>, FusedCfg<N1>::kSmem, attr_done)) return rc;
  AI4E_CUDA_TRY(launch_pdl(bottleneck_tail_kernel<N1>, dim3(grid), dim3(kIgemmThreads), (size_t)FusedCfg<N1>::kSmem, stream,
                           plan.tmap_a, plan.tmap_w2, plan.tmap_w3, plan.tmap_out, plan.tmap_res, prm));
  return AI4E_OK;
}

int fused_tail_launch(const FusedPlan& plan, int batch, cudaStream_t stream) {
  AI4E_REQUIRE(batch > 0 && batch <= plan.max_batch, "bottleneck tail: batch %d outside plan (max %d)", batch,
               plan.max_batch);
  FusedParams prm = plan.prm;
  const long long M = (long long)batch * prm.PQ;
  prm.M = (int)M;
  prm.num_tiles = (int)((M + kBlockM - 1) / kBlockM);
  const int grid = std::min(prm.num_tiles, plan.grid_sms > 0 ? std::min(plan.grid_sms, sm_count()) : sm_count());
  if (prm.Cmid == 64) return fused_launch_n1<64>(plan, prm, grid, stream);
  return fused_launch_n1<128>(plan, prm, grid, stream);
}

}  // namespace ai4e
