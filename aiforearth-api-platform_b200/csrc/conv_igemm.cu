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
// Structure (one persistent CTA per SM, 384 threads, warp-specialised; tiles are dealt round-robin):
//   warp 0   A producer.  A tiles come straight from the NHWC activation tensor through an
//            IM2COL-mode tensor map: one `cp.async.bulk.tensor.4d...im2col` per K block
//            fetches 128 consecutive output pixels x 64 channels for filter tap (r, s), with
//            stride, padding (zero fill) and the wrap across rows / images done by the TMA
//            unit.  128-byte-swizzled shared memory, as many stages as fit (3-8), mbarrier
//            `full`/`empty` rings.  Waits for the previous layer (griddepcontrol.wait) first.
//   warp 10  B producer: 2-D tiles of the weight matrix into the same stages.  Weights are
//            constants, so it does NOT wait for the previous layer: the first stages of B fill
//            while that kernel drains (programmatic dependent launch).
//   warp 1   allocates TMEM and issues `tcgen05.mma.kind::f16` (M = 128, N = BLOCK_N, K = 16)
//            from one elected lane of the convergent warp; `tcgen05.commit` releases smem stages
//            and signals the epilogue.  Two accumulator stages in TMEM (2 x BLOCK_N fp32
//            columns) let the epilogue of tile i overlap the MMAs of tile i+1.
//   warps 2-9  epilogue: eight INDEPENDENT warps, no block-level barriers.  A warp reads only its
//            TMEM lane quadrant (warp % 4); the 64-channel chunks of the CTA's tiles alternate
//            between the two warps of a quadrant.  Per chunk: `tcgen05.ld` (thread = pixel row),
//            + bias (staged in 256 warp-private bytes of shared memory), + residual, ReLU, pack to
//            bf16 into the warp's own 4 KB swizzled staging slice, `cp.async.bulk.tensor` store of
//            32 rows x 64 channels.  HBM sees full 128-byte lines, rows past the tensor end are
//            clipped by hardware.  Arithmetic is compiled per (residual, ReLU) combination.
//   warp 11  residual loader: 128-row x 64-channel residual chunks by TMA, two buffers per epilogue
//            side released through `res_empty` barriers, and an L2 prefetch four chunks ahead.
//
// CTA-pair mode (kPair): the two CTAs of a 2-CTA cluster share one 256 x BLOCK_N tile through
// `tcgen05.mma.cta_group::2`: each loads its own 128 rows of A and HALF of the weight tile, the
// leader CTA issues the MMAs for both and commits (multicast) to the barriers of both; the TMA
// loads of both CTAs signal the leader's `full` barrier.  Fewer operand bytes per MMA is what
// matters: L2 -> shared-memory ingest is the K loop's bound (tools/l2_probe.cu).
//
// The tiling (BLOCK_N 64 / 128 / 256, single CTA or pair) comes from a cost model
// (igemm_pick_tiling) or, in the engine, from timing the candidates once per layer
// (igemm_autotune).  The stem (7x7, 3-4 input channels) has two modes of its own, below.
#include "common.cuh"
#include "kernels.h"
#include "igemm_tiles.cuh"

// Developer aid: -DAI4E_PIPE_TRACE makes CTA 0 record clock64 stamps of its pipeline into
// IgemmParams::dbg (dumped by ai4e_conv_layer when AI4E_B200_PIPE_TRACE=<file> is set).  Compiled out
// by default: even the dormant `if (p.dbg ...)` checks cost 5 % of the convolution time.
#ifdef AI4E_PIPE_TRACE
#define AI4E_TRACE(cond, slot) do { if (p.dbg && blockIdx.x == 0 && (cond)) p.dbg[(slot)] = clock64(); } while (0)
#else
#define AI4E_TRACE(cond, slot) do { } while (0)
#endif

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace ai4e {


// Stem (7x7 stride 2, 3/4 input channels -> 64): the NHWC4 image is read as an [H][W/2][8] grid
// of "super-pixels" (two adjacent pixels = 8 channels = 16 bytes, the smallest unit im2col-mode
// TMA can fetch).  Output pixel q needs pixels 2q-3 .. 2q+3 = super-pixels q-2 .. q+1, so the
// 7-tap row becomes 4 super-pixel taps (the first carries one zero-weight pixel).  Each K block
// is one filter row r: 4 taps x 8 channels = 32 K values, held as four un-swizzled
// [128 rows][16 bytes] slabs (the canonical no-swizzle K-major layout: 8x16-byte core
// matrices, LBO = slab size, SBO = 128 bytes).  K = 7 x 32 = 224 for 147 useful taps.
constexpr int kStemTaps = 4;
constexpr int kStemSlabA = kBlockM * 16;  // 2 KiB
constexpr int kStemKBlocks = 7;

// Stem, row mode (W % 16 == 0, at most 128 output columns): one tile = one output row.  The seven
// input rows a row of outputs needs are fetched whole, each by one tiled TMA load of full
// 128-byte lines (zero fill outside the image gives the padding), into plain row buffers where
// 16-byte unit u holds super-pixel u - 8.  Output pixel q, tap t reads unit q + t + 6, so the A
// operand of filter row r is that row buffer itself seen through an un-swizzled K-major
// descriptor whose core matrices OVERLAP: 16 bytes from one GEMM row to the next (rows of a core
// matrix), 16 bytes from one tap to the next (LBO), 128 bytes between 8-row groups (SBO).  The
// im2col expansion happens in the tensor core's address generator; shared memory and L2 see
// every input byte once per output row.  All 28 KiB of weights stay resident in shared memory.
constexpr int kRowsMaxGroups = 18;                       // 128-byte groups per row buffer (W <= 256)
constexpr int kRowsStage = 7 * kRowsMaxGroups * 128 + 256;  // 16 KiB incl. over-read slack
constexpr int kModeGeneric = 0, kModeStemIm2col = 1, kModeStemRows = 2;

// kPair: the CTAs of a 2-CTA cluster work on one 256-row tile with tcgen05 cta_group::2 - each
// brings its own 128 rows of A and HALF of the weight tile, the tensor cores of both SMs read both
// halves.  Operand bytes pulled from L2 per MMA drop from 16 KB + BLOCK_N * 128 to 16 KB +
// BLOCK_N * 64 per CTA, and that is the bound: with all 148 SMs streaming, L2 delivers ~73 bytes
// per clock per SM (tools/l2_probe.cu; 126 B/clk with one SM alone), i.e. 449 cycles for the 32 KB
// a 128 x 128 x 64 block needs against 268 cycles of MMA.
template <int BLOCK_N, int kMode = kModeGeneric, bool kPair = false>
struct IgemmCfg {
  static_assert(!kPair || (kMode == kModeGeneric && BLOCK_N >= 128), "pair mode: generic convolution, BLOCK_N >= 128");
  static constexpr bool kStem = kMode == kModeStemIm2col;
  static constexpr bool kRows = kMode == kModeStemRows;
  static constexpr int kCtas = kPair ? 2 : 1;
  static constexpr int kAStage = kStem ? kStemTaps * kStemSlabA : kRows ? kRowsStage : kABytes;
  static constexpr int kBBytes = kStem ? kStemTaps * BLOCK_N * 16 : kRows ? 0 : BLOCK_N / kCtas * kBlockK * 2;
  static constexpr int kStageBytes = kAStage + kBBytes;
  static constexpr int kResidentB = kRows ? kStemKBlocks * kStemTaps * BLOCK_N * 16 : 0;  // 28 KiB
  // The K loop is latency-bound (measured on B200: a pipeline slot takes ~3100 cycles to go round -
  // TMA latency ~2150, MMAs + barrier wake-ups ~570, issue ~370 - against 128-512 cycles of MMA work),
  // so throughput is proportional to the number of stages in flight: take every stage that fits.
  static constexpr int kMaxStages = 12;
  static constexpr int kTmemCols = 2 * BLOCK_N;  // 128 / 256 / 512: power of two >= 32
  // epilogue: a 4 KB output staging slice and 256 bytes of bias per epilogue warp, + two 16 KB residual
  // buffers per side when the layer has a residual
  static constexpr int kEpiFixed = kEpiWarps * (kSliceBytes + kChunkN * 4);
  __host__ __device__ static constexpr int epi_bytes(bool has_res) { return kEpiFixed + (has_res ? 4 * kChunkBytes : 0); }
  static constexpr int kBarBytes = (2 * kMaxStages + 4 + 8 + 2) * 8 + 16;  // full, empty, tmem_full/empty[2], res_full/empty[4], b_full, tmem ptr
  static constexpr int kSmemLimit = 227 * 1024;
  __host__ __device__ static constexpr int stages(bool has_res) {
    const int room = kSmemLimit - kResidentB - epi_bytes(has_res) - kBarBytes;
    const int n = room / kStageBytes;
    return n > kMaxStages ? kMaxStages : n;
  }
  __host__ __device__ static constexpr int smem_bytes(bool has_res) {
    return stages(has_res) * kStageBytes + kResidentB + epi_bytes(has_res) + kBarBytes;
  }
};

template <int BLOCK_N, int kMode, bool kPair>
__global__ void __launch_bounds__(kIgemmThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                  const IgemmParams p) {
  using Cfg = IgemmCfg<BLOCK_N, kMode, kPair>;
  constexpr bool kStem = Cfg::kStem;
  constexpr bool kRows = Cfg::kRows;
  constexpr int kCtas = Cfg::kCtas;
  // tiles are handed out to "units" (a CTA, or a CTA pair); in pair mode a tile is 256 rows and this
  // CTA owns rows [128 * rank, 128 * rank + 128) of it
  const uint32_t cta_rank = kPair ? cluster_ctarank() : 0u;
  const int unit = (int)blockIdx.x / kCtas;
  const int unit_stride = (int)gridDim.x / kCtas;
  const bool leader = cta_rank == 0;
  // Row-tile order.  The engine alternates it from layer to layer: a consumer that starts with the
  // rows its producer wrote last finds them still in L2 (write-back), which matters for the 103 MB
  // activations of layer1 that do not fit next to everything else.
  auto m_index = [&](int tile_m) { return p.reverse ? p.tiles_m - 1 - tile_m : tile_m; };
  const int kStages = Cfg::stages(p.residual != nullptr);  // runtime: depends on the epilogue's needs
  // SWIZZLE_128B operand tiles need 1024-byte alignment.  The kernel has no static shared memory, so
  // the dynamic window starts at the CTA's shared base; the budget above has no slack for rounding up.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  if (smem_base & 1023u) __trap();
  const uint32_t bres_base = smem_base + kStages * Cfg::kStageBytes;  // resident weights (row-mode stem)
  const uint32_t epi_base = bres_base + Cfg::kResidentB;              // [out0][out1][res0][res1]
  const uint32_t bar_base = epi_base + Cfg::epi_bytes(p.residual != nullptr);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  constexpr int kMS = Cfg::kMaxStages;
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMS + s); };
  auto tmem_full_bar = [&](int s) { return bar_base + 8u * (2 * kMS + s); };
  auto tmem_empty_bar = [&](int s) { return bar_base + 8u * (2 * kMS + 2 + s); };
  auto res_full_bar = [&](int s) { return bar_base + 8u * (2 * kMS + 4 + s); };
  auto res_empty_bar = [&](int s) { return bar_base + 8u * (2 * kMS + 8 + s); };  // res_full/empty: two per epilogue side
  const uint32_t b_full_bar = bar_base + 8u * (2 * kMS + 12);
  const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * kMS + 13);
  volatile uint32_t* tmem_ptr_gen =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_addr - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  AI4E_TRACE(threadIdx.x == 0, 1008);  // kernel entry

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      // generic: A and B come from two producers; pair: from two producers in each CTA, all
      // arriving on the leader's barrier
      mbar_init(full_bar(s), (kMode == kModeGeneric ? 2 : 1) * kCtas);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tmem_full_bar(s), 1);
      // one arrival per epilogue warp that reads the tile (pair: of both CTAs, on the leader's
      // barrier): a 64-wide tile is drained by one side
      mbar_init(tmem_empty_bar(s), (BLOCK_N == kChunkN ? 4 : kEpiWarps) * kCtas);
    }
    for (int s = 0; s < 4; ++s) {
      mbar_init(res_full_bar(s), 1);
      mbar_init(res_empty_bar(s), 4);
    }
    mbar_init(b_full_bar, 1);
    tma_prefetch_desc(&tmap_out);
    if (p.residual) tma_prefetch_desc(&tmap_res);
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (kPair) {
      tmem_alloc_pair(tmem_ptr_addr, Cfg::kTmemCols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_ptr_addr, Cfg::kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();  // the peer's barriers exist before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;
  AI4E_TRACE(threadIdx.x == 0, 1009);  // set-up done
  // everything above touched only this kernel's own shared memory / TMEM / parameters: it may
  // overlap the previous layer's tail.  From here on the roles that read the previous layer's
  // output or overwrite buffers it may still read (activation loads, residual loads, output
  // stores) first wait for it (griddepcontrol.wait is per thread); the weight producer does not -
  // weights are constants, so the first stages of B fill while the previous kernel drains.
  pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------- TMA producer -------------------------------
    // (whole warp convergent, one elected lane issues - see the MMA warp)
    {
      int stage = 0;
      uint32_t phase = 0;
      [[maybe_unused]] int ptrace = 0;
      if constexpr (kRows) {
        // weights: 7 rows x 4 taps of [BLOCK_N][16 B] slabs, once per CTA
        if (elect_one()) {
          mbar_expect_tx(b_full_bar, Cfg::kResidentB);
          for (int i = 0; i < kStemKBlocks * kStemTaps; ++i)
            tma_load_2d(bres_base + i * BLOCK_N * 16, &tmap_b, b_full_bar, i * 8, 0);
        }
        __syncwarp();
      }
      pdl_wait();
      AI4E_TRACE(threadIdx.x == 0, 1010);  // previous kernel finished
      if constexpr (kRows) {
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
          const int n_img = tile / p.P;
          const int y0 = (tile - n_img * p.P) * 2 - 3;
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t a_dst = smem_base + stage * Cfg::kStageBytes;
          if (elect_one()) {
            // one load for the seven rows (box 64 x groups x 7 x 1; row r lands at r * stem_rb): seven
            // separate loads cost the issuing thread more than the tile's MMAs.  Group -1 = super-pixels
            // -8..-1 and rows outside the image: zero padding by OOB fill
            mbar_expect_tx(full_bar(stage), 7u * p.stem_rb);
            tma_load_4d(a_dst, &tmap_a, full_bar(stage), 0, -1, y0, n_img);
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
      for (int tile = unit; !kRows && tile < p.num_tiles; tile += unit_stride) {
        const int tile_n = tile % p.tiles_n;
        const int tile_m = m_index(tile / p.tiles_n);
        const int m0 = (tile_m * kCtas + (int)cta_rank) * kBlockM;
        const int n_img = m0 / p.PQ;
        const int rem = m0 - n_img * p.PQ;
        const int p0 = rem / p.Q;
        const int q0 = rem - p0 * p.Q;
        // base pixel of the tile in input coordinates (stem: W in super-pixels, stride 1, pad 2)
        const int w_base = kStem ? q0 - 2 : q0 * p.stride - p.pad;
        const int h_base = p0 * p.stride - p.pad;
        [[maybe_unused]] int r = 0, s = 0, cb = 0;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          AI4E_TRACE(lane == 0 && ptrace < 192, ptrace);  // slot free
          const uint32_t a_dst = smem_base + stage * Cfg::kStageBytes;
          [[maybe_unused]] const uint32_t b_dst = a_dst + Cfg::kAStage;
          // generic mode: this warp brings A only; the B tile comes from the B-producer warp
          if (elect_one()) {
            if constexpr (kPair) mbar_expect_tx_leader(full_bar(stage), Cfg::kAStage);
            else mbar_expect_tx(full_bar(stage), kMode == kModeGeneric ? Cfg::kAStage : Cfg::kStageBytes);
            if constexpr (kStem) {
              // K block kb = filter row; one im2col load per super-pixel tap, one weight slab each
#pragma unroll
              for (int t = 0; t < kStemTaps; ++t) {
                tma_load_im2col_4d(a_dst + t * kStemSlabA, &tmap_a, full_bar(stage), 0, w_base, h_base, n_img,
                                   (uint16_t)t, (uint16_t)kb);
                tma_load_2d(b_dst + t * BLOCK_N * 16, &tmap_b, full_bar(stage), kb * 32 + t * 8, tile_n * BLOCK_N);
              }
            } else if constexpr (kPair) {
              tma_load_im2col_4d_pair(a_dst, &tmap_a, full_bar(stage), cb * kBlockK, w_base, h_base, n_img,
                                      (uint16_t)s, (uint16_t)r);
            } else {
              tma_load_im2col_4d(a_dst, &tmap_a, full_bar(stage), cb * kBlockK, w_base, h_base, n_img,
                                 (uint16_t)s, (uint16_t)r);
            }
          }
          __syncwarp();
          AI4E_TRACE(lane == 0 && ptrace < 192, 192 + ptrace); ++ptrace;  // A issued
          if constexpr (!kStem) {
            if (++cb == p.cin_blocks) { cb = 0; if (++s == p.S) { s = 0; ++r; } }
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 2 + kEpiWarps) {
    // ------------------------------- B (weights) producer, generic mode ----------
    // One thread needs ~190 cycles per TMA instruction and ~615 per stage for wait + expect_tx +
    // two loads (clock64 trace, B200): longer than a stage's MMAs.  Two threads halve that.
    if (kMode == kModeGeneric) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = unit; tile < p.num_tiles; tile += unit_stride) {
        // pair: this CTA brings weight rows [n0, n0 + BLOCK_N / 2) of the tile's BLOCK_N
        const int n0 = (tile % p.tiles_n) * BLOCK_N + (int)cta_rank * (BLOCK_N / kCtas);
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          if (elect_one()) {
            const uint32_t b_dst = smem_base + stage * Cfg::kStageBytes + Cfg::kAStage;
            if constexpr (kPair) {
              mbar_expect_tx_leader(full_bar(stage), Cfg::kBBytes);
              tma_load_2d_pair(b_dst, &tmap_b, full_bar(stage), kb * kBlockK, n0);
            } else {
              mbar_expect_tx(full_bar(stage), Cfg::kBBytes);
              tma_load_2d(b_dst, &tmap_b, full_bar(stage), kb * kBlockK, n0);
            }
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer ---------------------------------
    // The whole warp runs this loop convergently and one elected lane issues: every operand is then
    // provably warp-uniform and the compiler feeds tcgen05.mma straight from uniform registers.
    // (Under `if (lane == 0)` it wraps each MMA in a per-lane "waterfall" loop of ~12 instructions.)
    if (leader) {  // pair: the leader CTA issues for both
      constexpr uint32_t idesc = umma_idesc_bf16_f32(kBlockM * kCtas, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      [[maybe_unused]] int mtrace = 0;
      // generic mode: descriptor low word and full-barrier address of the current stage, advanced with it
      // (umma_desc_join in common.cuh: three adds per MMA instead of rebuilding both descriptors from `stage`)
      const uint32_t ring_lo0 = umma_desc_lo(smem_base);
      [[maybe_unused]] uint32_t ring_lo = ring_lo0, fbar = full_bar(0);
      constexpr uint32_t kALo = kABytes >> 4, kStageLo = Cfg::kStageBytes >> 4, kEmptyOff = 8u * kMS;
      if constexpr (kRows) mbar_wait(b_full_bar, 0);
      for (int tile = unit; tile < p.num_tiles; tile += unit_stride, ++it) {
        const uint32_t acc = it & 1u;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(tmem_empty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        if constexpr (kRows) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * Cfg::kStageBytes;
          if (elect_one()) {
#pragma unroll
            for (int r = 0; r < 7; ++r) {
#pragma unroll
              for (int h = 0; h < 2; ++h) {  // taps (2h, 2h+1) = one K = 16 step
                const uint64_t adesc = umma_smem_desc(a_addr + r * p.stem_rb + (6 + 2 * h) * 16, 16, 128, 0);
                const uint64_t bdesc =
                    umma_smem_desc(bres_base + (r * kStemTaps + 2 * h) * BLOCK_N * 16, BLOCK_N * 16, 128, 0);
                umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (r | h) != 0 ? 1u : 0u);
              }
            }
            umma_commit(empty_bar(stage));
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        for (int kb = 0; !kRows && kb < p.num_k_blocks; ++kb) {
          mbar_wait(kStem ? full_bar(stage) : fbar, phase);
          // (the two trace stores cost ~140 cycles per K block: traced periods overstate the loop)
          AI4E_TRACE(lane == 0 && mtrace < 192, 384 + mtrace);  // data seen
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * Cfg::kStageBytes;
          if (elect_one()) {
            if constexpr (kStem) {
              // un-swizzled slabs: one MMA (K = 16) spans two slabs; LBO = slab size, SBO = 128 bytes
#pragma unroll
              for (int k = 0; k < kStemTaps / 2; ++k) {
                const uint64_t adesc = umma_smem_desc(a_addr + 2 * k * kStemSlabA, kStemSlabA, 128, 0);
                const uint64_t bdesc =
                    umma_smem_desc(a_addr + Cfg::kAStage + 2 * k * BLOCK_N * 16, BLOCK_N * 16, 128, 0);
                umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
              }
            } else {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                // advancing 16 bf16 = 32 bytes along K inside the 128-byte swizzle row:
                // +2 in the (addr >> 4) start-address field
                const uint64_t adesc = umma_desc_join(ring_lo + 2u * k), bdesc = umma_desc_join(ring_lo + kALo + 2u * k);
                if constexpr (kPair) umma_bf16_ss_pair(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                else umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
              }
            }
            // smem stage reusable once these MMAs retire (pair: in both CTAs)
            const uint32_t ebar = kStem ? empty_bar(stage) : fbar + kEmptyOff;
            if constexpr (kPair) umma_commit_pair(ebar);
            else umma_commit(ebar);
          }
          __syncwarp();
          AI4E_TRACE(lane == 0 && mtrace < 192, 576 + mtrace); ++mtrace;  // MMAs issued
          ring_lo += kStageLo; fbar += 8u;
          if (++stage == kStages) { stage = 0; phase ^= 1u; ring_lo = ring_lo0; fbar = full_bar(0); }
        }
        if (elect_one()) {  // accumulator complete -> epilogue (pair: of both CTAs)
          if constexpr (kPair) umma_commit_pair(tmem_full_bar(acc));
          else umma_commit(tmem_full_bar(acc));
        }
        __syncwarp();
      }
      AI4E_TRACE(lane == 0, 1011);  // all MMAs issued
    }
  } else if (warp < 2 + kEpiWarps) {
    // ------------------------------- epilogue -----------------------------------
    // Eight independent warps, no block-level barriers.  A warp can read only its own TMEM lane
    // quadrant (warp id % 4), so quadrant and "side" (warp - 2) / 4 identify it; the 64-channel
    // chunks of the CTA's tiles, numbered q = tile_iteration * kChunks + c, alternate between the two
    // sides.  Each warp owns a 32-row x 64-channel output staging slice which it stores with its own
    // 4 KB TMA box; residual chunks (128 rows) arrive through the residual warp below.  Per chunk:
    // accumulator load, arithmetic into packed registers, wait until the previous store has read
    // the staging slice, STS, store.  Thread = one pixel row, 64 channels = one 128-byte swizzled row.
    // The bias of the chunk is staged in a warp-private 256 bytes of shared memory: with 227 KB of
    // the SM carved out as shared memory L1 is tiny, and sixteen serialised LDG.128 of the bias
    // were most of the ~850-cycle arithmetic phase (pipeline trace, B200).
    const int ew = warp - 2;
    const int side = ew >> 2;
    const int quad = warp & 3;
    const bool has_res = p.residual != nullptr;
    const bool relu = p.relu != 0;
    constexpr int kChunks = BLOCK_N / kChunkN;
    const uint32_t out_buf = epi_base + ew * kSliceBytes;                                   // [out x8]
    uint8_t* const out_gen = smem_raw + (out_buf - smem_u32(smem_raw));
    float* const bias_gen = reinterpret_cast<float*>(smem_raw + (epi_base + kEpiWarps * kSliceBytes - smem_u32(smem_raw))) + ew * kChunkN;
    const uint8_t* const res_gen0 =                                                         // [res x2] per side
        smem_raw + (epi_base + Cfg::kEpiFixed + (2 * side) * kChunkBytes - smem_u32(smem_raw)) + quad * kSliceBytes;
    const uint32_t sw = (uint32_t)(lane & 7);
    const uint32_t row_off = (uint32_t)lane * 128u;
    // A warp's chunks: q = side, side + 2, ... over the CTA's tiles (kChunks chunks each).  `step`
    // advances by two chunks without divisions (they cost ~290 cycles per chunk in the trace).
    struct Chunk { int tile, c, m0, n0c, tile_m, tile_n; uint32_t it; };
    const int stride_m = unit_stride / p.tiles_n, stride_n = unit_stride - stride_m * p.tiles_n;
    auto place = [&](Chunk& k) -> bool {  // fills the coordinates from (tile, tile_m, tile_n, c)
      if (k.tile >= p.num_tiles) return false;
      k.m0 = (kRows ? 0 : (m_index(k.tile_m) * kCtas + (int)cta_rank) * kBlockM) + quad * 32;
      k.n0c = (kRows ? 0 : k.tile_n * BLOCK_N) + k.c * kChunkN;
      return true;
    };
    auto next_tile = [&](Chunk& k) {
      ++k.it;
      k.tile += unit_stride;
      k.tile_m += stride_m;
      k.tile_n += stride_n;
      if (k.tile_n >= p.tiles_n) { k.tile_n -= p.tiles_n; ++k.tile_m; }
    };
    auto step = [&](const Chunk& from, Chunk& k) -> bool {
      k = from;
      k.c += 2;
      while (k.c >= kChunks) { k.c -= kChunks; next_tile(k); }  // at most twice (kChunks = 1)
      return place(k);
    };
    auto first = [&](Chunk& k) -> bool {
      k.it = 0; k.tile = unit; k.tile_m = unit / p.tiles_n; k.tile_n = unit - k.tile_m * p.tiles_n;
      k.c = side;
      while (k.c >= kChunks) { k.c -= kChunks; next_tile(k); }
      return place(k);
    };
    uint32_t v[64];
    uint32_t (&vlo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[0]);
    uint32_t (&vhi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[32]);
    auto issue_ld = [&](const Chunk& k) {
      const uint32_t t_addr = tmem_base + (k.it & 1u) * BLOCK_N + k.c * kChunkN + ((uint32_t)(quad * 32) << 16);
      tmem_ld_32x32b_x32(t_addr, vlo);
      tmem_ld_32x32b_x32(t_addr + 32, vhi);
    };
    pdl_wait();
    Chunk cur, nxt;
    bool have = first(cur);
    bool ld_issued = false;  // the accumulator load of `cur` was already issued at the end of the previous chunk
    float2 bias2 = make_float2(0.f, 0.f);
    if (have) bias2 = __ldg(reinterpret_cast<const float2*>(p.bias + cur.n0c) + lane);
    uint32_t gs = 0;  // chunks this warp has processed: selects the residual buffer and phase
    for (uint32_t q = side; have; q += 2, ++gs) {
      const uint32_t acc = cur.it & 1u;
      [[maybe_unused]] const bool etr = ew == 0 && lane == 0 && gs < 30;
      if (!ld_issued) {
        // first chunk of this tile for this warp: the tile's accumulator must be complete
        if (cur.c < 2) {
          mbar_wait(tmem_full_bar(acc), (cur.it >> 1) & 1u);
          tc_fence_after();
        }
        issue_ld(cur);
      }
      AI4E_TRACE(etr, 768 + 8 * gs + 0);  // chunk start (accumulator load issued)
      reinterpret_cast<float2*>(bias_gen)[lane] = bias2;  // (the previous chunk's readers passed a __syncwarp)
      const uint8_t* const res_row = res_gen0 + (gs & 1u) * kChunkBytes + row_off;
      if (has_res) mbar_wait(res_full_bar(2 * side + (gs & 1u)), (gs >> 1) & 1u);
      __syncwarp();
      const float4* const b4 = reinterpret_cast<const float4*>(bias_gen);
      tmem_wait_ld_pin(vlo);
      tmem_wait_ld_pin(vhi);
      AI4E_TRACE(etr, 768 + 8 * gs + 1);  // accumulators in registers, residual landed
      // last chunk of this tile for this warp: its TMEM reads are done, free the accumulator stage
      if (cur.c + 2 >= kChunks) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (kPair) mbar_arrive_leader(tmem_empty_bar(acc));
          else mbar_arrive(tmem_empty_bar(acc));
        }
      }
      uint4 o[8];
      if (has_res) {
        if (relu) epi_row<true, true>(v, b4, res_row, sw, o);
        else epi_row<true, false>(v, b4, res_row, sw, o);
      } else {
        if (relu) epi_row<false, true>(v, b4, nullptr, sw, o);
        else epi_row<false, false>(v, b4, nullptr, sw, o);
      }
      AI4E_TRACE(etr, 768 + 8 * gs + 2);  // arithmetic done
      // this warp's next chunk: fetch its bias and, if its accumulator is already complete, start
      // the TMEM load now so that it travels during the staging and the store below
      have = step(cur, nxt);
      ld_issued = false;
      AI4E_TRACE(etr, 768 + 8 * gs + 3);  // next chunk located
      if (have) {
        bias2 = __ldg(reinterpret_cast<const float2*>(p.bias + nxt.n0c) + lane);
        bool ready = true;
        if (nxt.c < 2) {
          ready = __all_sync(0xffffffffu, mbar_test_wait(tmem_full_bar(nxt.it & 1u), (nxt.it >> 1) & 1u));
          if (ready) tc_fence_after();
        }
        if (ready) {
          issue_ld(nxt);
          ld_issued = true;
        }
      }
      AI4E_TRACE(etr, 768 + 8 * gs + 4);  // next accumulator load issued (or not ready)
      // the TMA store of this warp's previous chunk must have read the staging slice by now
      if (lane == 0) bulk_wait_group_read<0>();
      __syncwarp();
      if (has_res && lane == 0) mbar_arrive(res_empty_bar(2 * side + (gs & 1u)));  // this warp is done with it
      AI4E_TRACE(etr, 768 + 8 * gs + 5);  // staging slice free
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<uint4*>(out_gen + row_off + ((((uint32_t)j) ^ sw) << 4)) = o[j];
      fence_proxy_async_smem();  // generic-proxy writes -> visible to the TMA (async proxy)
      __syncwarp();
      AI4E_TRACE(etr, 768 + 8 * gs + 6);  // staged
      if (lane == 0) {
        if constexpr (kRows) {
          if (quad * 32 < p.Q) tma_store_3d(&tmap_out, out_buf, 0, quad * 32, cur.tile);  // row `tile` of [N*P][Q][64]
        } else {
          if (cur.m0 < p.M) tma_store_2d(&tmap_out, out_buf, cur.n0c, cur.m0);
        }
        bulk_commit_group();
      }
      AI4E_TRACE(etr, 768 + 8 * gs + 7);  // store issued
      cur = nxt;
    }
    if (lane == 0) bulk_wait_group_read<0>();  // shared memory must outlive the last TMA store
    AI4E_TRACE(ew == 0 && lane == 0, 1012);  // epilogue done
  } else if (warp == 3 + kEpiWarps) {
    // ------------------------------- residual warp -------------------------------
    // One thread walks the CTA's chunks in order and keeps two 128-row x 64-channel residual
    // buffers per epilogue side filled; the four warps of a side release a buffer through
    // res_empty (count 4).  The residual was written three layers ago and has usually left L2,
    // so the chunk four ahead is also pulled into L2.  (Issued from the epilogue warps these
    // TMA instructions cost each of them ~600 cycles per chunk.)
    if (p.residual != nullptr && lane == 0) {
      pdl_wait();
      constexpr int kChunks = BLOCK_N / kChunkN;
      auto coords = [&](uint32_t q, int& m0, int& n0c) -> bool {
        const int tile = unit + (int)(q / kChunks) * unit_stride;
        if (tile >= p.num_tiles) return false;
        m0 = (m_index(tile / p.tiles_n) * kCtas + (int)cta_rank) * kBlockM;
        n0c = (tile % p.tiles_n) * BLOCK_N + (int)(q % kChunks) * kChunkN;
        return true;
      };
      int m0, n0c;
      for (uint32_t q = 0; q < 4; ++q)
        if (coords(q, m0, n0c)) tma_prefetch_l2_2d(&tmap_res, n0c, m0);
      for (uint32_t q = 0;; ++q) {
        if (!coords(q, m0, n0c)) break;
        const uint32_t sd = q & 1u, n = q >> 1, b = n & 1u;  // side, side-chunk index, buffer
        if (n >= 2) mbar_wait(res_empty_bar(2 * sd + b), ((n >> 1) - 1u) & 1u);
        mbar_expect_tx(res_full_bar(2 * sd + b), kChunkBytes);
        tma_load_2d(epi_base + Cfg::kEpiFixed + (2 * sd + b) * kChunkBytes, &tmap_res, res_full_bar(2 * sd + b), n0c, m0);
        if (coords(q + 4, m0, n0c)) tma_prefetch_l2_2d(&tmap_res, n0c, m0);
      }
    }
  }

  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();  // neither CTA leaves (or frees TMEM) while the other may still touch it
  else __syncthreads();
  AI4E_TRACE(threadIdx.x == 0, 1013);  // every warp done
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    if constexpr (kPair) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
    else tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ----------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------
// Cost model behind the automatic tiling (cycles of the slowest CTA), from measured constants:
//  * K loop: one 64-deep block costs max(MMA time, operand bytes / what L2 can feed, TMA latency /
//    stages in flight).  MMA: 4 x 67 cycles up to N = 128, 512 at N = 256.  L2 -> SM: 126 B/clk for
//    an SM alone, ~10.9 KB/clk for the whole GPU (tools/l2_probe.cu).  TMA latency ~2600 cycles.
//  * epilogue: ~1050 cycles per 64-channel chunk (1250 with a residual), overlapped with the next
//    tile's K loop, so a tile costs the longer of the two.
//  * tiles are dealt round-robin to min(tiles, SMs) CTAs (or CTA pairs).
static double igemm_model_cycles(int Cout, long long M, int num_k_blocks, bool has_res, int bn, bool pair, int sms) {
  const int ctas = pair ? 2 : 1;
  const long long tiles = ((M + kBlockM * ctas - 1) / (kBlockM * ctas)) * (Cout / bn);
  const long long units = std::min<long long>(tiles, sms / ctas);
  const double active = (double)units * ctas;
  const double per_unit = (double)((tiles + units - 1) / units);
  const double feed = std::min(120.0, 10900.0 / active);
  const double bytes_kb = 16384.0 + bn * 128.0 / ctas;
  const double mma = bn <= 128 ? 268.0 : 512.0;
  const double room = 232448.0 - 34816.0 - (has_res ? 65536.0 : 0.0) - 300.0;
  const double stages = std::max(1.0, std::min(12.0, std::floor(room / bytes_kb)));
  const double kb = std::max({bytes_kb / feed, mma, 2600.0 / stages}) + (pair ? 40.0 : 25.0);
  const double kloop = num_k_blocks * kb;
  const double epi = (bn / 64) * (has_res ? 1250.0 : 1050.0);
  return per_unit * std::max(kloop, epi) + std::min(kloop, epi) + 2500.0 + (pair ? 1500.0 : 0.0);
}

int igemm_tiling_candidates(int Cout, int out[8]) {
  int n = 0;
  for (int bn : {64, 128, 256}) {
    if (Cout % bn) continue;
    out[n++] = bn;
    if (bn >= 128) out[n++] = bn + kTilingPair;
  }
  return n;
}

int igemm_pick_tiling(int Cout, long long M, int num_k_blocks, bool has_res, int sms) {
  if (sms <= 0) sms = sm_count();
  int cand[8];
  const int n = igemm_tiling_candidates(Cout, cand);
  int best = 0;
  double best_c = 0.0;
  for (int i = 0; i < n; ++i) {
    const bool pair = cand[i] >= kTilingPair;
    const int bn = pair ? cand[i] - kTilingPair : cand[i];
    // ties go to the later candidate only when it is clearly better: single-CTA and narrower first
    const double c = igemm_model_cycles(Cout, M, num_k_blocks, has_res, bn, pair, sms) * (pair ? 1.02 : 1.0);
    if (best == 0 || c < best_c * 0.985) { best = cand[i]; best_c = c; }
  }
  return best;
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
  plan->stem_mode = kModeGeneric;
  plan->tuned_tiling = plan->tuned_batch = 0;
  plan->grid_sms = 0;
  q.stem_rb = 0; q.dbg = nullptr; q.reverse = 0;
  int rc = tmap_encode_im2col_nhwc(&plan->tmap_a, in, N, H, W, Cin, R, S, stride, pad, kBlockK, kBlockM);
  if (rc) return rc;
  const uint64_t ktot = (uint64_t)R * S * Cin;
  const uint64_t dims[2] = {ktot, (uint64_t)Cout};
  const uint64_t strides[1] = {ktot * 2};
  const uint32_t es[2] = {1, 1};
  // output / residual: [M rows][Cout] bf16, 32 rows x 64 channels per TMA (one epilogue warp), 128B swizzle
  {
    const uint64_t odims[2] = {(uint64_t)Cout, (uint64_t)N * P * Q};
    const uint64_t ostr[1] = {(uint64_t)Cout * 2};
    const uint32_t obox[2] = {(uint32_t)kChunkN, 32};
    rc = tmap_encode_tiled(&plan->tmap_out, out, 2, odims, ostr, obox, es, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    const uint32_t rbox[2] = {(uint32_t)kChunkN, (uint32_t)kBlockM};  // residual: a whole 128-row chunk per TMA
    rc = tmap_encode_tiled(&plan->tmap_res, residual ? residual : out, 2, odims, ostr, rbox, es,
                           CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
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

int stem_supported(int H, int W, int Cin, int Cout, int R, int S, int stride, int pad) {
  return Cin == 4 && Cout == 64 && R == 7 && S == 7 && stride == 2 && pad == 3 && W % 2 == 0 && H >= 7 && W >= 8;
}

// [64][7][7][4] KRSC bf16 weights -> [64][7][4 taps][8] "super-pixel" weights (see kStemTaps):
// tap t, channel c < 4  <- filter column 2t-1 (zero for t = 0); channel 4+c <- filter column 2t.
void stem_repack_weights(const uint16_t* w_krsc, uint16_t* w_sp) {
  for (int co = 0; co < 64; ++co)
    for (int r = 0; r < 7; ++r)
      for (int t = 0; t < kStemTaps; ++t)
        for (int c = 0; c < 8; ++c) {
          const int s = (c < 4) ? 2 * t - 1 : 2 * t;
          const int ch = c & 3;
          w_sp[((co * 7 + r) * kStemTaps + t) * 8 + c] = (s < 0) ? (uint16_t)0 : w_krsc[((co * 7 + r) * 7 + s) * 4 + ch];
        }
}

// `weight_sp` is the repacked [64][224] device tensor.  Output geometry as for any 7x7/2 pad 3.
// Row mode when W % 16 == 0 and the output row fits one tile (W <= 256), else im2col mode.
int igemm_make_plan_stem(IgemmPlan* plan, const void* in, int N, int H, int W, const void* weight_sp,
                         const float* bias, int relu, void* out) {
  AI4E_REQUIRE(stem_supported(H, W, 4, 64, 7, 7, 2, 3), "conv(stem): unsupported input %dx%d", H, W);
  AI4E_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(weight_sp) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(out) & 15) == 0 && (reinterpret_cast<uintptr_t>(bias) & 15) == 0,
               "conv(stem): pointers must be 16-byte aligned");
  const int P = (H + 6 - 7) / 2 + 1, Q = (W + 6 - 7) / 2 + 1;
  AI4E_REQUIRE((long long)N * P * Q < (1ll << 31) - kBlockM, "conv(stem): M too large");
  IgemmParams& q = plan->prm;
  q.M = 0; q.P = P; q.Q = Q; q.PQ = P * Q; q.Cout = 64; q.S = kStemTaps; q.stride = 2; q.pad = 3;
  q.cin_blocks = 1; q.num_k_blocks = kStemKBlocks; q.stem_rb = 0; q.dbg = nullptr; q.reverse = 0;
  q.tiles_m = q.tiles_n = q.num_tiles = 0; q.relu = relu; q.bias = bias; q.residual = nullptr; q.out = (bf16*)out;
  plan->max_batch = N;
  const bool rows = (W % 16 == 0) && Q <= kBlockM && !getenv("AI4E_B200_STEM_IM2COL");
  plan->stem_mode = rows ? kModeStemRows : kModeStemIm2col;
  plan->tuned_tiling = plan->tuned_batch = 0;
  plan->grid_sms = 0;
  const uint32_t es4[4] = {1, 1, 1, 1};
  int rc;
  if (rows) {
    // input rows as [N][H][W/16 groups][64 elements]; the box is one row, `groups` x 128 bytes,
    // starting one group left of the image (zero fill) so that unit u holds super-pixel u - 8
    const int groups = (Q + 10 + 7) / 8;  // units 0 .. Q-1 + 3 + 6 must exist
    AI4E_REQUIRE(groups <= kRowsMaxGroups, "conv(stem): row too wide");
    q.stem_rb = groups * 128;
    const uint64_t adims[4] = {64, (uint64_t)W / 16, (uint64_t)H, (uint64_t)N};
    const uint64_t astr[3] = {128, (uint64_t)W * 8, (uint64_t)H * W * 8};
    const uint32_t abox[4] = {64, (uint32_t)groups, 7, 1};
    rc = tmap_encode_tiled(&plan->tmap_a, in, 4, adims, astr, abox, es4, CU_TENSOR_MAP_SWIZZLE_NONE);
  } else {
    // input as [N][H][W/2][8]: 4 taps from super-pixel q-2, traversal stride 1 in W and 2 in H.
    // lower corner {-2, -3}; upper corner = right/bottom padding - (taps - 1) = {1 - 3, 3 - 6}.
    const int lower[2] = {-2, -3}, upper[2] = {-2, -3}, st[2] = {1, 2};
    rc = tmap_encode_im2col(&plan->tmap_a, in, N, H, W / 2, 8, lower, upper, st, 8, kBlockM,
                            CU_TENSOR_MAP_SWIZZLE_NONE);
  }
  if (rc) return rc;
  const uint32_t es[2] = {1, 1};
  const uint64_t wdims[2] = {(uint64_t)kStemKBlocks * 32, 64};
  const uint64_t wstr[1] = {(uint64_t)kStemKBlocks * 32 * 2};
  const uint32_t wbox[2] = {8, 64};
  for (int i = 0; i < 3; ++i) plan->has_bn[i] = false;
  plan->has_bn[0] = true;
  rc = tmap_encode_tiled(&plan->tmap_b[0], weight_sp, 2, wdims, wstr, wbox, es, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc) return rc;
  if (rows) {
    const uint64_t odims[3] = {64, (uint64_t)Q, (uint64_t)N * P};
    const uint64_t ostr[2] = {64 * 2, (uint64_t)Q * 64 * 2};
    const uint32_t obox[3] = {(uint32_t)kChunkN, 32, 1};
    rc = tmap_encode_tiled(&plan->tmap_out, out, 3, odims, ostr, obox, es4, CU_TENSOR_MAP_SWIZZLE_128B);
  } else {
    const uint64_t odims[2] = {64, (uint64_t)N * P * Q};
    const uint64_t ostr[1] = {64 * 2};
    const uint32_t obox[2] = {(uint32_t)kChunkN, 32};
    rc = tmap_encode_tiled(&plan->tmap_out, out, 2, odims, ostr, obox, es, CU_TENSOR_MAP_SWIZZLE_128B);
  }
  if (rc) return rc;
  plan->tmap_res = plan->tmap_out;
  return AI4E_OK;
}

template <int BN, int kMode, bool kPair = false>
static int igemm_launch_bn(const IgemmPlan& plan, const CUtensorMap& tb, const IgemmParams& prm, int grid,
                           cudaStream_t stream) {
  static std::atomic<unsigned long long> attr_done{0};  // bit per device
  if (int rc = ensure_dyn_smem(conv_igemm_kernel<BN, kMode, kPair>, IgemmCfg<BN, kMode, kPair>::kSmemLimit, attr_done))
    return rc;
  AI4E_CUDA_TRY(launch_pdl_cluster(conv_igemm_kernel<BN, kMode, kPair>, dim3(grid), dim3(kIgemmThreads),
                                   (size_t)IgemmCfg<BN, kMode, kPair>::smem_bytes(prm.residual != nullptr), stream,
                                   kPair ? 2 : 1, plan.tmap_a, tb, plan.tmap_out, plan.tmap_res, prm));
  return AI4E_OK;
}

// `tiling`: 0 = automatic; otherwise BLOCK_N (64 / 128 / 256), + kTilingPair for the CTA-pair kernel.
int igemm_launch(const IgemmPlan& plan, int batch, int tiling, cudaStream_t stream) {
  AI4E_REQUIRE(batch > 0 && batch <= plan.max_batch, "conv(igemm): batch %d outside plan (max %d)", batch,
               plan.max_batch);
  IgemmParams prm = plan.prm;
  const long long M = (long long)batch * prm.PQ;
  // CTAs of a persistent grid: every SM, or fewer when the engine leaves room for its other lane
  const int sms = plan.grid_sms > 0 ? std::min(plan.grid_sms, sm_count()) : sm_count();
  if (plan.stem_mode == kModeStemRows) {
    prm.M = (int)M;
    prm.tiles_m = prm.num_tiles = batch * prm.P;  // one tile per output row
    prm.tiles_n = 1;
    return igemm_launch_bn<64, kModeStemRows>(plan, plan.tmap_b[0], prm, std::min(prm.num_tiles, sms), stream);
  }
  if (plan.stem_mode == kModeStemIm2col) {
    prm.M = (int)M;
    prm.tiles_m = (int)((M + kBlockM - 1) / kBlockM);
    prm.tiles_n = 1;
    prm.num_tiles = prm.tiles_m;
    return igemm_launch_bn<64, kModeStemIm2col>(plan, plan.tmap_b[0], prm, std::min(prm.num_tiles, sms),
                                                stream);
  }
  if (tiling == 0) {
    if (const char* v = getenv("AI4E_B200_TILE")) {  // developer aid, e.g. "128" or "256p"
      tiling = atoi(v);
      if (strchr(v, 'p')) tiling += kTilingPair;
    }
  }
  if (tiling == 0 && plan.tuned_tiling && batch == plan.tuned_batch) tiling = plan.tuned_tiling;
  if (tiling == 0) tiling = igemm_pick_tiling(prm.Cout, M, prm.num_k_blocks, prm.residual != nullptr, sms);
  const bool pair = tiling >= kTilingPair;
  const int bn = pair ? tiling - kTilingPair : tiling;
  const int bi = bn == 64 ? 0 : bn == 128 ? 1 : bn == 256 ? 2 : -1;
  AI4E_REQUIRE(bi >= 0 && plan.has_bn[bi] && (!pair || bi > 0), "conv(igemm): tiling %d not available for Cout %d", tiling,
               prm.Cout);
  prm.M = (int)M;
  const int rows = kBlockM * (pair ? 2 : 1);
  prm.tiles_m = (int)((M + rows - 1) / rows);
  prm.tiles_n = prm.Cout / bn;
  prm.num_tiles = prm.tiles_m * prm.tiles_n;
  if (pair) {
    // a pair tile's halves are loaded with the weight map whose box is BLOCK_N / 2 rows
    const int grid = 2 * std::min(prm.num_tiles, sms / 2);
    if (bn == 128) return igemm_launch_bn<128, kModeGeneric, true>(plan, plan.tmap_b[bi - 1], prm, grid, stream);
    return igemm_launch_bn<256, kModeGeneric, true>(plan, plan.tmap_b[bi - 1], prm, grid, stream);
  }
  const int grid = std::min(prm.num_tiles, sms);
  switch (bn) {
    case 64: return igemm_launch_bn<64, kModeGeneric>(plan, plan.tmap_b[bi], prm, grid, stream);
    case 128: return igemm_launch_bn<128, kModeGeneric>(plan, plan.tmap_b[bi], prm, grid, stream);
    default: return igemm_launch_bn<256, kModeGeneric>(plan, plan.tmap_b[bi], prm, grid, stream);
  }
}

// Times every tiling of this layer on the plan's own buffers (their contents do not matter, the
// work is data-independent) and keeps the fastest for launches with this batch.
int igemm_autotune(IgemmPlan* plan, int batch, cudaStream_t stream, float* best_us_out) {
  if (plan->stem_mode != kModeGeneric) return AI4E_OK;
  int cand[8];
  const int n = igemm_tiling_candidates(plan->prm.Cout, cand);
  cudaEvent_t e0, e1;
  AI4E_CUDA_TRY(cudaEventCreate(&e0));
  AI4E_CUDA_TRY(cudaEventCreate(&e1));
  constexpr int kReps = 4;
  int best = 0, rc = AI4E_OK;
  float best_ms = 0.f;
  for (int i = 0; i < n && rc == AI4E_OK; ++i) {
    for (int pass = 0; pass < 2 && rc == AI4E_OK; ++pass) {  // pass 0 warms up (attributes, instruction cache)
      cudaEventRecord(e0, stream);
      for (int r = 0; r < (pass ? kReps : 1) && rc == AI4E_OK; ++r) rc = igemm_launch(*plan, batch, cand[i], stream);
      cudaEventRecord(e1, stream);
      if (cudaEventSynchronize(e1) != cudaSuccess) rc = AI4E_ERR_CUDA;
    }
    float ms = 0.f;
    if (rc == AI4E_OK && cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess && (best == 0 || ms < best_ms)) {
      best = cand[i];
      best_ms = ms;
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (rc != AI4E_OK) return rc;
  plan->tuned_tiling = best;
  plan->tuned_batch = batch;
  if (best_us_out) *best_us_out = best_ms * 1e3f / kReps;
  return AI4E_OK;
}

}  // namespace ai4e
