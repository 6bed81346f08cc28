// Tile geometry shared by the tcgen05 convolution kernels (conv_igemm.cu, conv_fused.cu).
#pragma once

namespace ai4e {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                               // 64 bf16 = 128 bytes = one swizzle row
constexpr int kEpiWarps = 8;                              // two warps per TMEM lane quadrant
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kIgemmThreads = 64 + kEpiThreads + 64;      // + A-producer warp, MMA warp, B-producer warp, residual warp
constexpr int kABytes = kBlockM * kBlockK * 2;            // 16 KiB
constexpr int kChunkN = 64;                               // epilogue works on 64 channels at a time
constexpr int kChunkBytes = kBlockM * kChunkN * 2;        // 16 KiB: 128 rows x 128 bytes, swizzled
constexpr int kSliceBytes = 32 * kChunkN * 2;             // 4 KiB: one epilogue warp's 32 rows of a chunk

}  // namespace ai4e
