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

#ifdef __CUDACC__
// One 64-channel row of an accumulator chunk: acc + bias (+ residual) (ReLU) -> eight 16-byte units
// of packed bf16.  `v` = the 64 fp32 accumulators of the thread's row (tcgen05.ld 32x32b), `b4` = the
// chunk's bias in shared memory, `res_row` = the thread's 128-byte row of the swizzled residual
// tile, `sw` = row & 7.  kRes / kRelu are compile-time: with run-time flags every 8-channel unit
// became its own basic block (16 branches per row) and the arithmetic ran at 0.35 IPC.
// Rounding is monotonic and keeps the sign, so ReLU on the packed bf16 pair equals ReLU before it.
template <bool kRes, bool kRelu>
__device__ __forceinline__ void epi_row(const uint32_t (&v)[64], const float4* b4, const uint8_t* res_row, uint32_t sw,
                                        uint4 (&o)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 ba = b4[2 * j], bb = b4[2 * j + 1];
    float x[8];
    x[0] = __uint_as_float(v[8 * j + 0]) + ba.x; x[1] = __uint_as_float(v[8 * j + 1]) + ba.y;
    x[2] = __uint_as_float(v[8 * j + 2]) + ba.z; x[3] = __uint_as_float(v[8 * j + 3]) + ba.w;
    x[4] = __uint_as_float(v[8 * j + 4]) + bb.x; x[5] = __uint_as_float(v[8 * j + 5]) + bb.y;
    x[6] = __uint_as_float(v[8 * j + 6]) + bb.z; x[7] = __uint_as_float(v[8 * j + 7]) + bb.w;
    if constexpr (kRes) {
      const uint4 r = *reinterpret_cast<const uint4*>(res_row + ((((uint32_t)j) ^ sw) << 4));
      x[0] += bf16_lo(r.x); x[1] += bf16_hi(r.x); x[2] += bf16_lo(r.y); x[3] += bf16_hi(r.y);
      x[4] += bf16_lo(r.z); x[5] += bf16_hi(r.z); x[6] += bf16_lo(r.w); x[7] += bf16_hi(r.w);
    }
    o[j] = make_uint4(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
    if constexpr (kRelu) {
      o[j].x = relu_bf16x2(o[j].x); o[j].y = relu_bf16x2(o[j].y);
      o[j].z = relu_bf16x2(o[j].z); o[j].w = relu_bf16x2(o[j].w);
    }
  }
}
#endif

}  // namespace ai4e
