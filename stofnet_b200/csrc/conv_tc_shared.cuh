// Constants and device helpers shared by the chain kernels (conv_tc.cu: one CTA per tile; conv_tc2.cu: CTA pairs with
// cta_group::2 MMAs): the shared-memory operand geometry and the fp32 <-> fp16 hi/lo split.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace stof {

constexpr int TC_THREADS = 352;                         // 1 producer + 2 MMA issuers + 8 epilogue warps
constexpr int TC_EPI_WARPS = 8;
constexpr int TC_R = 280;                              // smem rows per (plane, c8) slab
constexpr int TC_SLAB = TC_R * 16;                     // bytes of one (plane, c8) slab
constexpr int TC_PLANE = 8 * TC_SLAB;                  // one fp16 plane (hi or lo): 64 channels
constexpr int TC_BUF = 2 * TC_PLANE;                   // hi + lo
constexpr int TC_STAGES = 5;                          // weight ring of the fp32-faithful mode: 5 x 16 KB taps
constexpr int TC_STAGE_BYTES = 16384;                  // one k7 tap: [8 c8][128 n][8] fp16
constexpr int TC_MAX_STAGES = 10;                     // one-pass fp16 mode: 10 x 8 KB (W_hi only) in the same 80 KB
constexpr float TC_LO_SCALE = 2048.0f;                 // 2^11
constexpr float TC_LO_INV = 1.0f / 2048.0f;

__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

// two fp32 -> packed fp16x2 (round to nearest, saturated to the finite fp16 range); a in the low half
__device__ __forceinline__ uint32_t cvt_sat_half2(float a, float b) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
// (v0, v1) -> packed hi = fp16(v), packed lo = fp16((v - hi) * 2^11); the subtraction and the scaling run as packed fp32 ops
// (both steps are exact, so the result equals the scalar form bit for bit)
__device__ __forceinline__ void split_pair(float v0, float v1, uint32_t &hi, uint32_t &lo) {
    hi = cvt_sat_half2(v0, v1);
    const float2 h = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    const float2 d = __fmul2_rn(__ffma2_rn(h, make_float2(-1.0f, -1.0f), make_float2(v0, v1)), make_float2(TC_LO_SCALE, TC_LO_SCALE));
    lo = cvt_sat_half2(d.x, d.y);
}

// v -> (hi, lo) with hi = fp16(v) (saturated to the finite range), lo = fp16((v - hi) * 2^11)
__device__ __forceinline__ void split_hi_lo(float v, float &hi, float &lo) {
    const float c = fminf(fmaxf(v, -65504.0f), 65504.0f);
    hi = __half2float(__float2half_rn(c));
    lo = (c - hi) * TC_LO_SCALE;
}

__device__ __forceinline__ void unpack8(const uint4 &u, float (&f)[8]) {
    const __half2 *h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __half22float2(h[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}

}  // namespace stof
