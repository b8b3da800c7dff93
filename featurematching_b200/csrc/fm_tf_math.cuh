// Arithmetic shared by the fine and the coarse LocalFeatureTransformer kernels.
#pragma once
#include <cuda_bf16.h>

#include "fm_ptx.cuh"

namespace fm {

// (a, b) -> packed bf16 pair of the high parts and of the residuals a - ah, b - bh
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    float ra, rb;   // one packed subtraction for both residuals
    unpack2(sub2(pack2(a, b), pack2(__uint_as_float(hi << 16), __uint_as_float(hi & 0xffff0000u))), ra, rb);
    const __nv_bfloat162 l = __floats2bfloat162_rn(ra, rb);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
// elu(x) + 1   (attentions.py:11-12)
__device__ __forceinline__ float elu1(float x) { return x > 0.f ? x + 1.f : ex2_approx(x * 1.4426950408889634f); }

// mean and centred sum of squares of 32 values
__device__ __forceinline__ float2 local_stats(const float (&v)[32]) {
    uint64_t s = pack2(v[0], v[1]);   // packed fp32x2 arithmetic: one instruction per channel pair
#pragma unroll
    for (int i = 2; i < 32; i += 2) s = add2(s, pack2(v[i], v[i + 1]));
    float s0, s1;
    unpack2(s, s0, s1);
    const float mean = (s0 + s1) * (1.f / 32.f);
    const uint64_t nm = pack2(-mean, -mean);
    uint64_t m2 = pack2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
        const uint64_t d = add2(pack2(v[i], v[i + 1]), nm);
        m2 = fma2(d, d, m2);
    }
    unpack2(m2, s0, s1);
    return make_float2(mean, s0 + s1);
}

// 16 x 32-bit (32 bf16) -> global, two 256-bit stores; p must be 32-byte aligned
__device__ __forceinline__ void stg256_b32(void* p, const uint32_t (&v)[16]) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(static_cast<uint8_t*>(p) + 32), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]),
                   "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}

}  // namespace fm
