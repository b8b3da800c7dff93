// K0 -- coarse feature preparation: token norms (fp32, of the bf16-rounded values), per-pair
// maximum norm, and the bf16 copy when the caller's features are fp32.
//
// The norms give every image-0 token a Cauchy-Schwarz bound on its scores, which the statistics
// kernel uses as a fixed, overflow-proof exponent reference instead of an online running max
// (network/utils/coarse_matching_new.py:64-68 is the arithmetic being restated).
#include "fm_common.cuh"

namespace fm {

namespace {

constexpr int kRowsPerBlock = 32;  // 8 warps x 4 rows

template <bool kF32In>
__global__ void __launch_bounds__(256) coarse_prep_kernel(const void* __restrict__ f0,
                                                          const void* __restrict__ f1,
                                                          __nv_bfloat16* __restrict__ f0b,
                                                          __nv_bfloat16* __restrict__ f1b,
                                                          float* __restrict__ norm0,
                                                          float* __restrict__ maxnorm, int N0, int N1,
                                                          int N0pad, int C) {
    const int b = blockIdx.y >> 1, side = blockIdx.y & 1;
    const int N = side ? N1 : N0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const void* src = side ? f1 : f0;
    [[maybe_unused]] __nv_bfloat16* dst = side ? f1b : f0b;
    float wmax = 0.f;
    // the warp's four rows: all loads first (one 16-byte load per lane and row; issued one row after the other the
    // kernel ran at 39 us for 168 MB = 65 % of the copy bandwidth)
    const int row_base = blockIdx.x * kRowsPerBlock + warp * 4;
    const bool lane_ok = lane * 8 < C;
    __nv_bfloat162 v[4][4];
    if constexpr (kF32In) {
        float4 a[4], c[4];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
            const int row = row_base + rr;
            if (row < N && lane_ok) {
                const float4* p = reinterpret_cast<const float4*>(static_cast<const float*>(src) + (static_cast<size_t>(b) * N + row) * C) + lane * 2;
                a[rr] = __ldg(p); c[rr] = __ldg(p + 1);
            } else {
                a[rr] = make_float4(0.f, 0.f, 0.f, 0.f); c[rr] = a[rr];
            }
        }
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
            v[rr][0] = __floats2bfloat162_rn(a[rr].x, a[rr].y);
            v[rr][1] = __floats2bfloat162_rn(a[rr].z, a[rr].w);
            v[rr][2] = __floats2bfloat162_rn(c[rr].x, c[rr].y);
            v[rr][3] = __floats2bfloat162_rn(c[rr].z, c[rr].w);
            const int row = row_base + rr;
            if (row < N && lane_ok)
                *reinterpret_cast<uint4*>(dst + (static_cast<size_t>(b) * N + row) * C + lane * 8) = *reinterpret_cast<uint4*>(v[rr]);
        }
    } else {
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
            const int row = row_base + rr;
            uint4 raw = make_uint4(0u, 0u, 0u, 0u);
            if (row < N && lane_ok)
                raw = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(src) + (static_cast<size_t>(b) * N + row) * C) + lane);
            *reinterpret_cast<uint4*>(v[rr]) = raw;
        }
    }
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
        const int row = row_base + rr;
        float ss = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float2 f = __bfloat1622float2(v[rr][k]);
            ss = fmaf(f.x, f.x, ss);
            ss = fmaf(f.y, f.y, ss);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        const float n = sqrtf(ss);
        if (row < N) {
            if (side == 0 && lane == 0) norm0[static_cast<size_t>(b) * N0pad + row] = n;
            wmax = fmaxf(wmax, n);
        }
    }
    __shared__ float smax[8];
    if (lane == 0) smax[warp] = wmax;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = smax[0];
#pragma unroll
        for (int k = 1; k < 8; ++k) m = fmaxf(m, smax[k]);
        // norms are non-negative, so the float ordering equals the unsigned-bit ordering
        atomicMax(reinterpret_cast<unsigned*>(maxnorm) + b * 2 + side, __float_as_uint(m));
    }
}

}  // namespace

cudaError_t launch_coarse_prep(const void* f0, const void* f1, int dtype, const CoarseDims& d,
                               const CoarseWorkspace& w, cudaStream_t st) {
    const int nmax = d.N0 > d.N1 ? d.N0 : d.N1;
    dim3 grid((nmax + kRowsPerBlock - 1) / kRowsPerBlock, d.B * 2);
    if (dtype == FM_DTYPE_F32)
        coarse_prep_kernel<true><<<grid, 256, 0, st>>>(f0, f1, w.f0b, w.f1b, w.norm0, w.maxnorm,
                                                        d.N0, d.N1, d.N0pad, d.C);
    else
        coarse_prep_kernel<false><<<grid, 256, 0, st>>>(f0, f1, nullptr, nullptr, w.norm0, w.maxnorm,
                                                         d.N0, d.N1, d.N0pad, d.C);
    count_launch();
    return cudaGetLastError();
}

}  // namespace fm
