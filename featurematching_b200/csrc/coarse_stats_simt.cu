// Debug twin of K1 (coarse_stats_tc.cu): the same workspace outputs -- row / column partial sums
// of 2^(t - ref) and the candidate list -- from a plain SIMT kernel with fp32 FMA dot products.
// Only fm_debug_coarse_match_simt reaches it; it lets a test tell a tensor-core plumbing bug
// from a resolve bug.  Not a fallback: the product entry point never calls it.
#include "fm_common.cuh"
#include "fm_ptx.cuh"

namespace fm {

namespace {

constexpr int kT = 64;   // tile edge

// phase 0: accumulate rowpart / colpart (atomics; the launcher clears them first)
// phase 1: emit candidates, using the finished row sums
template <int kPhase>
__global__ void __launch_bounds__(256) coarse_stats_simt_kernel(const __nv_bfloat16* __restrict__ f0,
                                                                const __nv_bfloat16* __restrict__ f1,
                                                                CoarseStatsParams p) {
    extern __shared__ __nv_bfloat16 sm[];
    const CoarseDims& d = p.d;
    const int C = d.C;
    __nv_bfloat16* s0 = sm;               // [kT][C + 2]
    __nv_bfloat16* s1 = sm + kT * (C + 2);
    const int b = blockIdx.z, i0 = blockIdx.y * kT, j0 = blockIdx.x * kT;
    for (int e = threadIdx.x; e < kT * C; e += blockDim.x) {
        const int r = e / C, c = e - r * C;
        s0[r * (C + 2) + c] = (i0 + r < d.N0) ? f0[(static_cast<size_t>(b) * d.N0 + i0 + r) * C + c] : __float2bfloat16(0.f);
        s1[r * (C + 2) + c] = (j0 + r < d.N1) ? f1[(static_cast<size_t>(b) * d.N1 + j0 + r) * C + c] : __float2bfloat16(0.f);
    }
    __syncthreads();
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    float acc[4][4] = {};
    for (int k = 0; k < C; ++k) {
        float a[4], bb[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) a[r] = __bfloat162float(s0[(ty * 4 + r) * (C + 2) + k]);
#pragma unroll
        for (int c = 0; c < 4; ++c) bb[c] = __bfloat162float(s1[(tx * 4 + c) * (C + 2) + k]);
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], bb[c], acc[r][c]);
    }
    const float maxn0 = p.maxnorm[b * 2 + 0], maxn1 = p.maxnorm[b * 2 + 1];
    const float cref = score_bound2(maxn0, maxn1, p.a2) - kHead2;
    float colsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = i0 + ty * 4 + r;
        if (i >= d.N0) continue;
        const float ci = score_bound2(p.norm0[static_cast<size_t>(b) * d.N0pad + i], maxn1, p.a2) - kHead2;
        const float wi = ex2_approx(ci - cref);
        const int ms = (i / kTileM) / d.tiles_per_split;
        float rowsum = 0.f;
        float full = 0.f;
        if (kPhase == 1) {
            const float* rp = p.rowpart + static_cast<size_t>(b) * d.NB * kColGroups * d.N0pad + i;
            for (int k = 0; k < d.NB * kColGroups; ++k) full += rp[static_cast<size_t>(k) * d.N0pad];
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = j0 + tx * 4 + c;
            if (j >= d.N1) continue;
            const float pij = ex2_approx(fmaf(acc[r][c], p.a2, -ci));
            if (kPhase == 0) {
                rowsum += pij;
                colsum[c] += pij * wi;
            } else if (pij > p.cand_thr * full) {
                // one record per entry: the other 31 dots are -inf (confidence 0 in K2)
                const int pool = static_cast<int>((blockIdx.x + blockIdx.y * gridDim.x + blockIdx.z * 7u) % d.num_pools);
                const int slot = atomicAdd(p.hot_count + pool, 1);
                if (slot < d.hot_cap) {
                    const size_t rec = static_cast<size_t>(pool) * d.hot_cap + slot;
                    p.hot_meta[rec] = make_uint2(static_cast<uint32_t>(b * d.N0 + i), static_cast<uint32_t>(j / kHotChunk));
                    for (int k = 0; k < kHotChunk; ++k)
                        p.hot_dots[rec * kHotChunk + k] = (k == j % kHotChunk) ? acc[r][c] : -INFINITY;
                } else {
                    atomicOr(p.status + kStFlags, FM_FLAG_CAND_OVERFLOW);
                }
                atomicMax(p.status + kStMaxUnitCand, slot + 1);
            }
        }
        if (kPhase == 0) {
            const int slot = (j0 + tx * 4) / kGroupCols;   // = kColGroups * nblk + column group, as K1 lays it out
            atomicAdd(p.rowpart + (static_cast<size_t>(b) * d.NB * kColGroups + slot) * d.N0pad + i, rowsum);
            // column sums of this row go to the split its tile belongs to
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = j0 + tx * 4 + c;
                if (j < d.N1) atomicAdd(p.colpart + (static_cast<size_t>(b) * d.S + ms) * d.N1pad + j, colsum[c]);
                colsum[c] = 0.f;
            }
        }
    }
}

}  // namespace

cudaError_t launch_coarse_stats_simt(const __nv_bfloat16* f0, const __nv_bfloat16* f1,
                                     const CoarseStatsParams& p, cudaStream_t st) {
    const CoarseDims& d = p.d;
    const size_t smem = static_cast<size_t>(2) * kT * (d.C + 2) * sizeof(__nv_bfloat16);
    static unsigned long long attr_set = 0;
    if (needs_attr_on_this_device(&attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(coarse_stats_simt_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(coarse_stats_simt_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
        if (e != cudaSuccess) return e;
    }
    cudaError_t e = cudaMemsetAsync(p.rowpart, 0, sizeof(float) * d.B * d.NB * kColGroups * d.N0pad, st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(p.colpart, 0, sizeof(float) * d.B * d.S * d.N1pad, st);
    if (e != cudaSuccess) return e;
    dim3 grid((d.N1 + kT - 1) / kT, (d.N0 + kT - 1) / kT, d.B);
    coarse_stats_simt_kernel<0><<<grid, 256, smem, st>>>(f0, f1, p); count_launch();
    coarse_stats_simt_kernel<1><<<grid, 256, smem, st>>>(f0, f1, p); count_launch();
    return cudaGetLastError();
}

}  // namespace fm
