// K2a on tensor cores -- FinePreprocess with fine_concat_coarse_feat (the shipped configuration).
//
//   out_k[m, ww, :] = merge_w[:, :Cf] . window_k[m, ww, :] + cvec_k[m]                fine_preprocess.py:56-60
//   cvec_k[m]       = merge_w[:, Cf:] . (down_w . tok_k[m] + down_b) + merge_b        fine_preprocess.py:54-55
//                   = Wc . tok_k[m] + bc,   Wc = merge_w[:, Cf:] . down_w,  bc = merge_w[:, Cf:] . down_b + merge_b
//
// Three kernels per call:
//   fold_kernel     Wc, bc and the 3-way bf16 splits of merge_w[:, :Cf] and Wc (a fp32 weight w is
//                   hi + mid + lo with bf16 terms, exact to 2^-25: fp32-grade products on the bf16 pipe)
//   ctx_gemm_kernel cvec for EVERY coarse token of both images: dense [B N, Cc] x [Cc, Cf] tcgen05 GEMM,
//                   tokens streamed by TMA (5.4 GFLOP per side at 32 pairs: cheaper than gathering
//                   the matched tokens, and independent of the match list)
//   window_merge_kernel   per tile = two windows: each W x W x Cf window is ONE 4-D TMA box out of
//                   the channels-last bf16 fine map (zero fill outside the map reproduces unfold's
//                   padding), lands as 49 K-major swizzled rows = the A operand of a
//                   128 x 64 x 64 tcgen05.mma against the resident split weights; the epilogue adds
//                   cvec and leaves through a swizzled staging buffer and TMA stores.
// The 49x unfolded map (fine_preprocess.py:43-46) is never built.
#include <algorithm>

#include "fm_common.cuh"
#include "fm_ptx.cuh"
#include "fm_tmap.cuh"

namespace fm {

namespace {

constexpr int kCf = 64;
constexpr int kSplit = 3;
constexpr int kGStages = 4;                   // A-operand ring, 16 KB per stage
constexpr int kGThreads = 384;                // control warpgroup + 2 epilogue warpgroups
constexpr int kWinRows = 64;                  // smem rows reserved per window (W*W <= 64)
constexpr int kStageBytes = 128 * 128;        // 128 rows x 64 bf16
constexpr int kWsplitBytes = kCf * 128;       // one split part of a [64 x 64] bf16 operand: 8 KB

// ---------------------------------------------------------------- 1. weight folding
// grid = Cf blocks (one output row o), Cc threads.
__global__ void fold_kernel(const float* __restrict__ down_w, const float* __restrict__ down_b,
                            const float* __restrict__ merge_w, const float* __restrict__ merge_b, int Cc,
                            __nv_bfloat16* __restrict__ wa_split,   // [3][Cf][Cf]
                            __nv_bfloat16* __restrict__ wc_split,   // [3][Cf][Cc]
                            float* __restrict__ bc) {               // [Cf]
    const int o = blockIdx.x, c = threadIdx.x;
    const float* wb = merge_w + static_cast<size_t>(o) * 2 * kCf + kCf;   // merge_w[o, Cf:]
    auto split3 = [](float w, __nv_bfloat16& hi, __nv_bfloat16& mid, __nv_bfloat16& lo) {
        hi = __float2bfloat16_rn(w);
        const float r1 = w - __bfloat162float(hi);
        mid = __float2bfloat16_rn(r1);
        lo = __float2bfloat16_rn(r1 - __bfloat162float(mid));
    };
    if (c < Cc) {
        float acc = 0.f;
        for (int k = 0; k < kCf; ++k) acc = fmaf(wb[k], down_w[static_cast<size_t>(k) * Cc + c], acc);
        __nv_bfloat16 h, m, l;
        split3(acc, h, m, l);
        wc_split[(0 * kCf + o) * static_cast<size_t>(Cc) + c] = h;
        wc_split[(1 * kCf + o) * static_cast<size_t>(Cc) + c] = m;
        wc_split[(2 * kCf + o) * static_cast<size_t>(Cc) + c] = l;
    }
    if (c < kCf) {
        __nv_bfloat16 h, m, l;
        split3(merge_w[static_cast<size_t>(o) * 2 * kCf + c], h, m, l);
        wa_split[(0 * kCf + o) * kCf + c] = h;
        wa_split[(1 * kCf + o) * kCf + c] = m;
        wa_split[(2 * kCf + o) * kCf + c] = l;
    }
    if (c == 0) {
        float acc = merge_b[o];
        for (int k = 0; k < kCf; ++k) acc = fmaf(wb[k], down_b[k], acc);
        bc[o] = acc;
    }
}

__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                        size_t n8) {
    for (size_t e = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; e < n8;
         e += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * e);
        const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * e + 1);
        __nv_bfloat162 v[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w),
                               __floats2bfloat162_rn(b.x, b.y), __floats2bfloat162_rn(b.z, b.w)};
        reinterpret_cast<uint4*>(dst)[e] = *reinterpret_cast<uint4*>(v);
    }
}

// 64 fp32 accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld_row64(uint32_t taddr, uint32_t (&lo)[32], uint32_t (&hi)[32]) {
    tmem_ld_32x32(taddr, lo);
    tmem_ld_32x32(taddr + 32, hi);
    tmem_ld_wait();
}

// ---------------------------------------------------------------- 2. cvec of every coarse token
struct CtxParams {
    int rows[2];          // B * N0, B * N1
    int tiles0, tiles;    // 128-row tiles of side 0, of both sides
    int KC;               // Cc / 64
    const float* bc;      // [Cf]
    float* ctx[2];        // [rows, Cf] fp32
};

struct __align__(16) GemmCtl {
    uint64_t a_full[kGStages];
    uint64_t a_empty[kGStages];
    uint64_t b_full;
    uint64_t t_full[2];
    uint64_t t_empty[2];
    uint32_t tmem_base;
    float bias[kCf];
};

__global__ void __launch_bounds__(kGThreads, 1)
ctx_gemm_kernel(const __grid_constant__ CUtensorMap tmap_tok0, const __grid_constant__ CUtensorMap tmap_tok1,
                const __grid_constant__ CUtensorMap tmap_w, const CtxParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_b = smem;                                   // KC x [hi | mid | lo] x 8 KB
    uint8_t* smem_a = smem + static_cast<size_t>(p.KC) * kSplit * kWsplitBytes;
    GemmCtl* ctl = reinterpret_cast<GemmCtl*>(smem_a + static_cast<size_t>(kGStages) * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KC = p.KC;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kGStages; ++s) { mbar_init(&ctl->a_full[s], 1); mbar_init(&ctl->a_empty[s], 1); }
        mbar_init(&ctl->b_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&ctl->t_full[s], 1); mbar_init(&ctl->t_empty[s], 4); }
        fence_mbar_init();
        tma_prefetch_desc(&tmap_tok0);
        tma_prefetch_desc(&tmap_tok1);
        tma_prefetch_desc(&tmap_w);
    }
    if (threadIdx.x < kCf) ctl->bias[threadIdx.x] = p.bc[threadIdx.x];
    if (warp == 2) tmem_alloc<128>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 0) {
        if (lane == 0) {   // ------------------------------------------------ TMA producer
            mbar_arrive_expect_tx(&ctl->b_full, static_cast<uint32_t>(KC) * kSplit * kWsplitBytes);
            for (int kc = 0; kc < KC; ++kc)
                tma_load_2d(smem_b + static_cast<size_t>(kc) * kSplit * kWsplitBytes, &tmap_w, &ctl->b_full, kc * 64, 0);
            uint32_t a_it = 0;
            for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
                const int side = t >= p.tiles0 ? 1 : 0;
                const int row0 = (side ? t - p.tiles0 : t) * 128;
                const CUtensorMap* tm = side ? &tmap_tok1 : &tmap_tok0;
                for (int kc = 0; kc < KC; ++kc, ++a_it) {
                    const uint32_t s = a_it % kGStages, ph = (a_it / kGStages) & 1;
                    mbar_wait(&ctl->a_empty[s], ph ^ 1);
                    mbar_arrive_expect_tx(&ctl->a_full[s], kStageBytes);
                    tma_load_2d(smem_a + static_cast<size_t>(s) * kStageBytes, tm, &ctl->a_full[s], kc * 64, row0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {   // ------------------------------------------------ MMA issuer
            constexpr uint32_t idesc = umma_idesc_bf16_f32(128, kCf);
            const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
            uint32_t a_it = 0, tile_it = 0;
            mbar_wait(&ctl->b_full, 0);
            for (int t = blockIdx.x; t < p.tiles; t += gridDim.x, ++tile_it) {
                const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
                mbar_wait(&ctl->t_empty[buf], tph ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * kCf;
                for (int kc = 0; kc < KC; ++kc, ++a_it) {
                    const uint32_t s = a_it % kGStages, ph = (a_it / kGStages) & 1;
                    mbar_wait(&ctl->a_full[s], ph);
                    tc_fence_after();
                    const uint32_t a_addr = a_base + s * kStageBytes;
#pragma unroll
                    for (int sp = 0; sp < kSplit; ++sp) {
                        const uint32_t b_addr = b_base + (kc * kSplit + sp) * kWsplitBytes;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_ss(d_tmem, umma_desc_k_sw128(a_addr + k * 32), umma_desc_k_sw128(b_addr + k * 32),
                                         idesc, (kc | sp | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&ctl->a_empty[s]);
                }
                umma_commit(&ctl->t_full[buf]);
            }
        }
    } else if (warp >= 4) {   // ---------------------------------------------- epilogue: WG g takes tiles of parity g
        const int g = (warp - 4) >> 2, q = warp & 3;
        uint32_t tile_it = 0;
        for (int t = blockIdx.x; t < p.tiles; t += gridDim.x, ++tile_it) {
            if ((tile_it & 1) != static_cast<uint32_t>(g)) continue;
            const int side = t >= p.tiles0 ? 1 : 0;
            const int row = (side ? t - p.tiles0 : t) * 128 + q * 32 + lane;
            mbar_wait(&ctl->t_full[g], (tile_it >> 1) & 1);
            tc_fence_after();
            uint32_t lo[32], hi[32];
            tmem_ld_row64(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + g * kCf, lo, hi);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->t_empty[g]);
            if (row < p.rows[side]) {
                float4* dst = reinterpret_cast<float4*>(p.ctx[side] + static_cast<size_t>(row) * kCf);
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    dst[c] = make_float4(__uint_as_float(lo[4 * c]) + ctl->bias[4 * c], __uint_as_float(lo[4 * c + 1]) + ctl->bias[4 * c + 1],
                                         __uint_as_float(lo[4 * c + 2]) + ctl->bias[4 * c + 2], __uint_as_float(lo[4 * c + 3]) + ctl->bias[4 * c + 3]);
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    dst[8 + c] = make_float4(__uint_as_float(hi[4 * c]) + ctl->bias[32 + 4 * c], __uint_as_float(hi[4 * c + 1]) + ctl->bias[32 + 4 * c + 1],
                                             __uint_as_float(hi[4 * c + 2]) + ctl->bias[32 + 4 * c + 2], __uint_as_float(hi[4 * c + 3]) + ctl->bias[32 + 4 * c + 3]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<128>(tmem_base);
}

// ---------------------------------------------------------------- 3. windows x merge weights
struct WindowParams {
    const int64_t* b_ids;
    const int64_t* ids[2];   // i_ids, j_ids
    const int* count;
    int capacity;
    int N[2], wc[2];         // tokens per image, coarse grid width
    int stride, pad, W, WW;
    const float* ctx[2];     // [B * N, Cf] fp32 (ctx_gemm_kernel)
};

template <typename TO> struct OutTraits;
template <> struct OutTraits<float> { static constexpr int kHalves = 2; };           // 64 fp32 = 2 x 128 B
template <> struct OutTraits<__nv_bfloat16> { static constexpr int kHalves = 1; };   // 64 bf16 = 128 B
constexpr int kStgBytes = 7 * 1024;   // one (window, half): up to 56 rows x 128 B, 1024-aligned

template <typename TO>
__global__ void __launch_bounds__(kGThreads, 1)
window_merge_kernel(const __grid_constant__ CUtensorMap tmap_f0, const __grid_constant__ CUtensorMap tmap_f1,
                    const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_o0,
                    const __grid_constant__ CUtensorMap tmap_o1, const WindowParams p) {
    constexpr int kHalves = OutTraits<TO>::kHalves;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_b = smem;                                        // [hi | mid | lo] x 8 KB
    uint8_t* smem_a = smem_b + kSplit * kWsplitBytes;              // ring of 128-row tiles
    uint8_t* smem_o = smem_a + kGStages * kStageBytes;             // staging: [2 WG][2 windows][kHalves]
    GemmCtl* ctl = reinterpret_cast<GemmCtl*>(smem_o + 2 * 2 * kHalves * kStgBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kGStages; ++s) { mbar_init(&ctl->a_full[s], 1); mbar_init(&ctl->a_empty[s], 1); }
        mbar_init(&ctl->b_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&ctl->t_full[s], 1); mbar_init(&ctl->t_empty[s], 4); }
        fence_mbar_init();
        tma_prefetch_desc(&tmap_f0); tma_prefetch_desc(&tmap_f1); tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_o0); tma_prefetch_desc(&tmap_o1);
    }
    if (warp == 2) tmem_alloc<128>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;
    const int M = p.count ? min(*p.count, p.capacity) : p.capacity;
    const int tiles = 2 * ((M + 1) / 2);   // tile T: side T & 1, matches 2 (T >> 1) and 2 (T >> 1) + 1
    const uint32_t win_bytes = static_cast<uint32_t>(p.WW) * 128;

    if (warp == 0) {
        if (lane == 0) {   // ------------------------------------------------ TMA producer
            mbar_arrive_expect_tx(&ctl->b_full, kSplit * kWsplitBytes);
            tma_load_2d(smem_b, &tmap_w, &ctl->b_full, 0, 0);
            uint32_t it = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
                const int side = t & 1, m0 = (t >> 1) * 2;
                const uint32_t s = it % kGStages, ph = (it / kGStages) & 1;
                const int nvalid = min(2, M - m0);
                mbar_wait(&ctl->a_empty[s], ph ^ 1);
                mbar_arrive_expect_tx(&ctl->a_full[s], nvalid * win_bytes);
                for (int w = 0; w < nvalid; ++w) {
                    const int m = m0 + w;
                    const int b = static_cast<int>(__ldg(p.b_ids + m));
                    const int cell = static_cast<int>(__ldg((side ? p.ids[1] : p.ids[0]) + m));
                    const int cr = cell / (side ? p.wc[1] : p.wc[0]), cc = cell - cr * (side ? p.wc[1] : p.wc[0]);
                    tma_load_4d(smem_a + static_cast<size_t>(s) * kStageBytes + w * kWinRows * 128,
                                side ? &tmap_f1 : &tmap_f0, &ctl->a_full[s], 0, cc * p.stride - p.pad,
                                cr * p.stride - p.pad, b);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {   // ------------------------------------------------ MMA issuer
            constexpr uint32_t idesc = umma_idesc_bf16_f32(128, kCf);
            const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
            uint32_t it = 0;
            mbar_wait(&ctl->b_full, 0);
            for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
                const uint32_t buf = it & 1, tph = (it >> 1) & 1;
                const uint32_t s = it % kGStages, ph = (it / kGStages) & 1;
                mbar_wait(&ctl->t_empty[buf], tph ^ 1);
                mbar_wait(&ctl->a_full[s], ph);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * kCf;
                const uint32_t a_addr = a_base + s * kStageBytes;
#pragma unroll
                for (int sp = 0; sp < kSplit; ++sp)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16_ss(d_tmem, umma_desc_k_sw128(a_addr + k * 32),
                                     umma_desc_k_sw128(b_base + sp * kWsplitBytes + k * 32), idesc, (sp | k) != 0 ? 1u : 0u);
                umma_commit(&ctl->a_empty[s]);
                umma_commit(&ctl->t_full[buf]);
            }
        }
    } else if (warp >= 4) {   // ---------------------------------------------- epilogue: WG g takes tiles of parity g
        const int g = (warp - 4) >> 2, q = warp & 3;
        const bool leader = (warp & 3) == 0 && lane == 0;
        const int w = q >> 1;                       // window slot of this thread's row
        const int ww = (q & 1) * 32 + lane;         // window cell
        uint8_t* stg = smem_o + static_cast<size_t>(g) * 2 * kHalves * kStgBytes + static_cast<size_t>(w) * kHalves * kStgBytes;
        // match / cell ids of this warpgroup's NEXT tile are fetched one tile ahead and the 64
        // context values of the current tile before the accumulator wait, so neither global
        // latency sits between the TMEM read and the stores
        auto tile_ids = [&](int t, int& b, int& cell) {
            const int side = t & 1, m = (t >> 1) * 2 + w;
            b = 0; cell = 0;
            if (t < tiles && m < M && ww < p.WW) {
                b = static_cast<int>(__ldg(p.b_ids + m));
                cell = static_cast<int>(__ldg((side ? p.ids[1] : p.ids[0]) + m));
            }
        };
        int nb, ncell;
        tile_ids(blockIdx.x + g * gridDim.x, nb, ncell);
        uint32_t it = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
            if ((it & 1) != static_cast<uint32_t>(g)) continue;
            const int side = t & 1, m0 = (t >> 1) * 2;
            const int m = m0 + w;
            const bool valid = m < M && ww < p.WW;
            float4 cx[16];
            if (valid) {
                const float4* cv = reinterpret_cast<const float4*>((side ? p.ctx[1] : p.ctx[0]) + (static_cast<size_t>(nb) * (side ? p.N[1] : p.N[0]) + ncell) * kCf);
#pragma unroll
                for (int c = 0; c < 16; ++c) cx[c] = __ldg(cv + c);
            } else {
#pragma unroll
                for (int c = 0; c < 16; ++c) cx[c] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            tile_ids(t + 2 * gridDim.x, nb, ncell);
            mbar_wait(&ctl->t_full[g], (it >> 1) & 1);
            tc_fence_after();
            uint32_t lo[32], hi[32];
            tmem_ld_row64(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + g * kCf, lo, hi);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->t_empty[g]);

            // the previous tile's TMA stores must have finished reading the staging buffer
            if (leader) bulk_wait_group_read<0>();
            named_bar_sync(1 + g, 128);
            if (valid) {
                uint8_t* rowp = stg + ww * 128;
                const int sw = ww & 7;
                if constexpr (kHalves == 2) {
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float4 a = cx[c];
                        *reinterpret_cast<float4*>(rowp + ((c ^ sw) << 4)) =
                            make_float4(__uint_as_float(lo[4 * c]) + a.x, __uint_as_float(lo[4 * c + 1]) + a.y,
                                        __uint_as_float(lo[4 * c + 2]) + a.z, __uint_as_float(lo[4 * c + 3]) + a.w);
                    }
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float4 a = cx[8 + c];
                        *reinterpret_cast<float4*>(rowp + kStgBytes + ((c ^ sw) << 4)) =
                            make_float4(__uint_as_float(hi[4 * c]) + a.x, __uint_as_float(hi[4 * c + 1]) + a.y,
                                        __uint_as_float(hi[4 * c + 2]) + a.z, __uint_as_float(hi[4 * c + 3]) + a.w);
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 8; ++c) {   // 8 channels = 16 B of bf16 per chunk
                        const uint32_t* src = c < 4 ? lo + 8 * c : hi + 8 * (c - 4);
                        const float4 a0 = cx[2 * c], a1 = cx[2 * c + 1];
                        __nv_bfloat162 v[4] = {
                            __floats2bfloat162_rn(__uint_as_float(src[0]) + a0.x, __uint_as_float(src[1]) + a0.y),
                            __floats2bfloat162_rn(__uint_as_float(src[2]) + a0.z, __uint_as_float(src[3]) + a0.w),
                            __floats2bfloat162_rn(__uint_as_float(src[4]) + a1.x, __uint_as_float(src[5]) + a1.y),
                            __floats2bfloat162_rn(__uint_as_float(src[6]) + a1.z, __uint_as_float(src[7]) + a1.w)};
                        *reinterpret_cast<uint4*>(rowp + ((c ^ sw) << 4)) = *reinterpret_cast<uint4*>(v);
                    }
                }
            }
            fence_proxy_async_smem();
            named_bar_sync(1 + g, 128);
            if (leader) {
                const CUtensorMap* tmo = side ? &tmap_o1 : &tmap_o0;
                uint8_t* base = smem_o + static_cast<size_t>(g) * 2 * kHalves * kStgBytes;
                const int nvalid = min(2, M - m0);
                for (int v = 0; v < nvalid; ++v)
#pragma unroll
                    for (int h = 0; h < kHalves; ++h)
                        tma_store_2d(tmo, base + (v * kHalves + h) * kStgBytes, h * 32, (m0 + v) * p.WW);
                bulk_commit_group();
            }
        }
        if (leader) bulk_wait_group<0>();   // all stores complete before the CTA exits
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<128>(tmem_base);
}

// ---------------------------------------------------------------- host side
size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

struct GatherWorkspace {
    __nv_bfloat16* wa_split;   // [3][Cf][Cf]
    __nv_bfloat16* wc_split;   // [3][Cf][Cc]
    float* bc;                 // [Cf]
    float* ctx[2];             // [B * N, Cf]
    __nv_bfloat16* tok[2];     // bf16 copies of fp32 coarse tokens
    size_t total;
};

GatherWorkspace carve(int B, int N0, int N1, int Cc, bool f32_tokens, void* base) {
    GatherWorkspace w{};
    uint8_t* p = static_cast<uint8_t*>(base);
    size_t off = 0;
    auto take = [&](size_t bytes) { uint8_t* r = p ? p + off : nullptr; off = align_up(off + bytes); return r; };
    w.wa_split = reinterpret_cast<__nv_bfloat16*>(take(sizeof(__nv_bfloat16) * kSplit * kCf * kCf));
    w.wc_split = reinterpret_cast<__nv_bfloat16*>(take(sizeof(__nv_bfloat16) * kSplit * kCf * Cc));
    w.bc = reinterpret_cast<float*>(take(sizeof(float) * kCf));
    w.ctx[0] = reinterpret_cast<float*>(take(sizeof(float) * static_cast<size_t>(B) * N0 * kCf));
    w.ctx[1] = reinterpret_cast<float*>(take(sizeof(float) * static_cast<size_t>(B) * N1 * kCf));
    if (f32_tokens) {
        w.tok[0] = reinterpret_cast<__nv_bfloat16*>(take(sizeof(__nv_bfloat16) * static_cast<size_t>(B) * N0 * Cc));
        w.tok[1] = reinterpret_cast<__nv_bfloat16*>(take(sizeof(__nv_bfloat16) * static_cast<size_t>(B) * N1 * Cc));
    }
    w.total = off;
    return w;
}

template <typename TO>
cudaError_t launch_window_merge(const CUtensorMap& tf0, const CUtensorMap& tf1, const CUtensorMap& tw,
                                const CUtensorMap& to0, const CUtensorMap& to1, const WindowParams& wp, int grid,
                                cudaStream_t st) {
    constexpr size_t smem = kSplit * kWsplitBytes + kGStages * kStageBytes +
                            2 * 2 * OutTraits<TO>::kHalves * kStgBytes + sizeof(GemmCtl) + 1024;
    cudaError_t e = cudaFuncSetAttribute(window_merge_kernel<TO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    window_merge_kernel<TO><<<grid, kGThreads, smem, st>>>(tf0, tf1, tw, to0, to1, wp); count_launch();
    return cudaGetLastError();
}

}  // namespace

bool fine_gather_tc_supported(int Cf, int Cc, int W) {
    return Cf == kCf && Cc >= 64 && Cc <= 256 && Cc % 64 == 0 && W >= 2 && W * W <= 56;
}

size_t fine_gather_tc_workspace_bytes(int B, int N0, int N1, int Cc, bool f32_tokens) {
    return carve(B, N0, N1, Cc, f32_tokens, nullptr).total;
}

cudaError_t launch_fine_gather_tc(const void* ff0, const void* ff1, const void* fc0, const void* fc1, int featc_dtype,
                                  const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const int* count,
                                  int capacity, int B, int Hf, int Wf, int N0, int N1, int Cc, int w0c, int w1c,
                                  int stride, int pad, int W, const float* down_w, const float* down_b,
                                  const float* merge_w, const float* merge_b, void* win0, void* win1, int out_dtype,
                                  void* workspace, size_t workspace_bytes, int num_sms, cudaStream_t st) {
    const bool f32_tok = featc_dtype == FM_DTYPE_F32;
    const GatherWorkspace ws = carve(B, N0, N1, Cc, f32_tok, workspace);
    if (!workspace || ws.total > workspace_bytes) {
        set_error("fine gather workspace too small: need %zu bytes, got %zu", ws.total, workspace_bytes);
        return cudaErrorInvalidValue;
    }
    // ---- 1. fold the weights; bf16 tokens
    fold_kernel<<<kCf, Cc < kCf ? kCf : Cc, 0, st>>>(down_w, down_b, merge_w, merge_b, Cc, ws.wa_split, ws.wc_split, ws.bc); count_launch();
    const __nv_bfloat16* tok[2] = {static_cast<const __nv_bfloat16*>(fc0), static_cast<const __nv_bfloat16*>(fc1)};
    if (f32_tok) {
        const size_t n8[2] = {static_cast<size_t>(B) * N0 * Cc / 8, static_cast<size_t>(B) * N1 * Cc / 8};
        const void* src[2] = {fc0, fc1};
        for (int s = 0; s < 2; ++s) {
            const int grid = static_cast<int>(std::min<size_t>((n8[s] + 255) / 256, static_cast<size_t>(num_sms) * 8));
            cast_bf16_kernel<<<grid, 256, 0, st>>>(static_cast<const float*>(src[s]), ws.tok[s], n8[s]); count_launch();
            tok[s] = ws.tok[s];
        }
    }
    // ---- 2. cvec of every token
    CUtensorMap t_tok[2], t_wc, t_wa, t_f[2], t_o[2];
    const int rows[2] = {B * N0, B * N1};
    bool ok = true;
    for (int s = 0; s < 2; ++s) {
        const uint64_t dims[2] = {static_cast<uint64_t>(Cc), static_cast<uint64_t>(rows[s])};
        const uint64_t str[1] = {static_cast<uint64_t>(Cc) * 2};
        const uint32_t box[2] = {64, 128};
        ok = ok && encode_tiled(&t_tok[s], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, tok[s], dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    }
    {
        const uint64_t dims[2] = {static_cast<uint64_t>(Cc), kSplit * kCf};
        const uint64_t str[1] = {static_cast<uint64_t>(Cc) * 2};
        const uint32_t box[2] = {64, kSplit * kCf};
        ok = ok && encode_tiled(&t_wc, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ws.wc_split, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
        const uint64_t dims_a[2] = {kCf, kSplit * kCf};
        const uint64_t str_a[1] = {kCf * 2};
        ok = ok && encode_tiled(&t_wa, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ws.wa_split, dims_a, str_a, box, CU_TENSOR_MAP_SWIZZLE_128B);
    }
    const void* ff[2] = {ff0, ff1};
    void* out[2] = {win0, win1};
    const bool o32 = out_dtype == FM_DTYPE_F32;
    const int WW = W * W;
    for (int s = 0; s < 2; ++s) {
        const uint64_t dims[4] = {kCf, static_cast<uint64_t>(Wf), static_cast<uint64_t>(Hf), static_cast<uint64_t>(B)};
        const uint64_t str[3] = {kCf * 2, static_cast<uint64_t>(Wf) * kCf * 2, static_cast<uint64_t>(Hf) * Wf * kCf * 2};
        const uint32_t box[4] = {kCf, static_cast<uint32_t>(W), static_cast<uint32_t>(W), 1};
        ok = ok && encode_tiled(&t_f[s], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ff[s], dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
        const uint64_t odims[2] = {kCf, static_cast<uint64_t>(capacity) * WW};
        const uint64_t ostr[1] = {static_cast<uint64_t>(kCf) * (o32 ? 4 : 2)};
        const uint32_t obox[2] = {o32 ? 32u : 64u, static_cast<uint32_t>(WW)};
        ok = ok && encode_tiled(&t_o[s], o32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out[s],
                                odims, ostr, obox, CU_TENSOR_MAP_SWIZZLE_128B);
    }
    if (!ok) {
        set_error("cuTensorMapEncodeTiled failed in fine gather (pointers must be 16-byte aligned)");
        return cudaErrorInvalidValue;
    }
    CtxParams cp{};
    cp.rows[0] = rows[0]; cp.rows[1] = rows[1];
    cp.tiles0 = (rows[0] + 127) / 128;
    cp.tiles = cp.tiles0 + (rows[1] + 127) / 128;
    cp.KC = Cc / 64;
    cp.bc = ws.bc;
    cp.ctx[0] = ws.ctx[0]; cp.ctx[1] = ws.ctx[1];
    const size_t smem_ctx = static_cast<size_t>(cp.KC) * kSplit * kWsplitBytes + kGStages * kStageBytes + sizeof(GemmCtl) + 1024;
    cudaError_t e = cudaFuncSetAttribute(ctx_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_ctx));
    if (e != cudaSuccess) return e;
    ctx_gemm_kernel<<<cp.tiles < num_sms ? cp.tiles : num_sms, kGThreads, smem_ctx, st>>>(t_tok[0], t_tok[1], t_wc, cp); count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // ---- 3. windows
    WindowParams wp{};
    wp.b_ids = b_ids; wp.ids[0] = i_ids; wp.ids[1] = j_ids; wp.count = count; wp.capacity = capacity;
    wp.N[0] = N0; wp.N[1] = N1; wp.wc[0] = w0c; wp.wc[1] = w1c;
    wp.stride = stride; wp.pad = pad; wp.W = W; wp.WW = WW;
    wp.ctx[0] = ws.ctx[0]; wp.ctx[1] = ws.ctx[1];
    const int tiles_cap = 2 * ((capacity + 1) / 2);
    const int grid = tiles_cap < num_sms ? tiles_cap : num_sms;
    if (o32) return launch_window_merge<float>(t_f[0], t_f[1], t_wa, t_o[0], t_o[1], wp, grid, st);
    return launch_window_merge<__nv_bfloat16>(t_f[0], t_f[1], t_wa, t_o[0], t_o[1], wp, grid, st);
}

}  // namespace fm
