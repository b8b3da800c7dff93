// K1 -- coarse correlation fused with the dual-softmax statistics (sm_100a, tcgen05 + TMEM + TMA).
//
// Replaces, without ever writing the N0 x N1 matrix to memory,
//   sim  = einsum('nlc,nsc->nls', f0/sqrt(C), f1/sqrt(C)) / temperature   coarse_matching_new.py:64-66
//   conf = softmax(sim, 1) * softmax(sim, 2)                              coarse_matching_new.py:68
// by producing, per pair, the two soft-max denominators of every token (as partial sums that the
// resolve kernels finish) plus the short list of entries that can exceed the threshold.
//
// Work unit = (pair b, block of 256 image-1 tokens, contiguous range of 128-token image-0 tiles).
// The image-1 block is the RESIDENT MMA operand (B, K-major, 128 KB of smem for C = 256); the
// image-0 tiles stream through a 5-stage TMA ring as the A operand.  Each tile is one
// 128 x 256 x C tcgen05.mma accumulation into one of two 256-column TMEM buffers.
//
// Threads: 12 warps = 3 warpgroups.  Warpgroup 0 is control (warp 0 TMA producer, warp 1 MMA
// issuer, warp 2 TMEM allocation); it gives its registers to the two epilogue warpgroups.
// Epilogue (8 warps = 2 warpgroups, warpgroup g owns columns [128 g, 128 g + 128) of every tile;
// a thread owns one accumulator row = one image-0 token):
//   t    = a2 * dot                      a2 = log2(e) / (C * temperature)
//   p    = 2^(t - c_i)                   c_i = U_i - kHead2, U_i >= every score of row i (norm bound)
//   row  : rowsum += p                   -> rowpart[b][2 nblk + g][i]          (one store per tile)
//   col  : colacc[j] += p * 2^(c_i - cmax)   thread-local over ALL streamed tiles of the unit,
//                                        reduced across the 128 rows once per unit -> colpart
//   cand : an entry with softmax_j share > kCandMargin * thr of its own 128-column partial sum
//          is pushed to the unit's candidate list (found by re-reading the hot 32-column chunk
//          from TMEM; rare).  conf > thr implies share > thr of the FULL row sum, so the list
//          is a superset of everything the resolve step needs.
// One MUFU.EX2 per score; scale/bias, row sum and column accumulate are packed f32x2 ops.
#include <cuda.h>

#include "fm_common.cuh"
#include "fm_ptx.cuh"
#include "fm_tmap.cuh"

namespace fm {

namespace {

struct __align__(16) SmemCtl {
    float colscratch[kEpiWarps][128];   // first: float4 stores need 16-byte alignment
    uint64_t a_full[kAStages];
    uint64_t a_empty[kAStages];
    uint64_t b_full;
    uint64_t b_empty;
    uint64_t t_full[2];
    uint64_t t_empty[2];
    uint32_t tmem_base;
    int cand_count;
};

constexpr size_t kSmemOperands = static_cast<size_t>(kBChunkBytes) * (kMaxC / kChunkK) +
                                 static_cast<size_t>(kAStageBytes) * kAStages;
constexpr size_t kSmemBytes = kSmemOperands + sizeof(SmemCtl) + 1024;  // + alignment slack
static_assert(kSmemBytes <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

// One 32-column chunk of the thread's accumulator row.
template <bool kMasked>
__device__ __forceinline__ void process_chunk(const uint32_t (&r)[32], uint64_t a2p, uint64_t biasp,
                                              uint64_t wp, uint64_t* colacc, uint64_t& racc0,
                                              uint64_t& racc1, float& cmax, int nvalid_in_chunk) {
    float m = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        uint64_t t2 = fma2(pack2(__uint_as_float(r[2 * k]), __uint_as_float(r[2 * k + 1])), a2p, biasp);
        float tl, th;
        unpack2(t2, tl, th);
        float pl = ex2_approx(tl), ph = ex2_approx(th);
        if constexpr (kMasked) {
            pl = (2 * k < nvalid_in_chunk) ? pl : 0.f;
            ph = (2 * k + 1 < nvalid_in_chunk) ? ph : 0.f;
        }
        uint64_t p2 = pack2(pl, ph);
        if (k & 1) racc1 = add2(racc1, p2); else racc0 = add2(racc0, p2);
        colacc[k] = fma2(p2, wp, colacc[k]);
        m = max3(m, pl, ph);
    }
    cmax = m;
}

// One butterfly step of the transposing reduction of 2*kHalf per-thread column sums across
// lanes: afterwards a lane keeps kHalf columns (the upper half if its `kBit` lane bit is set),
// each summed over the lane pair.
template <int kHalf, int kBit>
__device__ __forceinline__ void reduce_step(float* v, int lane) {
    const bool up = (lane & kBit) != 0;
#pragma unroll
    for (int k = 0; k < kHalf; ++k) {
        const float send = up ? v[k] : v[k + kHalf];
        const float keep = up ? v[k + kHalf] : v[k];
        v[k] = keep + __shfl_xor_sync(0xffffffffu, send, kBit);
    }
}

__global__ void __launch_bounds__(kCoarseThreads, 1)
coarse_stats_tc_kernel(const __grid_constant__ CUtensorMap tmap0,
                       const __grid_constant__ CUtensorMap tmap1, const CoarseStatsParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_b = smem;                                            // KC x 32 KB, resident
    uint8_t* smem_a = smem + static_cast<size_t>(kBChunkBytes) * (kMaxC / kChunkK);  // ring
    SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem + kSmemOperands);

    const CoarseDims& d = p.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KC = d.KC;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kAStages; ++s) { mbar_init(&ctl->a_full[s], 1); mbar_init(&ctl->a_empty[s], 1); }
        mbar_init(&ctl->b_full, 1);
        mbar_init(&ctl->b_empty, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&ctl->t_full[s], 1); mbar_init(&ctl->t_empty[s], kEpiWarps); }
        ctl->cand_count = 0;
        fence_mbar_init();
        tma_prefetch_desc(&tmap0);
        tma_prefetch_desc(&tmap1);
    }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    // Register re-balancing between warpgroups: the control warpgroup keeps 40 registers per
    // thread, the two epilogue warpgroups grow to 232 (128 live column accumulators each).
    if (warp < kCtlWarps) {
    reg_dealloc<40>();
    if (warp == 0) {
        // ===================================================== TMA producer (one lane)
        if (lane == 0) {
            uint32_t a_it = 0, unit_it = 0;
            for (int u = blockIdx.x; u < d.num_units; u += gridDim.x, ++unit_it) {
                const int ms = u % d.S, nblk = (u / d.S) % d.NB, b = u / (d.S * d.NB);
                const int mt0 = ms * d.tiles_per_split;
                const int mt1 = min(mt0 + d.tiles_per_split, d.MT);
                mbar_wait(&ctl->b_empty, (unit_it & 1) ^ 1);
                mbar_arrive_expect_tx(&ctl->b_full, static_cast<uint32_t>(kBChunkBytes) * KC);
                for (int kc = 0; kc < KC; ++kc)
                    tma_load_3d(smem_b + static_cast<size_t>(kc) * kBChunkBytes, &tmap1, &ctl->b_full,
                                kc * kChunkK, nblk * kBlockN, b);
                for (int mt = mt0; mt < mt1; ++mt) {
                    for (int kc = 0; kc < KC; ++kc, ++a_it) {
                        const uint32_t s = a_it % kAStages, ph = (a_it / kAStages) & 1;
                        mbar_wait(&ctl->a_empty[s], ph ^ 1);
                        mbar_arrive_expect_tx(&ctl->a_full[s], kAStageBytes);
                        tma_load_3d(smem_a + static_cast<size_t>(s) * kAStageBytes, &tmap0, &ctl->a_full[s],
                                    kc * kChunkK, mt * kTileM, b);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one lane)
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_f32(kTileM, kBlockN);
            const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
            uint32_t a_it = 0, tile_it = 0, unit_it = 0;
            for (int u = blockIdx.x; u < d.num_units; u += gridDim.x, ++unit_it) {
                const int ms = u % d.S;
                const int mt0 = ms * d.tiles_per_split;
                const int mt1 = min(mt0 + d.tiles_per_split, d.MT);
                mbar_wait(&ctl->b_full, unit_it & 1);
                for (int mt = mt0; mt < mt1; ++mt, ++tile_it) {
                    const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
                    mbar_wait(&ctl->t_empty[buf], tph ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * kBlockN;
                    for (int kc = 0; kc < KC; ++kc, ++a_it) {
                        const uint32_t s = a_it % kAStages, ph = (a_it / kAStages) & 1;
                        mbar_wait(&ctl->a_full[s], ph);
                        tc_fence_after();
                        const uint32_t a_addr = a_base + s * kAStageBytes;
                        const uint32_t b_addr = b_base + kc * kBChunkBytes;
#pragma unroll
                        for (int k = 0; k < kChunkK / 16; ++k) {
                            umma_bf16_ss(d_tmem, umma_desc_k_sw128(a_addr + k * 32),
                                         umma_desc_k_sw128(b_addr + k * 32), idesc,
                                         (kc | k) != 0 ? 1u : 0u);
                        }
                        umma_commit(&ctl->a_empty[s]);   // frees the A stage once these MMAs retire
                    }
                    umma_commit(&ctl->t_full[buf]);      // accumulator complete
                }
                umma_commit(&ctl->b_empty);              // resident block may be overwritten
            }
        }
    }
    } else {
        reg_alloc<232>();
        // ===================================================== epilogue: 8 warps, 2 warpgroups
        const int ew = warp - kCtlWarps;      // 0..7
        const int wg = ew >> 2;               // column half
        const int q = warp & 3;               // TMEM lane quarter this warp may access
        const int et = threadIdx.x - kCtlWarps * 32;   // 0..255
        const uint64_t a2p = pack2(p.a2, p.a2);
        uint32_t tile_it = 0;
        for (int u = blockIdx.x; u < d.num_units; u += gridDim.x) {
            const int ms = u % d.S, nblk = (u / d.S) % d.NB, b = u / (d.S * d.NB);
            const int mt0 = ms * d.tiles_per_split;
            const int mt1 = min(mt0 + d.tiles_per_split, d.MT);
            const float maxn0 = p.maxnorm[b * 2 + 0], maxn1 = p.maxnorm[b * 2 + 1];
            const float cref = score_bound2(maxn0, maxn1, p.a2) - kHead2;
            // valid columns of this warpgroup's half of the block
            const int nvalid = max(0, min(128, d.N1 - nblk * kBlockN - wg * 128));
            float* rowpart = p.rowpart + (static_cast<size_t>(b) * d.NB * 2 + nblk * 2 + wg) * d.N0pad;
            uint2* cand = p.cand + static_cast<size_t>(u) * d.cand_cap;

            uint64_t colacc[64];
#pragma unroll
            for (int k = 0; k < 64; ++k) colacc[k] = 0ull;

            for (int mt = mt0; mt < mt1; ++mt, ++tile_it) {
                const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
                const int row = mt * kTileM + q * 32 + lane;
                const bool rvalid = row < d.N0;
                const float n0 = rvalid ? __ldg(p.norm0 + static_cast<size_t>(b) * d.N0pad + row) : 0.f;
                const float ci = score_bound2(n0, maxn1, p.a2) - kHead2;
                const float bias = rvalid ? -ci : -INFINITY;          // invalid row -> p = 0
                const float wi = rvalid ? ex2_approx(ci - cref) : 0.f;  // <= 1
                const uint64_t biasp = pack2(bias, bias), wp = pack2(wi, wi);

                mbar_wait(&ctl->t_full[buf], tph);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kBlockN + wg * 128;

                uint64_t racc0 = 0ull, racc1 = 0ull;
                float cm[4];
                uint32_t r[32];
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    const int nv = nvalid - ch * 32;
                    if (nv <= 0) { cm[ch] = 0.f; continue; }   // warp-uniform
                    tmem_ld_32x32(taddr + ch * 32, r);
                    tmem_ld_wait();
                    if (nv >= 32) process_chunk<false>(r, a2p, biasp, wp, colacc + ch * 16, racc0, racc1, cm[ch], 32);
                    else          process_chunk<true>(r, a2p, biasp, wp, colacc + ch * 16, racc0, racc1, cm[ch], nv);
                }
                float s0, s1, s2, s3;
                unpack2(racc0, s0, s1);
                unpack2(racc1, s2, s3);
                const float rowsum = (s0 + s1) + (s2 + s3);
                rowpart[mt * kTileM + q * 32 + lane] = rowsum;

                // ---- candidates: chunks whose maximum exceeds the share threshold (rare)
                const float share = p.cand_thr * rowsum;
                uint32_t hot = 0;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) hot |= (cm[ch] > share) ? (1u << ch) : 0u;
                const uint32_t warp_hot = __reduce_or_sync(0xffffffffu, hot);
                if (warp_hot) {
                    for (int ch = 0; ch < 4; ++ch) {
                        if (!((warp_hot >> ch) & 1)) continue;   // warp-uniform
                        tmem_ld_32x32(taddr + ch * 32, r);
                        tmem_ld_wait();
                        if ((hot >> ch) & 1) {
                            const int nv = min(32, nvalid - ch * 32);
#pragma unroll
                            for (int k = 0; k < 32; ++k) {
                                const float dot = __uint_as_float(r[k]);
                                const float pk = ex2_approx(fmaf(dot, p.a2, bias));
                                if (k < nv && pk > share) {
                                    const int slot = atomicAdd(&ctl->cand_count, 1);
                                    if (slot < d.cand_cap) {
                                        const uint32_t rin = static_cast<uint32_t>((mt - mt0) * kTileM + q * 32 + lane);
                                        cand[slot] = make_uint2((rin << 8) | static_cast<uint32_t>(wg * 128 + ch * 32 + k),
                                                                __float_as_uint(dot * p.a2));
                                    }
                                }
                            }
                        }
                    }
                }
                // release the accumulator buffer to the MMA warp
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl->t_empty[buf]);
            }

            // ---- unit end: reduce the thread-local column sums over the 32 rows of the warp
            float v[128];
#pragma unroll
            for (int k = 0; k < 64; ++k) unpack2(colacc[k], v[2 * k], v[2 * k + 1]);
            reduce_step<64, 16>(v, lane);
            reduce_step<32, 8>(v, lane);
            reduce_step<16, 4>(v, lane);
            reduce_step<8, 2>(v, lane);
            reduce_step<4, 1>(v, lane);
            {
                // lane now holds columns base .. base + 3 summed over the warp's 32 rows
                const int base = ((lane >> 4) & 1) * 64 + ((lane >> 3) & 1) * 32 + ((lane >> 2) & 1) * 16 +
                                 ((lane >> 1) & 1) * 8 + (lane & 1) * 4;
                *reinterpret_cast<float4*>(&ctl->colscratch[ew][base]) = make_float4(v[0], v[1], v[2], v[3]);
            }
            named_bar_sync(1, kEpiWarps * 32);
            {
                const int g = et >> 7, c = et & 127;
                const float s = (ctl->colscratch[g * 4 + 0][c] + ctl->colscratch[g * 4 + 1][c]) +
                                (ctl->colscratch[g * 4 + 2][c] + ctl->colscratch[g * 4 + 3][c]);
                p.colpart[(static_cast<size_t>(b) * d.S + ms) * d.N1pad + nblk * kBlockN + et] = s;
                if (et == 0) {
                    const int cnt = ctl->cand_count;
                    p.unit_count[u] = min(cnt, d.cand_cap);
                    if (cnt > d.cand_cap) atomicOr(p.status + kStFlags, FM_FLAG_CAND_OVERFLOW);
                    atomicMax(p.status + kStMaxUnitCand, cnt);
                    ctl->cand_count = 0;
                }
            }
            named_bar_sync(1, kEpiWarps * 32);
        }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// [B, N, C] bf16 row-major viewed as a 3-D tensor (C, N, B); box = 64 channels x `rows` tokens,
// 128-byte swizzle; rows past N are zero-filled and never bleed into the next pair.
bool make_feature_map(CUtensorMap* m, const __nv_bfloat16* ptr, int B, int N, int C, int rows) {
    const uint64_t dims[3] = {static_cast<uint64_t>(C), static_cast<uint64_t>(N), static_cast<uint64_t>(B)};
    const uint64_t strides[2] = {static_cast<uint64_t>(C) * 2, static_cast<uint64_t>(N) * C * 2};
    const uint32_t box[3] = {static_cast<uint32_t>(kChunkK), static_cast<uint32_t>(rows), 1};
    return encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, ptr, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace

// ---------------------------------------------------------------- host: tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

bool encode_tiled(CUtensorMap* m, CUtensorMapDataType dt, int rank, const void* ptr, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swizzle) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return false;
        fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    cuuint64_t d[5], st[4];
    cuuint32_t bx[5], es[5];
    for (int k = 0; k < rank; ++k) { d[k] = dims[k]; bx[k] = box[k]; es[k] = 1; }
    for (int k = 0; k + 1 < rank; ++k) st[k] = strides_bytes[k];
    return fn(m, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(ptr), d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

cudaError_t launch_coarse_stats_tc(const __nv_bfloat16* f0, const __nv_bfloat16* f1,
                                   const CoarseStatsParams& p, int num_sms, cudaStream_t st) {
    CUtensorMap tm0, tm1;
    if (!make_feature_map(&tm0, f0, p.d.B, p.d.N0, p.d.C, kTileM) ||
        !make_feature_map(&tm1, f1, p.d.B, p.d.N1, p.d.C, kBlockN)) {
        set_error("cuTensorMapEncodeTiled failed (feature pointers must be 16-byte aligned)");
        return cudaErrorInvalidValue;
    }
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(coarse_stats_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int grid = p.d.num_units < num_sms ? p.d.num_units : num_sms;
    coarse_stats_tc_kernel<<<grid, kCoarseThreads, kSmemBytes, st>>>(tm0, tm1, p);
    return cudaGetLastError();
}

}  // namespace fm
