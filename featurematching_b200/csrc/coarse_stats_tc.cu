// K1 -- coarse correlation fused with the dual-softmax statistics (sm_100a, tcgen05 + TMEM + TMA).
//
// Replaces, without ever writing the N0 x N1 matrix to memory,
//   sim  = einsum('nlc,nsc->nls', f0/sqrt(C), f1/sqrt(C)) / temperature   coarse_matching_new.py:64-66
//   conf = softmax(sim, 1) * softmax(sim, 2)                              coarse_matching_new.py:68
// by producing, per pair, the two soft-max denominators of every token (as partial sums that the
// resolve kernels finish) plus the short list of entries that can exceed the threshold.
//
// Work unit = (pair b, block of 256 image-1 tokens, contiguous range of 128-token image-0 tiles).
// The image-1 block is the RESIDENT MMA operand (B, K-major, 4 x 32 KB channel chunks, each with
// its own full/empty barrier so the next unit's block streams in behind the last tile's MMAs);
// the image-0 tiles stream through a 5-stage TMA ring as the A operand.  Each tile is one
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
//   hot  : a 32-column chunk whose largest p exceeds kCandMargin * thr of the row's own
//          128-column partial sum is a "hot chunk": its 32 raw dot products are re-read from
//          TMEM and stored (128 B) with (row, chunk) in the CTA's record pool.  conf > thr implies
//          share > thr of the FULL row sum, so the pools hold a superset of everything the
//          resolve step needs; K2 evaluates the 32 entries exactly.
// One MUFU.EX2 per score; scale/bias, row sum and column accumulate are packed f32x2 ops; the
// TMEM loads are double-buffered in registers (chunk k+1 in flight while chunk k is processed).
#include <cuda.h>

#include "fm_common.cuh"
#include "fm_ptx.cuh"
#include "fm_tmap.cuh"

namespace fm {

namespace {

constexpr int kKC = kMaxC / kChunkK;   // channel chunks of the resident operand (4)

struct __align__(16) SmemCtl {
    float colscratch[kEpiWarps][128];   // first: float4 stores need 16-byte alignment
    uint64_t a_full[kAStages];
    uint64_t a_empty[kAStages];
    uint64_t b_full[kKC];
    uint64_t b_empty[kKC];
    uint64_t t_full[2];
    uint64_t t_empty[2];
    uint32_t tmem_base;
    int hot_count;
};

constexpr size_t kSmemOperands = static_cast<size_t>(kBChunkBytes) * kKC +
                                 static_cast<size_t>(kAStageBytes) * kAStages;
constexpr size_t kSmemBytes = kSmemOperands + sizeof(SmemCtl) + 1024;  // + alignment slack
static_assert(kSmemBytes <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

// One 32-column chunk of the thread's accumulator row.  Returns the chunk's largest p (first pass) or, in the
// exact pass (kExact; `lc` = log2 column soft-max sums of the chunk's 32 columns), its largest
// 2 log2(p) - l2col_j = log2(conf_ij) + (l2row_i - 2 c_i): the confidence itself up to a per-row constant.
template <bool kExact>
__device__ __forceinline__ float process_chunk(const uint32_t (&r)[32], uint64_t a2p, uint64_t biasp,
                                               uint64_t wp, uint64_t* colacc, uint64_t& racc0,
                                               uint64_t& racc1, const float* __restrict__ lc) {
    float m = kExact ? -INFINITY : 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        uint64_t t2 = fma2(pack2(__uint_as_float(r[2 * k]), __uint_as_float(r[2 * k + 1])), a2p, biasp);
        float tl, th;
        unpack2(t2, tl, th);
        const float pl = ex2_approx(tl), ph = ex2_approx(th);
        const uint64_t p2 = pack2(pl, ph);
        if (k & 1) racc1 = add2(racc1, p2); else racc0 = add2(racc0, p2);
        colacc[k] = fma2(p2, wp, colacc[k]);
        if (kExact) m = max3(m, fmaf(2.f, tl, -__ldg(lc + 2 * k)), fmaf(2.f, th, -__ldg(lc + 2 * k + 1)));
        else m = max3(m, pl, ph);
    }
    return m;
}

// Columns >= nv of the chunk do not exist (N1 is not a multiple of the block): make their
// scores -inf so that p = 0.  Warp-uniform, only ever taken in the last block of a pair.
__device__ __forceinline__ void mask_chunk(uint32_t (&r)[32], int nv) {
#pragma unroll
    for (int k = 0; k < 32; ++k) r[k] = (k < nv) ? r[k] : 0xff800000u;
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

// kExact = the second extraction pass (see CoarseStatsParams::l2row): same sums, exact hot-chunk test.
template <bool kExact>
__global__ void __launch_bounds__(kCoarseThreads, 1)
coarse_stats_tc_kernel(const __grid_constant__ CUtensorMap tmap0,
                       const __grid_constant__ CUtensorMap tmap1, const CoarseStatsParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_b = smem;                                            // KC x 32 KB, resident
    uint8_t* smem_a = smem + static_cast<size_t>(kBChunkBytes) * kKC;  // ring
    SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem + kSmemOperands);

    const CoarseDims& d = p.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KC = d.KC;
    if (kExact && p.redo && *p.redo == 0) return;   // exact pass not needed (uniform over the grid)

    if (threadIdx.x == 0) {
        for (int s = 0; s < kAStages; ++s) { mbar_init(&ctl->a_full[s], 1); mbar_init(&ctl->a_empty[s], 1); }
        for (int s = 0; s < kKC; ++s) { mbar_init(&ctl->b_full[s], 1); mbar_init(&ctl->b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&ctl->t_full[s], 1); mbar_init(&ctl->t_empty[s], kEpiWarps); }
        ctl->hot_count = 0;
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
    // thread, the two epilogue warpgroups grow to 232 (128 column accumulators + two 32-word
    // TMEM load buffers each).
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
                for (int mt = mt0; mt < mt1; ++mt) {
                    for (int kc = 0; kc < KC; ++kc, ++a_it) {
                        if (mt == mt0) {
                            // chunk kc of the new resident block, as soon as the previous unit's
                            // last MMAs on that chunk have retired
                            mbar_wait(&ctl->b_empty[kc], (unit_it & 1) ^ 1);
                            mbar_arrive_expect_tx(&ctl->b_full[kc], kBChunkBytes);
                            tma_load_3d(smem_b + static_cast<size_t>(kc) * kBChunkBytes, &tmap1, &ctl->b_full[kc],
                                        kc * kChunkK, nblk * kBlockN, b);
                        }
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
                for (int mt = mt0; mt < mt1; ++mt, ++tile_it) {
                    const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
                    mbar_wait(&ctl->t_empty[buf], tph ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * kBlockN;
                    for (int kc = 0; kc < KC; ++kc, ++a_it) {
                        const uint32_t s = a_it % kAStages, ph = (a_it / kAStages) & 1;
                        if (mt == mt0) mbar_wait(&ctl->b_full[kc], unit_it & 1);
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
                        if (mt == mt1 - 1) umma_commit(&ctl->b_empty[kc]);   // chunk may be overwritten
                    }
                    umma_commit(&ctl->t_full[buf]);      // accumulator complete
                }
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
        uint2* const pool_meta = p.hot_meta + static_cast<size_t>(blockIdx.x) * d.hot_cap;
        float* const pool_dots = p.hot_dots + static_cast<size_t>(blockIdx.x) * d.hot_cap * kHotChunk;
        uint32_t tile_it = 0;
        for (int u = blockIdx.x; u < d.num_units; u += gridDim.x) {
            const int ms = u % d.S, nblk = (u / d.S) % d.NB, b = u / (d.S * d.NB);
            const int mt0 = ms * d.tiles_per_split;
            const int mt1 = min(mt0 + d.tiles_per_split, d.MT);
            const float maxn0 = p.maxnorm[b * 2 + 0], maxn1 = p.maxnorm[b * 2 + 1];
            const float cref = score_bound2(maxn0, maxn1, p.a2) - kHead2;
            // valid columns of this warpgroup's half of the block
            const int col0 = nblk * kBlockN + wg * 128;
            const int nvalid = max(0, min(128, d.N1 - col0));
            float* rowpart = p.rowpart + (static_cast<size_t>(b) * d.NB * 2 + nblk * 2 + wg) * d.N0pad;

            uint64_t colacc[64];
#pragma unroll
            for (int k = 0; k < 64; ++k) colacc[k] = 0ull;

            // token norm of this thread's row, fetched one tile ahead (N0pad rows exist)
            const float* norm_b = p.norm0 + static_cast<size_t>(b) * d.N0pad + q * 32 + lane;
            float n0_next = __ldg(norm_b + mt0 * kTileM);
            // exact pass: log2 soft-max sums of this pair's rows and of this warpgroup's 128 columns (the l2col
            // buffer has kBlockN floats of slack, columns >= N1 carry p = 0 and can never be hot)
            const float* l2row_b = kExact ? p.l2row + static_cast<size_t>(b) * d.N0 : nullptr;
            const float* lcol = kExact ? p.l2col + static_cast<size_t>(b) * d.N1 + col0 : nullptr;
            const float log2_cand = kExact ? log2f(p.cand_thr) : 0.f;
            for (int mt = mt0; mt < mt1; ++mt, ++tile_it) {
                const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
                const int row = mt * kTileM + q * 32 + lane;
                const bool rvalid = row < d.N0;
                const float n0 = rvalid ? n0_next : 0.f;
                if (mt + 1 < mt1) n0_next = __ldg(norm_b + (mt + 1) * kTileM);
                const float ci = score_bound2(n0, maxn1, p.a2) - kHead2;
                const float bias = rvalid ? -ci : -INFINITY;          // invalid row -> p = 0
                const float wi = rvalid ? ex2_approx(ci - cref) : 0.f;  // <= 1
                const uint64_t biasp = pack2(bias, bias), wp = pack2(wi, wi);

                mbar_wait(&ctl->t_full[buf], tph);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kBlockN + wg * 128;

                uint64_t racc0 = 0ull, racc1 = 0ull;
                float cm[4];
                uint32_t ra[32], rb[32];
                tmem_ld_32x32(taddr, ra);
                // chunk 0
                tmem_ld_wait_dep(ra);
                tmem_ld_32x32(taddr + 32, rb);
                if (nvalid < 32) mask_chunk(ra, nvalid);
                cm[0] = process_chunk<kExact>(ra, a2p, biasp, wp, colacc, racc0, racc1, lcol + 0);
                // chunk 1
                tmem_ld_wait_dep(rb);
                tmem_ld_32x32(taddr + 64, ra);
                if (nvalid < 64) mask_chunk(rb, nvalid - 32);
                cm[1] = process_chunk<kExact>(rb, a2p, biasp, wp, colacc + 16, racc0, racc1, lcol + 32);
                // chunk 2
                tmem_ld_wait_dep(ra);
                tmem_ld_32x32(taddr + 96, rb);
                if (nvalid < 96) mask_chunk(ra, nvalid - 64);
                cm[2] = process_chunk<kExact>(ra, a2p, biasp, wp, colacc + 32, racc0, racc1, lcol + 64);
                // chunk 3
                tmem_ld_wait_dep(rb);
                if (nvalid < 128) mask_chunk(rb, nvalid - 96);
                cm[3] = process_chunk<kExact>(rb, a2p, biasp, wp, colacc + 48, racc0, racc1, lcol + 96);

                float s0, s1, s2, s3;
                unpack2(racc0, s0, s1);
                unpack2(racc1, s2, s3);
                const float rowsum = (s0 + s1) + (s2 + s3);
                rowpart[mt * kTileM + q * 32 + lane] = rowsum;

                // ---- hot chunks: largest p above the share threshold of the row's partial sum
                // first pass: the chunk's best p against cand_thr x the row's own 128-column partial sum (weaker
                // than the share of the full row sum, hence a superset).  Exact pass: log2 conf_ij =
                // (2 log2 p - l2col_j) + 2 c_i - l2row_i against log2(cand_thr); a lost row (l2row = INF) is never hot
                const float share = kExact ? (rvalid ? log2_cand + __ldg(l2row_b + row) - 2.f * ci : INFINITY)
                                           : p.cand_thr * rowsum;
                uint32_t hot = 0;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) hot |= (cm[ch] > share) ? (1u << ch) : 0u;
                uint32_t warp_hot = __reduce_or_sync(0xffffffffu, hot);
                while (warp_hot) {                                   // warp-uniform, usually 0 or 1 trips
                    const int ch = __ffs(warp_hot) - 1;
                    warp_hot &= warp_hot - 1;
                    tmem_ld_32x32(taddr + ch * 32, ra);
                    tmem_ld_wait_dep(ra);
                    if ((hot >> ch) & 1) {
                        const int slot = atomicAdd(&ctl->hot_count, 1);
                        if (slot < d.hot_cap) {
                            pool_meta[slot] = make_uint2(static_cast<uint32_t>(b * d.N0 + row),
                                                         static_cast<uint32_t>((col0 >> 5) + ch));
                            uint4* dst = reinterpret_cast<uint4*>(pool_dots + static_cast<size_t>(slot) * kHotChunk);
#pragma unroll
                            for (int k = 0; k < 8; ++k)
                                dst[k] = make_uint4(ra[4 * k], ra[4 * k + 1], ra[4 * k + 2], ra[4 * k + 3]);
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
            }
            named_bar_sync(1, kEpiWarps * 32);
        }
        // ---- kernel end: publish the pool's record count
        named_bar_sync(1, kEpiWarps * 32);
        if (et == 0) {
            const int cnt = ctl->hot_count;
            p.hot_count[blockIdx.x] = min(cnt, d.hot_cap);
            if (cnt > d.hot_cap) atomicOr(p.status + kStFlags, FM_FLAG_CAND_OVERFLOW);
            atomicMax(p.status + kStMaxUnitCand, cnt);
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
    static unsigned long long attr_set = 0;
    if (needs_attr_on_this_device(&attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(coarse_stats_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(coarse_stats_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
    }
    (void)num_sms;
    const int grid = p.d.num_pools;   // one record pool per CTA
    if (p.l2row) coarse_stats_tc_kernel<true><<<grid, kCoarseThreads, kSmemBytes, st>>>(tm0, tm1, p);
    else coarse_stats_tc_kernel<false><<<grid, kCoarseThreads, kSmemBytes, st>>>(tm0, tm1, p);
    count_launch();
    return cudaGetLastError();
}

}  // namespace fm
