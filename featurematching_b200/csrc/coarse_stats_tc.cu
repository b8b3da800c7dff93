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
constexpr int kChunks = kGroupCols / 32;   // 32-column chunks of a thread's accumulator row
// registers per epilogue thread after the control warpgroup gave its share back (setmaxnreg): wide 8 warps x 232,
// narrow 12 warps x 152 (64 column accumulators + two 32-word TMEM load buffers + the rest)
constexpr int kEpiRegs = kColGroups == 2 ? 232 : 152;

// -DK1_TRACE: clock64 stamps of CTA 0 (epilogue warp 0 of column group 0, the MMA thread, the TMA thread) per tile
#ifdef K1_TRACE
constexpr int kTrTiles = 96, kTrPts = 8;
__device__ long long g_k1_trace[3][kTrTiles][kTrPts];
#define K1T(role, n, id) do { if (blockIdx.x == 0 && (n) < kTrTiles) g_k1_trace[role][n][id] = clock64(); } while (0)
#else
#define K1T(role, n, id) do {} while (0)
#endif

struct __align__(16) SmemCtl {
    float colscratch[kEpiWarps][kGroupCols];   // first: vector stores need 16-byte alignment
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

// 2^x for two scores on the FMA pipe (an OPTION, off by default): x = n + f, n = round(x), f in [-0.5, 0.5]; 2^f by a
// degree-4 minimax polynomial (max relative error 2.7e-6), 2^n by an integer add into the exponent field; scores
// clamped at -126.  K1_POLY_MASK: bit k = pair k of a 32-column chunk goes through here instead of MUFU.EX2.
// Measured (r02, 32 pairs 640x512): 0 / 12.5 / 25 / 50 % of the scores -> 0.412 / 0.404 / 0.414 / 0.443 ms: inside the
// chunk phase the XU pipe is 80 % busy, but the epilogue as a whole is not MUFU-bound -- 2 % at best,
// not worth a second exponential in the numerics.
#ifndef K1_POLY_MASK
#define K1_POLY_MASK 0x0000u
#endif
__device__ __forceinline__ void ex2_poly2(uint64_t x2, float& pl, float& ph) {
    float xl, xh;
    unpack2(x2, xl, xh);
    const uint64_t xc = pack2(fmaxf(xl, -126.f), fmaxf(xh, -126.f));
    const uint64_t magic = pack2(12582912.f, 12582912.f);          // 1.5 * 2^23: the sum's low mantissa bits = round(x)
    const uint64_t t = add2(xc, magic);
    const uint64_t f = sub2(xc, sub2(t, magic));
    uint64_t q = fma2(pack2(0.009570082649588585f, 0.009570082649588585f), f, pack2(0.05591784417629242f, 0.05591784417629242f));
    q = fma2(q, f, pack2(0.240247443318367f, 0.240247443318367f));
    q = fma2(q, f, pack2(0.6931217908859253f, 0.6931217908859253f));
    q = fma2(q, f, pack2(0.9999992847442627f, 0.9999992847442627f));
    float ql, qh, tl, th;
    unpack2(q, ql, qh);
    unpack2(t, tl, th);
    pl = __uint_as_float(__float_as_uint(ql) + (__float_as_uint(tl) << 23));
    ph = __uint_as_float(__float_as_uint(qh) + (__float_as_uint(th) << 23));
}

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
        float tl, th, pl, ph;
        unpack2(t2, tl, th);
        if ((K1_POLY_MASK >> k) & 1u) {
            ex2_poly2(t2, pl, ph);
        } else {
            pl = ex2_approx(tl); ph = ex2_approx(th);
        }
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
                    K1T(1, tile_it, 0);
                    mbar_wait(&ctl->t_empty[buf], tph ^ 1);
                    tc_fence_after();
                    K1T(1, tile_it, 1);
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
                    K1T(1, tile_it, 2);
                }
            }
        }
    }
    } else {
        reg_alloc<kEpiRegs>();
        // ===================================================== epilogue: one warpgroup per column group
        const int ew = warp - kCtlWarps;      // 0 .. kEpiWarps - 1
        const int wg = ew >> 2;               // column group
        const int q = warp & 3;               // TMEM lane quarter this warp may access
        const int et = threadIdx.x - kCtlWarps * 32;   // 0 .. kEpiWarps * 32 - 1
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
            // valid columns of this warpgroup's column group
            const int col0 = nblk * kBlockN + wg * kGroupCols;
            const int nvalid = max(0, min(kGroupCols, d.N1 - col0));
            float* rowpart = p.rowpart + (static_cast<size_t>(b) * d.NB * kColGroups + nblk * kColGroups + wg) * d.N0pad;

            uint64_t colacc[kGroupCols / 2];
#pragma unroll
            for (int k = 0; k < kGroupCols / 2; ++k) colacc[k] = 0ull;

            // token norm of this thread's row, fetched one tile ahead (N0pad rows exist)
            const float* norm_b = p.norm0 + static_cast<size_t>(b) * d.N0pad + q * 32 + lane;
            float n0_next = __ldg(norm_b + mt0 * kTileM);
            // exact pass: log2 soft-max sums of this pair's rows and of this warpgroup's columns (the l2col
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

                const bool k1tr = ew == 0 && lane == 0;
                if (k1tr) K1T(0, tile_it, 0);
                mbar_wait(&ctl->t_full[buf], tph);
                tc_fence_after();
                if (k1tr) K1T(0, tile_it, 1);
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kBlockN + wg * kGroupCols;

                uint64_t racc0 = 0ull, racc1 = 0ull;
                float cm[kChunks];
                uint32_t ra[32], rb[32];
                tmem_ld_32x32(taddr, ra);
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) {   // chunk ch + 1 is in flight while chunk ch is processed
                    uint32_t (&cur)[32] = (ch & 1) ? rb : ra;
                    uint32_t (&nxt)[32] = (ch & 1) ? ra : rb;
                    tmem_ld_wait_dep(cur);
                    if (ch + 1 < kChunks) tmem_ld_32x32(taddr + 32 * (ch + 1), nxt);
                    if (nvalid < 32 * (ch + 1)) mask_chunk(cur, nvalid - 32 * ch);
                    cm[ch] = process_chunk<kExact>(cur, a2p, biasp, wp, colacc + 16 * ch, racc0, racc1, lcol + 32 * ch);
                }

                if (k1tr) K1T(0, tile_it, 2);
                float s0, s1, s2, s3;
                unpack2(racc0, s0, s1);
                unpack2(racc1, s2, s3);
                const float rowsum = (s0 + s1) + (s2 + s3);
                rowpart[mt * kTileM + q * 32 + lane] = rowsum;

                // ---- hot chunks: largest p above the share threshold of the row's partial sum
                // first pass: the chunk's best p against cand_thr x the row's own column-group partial sum (weaker
                // than the share of the full row sum, hence a superset).  Exact pass: log2 conf_ij =
                // (2 log2 p - l2col_j) + 2 c_i - l2row_i against log2(cand_thr); a lost row (l2row = INF) is never hot
                const float share = kExact ? (rvalid ? log2_cand + __ldg(l2row_b + row) - 2.f * ci : INFINITY)
                                           : p.cand_thr * rowsum;
                uint32_t hot = 0;
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) hot |= (rvalid && cm[ch] > share) ? (1u << ch) : 0u;
                uint32_t warp_hot = __reduce_or_sync(0xffffffffu, hot);
                auto store_record = [&](int ch, const uint32_t (&raw)[32]) {
                    if ((hot >> ch) & 1) {
                        const int slot = atomicAdd(&ctl->hot_count, 1);
                        if (slot < d.hot_cap) {
                            pool_meta[slot] = make_uint2(static_cast<uint32_t>(b * d.N0 + row),
                                                         static_cast<uint32_t>((col0 >> 5) + ch));
                            uint4* dst = reinterpret_cast<uint4*>(pool_dots + static_cast<size_t>(slot) * kHotChunk);
#pragma unroll
                            for (int k = 0; k < 8; ++k)
                                dst[k] = make_uint4(raw[4 * k], raw[4 * k + 1], raw[4 * k + 2], raw[4 * k + 3]);
                        }
                    }
                };
                // The LAST hot chunk (usually the only one) is only re-read from TMEM here; its record (a shared-memory
                // atomic and 136 bytes of global stores from one or two lanes) is written AFTER the accumulator has
                // been handed back: every tile has some warp with a hot chunk, and the MMA thread waits for the
                // slowest of the eight (r02 timeline: the epilogue, not the MMA, sets the tile period).
                int last_ch = -1;
                while (warp_hot) {                                   // warp-uniform, usually 0 or 1 trips
                    const int ch = __ffs(warp_hot) - 1;
                    warp_hot &= warp_hot - 1;
                    tmem_ld_32x32(taddr + ch * 32, ra);
                    tmem_ld_wait_dep(ra);
                    if (warp_hot) store_record(ch, ra); else last_ch = ch;
                }
                if (k1tr) K1T(0, tile_it, 3);
                // release the accumulator buffer to the MMA warp
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl->t_empty[buf]);
                if (last_ch >= 0) store_record(last_ch, ra);
                if (k1tr) K1T(0, tile_it, 4);
            }

            // ---- unit end: reduce the thread-local column sums over the 32 rows of the warp
            float v[kGroupCols];
#pragma unroll
            for (int k = 0; k < kGroupCols / 2; ++k) unpack2(colacc[k], v[2 * k], v[2 * k + 1]);
            reduce_step<kGroupCols / 2, 16>(v, lane);
            reduce_step<kGroupCols / 4, 8>(v, lane);
            reduce_step<kGroupCols / 8, 4>(v, lane);
            reduce_step<kGroupCols / 16, 2>(v, lane);
            reduce_step<kGroupCols / 32, 1>(v, lane);
            {
                // lane now holds columns base .. base + kGroupCols / 32 - 1 summed over the warp's 32 rows
                constexpr int kPer = kGroupCols / 32;
                const int base = (((lane >> 4) & 1) * 16 + ((lane >> 3) & 1) * 8 + ((lane >> 2) & 1) * 4 +
                                  ((lane >> 1) & 1) * 2 + (lane & 1)) * kPer;
#pragma unroll
                for (int k = 0; k < kPer; ++k) ctl->colscratch[ew][base + k] = v[k];
            }
            named_bar_sync(1, kEpiWarps * 32);
            if (et < kBlockN) {
                const int g = et / kGroupCols, c = et % kGroupCols;
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

#ifdef K1_TRACE
extern "C" __attribute__((visibility("default"))) int fm_debug_k1_trace(long long* host_out) {
    return static_cast<int>(cudaMemcpyFromSymbol(host_out, g_k1_trace, sizeof(g_k1_trace)));
}
#endif

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
