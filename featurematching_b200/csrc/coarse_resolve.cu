// K2 -- finishes the coarse stage from the statistics K1 left in the workspace:
//   * soft-max denominators of every token (log2 domain)              coarse_matching_new.py:68
//   * confidence of every candidate, threshold, per-row / per-column maxima (mutual nearest
//     neighbour on the CONFIDENCE, equality semantics)                coarse_matching_new.py:99-106
//   * border removal                                                  coarse_matching_new.py:11-28,101
//   * match list in torch.where order (b, i, j) + coarse pixel coords coarse_matching_new.py:109-134
// All of it touches O(N) data; the N0 x N1 matrix does not exist.
#include <algorithm>

#include <cooperative_groups.h>

#include "fm_common.cuh"
#include "fm_ptx.cuh"

namespace fm {

namespace cg = cooperative_groups;

namespace {

#ifdef RESOLVE_TRACE
__device__ unsigned long long g_blk[4096][4];
#endif


// ---------------------------------------------------------------- 1. log-sum-exp of every token
__global__ void __launch_bounds__(256) finish_lse_kernel(CoarseDims d, float a2, const float* __restrict__ norm0,
                                                         const float* __restrict__ maxnorm,
                                                         const float* __restrict__ rowpart,
                                                         const float* __restrict__ colpart,
                                                         float* __restrict__ l2row, float* __restrict__ l2col,
                                                         int* __restrict__ status, int* __restrict__ redo,
                                                         const float* __restrict__ hot_dots,
                                                         const int* __restrict__ hot_count) {
    const int b = blockIdx.y;
    {   // L2 prefetch of the candidate records the resolve kernel walks next: K1 wrote them while 168 MB of
        // features streamed through the cache, and the walk is a latency-bound gather (a hint: partial is fine)
        const int g = (blockIdx.y * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        const int per_pool = max(1, static_cast<int>(gridDim.x * gridDim.y * blockDim.x) / d.num_pools);
        const int pool = g / per_pool, r = g - pool * per_pool;
        if (pool < d.num_pools)
            for (int q = r; q < min(hot_count[pool], d.hot_cap); q += per_pool)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(hot_dots + (static_cast<size_t>(pool) * d.hot_cap + q) * kHotChunk));
    }
    // Gate of the exact extraction pass (K1 has finished: its flag is final).  If the single-pass candidate
    // filter overflowed its record pools, the statistics kernel runs once more with the full row sums computed
    // below as its hot-chunk reference; the flag is cleared so that only an overflow of THAT pass reaches the host.
    if (redo && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
        const int need = (status[kStFlags] & FM_FLAG_CAND_OVERFLOW) ? 1 : 0;
        *redo = need;
        if (need) {
            atomicAnd(status + kStFlags, ~FM_FLAG_CAND_OVERFLOW);
            status[kStRedo] = 1;
        }
    }
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const float maxn0 = maxnorm[b * 2 + 0], maxn1 = maxnorm[b * 2 + 1];
    int lost = 0;
    if (t < d.N0) {
        const float* rp = rowpart + static_cast<size_t>(b) * d.NB * kColGroups * d.N0pad + t;
        float s = 0.f;
        for (int k = 0; k < d.NB * kColGroups; ++k) s += rp[static_cast<size_t>(k) * d.N0pad];
        const float ci = score_bound2(norm0[static_cast<size_t>(b) * d.N0pad + t], maxn1, a2) - kHead2;
        // a sum below 2^-100 has lost relative precision to flush-to-zero: the token is "lost"
        const bool ok = s >= 7.8886e-31f && s < INFINITY;
        l2row[static_cast<size_t>(b) * d.N0 + t] = ok ? ci + log2f(s) : INFINITY;
        lost += ok ? 0 : 1;
    }
    if (t < d.N1) {
        const float* cp = colpart + static_cast<size_t>(b) * d.S * d.N1pad + t;
        float s = 0.f;
        for (int k = 0; k < d.S; ++k) s += cp[static_cast<size_t>(k) * d.N1pad];
        const float cref = score_bound2(maxn0, maxn1, a2) - kHead2;
        const bool ok = s >= 7.8886e-31f && s < INFINITY;
        l2col[static_cast<size_t>(b) * d.N1 + t] = ok ? cref + log2f(s) : INFINITY;
        lost += ok ? 0 : 1;
    }
    lost = __reduce_add_sync(0xffffffffu, lost);
    if ((threadIdx.x & 31) == 0 && lost) {
        atomicAdd(status + kStLost, lost);
        atomicOr(status + kStFlags, FM_FLAG_RANGE);
    }
}

// ---------------------------------------------------------------- 2. candidates
// K1 leaves "hot chunk" records in one pool per CTA: (row, first column / 32) + the 32 raw dot
// products.  One warp per record, one lane per entry; entries that do not exist (j >= N1) or
// whose confidence is below the threshold drop out here.
// conf = softmax_j * softmax_i = 2^(t - lr) * 2^(t - lc), t = a2 * dot.  ONE definition (operations pinned against
// contraction) for every walk: the mutual-NN test compares confidences for bit equality.
__device__ __forceinline__ float cand_conf(float dot, float a2, float lr, float lc) {
    const float t = __fmul_rn(dot, a2);
    return ex2_approx(__fadd_rn(__fsub_rn(t, lr), __fsub_rn(t, lc)));
}

__device__ __forceinline__ bool interior(int idx, int h, int w, int border) {
    if (border <= 0) return true;
    const int r = idx / w, c = idx - r * w;
    return r >= border && r < h - border && c >= border && c < w - border;
}

// ---------------------------------------------------------------- 3. ordered compaction helpers
__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
    __shared__ int warp_sums[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    __syncthreads();   // warp_sums of the previous call have been read by everyone
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = (lane < (blockDim.x >> 5)) ? warp_sums[lane] : 0;
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += n;
        }
        warp_sums[lane] = wi - w;   // exclusive warp offsets
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    return incl - v + warp_sums[warp];
}

// matches a row emits: all of them, unless the shared tie list itself overflowed (flagged): then only the slots
__device__ __forceinline__ int row_emit_count(int cnt, int ties) { return ties > kTieCap ? min(cnt, kRowSlots) : cnt; }

__device__ __forceinline__ void emit_match(const CoarseResolveParams& p, int pos, int b, int i, int j, int conf_bits,
                                           int i_r, int i_c, float s0x, float s0y, float s1x, float s1y) {
    const int j_r = j / p.w1c, j_c = j - j_r * p.w1c;
    p.b_ids[pos] = b;
    p.i_ids[pos] = i;
    p.j_ids[pos] = j;
    p.mconf[pos] = __int_as_float(conf_bits);
    p.mkpts0_c[2 * pos] = static_cast<float>(i_c) * s0x;
    p.mkpts0_c[2 * pos + 1] = static_cast<float>(i_r) * s0y;
    p.mkpts1_c[2 * pos] = static_cast<float>(j_c) * s1x;
    p.mkpts1_c[2 * pos + 1] = static_cast<float>(j_r) * s1y;
}

// the n matches of row r (ascending j) to positions pos0 .. pos0 + n - 1
__device__ __forceinline__ void emit_row(const CoarseResolveParams& p, int r, int n, int pos0, int ties,
                                         const int2* __restrict__ rowslot, const int4* __restrict__ tie_list) {
    const int b = r / p.d.N0, i = r - b * p.d.N0;
    float s0x = p.scale, s0y = p.scale, s1x = p.scale, s1y = p.scale;
    if (p.scale0) { s0x = p.scale * p.scale0[b * 2]; s0y = p.scale * p.scale0[b * 2 + 1]; }
    if (p.scale1) { s1x = p.scale * p.scale1[b * 2]; s1y = p.scale * p.scale1[b * 2 + 1]; }
    const int i_r = i / p.w0c, i_c = i - i_r * p.w0c;
    if (n > kRowSlots) {
        // rare: this row's matches are its slots plus its entries of the shared list; emit in ascending j by
        // repeated selection (no j repeats within a row)
        int last = -1;
        for (int k = 0; k < n; ++k) {
            int best_j = 0x7fffffff, best_bits = 0;
            for (int q = 0; q < kRowSlots; ++q) {
                const int2 e = rowslot[static_cast<size_t>(r) * kRowSlots + q];
                if (e.x > last && e.x < best_j) { best_j = e.x; best_bits = e.y; }
            }
            for (int q = 0; q < ties; ++q) {
                const int4 e = tie_list[q];
                if (e.x == r && e.y > last && e.y < best_j) { best_j = e.y; best_bits = e.z; }
            }
            if (best_j == 0x7fffffff) break;
            last = best_j;
            if (pos0 + k < p.capacity) emit_match(p, pos0 + k, b, i, best_j, best_bits, i_r, i_c, s0x, s0y, s1x, s1y);
        }
        return;
    }
    int2 s[kRowSlots];
#pragma unroll
    for (int k = 0; k < kRowSlots; ++k) s[k] = k < n ? rowslot[static_cast<size_t>(r) * kRowSlots + k] : make_int2(0x7fffffff, 0);
    // sort by j (at most 4 entries)
#pragma unroll
    for (int a = 0; a < kRowSlots - 1; ++a)
#pragma unroll
        for (int c = 0; c < kRowSlots - 1 - a; ++c)
            if (s[c].x > s[c + 1].x) { const int2 t = s[c]; s[c] = s[c + 1]; s[c + 1] = t; }
#pragma unroll
    for (int k = 0; k < kRowSlots; ++k) {
        if (k >= n) break;
        const int pos = pos0 + k;
        if (pos >= p.capacity) break;
        emit_match(p, pos, b, i, s[k].x, s[k].y, i_r, i_c, s0x, s0y, s1x, s1y);
    }
}

// ---------------------------------------------------------------- the resolve: ONE cooperative launch
// Four phases separated by grid-wide barriers (r01 ran them as five launches: 73 us of kernel time plus the launch
// gaps for a few MB of data; the candidate walks were chains of dependent loads, 6.6 records per warp one after
// the other -- here four records are in flight per warp):
//   1  every candidate's confidence -> per-row / per-column maxima (atomicMax on the float bits)
//   2  the same walk: threshold, equality with both maxima (mutual nearest neighbour on the confidence, all exact
//      ties survive), border -> up to kRowSlots (j, conf) per row, the rest to the shared tie list
//   3  per-block sums of the rows' match counts (rows in contiguous ranges, one range per block)
//   4  exclusive offsets (every block adds up the sums before it) -> matches written in torch.where order
constexpr int kResolveThreads = 512;

struct ResolveWs {
    const uint2* meta; const float* dots; const int* hot_count;
    const float* l2row; const float* l2col;
    unsigned* rowbest; unsigned* colbest;
    int* rowcnt; int2* rowslot; int* tie_count; int4* tie_list; int* blocksum;
};

// entries with conf > thr found by the first walk, kept in shared memory for the second phase: (row, col, j, bits),
// one segment per warp (filled through a warp-uniform register counter: no atomics)
constexpr int kWarpSurvivors = 128;
constexpr int kResolveWarps = kResolveThreads / 32;

__device__ __forceinline__ void select_one(const CoarseResolveParams& p, const ResolveWs& w, int row, int j, unsigned bits,
                                           int k) {
    if (k < kRowSlots) {
        w.rowslot[static_cast<size_t>(row) * kRowSlots + k] = make_int2(j, static_cast<int>(bits));
    } else {
        // more exactly tied matches in this row than slots: the reference's equality test keeps them all
        // (coarse_matching_new.py:105-109), so do we, through one shared list
        const int t = atomicAdd(w.tie_count, 1);
        if (t < kTieCap) w.tie_list[t] = make_int4(row, j, static_cast<int>(bits), 0);
        else atomicOr(p.status + kStFlags, FM_FLAG_TIE_OVERFLOW);
    }
}

// The records of all pools form one index space (prefix sums of the pools' counts in shared memory), dealt evenly to
// the warps of the grid.  A warp takes at most 32 consecutive records at a time: lane l decodes record l (pool,
// row, first column, the row's soft-max sum: one binary search, one division and two loads per RECORD -- the
// lane-redundant walk issued ~80 instructions per lane and record and was issue-bound at 19 us), then the records'
// 32 entries are evaluated for the whole chunk side by side, the per-record words broadcast by shuffles.
//   kSelect = false (phase 1): confidences above the threshold -> row / column maxima + the warp's survivor list
//   kSelect = true  (phase 2, only for a block whose survivor list overflowed: diffuse rows under a tiny
//                    threshold): the same walk again, equality with both maxima + border -> row slots
constexpr int kMaxPools = kResolveThreads;
constexpr int kWalkChunk = 12;   // records a warp evaluates side by side (2 registers per record)

template <bool kSelect>
__device__ __forceinline__ void walk_records(const CoarseResolveParams& p, const ResolveWs& w, const int* s_pref,
                                             int4* surv, int* surv_cnt) {
    const CoarseDims& d = p.d;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = s_pref[d.num_pools];
    const int nwarps = static_cast<int>(gridDim.x) * kResolveWarps;
    const int per = (total + nwarps - 1) / nwarps;                 // records of this warp: [g_lo, g_hi)
    const int nchunks = (per + kWalkChunk - 1) / kWalkChunk, chunk = nchunks ? (per + nchunks - 1) / nchunks : 0;
    const int gw = static_cast<int>(blockIdx.x) * kResolveWarps + warp;
    const int g_lo = min(total, gw * per), g_hi = min(total, g_lo + per);
    int nsurv = 0;   // warp-uniform
    int4* seg = surv + warp * kWarpSurvivors;
    for (int g0 = g_lo; g0 < g_hi; g0 += chunk) {
        // Three load levels per chunk of <= kWalkChunk records:
        //   0  the record words (one per lane) and, their addresses needing nothing but the record index, every
        //      record's 32 dot products
        //   1  the rows' soft-max sums (one per lane) and every record's 32 column sums
        //   2  nothing left to load: confidences
        const int cnt = min(chunk, g_hi - g0);
        const int g = g0 + min(lane, cnt - 1);
        int lo = 0, hi = d.num_pools;              // last pool with s_pref[pool] <= g
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (s_pref[mid] <= g) lo = mid; else hi = mid;
        }
        const int my_rec = lo * d.hot_cap + (g - s_pref[lo]);   // fits int: the pools hold < 2^31 / 32 records
        const uint2 m = w.meta[my_rec];
        float dot[kWalkChunk], lc[kWalkChunk];
#pragma unroll
        for (int u = 0; u < kWalkChunk; ++u)
            dot[u] = w.dots[static_cast<size_t>(__shfl_sync(0xffffffffu, my_rec, min(u, cnt - 1))) * kHotChunk + lane];
        const int b = min(static_cast<int>(m.x / static_cast<uint32_t>(d.N0)), d.B - 1);
        const int i = min(static_cast<int>(m.x - static_cast<uint32_t>(b) * d.N0), d.N0 - 1);
        const int my_row = b * d.N0 + i, my_j0 = static_cast<int>(m.y) * kHotChunk;
        const int my_col0 = b * d.N1 + my_j0;
        const float my_lr = w.l2row[my_row];
#pragma unroll
        for (int u = 0; u < kWalkChunk; ++u) {
            const int src = min(u, cnt - 1);
            const int jlim = d.N1 - __shfl_sync(0xffffffffu, my_j0, src);   // entries lane < jlim exist
            lc[u] = w.l2col[__shfl_sync(0xffffffffu, my_col0, src) + min(lane, jlim - 1)];
        }
#pragma unroll
        for (int u = 0; u < kWalkChunk; ++u) {
            const int src = min(u, cnt - 1);
            const float lr = __shfl_sync(0xffffffffu, my_lr, src);
            const int jlim = d.N1 - __shfl_sync(0xffffffffu, my_j0, src);
            const float conf = cand_conf(dot[u], p.a2, lr, lc[u]);
            const bool pass = u < cnt && lane < jlim && conf > p.thr;   // positive floats order like their bits
            const unsigned ballot = __ballot_sync(0xffffffffu, pass);
            if (ballot) {   // warp-uniform
                const int row = __shfl_sync(0xffffffffu, my_row, src);
                const int col = __shfl_sync(0xffffffffu, my_col0, src) + lane;
                const int j = __shfl_sync(0xffffffffu, my_j0, src) + lane;
                const unsigned bits = __float_as_uint(conf);
                if (!kSelect) {
                    if (pass) {
                        atomicMax(w.rowbest + row, bits);
                        atomicMax(w.colbest + col, bits);
                        const int q = nsurv + __popc(ballot & ((1u << lane) - 1u));
                        if (q < kWarpSurvivors) seg[q] = make_int4(row, col, j, static_cast<int>(bits));
                    }
                    nsurv += __popc(ballot);
                } else if (pass && bits == w.rowbest[row] && bits == w.colbest[col] &&
                           interior(row - (row / d.N0) * d.N0, p.h0c, p.w0c, p.border) && interior(j, p.h1c, p.w1c, p.border)) {
                    select_one(p, w, row, j, bits, atomicAdd(w.rowcnt + row, 1));
                }
            }
        }
    }
    if (!kSelect && lane == 0) surv_cnt[warp] = nsurv;
}

__global__ void __launch_bounds__(kResolveThreads, 2) coarse_resolve_kernel(const CoarseResolveParams p, const ResolveWs w) {
    cg::grid_group grid = cg::this_grid();
    __shared__ int s_total;
    __shared__ int s_carry;
#ifdef RESOLVE_TRACE
    unsigned long long tr[8];
#define RTR(k) do { if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) tr[k] = global_timer_ns(); } while (0)
#else
#define RTR(k) do {} while (0)
#endif
    RTR(0);
#ifdef RESOLVE_TRACE
    if (threadIdx.x == 0) g_blk[blockIdx.x][0] = global_timer_ns();
#endif
    __shared__ int4 s_surv[kResolveWarps * kWarpSurvivors];
    __shared__ int s_surv_cnt[kResolveWarps];
    __shared__ int s_pref[kMaxPools + 1];
    {   // prefix sums of the pools' record counts
        const int mine = static_cast<int>(threadIdx.x) < p.d.num_pools ? min(w.hot_count[threadIdx.x], p.d.hot_cap) : 0;
        const int ex = block_exclusive_scan(mine, &s_total);
        if (static_cast<int>(threadIdx.x) < p.d.num_pools) s_pref[threadIdx.x] = ex;
        if (threadIdx.x == 0) {
            s_pref[p.d.num_pools] = s_total;
            if (blockIdx.x == 0) p.status[kStCandTotal] = s_total;
        }
        __syncthreads();
    }
#ifdef RESOLVE_TRACE
    if (threadIdx.x == 0) g_blk[blockIdx.x][1] = global_timer_ns();
#endif
    walk_records<false>(p, w, s_pref, s_surv, s_surv_cnt);
#ifdef RESOLVE_TRACE
    __syncthreads();
    if (threadIdx.x == 0) g_blk[blockIdx.x][2] = global_timer_ns();
#endif
    RTR(1);
    grid.sync();
    RTR(2);
    // ---- 2. the block's own survivors of phase 1, one per thread: two load levels instead of a second walk of
    // the records (which only runs if a warp's list overflowed)
    {
        bool over = false;
        for (int k = 0; k < kResolveWarps; ++k) over = over || s_surv_cnt[k] > kWarpSurvivors;
        if (!over) {
            for (int q = threadIdx.x; q < kResolveWarps * kWarpSurvivors; q += kResolveThreads) {
                if ((q % kWarpSurvivors) >= s_surv_cnt[q / kWarpSurvivors]) continue;
                const int4 e = s_surv[q];
                const unsigned bits = static_cast<unsigned>(e.w);
                if (bits != w.rowbest[e.x] || bits != w.colbest[e.y]) continue;
                const int b = e.x / p.d.N0, i = e.x - b * p.d.N0;
                if (!interior(i, p.h0c, p.w0c, p.border) || !interior(e.z, p.h1c, p.w1c, p.border)) continue;
                select_one(p, w, e.x, e.z, bits, atomicAdd(w.rowcnt + e.x, 1));
            }
        } else {
            walk_records<true>(p, w, s_pref, s_surv, s_surv_cnt);
        }
    }
    RTR(3);
    grid.sync();
    RTR(4);
    // ---- 3. this block's rows: [row0, row1)
    const int nrows = p.d.B * p.d.N0;
    const int rpb = (nrows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const int row0 = min(nrows, static_cast<int>(blockIdx.x) * rpb), row1 = min(nrows, row0 + rpb);
    const int ties = *w.tie_count;
    {
        int v = 0;
        for (int r = row0 + threadIdx.x; r < row1; r += kResolveThreads) v += row_emit_count(w.rowcnt[r], ties);
        block_exclusive_scan(v, &s_total);
        if (threadIdx.x == 0) w.blocksum[blockIdx.x] = s_total;
    }
    RTR(5);
    grid.sync();
    RTR(6);
    // ---- 4. offsets and emission
    {
        int before = 0, all = 0;
        for (int k = threadIdx.x; k < static_cast<int>(gridDim.x); k += kResolveThreads) {
            const int bs = w.blocksum[k];
            all += bs;
            if (k < static_cast<int>(blockIdx.x)) before += bs;
        }
        block_exclusive_scan(before, &s_total);
        if (threadIdx.x == 0) s_carry = s_total;
        if (blockIdx.x == 0) {   // the total match count
            block_exclusive_scan(all, &s_total);
            if (threadIdx.x == 0) {
                p.status[kStCount] = min(s_total, p.capacity);
                if (s_total > p.capacity) atomicOr(p.status + kStFlags, FM_FLAG_OUT_OVERFLOW);
            }
        }
        __syncthreads();
    }
    for (int base = row0; base < row1; base += kResolveThreads) {
        const int r = base + threadIdx.x;
        const int n = r < row1 ? row_emit_count(w.rowcnt[r], ties) : 0;
        const int pos0 = s_carry + block_exclusive_scan(n, &s_total);
        if (n) emit_row(p, r, n, pos0, ties, w.rowslot, w.tie_list);
        __syncthreads();
        if (threadIdx.x == 0) s_carry += s_total;
        __syncthreads();
    }
    RTR(7);
#ifdef RESOLVE_TRACE
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    {
        unsigned long long s0 = ~0ull, s1 = 0, wmin = ~0ull, wmax = 0, pmax = 0, e1 = 0;
        for (unsigned k = 0; k < gridDim.x; ++k) {
            s0 = min(s0, g_blk[k][0]); s1 = max(s1, g_blk[k][0]);
            const unsigned long long wk = g_blk[k][2] - g_blk[k][1];
            wmin = min(wmin, wk); wmax = max(wmax, wk);
            pmax = max(pmax, g_blk[k][1] - g_blk[k][0]); e1 = max(e1, g_blk[k][2]);
        }
        printf("blocks: start spread %llu ns, prefix max %llu, walk min %llu max %llu, last end %llu after first start\n", s1 - s0, pmax, wmin, wmax, e1 - s0);
    }
        printf("resolve ns: walk1 %llu sync %llu walk2 %llu sync %llu rows %llu sync %llu emit %llu (grid %d)\n", tr[1] - tr[0],
               tr[2] - tr[1], tr[3] - tr[2], tr[4] - tr[3], tr[5] - tr[4], tr[6] - tr[5], tr[7] - tr[6], gridDim.x);
    }
#endif
}

}  // namespace

// Step 1 of the resolve: soft-max denominators of every token + the gate of the exact extraction pass
// (`with_gate`: w.redo is written; the caller then launches the statistics kernel once more, conditionally).
cudaError_t launch_coarse_finish_lse(const CoarseResolveParams& p, const CoarseWorkspace& w, bool with_gate,
                                     cudaStream_t st) {
    const CoarseDims& d = p.d;
    const int nmax = d.N0 > d.N1 ? d.N0 : d.N1;
    finish_lse_kernel<<<dim3((nmax + 255) / 256, d.B), 256, 0, st>>>(d, p.a2, w.norm0, w.maxnorm, w.rowpart,
                                                                     w.colpart, w.l2row, w.l2col, p.status,
                                                                     with_gate ? w.redo : nullptr, w.hot_dots, w.hot_count); count_launch();
    return cudaGetLastError();
}

cudaError_t launch_coarse_resolve(const CoarseResolveParams& p, const CoarseWorkspace& w, cudaStream_t st) {
    const CoarseDims& d = p.d;
    // co-resident grid (cooperative launch): as many virtual blocks per record pool as fit
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, coarse_resolve_kernel, kResolveThreads, 0);
    if (e != cudaSuccess) return e;
    if (d.num_pools > kMaxPools) { set_error("too many candidate pools for the resolve kernel"); return cudaErrorInvalidValue; }
    // the kernel indexes records, rows and columns with 32-bit integers
    if (1LL * d.num_pools * d.hot_cap >= (1LL << 31) || 1LL * d.B * d.N0 >= (1LL << 31) || 1LL * d.B * d.N1 >= (1LL << 31)) {
        set_error("coarse problem too large for the resolve kernel's 32-bit indices (B * N, pools * hot_cap must be < 2^31)");
        return cudaErrorInvalidValue;
    }
    const int resident = std::max(1, sms * per_sm);
    const int grid = std::min(resident, kMaxResolveBlocks);
    ResolveWs rw{w.hot_meta, w.hot_dots, w.hot_count, w.l2row, w.l2col, w.rowbest, w.colbest,
                 w.rowcnt, w.rowslot, w.tie_count, w.tie_list, w.blocksum};
    CoarseResolveParams pp = p;
    void* args[] = {&pp, &rw};
    e = cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(coarse_resolve_kernel), dim3(grid), dim3(kResolveThreads),
                                    args, 0, st);
    count_launch();
    return e != cudaSuccess ? e : cudaGetLastError();
}

}  // namespace fm
