// K2 -- finishes the coarse stage from the statistics K1 left in the workspace:
//   * soft-max denominators of every token (log2 domain)              coarse_matching_new.py:68
//   * confidence of every candidate, threshold, per-row / per-column maxima (mutual nearest
//     neighbour on the CONFIDENCE, equality semantics)                coarse_matching_new.py:99-106
//   * border removal                                                  coarse_matching_new.py:11-28,101
//   * match list in torch.where order (b, i, j) + coarse pixel coords coarse_matching_new.py:109-134
// All of it touches O(N) data; the N0 x N1 matrix does not exist.
#include "fm_common.cuh"
#include "fm_ptx.cuh"

namespace fm {

namespace {

constexpr int kScanBlock = 1024;

// ---------------------------------------------------------------- 1. log-sum-exp of every token
__global__ void __launch_bounds__(256) finish_lse_kernel(CoarseDims d, float a2, const float* __restrict__ norm0,
                                                         const float* __restrict__ maxnorm,
                                                         const float* __restrict__ rowpart,
                                                         const float* __restrict__ colpart,
                                                         float* __restrict__ l2row, float* __restrict__ l2col,
                                                         int* __restrict__ status, int* __restrict__ redo) {
    const int b = blockIdx.y;
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
        const float* rp = rowpart + static_cast<size_t>(b) * d.NB * 2 * d.N0pad + t;
        float s = 0.f;
        for (int k = 0; k < d.NB * 2; ++k) s += rp[static_cast<size_t>(k) * d.N0pad];
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
struct CandView {
    int b, i, j;
    float conf;
};

__device__ __forceinline__ CandView load_candidate(const CoarseDims& d, float a2, const uint2* __restrict__ meta,
                                                   const float* __restrict__ dots,
                                                   const float* __restrict__ l2row,
                                                   const float* __restrict__ l2col, size_t rec, int lane) {
    CandView c;
    const uint2 m = meta[rec];
    c.b = static_cast<int>(m.x / static_cast<uint32_t>(d.N0));
    c.i = static_cast<int>(m.x - static_cast<uint32_t>(c.b) * d.N0);
    c.j = static_cast<int>(m.y) * kHotChunk + lane;
    c.conf = 0.f;
    if (c.j < d.N1) {
        const float t = dots[rec * kHotChunk + lane] * a2;
        const float lr = l2row[static_cast<size_t>(c.b) * d.N0 + c.i];
        const float lc = l2col[static_cast<size_t>(c.b) * d.N1 + c.j];
        // conf = softmax_j * softmax_i = 2^(t - lr) * 2^(t - lc)
        c.conf = exp2f((t - lr) + (t - lc));
    }
    return c;
}

constexpr int kCandBlocksPerPool = 16;
constexpr int kCandWarps = 8;

__global__ void __launch_bounds__(kCandWarps * 32) cand_best_kernel(CoarseDims d, float a2, float thr,
                                                        const uint2* __restrict__ meta,
                                                        const float* __restrict__ dots,
                                                        const int* __restrict__ hot_count,
                                                        const float* __restrict__ l2row,
                                                        const float* __restrict__ l2col,
                                                        unsigned* __restrict__ rowbest,
                                                        unsigned* __restrict__ colbest, int* __restrict__ status) {
    const int pool = blockIdx.x / kCandBlocksPerPool, sub = blockIdx.x % kCandBlocksPerPool;
    const int n = min(hot_count[pool], d.hot_cap);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0 && sub == 0 && n) atomicAdd(status + kStCandTotal, n);
    for (int r = sub * kCandWarps + warp; r < n; r += kCandBlocksPerPool * kCandWarps) {
        const CandView c = load_candidate(d, a2, meta, dots, l2row, l2col, static_cast<size_t>(pool) * d.hot_cap + r, lane);
        if (c.conf > thr) {   // positive floats order like their bit patterns
            atomicMax(rowbest + static_cast<size_t>(c.b) * d.N0 + c.i, __float_as_uint(c.conf));
            atomicMax(colbest + static_cast<size_t>(c.b) * d.N1 + c.j, __float_as_uint(c.conf));
        }
    }
}

__device__ __forceinline__ bool interior(int idx, int h, int w, int border) {
    if (border <= 0) return true;
    const int r = idx / w, c = idx - r * w;
    return r >= border && r < h - border && c >= border && c < w - border;
}

__global__ void __launch_bounds__(kCandWarps * 32) cand_select_kernel(CoarseDims d, float a2, float thr, int h0c, int w0c, int h1c,
                                                          int w1c, int border, const uint2* __restrict__ meta,
                                                          const float* __restrict__ dots,
                                                          const int* __restrict__ hot_count,
                                                          const float* __restrict__ l2row,
                                                          const float* __restrict__ l2col,
                                                          const unsigned* __restrict__ rowbest,
                                                          const unsigned* __restrict__ colbest,
                                                          int* __restrict__ rowcnt, int2* __restrict__ rowslot,
                                                          int* __restrict__ tie_count, int4* __restrict__ tie_list,
                                                          int* __restrict__ status) {
    const int pool = blockIdx.x / kCandBlocksPerPool, sub = blockIdx.x % kCandBlocksPerPool;
    const int n = min(hot_count[pool], d.hot_cap);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = sub * kCandWarps + warp; r < n; r += kCandBlocksPerPool * kCandWarps) {
        const CandView c = load_candidate(d, a2, meta, dots, l2row, l2col, static_cast<size_t>(pool) * d.hot_cap + r, lane);
        if (!(c.conf > thr)) continue;
        const unsigned bits = __float_as_uint(c.conf);
        const size_t row = static_cast<size_t>(c.b) * d.N0 + c.i;
        if (bits != rowbest[row] || bits != colbest[static_cast<size_t>(c.b) * d.N1 + c.j]) continue;
        if (!interior(c.i, h0c, w0c, border) || !interior(c.j, h1c, w1c, border)) continue;
        const int k = atomicAdd(rowcnt + row, 1);
        if (k < kRowSlots) {
            rowslot[row * kRowSlots + k] = make_int2(c.j, static_cast<int>(bits));
        } else {
            // more exactly tied matches in this row than slots: the reference's equality test keeps them all
            // (coarse_matching_new.py:105-109), so do we, through one shared list
            const int t = atomicAdd(tie_count, 1);
            if (t < kTieCap) tie_list[t] = make_int4(static_cast<int>(row), c.j, static_cast<int>(bits), 0);
            else atomicOr(status + kStFlags, FM_FLAG_TIE_OVERFLOW);
        }
    }
}

// ---------------------------------------------------------------- 3. ordered compaction
__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
    __shared__ int warp_sums[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
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

__global__ void __launch_bounds__(kScanBlock) row_blocksum_kernel(const int* __restrict__ rowcnt, int nrows,
                                                                  const int* __restrict__ tie_count,
                                                                  int* __restrict__ blocksum) {
    __shared__ int total;
    const int r = blockIdx.x * kScanBlock + threadIdx.x;
    const int v = r < nrows ? row_emit_count(rowcnt[r], *tie_count) : 0;
    block_exclusive_scan(v, &total);
    if (threadIdx.x == 0) blocksum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanBlock) blocksum_scan_kernel(int* __restrict__ blocksum, int nblocks,
                                                                   int capacity, int* __restrict__ status) {
    __shared__ int total;
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += kScanBlock) {
        const int k = base + threadIdx.x;
        const int v = k < nblocks ? blocksum[k] : 0;
        const int ex = block_exclusive_scan(v, &total);
        const int carry = carry_s;
        if (k < nblocks) blocksum[k] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const int m = carry_s;
        status[kStCount] = min(m, capacity);
        if (m > capacity) atomicOr(status + kStFlags, FM_FLAG_OUT_OVERFLOW);
    }
}

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

__global__ void __launch_bounds__(kScanBlock) emit_kernel(CoarseResolveParams p, const int* __restrict__ rowcnt,
                                                          const int2* __restrict__ rowslot,
                                                          const int* __restrict__ tie_count,
                                                          const int4* __restrict__ tie_list,
                                                          const int* __restrict__ blockoff) {
    __shared__ int total;
    const int nrows = p.d.B * p.d.N0;
    const int r = blockIdx.x * kScanBlock + threadIdx.x;
    const int ties = *tie_count;
    const int n = r < nrows ? row_emit_count(rowcnt[r], ties) : 0;
    const int pos0 = blockoff[blockIdx.x] + block_exclusive_scan(n, &total);
    if (n == 0) return;
    if (n > kRowSlots) {
        // rare: this row's matches are its slots plus its entries of the shared list; emit in ascending j by
        // repeated selection (no j repeats within a row)
        const int b = r / p.d.N0, i = r - b * p.d.N0;
        float s0x = p.scale, s0y = p.scale, s1x = p.scale, s1y = p.scale;
        if (p.scale0) { s0x = p.scale * p.scale0[b * 2]; s0y = p.scale * p.scale0[b * 2 + 1]; }
        if (p.scale1) { s1x = p.scale * p.scale1[b * 2]; s1y = p.scale * p.scale1[b * 2 + 1]; }
        const int i_r = i / p.w0c, i_c = i - i_r * p.w0c;
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
    const int b = r / p.d.N0, i = r - b * p.d.N0;
    float s0x = p.scale, s0y = p.scale, s1x = p.scale, s1y = p.scale;
    if (p.scale0) { s0x = p.scale * p.scale0[b * 2]; s0y = p.scale * p.scale0[b * 2 + 1]; }
    if (p.scale1) { s1x = p.scale * p.scale1[b * 2]; s1y = p.scale * p.scale1[b * 2 + 1]; }
    const int i_r = i / p.w0c, i_c = i - i_r * p.w0c;
#pragma unroll
    for (int k = 0; k < kRowSlots; ++k) {
        if (k >= n) break;
        const int pos = pos0 + k;
        if (pos >= p.capacity) break;
        emit_match(p, pos, b, i, s[k].x, s[k].y, i_r, i_c, s0x, s0y, s1x, s1y);
    }
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
                                                                     with_gate ? w.redo : nullptr); count_launch();
    return cudaGetLastError();
}

cudaError_t launch_coarse_resolve(const CoarseResolveParams& p, const CoarseWorkspace& w, cudaStream_t st) {
    const CoarseDims& d = p.d;
    const unsigned cgrid = static_cast<unsigned>(d.num_pools) * kCandBlocksPerPool;
    cand_best_kernel<<<cgrid, kCandWarps * 32, 0, st>>>(d, p.a2, p.thr, w.hot_meta, w.hot_dots, w.hot_count, w.l2row,
                                                        w.l2col, w.rowbest, w.colbest, p.status); count_launch();
    cand_select_kernel<<<cgrid, kCandWarps * 32, 0, st>>>(d, p.a2, p.thr, p.h0c, p.w0c, p.h1c, p.w1c, p.border,
                                                          w.hot_meta, w.hot_dots, w.hot_count, w.l2row, w.l2col,
                                                          w.rowbest, w.colbest, w.rowcnt, w.rowslot, w.tie_count, w.tie_list,
                                                          p.status); count_launch();
    const int nrows = d.B * d.N0;
    const int nblocks = (nrows + kScanBlock - 1) / kScanBlock;
    row_blocksum_kernel<<<nblocks, kScanBlock, 0, st>>>(w.rowcnt, nrows, w.tie_count, w.blocksum); count_launch();
    blocksum_scan_kernel<<<1, kScanBlock, 0, st>>>(w.blocksum, nblocks, p.capacity, p.status); count_launch();
    emit_kernel<<<nblocks, kScanBlock, 0, st>>>(p, w.rowcnt, w.rowslot, w.tie_count, w.tie_list, w.blocksum); count_launch();
    return cudaGetLastError();
}

}  // namespace fm
