// Coarse LocalFeatureTransformer: the producer of the coarse matcher's
// inputs -- network/net.py:74 -> network/module/transformer.py:59-97, EncoderLayer
// transformer.py:34-57, LinearAttention attentions.py:22-46 -- on [B, L, 256] token sequences,
// d_model 256, 8 heads (head dimension 32), attention 'linear', layer_names any sequence of
// 'self' / 'cross' (shipped: ['self', 'cross'] * 4).
//
// The layer is five token-parallel GEMMs around one cross-token reduction (the per-image, per-head
// 32 x 32 matrices KV = K'^T V of the linear attention).  All GEMMs (six with the block-diagonal message product)
// run in ONE persistent tcgen05 kernel template (ct_gemm_kernel) that differs only in its epilogue:
//
//   half-pass (x rows of one image side, source rows of the same or the other side)
//     1. KV   : source . [Wk; Wv]^T       -> K' = elu + 1, V / S (fp32)          epilogue kEpiKV
//     2.        K'^T V and sum_s K' per image and head (SIMT, deterministic partials) + finalize into
//               a block-diagonal 256 x 256 matrix = the B operand of step 4
//     3. Q    : x . Wq^T                  -> Q' = elu + 1, Z = 1 / (Q' . sum K' + eps)        kEpiQ
//     4. MSG  : Q' . blockdiag(KV)        -> message = acc * Z * S                            kEpiMsg
//     5. MERGE: message . Wm^T            -> LayerNorm 1                                      kEpiLN
//     6. MLP1 : cat[x, message] . W1^T    -> relu                                             kEpiRelu
//     7. MLP2 : hidden . W2^T             -> LayerNorm 2, x += .                              kEpiLNRes
//
// fp32-grade products on the bf16 tensor pipe: every operand is kept as two bf16 planes (hi = rn(a),
// lo = rn(a - hi)) and a product is three MMAs, ah*bh + ah*bl + al*bh (error 2^-17 relative, the same
// split the fine transformer uses).  Every epilogue writes its result directly as the hi / lo planes
// the next GEMM's TMA loads read; fp32 copies exist only where SIMT code needs them (K', V, the
// residual stream x).  The hi plane of the final x IS the bf16 feature tensor the coarse matcher
// (K1) consumes -- there is no separate fp32 -> bf16 cast pass.
//
// GEMM kernel: persistent CTA PAIRS (__cluster_dims__(2, 1, 1), tcgen05.mma.cta_group::2).  A pair owns a 256-row
// x 256-column tile of one image; K in blocks of 64.  Each CTA TMA-loads its own 128 A rows and HALF of the W tile
// (its 128 of the 256 output columns' rows): a stage = {A hi, A lo, W/2 hi, W/2 lo} = 4 x 16 KB = 64 KB per CTA,
// three stages; both CTAs' loads complete on the leader's barrier; the leader's MMA thread issues 12 MMAs (3 terms
// x 4 k-steps, M 256 N 256 K 16) per stage and frees the stage in both CTAs with a multicast commit.  Two 256-column
// TMEM accumulators per CTA so tile i's epilogue overlaps tile i+1's MMAs.  10 warps per CTA: TMA producer, MMA
// issuer (idle in the peer), 8 epilogue warps (2 per TMEM lane quarter, each owning 128 of the 256 columns; a
// thread owns half a row of its CTA's 128 rows).  A tile never straddles two images (4-D tensor maps with the image
// as its own dimension; rows past L -- including a pair's whole second half when an image has an odd number of
// 128-row tiles -- are zero-filled by TMA and never stored).  tools/probe_2cta.cu pins the operand split.
#include <cuda.h>

#include <algorithm>
#include <vector>

#include "fm_common.cuh"
#include "fm_ptx.cuh"
#include "fm_tf_math.cuh"
#include "fm_tmap.cuh"

namespace fm {

namespace {

constexpr int kD = 256;                // d_model
constexpr int kH = 8;                  // heads
constexpr int kHd = 32;                // head dimension
constexpr int kTM = 128;               // rows per tile
constexpr int kTN = 256;               // columns per tile
constexpr int kBK = 64;                // K per stage (one 128-byte swizzle row of bf16)
constexpr int kStages = 3;
constexpr int kAPlaneBytes = kTM * kBK * 2;          // 16 KB: this CTA's 128 rows of the pair's 256-row A tile
constexpr int kWPlaneBytes = (kTN / 2) * kBK * 2;    // 16 KB: this CTA's 128 of the W tile's 256 rows (its N half)
constexpr int kStageBytes = 2 * kAPlaneBytes + 2 * kWPlaneBytes;   // 64 KB per CTA
constexpr int kGemmThreads = 320;      // producer warp, MMA warp, 8 epilogue warps
constexpr int kEpiThreads = 256;
constexpr int kKvChunkMin = 64;        // fewest tokens per K'^T V partial (bounds the partial buffers)

// ---- per-layer weight blob (bf16 planes [hi, lo], plain row-major [N][K]; TMA applies the swizzle)
constexpr size_t kWqOff = 0;                                   // [2][256][256]
constexpr size_t kWkvOff = kWqOff + 2ull * 256 * 256 * 2;      // [2][512][256]   rows 0..255 k_proj, 256..511 v_proj
constexpr size_t kWmOff = kWkvOff + 2ull * 512 * 256 * 2;      // [2][256][256]
constexpr size_t kW1Off = kWmOff + 2ull * 256 * 256 * 2;       // [2][512][512]
constexpr size_t kW2Off = kW1Off + 2ull * 512 * 512 * 2;       // [2][256][512]
constexpr size_t kLnOff = kW2Off + 2ull * 256 * 512 * 2;       // norm1.w, norm1.b, norm2.w, norm2.b: 4 x 256 fp32
constexpr size_t kLayerBlobBytes = kLnOff + 4ull * 256 * 4;

struct CtLayerParams {
    const float* q; const float* k; const float* v; const float* merge;
    const float* mlp0; const float* mlp2;
    const float* n1w; const float* n1b; const float* n2w; const float* n2b;
};

__device__ __forceinline__ void put_planes(__nv_bfloat16* hi, size_t plane, size_t idx, float w) {
    const __nv_bfloat16 h = __float2bfloat16_rn(w);
    hi[idx] = h;
    hi[plane + idx] = __float2bfloat16_rn(w - __bfloat162float(h));
}

// one layer per launch, grid-stride over its 655 360 weights
__global__ void __launch_bounds__(256) ct_pack_kernel(CtLayerParams p, uint8_t* __restrict__ blob) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    __nv_bfloat16* wq = reinterpret_cast<__nv_bfloat16*>(blob + kWqOff);
    __nv_bfloat16* wkv = reinterpret_cast<__nv_bfloat16*>(blob + kWkvOff);
    __nv_bfloat16* wm = reinterpret_cast<__nv_bfloat16*>(blob + kWmOff);
    __nv_bfloat16* w1 = reinterpret_cast<__nv_bfloat16*>(blob + kW1Off);
    __nv_bfloat16* w2 = reinterpret_cast<__nv_bfloat16*>(blob + kW2Off);
    for (int e = tid; e < 256 * 256; e += nthr) {
        put_planes(wq, 256 * 256, e, p.q[e]);
        put_planes(wm, 256 * 256, e, p.merge[e]);
        put_planes(wkv, 512 * 256, e, p.k[e]);
        put_planes(wkv, 512 * 256, 256 * 256 + e, p.v[e]);
    }
    for (int e = tid; e < 512 * 512; e += nthr) put_planes(w1, 512 * 512, e, p.mlp0[e]);
    for (int e = tid; e < 256 * 512; e += nthr) put_planes(w2, 256 * 512, e, p.mlp2[e]);
    float* ln = reinterpret_cast<float*>(blob + kLnOff);
    if (tid < 256) {
        ln[tid] = p.n1w[tid];
        ln[256 + tid] = p.n1b[tid];
        ln[512 + tid] = p.n2w[tid];
        ln[768 + tid] = p.n2b[tid];
    }
}

// fp32 tokens -> residual stream copy + operand planes
__global__ void __launch_bounds__(256) ct_split_kernel(const float4* __restrict__ in, float4* __restrict__ x,
                                                       uint2* __restrict__ hi, uint2* __restrict__ lo, size_t n4) {
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const float4 v = in[i];
        x[i] = v;
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
        const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
        const __nv_bfloat162 l0 = __floats2bfloat162_rn(v.x - f0.x, v.y - f0.y);
        const __nv_bfloat162 l1 = __floats2bfloat162_rn(v.z - f1.x, v.w - f1.y);
        hi[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
        lo[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1));
    }
}

// ---------------------------------------------------------------- K'^T V (attentions.py:42-43)
// grid (chunks, B); one warp per head; a lane owns a 4 (d) x 8 (v) patch of the head's 32 x 32
// matrix.  Tokens stream through a 3-stage cp.async ring in shared memory, 16 tokens (32 KB of K' | V) per
// stage, so two stages are always in flight per CTA (an HBM-bound streaming read); the products are
// packed f32x2 FMAs.  Partial sums per (image, chunk) are written, not accumulated atomically, so the result
// does not depend on scheduling.
constexpr int kKvStages = 3;
constexpr int kKvStageTokens = 16;
constexpr int kKvSmem = kKvStages * 2 * kKvStageTokens * kD * 4;   // 96 KB

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory"); }

__global__ void __launch_bounds__(256, 2) ct_kv_partial_kernel(const float* __restrict__ kf, const float* __restrict__ vf,
                                                               int L, int chunk_tokens, float* __restrict__ kvpart,
                                                               float* __restrict__ kspart) {
    extern __shared__ __align__(16) uint8_t kv_smem[];
    float* ring = reinterpret_cast<float*>(kv_smem);   // [stage][K' | V][16 tokens][256]
    const int chunk = blockIdx.x, b = blockIdx.y, nchunks = gridDim.x;
    const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int dg = lane >> 2, vg = lane & 3;
    const int t0 = min(chunk * chunk_tokens, L), t1 = min(t0 + chunk_tokens, L);
    const int nblk = (t1 - t0 + kKvStageTokens - 1) / kKvStageTokens;
    uint64_t acc[4][4];   // [d][v pair]
    uint64_t ks[2] = {0ull, 0ull};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;
    auto issue = [&](int blk) {   // one commit group per block, empty past the end (keeps the group count uniform)
        if (blk < nblk) {
            const int t = t0 + blk * kKvStageTokens;
            const int n4 = min(kKvStageTokens, t1 - t) * (kD / 4);
            float* sk = ring + static_cast<size_t>(blk % kKvStages) * 2 * kKvStageTokens * kD;
            float* sv = sk + kKvStageTokens * kD;
            const float4* gk = reinterpret_cast<const float4*>(kf + (static_cast<size_t>(b) * L + t) * kD);
            const float4* gv = reinterpret_cast<const float4*>(vf + (static_cast<size_t>(b) * L + t) * kD);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = threadIdx.x + i * 256;
                if (e < n4) { cp_async16(sk + e * 4, gk + e); cp_async16(sv + e * 4, gv + e); }
            }
        }
        cp_async_commit();
    };
    issue(0);
    issue(1);
    for (int blk = 0; blk < nblk; ++blk) {
        cp_async_wait<1>();   // block blk has landed (this thread's part); the barrier publishes everyone's
        __syncthreads();      // ... and everyone is done with the stage block blk + 2 will overwrite
        issue(blk + 2);
        const int nt = min(kKvStageTokens, t1 - (t0 + blk * kKvStageTokens));
        const float* sk = ring + static_cast<size_t>(blk % kKvStages) * 2 * kKvStageTokens * kD;
        const float* sv = sk + kKvStageTokens * kD;
        for (int r = 0; r < nt; ++r) {
            const float4 k4 = *reinterpret_cast<const float4*>(sk + r * kD + h * kHd + dg * 4);
            const ulonglong2 va = *reinterpret_cast<const ulonglong2*>(sv + r * kD + h * kHd + vg * 8);
            const ulonglong2 vb = *reinterpret_cast<const ulonglong2*>(sv + r * kD + h * kHd + vg * 8 + 4);
            ks[0] = add2(ks[0], pack2(k4.x, k4.y));
            ks[1] = add2(ks[1], pack2(k4.z, k4.w));
            const float kk[4] = {k4.x, k4.y, k4.z, k4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint64_t k2 = pack2(kk[i], kk[i]);
                acc[i][0] = fma2(k2, va.x, acc[i][0]);
                acc[i][1] = fma2(k2, va.y, acc[i][1]);
                acc[i][2] = fma2(k2, vb.x, acc[i][2]);
                acc[i][3] = fma2(k2, vb.y, acc[i][3]);
            }
        }
    }
    cp_async_wait<0>();
    float* out = kvpart + ((static_cast<size_t>(b) * nchunks + chunk) * kH + h) * (kHd * kHd);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        ulonglong2* row = reinterpret_cast<ulonglong2*>(out + (dg * 4 + i) * kHd + vg * 8);
        row[0] = make_ulonglong2(acc[i][0], acc[i][1]);
        row[1] = make_ulonglong2(acc[i][2], acc[i][3]);
    }
    if (vg == 0)
        *reinterpret_cast<ulonglong2*>(kspart + (static_cast<size_t>(b) * nchunks + chunk) * kD + h * kHd + dg * 4) =
            make_ulonglong2(ks[0], ks[1]);
}

// grid (heads, B): chunk partials summed in chunk order; KV[h][d][v] -> planes of the block-diagonal B
// operand kvbd[b][plane][n = h*32 + v][k = h*32 + d] (off-diagonal blocks are zeroed once per call),
// sum K' -> ksum.  A thread owns four consecutive v of one d (coalesced float4 partial reads).
__global__ void __launch_bounds__(256) ct_kv_finalize_kernel(const float* __restrict__ kvpart,
                                                             const float* __restrict__ kspart, int nchunks,
                                                             __nv_bfloat16* __restrict__ kvbd, float* __restrict__ ksum) {
    const int h = blockIdx.x, b = blockIdx.y;
    const int d = threadIdx.x >> 3, v0 = (threadIdx.x & 7) * 4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = 0; c < nchunks; ++c) {
        const float4 t = *reinterpret_cast<const float4*>(
            kvpart + ((static_cast<size_t>(b) * nchunks + c) * kH + h) * (kHd * kHd) + threadIdx.x * 4);
        s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
    }
    __nv_bfloat16* dst = kvbd + static_cast<size_t>(b) * 2 * kD * kD;
    const float sv[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
        put_planes(dst, kD * kD, static_cast<size_t>(h * kHd + v0 + i) * kD + h * kHd + d, sv[i]);
    if (threadIdx.x < kHd) {
        float t = 0.f;
        for (int c = 0; c < nchunks; ++c) t += kspart[(static_cast<size_t>(b) * nchunks + c) * kD + h * kHd + threadIdx.x];
        ksum[static_cast<size_t>(b) * kD + h * kHd + threadIdx.x] = t;
    }
}

// ---------------------------------------------------------------- the GEMM kernel
enum { kEpiKV = 0, kEpiQ = 1, kEpiMsg = 2, kEpiLN = 3, kEpiRelu = 4, kEpiLNRes = 5 };

struct CtGemmParams {
    int L, B;                 // rows per image, images
    int KB;                   // K / 64
    int kb_split;             // K blocks [0, kb_split) come from tensor map A0, the rest from A1 (cat[x, message])
    int n_tiles;              // N / 256
    int tiles_per_img;        // pairs of 128-row tiles per image
    int num_units;            // B * tiles_per_img * n_tiles (one per CTA pair), the column tile innermost
    int w_per_image;          // 1: the W operand is per image (blockdiag KV); 0: layer weights
    float* f32_a;             // kEpiKV: K' [B*L][256]
    float* f32_b;             // kEpiKV: V / S
    __nv_bfloat16* planes;    // output hi plane; lo plane = planes + plane_stride
    size_t plane_stride;
    int ld;                   // output row pitch in elements
    const float* ksum;        // kEpiQ: sum_s K' [B][256]
    float* z;                 // kEpiQ out, kEpiMsg in: [B*L][8]
    const float* ln;          // kEpiLN / kEpiLNRes: weight[256], bias[256]
    float* x;                 // kEpiLNRes: residual stream [B*L][256], updated in place
    float inv_len, len;       // 1 / S, S
};

struct __align__(16) CtCtl {
    float2 lnx[2][kTM][2];    // LayerNorm statistics of a row's two column halves, double-buffered by tile
    float lnp[2 * kD];
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t t_full[2];
    uint64_t t_empty[2];
    uint32_t tmem_base;
};

constexpr size_t kGemmSmem = static_cast<size_t>(kStages) * kStageBytes + sizeof(CtCtl) + 1024;
static_assert(kGemmSmem <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

__device__ __forceinline__ void load_chunk(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    tmem_ld_32x32(taddr, r);
    tmem_ld_wait_dep(r);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// 32 consecutive values of one row -> the hi and lo planes
__device__ __forceinline__ void store_planes32(__nv_bfloat16* dst, size_t plane_stride, const float (&v)[32]) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
    stg256_b32(dst, hi);
    stg256_b32(dst + plane_stride, lo);
}

template <int kEpi>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
ct_gemm_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
               const __grid_constant__ CUtensorMap mapW, const CtGemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    CtCtl* ctl = reinterpret_cast<CtCtl*>(smem + static_cast<size_t>(kStages) * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr bool kNorm = kEpi == kEpiLN || kEpi == kEpiLNRes;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        // t_empty lives in the leader CTA: the epilogue warps of BOTH CTAs arrive on it
        for (int s = 0; s < 2; ++s) { mbar_init(&ctl->t_full[s], 1); mbar_init(&ctl->t_empty[s], 2 * kEpiThreads / 32); }
        fence_mbar_init();
        tma_prefetch_desc(&mapA0);
        tma_prefetch_desc(&mapA1);
        tma_prefetch_desc(&mapW);
    }
    if (kNorm)
        for (int i = threadIdx.x; i < 2 * kD; i += kGemmThreads) ctl->lnp[i] = p.ln[i];
    if (warp == 0) tmem_alloc_pair<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the peer's barriers are initialised before anything signals them
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;
    // A CTA PAIR works on one unit: 256 rows (two row tiles of ONE image; rank r takes tile 2 i + r, which
    // may lie past the image's end: all its rows are then zero-filled by TMA and never stored) x 256 columns.
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (warp == 0) {
        // ===================================================== TMA producer (one lane, both CTAs)
        if (lane == 0) {
            uint32_t it = 0;
            for (int u = pair; u < p.num_units; u += npairs) {
                const int nt = u % p.n_tiles, t = u / p.n_tiles;
                const int b = t / p.tiles_per_img, r0 = ((t % p.tiles_per_img) * 2 + rank) * kTM;
                const int g = p.w_per_image ? b : 0;
                for (int kb = 0; kb < p.KB; ++kb, ++it) {
                    const uint32_t s = it % kStages, ph = (it / kStages) & 1;
                    mbar_wait(&ctl->empty[s], ph ^ 1);
                    // both CTAs' loads complete on the LEADER's barrier, which expects the pair's bytes
                    if (rank == 0) mbar_arrive_expect_tx(&ctl->full[s], 2 * kStageBytes);
                    uint8_t* st = smem + static_cast<size_t>(s) * kStageBytes;
                    const bool first = kb < p.kb_split;
                    const CUtensorMap* ma = first ? &mapA0 : &mapA1;
                    const int kk = (first ? kb : kb - p.kb_split) * kBK;
                    const int n0 = nt * kTN + static_cast<int>(rank) * (kTN / 2);
                    tma_load_4d_pair(st, ma, &ctl->full[s], kk, r0, b, 0);
                    tma_load_4d_pair(st + kAPlaneBytes, ma, &ctl->full[s], kk, r0, b, 1);
                    tma_load_4d_pair(st + 2 * kAPlaneBytes, &mapW, &ctl->full[s], kb * kBK, n0, 0, g);
                    tma_load_4d_pair(st + 2 * kAPlaneBytes + kWPlaneBytes, &mapW, &ctl->full[s], kb * kBK, n0, 1, g);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one lane)
        if (lane == 0 && rank == 0) {   // the leader CTA issues the pair's MMAs
            constexpr uint32_t idesc = umma_idesc_bf16_f32(2 * kTM, kTN);
            const uint32_t smem_base = smem_u32(smem);
            uint32_t it = 0, tile_it = 0;
            for (int u = pair; u < p.num_units; u += npairs, ++tile_it) {
                const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
                mbar_wait(&ctl->t_empty[buf], tph ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * kTN;
                for (int kb = 0; kb < p.KB; ++kb, ++it) {
                    const uint32_t s = it % kStages, ph = (it / kStages) & 1;
                    mbar_wait(&ctl->full[s], ph);
                    tc_fence_after();
                    const uint32_t a_hi = smem_base + s * kStageBytes, a_lo = a_hi + kAPlaneBytes;
                    const uint32_t w_hi = a_hi + 2 * kAPlaneBytes, w_lo = w_hi + kWPlaneBytes;
#pragma unroll 1
                    for (int term = 0; term < 3; ++term) {   // ah*wh + ah*wl + al*wh
                        const uint32_t a = term == 2 ? a_lo : a_hi, w = term == 1 ? w_lo : w_hi;
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k)
                            umma_bf16_ss_pair(d_tmem, umma_desc_k_sw128(a + k * 32), umma_desc_k_sw128(w + k * 32), idesc,
                                              (kb | term | k) != 0 ? 1u : 0u);
                    }
                    umma_commit_pair(&ctl->empty[s]);   // frees the stage in both CTAs once these MMAs retire
                }
                umma_commit_pair(&ctl->t_full[buf]);    // accumulator complete (both CTAs' epilogues)
            }
        }
    } else {
        // ===================================================== epilogue: 8 warps
        const int q = warp & 3;               // TMEM lane quarter this warp may access
        const int g = (warp - 2) >> 2;        // column half
        const int rl = q * 32 + lane;         // row within the tile
        uint32_t tile_it = 0;
        for (int u = pair; u < p.num_units; u += npairs, ++tile_it) {
            const int nt = u % p.n_tiles, t = u / p.n_tiles;
            const int b = t / p.tiles_per_img, row = ((t % p.tiles_per_img) * 2 + static_cast<int>(rank)) * kTM + rl;
            const bool valid = row < p.L;
            const size_t gr = static_cast<size_t>(b) * p.L + row;
            const uint32_t buf = tile_it & 1, tph = (tile_it >> 1) & 1;
            if (kEpi == kEpiLNRes && valid) {   // the residual row: in L2 by the time the second pass wants it
                const float* xr = p.x + gr * kD + g * 128;
#pragma unroll
                for (int i = 0; i < 4; ++i) asm volatile("prefetch.global.L2 [%0];" ::"l"(xr + i * 32));
            }
            mbar_wait(&ctl->t_full[buf], tph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kTN + g * 128;
            uint32_t ra[32], rb[32];   // two chunks of the accumulator row: chunk c+1 loads while chunk c is processed

            // ---- one 32-column chunk, everything but LayerNorm
            auto chunk = [&](int c, const uint32_t (&r)[32]) {
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                if (kEpi == kEpiKV) {
                    if (nt == 0) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = elu1(v[i]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] *= p.inv_len;
                    }
                    if (valid) {
                        float* dst = (nt == 0 ? p.f32_a : p.f32_b) + gr * kD + g * 128 + c * 32;
#pragma unroll
                        for (int i = 0; i < 32; i += 8) stg256(dst, v, i);
                    }
                } else if (kEpi == kEpiQ) {
                    const int h = g * 4 + c;   // a 32-column chunk is one head
                    const float4* ks = reinterpret_cast<const float4*>(p.ksum + static_cast<size_t>(b) * kD + h * kHd);
                    float den = 0.f;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 k4 = __ldg(ks + i);
                        v[4 * i] = elu1(v[4 * i]); v[4 * i + 1] = elu1(v[4 * i + 1]);
                        v[4 * i + 2] = elu1(v[4 * i + 2]); v[4 * i + 3] = elu1(v[4 * i + 3]);
                        den = fmaf(v[4 * i], k4.x, den); den = fmaf(v[4 * i + 1], k4.y, den);
                        den = fmaf(v[4 * i + 2], k4.z, den); den = fmaf(v[4 * i + 3], k4.w, den);
                    }
                    if (valid) {
                        p.z[gr * kH + h] = 1.0f / (den + 1e-6f);
                        store_planes32(p.planes + gr * p.ld + h * kHd, p.plane_stride, v);
                    }
                } else if (kEpi == kEpiMsg) {
                    const int h = g * 4 + c;
                    const float zz = valid ? p.z[gr * kH + h] : 0.f;
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = v[i] * zz * p.len;
                    if (valid) store_planes32(p.planes + gr * p.ld + h * kHd, p.plane_stride, v);
                } else if (kEpi == kEpiRelu) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
                    if (valid) store_planes32(p.planes + gr * p.ld + nt * kTN + g * 128 + c * 32, p.plane_stride, v);
                }
            };
            // ---- LayerNorm, pass 1: Chan's combination of the chunk statistics
            float mean = 0.f, m2 = 0.f, mu = 0.f, rstd = 0.f;
            auto stats = [&](int c, const uint32_t (&r)[32]) {
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                const float2 st = local_stats(v);
                const float na = 32.f * c, n = na + 32.f, dlt = st.x - mean;
                mean += dlt * (32.f / n);
                m2 += st.y + dlt * dlt * (na * 32.f / n);
            };
            // ---- LayerNorm, pass 2: normalise, (residual,) store
            auto norm = [&](int c, const uint32_t (&r)[32]) {
                const int col = g * 128 + c * 32;
                float v[32];
                const uint64_t r2 = pack2(rstd, rstd), nm = pack2(-mu, -mu);
#pragma unroll
                for (int i = 0; i < 32; i += 4) {   // y = v * (rstd w) + (b - mu rstd w), two channels per instruction
                    const float4 ww = *reinterpret_cast<const float4*>(&ctl->lnp[col + i]);
                    const float4 bb = *reinterpret_cast<const float4*>(&ctl->lnp[kD + col + i]);
                    const uint64_t a0 = mul2(pack2(ww.x, ww.y), r2), a1 = mul2(pack2(ww.z, ww.w), r2);
                    unpack2(fma2(pack2(__uint_as_float(r[i]), __uint_as_float(r[i + 1])), a0, fma2(a0, nm, pack2(bb.x, bb.y))),
                            v[i], v[i + 1]);
                    unpack2(fma2(pack2(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3])), a1, fma2(a1, nm, pack2(bb.z, bb.w))),
                            v[i + 2], v[i + 3]);
                }
                if (valid) {
                    if (kEpi == kEpiLNRes) {
                        float xv[32];   // prefetched into L2 at the top of the tile
#pragma unroll
                        for (int i = 0; i < 32; i += 8) ldg256(p.x + gr * kD + col, xv, i);
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] += xv[i];
#pragma unroll
                        for (int i = 0; i < 32; i += 8) stg256(p.x + gr * kD + col, v, i);
                    }
                    store_planes32(p.planes + gr * p.ld + col, p.plane_stride, v);
                }
            };
            // chunk c+1 is in flight while chunk c is processed
#define CT_PIPE4(F)                                                        \
    tmem_ld_32x32(taddr, ra);                                              \
    tmem_ld_wait_dep(ra); tmem_ld_32x32(taddr + 32, rb); F(0, ra);         \
    tmem_ld_wait_dep(rb); tmem_ld_32x32(taddr + 64, ra); F(1, rb);         \
    tmem_ld_wait_dep(ra); tmem_ld_32x32(taddr + 96, rb); F(2, ra);         \
    tmem_ld_wait_dep(rb); F(3, rb);

#ifdef CT_EPI_NOP   // mainloop-only timing (tools/dbg_build.sh)
            if (p.L >= 0) {
            } else
#endif
            if (!kNorm) {
                CT_PIPE4(chunk)
            } else {
                CT_PIPE4(stats)
                const uint32_t lb = tile_it & 1;
                ctl->lnx[lb][rl][g] = make_float2(mean, m2);
                tmem_ld_32x32(taddr, ra);   // pass 2's first chunk loads across the barrier
                named_bar_sync(1, kEpiThreads);
                const float2 other = ctl->lnx[lb][rl][g ^ 1];
                const float dm = mean - other.x;
                mu = 0.5f * (mean + other.x);
                rstd = 1.0f / sqrtf((m2 + other.y + dm * dm * 64.f) * (1.f / 256.f) + 1e-5f);
                tmem_ld_wait_dep(ra); tmem_ld_32x32(taddr + 32, rb); norm(0, ra);
                tmem_ld_wait_dep(rb); tmem_ld_32x32(taddr + 64, ra); norm(1, rb);
                tmem_ld_wait_dep(ra); tmem_ld_32x32(taddr + 96, rb); norm(2, ra);
                tmem_ld_wait_dep(rb); norm(3, rb);
            }
#undef CT_PIPE4
            // release the accumulator buffer to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&ctl->t_empty[buf], 0);
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the peer may still be reading operands / signalling barriers of this CTA
    tc_fence_after();
    if (warp == 0) tmem_dealloc_pair<512>(tmem_base);
}

// ---------------------------------------------------------------- host: workspace plan and launches
inline size_t align_up(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

struct CtPlan {
    size_t blobs, xp[2], kf, vf, qp, z, mp, m2p, hp, kvpart, kspart, kvbd, ksum, total;
    int chunks_max;
};

CtPlan ct_plan(int B, int L0, int L1, int n_layers) {
    CtPlan pl{};
    const size_t Lm = static_cast<size_t>(std::max(L0, L1)), R = static_cast<size_t>(B) * Lm;
    pl.chunks_max = static_cast<int>((Lm + kKvChunkMin - 1) / kKvChunkMin);
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off = align_up(off + bytes); return o; };
    pl.blobs = take(static_cast<size_t>(n_layers) * kLayerBlobBytes);
    pl.xp[0] = take(2ull * B * L0 * kD * 2);
    pl.xp[1] = take(2ull * B * L1 * kD * 2);
    pl.kf = take(R * kD * 4);
    pl.vf = take(R * kD * 4);
    pl.qp = take(2 * R * kD * 2);
    pl.z = take(R * kH * 4);
    pl.mp = take(2 * R * kD * 2);
    pl.m2p = take(2 * R * kD * 2);
    pl.hp = take(2 * R * 2 * kD * 2);
    pl.kvpart = take(static_cast<size_t>(B) * pl.chunks_max * kH * kHd * kHd * 4);
    pl.kspart = take(static_cast<size_t>(B) * pl.chunks_max * kD * 4);
    pl.kvbd = take(static_cast<size_t>(B) * 2 * kD * kD * 2);
    pl.ksum = take(static_cast<size_t>(B) * kD * 4);
    pl.total = off;
    return pl;
}

// activation planes [2][B][L][K] bf16 as (K, L, B, plane); box 64 x 128 rows
bool make_act_map(CUtensorMap* m, const void* ptr, int K, int L, int B) {
    const uint64_t dims[4] = {static_cast<uint64_t>(K), static_cast<uint64_t>(L), static_cast<uint64_t>(B), 2};
    const uint64_t strides[3] = {static_cast<uint64_t>(K) * 2, static_cast<uint64_t>(L) * K * 2,
                                 static_cast<uint64_t>(B) * L * K * 2};
    const uint32_t box[4] = {kBK, kTM, 1, 1};
    return encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}
// weight planes [G][2][N][K] bf16 as (K, N, plane, G); box 64 x 128 rows
bool make_w_map(CUtensorMap* m, const void* ptr, int K, int N, int G) {
    const uint64_t dims[4] = {static_cast<uint64_t>(K), static_cast<uint64_t>(N), 2, static_cast<uint64_t>(G)};
    const uint64_t strides[3] = {static_cast<uint64_t>(K) * 2, static_cast<uint64_t>(N) * K * 2,
                                 2ull * N * K * 2};
    const uint32_t box[4] = {kBK, kTN / 2, 1, 1};   // a CTA of the pair loads its half of the 256-row tile
    return encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

template <int kEpi>
cudaError_t launch_gemm(const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& w, CtGemmParams p, int K,
                        int N, int num_sms, cudaStream_t st) {
    static unsigned long long attr_set = 0;
    if (needs_attr_on_this_device(&attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(ct_gemm_kernel<kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kGemmSmem));
        if (e != cudaSuccess) return e;
    }
    p.KB = K / kBK;
    p.n_tiles = N / kTN;
    p.tiles_per_img = ((p.L + kTM - 1) / kTM + 1) / 2;   // PAIRS of row tiles per image (the last may be half empty)
    p.num_units = p.B * p.tiles_per_img * p.n_tiles;
    if (p.num_units == 0) return cudaSuccess;
    const int grid = 2 * std::min(num_sms / 2, p.num_units);   // whole CTA pairs
    ct_gemm_kernel<kEpi><<<grid, kGemmThreads, kGemmSmem, st>>>(a0, a1, w, p); count_launch();
    return cudaGetLastError();
}

}  // namespace

bool coarse_transformer_supported(int C, int nhead) { return C == kD && nhead == kH; }

size_t coarse_transformer_workspace_bytes(int B, int L0, int L1, int n_layers) {
    return ct_plan(B, L0, L1, n_layers).total;
}

// layer_params: n_layers x 10 device pointers (q, k, v, merge, mlp.0, mlp.2, norm1.w, norm1.b, norm2.w, norm2.b);
// out0 / out1: the updated fp32 tokens; bf0 / bf1 (optional): the same tokens as bf16 (what K1 reads)
cudaError_t launch_coarse_transformer(const float* feat0, const float* feat1, int B, int L0, int L1, int n_layers,
                                      const int* layer_kinds, const void* const* layer_params, float* out0,
                                      float* out1, void* bf0, void* bf1, void* workspace, int num_sms,
                                      cudaStream_t st) {
    const CtPlan pl = ct_plan(B, L0, L1, n_layers);
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    const int Ls[2] = {L0, L1};
    float* xs[2] = {out0, out1};
    const float* ins[2] = {feat0, feat1};
    __nv_bfloat16* xp[2] = {reinterpret_cast<__nv_bfloat16*>(ws + pl.xp[0]), reinterpret_cast<__nv_bfloat16*>(ws + pl.xp[1])};
    float* kf = reinterpret_cast<float*>(ws + pl.kf);
    float* vf = reinterpret_cast<float*>(ws + pl.vf);
    __nv_bfloat16* qp = reinterpret_cast<__nv_bfloat16*>(ws + pl.qp);
    float* z = reinterpret_cast<float*>(ws + pl.z);
    __nv_bfloat16* mp = reinterpret_cast<__nv_bfloat16*>(ws + pl.mp);
    __nv_bfloat16* m2p = reinterpret_cast<__nv_bfloat16*>(ws + pl.m2p);
    __nv_bfloat16* hp = reinterpret_cast<__nv_bfloat16*>(ws + pl.hp);
    float* kvpart = reinterpret_cast<float*>(ws + pl.kvpart);
    float* kspart = reinterpret_cast<float*>(ws + pl.kspart);
    __nv_bfloat16* kvbd = reinterpret_cast<__nv_bfloat16*>(ws + pl.kvbd);
    float* ksum = reinterpret_cast<float*>(ws + pl.ksum);
    cudaError_t e;

    for (int l = 0; layer_params && l < n_layers; ++l) {   // NULL: packed by an earlier call into this workspace
        CtLayerParams lp;
        const void* const* src = layer_params + static_cast<size_t>(l) * 10;
        lp.q = static_cast<const float*>(src[0]); lp.k = static_cast<const float*>(src[1]);
        lp.v = static_cast<const float*>(src[2]); lp.merge = static_cast<const float*>(src[3]);
        lp.mlp0 = static_cast<const float*>(src[4]); lp.mlp2 = static_cast<const float*>(src[5]);
        lp.n1w = static_cast<const float*>(src[6]); lp.n1b = static_cast<const float*>(src[7]);
        lp.n2w = static_cast<const float*>(src[8]); lp.n2b = static_cast<const float*>(src[9]);
        ct_pack_kernel<<<296, 256, 0, st>>>(lp, ws + pl.blobs + static_cast<size_t>(l) * kLayerBlobBytes); count_launch();
    }
    if ((e = cudaMemsetAsync(kvbd, 0, static_cast<size_t>(B) * 2 * kD * kD * 2, st)) != cudaSuccess) return e;
    for (int s = 0; s < 2; ++s) {
        const size_t n4 = static_cast<size_t>(B) * Ls[s] * kD / 4;
        if (n4 == 0) continue;
        const int grid = static_cast<int>(std::min<size_t>((n4 + 255) / 256, static_cast<size_t>(num_sms) * 16));
        ct_split_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(ins[s]), reinterpret_cast<float4*>(xs[s]),
                                              reinterpret_cast<uint2*>(xp[s]),
                                              reinterpret_cast<uint2*>(xp[s] + static_cast<size_t>(B) * Ls[s] * kD), n4); count_launch();
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;

    // tensor maps of the activation buffers (the intermediate ones depend on the x side's length)
    CUtensorMap m_x[2], m_q[2], m_m[2], m_m2[2], m_h[2], m_kv;
    bool ok = make_w_map(&m_kv, kvbd, kD, kD, B);
    for (int s = 0; s < 2; ++s) {
        ok = ok && make_act_map(&m_x[s], xp[s], kD, Ls[s], B) && make_act_map(&m_q[s], qp, kD, Ls[s], B) &&
             make_act_map(&m_m[s], mp, kD, Ls[s], B) && make_act_map(&m_m2[s], m2p, kD, Ls[s], B) &&
             make_act_map(&m_h[s], hp, 2 * kD, Ls[s], B);
    }
    static unsigned long long kv_attr = 0;
    if (needs_attr_on_this_device(&kv_attr)) {
        if ((e = cudaFuncSetAttribute(ct_kv_partial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kKvSmem)) != cudaSuccess) return e;
    }
    if (!ok) {
        set_error("cuTensorMapEncodeTiled failed (coarse transformer)");
        return cudaErrorInvalidValue;
    }

    for (int l = 0; l < n_layers; ++l) {
        const uint8_t* blob = ws + pl.blobs + static_cast<size_t>(l) * kLayerBlobBytes;
        CUtensorMap w_q, w_kv, w_m, w_1, w_2;
        if (!(make_w_map(&w_q, blob + kWqOff, kD, kD, 1) && make_w_map(&w_kv, blob + kWkvOff, kD, 2 * kD, 1) &&
              make_w_map(&w_m, blob + kWmOff, kD, kD, 1) && make_w_map(&w_1, blob + kW1Off, 2 * kD, 2 * kD, 1) &&
              make_w_map(&w_2, blob + kW2Off, 2 * kD, kD, 1))) {
            set_error("cuTensorMapEncodeTiled failed (coarse transformer weights)");
            return cudaErrorInvalidValue;
        }
        const float* ln = reinterpret_cast<const float*>(blob + kLnOff);
        for (int half = 0; half < 2; ++half) {
            // 'self': (x, source) = (0, 0), (1, 1); 'cross': (0, 1) then (1, 0) -- the second half-pass
            // sees the first one's update (transformer.py:94-95)
            const int xs_ = half, ss = layer_kinds[l] == 0 ? half : 1 - half;
            const int Lx = Ls[xs_], S = Ls[ss];
            const size_t Rx = static_cast<size_t>(B) * Lx;
            CtGemmParams p{};
            p.B = B;
            p.inv_len = 1.0f / static_cast<float>(S);
            p.len = static_cast<float>(S);
            // 1. K', V of the source
            p.L = S; p.kb_split = 4; p.w_per_image = 0; p.f32_a = kf; p.f32_b = vf;
            if ((e = launch_gemm<kEpiKV>(m_x[ss], m_x[ss], w_kv, p, kD, 2 * kD, num_sms, st)) != cudaSuccess) return e;
            // 2. K'^T V
            // chunks of ~192 tokens, their number trimmed so that B x chunks fills whole waves of 2 CTAs per SM
            int chunks = (S + 191) / 192;
            {
                const int slots = 2 * num_sms, waves = std::max(1, (B * chunks + slots / 2) / slots);
                chunks = std::max(1, std::min((S + kKvChunkMin - 1) / kKvChunkMin, std::max(chunks / 2, waves * slots / B)));
            }
            const int chunk_tokens = (S + chunks - 1) / chunks;
            chunks = (S + chunk_tokens - 1) / chunk_tokens;
            ct_kv_partial_kernel<<<dim3(chunks, B), 256, kKvSmem, st>>>(kf, vf, S, chunk_tokens, kvpart, kspart); count_launch();
            ct_kv_finalize_kernel<<<dim3(kH, B), 256, 0, st>>>(kvpart, kspart, chunks, kvbd, ksum); count_launch();
            // 3. Q', Z
            p.L = Lx; p.planes = qp; p.plane_stride = Rx * kD; p.ld = kD; p.ksum = ksum; p.z = z;
            if ((e = launch_gemm<kEpiQ>(m_x[xs_], m_x[xs_], w_q, p, kD, kD, num_sms, st)) != cudaSuccess) return e;
            // 4. message
            p.planes = mp; p.w_per_image = 1;
            if ((e = launch_gemm<kEpiMsg>(m_q[xs_], m_q[xs_], m_kv, p, kD, kD, num_sms, st)) != cudaSuccess) return e;
            // 5. merge + LayerNorm 1
            p.planes = m2p; p.w_per_image = 0; p.ln = ln;
            if ((e = launch_gemm<kEpiLN>(m_m[xs_], m_m[xs_], w_m, p, kD, kD, num_sms, st)) != cudaSuccess) return e;
            // 6. MLP layer 1 on cat[x, message]
            p.planes = hp; p.plane_stride = Rx * 2 * kD; p.ld = 2 * kD;
            if ((e = launch_gemm<kEpiRelu>(m_x[xs_], m_m2[xs_], w_1, p, 2 * kD, 2 * kD, num_sms, st)) != cudaSuccess) return e;
            // 7. MLP layer 2 + LayerNorm 2 + residual -> x (fp32 and planes)
            p.planes = xp[xs_]; p.plane_stride = Rx * kD; p.ld = kD; p.ln = ln + 2 * kD; p.x = xs[xs_]; p.kb_split = 8;
            if ((e = launch_gemm<kEpiLNRes>(m_h[xs_], m_h[xs_], w_2, p, 2 * kD, kD, num_sms, st)) != cudaSuccess) return e;
        }
    }
    for (int s = 0; s < 2; ++s) {
        void* dst = s == 0 ? bf0 : bf1;
        if (dst && (e = cudaMemcpyAsync(dst, xp[s], static_cast<size_t>(B) * Ls[s] * kD * 2, cudaMemcpyDeviceToDevice, st)) != cudaSuccess)
            return e;
    }
    return cudaGetLastError();
}

}  // namespace fm
