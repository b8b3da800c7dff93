// Fine LocalFeatureTransformer: the step between FinePreprocess and
// FineMatching -- network/net.py:79-80 -> network/module/transformer.py:78-97, EncoderLayer
// transformer.py:34-57, LinearAttention attentions.py:22-46 -- on [M, WW, 64] windows,
// d_model 64, 8 heads, attention 'linear', layer_names any sequence of 'self' / 'cross'.
//
// One launch per encoder layer (the layer's weights, bf16 hi + lo parts, stay resident in shared
// memory: 160 KB).  A tile is ONE match: TMEM lanes 0..WW-1 = its image-0 window, lanes
// 64..64+WW-1 = its image-1 window, one lane per window cell.  Two tiles are in flight per SM
// ("slots"); a slot has 8 epilogue warps (two warpgroups, each owning 32 of a row's 64 channels /
// 4 of its 8 heads) and its own MMA-issuing warp, so the two slots overlap each other's tensor
// and SIMT phases.
//
// Every GEMM is a tcgen05.mma with the ACTIVATIONS as the A operand IN TENSOR MEMORY (written by
// the epilogue threads with tcgen05.st as packed bf16 pairs -- tools/probe_ts_mma.cu pins the
// layout) and the weights as the K-major SW128 B operand in shared memory; fp32-grade products come
// from the split a*b ~= ah*bh + ah*bl + al*bh (a = ah + al, bf16 parts; error 2^-17 relative).
// Activations never touch shared memory; the only SIMT shared-memory traffic is the K^T V
// reduction of the linear attention (cross-row) and the LayerNorm statistics exchange.
//
// TMEM columns of a slot (256):
//   [  0,192) accumulators: Q|K|V (192) -> merge out (64) -> MLP hidden (128) -> MLP out (64)
//   [128,192) after K/V are consumed: attention message, then LayerNorm-1 output (A operand:
//             hi 32 columns + lo 32 columns), then the lo half of the hidden activations
//   [192,256) X operand (hi 32 + lo 32); overwritten by the hi half of the hidden activations
//             and rewritten with the updated x at the end of every pass
#include <algorithm>
#include <atomic>

#include "fm_common.cuh"
#include "fm_ptx.cuh"
#include "fm_tf_math.cuh"

namespace fm {

namespace {

constexpr int kC = 64;                 // d_model
constexpr int kHeads = 8;
constexpr int kMaxWW = 49;             // window cells per window (scratch sizing)
constexpr int kTfThreads = 640;        // 16 epilogue warps + 2 MMA warps + loader + spare
constexpr int kSlotThreads = 256;
constexpr int kSlotCols = 256;

// ---- per-layer weight blob: B operands (K-major, 128-byte swizzle, [term: hi, lo]) + LayerNorm
constexpr int kOffQkv = 0;             // [2][192 rows x 128 B]      rows: q 0..63, k 64..127, v 128..191
constexpr int kOffMerge = 49152;       // [2][64 x 128 B]
constexpr int kOffMlp0 = 65536;        // [2][2 K chunks][128 x 128 B]
constexpr int kOffMlp2 = 131072;       // [2][2 K chunks][64 x 128 B]
constexpr int kOffLn = 163840;         // norm1.weight, norm1.bias, norm2.weight, norm2.bias: 4 x 64 fp32
constexpr int kBlobBytes = 164864;
constexpr int kScrRow = 528;          // K | V fp32 of one cell (512 B) + 16 B pad: rows r and r + 8 share banks, so
                                      // a warp's 32 row-strided 16-byte stores take the minimum 4 wavefronts and
                                      // the reduction reads rows at compile-time offsets
constexpr int kScratchBytes = 2 * kMaxWW * kScrRow;

struct TfLayerParams {
    const float* q; const float* k; const float* v; const float* merge;
    const float* mlp0; const float* mlp2;
    const float* n1w; const float* n1b; const float* n2w; const float* n2b;
};

__device__ __forceinline__ void put_split(uint8_t* tile_hi, uint8_t* tile_lo, int n, int kk, float w) {
    const __nv_bfloat16 h = __float2bfloat16_rn(w);
    const __nv_bfloat16 l = __float2bfloat16_rn(w - __bfloat162float(h));
    const int off = n * 128 + ((((kk >> 3) ^ (n & 7))) << 4) + (kk & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(tile_hi + off) = h;
    *reinterpret_cast<__nv_bfloat16*>(tile_lo + off) = l;
}

// one layer per launch, grid-stride over its 40 960 weights
__global__ void __launch_bounds__(256) tf_pack_kernel(TfLayerParams p, uint8_t* __restrict__ blob) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    for (int e = tid; e < 192 * 64; e += nthr) {
        const int n = e >> 6, kk = e & 63;
        const float* src = n < 64 ? p.q : (n < 128 ? p.k : p.v);
        put_split(blob + kOffQkv, blob + kOffQkv + 24576, n, kk, src[(n & 63) * 64 + kk]);
    }
    for (int e = tid; e < 64 * 64; e += nthr) {
        const int n = e >> 6, kk = e & 63;
        put_split(blob + kOffMerge, blob + kOffMerge + 8192, n, kk, p.merge[e]);
    }
    for (int e = tid; e < 128 * 128; e += nthr) {
        const int n = e >> 7, k = e & 127, kc = k >> 6;
        put_split(blob + kOffMlp0 + kc * 16384, blob + kOffMlp0 + (2 + kc) * 16384, n, k & 63, p.mlp0[e]);
    }
    for (int e = tid; e < 64 * 128; e += nthr) {
        const int n = e >> 7, k = e & 127, kc = k >> 6;
        put_split(blob + kOffMlp2 + kc * 8192, blob + kOffMlp2 + (2 + kc) * 8192, n, k & 63, p.mlp2[e]);
    }
    float* ln = reinterpret_cast<float*>(blob + kOffLn);
    if (tid < 64) {
        ln[tid] = p.n1w[tid];
        ln[64 + tid] = p.n1b[tid];
        ln[128 + tid] = p.n2w[tid];
        ln[192 + tid] = p.n2b[tid];
    }
}

// ---------------------------------------------------------------- device helpers
// 32 values of this thread's row -> A operand columns (hi at `thi`, lo at `tlo`, 16 columns each)
__device__ __forceinline__ void store_operand32(uint32_t thi, uint32_t tlo, const float (&v)[32]) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
    tmem_st_32x16(thi, hi);
    tmem_st_32x16(tlo, lo);
}
// LayerNorm over 64 channels from the two halves' statistics (Chan's combination), eps 1e-5
// The LayerNorm parameters of the layer being run sit in the CONSTANT bank (copied there from the packed blob,
// stream-ordered, before the launch): read from shared memory they cost every warp 128 crossbar cycles per pass on a
// kernel whose shared-memory data pipe is 61 % busy (r02 ncu) -- 4.8 % of the kernel's time (A/B with the loads
// compiled out).  kLnSlots rotating slots: a launch on another stream must be kLnSlots launches behind before it
// could overwrite the parameters of a kernel still running.
constexpr int kLnSlots = 16;
__constant__ float c_tf_ln[kLnSlots][256];   // norm1.weight, norm1.bias, norm2.weight, norm2.bias (4 x 64)

// w, b: offsets of this thread's 32 weights / biases in c_tf_ln[slot]
__device__ __forceinline__ void layer_norm32(float (&v)[32], float2 mine, float2 other, int slot, int w, int b) {
    const float mean = 0.5f * (mine.x + other.x);
    const float dm = mine.x - other.x;
    const float var = (mine.y + other.y + dm * dm * 16.f) * (1.f / 64.f);
    const float rstd = 1.0f / sqrtf(var + 1e-5f);
    const uint64_t r2 = pack2(rstd, rstd), nm = pack2(-mean, -mean);
#pragma unroll
    for (int i = 0; i < 32; i += 4) {   // y = v * (rstd w) + (b - mean rstd w), two channels per instruction
        const float4 ww = make_float4(c_tf_ln[slot][w + i], c_tf_ln[slot][w + i + 1], c_tf_ln[slot][w + i + 2], c_tf_ln[slot][w + i + 3]);
        const float4 bb = make_float4(c_tf_ln[slot][b + i], c_tf_ln[slot][b + i + 1], c_tf_ln[slot][b + i + 2], c_tf_ln[slot][b + i + 3]);
        const uint64_t a0 = mul2(pack2(ww.x, ww.y), r2), a1 = mul2(pack2(ww.z, ww.w), r2);
        unpack2(fma2(pack2(v[i], v[i + 1]), a0, fma2(a0, nm, pack2(bb.x, bb.y))), v[i], v[i + 1]);
        unpack2(fma2(pack2(v[i + 2], v[i + 3]), a1, fma2(a1, nm, pack2(bb.z, bb.w))), v[i + 2], v[i + 3]);
    }
}

// -DTF_TRACE: clock64 stamps of CTA 0 (warp 0 of each slot + the slot's MMA thread) for tools/dbg_tf_trace.py
#ifdef TF_TRACE
constexpr int kTrTiles = 64, kTrPts = 32;
__device__ long long g_tf_trace[2][2][kTrTiles][kTrPts];   // [role: epilogue, mma][slot][tile][point]
#define TFT(role, slot, n, id) do { if (blockIdx.x == 0 && (n) < kTrTiles) g_tf_trace[role][slot][n][id] = clock64(); } while (0)
#else
#define TFT(role, slot, n, id) do {} while (0)
#endif

// Hand-off waits: mbarrier.try_wait with a suspend-time hint -- the hardware parks the warp until the phase completes
// or the hint expires, so a waiting warp neither burns issue slots nor oversleeps (try_wait + nanosleep(64) back-off
// measured 2 % slower: r02 A/B, tools/gpu_ab_tf.sh).  -DTF_WAIT_HINT=0 restores the back-off loop.
#ifndef TF_WAIT_HINT
#define TF_WAIT_HINT 1000
#endif
template <int kSleepNs>
__device__ __forceinline__ void tf_wait(uint64_t* bar, uint32_t parity) {
#if TF_WAIT_HINT > 0
    uint32_t spins = 0;
    while (!mbar_try_wait_hint(bar, parity, TF_WAIT_HINT)) {
        if (++spins > 20000000u) __trap();   // protocol bug -> launch failure, not a hung GPU
    }
#else
    mbar_wait_backoff<kSleepNs>(bar, parity);
#endif
}

struct __align__(16) TfCtl {
    uint64_t w_full;
    uint64_t to_epi[2];     // MMA warp -> the slot's epilogue warps (tcgen05.commit)
    uint64_t to_mma[2];     // epilogue warps -> the slot's MMA warp (8 arrivals)
    uint64_t q_done[2];     // cross: Q projection complete (own barrier, own phase counter)
    uint32_t tmem_base;
    int lock_owner;         // attention scratch: 0 free, else slot + 1
    int lock_left;          // warps of the owner still using it
};

struct TfParams {
    const float* in[2];
    float* out[2];
    const int* count;
    int capacity;
    int WW;
    const uint8_t* blob;
    int ln_slot;            // c_tf_ln slot holding this layer's LayerNorm parameters
};

// ---------------------------------------------------------------- one encoder layer
// kCross = false, 'self': a work item is ONE match; TMEM lanes 0.. = its image-0 window, lanes 64.. =
//   its image-1 window; one pass updates both (each window attends to itself).
// kCross = true, 'cross': a work item is a PAIR of matches (a, b) and a tile holds the SAME side of
//   both: lanes 0.. = match a, lanes 64.. = match b.  Pass 0 (transformer.py:94): the image-1 windows
//   go through the K|V projection only (their keys/values -> the per-window KV matrices), then the
//   image-0 windows take the tile, get their Q and run the rest of the layer against those matrices.
//   Pass 1 (transformer.py:95): the UPDATED image-0 windows are still in the tile's operand columns and
//   supply K|V; the image-1 windows are read again (L2) and updated.  Every lane of a pass is a row
//   that is being updated -- no half of the tile idles (the first version kept one match per tile and
//   spent two half-empty passes on it: 12.4 ms per 125 k matches).
// kWW = window cells when known at compile time (49: the shipped 7 x 7 window), 0 = read p.WW
template <bool kCross, int kWW>
__global__ void __launch_bounds__(kTfThreads, 1) tf_layer_kernel(const TfParams p) {
    extern __shared__ uint8_t smem_raw[];
    // aligned by OFFSET (not through an integer cast) so the compiler keeps the shared address space
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* blob = smem;
    uint8_t* scr = smem + kBlobBytes;                                         // K | V rows (both slots, locked)
    float* kvm_all = reinterpret_cast<float*>(scr + kScratchBytes);           // [2 slots][2 windows][64 d][8 v]
    float* ksum_all = kvm_all + 2 * 2 * 64 * 8;                               // [2 slots][2][64]
    float2* lnx = reinterpret_cast<float2*>(ksum_all + 2 * 2 * 64);           // [2 slots][128 rows][2 halves]
    TfCtl* ctl = reinterpret_cast<TfCtl*>(lnx + 2 * 128 * 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int kPasses = kCross ? 2 : 1;

    if (threadIdx.x == 0) {
        mbar_init(&ctl->w_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&ctl->to_epi[s], 1); mbar_init(&ctl->to_mma[s], 8); mbar_init(&ctl->q_done[s], 1); }
        ctl->lock_owner = 0;
        ctl->lock_left = 0;
        fence_mbar_init();
    }
    if (warp == 18) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;
    const int M = p.count ? min(*p.count, p.capacity) : p.capacity;
    const int WW = kWW ? kWW : p.WW;
    const int items = kCross ? (M + 1) / 2 : M;

    if (warp >= 16) {
        reg_dealloc<56>();
        if (warp == 18 && lane == 0) {   // -------------------------------- weights -> shared memory
            mbar_arrive_expect_tx(&ctl->w_full, kBlobBytes);
            for (int off = 0; off < kBlobBytes; off += 32768)
                bulk_load_1d(blob + off, p.blob + off, min(32768, kBlobBytes - off), &ctl->w_full);
        } else if (warp < 18 && lane == 0) {   // -------------------------- MMA issuer of slot `warp - 16`
            const int slot = warp - 16;
            const uint32_t ts = tmem_base + slot * kSlotCols;
            const uint32_t wb = smem_u32(blob);
            constexpr uint32_t id192 = umma_idesc_bf16_f32(128, 192), id128 = umma_idesc_bf16_f32(128, 128),
                               id64 = umma_idesc_bf16_f32(128, 64);
            const uint64_t d_qkv = umma_desc_k_sw128(wb + kOffQkv), d_merge = umma_desc_k_sw128(wb + kOffMerge),
                           d_mlp0 = umma_desc_k_sw128(wb + kOffMlp0), d_mlp2 = umma_desc_k_sw128(wb + kOffMlp2);
            uint32_t it = 0;
            auto wait_ops = [&]() { tf_wait<20>(&ctl->to_mma[slot], it & 1); ++it; tc_fence_after(); };
            // D[., n0 .. n0 + N) (+)= X operand (K = 64) x rows [n0, n0 + N) of Wqkv, three split terms.
            // term 0: a_hi b_hi, term 1: a_hi b_lo, term 2: a_lo b_hi.  The loops over terms are NOT
            // unrolled and descriptors are base + offset: fully unrolled, the 72 descriptors of a pass
            // become loop invariants that spill (the issuing thread then waits on local memory for
            // every MMA: 280 cycles per instruction instead of 50).
            auto project = [&](uint32_t n0, uint32_t idesc, uint64_t* done) {
#pragma unroll 1
                for (int term = 0; term < 3; ++term) {
                    const uint32_t a = ts + (term == 2 ? 224 : 192);
                    const uint64_t b = d_qkv + (term == 1 ? 24576 / 16 : 0) + n0 * (128 / 16);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16_ts(ts + n0, a + 8 * k, b + 2 * k, idesc, (term | k) != 0);
                }
                umma_commit(done);
            };
            mbar_wait(&ctl->w_full, 0);
            [[maybe_unused]] int tn = 0;
            for (int item = blockIdx.x * 2 + slot; item < items; item += 2 * gridDim.x, ++tn) {
                for (int pass = 0; pass < kPasses; ++pass) {
                    [[maybe_unused]] const int tb = 8 * pass;
                    if (kCross) {
                        // K | V of the source rows, then -- after the operand swap -- Q of the rows to
                        // update: the Q batch runs while the epilogue reduces K^T V
                        wait_ops(); project(64, id128, &ctl->to_epi[slot]);
                        wait_ops(); project(0, id64, &ctl->q_done[slot]);
                    } else {
                        wait_ops(); TFT(1, slot, tn, tb + 0); project(0, id192, &ctl->to_epi[slot]);   // Q | K | V
                        TFT(1, slot, tn, tb + 1);
                    }
                    wait_ops();   // message ready: merge
                    TFT(1, slot, tn, tb + 2);
#pragma unroll 1
                    for (int term = 0; term < 3; ++term) {
                        const uint32_t a = ts + (term == 2 ? 160 : 128);
                        const uint64_t b = d_merge + (term == 1 ? 8192 / 16 : 0);
#pragma unroll
                        for (int k = 0; k < 4; ++k) umma_bf16_ts(ts, a + 8 * k, b + 2 * k, id64, (term | k) != 0);
                    }
                    umma_commit(&ctl->to_epi[slot]);
                    TFT(1, slot, tn, tb + 3);
                    wait_ops();   // LayerNorm-1 output ready: hidden = [x | msg] Wmlp0^T
                    TFT(1, slot, tn, tb + 4);
#pragma unroll 1
                    for (int term = 0; term < 3; ++term)
#pragma unroll 1
                        for (int kc = 0; kc < 2; ++kc) {
                            const uint32_t a = ts + (kc == 0 ? (term == 2 ? 224 : 192) : (term == 2 ? 160 : 128));
                            const uint64_t b = d_mlp0 + ((term == 1 ? 2 : 0) + kc) * (16384 / 16);
#pragma unroll
                            for (int k = 0; k < 4; ++k) umma_bf16_ts(ts, a + 8 * k, b + 2 * k, id128, (term | kc | k) != 0);
                        }
                    umma_commit(&ctl->to_epi[slot]);
                    TFT(1, slot, tn, tb + 5);
                    wait_ops();   // relu(hidden) ready: out = hidden Wmlp2^T
                    TFT(1, slot, tn, tb + 6);
#pragma unroll 1
                    for (int term = 0; term < 3; ++term)
#pragma unroll 1
                        for (int kc = 0; kc < 2; ++kc) {
                            const uint32_t a = ts + (term == 2 ? 128 : 192) + 32 * kc;
                            const uint64_t b = d_mlp2 + ((term == 1 ? 2 : 0) + kc) * (8192 / 16);
#pragma unroll
                            for (int k = 0; k < 4; ++k) umma_bf16_ts(ts, a + 8 * k, b + 2 * k, id64, (term | kc | k) != 0);
                        }
                    umma_commit(&ctl->to_epi[slot]);
                    TFT(1, slot, tn, tb + 7);
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue warps
        reg_alloc<104>();   // pool = 640 x 96 (launch) - 128 x 56 (control warps) = 512 x 106
        const int slot = warp >> 3, g = (warp >> 2) & 1, q = warp & 3;
        const int st = threadIdx.x & (kSlotThreads - 1);          // thread within the slot
        const int w = q >> 1, r = (q & 1) * 32 + lane;            // window slot of the tile, window cell
        const int row = q * 32 + lane;                             // TMEM lane
        const uint32_t tl = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + slot * kSlotCols;
        const uint32_t bar_id = 1 + slot;
        const int ln_slot = p.ln_slot, ln_off = 32 * g;
        float2* lnrow = lnx + (slot * 128 + row) * 2;
        float* kvm = kvm_all + slot * 2 * 64 * 8;                  // this slot's KV matrices / key sums
        float* ksum = ksum_all + slot * 2 * 64;
        uint32_t it = 0, itq = 0;
        auto wait_acc = [&]() { tf_wait<64>(&ctl->to_epi[slot], it & 1); ++it; tc_fence_after(); };
        auto signal_mma = [&]() {
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->to_mma[slot]);
        };
        // this thread's 32 channels of (side, match m, cell r); zeros for padding lanes / matches
        auto load_rows = [&](float (&dst)[32], int side, int m) {
            const bool ok = r < WW && m < M;
            const float* src = p.in[side] + (static_cast<size_t>(ok ? m : 0) * WW + (ok ? r : 0)) * kC + 32 * g;
            if (ok) {
#pragma unroll
                for (int c = 0; c < 4; ++c) ldg256(src, dst, 8 * c);
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) dst[i] = 0.f;
            }
        };
        auto prefetch_rows = [&](int side, int m) {
            if (r < WW && m < M)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p.in[side] + (static_cast<size_t>(m) * WW + r) * kC + 32 * g));
        };
        // (the LayerNorm parameters come from the constant bank: the epilogue warps do not wait for the weight blob)

        [[maybe_unused]] int tn = 0;
        [[maybe_unused]] const bool tr = (warp & 7) == 0 && lane == 0;
#define TFE(id) do { if (tr) TFT(0, slot, tn, 16 * pass + (id)); } while (0)
        for (int item = blockIdx.x * 2 + slot; item < items; item += 2 * gridDim.x, ++tn) {
            const int next = item + 2 * gridDim.x;
            { [[maybe_unused]] const int pass = 0; TFE(0); }
            // match of this thread's row; self: both windows of one match, cross: same side of two matches
            const int m = kCross ? 2 * item + w : item;
            const bool valid = r < WW && m < M;
            float x[32];
            if (kCross) {
                load_rows(x, 1, m);                                   // the image-1 windows: K | V source of pass 0
                store_operand32(tl + 192 + 16 * g, tl + 224 + 16 * g, x);
                signal_mma();
                load_rows(x, 0, m);                                   // in flight under the K | V phase
                if (next < items) { prefetch_rows(1, 2 * next + w); prefetch_rows(0, 2 * next + w); }
            } else {
                load_rows(x, w, m);
                store_operand32(tl + 192 + 16 * g, tl + 224 + 16 * g, x);
                signal_mma();
                if (next < items) prefetch_rows(w, next);
            }
            { [[maybe_unused]] const int pass = 0; TFE(1); }

#pragma unroll 1
            for (int pass = 0; pass < kPasses; ++pass) {
                // ================= attention: K, V -> per-window KV matrices =================
                wait_acc();   // self: Q | K | V ready; cross: K | V of the source rows
                TFE(2);
                float qf[32];
                uint32_t v[32];
                if (kCross) {
                    // K | V of the source rows are in the accumulators, their operand is free: the rows
                    // to update take the tile now, and their Q projection runs under the K^T V phase
                    store_operand32(tl + 192 + 16 * g, tl + 224 + 16 * g, x);
                    signal_mma();
                }
                // (nothing of the residual stream is held in registers across the attention phases: the K^T V
                // reduction and the message are chains of shared-memory loads whose latency only hides if the
                // compiler has registers to keep a dozen of them in flight -- with x[32] and Q[32] live they ran
                // at 45 and 72 cycles per iteration (tools/dbg_tf_trace.py); x is re-read from L2 before LayerNorm 2)
                // The K | V scratch is shared by the two slots: held until this slot's K^T V sums are
                // done (every warp signs off after its last scratch read).
                TFE(3);
                if (st == 0) {
                    uint32_t spins = 0;
                    while (atomicCAS(&ctl->lock_owner, 0, slot + 1) != 0) {
                        __nanosleep(32);
                        if (++spins > 40000000u) __trap();   // protocol bug -> launch failure, not a hung GPU
                    }
                    ctl->lock_left = 8;
                }
                named_bar_sync(bar_id, kSlotThreads);
                TFE(4);
                {
                    uint8_t* rowp = scr + (w * kMaxWW + (r < WW ? r : 0)) * kScrRow;
                    tmem_ld_32x32(tl + 64 + 32 * g, v);
                    tmem_ld_wait();
                    if (r < WW) {
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            *reinterpret_cast<float4*>(rowp + ((8 * g + c) << 4)) =
                                make_float4(elu1(__uint_as_float(v[4 * c])), elu1(__uint_as_float(v[4 * c + 1])),
                                            elu1(__uint_as_float(v[4 * c + 2])), elu1(__uint_as_float(v[4 * c + 3])));
                    }
                    tmem_ld_32x32(tl + 128 + 32 * g, v);
                    tmem_ld_wait();
                    if (r < WW) {
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            *reinterpret_cast<uint4*>(rowp + ((16 + 8 * g + c) << 4)) =
                                make_uint4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
                    }
                }
                named_bar_sync(bar_id, kSlotThreads);
                TFE(5);
                {   // KV[d][v] = sum_r K[r][d] V[r][v] within a head; thread = (window, d, half of the 8 v)
                    const int w2 = st >> 7, d = (st & 127) >> 1, vh = st & 1;
                    const uint8_t* kp = scr + w2 * kMaxWW * kScrRow + d * 4;
                    const uint8_t* vp = scr + w2 * kMaxWW * kScrRow + 256 + (d >> 3) * 32 + vh * 16;
                    uint64_t a01 = pack2(0.f, 0.f), a23 = a01;
                    float ks = 0.f;
#pragma unroll
                    for (int rr = 0; rr < kMaxWW; ++rr) {
                        if (kWW ? rr < kWW : rr < WW) {
                            const float kv = *reinterpret_cast<const float*>(kp + rr * kScrRow);
                            const float4 vv = *reinterpret_cast<const float4*>(vp + rr * kScrRow);
                            const uint64_t k2 = pack2(kv, kv);
                            a01 = fma2(k2, pack2(vv.x, vv.y), a01);
                            a23 = fma2(k2, pack2(vv.z, vv.w), a23);
                            ks += kv;
                        }
                    }
                    float a0, a1, a2, a3;
                    unpack2(a01, a0, a1);
                    unpack2(a23, a2, a3);
                    *reinterpret_cast<float4*>(kvm + (w2 * 64 + d) * 8 + vh * 4) = make_float4(a0, a1, a2, a3);
                    if (vh == 0) ksum[w2 * 64 + d] = ks;
                }
                __syncwarp();
                if (lane == 0) {   // this warp is done with the shared scratch
                    __threadfence_block();
                    if (atomicSub(&ctl->lock_left, 1) == 1) atomicExch(&ctl->lock_owner, 0);
                }
                TFE(6);
                if (kCross) {
                    tf_wait<64>(&ctl->q_done[slot], itq & 1);   // Q (ran under the K^T V phase)
                    ++itq;
                    tc_fence_after();
                }
                tmem_ld_32x32(tl + 32 * g, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) qf[i] = elu1(__uint_as_float(v[i]));
                named_bar_sync(bar_id, kSlotThreads);   // KV matrices visible
                TFE(7);
                // ================= message = (Q KV) / (Q sum K + eps)  (attentions.py:42-44) =================
#pragma unroll
                for (int hh = 0; hh < 4; ++hh) {
                    const int h = 4 * g + hh;
                    const float* kvp = kvm + (w * 64 + h * 8) * 8;
                    const float* ksp = ksum + w * 64 + h * 8;
                    float den = 1e-6f;
                    uint64_t mm[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) mm[j] = pack2(0.f, 0.f);
#pragma unroll
                    for (int dd = 0; dd < 8; ++dd) {
                        const float qd = qf[hh * 8 + dd];
                        const float4 ka = *reinterpret_cast<const float4*>(kvp + dd * 8);
                        const float4 kb = *reinterpret_cast<const float4*>(kvp + dd * 8 + 4);
                        const uint64_t q2 = pack2(qd, qd);
                        den = fmaf(qd, ksp[dd], den);
                        mm[0] = fma2(q2, pack2(ka.x, ka.y), mm[0]);
                        mm[1] = fma2(q2, pack2(ka.z, ka.w), mm[1]);
                        mm[2] = fma2(q2, pack2(kb.x, kb.y), mm[2]);
                        mm[3] = fma2(q2, pack2(kb.z, kb.w), mm[3]);
                    }
                    const float z = 1.0f / den;
                    const uint64_t z2 = pack2(z, z);
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float ma, mb;
                        unpack2(mul2(mm[j], z2), ma, mb);
                        split2(ma, mb, hi[j], lo[j]);
                    }
                    tmem_st_32x4(tl + 128 + 16 * g + 4 * hh, hi);
                    tmem_st_32x4(tl + 160 + 16 * g + 4 * hh, lo);
                }
                signal_mma();
                TFE(8);
                // ================= merge -> LayerNorm 1 =================
                wait_acc();
                TFE(9);
                float y[32];
                tmem_ld_32x32(tl + 32 * g, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) y[i] = __uint_as_float(v[i]);
                float2 mine = local_stats(y);
                lnrow[g] = mine;
                named_bar_sync(bar_id, kSlotThreads);
                layer_norm32(y, mine, lnrow[g ^ 1], ln_slot, ln_off, ln_off + 64);
                store_operand32(tl + 128 + 16 * g, tl + 160 + 16 * g, y);
                signal_mma();
                TFE(10);
                // ================= MLP hidden: relu =================
                wait_acc();
                TFE(11);
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    tmem_ld_32x32(tl + 64 * g + 32 * half, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) y[i] = fmaxf(__uint_as_float(v[i]), 0.f);
                    store_operand32(tl + 192 + 32 * g + 16 * half, tl + 128 + 32 * g + 16 * half, y);
                }
                signal_mma();
                TFE(12);
                // ================= MLP out -> LayerNorm 2 -> residual =================
                load_rows(x, kCross ? pass : w, m);   // the rows being updated, again (L2): in flight under the MLP-2 batch
                wait_acc();
                TFE(13);
                tmem_ld_32x32(tl + 32 * g, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) y[i] = __uint_as_float(v[i]);
                mine = local_stats(y);
                lnrow[g] = mine;
                named_bar_sync(bar_id, kSlotThreads);
                layer_norm32(y, mine, lnrow[g ^ 1], ln_slot, ln_off + 128, ln_off + 192);
#pragma unroll
                for (int i = 0; i < 32; ++i) x[i] += y[i];
                TFE(14);
                if (kCross && pass == 0) {
                    // the updated image-0 windows are the K | V source of pass 1 (transformer.py:95)
                    store_operand32(tl + 192 + 16 * g, tl + 224 + 16 * g, x);
                    signal_mma();
                }
                if (valid) {   // four full 32-byte sectors per thread
                    float* dst = p.out[kCross ? pass : w] + (static_cast<size_t>(m) * WW + r) * kC + 32 * g;
#pragma unroll
                    for (int c = 0; c < 4; ++c) stg256(dst, x, 8 * c);
                }
                if (kCross && pass == 0) load_rows(x, 1, m);   // pass 1 updates the image-1 windows
                TFE(15);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 18) tmem_dealloc<512>(tmem_base);
}

constexpr size_t kTfSmem = 1024 + kBlobBytes + kScratchBytes + 2 * (2 * 64 * 8 * 4 + 2 * 64 * 4) + 2 * 128 * 2 * 8 + sizeof(TfCtl);
static_assert(kTfSmem <= 232448, "fine transformer shared memory budget");

}  // namespace

bool fine_transformer_supported(int C, int nhead, int WW) { return C == kC && nhead == kHeads && WW >= 1 && WW <= kMaxWW; }

size_t fine_transformer_workspace_bytes(int n_layers) { return static_cast<size_t>(n_layers) * kBlobBytes; }

// layer_ptrs: n_layers x 10 device pointers in the order q, k, v, merge, mlp.0, mlp.2, norm1.weight,
// norm1.bias, norm2.weight, norm2.bias;  kinds[l]: 0 = 'self', 1 = 'cross'.
cudaError_t launch_fine_transformer(const float* win0, const float* win1, const int* count, int capacity, int WW,
                                    int n_layers, const int* kinds, const float* const* layer_ptrs, float* out0,
                                    float* out1, void* workspace, int num_sms, cudaStream_t st) {
    uint8_t* blobs = static_cast<uint8_t*>(workspace);
    for (int l = 0; layer_ptrs && l < n_layers; ++l) {   // NULL: packed by an earlier call into this workspace
        const float* const* q = layer_ptrs + 10 * l;
        TfLayerParams lp{q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7], q[8], q[9]};
        tf_pack_kernel<<<20, 256, 0, st>>>(lp, blobs + static_cast<size_t>(l) * kBlobBytes); count_launch();
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    using Kern = void (*)(const TfParams);
    const Kern kern[2] = {WW == 49 ? tf_layer_kernel<false, 49> : tf_layer_kernel<false, 0>,
                          WW == 49 ? tf_layer_kernel<true, 49> : tf_layer_kernel<true, 0>};
    for (int k = 0; k < 2; ++k) {
        e = cudaFuncSetAttribute(kern[k], cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kTfSmem));
        if (e != cudaSuccess) return e;
    }
    const int grid = std::max(1, std::min(num_sms, (capacity + 1) / 2));
    for (int l = 0; l < n_layers; ++l) {
        TfParams p{};
        // layer 0 reads the caller's windows; later layers update the output in place
        p.in[0] = l == 0 ? win0 : out0; p.in[1] = l == 0 ? win1 : out1;
        p.out[0] = out0; p.out[1] = out1;
        p.count = count; p.capacity = capacity; p.WW = WW;
        p.blob = blobs + static_cast<size_t>(l) * kBlobBytes;
        static std::atomic<unsigned> next_slot{0};
        p.ln_slot = static_cast<int>(next_slot.fetch_add(1, std::memory_order_relaxed) % kLnSlots);
        e = cudaMemcpyToSymbolAsync(c_tf_ln, p.blob + kOffLn, 256 * sizeof(float), static_cast<size_t>(p.ln_slot) * 256 * sizeof(float),
                                    cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return e;
        kern[kinds[l] ? 1 : 0]<<<grid, kTfThreads, kTfSmem, st>>>(p); count_launch();
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

#ifdef TF_TRACE
extern "C" __attribute__((visibility("default"))) int fm_debug_tf_trace(long long* host_out) {
    return static_cast<int>(cudaMemcpyFromSymbol(host_out, g_tf_trace, sizeof(g_tf_trace)));
}
#endif

}  // namespace fm
