// Shared definitions for the coarse/fine kernels and the C-ABI layer.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/fm_b200.h"

namespace fm {

// ---------------------------------------------------------------- coarse-stage tiling
constexpr int kTileM = 128;      // image-0 tokens per MMA tile (TMEM lanes)
// K1's block of image-1 tokens = kColGroups column groups of kGroupCols columns; one epilogue warpgroup (4 warps, one
// per TMEM lane quarter) per group, a thread owning one accumulator row x kGroupCols columns.  Shipped: 256 = 2 x 128
// (8 epilogue warps, 232 registers each).  -DFM_K1_NARROW builds 192 = 3 x 64 (12 epilogue warps = 3 per scheduler,
// 152 registers): measured SLOWER (0.49 vs 0.41 ms, r02) -- the MMA phase of this kernel runs at the rate the
// shared-memory operand reads allow (SS MMAs: 4 KB of A + 32 N bytes of B per 128 x N x 16 instruction), and a
// narrower block reads the same A for fewer flops (tools/dbg_k1_trace.py: 3 180 cycles per 128 x 192 tile against
// 3 250 per 128 x 256 tile on the MMA thread).
#ifdef FM_K1_NARROW
constexpr int kColGroups = 3, kGroupCols = 64;
#else
constexpr int kColGroups = 2, kGroupCols = 128;
#endif
constexpr int kBlockN = kColGroups * kGroupCols;     // image-1 tokens resident per work unit (TMEM columns)
constexpr int kChunkK = 64;      // bf16 elements per 128-byte swizzle row
constexpr int kMaxC = 256;       // descriptor channels supported by the resident-B layout
#ifndef FM_K1_ASTAGES
#define FM_K1_ASTAGES 5
#endif
constexpr int kAStages = FM_K1_ASTAGES;      // A-operand ring (16 KB per stage)
constexpr int kAStageBytes = kTileM * kChunkK * 2;    // 16384
constexpr int kBChunkBytes = kBlockN * kChunkK * 2;   // 24576 (32768 wide)
constexpr int kEpiWarps = 4 * kColGroups;
constexpr int kCtlWarps = 4;     // warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM owner, warp 3 idle
constexpr int kCoarseThreads = (kCtlWarps + kEpiWarps) * 32;   // control warpgroup + one epilogue warpgroup per column group
constexpr int kRowSlots = 4;     // in-row slots for exactly tied matches of one image-0 token ...
constexpr int kTieCap = 16384;   // ... any further ones go to one shared list of (row, j, conf) entries

// log2-domain head-room of the per-row exponent reference: p = 2^(t - c_i) with
// c_i = U_i - kHead2 and U_i >= every score of the row, so p <= 2^kHead2 and sums over up to
// 2^17 tokens stay finite; a row (column) is "lost" only if its best score is more than
// 126 + kHead2 log2-units (157 nats) below its Cauchy-Schwarz bound.
constexpr float kHead2 = 100.0f;
constexpr float kCandMargin = 0.9f;  // candidate pre-filter uses kCandMargin * thr

// status word layout (int32[8], device): written by the coarse kernels
enum StatusSlot { kStCount = 0, kStFlags = 1, kStMaxUnitCand = 2, kStLost = 3, kStCandTotal = 4, kStRedo = 5 };

// Upper bound (log2 domain, slightly inflated) on every score of a token with norm n0 against
// tokens with norm <= maxn1; used identically by the stats and resolve kernels.
__device__ __forceinline__ float score_bound2(float n0, float maxn1, float a2) {
    return __fmul_rn(__fmul_rn(__fmul_rn(n0, maxn1), a2), 1.001953125f);
}

struct CoarseDims {
    int B, N0, N1, C, KC;      // KC = C / 64
    int MT, NB;                // M tiles (128 rows), N blocks (kBlockN columns)
    int S, tiles_per_split;    // M range split into S contiguous pieces per (pair, N block)
    int num_units;             // B * NB * S
    int N0pad, N1pad;          // MT * 128, NB * kBlockN
    int num_pools;             // hot-chunk record pools (= CTAs of the statistics kernel)
    int hot_cap;               // records per pool
};

// A "hot chunk" = 32 consecutive scores of one image-0 token, at least one of which may exceed
// the confidence threshold.  K1 stores the raw dot products; K2 evaluates all 32 exactly.
constexpr int kHotChunk = 32;

constexpr int kMaxResolveBlocks = 4096;   // grid bound of the cooperative resolve kernel (co-resident blocks)

struct CoarseWorkspace {
    // all pointers into the caller's workspace
    __nv_bfloat16* f0b;   // [B, N0, C] bf16 copy (only if the input was fp32)
    __nv_bfloat16* f1b;
    float* norm0;         // [B, N0pad]
    float* maxnorm;       // [B, 2] (float bits, atomicMax)
    float* rowpart;       // [B, NB * kColGroups, N0pad]
    float* colpart;       // [B, S, N1pad]
    uint2* hot_meta;      // [num_pools, hot_cap]  x = b * N0 + i, y = first column / 32
    float* hot_dots;      // [num_pools, hot_cap, 32]  raw <f0_i, f1_j> of the chunk's 32 columns
    int* hot_count;       // [num_pools]
    float* l2row;         // [B, N0]  log2 sum_j 2^t_ij
    float* l2col;         // [B, N1]
    unsigned* rowbest;    // [B, N0]  float bits of best strong confidence
    unsigned* colbest;    // [B, N1]
    int* rowcnt;          // [B * N0]
    int2* rowslot;        // [B * N0, kRowSlots]  (j, conf bits)
    int* tie_count;       // entries in tie_list (may exceed kTieCap: then FM_FLAG_TIE_OVERFLOW)
    int4* tie_list;       // [kTieCap] (row, j, conf bits, 0): matches beyond a row's kRowSlots
    int* redo;            // 1 = the single-pass candidate filter overflowed: re-extract with exact row sums
    int* blocksum;        // [kMaxResolveBlocks] per-block match counts of the resolve kernel
    size_t zero_begin, zero_bytes;   // region cleared at the start of every call
    size_t total_bytes;
};

struct CoarseStatsParams {
    CoarseDims d;
    float a2;          // log2(e) / (C * temperature)
    float cand_thr;    // kCandMargin * thr
    const float* norm0;
    const float* maxnorm;
    float* rowpart;
    float* colpart;
    uint2* hot_meta;
    float* hot_dots;
    int* hot_count;
    int* status;
    // Second ("exact") extraction pass, taken only when the first one overflowed its record pools: with both
    // soft-max denominators known (finish_lse) a chunk is hot iff one of its entries has conf > cand_thr.
    // Diffuse rows (random-init networks, thr << 1/128) make every chunk hot against a 128-column partial
    // sum; against the confidence itself only the handful of real candidates remain.
    const float* l2row;   // [B, N0] from finish_lse, or NULL in the first pass
    const float* l2col;   // [B, N1] (+ kBlockN floats of slack)
    const int* redo;      // the pass runs only if *redo != 0 (NULL: always)
};

CoarseDims coarse_dims(int B, int N0, int N1, int C, int cand_per_row, int num_sms);
CoarseWorkspace coarse_workspace(const CoarseDims& d, bool fp32_input, void* base);

// kernels' host launchers (each returns a cudaError_t from the launch)
cudaError_t launch_coarse_prep(const void* f0, const void* f1, int dtype, const CoarseDims& d,
                               const CoarseWorkspace& w, cudaStream_t st);
cudaError_t launch_coarse_stats_tc(const __nv_bfloat16* f0, const __nv_bfloat16* f1,
                                   const CoarseStatsParams& p, int num_sms, cudaStream_t st);
cudaError_t launch_coarse_stats_simt(const __nv_bfloat16* f0, const __nv_bfloat16* f1,
                                     const CoarseStatsParams& p, cudaStream_t st);

struct CoarseResolveParams {
    CoarseDims d;
    float a2, thr;
    int h0c, w0c, h1c, w1c, border;
    float scale;
    const float* scale0;   // optional [B, 2]
    const float* scale1;
    int64_t* b_ids; int64_t* i_ids; int64_t* j_ids;
    float* mconf; float* mkpts0_c; float* mkpts1_c;
    int capacity;
    int* status;
};
cudaError_t launch_coarse_finish_lse(const CoarseResolveParams& p, const CoarseWorkspace& w, bool with_gate,
                                     cudaStream_t st);
cudaError_t launch_coarse_resolve(const CoarseResolveParams& p, const CoarseWorkspace& w,
                                  cudaStream_t st);

bool backbone_emit_supported(int Cin, int Cout);
cudaError_t launch_backbone_emit(const float* x, long long rows, int Cin, int Cout, const float* ln_w, const float* ln_b,
                                 float eps, const float* lin_w, float* out_f32, void* out_bf16, int num_sms,
                                 cudaStream_t st);

void set_error(const char* fmt, ...);

// Kernel launches issued by this library since load (fm_launch_count): every launcher bumps it right after a
// <<<>>>.  Host-side bookkeeping only; bench.py turns the difference over its timed region into `gpu_launches`.
void count_launch(int n = 1);

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) belongs to a device's context: `mask` (one static per
// kernel) remembers the devices it was set on.  True = the caller must set it for the current device.
inline bool needs_attr_on_this_device(unsigned long long* mask) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
    if ((*mask >> dev) & 1ull) return false;
    *mask |= 1ull << dev;
    return true;
}

}  // namespace fm
