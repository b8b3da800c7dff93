// C ABI of libfm_b200.so (declared in include/fm_b200.h): argument checks, workspace carving and
// kernel sequencing.  No device allocation, no host synchronisation, no CPU fallback.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>
#include <vector>

#include "fm_common.cuh"

namespace fm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static int cuda_fail(cudaError_t e, const char* what) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return FM_ERR_CUDA;
}

// ---------------------------------------------------------------- stage timers (fm_profile_*)
struct StageTimers {
    std::mutex mu;
    bool on = false;
    std::vector<cudaEvent_t> pool;   // pairs: [2k] start, [2k + 1] stop
    size_t used = 0;                 // events handed out since the last collect
};
static StageTimers g_timers;
constexpr size_t kMaxTimedCalls = 4096;

// Hands out the next (start, stop) pair, or false when disabled / full.
static bool timer_pair(cudaEvent_t* a, cudaEvent_t* b) {
    std::lock_guard<std::mutex> lk(g_timers.mu);
    if (!g_timers.on || g_timers.used >= 2 * kMaxTimedCalls) return false;
    while (g_timers.pool.size() < g_timers.used + 2) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return false;
        g_timers.pool.push_back(e);
    }
    *a = g_timers.pool[g_timers.used];
    *b = g_timers.pool[g_timers.used + 1];
    g_timers.used += 2;
    return true;
}

static int sm_count() {   // of the CURRENT device (cached per device)
    static int cached[64] = {0};
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (dev >= 0 && dev < 64 && cached[dev] > 0) return cached[dev];
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    if (dev >= 0 && dev < 64) cached[dev] = n;
    return n;
}

// ---------------------------------------------------------------- coarse work decomposition
// Work unit = (pair, 256-column block, contiguous run of `tiles_per_split` 128-row tiles); units
// are dealt round-robin to a persistent grid.  S (runs per column block) is chosen to minimise
// the makespan of that static schedule, charging each unit a fixed overhead (loading the
// resident block, draining the pipeline, reducing the column sums) of ~1 tile.
CoarseDims coarse_dims(int B, int N0, int N1, int C, int cand_per_row, int num_sms) {
    // the schedule search costs a fraction of a millisecond: remember the last shape per thread
    static thread_local int key[6] = {0, 0, 0, 0, 0, 0};
    static thread_local CoarseDims cached{};
    if (key[0] == B && key[1] == N0 && key[2] == N1 && key[3] == C && key[4] == cand_per_row && key[5] == num_sms)
        return cached;
    CoarseDims d{};
    d.B = B; d.N0 = N0; d.N1 = N1; d.C = C; d.KC = C / kChunkK;
    d.MT = (N0 + kTileM - 1) / kTileM;
    d.NB = (N1 + kBlockN - 1) / kBlockN;
    d.N0pad = d.MT * kTileM;
    d.N1pad = d.NB * kBlockN;
    if (num_sms < 1) num_sms = 148;
    // cost of a unit boundary in tiles: new resident block + column-sum reduction = 3 300 of a tile's 3 850 cycles,
    // plus the pipeline refill (r02 timeline, tools/dbg_k1_trace.py); 1.0 under-counted it and chose too many units
    const double ovh = 2.5;
    double best = 1e300;
    int best_tps = d.MT;
    int prev_tps = -1;
    for (int S = 1; S <= d.MT; ++S) {
        const int tps = (d.MT + S - 1) / S;
        if (tps == prev_tps) continue;
        prev_tps = tps;
        const int Se = (d.MT + tps - 1) / tps;
        const long long units = 1LL * B * d.NB * Se;
        const int grid = units < num_sms ? static_cast<int>(units) : num_sms;
        // CTA c runs units c, c + grid, ...; unit u has tiles(u % Se)
        const int last_tiles = d.MT - (Se - 1) * tps;
        double makespan = 0.0;
        if (units <= 262144) {
            for (int c = 0; c < grid; ++c) {
                double load = 0.0;
                for (long long u = c; u < units; u += grid) load += ovh + ((u % Se) == Se - 1 ? last_tiles : tps);
                if (load > makespan) makespan = load;
            }
        } else {   // many rounds: the average is accurate enough
            const double per_unit = ovh + static_cast<double>(d.MT) / Se;
            makespan = per_unit * static_cast<double>((units + grid - 1) / grid);
        }
        if (makespan < best * 0.999) { best = makespan; best_tps = tps; }
    }
    d.tiles_per_split = best_tps;
    d.S = (d.MT + best_tps - 1) / best_tps;
    d.num_units = B * d.NB * d.S;
    d.num_pools = d.num_units < num_sms ? d.num_units : num_sms;
    // planted / real data give about one hot chunk per image-0 token; cand_per_row = 4 leaves 2x
    const long long want = (1LL * cand_per_row * B * N0 + 2LL * d.num_pools - 1) / (2LL * d.num_pools) + 64;
    d.hot_cap = want > (1 << 24) ? (1 << 24) : static_cast<int>(want);
    key[0] = B; key[1] = N0; key[2] = N1; key[3] = C; key[4] = cand_per_row; key[5] = num_sms;
    cached = d;
    return d;
}

static size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

CoarseWorkspace coarse_workspace(const CoarseDims& d, bool fp32_input, void* base) {
    CoarseWorkspace w{};
    uint8_t* p = static_cast<uint8_t*>(base);
    size_t off = 0;
    auto take = [&](size_t bytes) { uint8_t* r = p ? p + off : nullptr; off = align_up(off + bytes); return r; };
    const size_t B = d.B;
    // ---- region cleared at the start of every call
    w.zero_begin = off;
    w.maxnorm = reinterpret_cast<float*>(take(sizeof(float) * B * 2));
    w.hot_count = reinterpret_cast<int*>(take(sizeof(int) * d.num_pools));
    w.rowbest = reinterpret_cast<unsigned*>(take(sizeof(unsigned) * B * d.N0));
    w.colbest = reinterpret_cast<unsigned*>(take(sizeof(unsigned) * B * d.N1));
    w.rowcnt = reinterpret_cast<int*>(take(sizeof(int) * B * d.N0));
    w.tie_count = reinterpret_cast<int*>(take(sizeof(int)));
    w.redo = reinterpret_cast<int*>(take(sizeof(int)));
    w.zero_bytes = off - w.zero_begin;
    // ---- fully rewritten by every call
    w.norm0 = reinterpret_cast<float*>(take(sizeof(float) * B * d.N0pad));
    w.rowpart = reinterpret_cast<float*>(take(sizeof(float) * B * d.NB * kColGroups * d.N0pad));
    w.colpart = reinterpret_cast<float*>(take(sizeof(float) * B * d.S * d.N1pad));
    w.l2row = reinterpret_cast<float*>(take(sizeof(float) * B * d.N0));
    w.l2col = reinterpret_cast<float*>(take(sizeof(float) * (B * d.N1 + kBlockN)));   // slack: the exact pass reads whole blocks
    w.rowslot = reinterpret_cast<int2*>(take(sizeof(int2) * B * d.N0 * kRowSlots));
    w.tie_list = reinterpret_cast<int4*>(take(sizeof(int4) * kTieCap));
    w.blocksum = reinterpret_cast<int*>(take(sizeof(int) * kMaxResolveBlocks));
    w.hot_meta = reinterpret_cast<uint2*>(take(sizeof(uint2) * static_cast<size_t>(d.num_pools) * d.hot_cap));
    w.hot_dots = reinterpret_cast<float*>(take(sizeof(float) * kHotChunk * static_cast<size_t>(d.num_pools) * d.hot_cap));
    if (fp32_input) {
        w.f0b = reinterpret_cast<__nv_bfloat16*>(take(sizeof(__nv_bfloat16) * B * d.N0 * d.C));
        w.f1b = reinterpret_cast<__nv_bfloat16*>(take(sizeof(__nv_bfloat16) * B * d.N1 * d.C));
    }
    w.total_bytes = off;
    return w;
}

cudaError_t launch_fine_gather(const void* ff0, const void* ff1, const void* fc0, const void* fc1, int featc_dtype,
                               const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const int* count,
                               int capacity, int B, int Hf, int Wf, int N0, int N1, int Cc, int w0c, int w1c,
                               int stride, int pad, int W, const float* down_w, const float* down_b,
                               const float* merge_w, const float* merge_b, void* win0, void* win1, int out_dtype,
                               int num_sms, cudaStream_t st);
bool fine_gather_tc_supported(int Cf, int Cc, int W);
size_t fine_gather_tc_workspace_bytes(int B, int N0, int N1, int Cc, bool f32_tokens);
cudaError_t launch_fine_gather_tc(const void* ff0, const void* ff1, const void* fc0, const void* fc1, int featc_dtype,
                                  const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const int* count,
                                  int capacity, int B, int Hf, int Wf, int N0, int N1, int Cc, int w0c, int w1c,
                                  int stride, int pad, int W, const float* down_w, const float* down_b,
                                  const float* merge_w, const float* merge_b, void* win0, void* win1, int out_dtype,
                                  void* workspace, size_t workspace_bytes, int num_sms, cudaStream_t st);
cudaError_t launch_to_nhwc(const void* src, int src_dtype, int B, int C, int H, int W, void* dst, cudaStream_t st);
cudaError_t launch_pack_matches(const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const float* mconf,
                                const float* mk0, const float* mk1, const int* count, int capacity, int pair_offset,
                                void* records, int* chunk_counts, int chunk_index, int total_capacity, int num_sms,
                                cudaStream_t st);
cudaError_t launch_fine_match(const void* win0, const void* win1, int dtype, const float* mk0, const float* mk1,
                              const float* w0, const float* b0, const float* w1, const float* b1, const int* count,
                              int capacity, int WW, float scale, float* out0, float* out1, int num_sms,
                              cudaStream_t st);

bool coarse_transformer_supported(int C, int nhead);
size_t coarse_transformer_workspace_bytes(int B, int L0, int L1, int n_layers);
cudaError_t launch_coarse_transformer(const float* feat0, const float* feat1, int B, int L0, int L1, int n_layers,
                                      const int* layer_kinds, const void* const* layer_params, float* out0,
                                      float* out1, void* bf0, void* bf1, void* workspace, int num_sms,
                                      cudaStream_t st);
bool fine_transformer_supported(int C, int nhead, int WW);
size_t fine_transformer_workspace_bytes(int n_layers);
cudaError_t launch_fine_transformer(const float* win0, const float* win1, const int* count, int capacity, int WW,
                                    int n_layers, const int* kinds, const float* const* layer_ptrs, float* out0,
                                    float* out1, void* workspace, int num_sms, cudaStream_t st);

size_t pose_workspace_bytes(int B, int capacity, int H);
cudaError_t launch_pose_from_matches(const int64_t* b_ids, const float* mk0, const float* mk1, int stride, const int* count,
                                     int capacity, int B, const float* K, float thr_px, int H, int flags, uint32_t seed,
                                     float* R, float* t, float* E, int* n_inliers, int* status, uint8_t* mask,
                                     void* workspace, cudaStream_t st);

cudaError_t launch_pose_count_valid(const void* workspace, int B, int capacity, int H, int* valid, cudaStream_t st);

static bool dtype_ok(int t) { return t == FM_DTYPE_BF16 || t == FM_DTYPE_F32; }

static int check_coarse_args(const void* f0, const void* f1, int dt, int B, int N0, int N1, int C, int h0c, int w0c,
                             int h1c, int w1c, float temperature, float thr, int capacity, const void* ws,
                             int cand_per_row) {
    if (!f0 || !f1 || !ws) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    if (!dtype_ok(dt)) { set_error("feat_dtype must be FM_DTYPE_BF16 or FM_DTYPE_F32"); return FM_ERR_BAD_ARG; }
    if (B < 1 || N0 < 1 || N1 < 1 || capacity < 0) { set_error("B, N0, N1 must be >= 1 and capacity >= 0"); return FM_ERR_BAD_ARG; }
    if (C < 64 || C > kMaxC || C % 64) { set_error("C = %d unsupported: need a multiple of 64 in [64, %d]", C, kMaxC); return FM_ERR_UNSUPPORTED; }
    if (h0c * w0c != N0 || h1c * w1c != N1) { set_error("N0 != h0c * w0c or N1 != h1c * w1c"); return FM_ERR_BAD_ARG; }
    if (!(temperature > 0.f)) { set_error("temperature must be > 0"); return FM_ERR_BAD_ARG; }
    if (!(thr > 0.f) || !(thr < 1.f)) { set_error("thr must be in (0, 1): the candidate bound needs thr > 0"); return FM_ERR_BAD_ARG; }
    if (cand_per_row < 1 || cand_per_row > 256) { set_error("cand_per_row must be in [1, 256]"); return FM_ERR_BAD_ARG; }
    if (N0 > (1 << 22) || N1 > (1 << 22)) { set_error("token count too large"); return FM_ERR_UNSUPPORTED; }
    if ((reinterpret_cast<uintptr_t>(f0) | reinterpret_cast<uintptr_t>(f1) | reinterpret_cast<uintptr_t>(ws)) & 15) {
        set_error("feature and workspace pointers must be 16-byte aligned");
        return FM_ERR_BAD_ARG;
    }
    return FM_OK;
}

static int coarse_match_impl(bool simt, const void* feat_c0, const void* feat_c1, int feat_dtype, int B, int N0,
                             int N1, int C, int h0c, int w0c, int h1c, int w1c, float temperature, float thr,
                             int border_rm, float scale, const float* scale0, const float* scale1, int64_t* b_ids,
                             int64_t* i_ids, int64_t* j_ids, float* mconf, float* mkpts0_c, float* mkpts1_c,
                             int32_t* status, int capacity, void* workspace, size_t workspace_bytes,
                             int cand_per_row, void* stream) {
    int rc = check_coarse_args(feat_c0, feat_c1, feat_dtype, B, N0, N1, C, h0c, w0c, h1c, w1c, temperature, thr,
                               capacity, workspace, cand_per_row);
    if (rc != FM_OK) return rc;
    if (!status || (capacity > 0 && (!b_ids || !i_ids || !j_ids || !mconf || !mkpts0_c || !mkpts1_c))) {
        set_error("null output pointer");
        return FM_ERR_BAD_ARG;
    }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool f32 = feat_dtype == FM_DTYPE_F32;
    const CoarseDims d = coarse_dims(B, N0, N1, C, cand_per_row, sms);
    if (d.tiles_per_split * kTileM >= (1 << 24)) { set_error("row range per unit too large"); return FM_ERR_UNSUPPORTED; }
    const CoarseWorkspace w = coarse_workspace(d, f32, workspace);
    if (w.total_bytes > workspace_bytes) {
        set_error("workspace too small: need %zu bytes, got %zu", w.total_bytes, workspace_bytes);
        return FM_ERR_WORKSPACE;
    }
    cudaError_t e = cudaMemsetAsync(static_cast<uint8_t*>(workspace) + w.zero_begin, 0, w.zero_bytes, st);
    if (e != cudaSuccess) return cuda_fail(e, "memset(workspace)");
    e = cudaMemsetAsync(status, 0, sizeof(int32_t) * FM_STATUS_WORDS, st);
    if (e != cudaSuccess) return cuda_fail(e, "memset(status)");

    e = launch_coarse_prep(feat_c0, feat_c1, feat_dtype, d, w, st);
    if (e != cudaSuccess) return cuda_fail(e, "coarse_prep launch");
    const __nv_bfloat16* f0 = f32 ? w.f0b : static_cast<const __nv_bfloat16*>(feat_c0);
    const __nv_bfloat16* f1 = f32 ? w.f1b : static_cast<const __nv_bfloat16*>(feat_c1);

    CoarseStatsParams sp{};
    sp.d = d;
    sp.a2 = 1.4426950408889634f / (static_cast<float>(C) * temperature);
    sp.cand_thr = kCandMargin * thr;
    sp.norm0 = w.norm0; sp.maxnorm = w.maxnorm; sp.rowpart = w.rowpart; sp.colpart = w.colpart;
    sp.hot_meta = w.hot_meta; sp.hot_dots = w.hot_dots; sp.hot_count = w.hot_count; sp.status = status;
    cudaEvent_t ev0, ev1;
    const bool timed = !simt && timer_pair(&ev0, &ev1);
    if (timed) cudaEventRecord(ev0, st);
    e = simt ? launch_coarse_stats_simt(f0, f1, sp, st) : launch_coarse_stats_tc(f0, f1, sp, sms, st);
    if (timed) cudaEventRecord(ev1, st);
    if (e != cudaSuccess) return cuda_fail(e, simt ? "coarse_stats_simt launch" : "coarse_stats_tc launch");

    CoarseResolveParams rp{};
    rp.d = d; rp.a2 = sp.a2; rp.thr = thr;
    rp.h0c = h0c; rp.w0c = w0c; rp.h1c = h1c; rp.w1c = w1c; rp.border = border_rm;
    rp.scale = scale; rp.scale0 = scale0; rp.scale1 = scale1;
    rp.b_ids = b_ids; rp.i_ids = i_ids; rp.j_ids = j_ids; rp.mconf = mconf;
    rp.mkpts0_c = mkpts0_c; rp.mkpts1_c = mkpts1_c; rp.capacity = capacity; rp.status = status;
    // soft-max denominators; then -- only if the single-pass candidate filter overflowed (decided on the
    // device, no host round trip) -- the statistics kernel once more, testing hot chunks against the full row sums
    e = launch_coarse_finish_lse(rp, w, !simt, st);
    if (e != cudaSuccess) return cuda_fail(e, "coarse_finish_lse launch");
    if (!simt) {
        sp.l2row = w.l2row;
        sp.l2col = w.l2col;
        sp.redo = w.redo;
        e = launch_coarse_stats_tc(f0, f1, sp, sms, st);
        if (e != cudaSuccess) return cuda_fail(e, "coarse_stats_tc (exact pass) launch");
    }
    e = launch_coarse_resolve(rp, w, st);
    if (e != cudaSuccess) return cuda_fail(e, "coarse_resolve launch");
    return FM_OK;
}

}  // namespace fm

using namespace fm;

extern "C" {

int fm_version(void) { return FM_B200_VERSION; }

const char* fm_last_error_string(void) { return g_err; }

int fm_device_sm_count(void) { return sm_count(); }

long long fm_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

size_t fm_coarse_workspace_bytes(int B, int N0, int N1, int C, int feat_dtype, int cand_per_row) {
    if (B < 1 || N0 < 1 || N1 < 1 || C < 64 || C > kMaxC || C % 64 || !dtype_ok(feat_dtype) || cand_per_row < 1)
        return 0;
    const int sms = sm_count();
    const CoarseDims d = coarse_dims(B, N0, N1, C, cand_per_row, sms > 0 ? sms : 148);
    return coarse_workspace(d, feat_dtype == FM_DTYPE_F32, nullptr).total_bytes;
}

int fm_coarse_match(const void* feat_c0, const void* feat_c1, int feat_dtype, int B, int N0, int N1, int C, int h0c,
                    int w0c, int h1c, int w1c, float temperature, float thr, int border_rm, float scale,
                    const float* scale0, const float* scale1, int64_t* b_ids, int64_t* i_ids, int64_t* j_ids,
                    float* mconf, float* mkpts0_c, float* mkpts1_c, int32_t* status, int capacity, void* workspace,
                    size_t workspace_bytes, int cand_per_row, void* stream) {
    return coarse_match_impl(false, feat_c0, feat_c1, feat_dtype, B, N0, N1, C, h0c, w0c, h1c, w1c, temperature, thr,
                             border_rm, scale, scale0, scale1, b_ids, i_ids, j_ids, mconf, mkpts0_c, mkpts1_c, status,
                             capacity, workspace, workspace_bytes, cand_per_row, stream);
}

int fm_debug_coarse_match_simt(const void* feat_c0, const void* feat_c1, int feat_dtype, int B, int N0, int N1, int C,
                               int h0c, int w0c, int h1c, int w1c, float temperature, float thr, int border_rm,
                               float scale, const float* scale0, const float* scale1, int64_t* b_ids, int64_t* i_ids,
                               int64_t* j_ids, float* mconf, float* mkpts0_c, float* mkpts1_c, int32_t* status,
                               int capacity, void* workspace, size_t workspace_bytes, int cand_per_row, void* stream) {
    return coarse_match_impl(true, feat_c0, feat_c1, feat_dtype, B, N0, N1, C, h0c, w0c, h1c, w1c, temperature, thr,
                             border_rm, scale, scale0, scale1, b_ids, i_ids, j_ids, mconf, mkpts0_c, mkpts1_c, status,
                             capacity, workspace, workspace_bytes, cand_per_row, stream);
}

int fm_debug_coarse_lse(const void* workspace, int B, int N0, int N1, int C, int feat_dtype, int cand_per_row,
                        float* l2row, float* l2col, void* stream) {
    if (!workspace || !l2row || !l2col) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available"); return FM_ERR_CUDA; }
    const CoarseDims d = coarse_dims(B, N0, N1, C, cand_per_row, sms);
    const CoarseWorkspace w = coarse_workspace(d, feat_dtype == FM_DTYPE_F32, const_cast<void*>(workspace));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMemcpyAsync(l2row, w.l2row, sizeof(float) * B * N0, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(l2col, w.l2col, sizeof(float) * B * N1, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "copy lse");
    return FM_OK;
}

int fm_profile_coarse_stats(const void* feat_c0, const void* feat_c1, int B, int N0, int N1, int C, float temperature,
                            float thr, void* workspace, size_t workspace_bytes, int cand_per_row, int iters, float* ms,
                            void* stream) {
    if (!feat_c0 || !feat_c1 || !workspace || !ms || iters < 1) { set_error("bad argument"); return FM_ERR_BAD_ARG; }
    if (C < 64 || C > kMaxC || C % 64) { set_error("C unsupported"); return FM_ERR_UNSUPPORTED; }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available"); return FM_ERR_CUDA; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const CoarseDims d = coarse_dims(B, N0, N1, C, cand_per_row, sms);
    const CoarseWorkspace w = coarse_workspace(d, false, workspace);
    if (w.total_bytes > workspace_bytes) { set_error("workspace too small"); return FM_ERR_WORKSPACE; }
    int32_t* status = reinterpret_cast<int32_t*>(w.blocksum);   // scratch: results are not used
    cudaError_t e = cudaMemsetAsync(static_cast<uint8_t*>(workspace) + w.zero_begin, 0, w.zero_bytes, st);
    if (e != cudaSuccess) return cuda_fail(e, "memset");
    e = launch_coarse_prep(feat_c0, feat_c1, FM_DTYPE_BF16, d, w, st);
    if (e != cudaSuccess) return cuda_fail(e, "coarse_prep launch");
    CoarseStatsParams sp{};
    sp.d = d;
    sp.a2 = 1.4426950408889634f / (static_cast<float>(C) * temperature);
    sp.cand_thr = kCandMargin * thr;
    sp.norm0 = w.norm0; sp.maxnorm = w.maxnorm; sp.rowpart = w.rowpart; sp.colpart = w.colpart;
    sp.hot_meta = w.hot_meta; sp.hot_dots = w.hot_dots; sp.hot_count = w.hot_count; sp.status = status;
    cudaEvent_t ev0, ev1;
    cudaEventCreate(&ev0);
    cudaEventCreate(&ev1);
    const __nv_bfloat16* f0 = static_cast<const __nv_bfloat16*>(feat_c0);
    const __nv_bfloat16* f1 = static_cast<const __nv_bfloat16*>(feat_c1);
    e = launch_coarse_stats_tc(f0, f1, sp, sms, st);   // warm-up
    cudaEventRecord(ev0, st);
    for (int k = 0; k < iters && e == cudaSuccess; ++k) e = launch_coarse_stats_tc(f0, f1, sp, sms, st);
    cudaEventRecord(ev1, st);
    cudaError_t es = cudaEventSynchronize(ev1);
    float total = 0.f;
    cudaEventElapsedTime(&total, ev0, ev1);
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    if (e != cudaSuccess) return cuda_fail(e, "coarse_stats_tc launch");
    if (es != cudaSuccess) return cuda_fail(es, "coarse_stats_tc run");
    *ms = total / iters;
    return FM_OK;
}

int fm_fine_to_nhwc_bf16(const void* src, int src_dtype, int B, int C, int H, int W, void* dst, void* stream) {
    if (!src || !dst || B < 1 || C < 2 || C % 2 || H < 1 || W < 1 || !dtype_ok(src_dtype)) {
        set_error("bad argument to fm_fine_to_nhwc_bf16");
        return FM_ERR_BAD_ARG;
    }
    if (C * 33 * sizeof(float) > 48 * 1024) { set_error("C too large"); return FM_ERR_UNSUPPORTED; }
    cudaError_t e = launch_to_nhwc(src, src_dtype, B, C, H, W, dst, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "to_nhwc launch");
    return FM_OK;
}

size_t fm_fine_gather_workspace_bytes(int B, int N0, int N1, int Cc, int Cf, int featc_dtype) {
    if (B < 1 || N0 < 1 || N1 < 1 || Cc < 1 || !dtype_ok(featc_dtype)) return 0;
    // The size does not depend on the window: every window the tensor-core path takes (2 <= W <= 7) needs the same
    // carve, and for a window it does not take (W = 8) the workspace is merely unused (the SIMT path needs none).
    if (!fine_gather_tc_supported(Cf, Cc, 7)) return 0;
    return fine_gather_tc_workspace_bytes(B, N0, N1, Cc, featc_dtype == FM_DTYPE_F32);
}

int fm_fine_gather(const void* feat_f0, const void* feat_f1, const void* feat_c0, const void* feat_c1, int featc_dtype,
                   const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const int32_t* count, int capacity,
                   int B, int Hf, int Wf, int Cf, int N0, int N1, int Cc, int w0c, int w1c, int stride, int pad, int W,
                   const float* down_w, const float* down_b, const float* merge_w, const float* merge_b, void* win0,
                   void* win1, int out_dtype, void* workspace, size_t workspace_bytes, void* stream) {
    if (!feat_f0 || !feat_f1 || !b_ids || !i_ids || !j_ids || !win0 || !win1) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    if (Cf != 64) { set_error("Cf = %d unsupported (fine descriptors have 64 channels)", Cf); return FM_ERR_UNSUPPORTED; }
    if (W < 2 || W * W > 64) { set_error("window %d unsupported (need 2 <= W and W*W <= 64)", W); return FM_ERR_UNSUPPORTED; }
    if (capacity < 0 || B < 1 || Hf < 1 || Wf < 1 || w0c < 1 || w1c < 1 || stride < 1 || pad < 0) { set_error("bad size argument"); return FM_ERR_BAD_ARG; }
    if (!dtype_ok(out_dtype) || !dtype_ok(featc_dtype)) { set_error("bad dtype"); return FM_ERR_BAD_ARG; }
    const bool merge = down_w != nullptr;
    if (merge && (!down_b || !merge_w || !merge_b || !feat_c0 || !feat_c1)) { set_error("merge parameters must be all set or all NULL"); return FM_ERR_BAD_ARG; }
    if (merge && (Cc < 8 || Cc % 8 || Cc > 1024)) { set_error("Cc = %d unsupported", Cc); return FM_ERR_UNSUPPORTED; }
    if (capacity == 0) return FM_OK;
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e;
    if (merge && fine_gather_tc_supported(Cf, Cc, W)) {
        if (N0 < 1 || N1 < 1) { set_error("N0, N1 must be >= 1"); return FM_ERR_BAD_ARG; }
        if ((reinterpret_cast<uintptr_t>(feat_f0) | reinterpret_cast<uintptr_t>(feat_f1) | reinterpret_cast<uintptr_t>(feat_c0) |
             reinterpret_cast<uintptr_t>(feat_c1) | reinterpret_cast<uintptr_t>(win0) | reinterpret_cast<uintptr_t>(win1) |
             reinterpret_cast<uintptr_t>(workspace)) & 15) {
            set_error("feature, window and workspace pointers must be 16-byte aligned");
            return FM_ERR_BAD_ARG;
        }
        const size_t need = fine_gather_tc_workspace_bytes(B, N0, N1, Cc, featc_dtype == FM_DTYPE_F32);
        if (!workspace || workspace_bytes < need) {
            set_error("fine gather workspace too small: need %zu bytes, got %zu", need, workspace ? workspace_bytes : 0);
            return FM_ERR_WORKSPACE;
        }
        e = launch_fine_gather_tc(feat_f0, feat_f1, feat_c0, feat_c1, featc_dtype, b_ids, i_ids, j_ids, count, capacity, B,
                                  Hf, Wf, N0, N1, Cc, w0c, w1c, stride, pad, W, down_w, down_b, merge_w, merge_b, win0,
                                  win1, out_dtype, workspace, workspace_bytes, sms, static_cast<cudaStream_t>(stream));
    } else {
        e = launch_fine_gather(feat_f0, feat_f1, feat_c0, feat_c1, featc_dtype, b_ids, i_ids, j_ids, count, capacity, B,
                               Hf, Wf, N0, N1, Cc, w0c, w1c, stride, pad, W, down_w, down_b, merge_w, merge_b, win0,
                               win1, out_dtype, sms, static_cast<cudaStream_t>(stream));
    }
    if (e != cudaSuccess) return cuda_fail(e, "fine_gather launch");
    return FM_OK;
}

int fm_fine_match(const void* win0, const void* win1, int win_dtype, const float* mkpts0_c, const float* mkpts1_c,
                  const float* mix_w0, const float* mix_b0, const float* mix_w1, const float* mix_b1,
                  const int32_t* count, int capacity, int WW, int Cf, float scale, float* mkpts0_f, float* mkpts1_f,
                  void* stream) {
    if (!win0 || !win1 || !mkpts0_c || !mkpts1_c || !mix_w0 || !mix_b0 || !mix_w1 || !mix_b1 || !mkpts0_f || !mkpts1_f) {
        set_error("null pointer argument");
        return FM_ERR_BAD_ARG;
    }
    if (Cf != 64) { set_error("Cf = %d unsupported (fine descriptors have 64 channels)", Cf); return FM_ERR_UNSUPPORTED; }
    int W = 1;
    while (W * W < WW) ++W;
    if (W * W != WW || W < 2 || WW > 64) { set_error("WW = %d unsupported (need a square window, 4 <= WW <= 64)", WW); return FM_ERR_UNSUPPORTED; }
    if (!dtype_ok(win_dtype) || capacity < 0) { set_error("bad argument"); return FM_ERR_BAD_ARG; }
    if (capacity == 0) return FM_OK;
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_fine_match(win0, win1, win_dtype, mkpts0_c, mkpts1_c, mix_w0, mix_b0, mix_w1, mix_b1, count,
                                      capacity, WW, scale, mkpts0_f, mkpts1_f, sms, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "fine_match launch");
    return FM_OK;
}

size_t fm_fine_transformer_workspace_bytes(int n_layers) {
    return n_layers > 0 ? fine_transformer_workspace_bytes(n_layers) : 0;
}

int fm_fine_transformer(const float* win0, const float* win1, const int32_t* count, int capacity, int WW, int C,
                        int nhead, int n_layers, const int32_t* layer_kinds, const float* const* layer_params,
                        float* out0, float* out1, void* workspace, size_t workspace_bytes, void* stream) {
    if (!win0 || !win1 || !out0 || !out1 || !layer_kinds || !workspace) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    if (n_layers < 1 || n_layers > 16 || capacity < 0) { set_error("n_layers must be in [1, 16] and capacity >= 0"); return FM_ERR_BAD_ARG; }
    if (!fine_transformer_supported(C, nhead, WW)) {
        set_error("fine transformer supports d_model 64, nhead 8, WW <= 49 (got C=%d nhead=%d WW=%d)", C, nhead, WW);
        return FM_ERR_UNSUPPORTED;
    }
    for (int l = 0; l < n_layers; ++l) {
        if (layer_kinds[l] != 0 && layer_kinds[l] != 1) { set_error("layer_kinds[%d] must be 0 (self) or 1 (cross)", l); return FM_ERR_BAD_ARG; }
        for (int k = 0; layer_params && k < 10; ++k)   // layer_params == NULL: the workspace holds them packed already
            if (!layer_params[10 * l + k]) { set_error("layer_params[%d] is NULL", 10 * l + k); return FM_ERR_BAD_ARG; }
    }
    if ((reinterpret_cast<uintptr_t>(win0) | reinterpret_cast<uintptr_t>(win1) | reinterpret_cast<uintptr_t>(out0) |
         reinterpret_cast<uintptr_t>(out1) | reinterpret_cast<uintptr_t>(workspace)) & 31) {
        set_error("window and workspace pointers must be 32-byte aligned");
        return FM_ERR_BAD_ARG;
    }
    if (workspace_bytes < fine_transformer_workspace_bytes(n_layers)) {
        set_error("fine transformer workspace too small: need %zu bytes, got %zu", fine_transformer_workspace_bytes(n_layers), workspace_bytes);
        return FM_ERR_WORKSPACE;
    }
    if (capacity == 0) return FM_OK;
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_fine_transformer(win0, win1, count, capacity, WW, n_layers, layer_kinds, layer_params, out0, out1,
                                            workspace, sms, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "fine_transformer launch");
    return FM_OK;
}

size_t fm_coarse_transformer_workspace_bytes(int B, int L0, int L1, int n_layers) {
    if (B < 0 || L0 < 0 || L1 < 0 || n_layers < 1) return 0;
    return coarse_transformer_workspace_bytes(B, L0, L1, n_layers);
}

int fm_coarse_transformer(const float* feat0, const float* feat1, int B, int L0, int L1, int C, int nhead,
                          int n_layers, const int32_t* layer_kinds, const float* const* layer_params, float* out0,
                          float* out1, void* out0_bf16, void* out1_bf16, void* workspace, size_t workspace_bytes,
                          void* stream) {
    if (!layer_kinds || !workspace) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    if (n_layers < 1 || n_layers > 32 || B < 0 || L0 < 0 || L1 < 0) { set_error("n_layers must be in [1, 32] and B, L0, L1 >= 0"); return FM_ERR_BAD_ARG; }
    if ((static_cast<size_t>(B) * L0 && (!feat0 || !out0)) || (static_cast<size_t>(B) * L1 && (!feat1 || !out1))) {
        set_error("null token pointer");
        return FM_ERR_BAD_ARG;
    }
    if (!coarse_transformer_supported(C, nhead)) {
        set_error("coarse transformer supports d_model 256, nhead 8 (got C=%d nhead=%d)", C, nhead);
        return FM_ERR_UNSUPPORTED;
    }
    for (int l = 0; l < n_layers; ++l) {
        if (layer_kinds[l] != 0 && layer_kinds[l] != 1) { set_error("layer_kinds[%d] must be 0 (self) or 1 (cross)", l); return FM_ERR_BAD_ARG; }
        for (int k = 0; layer_params && k < 10; ++k)   // layer_params == NULL: the workspace holds them packed already
            if (!layer_params[10 * l + k]) { set_error("layer_params[%d] is NULL", 10 * l + k); return FM_ERR_BAD_ARG; }
    }
    if ((reinterpret_cast<uintptr_t>(feat0) | reinterpret_cast<uintptr_t>(feat1) | reinterpret_cast<uintptr_t>(out0) |
         reinterpret_cast<uintptr_t>(out1) | reinterpret_cast<uintptr_t>(out0_bf16) | reinterpret_cast<uintptr_t>(out1_bf16)) & 31) {
        set_error("token pointers must be 32-byte aligned");
        return FM_ERR_BAD_ARG;
    }
    if (reinterpret_cast<uintptr_t>(workspace) & 255) { set_error("workspace must be 256-byte aligned"); return FM_ERR_BAD_ARG; }
    const size_t need = coarse_transformer_workspace_bytes(B, L0, L1, n_layers);
    if (workspace_bytes < need) {
        set_error("coarse transformer workspace too small: need %zu bytes, got %zu", need, workspace_bytes);
        return FM_ERR_WORKSPACE;
    }
    if (B == 0 || (L0 == 0 && L1 == 0)) return FM_OK;
    if (L0 == 0 || L1 == 0) {   // attentions.py:40-41 divides by the source length: the reference yields NaN here
        set_error("coarse transformer: one sequence is empty (L0=%d, L1=%d)", L0, L1);
        return FM_ERR_BAD_ARG;
    }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_coarse_transformer(feat0, feat1, B, L0, L1, n_layers, layer_kinds,
                                              reinterpret_cast<const void* const*>(layer_params), out0, out1, out0_bf16,
                                              out1_bf16, workspace, sms, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "coarse_transformer launch");
    return FM_OK;
}

int fm_pack_matches(const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const float* mconf,
                    const float* mkpts0_f, const float* mkpts1_f, const int32_t* count, int capacity, int pair_offset,
                    void* records, void* stream) {
    if (!b_ids || !i_ids || !j_ids || !mconf || !mkpts0_f || !mkpts1_f || !count || !records || capacity < 0) {
        set_error("bad argument to fm_pack_matches");
        return FM_ERR_BAD_ARG;
    }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_pack_matches(b_ids, i_ids, j_ids, mconf, mkpts0_f, mkpts1_f, count, capacity, pair_offset,
                                        records, nullptr, 0, 0, sms, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "pack_matches launch");
    return FM_OK;
}

int fm_pack_matches_append(const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const float* mconf,
                           const float* mkpts0_f, const float* mkpts1_f, const int32_t* count, int capacity,
                           int pair_offset, int32_t* chunk_counts, int chunk_index, int total_capacity, void* records,
                           void* stream) {
    if (!b_ids || !i_ids || !j_ids || !mconf || !mkpts0_f || !mkpts1_f || !count || !records || !chunk_counts ||
        capacity < 0 || chunk_index < 0 || chunk_index >= FM_MAX_CHUNKS || total_capacity < 0) {
        set_error("bad argument to fm_pack_matches_append");
        return FM_ERR_BAD_ARG;
    }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_pack_matches(b_ids, i_ids, j_ids, mconf, mkpts0_f, mkpts1_f, count, capacity, pair_offset,
                                        records, chunk_counts, chunk_index, total_capacity, sms,
                                        static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "pack_matches launch");
    return FM_OK;
}

size_t fm_pose_workspace_bytes(int B, int capacity, int num_hypotheses) {
    if (B < 1 || capacity < 0 || num_hypotheses < 1 || num_hypotheses > (1 << 20)) return 0;
    return pose_workspace_bytes(B, capacity, num_hypotheses);
}

int fm_pose_from_matches(const int64_t* b_ids, const float* mkpts0_f, const float* mkpts1_f, int pts_stride,
                         const int32_t* count, int capacity, int B, const float* K0, float threshold_px,
                         int num_hypotheses, int flags, uint32_t seed, float* R, float* t, float* E, int32_t* n_inliers,
                         int32_t* pose_status, uint8_t* inlier_mask, void* workspace, size_t workspace_bytes, void* stream) {
    if (!b_ids || !mkpts0_f || !mkpts1_f || !K0 || !R || !t || !n_inliers || !pose_status || !workspace) {
        set_error("null pointer argument");
        return FM_ERR_BAD_ARG;
    }
    if (B < 1 || capacity < 0 || pts_stride < 2 || num_hypotheses < 1 || num_hypotheses > (1 << 20) || !(threshold_px > 0.f)) {
        set_error("bad argument to fm_pose_from_matches (B >= 1, capacity >= 0, pts_stride >= 2, 1 <= hypotheses <= 2^20, threshold > 0)");
        return FM_ERR_BAD_ARG;
    }
    if (reinterpret_cast<uintptr_t>(workspace) & 15) { set_error("workspace must be 16-byte aligned"); return FM_ERR_BAD_ARG; }
    const size_t need = pose_workspace_bytes(B, capacity, num_hypotheses);
    if (workspace_bytes < need) {
        set_error("pose workspace too small: need %zu bytes, got %zu", need, workspace_bytes);
        return FM_ERR_WORKSPACE;
    }
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_pose_from_matches(b_ids, mkpts0_f, mkpts1_f, pts_stride, count, capacity, B, K0, threshold_px,
                                             num_hypotheses, flags, seed, R, t, E, n_inliers, pose_status, inlier_mask,
                                             workspace, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "pose_from_matches launch");
    return FM_OK;
}

int fm_debug_pose_valid_hypotheses(const void* workspace, int B, int capacity, int num_hypotheses, int32_t* valid,
                                   void* stream) {
    if (!workspace || !valid || B < 1 || capacity < 0 || num_hypotheses < 1 || num_hypotheses > (1 << 20)) {
        set_error("bad argument to fm_debug_pose_valid_hypotheses");
        return FM_ERR_BAD_ARG;
    }
    if (sm_count() <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_pose_count_valid(workspace, B, capacity, num_hypotheses, valid, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "pose_count_valid launch");
    return FM_OK;
}

int fm_backbone_emit(const float* x, long long rows, int Cin, int Cout, const float* ln_weight, const float* ln_bias,
                     float ln_eps, const float* linear_weight, float* out_f32, void* out_bf16, void* stream) {
    if (!x || !ln_weight || !ln_bias || !linear_weight || (!out_f32 && !out_bf16)) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    if (rows < 0 || !(ln_eps > 0.f)) { set_error("bad argument to fm_backbone_emit (rows >= 0, ln_eps > 0)"); return FM_ERR_BAD_ARG; }
    if (!backbone_emit_supported(Cin, Cout)) {
        set_error("fm_backbone_emit supports the two heads of the backbone: 64 -> 64 and 128 -> 256 (got %d -> %d)", Cin, Cout);
        return FM_ERR_UNSUPPORTED;
    }
    if (rows == 0) return FM_OK;
    const int sms = sm_count();
    if (sms <= 0) { set_error("no CUDA device available (this library has no CPU path)"); return FM_ERR_CUDA; }
    cudaError_t e = launch_backbone_emit(x, rows, Cin, Cout, ln_weight, ln_bias, ln_eps, linear_weight, out_f32, out_bf16,
                                         sms, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "backbone_emit launch");
    return FM_OK;
}

int fm_profile_enable(int on) {
    std::lock_guard<std::mutex> lk(g_timers.mu);
    g_timers.on = on != 0;
    if (!on) g_timers.used = 0;
    return FM_OK;
}

int fm_profile_collect(float* ms_sum, int* launches) {
    if (!ms_sum || !launches) { set_error("null pointer argument"); return FM_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> lk(g_timers.mu);
    float total = 0.f;
    int n = 0;
    for (size_t k = 0; k + 1 < g_timers.used; k += 2) {
        cudaError_t e = cudaEventSynchronize(g_timers.pool[k + 1]);
        float ms = 0.f;
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, g_timers.pool[k], g_timers.pool[k + 1]);
        if (e != cudaSuccess) { g_timers.used = 0; return cuda_fail(e, "fm_profile_collect"); }
        total += ms;
        ++n;
    }
    g_timers.used = 0;
    *ms_sum = total;
    *launches = n;
    return FM_OK;
}

}  // extern "C"
