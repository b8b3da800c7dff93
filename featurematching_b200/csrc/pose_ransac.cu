// Relative pose from the match list, on the device.
//
// Replaces the reference's device->host copy + OpenCV calls
//   network/utils/pose_pred.py:130-151  estimate_pose:  cv2.findEssentialMat(RANSAC, 0.999, 1.0) + cv2.recoverPose
//   network/utils/pose_pred.py:170-176  get_opencv_pose (R = I, t = 0 when no pose can be estimated)
//   utils/metrics.py:79-109,124-159     the evaluation twin, per batch element (mask = m_bids == bs)
// by a batched RANSAC over all pairs of a step at once, consuming the head's outputs where they lie
// (b_ids sorted as torch.where leaves them, mkpts*_f [M, 3]).
//
// RANSAC is randomised; "the same result as OpenCV" is therefore defined on the OUTCOME: the inlier set
// (Sampson error <= (threshold / focal)^2, OpenCV's computeError) and the (R, t) recoverPose selects from it.
//   hypotheses   H per pair, one thread each: 8 matches (counter-based hash sampling) -> null vector of the 8 x 9
//                epipolar system (Gauss-Jordan, full pivoting) -> nearest essential matrix (3 x 3 SVD)
//   scoring      one warp per hypothesis over the pair's matches: MSAC cost sum min(err, (2 thr)^2)
//   polish       the 16 cheapest hypotheses of a pair, one CTA each: least-squares refit (9 x 9 normal equations in
//                fp64, Jacobi) on the matches within kappa x thr, kappa 3 -> 1, then while the MSAC cost falls
//                (a minimal 8-point sample is a poor estimator under pixel noise; OpenCV's 5-point sample is not
//                refitted at all -- this stage is why the poses land closer to the truth than OpenCV's)
//   finalize     per pair: the polished model with the lowest MSAC cost -> inlier mask -> decomposeEssentialMat ->
//                DLT triangulation of every inlier under the four (R, +-t) candidates -> recoverPose's ranking.
// The arithmetic lives in fm_pose_math.cuh, which also compiles for the host (tests/pose_hostcheck.cpp checks it
// against OpenCV on the CPU); this file is the parallel orchestration.
#include <algorithm>

#include "fm_common.cuh"
#include "fm_pose_math.cuh"

namespace fm {

namespace {

using namespace pose;

constexpr int kStarts = 16;        // hypotheses polished per pair
constexpr int kPolishIters = 8;
constexpr int kPoseThreads = 256;

struct PoseWs {
    int* range;        // [B + 1] first match of every pair (b_ids is sorted)
    float4* pts;       // [capacity] normalised (x0, y0, x1, y1)
    float* hyp;        // [B, H, 9]
    float* cost;       // [B, H]
    int* start;        // [B, kStarts] hypothesis index, -1 = none
    double* polE;      // [B, kStarts, 9]
    float* polC;       // [B, kStarts] MSAC cost of the polished model, INF = none
    size_t total;
};

PoseWs carve(int B, int capacity, int H, void* base) {
    PoseWs w{};
    uint8_t* p = static_cast<uint8_t*>(base);
    size_t off = 0;
    auto take = [&](size_t bytes) { uint8_t* r = p ? p + off : nullptr; off = (off + bytes + 255) / 256 * 256; return r; };
    w.range = reinterpret_cast<int*>(take(sizeof(int) * (B + 1)));
    w.pts = reinterpret_cast<float4*>(take(sizeof(float4) * (capacity > 0 ? capacity : 1)));
    w.hyp = reinterpret_cast<float*>(take(sizeof(float) * 9 * B * H));
    w.cost = reinterpret_cast<float*>(take(sizeof(float) * B * H));
    w.start = reinterpret_cast<int*>(take(sizeof(int) * B * kStarts));
    w.polE = reinterpret_cast<double*>(take(sizeof(double) * 9 * B * kStarts));
    w.polC = reinterpret_cast<float*>(take(sizeof(float) * B * kStarts));
    w.total = off;
    return w;
}

struct PoseParams {
    const int64_t* b_ids;
    const float* mk0;
    const float* mk1;
    int stride;            // floats per point row (3 for the head's [M, 3], 2 for plain coordinates)
    const int* count;
    int capacity, B, H;
    const float* K;        // [B, 9]
    float thr_px;
    int flags;
    uint32_t seed;
    float* R; float* t; float* E; int* n_inliers; int* status; uint8_t* mask;
};

__device__ __forceinline__ int match_count(const PoseParams& p) { return p.count ? min(*p.count, p.capacity) : p.capacity; }

// ---- 1. per-pair ranges (lower_bound on the sorted b_ids) and normalised points
__global__ void __launch_bounds__(kPoseThreads) pose_prepare_kernel(PoseParams p, PoseWs w) {
    const int M = match_count(p);
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid <= p.B) {
        int lo = 0, hi = M;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (p.b_ids[mid] < tid) lo = mid + 1; else hi = mid;
        }
        w.range[tid] = lo;
    }
    for (int m = tid; m < M; m += gridDim.x * blockDim.x) {
        const int b = static_cast<int>(p.b_ids[m]);
        const float* K = p.K + 9 * min(max(b, 0), p.B - 1);
        const float inv_f = 1.0f / (0.5f * (K[0] + K[4]));   // focal = (fx + fy) / 2 for both axes, pose_pred.py:134
        const float cx = K[2], cy = K[5];
        w.pts[m] = make_float4((p.mk0[m * p.stride] - cx) * inv_f, (p.mk0[m * p.stride + 1] - cy) * inv_f,
                               (p.mk1[m * p.stride] - cx) * inv_f, (p.mk1[m * p.stride + 1] - cy) * inv_f);
    }
}

// ---- 2. one thread = one hypothesis
// The three steps are kept as separate device functions: inlined into one body (sample8 -> null_vector_8x9 ->
// early return -> project_essential) nvcc 12.9 -O3 for sm_100a produced a kernel whose pivot test rejected EVERY
// sample on the device (tools/probe_pose_null.cu: 0 / 1024 models where the host build and every partial
// inlining give 1024 / 1024).
__device__ __noinline__ void sample8_dev(uint32_t seed, int b, int h, int n, int* idx) { sample8(seed, b, h, n, idx); }
__device__ __noinline__ bool null_vector_dev(float* A, float* e) { return null_vector_8x9(A, e); }
__device__ __noinline__ void project_essential_dev(float* e) { project_essential(e); }

__global__ void __launch_bounds__(128) pose_hypotheses_kernel(PoseParams p, PoseWs w) {
    const int b = blockIdx.y, h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= p.H) return;
    const int lo = w.range[b], n = w.range[b + 1] - lo;
    float* E = w.hyp + (static_cast<size_t>(b) * p.H + h) * 9;
    float* cost = w.cost + static_cast<size_t>(b) * p.H + h;
    *cost = INFINITY;                     // scored (overwritten) only if a model comes out
    if (n < 8) return;
    int idx[8];
    sample8_dev(p.seed, b, h, n, idx);
    float A[72], e[9];
    for (int k = 0; k < 8; ++k) {
        const float4 q = w.pts[lo + idx[k]];
        epipolar_row(q.x, q.y, q.z, q.w, A + 9 * k);
    }
    if (!null_vector_dev(A, e)) { E[0] = NAN; return; }
    project_essential_dev(e);
    for (int i = 0; i < 9; ++i) E[i] = e[i];
    *cost = 0.f;                          // marks "valid" for the scoring kernel
}

// ---- 3. one warp = one hypothesis over the pair's matches
__global__ void __launch_bounds__(kPoseThreads) pose_score_kernel(PoseParams p, PoseWs w) {
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x * (kPoseThreads / 32) + warp;
    if (h >= p.H) return;
    float* cost = w.cost + static_cast<size_t>(b) * p.H + h;
    if (!(*cost == 0.f)) return;          // no model (INF)
    const int lo = w.range[b], n = w.range[b + 1] - lo;
    float E[9];
    for (int i = 0; i < 9; ++i) E[i] = w.hyp[(static_cast<size_t>(b) * p.H + h) * 9 + i];
    const float K0 = p.K[9 * b] + p.K[9 * b + 4];
    const float thr = p.thr_px / (0.5f * K0);
    const float trunc2 = 4.f * thr * thr;
    float c = 0.f;
    for (int m = lane; m < n; m += 32) {
        const float4 q = w.pts[lo + m];
        const float e = sampson_error(E, q.x, q.y, q.z, q.w);
        c += e < trunc2 ? e : trunc2;     // NaN (degenerate denominators) counts as an outlier
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) *cost = c;
}

// ---- 4. the kStarts cheapest hypotheses of every pair (ties: lowest index)
__global__ void __launch_bounds__(kPoseThreads) pose_starts_kernel(PoseParams p, PoseWs w) {
    __shared__ float s_val[kPoseThreads / 32];
    __shared__ int s_idx[kPoseThreads / 32];
    __shared__ int s_pick;
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* cost = w.cost + static_cast<size_t>(b) * p.H;
    for (int s = 0; s < kStarts; ++s) {
        float best = INFINITY;
        int bi = 0x7fffffff;
        for (int h = threadIdx.x; h < p.H; h += kPoseThreads) {
            const float c = cost[h];
            if (c < best) { best = c; bi = h; }     // ascending h per thread: the first minimum stays
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = best; s_idx[warp] = bi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int k = 1; k < kPoseThreads / 32; ++k)
                if (s_val[k] < s_val[0] || (s_val[k] == s_val[0] && s_idx[k] < s_idx[0])) { s_val[0] = s_val[k]; s_idx[0] = s_idx[k]; }
            const int pick = s_val[0] < INFINITY ? s_idx[0] : -1;
            w.start[b * kStarts + s] = pick;
            if (pick >= 0) cost[pick] = INFINITY;   // taken
            s_pick = pick;
        }
        __syncthreads();
        (void)s_pick;
    }
}

// block-wide sums; results valid in every thread
__device__ __forceinline__ float block_sum_float(float v, float* scratch) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = 0.f;
    for (int k = 0; k < kPoseThreads / 32; ++k) s += scratch[k];
    return s;
}
__device__ __forceinline__ int block_sum_int(int v, int* scratch) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    int s = 0;
    for (int k = 0; k < kPoseThreads / 32; ++k) s += scratch[k];
    return s;
}

// Cyclic Jacobi of a symmetric 9 x 9 matrix by ONE WARP (a, v in shared memory, doubles): the same rotations in the
// same order as pose::jacobi_eigen, with the nine elements of each row / column update on nine lanes.  (Run by a
// single thread the eight refits of a polish took 22 ms per 32 pairs: ~100 k dependent fp64 operations each.)
__device__ void jacobi9_warp(double* a, double* v, int lane) {
    if (lane < 9)
        for (int j = 0; j < 9; ++j) v[lane * 9 + j] = lane == j ? 1.0 : 0.0;
    __syncwarp();
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0, dg = 0.0;
        if (lane < 9)
            for (int q = 0; q < 9; ++q) {
                const double x = a[lane * 9 + q];
                if (q == lane) dg = x * x; else off += x * x;
            }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            off += __shfl_xor_sync(0xffffffffu, off, o);
            dg += __shfl_xor_sync(0xffffffffu, dg, o);
        }
        if (off * 0.5 <= dg * 1e-30) break;   // warp-uniform
        for (int p = 0; p < 9; ++p)
            for (int q = p + 1; q < 9; ++q) {
                const double apq = a[p * 9 + q], app = a[p * 9 + p], aqq = a[q * 9 + q];
                __syncwarp();                  // everyone holds the pivot entries before anyone rewrites them
                if (apq == 0.0) continue;      // warp-uniform
                const double theta = (aqq - app) / (2.0 * apq);
                const double tt = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(tt * tt + 1.0), sn = tt * c;
                if (lane < 9) {                // A <- A J (columns p, q): lane = row
                    const double akp = a[lane * 9 + p], akq = a[lane * 9 + q];
                    a[lane * 9 + p] = c * akp - sn * akq;
                    a[lane * 9 + q] = sn * akp + c * akq;
                }
                __syncwarp();
                if (lane < 9) {                // A <- J' A (rows p, q): lane = column; V <- V J
                    const double apk = a[p * 9 + lane], aqk = a[q * 9 + lane];
                    a[p * 9 + lane] = c * apk - sn * aqk;
                    a[q * 9 + lane] = sn * apk + c * aqk;
                    const double vkp = v[lane * 9 + p], vkq = v[lane * 9 + q];
                    v[lane * 9 + p] = c * vkp - sn * vkq;
                    v[lane * 9 + q] = sn * vkp + c * vkq;
                }
                __syncwarp();
            }
    }
}

// ---- 5. one CTA = one start of one pair: local optimisation
__global__ void __launch_bounds__(kPoseThreads) pose_polish_kernel(PoseParams p, PoseWs w) {
    __shared__ double s_ata[kPoseThreads / 32][45];
    __shared__ double s_A[81], s_V[81];
    __shared__ float s_E[9];
    __shared__ double s_Ed[9];
    __shared__ float s_red[kPoseThreads / 32];
    const int s = blockIdx.x, b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h = w.start[b * kStarts + s];
    double* outE = w.polE + (static_cast<size_t>(b) * kStarts + s) * 9;
    float* outC = w.polC + b * kStarts + s;
    if (h < 0) { if (threadIdx.x == 0) *outC = INFINITY; return; }
    const int lo = w.range[b], n = w.range[b + 1] - lo;
    const float thr = p.thr_px / (0.5f * (p.K[9 * b] + p.K[9 * b + 4]));
    const float thr2 = thr * thr;
    if (threadIdx.x < 9) {
        const float e = w.hyp[(static_cast<size_t>(b) * p.H + h) * 9 + threadIdx.x];
        s_E[threadIdx.x] = e;
        s_Ed[threadIdx.x] = e;
    }
    __syncthreads();
    // consensus of a model: MSAC cost sum min(err, thr^2), lower is better (a pure inlier count lets a refit trade
    // a slightly worse fit of everything for one more borderline match)
    auto consensus = [&](const float* E) {
        float c = 0.f;
        for (int m = threadIdx.x; m < n; m += kPoseThreads) {
            const float4 q = w.pts[lo + m];
            const float e = sampson_error(E, q.x, q.y, q.z, q.w);
            c += e <= thr2 ? e : thr2;
        }
        return block_sum_float(c, s_red);
    };
    float Ecur[9];
    for (int i = 0; i < 9; ++i) Ecur[i] = s_E[i];
    float best_c = consensus(Ecur);
    double bestE[9];
    for (int i = 0; i < 9; ++i) bestE[i] = s_Ed[i];
    if (p.flags & 1) {
        const float kappa[kPolishIters] = {3.f, 2.5f, 2.f, 1.5f, 1.f, 1.f, 1.f, 1.f};
        for (int it = 0; it < kPolishIters; ++it) {
            const float t2 = thr2 * kappa[it] * kappa[it];
            double acc[45];
#pragma unroll
            for (int i = 0; i < 45; ++i) acc[i] = 0.0;
            for (int m = threadIdx.x; m < n; m += kPoseThreads) {
                const float4 q = w.pts[lo + m];
                if (!(sampson_error(Ecur, q.x, q.y, q.z, q.w) <= t2)) continue;
                double row[9];
                epipolar_row<double>(q.x, q.y, q.z, q.w, row);
                int k = 0;
#pragma unroll
                for (int i = 0; i < 9; ++i)
#pragma unroll
                    for (int j = i; j < 9; ++j) acc[k++] += row[i] * row[j];
            }
#pragma unroll
            for (int i = 0; i < 45; ++i) {
                double v = acc[i];
                for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) s_ata[warp][i] = v;
            }
            __syncthreads();
            if (warp == 0) {
                for (int e = lane; e < 81; e += 32) {          // the full symmetric matrix from the warps' packed partials
                    const int r = e / 9, c = e - r * 9, i = min(r, c), j = max(r, c);
                    const int k = i * 9 - (i * (i - 1)) / 2 + (j - i);
                    double v = 0.0;
                    for (int ww = 0; ww < kPoseThreads / 32; ++ww) v += s_ata[ww][k];
                    s_A[e] = v;
                }
                __syncwarp();
                jacobi9_warp(s_A, s_V, lane);
                if (lane == 0) {
                    int kmin = 0;
                    for (int i = 1; i < 9; ++i) if (s_A[i * 10] < s_A[kmin * 10]) kmin = i;
                    double Er[9];
                    for (int i = 0; i < 9; ++i) Er[i] = s_V[i * 9 + kmin];
                    project_essential(Er);
                    for (int i = 0; i < 9; ++i) { s_Ed[i] = Er[i]; s_E[i] = static_cast<float>(Er[i]); }
                }
            }
            __syncthreads();
            float Ef[9];
            for (int i = 0; i < 9; ++i) Ef[i] = s_E[i];
            const float c2 = consensus(Ef);
            if (kappa[it] == 1.f && !(c2 < best_c)) break;     // uniform over the CTA
            for (int i = 0; i < 9; ++i) Ecur[i] = Ef[i];
            if (c2 < best_c) {
                best_c = c2;
                for (int i = 0; i < 9; ++i) bestE[i] = s_Ed[i];
            }
            __syncthreads();   // s_E / s_Ed are rewritten by thread 0 in the next round
        }
    }
    if (threadIdx.x == 0) {
        *outC = best_c;
        for (int i = 0; i < 9; ++i) outE[i] = bestE[i];
    }
}

// ---- 6. one CTA = one pair: best polished model, inlier mask, recoverPose
__global__ void __launch_bounds__(kPoseThreads) pose_finalize_kernel(PoseParams p, PoseWs w) {
    __shared__ double s_R[2][9];
    __shared__ double s_t[3];
    __shared__ float s_E[9];
    __shared__ int s_cnt[kPoseThreads / 32];
    __shared__ int s_best;
    const int b = blockIdx.x;
    const int lo = w.range[b], n = w.range[b + 1] - lo;
    float* R = p.R + 9 * b;
    float* t = p.t + 3 * b;
    if (threadIdx.x == 0) {
        int best = -1;
        float bc = INFINITY;
        for (int s = 0; s < kStarts; ++s) {
            const float c = w.polC[b * kStarts + s];
            if (c < bc) { bc = c; best = s; }
        }
        s_best = best;
        if (best >= 0) {
            const double* E = w.polE + (static_cast<size_t>(b) * kStarts + best) * 9;
            for (int i = 0; i < 9; ++i) s_E[i] = static_cast<float>(E[i]);
            decompose_essential(E, s_R[0], s_R[1], s_t);
        }
    }
    __syncthreads();
    if (s_best < 0) {
        // fewer than 8 matches, or no sample produced a model: the reference's `ret is None` branch
        // (pose_pred.py:177-179): R = I, t = 0
        if (threadIdx.x < 9) { R[threadIdx.x] = (threadIdx.x % 4 == 0) ? 1.f : 0.f; if (p.E) p.E[9 * b + threadIdx.x] = 0.f; }
        if (threadIdx.x < 3) t[threadIdx.x] = 0.f;
        if (threadIdx.x == 0) { p.n_inliers[b] = 0; p.status[b] = n < 8 ? 1 : 2; }
        if (p.mask) for (int m = threadIdx.x; m < n; m += kPoseThreads) p.mask[lo + m] = 0;
        return;
    }
    const float thr = p.thr_px / (0.5f * (p.K[9 * b] + p.K[9 * b + 4]));
    const float thr2 = thr * thr;
    float E[9];
    for (int i = 0; i < 9; ++i) E[i] = s_E[i];
    double tn[3] = {-s_t[0], -s_t[1], -s_t[2]}, tp[3] = {s_t[0], s_t[1], s_t[2]};
    const bool identity_k = (p.flags & 2) != 0;   // the reference hands recoverPose PIXEL coordinates and np.eye(3)
    int good[4] = {0, 0, 0, 0}, ninl = 0;
    for (int m = threadIdx.x; m < n; m += kPoseThreads) {
        const float4 q = w.pts[lo + m];
        const bool inl = sampson_error(E, q.x, q.y, q.z, q.w) <= thr2;
        if (p.mask) p.mask[lo + m] = inl ? 1 : 0;
        if (!inl) continue;
        ++ninl;
        double a0 = q.x, b0 = q.y, a1 = q.z, b1 = q.w;
        if (identity_k) {
            const size_t g = static_cast<size_t>(lo + m) * p.stride;
            a0 = p.mk0[g]; b0 = p.mk0[g + 1]; a1 = p.mk1[g]; b1 = p.mk1[g + 1];
        }
        good[0] += in_front_of_both(s_R[0], tp, a0, b0, a1, b1, 1e9) ? 1 : 0;
        good[1] += in_front_of_both(s_R[1], tp, a0, b0, a1, b1, 1e9) ? 1 : 0;
        good[2] += in_front_of_both(s_R[0], tn, a0, b0, a1, b1, 1e9) ? 1 : 0;
        good[3] += in_front_of_both(s_R[1], tn, a0, b0, a1, b1, 1e9) ? 1 : 0;
    }
    for (int k = 0; k < 4; ++k) good[k] = block_sum_int(good[k], s_cnt);
    ninl = block_sum_int(ninl, s_cnt);
    if (threadIdx.x == 0) {
        const int c = best_candidate(good);
        const double* Rs = s_R[c & 1];
        // no inlier in front of both cameras under any of the four candidates: recoverPose returns n = 0, the
        // reference's `n > best_num_inliers` never holds and it hands back None (pose_pred.py:143-151) -> R = I, t = 0;
        // E, the inlier count and the mask stay (they describe the model that was found)
        const bool none_in_front = good[c] == 0;
        for (int i = 0; i < 9; ++i) {
            R[i] = none_in_front ? ((i % 4 == 0) ? 1.f : 0.f) : static_cast<float>(Rs[i]);
            if (p.E) p.E[9 * b + i] = E[i];
        }
        for (int i = 0; i < 3; ++i) t[i] = none_in_front ? 0.f : static_cast<float>((c & 2) ? tn[i] : tp[i]);
        p.n_inliers[b] = ninl;
        p.status[b] = none_in_front ? 3 : 0;
    }
}

// ---- debug: how many of a pair's H samples produced a model (pose_hypotheses_kernel marks the others with E[0] = NaN).
// Reads the workspace a finished fm_pose_from_matches call left behind; pairs with fewer than 8 matches report 0.
__global__ void __launch_bounds__(kPoseThreads) pose_count_valid_kernel(PoseWs w, int H, int* valid) {
    __shared__ int s_cnt[kPoseThreads / 32];
    const int b = blockIdx.x;
    int c = 0;
    if (w.range[b + 1] - w.range[b] >= 8)
        for (int h = threadIdx.x; h < H; h += kPoseThreads) {
            const float e0 = w.hyp[(static_cast<size_t>(b) * H + h) * 9];
            c += (e0 == e0) ? 1 : 0;
        }
    c = block_sum_int(c, s_cnt);
    if (threadIdx.x == 0) valid[b] = c;
}

}  // namespace

size_t pose_workspace_bytes(int B, int capacity, int H) { return carve(B, capacity, H, nullptr).total; }

cudaError_t launch_pose_count_valid(const void* workspace, int B, int capacity, int H, int* valid, cudaStream_t st) {
    const PoseWs w = carve(B, capacity, H, const_cast<void*>(workspace));
    pose_count_valid_kernel<<<B, kPoseThreads, 0, st>>>(w, H, valid); count_launch();
    return cudaGetLastError();
}

cudaError_t launch_pose_from_matches(const int64_t* b_ids, const float* mk0, const float* mk1, int stride, const int* count,
                                     int capacity, int B, const float* K, float thr_px, int H, int flags, uint32_t seed,
                                     float* R, float* t, float* E, int* n_inliers, int* status, uint8_t* mask,
                                     void* workspace, cudaStream_t st) {
    const PoseWs w = carve(B, capacity, H, workspace);
    PoseParams p{b_ids, mk0, mk1, stride, count, capacity, B, H, K, thr_px, flags, seed, R, t, E, n_inliers, status, mask};
    const int prep_blocks = std::max(1, std::min(1024, (std::max(capacity, B + 1) + kPoseThreads - 1) / kPoseThreads));
    pose_prepare_kernel<<<prep_blocks, kPoseThreads, 0, st>>>(p, w); count_launch();
    pose_hypotheses_kernel<<<dim3((H + 127) / 128, B), 128, 0, st>>>(p, w); count_launch();
    pose_score_kernel<<<dim3((H + kPoseThreads / 32 - 1) / (kPoseThreads / 32), B), kPoseThreads, 0, st>>>(p, w); count_launch();
    pose_starts_kernel<<<B, kPoseThreads, 0, st>>>(p, w); count_launch();
    pose_polish_kernel<<<dim3(kStarts, B), kPoseThreads, 0, st>>>(p, w); count_launch();
    pose_finalize_kernel<<<B, kPoseThreads, 0, st>>>(p, w); count_launch();
    return cudaGetLastError();
}

}  // namespace fm
