// K2a -- FinePreprocess: fine windows at the matched coarse cells (+ coarse-context merge).
//
//   F.unfold(feat_f, (W, W), stride, padding=pad) selected at (b, i) / (b, j)    fine_preprocess.py:43-50
//     == rows  stride * r - pad + dy, cols  stride * c - pad + dx  of the fine map, zero outside,
//        window cell ww = dy * W + dx, cell (r, c) = divmod(token, w_coarse)
//   ctx  = down_proj(coarse token)                                               fine_preprocess.py:54-55
//   out  = merge_feat(cat[window, repeat(ctx)])                                  fine_preprocess.py:56-60
//        = Wm[:, :Cf] . window[ww] + (Wm[:, Cf:] . ctx + bm)      (second term constant per match)
//
// The 49x unfolded map is never built: one CTA visit = one (match, image) pair, its W*W pixels are
// read straight from the channels-last bf16 map (one 128-byte line per pixel, 8 lanes x 16 B).
// Also here: the NCHW -> channels-last bf16 transpose for callers that hold the reference's layout.
#include "fm_common.cuh"

namespace fm {

namespace {

constexpr int kCf = 64;
constexpr int kThreads = 128;
constexpr int kXStride = kCf + 4;   // fp32 words per staged pixel (16-byte aligned, conflict-light)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <typename TC>
__device__ __forceinline__ void load8(const TC* p, float (&v)[8]);
template <>
__device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(p));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float2 f = __bfloat1622float2(h[k]);
        v[2 * k] = f.x; v[2 * k + 1] = f.y;
    }
}

__device__ __forceinline__ void store2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
__device__ __forceinline__ void store2(__nv_bfloat16* p, float a, float b) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}

struct GatherParams {
    const __nv_bfloat16* ff[2];   // channels-last fine maps [B, Hf, Wf, 64]
    const void* fc[2];            // coarse tokens [B, N, Cc]
    const int64_t* b_ids;
    const int64_t* ids[2];        // i_ids, j_ids
    const int* count;
    int capacity, B, Hf, Wf, N[2], Cc, wc[2], stride, pad, W;
    const float* down_w; const float* down_b; const float* merge_w; const float* merge_b;
    void* out[2];
};

template <typename TC, typename TO, bool kMerge>
__global__ void __launch_bounds__(kThreads) fine_gather_kernel(const GatherParams p) {
    extern __shared__ __align__(16) float sm[];
    float* WaT = sm;                          // [64 k][64 o]  merge_w[:, :64] transposed
    float* WbT = WaT + kCf * kCf;             // [64 k][64 o]  merge_w[:, 64:] transposed
    float* X = WbT + kCf * kCf;               // [WW][kXStride]
    const int WW = p.W * p.W;
    float* tok = X + WW * kXStride;           // [Cc]
    float* ctx = tok + p.Cc;                  // [64]
    float* cvec = ctx + kCf;                  // [64]
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

    if (kMerge) {
        for (int e = t; e < kCf * kCf; e += kThreads) {
            const int o = e >> 6, k = e & 63;
            WaT[k * kCf + o] = p.merge_w[o * 2 * kCf + k];
            WbT[k * kCf + o] = p.merge_w[o * 2 * kCf + kCf + k];
        }
    }
    const int M = p.count ? min(*p.count, p.capacity) : p.capacity;
    __syncthreads();

    for (int unit = blockIdx.x; unit < 2 * M; unit += gridDim.x) {
        const int m = unit >> 1, side = unit & 1;
        const int b = static_cast<int>(p.b_ids[m]);
        const int cell = static_cast<int>(p.ids[side][m]);
        const int cr = cell / p.wc[side], cc = cell - cr * p.wc[side];
        const int y0 = cr * p.stride - p.pad, x0 = cc * p.stride - p.pad;
        const __nv_bfloat16* fmap = p.ff[side] + static_cast<size_t>(b) * p.Hf * p.Wf * kCf;

        // ---- 1. stage the window (fp32) and the coarse token
        for (int e = t; e < WW * 8; e += kThreads) {
            const int px = e >> 3, ch = (e & 7) * 8;
            const int dy = px / p.W, dx = px - dy * p.W;
            const int y = y0 + dy, x = x0 + dx;
            float v[8];
            if (y >= 0 && y < p.Hf && x >= 0 && x < p.Wf) {
                load8<__nv_bfloat16>(fmap + (static_cast<size_t>(y) * p.Wf + x) * kCf + ch, v);
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = 0.f;
            }
            float4* dst = reinterpret_cast<float4*>(X + px * kXStride + ch);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
        if (kMerge) {
            const TC* tsrc = static_cast<const TC*>(p.fc[side]) + (static_cast<size_t>(b) * p.N[side] + cell) * p.Cc;
            for (int e = t * 8; e < p.Cc; e += kThreads * 8) {
                float v[8];
                load8<TC>(tsrc + e, v);
#pragma unroll
                for (int k = 0; k < 8; ++k) tok[e + k] = v[k];
            }
        }
        __syncthreads();

        if (kMerge) {
            // ---- 2. ctx = down_w . tok + down_b   (warp w: outputs 16 w .. 16 w + 15)
            for (int oo = 0; oo < 16; ++oo) {
                const int o = warp * 16 + oo;
                float acc = 0.f;
                for (int k = lane * 4; k < p.Cc; k += 128) {
                    const float4 w = __ldg(reinterpret_cast<const float4*>(p.down_w + static_cast<size_t>(o) * p.Cc + k));
                    const float4 x = *reinterpret_cast<const float4*>(tok + k);
                    acc = fmaf(w.x, x.x, acc); acc = fmaf(w.y, x.y, acc);
                    acc = fmaf(w.z, x.z, acc); acc = fmaf(w.w, x.w, acc);
                }
                acc = warp_sum(acc);
                if (lane == 0) ctx[o] = acc + p.down_b[o];
            }
            __syncthreads();
            // ---- 3. cvec = merge_w[:, 64:] . ctx + merge_b
            if (t < kCf) {
                float acc = 0.f;
#pragma unroll 8
                for (int k = 0; k < kCf; ++k) acc = fmaf(WbT[k * kCf + t], ctx[k], acc);
                cvec[t] = acc + p.merge_b[t];
            }
            __syncthreads();
        }

        // ---- 4. out[ww, o] = sum_k Wa[o, k] x[ww, k] + cvec[o]; thread: 2 channels x rows rg, rg+4, ..
        TO* out = static_cast<TO*>(p.out[side]) + static_cast<size_t>(m) * WW * kCf;
        const int rg = warp, o2 = lane * 2;
        if (kMerge) {
            constexpr int kRows = 16;   // rows per thread, covers W*W <= 64
            float acc[kRows][2];
#pragma unroll
            for (int n = 0; n < kRows; ++n) { acc[n][0] = 0.f; acc[n][1] = 0.f; }
            for (int k4 = 0; k4 < kCf; k4 += 4) {
                float2 w[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) w[q] = *reinterpret_cast<const float2*>(WaT + (k4 + q) * kCf + o2);
#pragma unroll
                for (int n = 0; n < kRows; ++n) {
                    const int row = rg + 4 * n;
                    if (row < WW) {
                        const float4 x = *reinterpret_cast<const float4*>(X + row * kXStride + k4);
                        acc[n][0] = fmaf(w[0].x, x.x, acc[n][0]); acc[n][1] = fmaf(w[0].y, x.x, acc[n][1]);
                        acc[n][0] = fmaf(w[1].x, x.y, acc[n][0]); acc[n][1] = fmaf(w[1].y, x.y, acc[n][1]);
                        acc[n][0] = fmaf(w[2].x, x.z, acc[n][0]); acc[n][1] = fmaf(w[2].y, x.z, acc[n][1]);
                        acc[n][0] = fmaf(w[3].x, x.w, acc[n][0]); acc[n][1] = fmaf(w[3].y, x.w, acc[n][1]);
                    }
                }
            }
            const float c0 = cvec[o2], c1 = cvec[o2 + 1];
#pragma unroll
            for (int n = 0; n < kRows; ++n) {
                const int row = rg + 4 * n;
                if (row < WW) store2(out + row * kCf + o2, acc[n][0] + c0, acc[n][1] + c1);
            }
        } else {
            for (int row = rg; row < WW; row += 4) {
                const float2 x = *reinterpret_cast<const float2*>(X + row * kXStride + o2);
                store2(out + row * kCf + o2, x.x, x.y);
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- NCHW -> channels-last bf16
template <typename TS>
__global__ void __launch_bounds__(256) to_nhwc_kernel(const TS* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                      int C, int H, int W) {
    // block: one (b, h) row, 32 pixels; smem tile [C][33]
    extern __shared__ float tile[];
    const int b = blockIdx.z, h = blockIdx.y, w0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int c = ty; c < C; c += 8) {
        const int w = w0 + tx;
        float v = 0.f;
        if (w < W) {
            const TS s = src[((static_cast<size_t>(b) * C + c) * H + h) * W + w];
            v = static_cast<float>(s);
        }
        tile[c * 33 + tx] = v;
    }
    __syncthreads();
    const int pairs = C / 2;
    for (int e = threadIdx.x; e < 32 * pairs; e += 256) {
        const int px = e / pairs, cp = e - px * pairs;
        const int w = w0 + px;
        if (w < W) {
            __nv_bfloat162 v = __floats2bfloat162_rn(tile[(2 * cp) * 33 + px], tile[(2 * cp + 1) * 33 + px]);
            *reinterpret_cast<__nv_bfloat162*>(dst + ((static_cast<size_t>(b) * H + h) * W + w) * C + 2 * cp) = v;
        }
    }
}

template <typename TC, typename TO>
cudaError_t launch_gather_t(const GatherParams& p, bool merge, int num_sms, cudaStream_t st) {
    const int WW = p.W * p.W;
    const size_t smem = sizeof(float) * (2 * kCf * kCf + WW * kXStride + p.Cc + 2 * kCf);
    auto kern = merge ? fine_gather_kernel<TC, TO, true> : fine_gather_kernel<TC, TO, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = static_cast<int>((227 * 1024) / (smem + 1024));
    if (per_sm > 8) per_sm = 8;
    int grid = num_sms * (per_sm > 0 ? per_sm : 1);
    if (grid > 2 * p.capacity) grid = 2 * p.capacity;
    if (grid < 1) grid = 1;
    kern<<<grid, kThreads, smem, st>>>(p); count_launch();
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_fine_gather(const void* ff0, const void* ff1, const void* fc0, const void* fc1, int featc_dtype,
                               const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids, const int* count,
                               int capacity, int B, int Hf, int Wf, int N0, int N1, int Cc, int w0c, int w1c,
                               int stride, int pad, int W, const float* down_w, const float* down_b,
                               const float* merge_w, const float* merge_b, void* win0, void* win1, int out_dtype,
                               int num_sms, cudaStream_t st) {
    GatherParams p;
    p.ff[0] = static_cast<const __nv_bfloat16*>(ff0);
    p.ff[1] = static_cast<const __nv_bfloat16*>(ff1);
    p.fc[0] = fc0; p.fc[1] = fc1;
    p.b_ids = b_ids; p.ids[0] = i_ids; p.ids[1] = j_ids;
    p.count = count; p.capacity = capacity; p.B = B; p.Hf = Hf; p.Wf = Wf;
    p.N[0] = N0; p.N[1] = N1; p.Cc = Cc; p.wc[0] = w0c; p.wc[1] = w1c;
    p.stride = stride; p.pad = pad; p.W = W;
    p.down_w = down_w; p.down_b = down_b; p.merge_w = merge_w; p.merge_b = merge_b;
    p.out[0] = win0; p.out[1] = win1;
    const bool merge = down_w != nullptr;
    const bool c32 = featc_dtype == FM_DTYPE_F32, o32 = out_dtype == FM_DTYPE_F32;
    if (c32 && o32) return launch_gather_t<float, float>(p, merge, num_sms, st);
    if (c32) return launch_gather_t<float, __nv_bfloat16>(p, merge, num_sms, st);
    if (o32) return launch_gather_t<__nv_bfloat16, float>(p, merge, num_sms, st);
    return launch_gather_t<__nv_bfloat16, __nv_bfloat16>(p, merge, num_sms, st);
}

cudaError_t launch_to_nhwc(const void* src, int src_dtype, int B, int C, int H, int W, void* dst, cudaStream_t st) {
    dim3 grid((W + 31) / 32, H, B);
    const size_t smem = sizeof(float) * C * 33;
    if (src_dtype == FM_DTYPE_F32)
        to_nhwc_kernel<float><<<grid, 256, smem, st>>>(static_cast<const float*>(src), static_cast<__nv_bfloat16*>(dst), C, H, W);
    else
        to_nhwc_kernel<__nv_bfloat16><<<grid, 256, smem, st>>>(static_cast<const __nv_bfloat16*>(src),
                                                               static_cast<__nv_bfloat16*>(dst), C, H, W);
    count_launch();
    return cudaGetLastError();
}

}  // namespace fm
