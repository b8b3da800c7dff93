// Debug probe: the hypothesis step of pose_ransac.cu (sample8 -> epipolar rows -> null_vector_8x9 ->
// project_essential) on the device vs the host, on noise-free synthetic two-view geometry.
#include <cstdio>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "../featurematching_b200/csrc/fm_pose_math.cuh"
using namespace fm::pose;
__global__ void k(const float4* pts, int n, int H, float* hyp, int* ok, int* idx_out) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    int idx[8];
    sample8(3u, 0, h, n, idx);
    if (h == 0) for (int i = 0; i < 8; ++i) idx_out[i] = idx[i];
    float A[72], e[9];
    for (int i = 0; i < 8; ++i) { const float4 q = pts[idx[i]]; epipolar_row(q.x, q.y, q.z, q.w, A + 9 * i); }
    if (!null_vector_8x9(A, e)) { ok[h] = 0; return; }
    project_essential(e);
    for (int i = 0; i < 9; ++i) hyp[9 * h + i] = e[i];
    ok[h] = 1;
}
__device__ __noinline__ void sample8_ni(uint32_t seed, int b, int h, int n, int* idx) { sample8(seed, b, h, n, idx); }
__global__ void k4(const float4* pts, int n, int H, int* ok) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    int idx[8];
    sample8_ni(3u, 0, h, n, idx);
    float A[72], e[9];
    for (int i = 0; i < 8; ++i) { const float4 q = pts[idx[i]]; epipolar_row(q.x, q.y, q.z, q.w, A + 9 * i); }
    ok[h] = null_vector_8x9(A, e) ? 1 : 0;
}
__global__ void k5(const float4* pts, int n, int H, int* ok) {   // points first, then the rows
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    int idx[8];
    sample8(3u, 0, h, n, idx);
    float4 q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) q[i] = pts[idx[i]];
    float A[72], e[9];
#pragma unroll
    for (int i = 0; i < 8; ++i) epipolar_row(q[i].x, q[i].y, q[i].z, q[i].w, A + 9 * i);
    ok[h] = null_vector_8x9(A, e) ? 1 : 0;
}
__global__ void k6(const float4* pts, int n, int H, int* ok, float* Aout) {   // like k, dumps A before elimination for h = 0
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    int idx[8];
    sample8(3u, 0, h, n, idx);
    float A[72], e[9];
    for (int i = 0; i < 8; ++i) { const float4 q = pts[idx[i]]; epipolar_row(q.x, q.y, q.z, q.w, A + 9 * i); }
    if (h == 0) for (int i = 0; i < 72; ++i) Aout[i] = A[i];
    const bool r = null_vector_8x9(A, e);
    if (h == 0) for (int i = 0; i < 72; ++i) Aout[72 + i] = A[i];
    ok[h] = r ? 1 : 0;
}
__device__ __noinline__ bool null_vector_ni(float* A, float* e) { return null_vector_8x9(A, e); }
__device__ __noinline__ void project_essential_ni(float* e) { project_essential(e); }
__global__ void k7(const float4* pts, int n, int H, float* hyp, int* ok) {   // the shipped structure
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    int idx[8];
    sample8_ni(3u, 0, h, n, idx);
    float A[72], e[9];
    for (int i = 0; i < 8; ++i) { const float4 q = pts[idx[i]]; epipolar_row(q.x, q.y, q.z, q.w, A + 9 * i); }
    if (!null_vector_ni(A, e)) { ok[h] = 0; return; }
    project_essential_ni(e);
    for (int i = 0; i < 9; ++i) hyp[9 * h + i] = e[i];
    ok[h] = 1;
}
__global__ void k2(const float4* pts, const int* idxs, int H, int* ok) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    float A[72], e[9];
    for (int i = 0; i < 8; ++i) { const float4 q = pts[idxs[8 * h + i]]; epipolar_row(q.x, q.y, q.z, q.w, A + 9 * i); }
    ok[h] = null_vector_8x9(A, e) ? 1 : 0;
}
__global__ void k3(const float* As, int H, int* ok, float* best_out) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    float A[72], e[9];
    for (int i = 0; i < 72; ++i) A[i] = As[72 * h + i];
    ok[h] = null_vector_8x9(A, e) ? 1 : 0;
    if (h == 0) for (int i = 0; i < 72; ++i) best_out[i] = A[i];
}
int main() {
    const int n = 2000, H = 1024;
    std::vector<float4> pts(n);
    unsigned s = 777;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xffff) / 65536.0f; };
    const double a = 0.05, R[9] = {cos(a), 0, sin(a), 0, 1, 0, -sin(a), 0, cos(a)}, t[3] = {0.3, 0.1, -0.05};
    for (int i = 0; i < n; ++i) {
        const double X = 4 * rnd() - 2, Y = 3.2 * rnd() - 1.6, Z = 3 + 6 * rnd();
        const double X1 = R[0] * X + R[1] * Y + R[2] * Z + t[0], Y1 = R[3] * X + R[4] * Y + R[5] * Z + t[1], Z1 = R[6] * X + R[7] * Y + R[8] * Z + t[2];
        pts[i] = make_float4(float(X / Z), float(Y / Z), float(X1 / Z1), float(Y1 / Z1));
    }
    float4* d; float* hyp; int *ok, *idx;
    cudaMalloc(&d, n * sizeof(float4)); cudaMalloc(&hyp, H * 36); cudaMalloc(&ok, H * 4); cudaMalloc(&idx, 32);
    cudaMemcpy(d, pts.data(), n * sizeof(float4), cudaMemcpyHostToDevice);
    k<<<H / 128, 128>>>(d, n, H, hyp, ok, idx);
    std::vector<int> hok(H); int hidx[8];
    cudaError_t err = cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(hidx, idx, 32, cudaMemcpyDeviceToHost);
    int dev_ok = 0, host_ok = 0;
    for (int h = 0; h < H; ++h) {
        dev_ok += hok[h];
        int id[8]; sample8(3u, 0, h, n, id);
        float A[72], e[9];
        for (int i = 0; i < 8; ++i) epipolar_row(pts[id[i]].x, pts[id[i]].y, pts[id[i]].z, pts[id[i]].w, A + 9 * i);
        host_ok += null_vector_8x9(A, e) ? 1 : 0;
        if (h == 0) { printf("idx host:"); for (int i = 0; i < 8; ++i) printf(" %d", id[i]); printf("\nidx dev :"); for (int i = 0; i < 8; ++i) printf(" %d", hidx[i]); printf("\n"); }
    }
    {
        std::vector<int> idxs(8 * H); std::vector<float> As(72 * H);
        for (int h = 0; h < H; ++h) {
            sample8(3u, 0, h, n, &idxs[8 * h]);
            for (int i = 0; i < 8; ++i) { const float4 q = pts[idxs[8 * h + i]]; epipolar_row(q.x, q.y, q.z, q.w, &As[72 * h + 9 * i]); }
        }
        int* di; float* dA; float* dB;
        cudaMalloc(&di, idxs.size() * 4); cudaMalloc(&dA, As.size() * 4); cudaMalloc(&dB, 72 * 4);
        cudaMemcpy(di, idxs.data(), idxs.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dA, As.data(), As.size() * 4, cudaMemcpyHostToDevice);
        k2<<<H / 128, 128>>>(d, di, H, ok);
        cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
        int c2 = 0; for (int h = 0; h < H; ++h) c2 += hok[h];
        k3<<<H / 128, 128>>>(dA, H, ok, dB);
        cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
        int c3 = 0; for (int h = 0; h < H; ++h) c3 += hok[h];
        float hb[72]; cudaMemcpy(hb, dB, 288, cudaMemcpyDeviceToHost);
        k4<<<H / 128, 128>>>(d, n, H, ok);
        cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
        int c4 = 0; for (int h = 0; h < H; ++h) c4 += hok[h];
        k5<<<H / 128, 128>>>(d, n, H, ok);
        cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
        int c5 = 0; for (int h = 0; h < H; ++h) c5 += hok[h];
        float* dC; cudaMalloc(&dC, 144 * 4);
        k6<<<H / 128, 128>>>(d, n, H, ok, dC);
        cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
        int c6 = 0; for (int h = 0; h < H; ++h) c6 += hok[h];
        float hc[144]; cudaMemcpy(hc, dC, 576, cudaMemcpyDeviceToHost);
        k7<<<H / 128, 128>>>(d, n, H, hyp, ok);
        cudaMemcpy(hok.data(), ok, H * 4, cudaMemcpyDeviceToHost);
        int c7 = 0; for (int h = 0; h < H; ++h) c7 += hok[h];
        printf("k7 (shipped structure, noinline steps): %d\n", c7);
        printf("k4 (noinline sample8): %d, k5 (points first): %d, k6 (dump): %d\n", c4, c5, c6);
        printf("k6 A before (dev, h = 0) vs host:\n"); for (int r = 0; r < 8; ++r) { for (int c = 0; c < 9; ++c) printf(" %9.2e", hc[r * 9 + c] - As[r * 9 + c]); printf("\n"); }
        printf("k6 A after (dev, h = 0):\n"); for (int r = 0; r < 8; ++r) { for (int c = 0; c < 9; ++c) printf(" %9.2e", hc[72 + r * 9 + c]); printf("\n"); }
        printf("k2 (host idx): %d, k3 (host A): %d\n", c2, c3);
        printf("A after (dev, h = 0):\n"); for (int r = 0; r < 8; ++r) { for (int c = 0; c < 9; ++c) printf(" %9.2e", hb[r * 9 + c]); printf("\n"); }
        float A[72], e[9]; for (int i = 0; i < 72; ++i) A[i] = As[i];
        null_vector_8x9(A, e);
        printf("A after (host, h = 0):\n"); for (int r = 0; r < 8; ++r) { for (int c = 0; c < 9; ++c) printf(" %9.2e", A[r * 9 + c]); printf("\n"); }
    }
    printf("cuda: %s; models: device %d / %d, host %d / %d\n", cudaGetErrorString(err), dev_ok, H, host_ok, H);
    return 0;
}
