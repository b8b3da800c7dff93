// Probe: where does tcgen05.mma with M = 64 (cta_group::1) put its accumulator rows in TMEM?
// A = 64 x 16 with A[i][0] = i + 1 (rest 0), B = 64 x 16 with B[n][0] = 1  ->  D[i][n] = i + 1.
#include <cstdio>
#include <vector>
#include <cuda_bf16.h>
#include "../featurematching_b200/csrc/fm_ptx.cuh"
using namespace fm;
__global__ void __launch_bounds__(128, 1) probe(float* out) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    const int t = threadIdx.x, warp = t >> 5;
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<128>(&tbase);
    for (int i = t; i < 16384 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    __syncthreads();
    if (t < 64) {   // A tile at smem, B tile at smem + 8192; row r: element k = 0 lives in chunk (0 ^ (r & 7))
        *reinterpret_cast<__nv_bfloat16*>(smem + t * 128 + ((0 ^ (t & 7)) << 4)) = __float2bfloat16(static_cast<float>(t + 1));
        *reinterpret_cast<__nv_bfloat16*>(smem + 8192 + t * 128 + ((0 ^ (t & 7)) << 4)) = __float2bfloat16(1.f);
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tbase;
    {   // clear the accumulator columns
        uint32_t z[16];
        for (int i = 0; i < 16; ++i) z[i] = 0;
        for (int c = 0; c < 64; c += 16) tmem_st_32x16(tm + (static_cast<uint32_t>(warp * 32) << 16) + c, z);
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {
        umma_bf16_ss(tm, umma_desc_k_sw128(smem_u32(smem)), umma_desc_k_sw128(smem_u32(smem) + 8192), umma_idesc_bf16_f32(64, 64), 0);
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    uint32_t lo[32];
    tmem_ld_32x32(tm + (static_cast<uint32_t>(warp * 32) << 16), lo);
    tmem_ld_wait();
    out[t] = __uint_as_float(lo[0]);
    out[128 + t] = __uint_as_float(lo[5]);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<128>(tm);
}
int main() {
    float* d; cudaMalloc(&d, 256 * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    probe<<<1, 128, 32768>>>(d);
    printf("launch: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    std::vector<float> h(256); cudaMemcpy(h.data(), d, 1024, cudaMemcpyDeviceToHost);
    printf("lane -> value of D column 0 (row index + 1, 0 = lane not written):\n");
    for (int l = 0; l < 128; ++l) printf("%3.0f%s", h[l], (l % 32 == 31) ? "\n" : " ");
    return 0;
}
