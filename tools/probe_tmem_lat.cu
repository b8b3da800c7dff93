// Probe: round-trip latencies the fine-transformer's serial chain is made of:
// tcgen05.ld x32 + wait, two loads + one wait, tcgen05.st x16 x2 + wait, named barrier (256 threads).
#include <cstdio>
#include "../featurematching_b200/csrc/fm_ptx.cuh"
using namespace fm;
__global__ void __launch_bounds__(256, 1) probe(long long* out) {
    __shared__ uint32_t tbase;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc<512>(&tbase);
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tl = tbase + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    uint32_t a[32], b[32];
    long long t[6];
    for (int rep = 0; rep < 3; ++rep) {
        t[0] = clock64();
        tmem_ld_32x32(tl, a); tmem_ld_wait_dep(a);
        t[1] = clock64();
        tmem_ld_32x32(tl + 32, a); tmem_ld_32x32(tl + 64, b); tmem_ld_wait_dep(a);
        t[2] = clock64();
        uint32_t h[16];
        for (int i = 0; i < 16; ++i) h[i] = a[i] ^ b[i];
        tmem_st_32x16(tl + 128, h); tmem_st_32x16(tl + 144, h); tmem_st_wait();
        t[3] = clock64();
        named_bar_sync(1, 256);
        t[4] = clock64();
        named_bar_sync(1, 256);
        t[5] = clock64();
    }
    if (threadIdx.x == 0) for (int i = 0; i < 5; ++i) out[i] = t[i + 1] - t[i];
    if (a[3] == 0x12345u) out[7] = b[5];
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tbase);
}
int main() {
    long long* d; cudaMalloc(&d, 64);
    probe<<<1, 256>>>(d);
    printf("launch: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    long long h[8]; cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
    printf("ld x32 + wait: %lld | 2 x ld x32 + wait: %lld | 2 x st x16 + wait: %lld | bar.sync 256 (1st, 2nd): %lld %lld cycles\n", h[0], h[1], h[2], h[3], h[4]);
    return 0;
}
