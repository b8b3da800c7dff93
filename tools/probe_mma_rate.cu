// Probe: latency of a batch of n tcgen05.mma (M=128, N, K=16) + commit + wait, A in TMEM (TS) or smem (SS).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/probe_mma_rate.bin tools/probe_mma_rate.cu
#include <cstdio>
#include <cuda_bf16.h>
#include "../featurematching_b200/csrc/fm_ptx.cuh"
using namespace fm;

// same batch, accumulating alternately into two independent regions (D0 at column 0, D1 at column 64)
template <int N>
__device__ long long run_alt(uint32_t tm, uint32_t sbase, uint64_t* bar, uint32_t& ph, int n) {
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, N);
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        const int k = i & 3;
        umma_bf16_ts(tm + 64 * (i & 1), tm + 256 + 8 * (i & 7), umma_desc_k_sw128(sbase + 32 * k), idesc, i > 1 ? 1u : 0u);
    }
    umma_commit(bar);
    mbar_wait(bar, ph); ph ^= 1;
    return clock64() - t0;
}

template <int N, bool TS>
__device__ long long run(uint32_t tm, uint32_t sbase, uint64_t* bar, uint32_t& ph, int n) {
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, N);
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        const int k = i & 3;
        if (TS) umma_bf16_ts(tm, tm + 256 + 8 * (i & 7), umma_desc_k_sw128(sbase + 32 * k), idesc, i ? 1u : 0u);
        else umma_bf16_ss(tm, umma_desc_k_sw128(sbase + 32768 + 32 * k), umma_desc_k_sw128(sbase + 32 * k), idesc, i ? 1u : 0u);
    }
    umma_commit(bar);
    mbar_wait(bar, ph); ph ^= 1;
    return clock64() - t0;
}

__global__ void __launch_bounds__(640, 1) probe(long long* out, int pollers) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint64_t bar2;
    __shared__ volatile int done;
    __shared__ uint32_t tbase;
    const int t = threadIdx.x, warp = t >> 5;
    if (t == 0) { mbar_init(&bar, 1); mbar_init(&bar2, 1); done = 0; fence_mbar_init(); }
    if (warp == 0) tmem_alloc<512>(&tbase);
    for (int i = t; i < 49152 / 4; i += 640) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tbase;
    if (t == 0) {
        uint32_t ph = 0;
        const uint32_t sb = smem_u32(smem);
        int o = 0;
        for (int rep = 0; rep < 2; ++rep) {
            o = 0;
            out[o++] = run<64, true>(tm, sb, &bar, ph, 12);
            out[o++] = run<64, true>(tm, sb, &bar, ph, 24);
            out[o++] = run<128, true>(tm, sb, &bar, ph, 24);
            out[o++] = run<192, true>(tm, sb, &bar, ph, 12);
            out[o++] = run<64, false>(tm, sb, &bar, ph, 12);
            out[o++] = run<64, false>(tm, sb, &bar, ph, 24);
            out[o++] = run<128, false>(tm, sb, &bar, ph, 24);
            out[o++] = run<192, false>(tm, sb, &bar, ph, 12);
            out[o++] = run<64, true>(tm, sb, &bar, ph, 1);
            out[o++] = run<64, true>(tm, sb, &bar, ph, 96);
            out[o++] = run_alt<64>(tm, sb, &bar, ph, 24);
            out[o++] = run_alt<64>(tm, sb, &bar, ph, 12);
        }
        done = 1;
        mbar_arrive(&bar2);
    } else if (warp >= 4 && warp < 4 + pollers) {
        // spin like the epilogue warps do while they wait for an accumulator
        while (!mbar_try_wait(&bar2, 0)) { }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tm);
}

int main() {
    long long* d; cudaMalloc(&d, 16 * 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560);
    for (int pollers = 0; pollers <= 0; pollers += 8) {
    probe<<<1, 640, 66560>>>(d, pollers);
    printf("launch: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    long long h[16]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const char* names[] = {"TS N=64 x12", "TS N=64 x24", "TS N=128 x24", "TS N=192 x12", "SS N=64 x12", "SS N=64 x24",
                           "SS N=128 x24", "SS N=192 x12", "TS N=64 x1", "TS N=64 x96", "TS N=64 x24 2 accumulators", "TS N=64 x12 2 accumulators"};
    printf("pollers = %d\n", pollers);
    for (int i = 0; i < 12; ++i) printf("%-14s %lld cycles\n", names[i], h[i]);
    }
    return 0;
}
