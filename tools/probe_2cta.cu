// Probe: tcgen05.mma.cta_group::2 (M = 256 across a CTA pair, N = 256, K = 64).  Pins what the coarse
// transformer's pair kernel relies on: CTA r supplies rows [128 r, 128 r + 128) of A and rows
// [128 r, 128 r + 128) of B (the N half), gets rows [128 r, ...) x all 256 columns of D in its own TMEM;
// alloc / dealloc with cta_group::2 in both CTAs; commit multicast to both CTAs' barriers.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/probe_2cta.bin tools/probe_2cta.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_bf16.h>
#include "../featurematching_b200/csrc/fm_ptx.cuh"

using namespace fm;

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// A [256][64], B [256 n][64 k] bf16 row-major, D [256][256] fp32
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    const int t = threadIdx.x, warp = t >> 5;
    const uint32_t rank = cluster_ctarank();
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    uint8_t* sa = smem;            // 128 rows x 128 B
    uint8_t* sb = smem + 16384;    // 128 rows x 128 B
    {
        const uint4* srca = reinterpret_cast<const uint4*>(A + (rank * 128 + t) * 64);
        const uint4* srcb = reinterpret_cast<const uint4*>(B + (rank * 128 + t) * 64);
        for (int c = 0; c < 8; ++c) {
            *reinterpret_cast<uint4*>(sa + t * 128 + ((c ^ (t & 7)) << 4)) = srca[c];
            *reinterpret_cast<uint4*>(sb + t * 128 + ((c ^ (t & 7)) << 4)) = srcb[c];
        }
    }
    fence_proxy_async_smem();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tbase)), "r"(256) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tm = tbase;
    if (rank == 0 && t == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16_f32(256, 256);
        for (int k = 0; k < 4; ++k) {
            const uint64_t ad = umma_desc_k_sw128(smem_u32(sa) + k * 32), bd = umma_desc_k_sw128(smem_u32(sb) + k * 32);
            const uint32_t acc = k ? 1u : 0u;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
                         ::"r"(tm), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     ::"r"(smem_u32(&bar)), "h"(static_cast<uint16_t>(3)) : "memory");
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    for (int c = 0; c < 8; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tm + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, r);
        tmem_ld_wait();
        for (int i = 0; i < 32; ++i) D[(rank * 128 + t) * 256 + c * 32 + i] = __uint_as_float(r[i]);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256) : "memory");
}

int main() {
    std::vector<__nv_bfloat16> A(256 * 64), B(256 * 64);
    srand(1);
    for (auto& v : A) v = __float2bfloat16((rand() % 2001 - 1000) / 500.f);
    for (auto& v : B) v = __float2bfloat16((rand() % 2001 - 1000) / 500.f);
    __nv_bfloat16 *dA, *dB; float* dD;
    cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, 256 * 256 * 4);
    cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, 256 * 256 * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40960);
    probe<<<2, 128, 40960>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("launch: %s\n", cudaGetErrorString(e));
    std::vector<float> D(256 * 256);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int m = 0; m < 256; ++m)
        for (int n = 0; n < 256; ++n) {
            double acc = 0;
            for (int k = 0; k < 64; ++k) acc += double(__bfloat162float(A[m * 64 + k])) * __bfloat162float(B[n * 64 + k]);
            maxerr = fmax(maxerr, fabs(acc - D[m * 256 + n]));
        }
    printf("2-CTA MMA max abs err vs fp64: %.3e  (%s)\n", maxerr, maxerr < 1e-3 ? "LAYOUT OK" : "MISMATCH");
    return maxerr < 1e-3 ? 0 : 1;
}
