// Probe: tcgen05.mma with the A operand in TMEM (written by tcgen05.st as packed bf16 pairs).
// Establishes the operand layout the fine-transformer kernel relies on:
//   lane = row, 32-bit column c of the operand holds K elements (2c, 2c+1), low half = even k.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/probe_ts_mma tools/probe_ts_mma.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_bf16.h>
#include "../featurematching_b200/csrc/fm_ptx.cuh"

using namespace fm;

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// A [128][64] bf16 row-major, B [N=64][K=64] bf16 row-major (K contiguous), D [128][64] fp32
__global__ void __launch_bounds__(128, 1) probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    const int t = threadIdx.x, warp = t >> 5;
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<128>(&tbase);
    // B tile: row n at n*128, 16-byte chunk c at ((c ^ (n & 7)) << 4)   (SW128 K-major)
    if (t < 64) {
        const uint4* src = reinterpret_cast<const uint4*>(B + t * 64);
        for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(smem + t * 128 + ((c ^ (t & 7)) << 4)) = src[c];
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tbase;
    // A operand: thread t = row t; columns [64, 96)
    uint32_t r[32];
    const uint32_t* arow = reinterpret_cast<const uint32_t*>(A + t * 64);
    for (int c = 0; c < 32; ++c) r[c] = arow[c];   // (k = 2c) low half, (k = 2c + 1) high half
    tmem_st_32x32(tm + (static_cast<uint32_t>(warp * 32) << 16) + 64, r);
    tmem_st_wait();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16_f32(128, 64);
        for (int k = 0; k < 4; ++k)
            umma_bf16_ts(tm, tm + 64 + k * 8, umma_desc_k_sw128(smem_u32(smem) + k * 32), idesc, k ? 1u : 0u);
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    uint32_t lo[32], hi[32];
    tmem_ld_32x32(tm + (static_cast<uint32_t>(warp * 32) << 16), lo);
    tmem_ld_32x32(tm + (static_cast<uint32_t>(warp * 32) << 16) + 32, hi);
    tmem_ld_wait();
    for (int c = 0; c < 32; ++c) { D[t * 64 + c] = __uint_as_float(lo[c]); D[t * 64 + 32 + c] = __uint_as_float(hi[c]); }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<128>(tm);
}

int main() {
    std::vector<__nv_bfloat16> A(128 * 64), B(64 * 64);
    srand(1);
    for (auto& v : A) v = __float2bfloat16((rand() % 2001 - 1000) / 500.f);
    for (auto& v : B) v = __float2bfloat16((rand() % 2001 - 1000) / 500.f);
    __nv_bfloat16 *dA, *dB; float* dD;
    cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, 128 * 64 * 4);
    cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
    probe<<<1, 128, 16384>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("launch: %s\n", cudaGetErrorString(e));
    std::vector<float> D(128 * 64);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 64; ++n) {
            double acc = 0;
            for (int k = 0; k < 64; ++k) acc += double(__bfloat162float(A[m * 64 + k])) * __bfloat162float(B[n * 64 + k]);
            maxerr = fmax(maxerr, fabs(acc - D[m * 64 + n]));
        }
    printf("TS-MMA max abs err vs fp64: %.3e  (%s)\n", maxerr, maxerr < 1e-3 ? "LAYOUT OK" : "MISMATCH");
    return maxerr < 1e-3 ? 0 : 1;
}
