// GPU tool: issue rate of tcgen05.mma on this chip for kind::i8 and kind::f16 — the tensor-pipe peaks the prefill roofline is
// quoted against (BASELINE.md §3 asks for a measured kind::i8 figure).  One CTA per SM; one thread issues back-to-back
// M128 x N256 MMAs whose operands sit in shared memory (K-major, 128-byte swizzle, zero-filled: the tensor pipe does the same work
// for any values), two TMEM accumulators alternating; the CTA commits and waits once at the end.  Reports MAC / clock / SM from
// clock64 and whole-chip T(FL)OPS from CUDA events.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I llama.cpp-s390x_b200/csrc -o tools/bin/tc_rate tools/tc_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "tc_ptx.cuh"

// KIND 0: kind::i8 (s8 x s8 -> s32, K = 32 per instruction); KIND 1: kind::f16 (f16 x f16 -> f32, K = 16 per instruction)
template <int KIND>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, long long * clocks) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    const uint32_t sA = tc::smem_u32(smem), sB = sA + 128 * 128, bar = sB + 256 * 128, slot = bar + 8;
    for (int i = threadIdx.x; i < (128 + 256) * 128 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_alloc(slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t *>(smem + (128 + 256) * 128 + 8);
    tc::fence_proxy_async();
    if (threadIdx.x == 0) {
        // D format: 1 = f32, 2 = s32; A / B format: kind::f16 0 = f16; kind::i8 1 = signed 8-bit
        constexpr uint32_t idesc = KIND == 0 ? tc::umma_idesc(2, 1, 1, 128, 256) : tc::umma_idesc(1, 0, 0, 128, 256);
        const uint64_t da = tc::umma_desc_sw128(sA), db = tc::umma_desc_sw128(sB);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem + (uint32_t) (it & 1) * 256;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {            // one 128-byte operand row: 4 x K = 32 bytes (i8: 32 values, f16: 16 values)
                if (KIND == 0) tc::umma_i8(d, da + 2 * kk, db + 2 * kk, idesc, (uint32_t) (it > 1 || kk > 0));
                else           tc::umma_f16(d, da + 2 * kk, db + 2 * kk, idesc, (uint32_t) (it > 1 || kk > 0));
            }
        }
        tc::tc_commit(bar);
        tc::mbar_wait(bar, 0);
        clocks[blockIdx.x] = clock64() - t0;
    }
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

template <int KIND>
static void run(const char * name, int sms, int iters, long long * d_clk) {
    const size_t smem = (128 + 256) * 128 + 64 + 1024;
    cudaFuncSetAttribute(rate_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    rate_kernel<KIND><<<sms, 128, smem>>>(64, d_clk);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    rate_kernel<KIND><<<sms, 128, smem>>>(iters, d_clk);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); return; }
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long clk[256] = {};
    cudaMemcpy(clk, d_clk, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    long long worst = 0; for (int i = 0; i < sms; ++i) worst = clk[i] > worst ? clk[i] : worst;
    const double k_per_mma = KIND == 0 ? 32.0 : 16.0;
    const double mac_sm = (double) iters * 4 * 128 * 256 * k_per_mma;
    printf("%-44s %d SMs x %d MMAs (M128 N256 K%d): %.1f MAC/clk/SM (slowest SM), %.3f ms, %.1f T%sOPS whole chip\n", name, sms, iters * 4, (int) k_per_mma,
           mac_sm / (double) worst, ms, 2.0 * mac_sm * sms / (ms * 1e-3) / 1e12, KIND == 0 ? "" : "FL");
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    long long * d_clk; cudaMalloc(&d_clk, sizeof(long long) * 256);
    printf("%s, %d SMs\n", prop.name, sms);
    for (int rep = 0; rep < 2; ++rep) {
        run<1>("tcgen05.mma kind::f16 (f16 x f16 -> f32)", sms, 4096, d_clk);
        run<0>("tcgen05.mma kind::i8  (s8 x s8 -> s32)", sms, 4096, d_clk);
    }
    return 0;
}
