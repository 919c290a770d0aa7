// GPU tool: tcgen05.mma with the A operand in TENSOR MEMORY (written by tcgen05.st from registers) — (1) checks the layout
// assumption of the prefill GEMM (lane = tile row, 32-bit column c of a K step = K elements 2c, 2c+1, a K = 16 step = 8 columns)
// against a host product, (2) measures the issue rate of that form with and without concurrent tcgen05.st traffic from 8 other
// warps, for N = 256 and N = 128, beside the shared-memory-A form.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I llama.cpp-s390x_b200/csrc -o tools/bin/tc_ts_probe tools/tc_ts_probe.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

#include "tc_ptx.cuh"

// ---- (1) layout --------------------------------------------------------------------------------------------------------
// A[128][64], B[128][64] small integers as fp16; D = A B^T (128 x 128) through 4 K steps
__global__ void __launch_bounds__(128, 1) layout_kernel(const __half * A, const __half * B, float * D) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    const uint32_t sB = tc::smem_u32(smem), bar = sB + 128 * 128, slot = bar + 8;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // B: K-major rows of 128 bytes, SWIZZLE_128B: 16-byte chunk j of row n sits at chunk j ^ (n & 7)
    for (int i = tid; i < 128 * 64; i += 128) {
        const int n = i >> 6, k = i & 63;
        const uint32_t off = (uint32_t) ((n >> 3) * 1024 + (n & 7) * 128 + (((k >> 3) ^ (n & 7)) << 4) + (k & 7) * 2);
        *reinterpret_cast<__half *>(smem + off) = B[i];
    }
    if (tid == 0) { tc::mbar_init(bar, 1); tc::mbar_fence_init(); }
    __syncthreads();
    if (warp == 0) tc::tmem_alloc(slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t *>(smem + 128 * 128 + 8);
    // A: thread = row; columns 256 .. 287 of its lane
    {
        const int row = 32 * warp + lane;
        const uint32_t * a32 = reinterpret_cast<const uint32_t *>(A + row * 64);
        uint32_t v0[16], v1[16];
        for (int c = 0; c < 16; ++c) { v0[c] = a32[c]; v1[c] = a32[16 + c]; }
        const uint32_t ta = tmem + ((uint32_t) (32 * warp) << 16) + 256u;
        tc::tmem_st16(ta, v0);
        tc::tmem_st16(ta + 16, v1);
        tc::tmem_st_wait();
    }
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    if (tid == 0) {
        constexpr uint32_t idesc = tc::umma_idesc(1, 0, 0, 128, 128);
        const uint64_t db = tc::umma_desc_sw128(sB);
        for (int kk = 0; kk < 4; ++kk) tc::umma_f16_ts(tmem, tmem + 256u + 8u * kk, db + 2 * kk, idesc, (uint32_t) (kk > 0));
        tc::tc_commit(bar);
        tc::mbar_wait(bar, 0);
    }
    __syncthreads();
    tc::tc_fence_after();
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t v[32];
        tc::tmem_ld32(tmem + ((uint32_t) (32 * warp) << 16) + (uint32_t) c0, v);
        tc::tmem_ld_wait();
        for (int c = 0; c < 32; ++c) D[(32 * warp + lane) * 128 + c0 + c] = __uint_as_float(v[c]);
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ---- (2) rate ----------------------------------------------------------------------------------------------------------
// warp 8 issues the MMAs; what else runs is chosen by LOAD bits:
//   1: warps 0-7 keep writing 32-column A stages into a 4-slot TMEM ring (tcgen05.st)
//   2: two tcgen05.commit per 4 MMAs (as the GEMM's stage loop does), on barriers nobody waits for
//   4: warps 0-7 keep writing shared memory with st.shared.v4 (a separate 32 KB region)
//   8: warp 9 keeps fetching 24 KB blocks from global memory into a 2-slot shared-memory ring (cp.async.bulk)
//  16: warps 0-7 keep READING shared memory with ld.shared.v4
//  32: operands hold random finite fp16 values instead of zeros (does the pipe's rate depend on the data?)
template <int N, bool TS, int LOAD>
__global__ void __launch_bounds__(320, 1) rate_kernel(int iters, long long * clocks, const uint8_t * gsrc) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    const uint32_t sA = tc::smem_u32(smem), sB = sA + 128 * 128, bar = sB + 256 * 128, slot = bar + 8, bar2 = bar + 16, bar3 = bar + 24;
    const uint32_t sScratch = sB + 256 * 128 + 1024, sBulk = sScratch + 32768, barBulk = bar + 32;
    __shared__ volatile int stop;
    for (int i = threadIdx.x; i < (128 + 256) * 128 / 4; i += blockDim.x) {
        uint32_t r = (uint32_t) i * 2654435761u + blockIdx.x * 40503u; r ^= r >> 13; r *= 2246822519u; r ^= r >> 16;
        reinterpret_cast<uint32_t *>(smem)[i] = (LOAD & 32) ? (r & 0xBBFFBBFFu) : 0u;       // exponent < 15: |x| < 2
    }
    if (threadIdx.x == 0) {
        tc::mbar_init(bar, 1); tc::mbar_init(bar2, 1); tc::mbar_init(bar3, 1); tc::mbar_init(barBulk, 1); tc::mbar_init(barBulk + 8, 1);
        tc::mbar_fence_init(); stop = 0;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 8) tc::tmem_alloc(slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t *>(smem + (128 + 256) * 128 + 8);
    tc::fence_proxy_async();
    if ((LOAD & 32) && warp < 4) {          // random A stages in tensor memory (columns 256 .. 383)
        uint32_t v[16];
        for (int j = 0; j < 8; ++j) {
            for (int c = 0; c < 16; ++c) { uint32_t r = (uint32_t) (threadIdx.x * 131 + j * 16 + c) * 2654435761u; r ^= r >> 13; r *= 2246822519u; r ^= r >> 16; v[c] = r & 0xBBFFBBFFu; }
            tc::tmem_st16(tmem + ((uint32_t) (32 * warp) << 16) + 256u + 16u * j, v);
        }
        tc::tmem_st_wait();
        tc::tc_fence_before();
    }
    __syncthreads();
    tc::tc_fence_after();
    if (warp < 8) {
        if (LOAD & 1) {
            uint32_t v[16];
            for (int c = 0; c < 16; ++c) v[c] = 0;
            const uint32_t ta = tmem + ((uint32_t) (32 * (warp & 3)) << 16) + 384u;
            int s = warp >> 2;
            while (!stop) {
                tc::tmem_st16(ta + 32u * s, v);
                tc::tmem_st16(ta + 32u * s + 16u, v);
                tc::tmem_st_wait();
                s = (s + 2) & 3;
            }
        }
        if (LOAD & 4) {
            const uint32_t a = sScratch + (uint32_t) (threadIdx.x * 16);
            while (!stop) {
#pragma unroll
                for (int j = 0; j < 8; ++j) asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" :: "r"(a + 4096u * j), "r"(j) : "memory");
            }
        }
        if (LOAD & 16) {
            const uint32_t a = sScratch + (uint32_t) (threadIdx.x * 16);
            uint32_t acc = 0;
            while (!stop) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint32_t x, y, z, w;
                    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(a + 4096u * j) : "memory");
                    acc += x ^ y ^ z ^ w;
                }
            }
            if (acc == 0x12345678u) clocks[200] = acc;
        }
    } else if (warp == 9) {
        if ((LOAD & 8) && lane == 0) {
            uint32_t ph[2] = {0, 0};
            int s = 0, blk = 0;
            // prime both slots, then refill each as it lands
            for (int i = 0; i < 2; ++i) {
                tc::mbar_arrive_expect_tx(barBulk + 8 * i, 24576);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(sBulk + 24576u * i), "l"(gsrc + (size_t) ((blockIdx.x * 64 + blk++) & 4095) * 24576), "r"(24576), "r"(barBulk + 8 * i) : "memory");
            }
            while (!stop) {
                tc::mbar_wait(barBulk + 8 * s, ph[s]); ph[s] ^= 1;
                tc::mbar_arrive_expect_tx(barBulk + 8 * s, 24576);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(sBulk + 24576u * s), "l"(gsrc + (size_t) ((blockIdx.x * 64 + blk++) & 4095) * 24576), "r"(24576), "r"(barBulk + 8 * s) : "memory");
                s ^= 1;
            }
            tc::mbar_wait(barBulk, ph[0]); tc::mbar_wait(barBulk + 8, ph[1]);
        }
    } else if (threadIdx.x == 256) {
        constexpr uint32_t idesc = tc::umma_idesc(1, 0, 0, 128, N);
        const uint64_t da = tc::umma_desc_sw128(sA), db = tc::umma_desc_sw128(sB);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem + (uint32_t) (it & 1) * (N > 128 ? 0 : 128);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                if (TS) tc::umma_f16_ts(d, tmem + 256u + 32u * (it & 3) + 8u * kk, db + 2 * kk, idesc, (uint32_t) (it > 1 || kk > 0));
                else    tc::umma_f16(d, da + 2 * kk, db + 2 * kk, idesc, (uint32_t) (it > 1 || kk > 0));
            }
            if (LOAD & 2) { tc::tc_commit(bar2); tc::tc_commit(bar3); }
        }
        tc::tc_commit(bar);
        tc::mbar_wait(bar, 0);
        clocks[blockIdx.x] = clock64() - t0;
        stop = 1;
    }
    __syncthreads();
    if (warp == 8) tc::tmem_dealloc(tmem, 512);
}

static uint8_t * g_src = nullptr;

template <int N, bool TS, int LOAD>
static void run(const char * name, int sms, int iters, long long * d_clk) {
    const size_t smem = (128 + 256) * 128 + 1024 + 32768 + 2 * 24576 + 1024;
    cudaFuncSetAttribute(rate_kernel<N, TS, LOAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    rate_kernel<N, TS, LOAD><<<sms, 320, smem>>>(64, d_clk, g_src);
    cudaDeviceSynchronize();
    rate_kernel<N, TS, LOAD><<<sms, 320, smem>>>(iters, d_clk, g_src);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); exit(1); }
    long long clk[256] = {};
    cudaMemcpy(clk, d_clk, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    long long worst = 0; for (int i = 0; i < sms; ++i) worst = clk[i] > worst ? clk[i] : worst;
    const double mac_sm = (double) iters * 4 * 128 * N * 16.0;
    printf("%-86s %6.1f MAC/clk/SM, %4.0f clk per 64-K stage\n", name, mac_sm / (double) worst, (double) worst / iters);
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    printf("%s, %d SMs\n", prop.name, sms);
    // (1)
    std::vector<__half> A(128 * 64), B(128 * 64);
    std::vector<float> Af(128 * 64), Bf(128 * 64), D(128 * 128);
    for (int r = 0; r < 128; ++r) for (int k = 0; k < 64; ++k) {
        Af[r * 64 + k] = (float) ((r * 3 + k * 5) % 7 - 3); Bf[r * 64 + k] = (float) ((r + 2 * k + k / 16) % 5 - 2);
        A[r * 64 + k] = __float2half(Af[r * 64 + k]); B[r * 64 + k] = __float2half(Bf[r * 64 + k]);
    }
    __half * dA, * dB; float * dD;
    cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, D.size() * 4);
    cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
    const size_t smem1 = 128 * 128 + 64 + 1024;
    layout_kernel<<<1, 128, smem1>>>(dA, dB, dD);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("layout kernel: %s\n", cudaGetErrorString(err)); return 1; }
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r = 0; r < 128; ++r) for (int n = 0; n < 128; ++n) {
        float ref = 0; for (int k = 0; k < 64; ++k) ref += Af[r * 64 + k] * Bf[n * 64 + k];
        if (ref != D[r * 128 + n]) { if (bad < 5) printf("  D[%d][%d] = %g, expected %g\n", r, n, D[r * 128 + n], ref); ++bad; }
    }
    printf("A operand in tensor memory, layout check: %s (%d of 16384 differ)\n", bad ? "MISMATCH" : "OK", bad);
    // (2)
    long long * d_clk; cudaMalloc(&d_clk, sizeof(long long) * 256);
    cudaMalloc(&g_src, (size_t) 4096 * 24576); cudaMemset(g_src, 0, (size_t) 4096 * 24576);
    run<256, false, 0>("A from shared memory, N = 256", sms, 4096, d_clk);
    run<128, false, 0>("A from shared memory, N = 128", sms, 4096, d_clk);
    run<256, true,  0>("A from tensor memory, N = 256", sms, 4096, d_clk);
    run<192, true,  0>("A from tensor memory, N = 192", sms, 4096, d_clk);
    run<128, true,  0>("A from tensor memory, N = 128", sms, 4096, d_clk);
    run<64,  true,  0>("A from tensor memory, N = 64", sms, 4096, d_clk);
    run<192, true,  1>("A from tensor memory, N = 192, 8 warps storing A stages (tcgen05.st)", sms, 4096, d_clk);
    run<192, true,  2>("A from tensor memory, N = 192, two commits per stage", sms, 4096, d_clk);
    run<64,  true,  2>("A from tensor memory, N = 64, two commits per stage", sms, 4096, d_clk);
    run<192, true,  4>("A from tensor memory, N = 192, 8 warps writing shared memory (st.shared.v4)", sms, 4096, d_clk);
    run<192, true, 16>("A from tensor memory, N = 192, 8 warps reading shared memory (ld.shared.v4)", sms, 4096, d_clk);
    run<192, true,  8>("A from tensor memory, N = 192, bulk copies global -> shared (24 KB blocks, 2 in flight)", sms, 4096, d_clk);
    run<64,  true,  8>("A from tensor memory, N = 64, bulk copies global -> shared (24 KB blocks, 2 in flight)", sms, 4096, d_clk);
    run<192, true, 11>("A from tensor memory, N = 192, tcgen05.st + commits + bulk copies", sms, 4096, d_clk);
    run<256, false, 8>("A from shared memory, N = 256, bulk copies global -> shared", sms, 4096, d_clk);
    run<256, false, 10>("A from shared memory, N = 256, commits + bulk copies", sms, 4096, d_clk);
    run<192, true, 32>("A from tensor memory, N = 192, RANDOM operand values", sms, 4096, d_clk);
    run<192, true, 32>("A from tensor memory, N = 192, RANDOM operand values, 16 x longer", sms, 65536, d_clk);
    run<256, false, 32>("A from shared memory, N = 256, RANDOM operand values, 16 x longer", sms, 65536, d_clk);
    run<192, true, 43>("A from tensor memory, N = 192, RANDOM values + tcgen05.st + commits + bulk copies", sms, 16384, d_clk);
    return 0;
}
