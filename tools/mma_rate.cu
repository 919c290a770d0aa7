// GPU tool: issue rate of the legacy warp-level MMA (mma.sync) and dp4a on this chip, per SM and clock.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/mma_rate tools/mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

template <int KIND> __global__ void __launch_bounds__(512) rate_kernel(int iters, int * out, long long * clocks) {
    int c[8][4] = {};
    float f[8][4] = {};
    uint32_t a0 = threadIdx.x * 0x01010101u, a1 = a0 ^ 0x55, a2 = a0 + 7, a3 = a0 ^ 0x33, b0 = a0 * 3, b1 = a0 * 5;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (KIND == 0) {
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                             : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else if (KIND == 1) {
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                             : "+f"(f[j][0]), "+f"(f[j][1]), "+f"(f[j][2]), "+f"(f[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else if (KIND == 2) {
#pragma unroll
                for (int e = 0; e < 4; ++e) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(c[j][e]) : "r"(a0 + e), "r"(b0 + j));
            } else if (KIND == 3) {
                asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
                             : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3]) : "r"(a0), "r"(a1), "r"(b0));
            } else {
                asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                             : "+f"(f[j][0]), "+f"(f[j][1]), "+f"(f[j][2]), "+f"(f[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            }
        }
    }
    const long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) for (int e = 0; e < 4; ++e) s += c[j][e] + (int) f[j][e];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int KIND> void run(const char * name, int per_iter_ops, double macs_per_op) {
    int * out; long long * clk;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&clk, 148 * 8);
    const int iters = 4096;
    rate_kernel<KIND><<<148, 512>>>(16, out, clk);
    rate_kernel<KIND><<<148, 512>>>(iters, out, clk);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    const double ops = 16.0 * iters * per_iter_ops;        // warp-level instructions per SM
    printf("%-28s %s: %.1f clocks per warp instruction per SM  (%.0f MAC / clk / SM)\n", name, cudaGetErrorString(e), h[0] / ops, ops * macs_per_op / h[0]);
}

int main() {
    run<0>("IMMA m16n8k32 u8.s8", 8, 16 * 8 * 32);
    run<3>("IMMA m16n8k16 u8.s8", 8, 16 * 8 * 16);
    run<1>("HMMA m16n8k16 f16 -> f32", 8, 16 * 8 * 16);
    run<4>("QMMA m16n8k32 e4m3 -> f32", 8, 16 * 8 * 32);
    run<2>("dp4a (x4 per item)", 32, 32 * 4);
    return 0;
}
