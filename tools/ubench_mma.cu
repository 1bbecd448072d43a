// Throughput of the legacy warp-level integer MMA (mma.sync.m16n8k32 u8*u8 -> s32) on B200, in warp-instructions per
// clock per SM and TMAC/s: decides whether ModUp/ModDown base conversion is worth moving onto byte-sliced IMMA.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_mma tools/ubench_mma.cu && /tmp/ubench_mma
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) imma_loop(int iters, unsigned *out, long long *cycles) {
    unsigned a0 = threadIdx.x, a1 = a0 * 3 + 1, a2 = a0 * 5 + 2, a3 = a0 * 7 + 3, b0 = a0 ^ 0x55, b1 = a0 ^ 0xaa;
    int c[8][4];
#pragma unroll
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    long long t1 = clock64();
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    unsigned *out;
    long long *cyc, h = 0;
    cudaMalloc(&out, sizeof(unsigned) * sms * 4 * 256);
    cudaMalloc(&cyc, sizeof(long long));
    const int iters = 20000;
    for (int blocks_per_sm = 1; blocks_per_sm <= 4; blocks_per_sm *= 2) {
        imma_loop<<<sms * blocks_per_sm, 256>>>(100, out, cyc);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0), cudaEventCreate(&e1);
        cudaEventRecord(e0);
        imma_loop<<<sms * blocks_per_sm, 256>>>(iters, out, cyc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        const double warp_mma = (double)blocks_per_sm * 8 /*warps*/ * 8 * iters; // per SM
        printf("IMMA m16n8k32 u8: %d warps/SM  %.3f warp-inst/clk/SM  %.1f TMAC/s (whole GPU, %d SMs)  %.3f ms\n",
               blocks_per_sm * 8, warp_mma / (double)h, warp_mma * sms * 4096.0 / (ms * 1e-3) / 1e12, sms, ms);
    }
    return 0;
}
