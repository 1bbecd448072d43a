// ubench_int.cu — integer-pipe throughput probes for sm_100a (results recorded in DESIGN.md, "Integer roofline").
// Each kernel runs ITER iterations of 8 independent dependency chains of one instruction; reports warp-instructions
// per cycle per SM from clock64().  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_int ubench_int.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
constexpr int ITER = 4096;
#define CHAINS 8
template <int OP> __global__ void probe(u64 *out, unsigned seed, long long *cycles) {
    unsigned a[CHAINS], b = seed | 1u, c = seed * 3u + 7u;
    u64 w[CHAINS];
    double d[CHAINS];
    for (int i = 0; i < CHAINS; i++) a[i] = threadIdx.x + i + seed, w[i] = a[i] * 0x9E3779B97F4A7C15ull, d[i] = a[i];
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) {
            if (OP == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 1) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b));
            if (OP == 2) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 4) asm volatile("mul.hi.u64 %0, %0, %1;" : "+l"(w[i]) : "l"((u64)b << 32 | c));
            if (OP == 5) asm volatile("mul.lo.u64 %0, %0, %1;" : "+l"(w[i]) : "l"((u64)b << 32 | c));
            if (OP == 6) asm volatile("add.u64 %0, %0, %1;" : "+l"(w[i]) : "l"((u64)b << 32 | c));
            if (OP == 7) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(0.5));
            if (OP == 8) { // IMAD.WIDE and IADD3 interleaved (dual issue across pipes?)
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(b), "r"(c));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            }
            if (OP == 9) { // 64-bit compare + select (conditional subtract)
                u64 t = w[i] - ((u64)b << 32);
                asm volatile("{ .reg .pred p; setp.ge.u64 p, %0, %1; selp.u64 %0, %2, %0, p; }" : "+l"(w[i]) : "l"((u64)b << 32), "l"(t));
            }
            if (OP == 10) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 11) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(b));
        }
    }
    long long t1 = clock64();
    u64 s = 0;
    for (int i = 0; i < CHAINS; i++) s += a[i] + w[i] + (u64)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int OP> void run(const char *name, int per_iter) {
    u64 *out; long long *cyc, h;
    const int threads = 1024, blocks = 148 * 2; // 2048 threads per SM: full occupancy
    cudaMalloc(&out, sizeof(u64) * threads * blocks); cudaMalloc(&cyc, 8);
    probe<OP><<<blocks, threads>>>(out, 12345u, cyc);
    probe<OP><<<blocks, threads>>>(out, 12345u, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double winst = (double)ITER * CHAINS * per_iter * (2048 / 32); // warp-instructions per SM
    printf("%-28s %8.3f warp-inst/clk/SM  (%6.1f lanes/clk/SM)  cycles=%lld\n", name, winst / h, 32 * winst / h, h);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<0>("mad.lo.u32 (IMAD)", 1);
    run<1>("mad.wide.u32 (IMAD.WIDE)", 1);
    run<2>("mad.hi.u32", 1);
    run<3>("add.u32 (IADD3)", 1);
    run<4>("mul.hi.u64", 1);
    run<5>("mul.lo.u64", 1);
    run<6>("add.u64", 1);
    run<7>("fma.f64 (DFMA)", 1);
    run<8>("IMAD.WIDE + IADD3 pair", 2);
    run<9>("setp.ge.u64 + selp.u64", 1);
    run<10>("lop3.b32", 1);
    run<11>("shf.l.wrap", 1);
    return 0;
}
