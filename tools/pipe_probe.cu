// Probe: issue rate of a few integer instructions on sm_100a (which pipe they share).  nvcc -arch=sm_100a -o pipe_probe pipe_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int OP> __global__ void k(unsigned *out, unsigned seed, int iters) {
    unsigned a[8];
    for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 7 + i;
    unsigned b = seed * 3 + 1;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(0x01ff01ff));
            if (OP == 1) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(a[(i + 1) & 7]));
            if (OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(a[(i + 1) & 7]));
            if (OP == 3) asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(a[(i + 1) & 7]));
            if (OP == 4) asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(a[i]) : "r"(b));
            if (OP == 5) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(a[(i + 1) & 7]));
            if (OP == 6) { asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(0x01ff01ff)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[(i + 4) & 7]) : "r"(b), "r"(a[(i + 1) & 7])); }
            if (OP == 7) { asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(a[(i + 1) & 7])); asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a[(i + 4) & 7]) : "r"(b), "r"(0x01ff01ff)); }
        }
    }
    unsigned s = 0;
    for (int i = 0; i < 8; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP> void run(const char *name, int per) {
    unsigned *out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 4096;
    k<OP><<<148 * 8, 256>>>(out, 1, 16);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a); k<OP><<<148 * 8, 256>>>(out, 2, iters); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double inst = 148.0 * 8 * 8 * iters * 8 * per;          // warp instructions
    printf("%-14s %.3f ms  %.2f warp-instr / clk / SM (1.965 GHz)\n", name, ms, inst / (ms * 1e-3) / 1.965e9 / 148);
    cudaFree(out);
}
int main() {
    run<0>("dp4a", 1); run<1>("imad", 1); run<2>("lop3", 1); run<3>("vabsdiff4", 1); run<4>("prmt", 1); run<5>("iadd", 1);
    run<6>("dp4a+lop3", 2); run<7>("imad+dp4a", 2);
    return 0;
}
