// Microbenchmark: per-SM throughput of the integer instructions the NTT butterflies are made of (sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipes int_pipes.cu ; run on a B200.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned int u32;
typedef unsigned long long u64;
#define ITERS 2048
#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)

template <int MODE>
__global__ void __launch_bounds__(256) k(u32* out, u32 a0, u32 b0) {
    u32 a = a0 + threadIdx.x, b = b0 | 1u;
    u32 r0 = a + 1, r1 = a + 2, r2 = a + 3, r3 = a + 4, r4 = a + 5, r5 = a + 6, r6 = a + 7, r7 = a + 8;
    u64 w0 = a + 1, w1 = a + 2, w2 = a + 3, w3 = a + 4, w4 = a + 5, w5 = a + 6, w6 = a + 7, w7 = a + 8;
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
        if (MODE == 0) {  // IMAD lo
#define X(n) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(r##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 1) {  // IMAD.WIDE
#define X(n) asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %2, %0;}" : "+l"(w##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 2) {  // IMAD.HI
#define X(n) asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(r##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 3) {  // IADD3
#define X(n) asm volatile("add.u32 %0, %0, %1;\n\txor.b32 %0, %0, %2;" : "+r"(r##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 4) {  // IMAD lo + IADD3 interleaved (8 + 8)
#define X(n) asm volatile("mad.lo.u32 %0, %0, %2, %0;\n\tadd.u32 %1, %1, %0;" : "+r"(r##n), "+r"(*((u32*)&w##n)) : "r"(a), "r"(b));
            REP8(X)
#undef X
        } else if (MODE == 5) {  // IMAD.WIDE + 2x IADD3 interleaved (8 + 16)
#define X(n) asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %3, %0;\n\tadd.u32 %1, %1, lo;\n\txor.b32 %1, %1, hi;}" : "+l"(w##n), "+r"(r##n) : "r"(a), "r"(b));
            REP8(X)
#undef X
        } else if (MODE == 6) {  // LOP3
#define X(n) asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(r##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 7) {  // 64-bit add = add.cc + addc
#define X(n) asm volatile("add.u64 %0, %0, %1;\n\txor.b64 %0, %0, %1;" : "+l"(w##n) : "l"((u64)a << 20 | b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 8) {  // setp + selp pair (csub flavour)
#define X(n) asm volatile("{.reg .pred p; setp.ge.u32 p, %0, %1; selp.u32 %0, %2, %0, p;}" : "+r"(r##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        }
    }
    u32 s = r0 ^ r1 ^ r2 ^ r3 ^ r4 ^ r5 ^ r6 ^ r7;
    u64 t = w0 ^ w1 ^ w2 ^ w3 ^ w4 ^ w5 ^ w6 ^ w7;
    if (a0 == 0xdeadbeefu) out[threadIdx.x] = s ^ (u32)t ^ (u32)(t >> 32);  // never true at run time, unknown to ptxas
}

template <int MODE>
static void run(const char* name, double ops_per_iter) {
    u32* d;
    cudaMalloc(&d, 4);
    int blocks = 148 * 8;
    k<MODE><<<blocks, 256>>>(d, 3, 5);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, 3, 5);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double inst = (double)blocks * 256 * ITERS * ops_per_iter;  // thread-instructions
    int clk;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    double per_sm_clk = inst / (ms * 1e-3) / 148.0 / (clk * 1e3);
    printf("%-28s %8.3f ms  %7.2f thread-inst/clk/SM (at %d MHz nominal)\n", name, ms, per_sm_clk, clk / 1000);
    cudaFree(d);
}

int main() {
    run<0>("IMAD (mad.lo.u32)", 16);
    run<1>("IMAD.WIDE (mad.wide.u32)", 16);
    run<2>("IMAD.HI (mad.hi.u32)", 16);
    run<3>("IADD3+LOP3 (add,xor)", 32);
    run<4>("IMAD + IADD3 (1:1)", 16);
    run<5>("IMAD.WIDE + 2 IADD3", 24);
    run<6>("LOP3", 16);
    run<7>("add.u64+xor.b64 (4 inst)", 64);
    run<8>("setp+selp", 32);
    return 0;
}
