// Microbenchmark: can the FP64 pipe of B200 (64 DFMA lanes / SM / clk by the data sheet) carry part of the
// modular-multiply work that saturates the fmaheavy (IMAD) pipe?  Measures DFMA / DADD alone, DFMA interleaved
// with IMAD.WIDE (do the two pipes issue concurrently?), the integer Shoup product of modarith.cuh against a
// hybrid whose two IMAD.HI partial products are DFMAs (bit-identical result, checked here), and base-conversion
// style multiply-accumulate terms on either pipe.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipes fp64_pipes.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned int u32;
typedef unsigned long long u64;
#define ITERS 1024
#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)

__device__ __forceinline__ u64 shoup_int(u64 x, u64 w, u64 ws, u64 q) {
    const u64 nq = 0ull - q;
    u64 r;
    asm("{\n\t"
        ".reg .u32 x0, x1, w0, w1, s0, s1, n0, n1, a, b, m0, m1, h0, h1, r0, r1;\n\t"
        ".reg .u64 h, t;\n\t"
        "mov.b64 {x0, x1}, %1;\n\t"
        "mov.b64 {w0, w1}, %2;\n\t"
        "mov.b64 {s0, s1}, %3;\n\t"
        "mov.b64 {n0, n1}, %4;\n\t"
        "mul.hi.u32 a, x0, s1;\n\t"
        "mul.hi.u32 b, x1, s0;\n\t"
        "add.cc.u32 m0, a, b;\n\t"
        "addc.u32 m1, 0, 0;\n\t"
        "mov.b64 h, {m0, m1};\n\t"
        "mad.wide.u32 h, x1, s1, h;\n\t"
        "mov.b64 {h0, h1}, h;\n\t"
        "mul.wide.u32 t, x0, w0;\n\t"
        "mad.wide.u32 t, h0, n0, t;\n\t"
        "mov.b64 {r0, r1}, t;\n\t"
        "mad.lo.u32 r1, x0, w1, r1;\n\t"
        "mad.lo.u32 r1, x1, w0, r1;\n\t"
        "mad.lo.u32 r1, h0, n1, r1;\n\t"
        "mad.lo.u32 r1, h1, n0, r1;\n\t"
        "mov.b64 %0, {r0, r1};\n\t"
        "}"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "l"(nq));
    return r;
}

// Same h = x1*s1 + floor(x0*s1/2^32) + floor(x1*s0/2^32): the two floors are formed on the FP64 pipe.
// (hi=0x41300000, lo=v) is the double 2^20 + v*2^-32, (hi=0x43300000, lo=v) is 2^52 + v; a round-down FMA onto
// 2^52 leaves floor(.) in the low mantissa bits (ulp = 1).
__device__ __forceinline__ u64 shoup_fp(u64 x, u64 w, u64 ws, u64 q) {
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32), s0 = (u32)ws, s1 = (u32)(ws >> 32);
    const double x0d = __hiloint2double(0x41300000, x0) - 0x1p20;
    const double x1d = __hiloint2double(0x41300000, x1) - 0x1p20;
    const double s0d = __hiloint2double(0x43300000, s0) - 0x1p52;
    const double s1d = __hiloint2double(0x43300000, s1) - 0x1p52;
    double t = __fma_rd(x1d, s0d, 0x1p52);
    t = __fma_rd(x0d, s1d, t);
    const u32 m0 = (u32)__double2loint(t), m1 = (u32)__double2hiint(t) & 1u;
    const u64 nq = 0ull - q;
    u64 r;
    asm("{\n\t"
        ".reg .u32 x0, x1, w0, w1, s0, s1, n0, n1, h0, h1, r0, r1;\n\t"
        ".reg .u64 h, t;\n\t"
        "mov.b64 {x0, x1}, %1;\n\t"
        "mov.b64 {w0, w1}, %2;\n\t"
        "mov.b64 {s0, s1}, %3;\n\t"
        "mov.b64 {n0, n1}, %4;\n\t"
        "mov.b64 h, {%5, %6};\n\t"
        "mad.wide.u32 h, x1, s1, h;\n\t"
        "mov.b64 {h0, h1}, h;\n\t"
        "mul.wide.u32 t, x0, w0;\n\t"
        "mad.wide.u32 t, h0, n0, t;\n\t"
        "mov.b64 {r0, r1}, t;\n\t"
        "mad.lo.u32 r1, x0, w1, r1;\n\t"
        "mad.lo.u32 r1, x1, w0, r1;\n\t"
        "mad.lo.u32 r1, h0, n1, r1;\n\t"
        "mad.lo.u32 r1, h1, n0, r1;\n\t"
        "mov.b64 %0, {r0, r1};\n\t"
        "}"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "l"(nq), "r"(m0), "r"(m1));
    return r;
}

template <int MODE>
__global__ void __launch_bounds__(256) k(u32* out, u32 a0, u32 b0, u64 q) {
    u32 a = a0 + threadIdx.x, b = b0 | 1u;
    u64 w0 = a + 1, w1 = a + 2, w2 = a + 3, w3 = a + 4, w4 = a + 5, w5 = a + 6, w6 = a + 7, w7 = a + 8;
    double d0 = a + 1, d1 = a + 2, d2 = a + 3, d3 = a + 4, d4 = a + 5, d5 = a + 6, d6 = a + 7, d7 = a + 8;
    double e0 = a + 1.5, e1 = a + 2.5, e2 = a + 3.5, e3 = a + 4.5, e4 = a + 5.5, e5 = a + 6.5, e6 = a + 7.5, e7 = a + 8.5;
    const double da = 1.0 + 1e-9 * a, db = 1e-7 * b;
    u64 tw = 0x0123456789abcdefull ^ a, tws = 0xfedcba9876543210ull ^ b;
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
        if (MODE == 0) {  // DFMA, 16 independent chains
#define X(n) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d##n) : "d"(da), "d"(db));
            REP8(X)
#undef X
#define X(n) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(e##n) : "d"(da), "d"(db));
            REP8(X)
#undef X
        } else if (MODE == 1) {  // DADD
#define X(n) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d##n) : "d"(db));
            REP8(X)
#undef X
#define X(n) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(e##n) : "d"(db));
            REP8(X)
#undef X
        } else if (MODE == 2) {  // 8 DFMA + 8 IMAD.WIDE interleaved
#define X(n)                                                                                               \
    asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d##n) : "d"(da), "d"(db));                            \
    asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %2, %0;}" : "+l"(w##n) : "r"(a), "r"(b));
            REP8(X)
#undef X
        } else if (MODE == 3) {  // 8 IMAD.WIDE only (reference for mode 2)
#define X(n) asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %2, %0;}" : "+l"(w##n) : "r"(a), "r"(b));
            REP8(X)
#undef X
        } else if (MODE == 4) {  // 16 DFMA + 8 IMAD.WIDE
#define X(n)                                                                                               \
    asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d##n) : "d"(da), "d"(db));                            \
    asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(e##n) : "d"(da), "d"(db));                            \
    asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %2, %0;}" : "+l"(w##n) : "r"(a), "r"(b));
            REP8(X)
#undef X
        } else if (MODE == 5) {  // integer Shoup product, 8 independent chains, twiddle changes every iteration
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
#define X(n) w##n = shoup_int(w##n, tw + n, tws + n, q);
            REP8(X)
#undef X
        } else if (MODE == 6) {  // hybrid Shoup product
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
#define X(n) w##n = shoup_fp(w##n, tw + n, tws + n, q);
            REP8(X)
#undef X
        } else if (MODE == 7) {  // hybrid, one twiddle shared by the 8 products (conversion of the twiddle amortised)
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
#define X(n) w##n = shoup_fp(w##n, tw, tws, q);
            REP8(X)
#undef X
        } else if (MODE == 8) {  // integer, one twiddle shared
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
#define X(n) w##n = shoup_int(w##n, tw, tws, q);
            REP8(X)
#undef X
        }
    }
    u64 t = w0 ^ w1 ^ w2 ^ w3 ^ w4 ^ w5 ^ w6 ^ w7;
    double s = d0 + d1 + d2 + d3 + d4 + d5 + d6 + d7 + e0 + e1 + e2 + e3 + e4 + e5 + e6 + e7;
    if (a0 == 0xdeadbeefu) out[threadIdx.x] = (u32)t ^ (u32)(t >> 32) ^ (u32)__double2loint(s);
}

// correctness: hybrid == integer product on pseudo-random inputs (x over the full 64-bit range)
__global__ void k_check(unsigned long long* bad, u64 q, u64 seed) {
    u64 s = seed + (blockIdx.x * (u64)blockDim.x + threadIdx.x) * 0x9e3779b97f4a7c15ull;
    unsigned long long nb = 0;
    for (int i = 0; i < 256; ++i) {
        s ^= s >> 30; s *= 0xbf58476d1ce4e5b9ull; s ^= s >> 27; s *= 0x94d049bb133111ebull; s ^= s >> 31;
        u64 x = s;
        s += 0x9e3779b97f4a7c15ull;
        u64 w = (s * 0xd6e8feb86659fd93ull) % q;
        u64 ws = (u64)((((unsigned __int128)w) << 64) / q);
        if (i & 1) x |= 0xffffffff00000000ull;
        if (i & 2) x |= 0x00000000ffffffffull;
        u64 r0 = shoup_int(x, w, ws, q), r1 = shoup_fp(x, w, ws, q);
        u64 ref = (u64)(((unsigned __int128)x * w) % q);
        if (r0 != r1 || r1 % q != ref || r1 >= 4 * q) ++nb;
    }
    if (nb) atomicAdd(bad, nb);
}

template <int MODE>
static void run(const char* name, double ops_per_iter) {
    u32* d;
    cudaMalloc(&d, 4);
    int blocks = 148 * 8;
    const u64 q = 1152921504606584833ull;  // 2^60 - 2^18 + 1 style 60-bit modulus (value irrelevant for timing)
    k<MODE><<<blocks, 256>>>(d, 3, 5, q);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, 3, 5, q);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double inst = (double)blocks * 256 * ITERS * ops_per_iter;
    int clk;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    double per_sm_clk = inst / (ms * 1e-3) / 148.0 / (clk * 1e3);
    printf("%-44s %8.3f ms  %7.2f ops/clk/SM (at %d MHz nominal)\n", name, ms, per_sm_clk, clk / 1000);
    cudaFree(d);
}

int main() {
    run<0>("DFMA x16", 16);
    run<1>("DADD x16", 16);
    run<3>("IMAD.WIDE x8", 8);
    run<2>("DFMA x8 + IMAD.WIDE x8 (count 8)", 8);
    run<4>("DFMA x16 + IMAD.WIDE x8 (count 8)", 8);
    run<5>("Shoup product, integer (products)", 8);
    run<6>("Shoup product, FP64 hybrid (products)", 8);
    run<8>("Shoup integer, shared twiddle", 8);
    run<7>("Shoup hybrid, shared twiddle", 8);
    unsigned long long* bad;
    cudaMalloc(&bad, 8);
    cudaMemset(bad, 0, 8);
    const u64 qs[3] = {1152921504606584833ull, 576460752303415297ull, 1099511922689ull};
    for (int i = 0; i < 3; ++i) k_check<<<1024, 256>>>(bad, qs[i], 12345 + i);
    unsigned long long hb = 0;
    cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost);
    printf("hybrid-vs-integer mismatches over %d products: %llu  (%s)\n", 3 * 1024 * 256 * 256, hb, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
