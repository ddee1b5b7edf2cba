// Microbenchmark 2: cost of each multiply FORM on the fmaheavy pipe of B200 (does a 64-bit addend make
// IMAD.WIDE dearer than a plain 32x32->64 product?), and register-resident throughput of candidate Shoup
// products, Cooley-Tukey butterflies and base-conversion MAC terms built from them.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipes2 int_pipes2.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned int u32;
typedef unsigned long long u64;
#define ITERS 1024
#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)

// ---- A: current product (modarith.cuh): 2 mul.hi + mad.wide(addend) + mul.wide + mad.wide(addend) + 4 mad.lo
__device__ __forceinline__ u64 shoup_a(u64 x, u64 w, u64 ws, u64 nq) {
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
// ---- B: no 64-bit addends on the multiply pipe: every wide product is a plain mul.wide, sums on the ALU
__device__ __forceinline__ u64 shoup_b(u64 x, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 x0, x1, w0, w1, s0, s1, n0, n1, a, b, p0, p1, h0, h1, t0, t1, u0, u1, r0, r1;\n\t"
        ".reg .u64 p, t, u;\n\t"
        "mov.b64 {x0, x1}, %1;\n\t"
        "mov.b64 {w0, w1}, %2;\n\t"
        "mov.b64 {s0, s1}, %3;\n\t"
        "mov.b64 {n0, n1}, %4;\n\t"
        "mul.hi.u32 a, x0, s1;\n\t"
        "mul.hi.u32 b, x1, s0;\n\t"
        "mul.wide.u32 p, x1, s1;\n\t"
        "mov.b64 {p0, p1}, p;\n\t"
        "add.cc.u32 h0, p0, a;\n\t"
        "addc.u32 h1, p1, 0;\n\t"
        "add.cc.u32 h0, h0, b;\n\t"
        "addc.u32 h1, h1, 0;\n\t"
        "mul.wide.u32 t, x0, w0;\n\t"
        "mul.wide.u32 u, h0, n0;\n\t"
        "mov.b64 {t0, t1}, t;\n\t"
        "mov.b64 {u0, u1}, u;\n\t"
        "add.cc.u32 r0, t0, u0;\n\t"
        "addc.u32 r1, t1, u1;\n\t"
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
// ---- C: as A but the four low products summed as two independent pairs (shorter dependent chain)
__device__ __forceinline__ u64 shoup_c(u64 x, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 x0, x1, w0, w1, s0, s1, n0, n1, a, b, m0, m1, h0, h1, r0, r1, v;\n\t"
        ".reg .u64 h, t;\n\t"
        "mov.b64 {x0, x1}, %1;\n\t"
        "mov.b64 {w0, w1}, %2;\n\t"
        "mov.b64 {s0, s1}, %3;\n\t"
        "mov.b64 {n0, n1}, %4;\n\t"
        "mul.hi.u32 a, x0, s1;\n\t"
        "mul.hi.u32 b, x1, s0;\n\t"
        "mul.lo.u32 v, x0, w1;\n\t"
        "mad.lo.u32 v, x1, w0, v;\n\t"
        "add.cc.u32 m0, a, b;\n\t"
        "addc.u32 m1, 0, 0;\n\t"
        "mov.b64 h, {m0, m1};\n\t"
        "mad.wide.u32 h, x1, s1, h;\n\t"
        "mov.b64 {h0, h1}, h;\n\t"
        "mul.wide.u32 t, x0, w0;\n\t"
        "mad.wide.u32 t, h0, n0, t;\n\t"
        "mov.b64 {r0, r1}, t;\n\t"
        "mad.lo.u32 v, h0, n1, v;\n\t"
        "mad.lo.u32 v, h1, n0, v;\n\t"
        "add.u32 r1, r1, v;\n\t"
        "mov.b64 %0, {r0, r1};\n\t"
        "}"
        : "=l"(r)
        : "l"(x), "l"(w), "l"(ws), "l"(nq));
    return r;
}

template <int V>
__device__ __forceinline__ u64 shoup(u64 x, u64 w, u64 ws, u64 nq) {
    return V == 0 ? shoup_a(x, w, ws, nq) : V == 1 ? shoup_b(x, w, ws, nq) : shoup_c(x, w, ws, nq);
}
template <int V>
__device__ __forceinline__ void ct_bfly(u64& X, u64& Y, u64 w, u64 ws, u64 nq, u64 q4) {
    const bool ge = X >= q4;
    const u64 s1 = ge ? q4 : 0ull, s2 = ge ? 0ull : q4;
    const u64 t = shoup<V>(Y, w, ws, nq);
    Y = X + s2 - t;
    X = X - s1 + t;
}
// butterfly with the conditional subtraction as a predicated subtract (fewer ALU instructions)
template <int V>
__device__ __forceinline__ void ct_bfly_p(u64& X, u64& Y, u64 w, u64 ws, u64 nq, u64 q4) {
    const u64 t = shoup<V>(Y, w, ws, nq);
    if (X >= q4) X -= q4;
    Y = X + q4 - t;
    X = X + t;
}

template <int MODE>
__global__ void __launch_bounds__(256) k(u32* out, u32 a0, u32 b0, u64 q) {
    u32 a = a0 + threadIdx.x, b = b0 | 1u;
    u32 r0 = a + 1, r1 = a + 2, r2 = a + 3, r3 = a + 4, r4 = a + 5, r5 = a + 6, r6 = a + 7, r7 = a + 8;
    u64 w0 = a + 1, w1 = a + 2, w2 = a + 3, w3 = a + 4, w4 = a + 5, w5 = a + 6, w6 = a + 7, w7 = a + 8;
    u64 v0 = a + 11, v1 = a + 12, v2 = a + 13, v3 = a + 14, v4 = a + 15, v5 = a + 16, v6 = a + 17, v7 = a + 18;
    u64 tw = 0x0123456789abcdefull ^ a, tws = 0xfedcba9876543210ull ^ b;
    const u64 nq = 0ull - q, q4 = 4 * q;
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
        if (MODE == 0) {  // mul.wide.u32 (no addend), dependent through the low word
#define X(n) asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mul.wide.u32 %0, lo, %2;}" : "+l"(w##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 1) {  // mad.wide.u32 with 64-bit addend
#define X(n) asm volatile("{.reg .u32 lo, hi; mov.b64 {lo,hi}, %0; mad.wide.u32 %0, lo, %2, %0;}" : "+l"(w##n) : "r"(a), "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 2) {  // mul.hi.u32 (no addend)
#define X(n) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(r##n) : "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 3) {  // mad.hi.u32 (32-bit addend)
#define X(n) asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(r##n) : "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 4) {  // mul.lo.u32
#define X(n) asm volatile("mul.lo.u32 %0, %0, %1;" : "+r"(r##n) : "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE == 5) {  // mul.wide + add.cc/addc into a separate accumulator (MAC with ALU adds)
#define X(n)                                                                                                  \
    asm volatile("{.reg .u32 lo, hi, p0, p1, c0, c1; .reg .u64 p; mov.b64 {lo,hi}, %0; mul.wide.u32 p, lo, %2;\n\t" \
                 "mov.b64 {p0,p1}, p; mov.b64 {c0,c1}, %1; add.cc.u32 c0, c0, p0; addc.u32 c1, c1, p1;\n\t"       \
                 "mov.b64 %1, {c0,c1}; xor.b32 lo, lo, p0; mov.b64 %0, {lo,hi};}"                               \
                 : "+l"(w##n), "+l"(v##n) : "r"(b));
            REP8(X) REP8(X)
#undef X
        } else if (MODE >= 10 && MODE < 13) {  // Shoup product variants, twiddle shared by the 8 products
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
#define X(n) w##n = shoup<MODE - 10>(w##n, tw, tws, nq);
            REP8(X)
#undef X
        } else if (MODE >= 20 && MODE < 23) {  // CT butterflies (4 per step, radix-8 group = 3 steps) select form
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
            ct_bfly<MODE - 20>(w0, w4, tw, tws, nq, q4);
            ct_bfly<MODE - 20>(w1, w5, tw, tws, nq, q4);
            ct_bfly<MODE - 20>(w2, w6, tw, tws, nq, q4);
            ct_bfly<MODE - 20>(w3, w7, tw, tws, nq, q4);
            ct_bfly<MODE - 20>(w0, w2, tw + 1, tws, nq, q4);
            ct_bfly<MODE - 20>(w1, w3, tw + 1, tws, nq, q4);
            ct_bfly<MODE - 20>(w4, w6, tw + 2, tws, nq, q4);
            ct_bfly<MODE - 20>(w5, w7, tw + 2, tws, nq, q4);
        } else if (MODE >= 30 && MODE < 33) {  // predicated form
            tw += 0x9e3779b97f4a7c15ull;
            tws ^= tw;
            ct_bfly_p<MODE - 30>(w0, w4, tw, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w1, w5, tw, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w2, w6, tw, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w3, w7, tw, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w0, w2, tw + 1, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w1, w3, tw + 1, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w4, w6, tw + 2, tws, nq, q4);
            ct_bfly_p<MODE - 30>(w5, w7, tw + 2, tws, nq, q4);
        }
    }
    u32 s = r0 ^ r1 ^ r2 ^ r3 ^ r4 ^ r5 ^ r6 ^ r7;
    u64 t = w0 ^ w1 ^ w2 ^ w3 ^ w4 ^ w5 ^ w6 ^ w7 ^ v0 ^ v1 ^ v2 ^ v3 ^ v4 ^ v5 ^ v6 ^ v7;
    if (a0 == 0xdeadbeefu) out[threadIdx.x] = s ^ (u32)t ^ (u32)(t >> 32);
}

__global__ void k_check(unsigned long long* bad, u64 q, u64 seed) {
    u64 s = seed + (blockIdx.x * (u64)blockDim.x + threadIdx.x) * 0x9e3779b97f4a7c15ull;
    unsigned long long nb = 0;
    const u64 nq = 0ull - q;
    for (int i = 0; i < 256; ++i) {
        s ^= s >> 30; s *= 0xbf58476d1ce4e5b9ull; s ^= s >> 27; s *= 0x94d049bb133111ebull; s ^= s >> 31;
        u64 x = s;
        s += 0x9e3779b97f4a7c15ull;
        u64 w = (s * 0xd6e8feb86659fd93ull) % q;
        u64 ws = (u64)((((unsigned __int128)w) << 64) / q);
        if (i & 1) x |= 0xffffffff00000000ull;
        if (i & 2) x |= 0x00000000ffffffffull;
        u64 ra = shoup_a(x, w, ws, nq), rb = shoup_b(x, w, ws, nq), rc = shoup_c(x, w, ws, nq);
        if (ra != rb || ra != rc) ++nb;
    }
    if (nb) atomicAdd(bad, nb);
}

template <int MODE>
static void run(const char* name, double ops_per_iter) {
    u32* d;
    cudaMalloc(&d, 4);
    int blocks = 148 * 8;
    const u64 q = 1152921504606584833ull;
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
    printf("%-44s %8.3f ms  %7.2f ops/clk/SM  = %6.2f clk per warp-op per SMSP\n", name, ms, per_sm_clk, 32.0 / per_sm_clk);
    cudaFree(d);
}

int main() {
    run<0>("mul.wide.u32", 16);
    run<1>("mad.wide.u32 (64-bit addend)", 16);
    run<2>("mul.hi.u32", 16);
    run<3>("mad.hi.u32", 16);
    run<4>("mul.lo.u32", 16);
    run<5>("mul.wide + add.cc/addc", 16);
    run<10>("Shoup A (current)", 8);
    run<11>("Shoup B (no wide addends)", 8);
    run<12>("Shoup C (split low chain)", 8);
    run<20>("CT butterfly, select form, Shoup A", 8);
    run<21>("CT butterfly, select form, Shoup B", 8);
    run<22>("CT butterfly, select form, Shoup C", 8);
    run<30>("CT butterfly, predicated, Shoup A", 8);
    run<31>("CT butterfly, predicated, Shoup B", 8);
    run<32>("CT butterfly, predicated, Shoup C", 8);
    unsigned long long* bad;
    cudaMalloc(&bad, 8);
    cudaMemset(bad, 0, 8);
    const u64 qs[3] = {1152921504606584833ull, 576460752303415297ull, 1099511922689ull};
    for (int i = 0; i < 3; ++i) k_check<<<1024, 256>>>(bad, qs[i], 12345 + i);
    unsigned long long hb = 0;
    cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost);
    printf("variant mismatches: %llu (%s)\n", hb, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
