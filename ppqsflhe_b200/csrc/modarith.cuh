// modarith.cuh -- 64-bit modular arithmetic for sm_100a (device side).
//
// B200 has no 64-bit integer multiplier: every 64x64 product is built from 32-bit IMAD / IMAD.WIDE on the
// FMA pipe, so the multiply count per butterfly is THE cost driver ("integer-pipe co-bound").
// Conventions:
//   * q < 2^60 for every limb (OpenFHE 60-bit NATIVEINT limbs), so 4q and even 8q fit in a u64.
//   * "lazy" values live in [0,4q) (inverse passes) or [0,8q) (forward passes); 8q < 2^63.  Every value that
//     crosses the C-ABI is canonical.
//   * Shoup multiplication by a precomputed constant w needs ws = floor(w * 2^64 / q) and accepts ANY 64-bit x.
//   * Barrett reduction of a 128-bit accumulator needs mu = floor(2^(64+sh) / q), sh = bitlen(q) - 1,
//     and is valid for x < 2^(63 + bitlen(q)) (e.g. a sum of up to 8 products y*c with y < 2^60, c < q).
// Replaces (restates for the GPU) OpenFHE NativeIntegerT::ModMulFastConst / ModMulFast and the
// unsigned __int128 accumulate + BarrettUint128ModUint64 of DCRTPolyImpl::ApproxSwitchCRTBasis, which the
// reference reaches through cc->ReEncrypt (server/src/changeCipherDomain.cpp:74,89,105).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ppq {

typedef unsigned long long u64;
typedef unsigned int u32;

// Per-limb constants, one 32-byte record per modulus of Q u P (device array, read through the RO path).
struct LimbC {
    u64 q;    // modulus
    u64 q2;   // 2q
    u64 mu;   // floor(2^(64+sh)/q)
    u32 sh;   // bitlen(q) - 1
    u32 pad;
};

// Conditional subtraction for q <= 2^63 and x < 2^63 + q (every caller: x < 16q, q < 2^60): the SIGN of x - q decides,
// so the compare is one 32-bit ISETP on the high word instead of a two-instruction 64-bit compare
// (5 instructions instead of 6; the NTT kernels' time follows their instruction count, profiles/r01_notes.md).
__device__ __forceinline__ u64 csub(u64 x, u64 q) {
#ifdef PPQ_PLAIN_CSUB
    return x >= q ? x - q : x;
#else
    const u64 d = x - q;
    return (int)(d >> 32) < 0 ? x : d;
#endif
}

// x * w mod q in [0, 2q) for any x < 2^64 (Shoup / Harvey).
__device__ __forceinline__ u64 mul_shoup_lazy(u64 x, u64 w, u64 ws, u64 q) {
    u64 h = __umul64hi(x, ws);
    return x * w - h * q;
}
__device__ __forceinline__ u64 mul_shoup(u64 x, u64 w, u64 ws, u64 q) { return csub(mul_shoup_lazy(x, w, ws, q), q); }

// x * w mod q in [0, 4q) for any x < 2^64, with a 3-multiply approximation of the high product:
// h' = x1*w1 + hi32(x0*w1) + hi32(x1*w0) underestimates floor(x*ws / 2^64) by at most 2 (the dropped terms are
// lo32(x0*w1) + lo32(x1*w0) < 2^33 units of 2^32 and x0*w0 < 2^64), so the remainder grows by at most 2q.
// 3 IMAD.WIDE + 2 adds instead of 4 IMAD.WIDE + a 4-deep carry chain: the butterflies use this one.
// Written as PTX so that the instruction selection is the minimal one measured on B200 (tools/ubench/int_pipes.cu:
// IMAD 2 clk, IMAD.WIDE / IMAD.HI ~5 clk per warp and SMSP, all on the fmaheavy pipe, which is the binding pipe of
// every NTT kernel): 5 wide/hi multiplies + 4 low multiplies + 2 adds, x*w - h*q evaluated as x*w + h*(2^64 - q) so
// that the products chain through the multiply-add addend instead of separate negate/add instructions.
__device__ __forceinline__ u64 mul_shoup_lazy4(u64 x, u64 w, u64 ws, u64 q) {
#ifdef PPQ_PLAIN_SHOUP
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32), s0 = (u32)ws, s1 = (u32)(ws >> 32);
    const u64 p01 = (u64)x0 * s1, p10 = (u64)x1 * s0;
    const u64 mid = (p01 >> 32) + (p10 >> 32);
    const u64 h = (u64)x1 * s1 + mid;
    return x * w - h * q;
#else
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
#endif
}

// 128-bit accumulator: (hi,lo) += a * b   (4 IMAD.WIDE + carry adds; nvcc's __int128 lowering is the tightest)
__device__ __forceinline__ void mac128(u64& hi, u64& lo, u64 a, u64 b) {
    unsigned __int128 acc = ((unsigned __int128)hi << 64) | lo;
    acc += (unsigned __int128)a * b;
    hi = (u64)(acc >> 64);
    lo = (u64)acc;
}

// (hi,lo) mod q, canonical.  Requires (hi,lo) < 2^(63 + bitlen(q)).
__device__ __forceinline__ u64 barrett128(u64 hi, u64 lo, const LimbC& c) {
    // x1 = bits [sh, sh+64) of (hi:lo).  For sh >= 32 (every limb of 33 bits and more; the branch is block-uniform)
    // two funnel shifts do it; the generic 128-bit shift costs seven instructions.
    u64 x1;
    if (c.sh >= 32u) {
        const u32 s = c.sh - 32u;
        const u32 l1 = (u32)(lo >> 32), h0 = (u32)hi, h1 = (u32)(hi >> 32);
        x1 = ((u64)__funnelshift_r(h0, h1, s) << 32) | __funnelshift_r(l1, h0, s);
    } else {
        x1 = (hi << (64u - c.sh)) | (lo >> c.sh);
    }
    u64 qh = __umul64hi(x1, c.mu);
    u64 r = lo - qh * c.q;  // true remainder + {0,1,2} * q
    r = csub(r, c.q2);
    return csub(r, c.q);
}

// The same without the two conditional subtractions: (hi,lo) mod q + {0,1,2} q, i.e. a value in [0, 3q).  For results
// that feed a forward COL pass (which accepts < 4q): ten instructions less per reduction.
__device__ __forceinline__ u64 barrett128_lazy3(u64 hi, u64 lo, const LimbC& c) {
    u64 x1;
    if (c.sh >= 32u) {
        const u32 s = c.sh - 32u;
        const u32 l1 = (u32)(lo >> 32), h0 = (u32)hi, h1 = (u32)(hi >> 32);
        x1 = ((u64)__funnelshift_r(h0, h1, s) << 32) | __funnelshift_r(l1, h0, s);
    } else {
        x1 = (hi << (64u - c.sh)) | (lo >> c.sh);
    }
    return lo - __umul64hi(x1, c.mu) * c.q;
}

// a * b mod q, canonical, for a < 2^63, b < q (no precomputed companion).
__device__ __forceinline__ u64 mulmod_barrett(u64 a, u64 b, const LimbC& c) {
    return barrett128(__umul64hi(a, b), a * b, c);
}

// [0, 8q) -> [0, q)
__device__ __forceinline__ u64 canon8(u64 x, const LimbC& c) { return csub(csub(csub(x, c.q2 << 1), c.q2), c.q); }
// [0, 4q) -> [0, q)
__device__ __forceinline__ u64 canon4(u64 x, const LimbC& c) { return csub(csub(x, c.q2), c.q); }

// [0, 16q) -> [0, q) with ONE quotient estimate instead of a chain of four conditional subtractions (12 instructions
// instead of 20).  Needs bitlen(q) >= 29.  x / q = (x >> (sh-27)) * floor(2^(32+sh)/q) / 2^59; the estimate is exact or
// one short (both truncations together lose < 2^-26), so the remainder lies in [0, 2q).  floor(2^(32+sh)/q) is the high
// word of the Barrett constant mu.
__device__ __forceinline__ u64 canon16_q(u64 x, const LimbC& c) {
    const u32 xs = (u32)(x >> (c.sh - 27u));
    const u32 k = __umulhi(xs, (u32)(c.mu >> 32)) >> 27;
    return csub(x - (u64)k * c.q, c.q);
}

// the same without the final conditional subtraction: [0, 16q) -> [0, 2q)
__device__ __forceinline__ u64 canon16_q_lazy2(u64 x, const LimbC& c) {
    const u32 xs = (u32)(x >> (c.sh - 27u));
    const u32 k = __umulhi(xs, (u32)(c.mu >> 32)) >> 27;
    return x - (u64)k * c.q;
}

__device__ __forceinline__ u64 addmod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
__device__ __forceinline__ u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }

// x mod q for an arbitrary 64-bit x (used for signed small -> residue and SwitchModulus).
__device__ __forceinline__ u64 reduce64(u64 x, const LimbC& c) { return barrett128(0ull, x, c); }

}  // namespace ppq
