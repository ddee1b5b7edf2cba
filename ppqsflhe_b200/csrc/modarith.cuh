// modarith.cuh -- 64-bit modular arithmetic for sm_100a (device side).
//
// B200 has no 64-bit integer multiplier: every 64x64 product is built from 32-bit IMAD / IMAD.WIDE on the
// FMA pipe, so the multiply count per butterfly is THE cost driver ("integer-pipe co-bound").
// Conventions:
//   * q < 2^60 for every limb (OpenFHE 60-bit NATIVEINT limbs), so 4q and even 8q fit in a u64.
//   * "lazy" values live in [0,2q) or [0,4q); every value written to HBM that crosses the C-ABI is canonical.
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

__device__ __forceinline__ u64 csub(u64 x, u64 q) { return x >= q ? x - q : x; }

// x * w mod q in [0, 2q) for any x < 2^64 (Shoup / Harvey).
__device__ __forceinline__ u64 mul_shoup_lazy(u64 x, u64 w, u64 ws, u64 q) {
    u64 h = __umul64hi(x, ws);
    return x * w - h * q;
}
__device__ __forceinline__ u64 mul_shoup(u64 x, u64 w, u64 ws, u64 q) { return csub(mul_shoup_lazy(x, w, ws, q), q); }

// 128-bit accumulator: (hi,lo) += a * b
__device__ __forceinline__ void mac128(u64& hi, u64& lo, u64 a, u64 b) {
    u64 pl = a * b;
    u64 ph = __umul64hi(a, b);
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(lo), "+l"(hi) : "l"(pl), "l"(ph));
}

// (hi,lo) mod q, canonical.  Requires (hi,lo) < 2^(63 + bitlen(q)).
__device__ __forceinline__ u64 barrett128(u64 hi, u64 lo, const LimbC& c) {
    u64 x1 = (hi << (64u - c.sh)) | (lo >> c.sh);
    u64 qh = __umul64hi(x1, c.mu);
    u64 r = lo - qh * c.q;  // true remainder + {0,1,2} * q
    r = csub(r, c.q2);
    return csub(r, c.q);
}

// a * b mod q, canonical, for a < 2^63, b < q (no precomputed companion).
__device__ __forceinline__ u64 mulmod_barrett(u64 a, u64 b, const LimbC& c) {
    return barrett128(__umul64hi(a, b), a * b, c);
}

__device__ __forceinline__ u64 addmod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
__device__ __forceinline__ u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }

// x mod q for an arbitrary 64-bit x (used for signed small -> residue and SwitchModulus).
__device__ __forceinline__ u64 reduce64(u64 x, const LimbC& c) { return barrett128(0ull, x, c); }

}  // namespace ppq
