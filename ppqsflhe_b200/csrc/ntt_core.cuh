// ntt_core.cuh -- register/shared-memory tile transforms that every NTT-type kernel is built from.
//
// Negacyclic NTT over Z_q[x]/(x^N+1), q < 2^60, EXACTLY the transform OpenFHE's
// ChineseRemainderTransformFTT computes (ForwardTransformToBitReverseInPlace: Cooley-Tukey, natural in,
// bit-reversed out, twiddles psi^brev(k); InverseTransformFromBitReverseInPlace: Gentleman-Sande), which the
// reference reaches from cc->ReEncrypt / EvalMult / Encrypt / Decrypt.
//
// B200 mapping.  One limb of N = 2^16 coefficients is 512 KiB and does not fit one SM, so a transform is
// two HBM/L2 passes over a (n1 x n2) view of the limb, N = n1*n2, element (row, col) at row*n2 + col:
//   forward  = COL pass (first  log n1 stages: n2 independent size-n1 transforms over rows, stride n2;
//                        twiddles psi_br[m + i], m < n1, identical for every column)
//            + ROW pass (last   log n2 stages: n1 independent size-n2 transforms over contiguous rows;
//                        row r uses twiddles psi_br[m'(n1 + r) + i'], m' < n2)
//   inverse  = ROW pass + COL pass with the inverse table (same index law).
// Inside a pass every thread owns 8 points of one size-2^S transform and runs radix-8 "groups" of three
// stages entirely in registers; between groups the 8 points are exchanged through shared memory
// (padded so both the store and the load side are bank-conflict free).  Prologue / epilogue work
// (base conversion, key inner product, ModDown combine, accumulate) is fused by the kernels around these
// functions, so a pass is the only time a coefficient travels through the memory system.
//
// Point <-> (thread, register) law.  u = thread index inside the transform (S-3 bits), k = register (3 bits).
// A group with register bits [LO, LO+3):  p = ((u >> LO) << (LO+3)) | (k << LO) | (u & (2^LO - 1)).
// Forward groups: LO = S-3, S-6, ..., last group LO = 0 (processing only the bits not done yet).
// Inverse groups: LO = 0, 3, ..., last group LO = S-3.
// Twiddle for the stage on point-bit b, pair containing point p:  tw[(rowfac << (S-1-b)) + (p >> (b+1))],
// rowfac = 1 (COL pass) or n1 + row (ROW pass).
#pragma once
#include "modarith.cuh"

namespace ppq {

// ---- lazy butterflies (Harvey, with the 3-multiply Shoup product whose result lies in [0,4q)).
// Forward keeps values in [0,8q); inverse keeps them in [0,4q).  q4 = 4q.
__device__ __forceinline__ void ct_bfly(u64& X, u64& Y, ulonglong2 tw, u64 q, u64 q4) {
    // conditional subtraction folded into three-input adds (IADD3 on the ALU pipe; a two-input 64-bit add is
    // otherwise placed on the fmaheavy pipe as IMAD.X, which is the binding pipe)
    const bool ge = X >= q4;
    const u64 s1 = ge ? q4 : 0ull, s2 = ge ? 0ull : q4;
    const u64 t = mul_shoup_lazy4(Y, tw.x, tw.y, q);
    Y = X + s2 - t;
    X = X - s1 + t;
}
__device__ __forceinline__ void gs_bfly(u64& X, u64& Y, ulonglong2 tw, u64 q, u64 q4) {
    u64 s = X + Y;
    u64 d = X - Y + q4;
    X = csub(s, q4);
    Y = mul_shoup_lazy4(d, tw.x, tw.y, q);
}

template <int LO>
__device__ __forceinline__ u32 point_of(u32 u, int k) {
    return ((u >> LO) << (LO + 3)) | ((u32)k << LO) | (u & ((1u << LO) - 1u));
}

// Forward stages on point bits BHI-1 .. LO (register bits BHI-1-LO .. 0).
template <int S, int LO, int BHI>
__device__ __forceinline__ void fwd_group(u64 (&a)[8], const ulonglong2* __restrict__ tw, u32 rowfac, u32 u, u64 q,
                                          u64 q4) {
    const u32 uhi = u >> LO;
#pragma unroll
    for (int rb = BHI - 1 - LO; rb >= 0; --rb) {
        const u32 base = rowfac << (S - 1 - LO - rb);
#pragma unroll
        for (int g = 0; g < (4 >> rb); ++g) {
#ifdef PPQ_EXP_NOTW
            const ulonglong2 t = make_ulonglong2(q ^ (u64)(base + uhi + g), q4 ^ (u64)(uhi * 77 + g));
#else
            const ulonglong2 t = __ldg(tw + (base + ((uhi << (2 - rb)) | (u32)g)));
#endif
#pragma unroll
            for (int lk = 0; lk < (1 << rb); ++lk) {
                const int k = (g << (rb + 1)) | lk;
                ct_bfly(a[k], a[k + (1 << rb)], t, q, q4);
            }
        }
    }
}

// Inverse stages on point bits BLO .. LO+2 (register bits BLO-LO .. 2), ascending.
template <int S, int LO, int BLO>
__device__ __forceinline__ void inv_group(u64 (&a)[8], const ulonglong2* __restrict__ tw, u32 rowfac, u32 u, u64 q,
                                          u64 q4) {
    const u32 uhi = u >> LO;
#pragma unroll
    for (int rb = BLO - LO; rb <= 2; ++rb) {
        const u32 base = rowfac << (S - 1 - LO - rb);
#pragma unroll
        for (int g = 0; g < (4 >> rb); ++g) {
#ifdef PPQ_EXP_NOTW
            const ulonglong2 t = make_ulonglong2(q ^ (u64)(base + uhi + g), q4 ^ (u64)(uhi * 77 + g));
#else
            const ulonglong2 t = __ldg(tw + (base + ((uhi << (2 - rb)) | (u32)g)));
#endif
#pragma unroll
            for (int lk = 0; lk < (1 << rb); ++lk) {
                const int k = (g << (rb + 1)) | lk;
                gs_bfly(a[k], a[k + (1 << rb)], t, q, q4);
            }
        }
    }
}

// Shared-memory position policies.  ROW tiles: one transform is contiguous, padded by one word per 8
// (stride-8 and stride-32 register patterns both become conflict free).  COL tiles: [point][column].
template <int S>
struct RowPos {
    static constexpr int PITCH = (1 << S) + (1 << (S - 3));
    u32 base;  // w * PITCH
    __device__ __forceinline__ u32 operator()(u32 p) const { return base + p + (p >> 3); }
};
template <int W>
struct ColPos {
    u32 w;
    __device__ __forceinline__ u32 operator()(u32 p) const { return p * W + w; }
};

template <int LO, class Pos>
__device__ __forceinline__ void sm_store(const u64 (&a)[8], u64* sm, const Pos& pos, u32 u) {
#pragma unroll
    for (int k = 0; k < 8; ++k) sm[pos(point_of<LO>(u, k))] = a[k];
}
template <int LO, class Pos>
__device__ __forceinline__ void sm_load(u64 (&a)[8], const u64* sm, const Pos& pos, u32 u) {
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = sm[pos(point_of<LO>(u, k))];
}

__host__ __device__ constexpr int cmax(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int cmin(int a, int b) { return a < b ? a : b; }

// Forward pass over one size-2^S transform.  In: a[k] = point point_of<S-3>(u,k), values in [0,8q).
// Out: a[k] = point point_of<0>(u,k) = 8u + k, values in [0,8q).  `guard` inserts a leading barrier so the
// shared buffer can be reused across loop iterations (write-after-read protection).
template <int S, class Pos>
__device__ __forceinline__ void fwd_pass(u64 (&a)[8], u64* sm, const Pos& pos, u32 u, const ulonglong2* __restrict__ tw,
                                         u32 rowfac, u64 q, u64 q4, bool guard) {
    constexpr int LO0 = S - 3;
    constexpr int LO1 = cmax(S - 6, 0);
    fwd_group<S, LO0, S>(a, tw, rowfac, u, q, q4);
    if (guard) __syncthreads();
    sm_store<LO0>(a, sm, pos, u);
    __syncthreads();
    sm_load<LO1>(a, sm, pos, u);
    fwd_group<S, LO1, LO0>(a, tw, rowfac, u, q, q4);
    if constexpr (S > 6) {
        sm_store<LO1>(a, sm, pos, u);  // same words this thread just read: no barrier needed before
        __syncthreads();
        sm_load<0>(a, sm, pos, u);
        fwd_group<S, 0, LO1>(a, tw, rowfac, u, q, q4);
    }
}

// Inverse pass.  In: a[k] = point 8u + k, values in [0,4q).  Out: a[k] = point_of<S-3>(u,k), in [0,4q).
template <int S, class Pos>
__device__ __forceinline__ void inv_pass(u64 (&a)[8], u64* sm, const Pos& pos, u32 u, const ulonglong2* __restrict__ tw,
                                         u32 rowfac, u64 q, u64 q4, bool guard) {
    constexpr int LOL = S - 3;           // last group
    constexpr int LO1 = cmin(3, S - 3);  // second group
    inv_group<S, 0, 0>(a, tw, rowfac, u, q, q4);
    if (guard) __syncthreads();
    sm_store<0>(a, sm, pos, u);
    __syncthreads();
    sm_load<LO1>(a, sm, pos, u);
    inv_group<S, LO1, 3>(a, tw, rowfac, u, q, q4);
    if constexpr (S > 6) {
        sm_store<LO1>(a, sm, pos, u);
        __syncthreads();
        sm_load<LOL>(a, sm, pos, u);
        inv_group<S, LOL, 6>(a, tw, rowfac, u, q, q4);
    }
}

}  // namespace ppq
