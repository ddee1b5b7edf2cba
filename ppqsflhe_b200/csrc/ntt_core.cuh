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

#ifndef PPQ_BFLY_V
#define PPQ_BFLY_V 2  // 2 = statically tracked lazy bounds (default); 0 = fold on every butterfly (round-1 form)
#endif

namespace ppq {

// ---- lazy butterflies (Harvey, with the 3-multiply Shoup product whose result lies in [0,4q)), round-1 form
// (PPQ_BFLY_V == 0): one fold per butterfly, forward values in [0,8q), inverse in [0,4q).  q4 = 4q.
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

// ---- forward butterflies with statically tracked bounds (PPQ_BFLY_V == 2).  q < 2^60, so 16q <= 2^64: a value may
// grow to < 16q before it has to be folded.  One CT stage maps a uniform bound b (units of q) to b + 4
// (X' = X + t, Y' = X - t + 4q with t < 4q; Y itself is only ever multiplied, and the Shoup product accepts any
// 64-bit input), so a conditional subtraction of 8q is needed only on stages entered with b > 12: about every
// second stage instead of every stage.
__host__ __device__ constexpr int fwd_bound_after(int b, int stages) {
    for (int i = 0; i < stages; ++i) {
        if (b + 4 > 16) b = 8;
        b += 4;
    }
    return b;
}
template <bool FOLD>
__device__ __forceinline__ void ct_bfly_lazy(u64& X, u64& Y, ulonglong2 tw, u64 q, u64 q4, u64 z) {
    const u64 t = mul_shoup_lazy4(Y, tw.x, tw.y, q);
    if (FOLD) X = csub(X, q4 << 1);
    Y = X + q4 - t;
    // z is a run-time zero (LimbC::pad): a THREE-input add cannot be turned into IMAD.X, which ptxas otherwise
    // places on the fmaheavy pipe -- the binding pipe of every NTT kernel (tools/sass_cost.py)
    X = X + t + z;
}

// ---- inverse (Gentleman-Sande) butterflies with statically tracked per-register bounds.  Invariant: every
// register is < 8q when a butterfly reads it, so X + Y and X - Y + 8q stay below 16q <= 2^64.  The difference goes
// through the Shoup product (< 4q); the sum is folded by 8q only if its bound exceeds 8q.  Registers that hold
// products (bound 4) pair with each other in the next stage (the two inputs of a GS butterfly always have the same
// history), so about a third of the folds of a radix-8 group disappear, and nearly all of them in the first group
// of a pass that reads canonical data.
struct InvPlan {
    unsigned fold;  // bit (8*stage + k): fold the sum written to register k in that stage
    int out;        // bound (units of q) of every register after the group
};
__host__ __device__ constexpr InvPlan inv_plan(int bin, int first_rb) {
    int b[8] = {bin, bin, bin, bin, bin, bin, bin, bin};
    unsigned fold = 0;
    int st = 0;
    for (int rb = first_rb; rb <= 2; ++rb, ++st)
        for (int k = 0; k < 8; ++k)
            if (!(k & (1 << rb))) {
                const int j = k | (1 << rb);
                const int sum = b[k] + b[j];
                if (sum > 8) fold |= 1u << (8 * st + k);
                b[k] = sum > 8 ? 8 : sum;
                b[j] = 4;
            }
    int out = 0;
    for (int k = 0; k < 8; ++k) out = b[k] > out ? b[k] : out;
    return InvPlan{fold, out};
}
template <bool FOLD>
__device__ __forceinline__ void gs_bfly_lazy(u64& X, u64& Y, ulonglong2 tw, u64 q, u64 q8) {
    u64 s = X + Y;
    const u64 d = X - Y + q8;
    if (FOLD) s = csub(s, q8);
    X = s;
    Y = mul_shoup_lazy4(d, tw.x, tw.y, q);
}

template <int LO>
__device__ __forceinline__ u32 point_of(u32 u, int k) {
    return ((u >> LO) << (LO + 3)) | ((u32)k << LO) | (u & ((1u << LO) - 1u));
}

// Forward stages on point bits BHI-1 .. LO (register bits BHI-1-LO .. 0).
// The (up to) seven twiddles of a radix-8 group, slot (4 >> rb) - 1 + g.  Loading them is separate from using them so
// that a pass can request the NEXT group's twiddles before the shared-memory exchange: their L2 latency (the top
// stall of the ROW passes, profiles/r04_ncu_full.md) then overlaps the exchange instead of following it.
template <int S, int LO, int BHI>
__device__ __forceinline__ void fwd_group_tw(ulonglong2 (&t)[7], const ulonglong2* __restrict__ tw, u32 rowfac, u32 u) {
    const u32 uhi = u >> LO;
#pragma unroll
    for (int rb = BHI - 1 - LO; rb >= 0; --rb) {
        const u32 base = rowfac << (S - 1 - LO - rb);
#pragma unroll
        for (int g = 0; g < (4 >> rb); ++g) t[(4 >> rb) - 1 + g] = __ldg(tw + (base + ((uhi << (2 - rb)) | (u32)g)));
    }
}
template <int S, int LO, int BLO>
__device__ __forceinline__ void inv_group_tw(ulonglong2 (&t)[7], const ulonglong2* __restrict__ tw, u32 rowfac, u32 u) {
    const u32 uhi = u >> LO;
#pragma unroll
    for (int rb = BLO - LO; rb <= 2; ++rb) {
        const u32 base = rowfac << (S - 1 - LO - rb);
#pragma unroll
        for (int g = 0; g < (4 >> rb); ++g) t[(4 >> rb) - 1 + g] = __ldg(tw + (base + ((uhi << (2 - rb)) | (u32)g)));
    }
}

template <int S, int LO, int BHI, int B = 8>
__device__ __forceinline__ void fwd_group(u64 (&a)[8], const ulonglong2 (&tws)[7], u64 q, u64 q4, u64 z = 0) {
#pragma unroll
    for (int rb = BHI - 1 - LO; rb >= 0; --rb) {
        const bool fold = fwd_bound_after(B, BHI - 1 - LO - rb) + 4 > 16;  // folds at compile time (unrolled)
#pragma unroll
        for (int g = 0; g < (4 >> rb); ++g) {
            const ulonglong2 t = tws[(4 >> rb) - 1 + g];
#pragma unroll
            for (int lk = 0; lk < (1 << rb); ++lk) {
                const int k = (g << (rb + 1)) | lk;
#if PPQ_BFLY_V == 2
                if (fold) ct_bfly_lazy<true>(a[k], a[k + (1 << rb)], t, q, q4, z);
                else ct_bfly_lazy<false>(a[k], a[k + (1 << rb)], t, q, q4, z);
#else
                (void)fold;
                ct_bfly(a[k], a[k + (1 << rb)], t, q, q4);
#endif
            }
        }
    }
}

// Inverse stages on point bits BLO .. LO+2 (register bits BLO-LO .. 2), ascending.
// FIN: this group ends the whole transform (COL pass, LO = S-3) and its last stage -- ONE twiddle, psi^-brev(1), for
// every butterfly -- also applies the per-limb output factor f (N^-1 times the base-conversion pre-factor): the
// difference is multiplied by fin_w = twiddle * f instead of the twiddle (free), the sum by fin_s = f: four extra lazy
// products per thread instead of eight exact ones in an epilogue.  Outputs are < 4q.
template <int S, int LO, int BLO, int B = 4, bool FIN = false>
__device__ __forceinline__ void inv_group(u64 (&a)[8], const ulonglong2 (&tws)[7], u64 q, u64 q4,
                                          ulonglong2 fin_s = ulonglong2{0, 0}, ulonglong2 fin_w = ulonglong2{0, 0}) {
    constexpr InvPlan plan = inv_plan(B, BLO - LO);
    (void)plan;
#pragma unroll
    for (int rb = BLO - LO; rb <= 2; ++rb) {
        if (FIN && rb == 2) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const u64 sum = a[k] + a[k + 4], dif = a[k] - a[k + 4] + (q4 << 1);
                a[k] = mul_shoup_lazy4(sum, fin_s.x, fin_s.y, q);
                a[k + 4] = mul_shoup_lazy4(dif, fin_w.x, fin_w.y, q);
            }
            continue;
        }
#pragma unroll
        for (int g = 0; g < (4 >> rb); ++g) {
            const ulonglong2 t = tws[(4 >> rb) - 1 + g];
#pragma unroll
            for (int lk = 0; lk < (1 << rb); ++lk) {
                const int k = (g << (rb + 1)) | lk;
#if PPQ_BFLY_V == 2
                if ((plan.fold >> (8 * (rb - (BLO - LO)) + k)) & 1u)
                    gs_bfly_lazy<true>(a[k], a[k + (1 << rb)], t, q, q4 << 1);
                else
                    gs_bfly_lazy<false>(a[k], a[k + (1 << rb)], t, q, q4 << 1);
#else
                gs_bfly(a[k], a[k + (1 << rb)], t, q, q4);
#endif
            }
        }
    }
}

// Shared-memory position policies.  ROW tiles: one transform is contiguous, padded by one word per 8
// (stride-8 and stride-32 register patterns both become conflict free).  COL tiles: [point][column].
template <int S>
struct RowPos {
    // the 2^(S-3) threads of one row transform are consecutive lanes of ONE warp for S <= 8: the exchanges between
    // radix-8 groups then need __syncwarp only, and the warps of a CTA never wait for each other
    static constexpr bool WARP_LOCAL = (S - 3) <= 5;
    static constexpr int PITCH = (1 << S) + (1 << (S - 3));
    u32 base;  // w * PITCH
    __device__ __forceinline__ u32 operator()(u32 p) const { return base + p + (p >> 3); }
};
template <int W>
struct ColPos {
    static constexpr bool WARP_LOCAL = false;  // a column transform spans every warp of the CTA
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

template <class Pos>
__device__ __forceinline__ void pass_sync() {
#ifndef PPQ_NO_WARPSYNC
    if (Pos::WARP_LOCAL) __syncwarp();
    else
#endif
        __syncthreads();
}

__host__ __device__ constexpr int cmax(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int cmin(int a, int b) { return a < b ? a : b; }

// Forward pass over one size-2^S transform.  In: a[k] = point point_of<S-3>(u,k), values in [0,8q).
// Out: a[k] = point point_of<0>(u,k) = 8u + k, values in [0,8q).  `guard` inserts a leading barrier so the
// shared buffer can be reused across loop iterations (write-after-read protection).
#if PPQ_BFLY_V == 2
constexpr int FWD_COL_IN = 4, FWD_ROW_IN = 16;  // COL passes read canonical (or < 4q) data, ROW passes a COL pass's output
template <int S, int BIN>
struct FwdOut {
    static constexpr int B = fwd_bound_after(BIN, S);
};
#else
constexpr int FWD_COL_IN = 8, FWD_ROW_IN = 8;
template <int S, int BIN>
struct FwdOut {
    static constexpr int B = 8;
};
#endif
// [0, B q) -> [0, q) for B <= 16
template <int B>
__device__ __forceinline__ u64 canon_b(u64 x, const LimbC& c) {
    if (B > 8) x = csub(x, c.q2 << 2);
    if (B > 4) x = csub(x, c.q2 << 1);
    if (B > 2) x = csub(x, c.q2);
    return csub(x, c.q);
}
// [0, B q) -> [0, 8q)
template <int B>
__device__ __forceinline__ u64 fold8_b(u64 x, const LimbC& c) {
    return B > 8 ? csub(x, c.q2 << 2) : x;
}

template <int S, class Pos, int BIN = 8>
__device__ __forceinline__ void fwd_pass(u64 (&a)[8], u64* sm, const Pos& pos, u32 u, const ulonglong2* __restrict__ tw,
                                         u32 rowfac, u64 q, u64 q4, bool guard, u64 z = 0) {
    constexpr int LO0 = S - 3;
    constexpr int LO1 = cmax(S - 6, 0);
    ulonglong2 t[7];
    fwd_group_tw<S, LO0, S>(t, tw, rowfac, u);
    fwd_group<S, LO0, S, BIN>(a, t, q, q4, z);
    fwd_group_tw<S, LO1, LO0>(t, tw, rowfac, u);  // in flight during the exchange
    if (guard) pass_sync<Pos>();
    sm_store<LO0>(a, sm, pos, u);
    pass_sync<Pos>();
    sm_load<LO1>(a, sm, pos, u);
    fwd_group<S, LO1, LO0, fwd_bound_after(BIN, 3)>(a, t, q, q4, z);
    if constexpr (S > 6) {
        fwd_group_tw<S, 0, LO1>(t, tw, rowfac, u);
        sm_store<LO1>(a, sm, pos, u);  // same words this thread just read: no barrier needed before
        pass_sync<Pos>();
        sm_load<0>(a, sm, pos, u);
        fwd_group<S, 0, LO1, fwd_bound_after(BIN, S - LO1)>(a, t, q, q4, z);
    }
}

// COL pass over TWO adjacent columns per thread.  Every column of a COL pass uses the same twiddles, so a thread that
// owns the same 8 rows of two neighbouring columns loads each twiddle once for 16 points, moves its data as 16-byte
// words (global and shared) and pays the address arithmetic of the pass once: the part of a pass that is not
// butterflies -- a quarter of k_fwd_col's instructions -- shrinks.
// Shared tile: two planes [column of the pair][point][pair], 8-byte words (a 16-byte store would want its two values
// in one aligned register quad, which costs ptxas about three MOVs per store -- the values come out of independent
// butterflies -- and ptxas fuses two adjacent 8-byte stores back into one).
template <int LO, int W2, int S>
__device__ __forceinline__ void sm_store2(const u64 (&a)[8], const u64 (&b)[8], u64* sm, u32 w, u32 u) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        u64* s = sm + (point_of<LO>(u, k) * W2 + w);
        s[0] = a[k];
        s[W2 << S] = b[k];
    }
}
template <int LO, int W2, int S>
__device__ __forceinline__ void sm_load2(u64 (&a)[8], u64 (&b)[8], const u64* sm, u32 w, u32 u) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const u64* s = sm + (point_of<LO>(u, k) * W2 + w);
        a[k] = s[0];
        b[k] = s[W2 << S];
    }
}
template <int S, int W2, int BIN>
__device__ __forceinline__ void fwd_pass_col2(u64 (&a)[8], u64 (&b)[8], u64* sm, u32 w, u32 u,
                                              const ulonglong2* __restrict__ tw, u64 q, u64 q4, u64 z) {
    static_assert(S > 6, "two-column COL passes are built for S >= 7");
    constexpr int LO0 = S - 3;
    constexpr int LO1 = S - 6;
    ulonglong2 t[7];
    fwd_group_tw<S, LO0, S>(t, tw, 1u, u);
    fwd_group<S, LO0, S, BIN>(a, t, q, q4, z);
    fwd_group<S, LO0, S, BIN>(b, t, q, q4, z);
    fwd_group_tw<S, LO1, LO0>(t, tw, 1u, u);
    sm_store2<LO0, W2, S>(a, b, sm, w, u);
    __syncthreads();
    sm_load2<LO1, W2, S>(a, b, sm, w, u);
    fwd_group<S, LO1, LO0, fwd_bound_after(BIN, 3)>(a, t, q, q4, z);
    fwd_group<S, LO1, LO0, fwd_bound_after(BIN, 3)>(b, t, q, q4, z);
    fwd_group_tw<S, 0, LO1>(t, tw, 1u, u);
    sm_store2<LO1, W2, S>(a, b, sm, w, u);  // same words this thread just read: no barrier needed before
    __syncthreads();
    sm_load2<0, W2, S>(a, b, sm, w, u);
    fwd_group<S, 0, LO1, fwd_bound_after(BIN, S - LO1)>(a, t, q, q4, z);
    fwd_group<S, 0, LO1, fwd_bound_after(BIN, S - LO1)>(b, t, q, q4, z);
}

// Forward pass over the SAME transform (same limb, same row) of two ciphertexts: the twiddles and their addresses are
// shared; the two tiles live in two shared-memory planes `plane` words apart.
template <int S, class Pos, int BIN>
__device__ __forceinline__ void fwd_pass_x2(u64 (&a)[8], u64 (&b)[8], u64* sm, u32 plane, const Pos& pos, u32 u,
                                            const ulonglong2* __restrict__ tw, u32 rowfac, u64 q, u64 q4, bool guard,
                                            u64 z) {
    constexpr int LO0 = S - 3;
    constexpr int LO1 = cmax(S - 6, 0);
    ulonglong2 t[7];
    fwd_group_tw<S, LO0, S>(t, tw, rowfac, u);
    fwd_group<S, LO0, S, BIN>(a, t, q, q4, z);
    fwd_group<S, LO0, S, BIN>(b, t, q, q4, z);
    fwd_group_tw<S, LO1, LO0>(t, tw, rowfac, u);
    if (guard) pass_sync<Pos>();
    sm_store<LO0>(a, sm, pos, u);
    sm_store<LO0>(b, sm + plane, pos, u);
    pass_sync<Pos>();
    sm_load<LO1>(a, sm, pos, u);
    sm_load<LO1>(b, sm + plane, pos, u);
    fwd_group<S, LO1, LO0, fwd_bound_after(BIN, 3)>(a, t, q, q4, z);
    fwd_group<S, LO1, LO0, fwd_bound_after(BIN, 3)>(b, t, q, q4, z);
    if constexpr (S > 6) {
        fwd_group_tw<S, 0, LO1>(t, tw, rowfac, u);
        sm_store<LO1>(a, sm, pos, u);
        sm_store<LO1>(b, sm + plane, pos, u);
        pass_sync<Pos>();
        sm_load<0>(a, sm, pos, u);
        sm_load<0>(b, sm + plane, pos, u);
        fwd_group<S, 0, LO1, fwd_bound_after(BIN, S - LO1)>(a, t, q, q4, z);
        fwd_group<S, 0, LO1, fwd_bound_after(BIN, S - LO1)>(b, t, q, q4, z);
    }
}

// Inverse pass.  In: a[k] = point 8u + k, values < BIN q.  Out: a[k] = point_of<S-3>(u,k), < 8q (< 4q for
// PPQ_BFLY_V == 0).
#if PPQ_BFLY_V == 2
constexpr int INV_ROW_IN = 1, INV_COL_IN = 8;  // ROW passes read canonical data, COL passes a ROW pass's output (< 8q)
#else
constexpr int INV_ROW_IN = 4, INV_COL_IN = 4;
#endif
template <int S, class Pos, int BIN = 4, bool FIN = false>
__device__ __forceinline__ void inv_pass(u64 (&a)[8], u64* sm, const Pos& pos, u32 u, const ulonglong2* __restrict__ tw,
                                         u32 rowfac, u64 q, u64 q4, bool guard, ulonglong2 fin_s = ulonglong2{0, 0},
                                         ulonglong2 fin_w = ulonglong2{0, 0}) {
    constexpr int LOL = S - 3;           // last group
    constexpr int LO1 = cmin(3, S - 3);  // second group
    constexpr int B1 = inv_plan(BIN, 0).out;
    constexpr int B2 = inv_plan(B1, 3 - LO1).out;
    ulonglong2 t[7];
    inv_group_tw<S, 0, 0>(t, tw, rowfac, u);
    inv_group<S, 0, 0, BIN>(a, t, q, q4);
    // (unlike the forward passes, requesting the next group's twiddles before the exchange does not pay here:
    // the inverse kernels are register-tighter and measured 2-3 % slower with it)
    if (guard) pass_sync<Pos>();
    sm_store<0>(a, sm, pos, u);
    pass_sync<Pos>();
    sm_load<LO1>(a, sm, pos, u);
    inv_group_tw<S, LO1, 3>(t, tw, rowfac, u);
    inv_group<S, LO1, 3, B1, (FIN && S <= 6)>(a, t, q, q4, fin_s, fin_w);
    if constexpr (S > 6) {
        sm_store<LO1>(a, sm, pos, u);
        pass_sync<Pos>();
        sm_load<LOL>(a, sm, pos, u);
        inv_group_tw<S, LOL, 6>(t, tw, rowfac, u);
        inv_group<S, LOL, 6, B2, FIN>(a, t, q, q4, fin_s, fin_w);
    }
}

// Inverse pass over the same transform of two ciphertexts (see fwd_pass_x2).
template <int S, class Pos, int BIN>
__device__ __forceinline__ void inv_pass_x2(u64 (&a)[8], u64 (&b)[8], u64* sm, u32 plane, const Pos& pos, u32 u,
                                            const ulonglong2* __restrict__ tw, u32 rowfac, u64 q, u64 q4, bool guard) {
    constexpr int LOL = S - 3;
    constexpr int LO1 = cmin(3, S - 3);
    constexpr int B1 = inv_plan(BIN, 0).out;
    constexpr int B2 = inv_plan(B1, 3 - LO1).out;
    ulonglong2 t[7];
    inv_group_tw<S, 0, 0>(t, tw, rowfac, u);
    inv_group<S, 0, 0, BIN>(a, t, q, q4);
    inv_group<S, 0, 0, BIN>(b, t, q, q4);
    if (guard) pass_sync<Pos>();
    sm_store<0>(a, sm, pos, u);
    sm_store<0>(b, sm + plane, pos, u);
    pass_sync<Pos>();
    sm_load<LO1>(a, sm, pos, u);
    sm_load<LO1>(b, sm + plane, pos, u);
    inv_group_tw<S, LO1, 3>(t, tw, rowfac, u);
    inv_group<S, LO1, 3, B1>(a, t, q, q4);
    inv_group<S, LO1, 3, B1>(b, t, q, q4);
    if constexpr (S > 6) {
        sm_store<LO1>(a, sm, pos, u);
        sm_store<LO1>(b, sm + plane, pos, u);
        pass_sync<Pos>();
        sm_load<LOL>(a, sm, pos, u);
        sm_load<LOL>(b, sm + plane, pos, u);
        inv_group_tw<S, LOL, 6>(t, tw, rowfac, u);
        inv_group<S, LOL, 6, B2>(a, t, q, q4);
        inv_group<S, LOL, 6, B2>(b, t, q, q4);
    }
}

}  // namespace ppq
