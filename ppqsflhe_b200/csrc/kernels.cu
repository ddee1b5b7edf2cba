// kernels.cu -- hand-written sm_100a kernels for the PPQSFLHE encrypted-aggregation hot path.
//
// Everything the reference does through OpenFHE on the CPU for
//   cc->ReEncrypt   (server/src/changeCipherDomain.cpp:74,89,105)
//   cc->EvalAdd / cc->EvalMult(.,0.5) (server/src/aggregateEncryptedWeights.cpp:82-83,91-92,106-107)
//   cc->Encrypt / cc->Decrypt (client/src/encryptModelWeights.cpp:83,91,110; decryptModelWeights.cpp:81,90,108)
// is expressed here as NTT passes (ntt_core.cuh) with the surrounding RNS arithmetic fused in as the
// prologue or epilogue of a pass.  Integer modular arithmetic only: no tensor cores (nothing here is a dense
// floating-point contraction).  Layouts are limb-major: a polynomial limb is N contiguous u64 residues.
#include <cstdlib>

#include "kernels.cuh"
#include "ckks.cuh"
#include "ntt_core.cuh"

#ifndef PPQ_ROW_MINB
#define PPQ_ROW_MINB 8  // 128-thread CTAs (PPQ_ROW_PTS 1024), 64 registers: best of 5..20 on B200 (profiles/r01_notes.md)
#endif
#ifndef PPQ_ROW_GPC
#define PPQ_ROW_GPC 3  // ciphertexts per ROW CTA (same rows, same twiddles: the later ones find the twiddles in L1); 1: 134.4, 2: 133.6, 3: 133.2, 6-18: 133.1-134.0 us/ct
#endif
#ifndef PPQ_ROW_PTS
#define PPQ_ROW_PTS 1024
#endif
#ifndef PPQ_COL_WDIV
#define PPQ_COL_WDIV 1
#endif
#ifndef PPQ_COL_MINB
#define PPQ_COL_MINB 5  // 51 registers: 8.4 -> 7.7 us/ct on the inverse COL passes, forward unchanged (profiles/r03_notes.md)
#endif
#ifndef PPQ_COL9_MINB
// 512-thread COL CTAs (n1 = 512, N = 2^17) at 2 CTAs/SM: 64 registers, no spills; C4 601 -> 528 us/ct against 1 CTA/SM
// (92 registers); 3 CTAs/SM spills 8-44 bytes for 521, 4 CTAs/SM 523 (tools/stage_times.py --config c4)
#define PPQ_COL9_MINB 2
#endif
namespace ppq {

#ifndef PPQ_NO_ZADD
#define PPQ_Z(lc) ((u64)(lc).pad)  // run-time zero, see ct_bfly_lazy
#else
#define PPQ_Z(lc) 0ull
#endif

// ---------------------------------------------------------------------------------------------------------
// streaming global accesses: data limbs are touched once per pass (evict-first), twiddles/keys stay cached
__device__ __forceinline__ u64 ld_stream(const u64* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(u64* p, u64 v) { __stcs(p, v); }

// bring a line that the epilogue will read into L2 while the NTT pass runs (costs no register)
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ void ld8(u64 (&a)[8], const u64* p) {  // 64 contiguous bytes, 16-byte aligned
    const ulonglong2* v = reinterpret_cast<const ulonglong2*>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        ulonglong2 t = v[i];
        a[2 * i] = t.x;
        a[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ void ld8_ro(u64 (&a)[8], const u64* p) {
    const ulonglong2* v = reinterpret_cast<const ulonglong2*>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        ulonglong2 t = __ldg(v + i);
        a[2 * i] = t.x;
        a[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ void st8(u64* p, const u64 (&a)[8]) {
    ulonglong2* v = reinterpret_cast<ulonglong2*>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = make_ulonglong2(a[2 * i], a[2 * i + 1]);
}

// =========================================================================================================
// COL passes
// =========================================================================================================
// MODE 0: plain (src poly = blockIdx.y, limb from limb_ids; a negative limb id skips the slot)
// MODE 2: centred SwitchModulus prologue (DropLastElementAndScale)
// MODE 3: small signed coefficients -> residues prologue (secret / noise / encoded plaintext polynomials): the int64
//         polynomial is read once per limb by the pass itself instead of being expanded to n_small * ell limbs first
// SAME: the ROW size equals the COL size (n2 = 2^S), else n2 = 2^(S-1) (split_of): with n2 known at compile time the
// eight strided accesses of a thread share one base register and use immediate offsets.
template <int S, int W, int MODE, bool SAME>
__global__ void __launch_bounds__(W << (S - 3), S == 9 ? PPQ_COL9_MINB : (MODE == 2 && S == 6) ? 4 : PPQ_COL_MINB)
    k_fwd_col(KCtx c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs,
              const BcJob* __restrict__ jobs, const u64* __restrict__ chat, const int* __restrict__ limb_ids,
              u64 q_src) {
    __shared__ u64 sm[W << S];
    constexpr u32 n2 = 1u << (SAME ? S : S - 1), N = n2 << S;
    const u32 w = threadIdx.x % W, u = threadIdx.x / W;
    const u32 col = blockIdx.x * W + w;
    const u32 y = blockIdx.y, g = blockIdx.z;

    int limb, ns = 1, chat_off = 0;
    const u64* in;
    u64* out;
    if (MODE == 0) {
        limb = limb_ids[y];
        if (limb < 0) return;  // block-uniform
        in = src + g * src_gs + (size_t)y * N;
        out = dst + g * dst_gs + (size_t)y * N;
    } else if (MODE == 3) {  // source: small signed coefficients [g][src_gs][N] (int64); slot y = s * ell + i, ell = q_src
        const u32 ell = (u32)q_src;
        limb = (int)(y % ell);
        in = src + (g * src_gs + y / ell) * N;
        out = dst + g * dst_gs + (size_t)y * N;
    } else {
        const BcJob jb = jobs[y];
        limb = jb.limb;
        ns = jb.ns;
        chat_off = jb.chat_off;
        in = src + g * src_gs + (size_t)jb.src_first * N;
        out = dst + g * dst_gs + (size_t)jb.dst_slot * N;
    }
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;

    u64 a[8];
    if (MODE == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = ld_stream(in + (size_t)point_of<S - 3>(u, k) * n2 + col);
    } else if (MODE == 3) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const long long v = reinterpret_cast<const long long*>(in)[(size_t)point_of<S - 3>(u, k) * n2 + col];
            const u64 av = v < 0 ? (u64)(-v) : (u64)v;
            const u64 m = av < lc.q ? av : reduce64(av, lc);  // samples are a few bits wide: no reduction (warp-uniform)
            a[k] = (v < 0 && m != 0) ? lc.q - m : m;
        }
        // fused encode + encrypt (ppq_encode_encrypt): slot 1 (e0) also receives the encoded plaintext coefficients
        // `chat` [g][N] (int64) times 2^approx, so that m + e0 costs ONE transform per limb (the NTT is linear)
        if (chat != nullptr && y / (u32)q_src == 1u) {  // block-uniform
            const long long* co = reinterpret_cast<const long long*>(chat) + (size_t)g * N;
            const u32 approx = (u32)(q_src >> 32);
            const u64 f = reduce64(1ull << approx, lc);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const long long v = co[(size_t)point_of<S - 3>(u, k) * n2 + col];
                u64 m = reduce64(v < 0 ? (u64)(-v) : (u64)v, lc);
                if (approx) m = mulmod_barrett(m, f, lc);
                a[k] = addmod(a[k], (v < 0 && m != 0) ? lc.q - m : m, lc.q);
            }
        }
    } else {
        const u64 half = q_src >> 1;
        const u64 qs_mod = reduce64(q_src, lc);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const u64 r = in[(size_t)point_of<S - 3>(u, k) * n2 + col];
            const u64 v = reduce64(r, lc);
            a[k] = r > half ? submod(v, qs_mod, lc.q) : v;
        }
    }
    ColPos<W> pos{w};
    fwd_pass<S, ColPos<W>, FWD_COL_IN>(a, sm, pos, u, tw, 1u, lc.q, lc.q2 << 1, false, PPQ_Z(lc));
#pragma unroll
    for (int k = 0; k < 8; ++k) out[(size_t)(8 * u + k) * n2 + col] = a[k];  // lazy (< 16q): consumed by a ROW pass
}

// Plain forward COL pass, two adjacent columns per thread (ntt_core.cuh fwd_pass_col2), in place.
// Measured on B200, C3 (tools/stage_times.py): forward 18.5 -> 17.9 and 12.1 -> 11.5 us/ct (768 -> 700 instructions
// per 8 points at 3 CTAs of 256 threads per SM); the inverse form (936 -> 888) was no faster (6.8 -> 7.0) and has been removed.
// Only for n1 = 256 (N = 2^15, 2^16): at n1 = 128 the form measured 3 % slower (N = 2^13), at n1 = 512 the same.
#ifndef PPQ_COL2
#define PPQ_COL2 1
#endif
#ifndef PPQ_COL2_W
#define PPQ_COL2_W 8  // column pairs per CTA
#endif
#ifndef PPQ_COL2_MINB
#define PPQ_COL2_MINB 3
#endif
template <int S, int W2, bool SAME>
__global__ void __launch_bounds__(W2 << (S - 3), PPQ_COL2_MINB)
    k_fwd_col2(KCtx c, u64* __restrict__ data, size_t gs, const int* __restrict__ limb_ids) {
    __shared__ u64 sm[2 * (W2 << S)];
    constexpr u32 n2 = 1u << (SAME ? S : S - 1), N = n2 << S;
    const u32 w = threadIdx.x % W2, u = threadIdx.x / W2;
    const u32 col = (blockIdx.x * W2 + w) * 2;
    const u32 y = blockIdx.y, g = blockIdx.z;
    const int limb = limb_ids[y];
    if (limb < 0) return;  // block-uniform
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    u64* p = data + g * gs + (size_t)y * N + col;
    u64 a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const ulonglong2 v = __ldcs(reinterpret_cast<const ulonglong2*>(p + (size_t)point_of<S - 3>(u, k) * n2));
        a[k] = v.x;
        b[k] = v.y;
    }
    fwd_pass_col2<S, W2, FWD_COL_IN>(a, b, sm, w, u, tw, lc.q, lc.q2 << 1, PPQ_Z(lc));
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // lazy (< 16q): consumed by a ROW pass
#if PPQ_COL2_ST128
        *reinterpret_cast<ulonglong2*>(p + (size_t)(8 * u + k) * n2) = make_ulonglong2(a[k], b[k]);
#else
        p[(size_t)(8 * u + k) * n2] = a[k];
        p[(size_t)(8 * u + k) * n2 + 1] = b[k];
#endif
    }
}

// EXPERIMENT (-DPPQ_COL2_PF=<ciphertexts per CTA>): the same pass over PPQ_COL2_PF consecutive ciphertexts per CTA with
// the NEXT ciphertext's tile on its way into a per-thread staging area of shared memory (cp.async, 16 bytes per copy,
// no registers held) while the current one is transformed: k_fwd_col2 shows 16 % long_scoreboard stall samples, the
// loads at the start of every CTA.
#ifndef PPQ_COL2_PF
#define PPQ_COL2_PF 0
#endif
#if PPQ_COL2_PF
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const u32 d = (u32)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
template <int S, int W2, bool SAME>
__global__ void __launch_bounds__(W2 << (S - 3), PPQ_COL2_MINB)
    k_fwd_col2p(KCtx c, u64* __restrict__ data, size_t gs, const int* __restrict__ limb_ids, int n_g) {
    extern __shared__ __align__(16) u64 dsm[];
    constexpr u32 T = W2 << (S - 3);
    u64* sm = dsm;                                                              // exchange tile, 2 * (W2 << S) words
    ulonglong2* stage = reinterpret_cast<ulonglong2*>(dsm + 2 * (W2 << S));  // [8][T], private to each thread
    constexpr u32 n2 = 1u << (SAME ? S : S - 1), N = n2 << S;
    const u32 w = threadIdx.x % W2, u = threadIdx.x / W2;
    const u32 col = (blockIdx.x * W2 + w) * 2;
    const u32 y = blockIdx.y;
    const int limb = limb_ids[y];
    if (limb < 0) return;  // block-uniform
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    const u32 g0 = blockIdx.z * PPQ_COL2_PF, g1 = min((u32)n_g, g0 + PPQ_COL2_PF);
    u64* p = data + g0 * gs + (size_t)y * N + col;
    ulonglong2* mine = stage + threadIdx.x;
#pragma unroll
    for (int k = 0; k < 8; ++k) cp_async16(mine + k * T, p + (size_t)point_of<S - 3>(u, k) * n2);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (u32 g = g0; g < g1; ++g, p += gs) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        u64 a[8], b[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const ulonglong2 v = mine[k * T];
            a[k] = v.x;
            b[k] = v.y;
        }
        if (g + 1 < g1) {
#pragma unroll
            for (int k = 0; k < 8; ++k) cp_async16(mine + k * T, p + gs + (size_t)point_of<S - 3>(u, k) * n2);
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        fwd_pass_col2<S, W2, FWD_COL_IN>(a, b, sm, w, u, tw, lc.q, lc.q2 << 1, PPQ_Z(lc));
#pragma unroll
        for (int k = 0; k < 8; ++k) {  // lazy (< 16q): consumed by a ROW pass
            p[(size_t)(8 * u + k) * n2] = a[k];
            p[(size_t)(8 * u + k) * n2 + 1] = b[k];
        }
        if (g + 1 < g1) __syncthreads();  // the exchange tile is reused
    }
}
#endif

// The two-column forward COL pass with the small-coefficient prologue of k_fwd_col<MODE 3> (signed samples -> residues,
// the encoded plaintext joining slot 1): what Encrypt / KeyGen / ReKeyGen run on their sampled polynomials at n1 = 256.
template <int S, int W2, bool SAME>
__global__ void __launch_bounds__(W2 << (S - 3), PPQ_COL2_MINB)
    k_fwd_col2_small(KCtx c, const long long* __restrict__ small, u32 n_small, u64* __restrict__ out, size_t gs, u32 ell,
                     const long long* __restrict__ pt_coeff, u32 approx) {
    __shared__ u64 sm[2 * (W2 << S)];
    constexpr u32 n2 = 1u << (SAME ? S : S - 1), N = n2 << S;
    const u32 w = threadIdx.x % W2, u = threadIdx.x / W2;
    const u32 col = (blockIdx.x * W2 + w) * 2;
    const u32 y = blockIdx.y, g = blockIdx.z;
    const u32 limb = y % ell, slot = y / ell;
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    const long long* in = small + ((size_t)g * n_small + slot) * N + col;
    auto res = [&](long long v) {
        const u64 av = v < 0 ? (u64)(-v) : (u64)v;
        const u64 m = av < lc.q ? av : reduce64(av, lc);
        return (v < 0 && m != 0) ? lc.q - m : m;
    };
    u64 a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const longlong2 v = *reinterpret_cast<const longlong2*>(in + (size_t)point_of<S - 3>(u, k) * n2);
        a[k] = res(v.x);
        b[k] = res(v.y);
    }
    if (pt_coeff != nullptr && slot == 1u) {  // block-uniform: m * 2^approx joins e0 before the transform
        const long long* co = pt_coeff + (size_t)g * N + col;
        const u64 f = reduce64(1ull << approx, lc);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const longlong2 v = *reinterpret_cast<const longlong2*>(co + (size_t)point_of<S - 3>(u, k) * n2);
            u64 mx = res(v.x), my = res(v.y);
            if (approx) {
                mx = mulmod_barrett(mx, f, lc);
                my = mulmod_barrett(my, f, lc);
            }
            a[k] = addmod(a[k], mx, lc.q);
            b[k] = addmod(b[k], my, lc.q);
        }
    }
    fwd_pass_col2<S, W2, FWD_COL_IN>(a, b, sm, w, u, tw, lc.q, lc.q2 << 1, PPQ_Z(lc));
    u64* p = out + (size_t)g * gs + (size_t)y * N + col;
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // lazy (< 16q): consumed by a ROW pass
        p[(size_t)(8 * u + k) * n2] = a[k];
        p[(size_t)(8 * u + k) * n2 + 1] = b[k];
    }
}

template <int S, int W, bool SAME>
__global__ void __launch_bounds__(W << (S - 3), S == 9 ? PPQ_COL9_MINB : PPQ_COL_MINB)
    k_inv_col(KCtx c, u64* __restrict__ data, size_t gs, const int* __restrict__ limb_map,
              const u64* __restrict__ scale) {
    __shared__ u64 sm[W << S];
    constexpr u32 n2 = 1u << (SAME ? S : S - 1), N = n2 << S;
    const u32 w = threadIdx.x % W, u = threadIdx.x / W;
    const u32 col = blockIdx.x * W + w;
    const u32 y = blockIdx.y, g = blockIdx.z;
    const int limb = limb_map[y];
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twi + (size_t)limb * N;
    u64* p = data + g * gs + (size_t)y * N + col;
    u64 a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = p[(size_t)(8 * u + k) * n2];
    ColPos<W> pos{w};
    // scale[] holds, per slot, (f, shoup(f), f * psi^-brev(1), shoup(.)): the output factor rides on the last stage
    const ulonglong2 fs = __ldg(reinterpret_cast<const ulonglong2*>(scale) + 2 * y);
    const ulonglong2 fw = __ldg(reinterpret_cast<const ulonglong2*>(scale) + 2 * y + 1);
    inv_pass<S, ColPos<W>, INV_COL_IN, true>(a, sm, pos, u, tw, 1u, lc.q, lc.q2 << 1, false, fs, fw);
#pragma unroll
    for (int k = 0; k < 8; ++k) p[(size_t)point_of<S - 3>(u, k) * n2] = canon4(a[k], lc);
}

// =========================================================================================================
// Approximate CRT basis switch (DCRTPolyImpl::ApproxSwitchCRTBasis), pointwise over coefficients:
//   dst[job.dst_slot][n] = ( sum_{i<NS} src[src_first+i][n] * chat[job.chat_off+i] ) mod q_{job.limb}
// WITHOUT the overflow correction (bit-exactness with OpenFHE needs the approximate form).  One thread owns
// one coefficient: its NS source residues are loaded once (coalesced) and reused for the TPB targets of this
// block.  All jobs of a launch share (src_first, NS).  Sources are canonical COEFF residues (< 2^60) already
// scaled by (D/q_i)^-1 (k_inv_col).
//
// The exact integer sum is formed with a 30-bit Karatsuba split, y = y1 2^30 + y0, c = c1 2^30 + c0:
//   sum y c = A0 + (A1 - A0 - A2) 2^30 + A2 2^60,  A0 = sum y0 c0, A2 = sum y1 c1, A1 = sum (y0+y1)(c0+c1)
// i.e. THREE multiply-adds per term, each a single IMAD.WIDE with a 64-bit addend (no carry chain), instead of
// the four IMAD.WIDE + carries of a 64x64->128 MAC: wide multiplies are the scarce resource on B200 (~5 clk
// per warp and SMSP on the fmaheavy pipe, tools/ubench/int_pipes.cu).  Up to 8 terms per accumulator set
// (A0, A2 < 8 2^60; A1 may wrap: only A1 - A0 - A2 = sum (y1 c0 + y0 c1) < 2^64 is used, and that is exact modulo
// 2^64), one Barrett reduction per 8 terms.  Unsigned products only: a signed IMAD.WIDE costs two passes on B200
// (measured, profiles/r01_notes.md).  chat3[.] = (c0, c1, c0 + c1, 0).
// =========================================================================================================
// 32 x 32 -> 64 product spelled in PTX: the C form (u64)a * b with b read from the kernel parameter goes through a
// zero-extended uniform register pair and leaves additions of zero behind in the SASS; ptxas fuses this product with
// the C addition that follows into one IMAD.WIDE
__device__ __forceinline__ u64 mulw(u32 a, u32 b) {
    u64 r;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(a), "r"(b));
    return r;
}
// 32 x 32 + 64 -> 64 multiply-add spelled in PTX (k_ip_tma: 17.1 -> 15.9 us/ct together with the pre-split keys)
__device__ __forceinline__ u64 madw(u32 a, u32 b, u64 c) {
    u64 r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
    return r;
}
template <bool LAZY = false>
__device__ __forceinline__ u64 kara_reduce(u64 A0, u64 A2, u64 A1, const LimbC& lc) {
    const u64 mid = A1 - A0 - A2;  // = sum (y1 c0 + y0 c1) < 2^64: exact modulo 2^64
    unsigned __int128 t = (unsigned __int128)A0 + ((unsigned __int128)mid << 30) + ((unsigned __int128)A2 << 60);
    return LAZY ? barrett128_lazy3((u64)(t >> 64), (u64)t, lc) : barrett128((u64)(t >> 64), (u64)t, lc);
}
// Conversions of narrow digits write values in [0, 3q): every consumer is a forward COL pass, which reads < 4q
// (FWD_COL_IN); two conditional subtractions less per output (k_bconv2: 89 -> 79 instructions per output).
#ifndef PPQ_BCONV_LAZY
#define PPQ_BCONV_LAZY 1
#endif
static_assert(!PPQ_BCONV_LAZY || FWD_COL_IN >= 3, "lazy conversions need a COL pass that accepts 3q");

// Constants of one launch travel as a kernel parameter (constant bank, read with LDC/ULDC by a block-uniform index):
// per-thread LDG of conversion constants saturated L1TEX (74 % in profiles/r01_notes.md) without doing useful work.
template <int NS>
struct BcPack {
    struct J {
        u64 q, mu;
        u32 sh;
        int dst_slot;
        u32 pad[2];
        uint4 c[NS];  // (c0, c1, c0 + c1, 0) per source limb
    };
    static constexpr int MAXJ = (12288 / (int)sizeof(J)) < 48 ? (12288 / (int)sizeof(J)) : 48;
    int n_jobs, src_first;
    int pad[2];
    J j[MAXJ];
};

template <int NS>
__global__ void __launch_bounds__(256)
    k_bconv(int logN, const u64* __restrict__ src, size_t src_gs, u64* __restrict__ dst, size_t dst_gs,
            const __grid_constant__ BcPack<NS> P) {
    const u32 N = 1u << logN;
    const u32 n = blockIdx.x * 256 + threadIdx.x;
    const u32 g = blockIdx.z;
    const u64* in = src + g * src_gs + (size_t)P.src_first * N + n;
    // narrow digits (NS <= 8) keep y0 + y1 in a register; wide ones re-form it per term (registers are the scarcer
    // resource there)
    constexpr bool KEEP = NS <= 8;
    u32 y0[NS], y1[NS], ys[KEEP ? NS : 1];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
        const u64 y = ld_stream(in + (size_t)i * N);
        y0[i] = (u32)y & 0x3fffffffu;
        y1[i] = (u32)(y >> 30);
        if (KEEP) ys[i] = y0[i] + y1[i];
    }
    u64* out = dst + g * dst_gs + n;
#pragma unroll(NS <= 8 ? 2 : 1)
    for (int t = 0; t < P.n_jobs; ++t) {
        const typename BcPack<NS>::J& J = P.j[t];
        LimbC lc;
        lc.q = J.q;
        lc.q2 = J.q << 1;
        lc.mu = J.mu;
        lc.sh = J.sh;
        u64 r = 0;
#pragma unroll
        for (int i0 = 0; i0 < NS; i0 += 8) {
            u64 A0, A2, A1;
            {
                const uint4 cc = J.c[i0];
                A0 = mulw(y0[i0], cc.x);
                A2 = mulw(y1[i0], cc.y);
                A1 = mulw(KEEP ? ys[i0] : y0[i0] + y1[i0], cc.z);
            }
#pragma unroll
            for (int i = i0 + 1; i < (i0 + 8 < NS ? i0 + 8 : NS); ++i) {
                const uint4 cc = J.c[i];
                A0 += mulw(y0[i], cc.x);  // mul.wide + C add: one IMAD.WIDE (a PTX mad.wide is split by ptxas, 3 % slower)
                A2 += mulw(y1[i], cc.y);
                A1 += mulw(KEEP ? ys[i] : y0[i] + y1[i], cc.z);
            }
            r += kara_reduce<(PPQ_BCONV_LAZY && NS <= 8)>(A0, A2, A1, lc);
        }
        if (NS > 8) r = canon4(r, lc);  // at most 4 chunks
        out[(size_t)J.dst_slot * N] = r;
    }
}

// Two adjacent coefficients per thread (narrow digits, NS <= 8): the per-target work that does not depend on the
// coefficient -- loading the conversion constants from the constant bank, loop control, address arithmetic -- is
// paid once per pair, and the residues travel as 16-byte accesses.  PPQ_BCONV_PAIR=0 selects the one-coefficient form.
#ifndef PPQ_BCONV_PAIR
#define PPQ_BCONV_PAIR 1
#endif
#ifndef PPQ_BCONV_E
#define PPQ_BCONV_E 2  // coefficients per thread (4: 128 registers, 14.9 + 10.2 -> 15.9 + 10.8 us/ct at C3)
#endif
template <int NS>
__global__ void __launch_bounds__(256)
    k_bconv2(int logN, const u64* __restrict__ src, size_t src_gs, u64* __restrict__ dst, size_t dst_gs,
             const __grid_constant__ BcPack<NS> P) {
    constexpr int E = PPQ_BCONV_E;
    const u32 N = 1u << logN;
    const u32 n = (blockIdx.x * 256 + threadIdx.x) * E;
    const u32 g = blockIdx.z;
    const u64* in = src + g * src_gs + (size_t)P.src_first * N + n;
    u32 y0[E][NS], y1[E][NS], ys[E][NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
#pragma unroll
        for (int h = 0; h < E / 2; ++h) {
            const ulonglong2 y = __ldcs(reinterpret_cast<const ulonglong2*>(in + (size_t)i * N) + h);
            y0[2 * h][i] = (u32)y.x & 0x3fffffffu;
            y1[2 * h][i] = (u32)(y.x >> 30);
            ys[2 * h][i] = y0[2 * h][i] + y1[2 * h][i];
            y0[2 * h + 1][i] = (u32)y.y & 0x3fffffffu;
            y1[2 * h + 1][i] = (u32)(y.y >> 30);
            ys[2 * h + 1][i] = y0[2 * h + 1][i] + y1[2 * h + 1][i];
        }
    }
    u64* out = dst + g * dst_gs + n;
#pragma unroll 1
    for (int t = 0; t < P.n_jobs; ++t) {
        const typename BcPack<NS>::J& J = P.j[t];
        LimbC lc;
        lc.q = J.q;
        lc.q2 = J.q << 1;
        lc.mu = J.mu;
        lc.sh = J.sh;
        u64 A0[E], A2[E], A1[E];
#pragma unroll
        for (int i = 0; i < NS; ++i) {
            const uint4 cc = J.c[i];
#pragma unroll
            for (int e = 0; e < E; ++e) {
                if (i == 0) {
                    A0[e] = mulw(y0[e][i], cc.x);
                    A2[e] = mulw(y1[e][i], cc.y);
                    A1[e] = mulw(ys[e][i], cc.z);
                } else {
                    A0[e] += mulw(y0[e][i], cc.x);
                    A2[e] += mulw(y1[e][i], cc.y);
                    A1[e] += mulw(ys[e][i], cc.z);
                }
            }
        }
#pragma unroll
        for (int h = 0; h < E / 2; ++h)
            reinterpret_cast<ulonglong2*>(out + (size_t)J.dst_slot * N)[h] =
                make_ulonglong2(kara_reduce<PPQ_BCONV_LAZY>(A0[2 * h], A2[2 * h], A1[2 * h], lc),
                                kara_reduce<PPQ_BCONV_LAZY>(A0[2 * h + 1], A2[2 * h + 1], A1[2 * h + 1], lc));
    }
}

// [0, B q) -> [0, q) for B <= 16 and any limb width (the quotient-estimate form needs bitlen(q) >= 29; block-uniform)
__device__ __forceinline__ u64 canon_any16(u64 x, const LimbC& lc) {
    return lc.sh >= 28u ? canon16_q(x, lc) : canon_b<16>(x, lc);
}

// =========================================================================================================
// Wide base conversion for the deferred ModDown (ppq_agg_*).  ApproxModDown's conversion P -> Q is linear in the
// INTEGER lifts u_k in [0, p_k):  sum_c sum_k u_{k,c} chat_k = sum_k (sum_c u_{k,c}) chat_k (mod q_i), so the
// conversions of up to 8 clients' ciphertexts of one slot cost one conversion of U_k = sum_c u_{k,c} < 2^63.
// U is split into 21-bit pieces, chat into its 30-bit halves: six IMAD.WIDE per term, partial sums < 16 2^52; the
// exact 128-bit total V = hi 2^64 + lo (< 16 2^63 2^60) is reduced as hi * (2^64 mod q) + lo: one Shoup product
// (any 64-bit operand) and one Barrett estimate, for every limb width.
// =========================================================================================================
template <int NS>
struct BwPack {
    struct J {
        u64 q, mu, r64, r64s;  // modulus, Barrett constant, 2^64 mod q and its Shoup companion
        u32 sh;
        int dst_slot;
        uint2 c[NS];  // (c0, c1): 30-bit halves of the conversion constant per source limb
    };
    static constexpr int MAXJ = (12288 / (int)sizeof(J)) < 48 ? (12288 / (int)sizeof(J)) : 48;
    int n_jobs, src_first;
    J j[MAXJ];
};
template <int NS>
__global__ void __launch_bounds__(256)
    k_bconv_wide(int logN, const u64* __restrict__ src, size_t src_gs, size_t in_stride, int n_in,
                 u64* __restrict__ dst, size_t dst_gs, int first, const __grid_constant__ BwPack<NS> P) {
    constexpr int E = NS <= 8 ? 2 : 1;  // coefficients per thread
    const u32 N = 1u << logN;
    const u32 n = (blockIdx.x * 256 + threadIdx.x) * E;
    const u32 g = blockIdx.z;
    const u64* in = src + g * src_gs + (size_t)P.src_first * N + n;
    u32 u0[E][NS], u1[E][NS], u2[E][NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
        u64 U[E];
#pragma unroll
        for (int e = 0; e < E; ++e) U[e] = 0;
        for (int cl = 0; cl < n_in; ++cl) {
            const u64* p = in + (size_t)cl * in_stride + (size_t)i * N;
            if (E == 2) {
                const ulonglong2 y = __ldcs(reinterpret_cast<const ulonglong2*>(p));
                U[0] += y.x;
                U[E - 1] += y.y;
            } else {
                U[0] += ld_stream(p);
            }
        }
#pragma unroll
        for (int e = 0; e < E; ++e) {
            u0[e][i] = (u32)U[e] & 0x1fffffu;
            u1[e][i] = (u32)(U[e] >> 21) & 0x1fffffu;
            u2[e][i] = (u32)(U[e] >> 42);
        }
    }
    u64* out = dst + g * dst_gs + n;
#pragma unroll 1
    for (int t = 0; t < P.n_jobs; ++t) {
        const typename BwPack<NS>::J& J = P.j[t];
        LimbC lc;
        lc.q = J.q;
        lc.q2 = J.q << 1;
        lc.mu = J.mu;
        lc.sh = J.sh;
        u64 S[E][6];
#pragma unroll
        for (int i = 0; i < NS; ++i) {
            const uint2 cc = J.c[i];
#pragma unroll
            for (int e = 0; e < E; ++e) {
                if (i == 0) {
                    S[e][0] = mulw(u0[e][i], cc.x);
                    S[e][1] = mulw(u1[e][i], cc.x);
                    S[e][2] = mulw(u2[e][i], cc.x);
                    S[e][3] = mulw(u0[e][i], cc.y);
                    S[e][4] = mulw(u1[e][i], cc.y);
                    S[e][5] = mulw(u2[e][i], cc.y);
                } else {
                    S[e][0] += mulw(u0[e][i], cc.x);
                    S[e][1] += mulw(u1[e][i], cc.x);
                    S[e][2] += mulw(u2[e][i], cc.x);
                    S[e][3] += mulw(u0[e][i], cc.y);
                    S[e][4] += mulw(u1[e][i], cc.y);
                    S[e][5] += mulw(u2[e][i], cc.y);
                }
            }
        }
        u64 r[E];
#pragma unroll
        for (int e = 0; e < E; ++e) {
            typedef unsigned __int128 u128;
            const u128 X0 = (u128)S[e][0] + ((u128)S[e][1] << 21) + ((u128)S[e][2] << 42);  // sum U c0
            const u128 X1 = (u128)S[e][3] + ((u128)S[e][4] << 21) + ((u128)S[e][5] << 42);  // sum U c1
            const u128 V = X0 + (X1 << 30);
            u64 v = mul_shoup_lazy((u64)(V >> 64), J.r64, J.r64s, lc.q) + barrett128_lazy3(0ull, (u64)V, lc);  // < 5q
            if (!first) v += out[(size_t)J.dst_slot * N + e];
            r[e] = canon_any16(v, lc);
        }
        if (E == 2) {
            *reinterpret_cast<ulonglong2*>(out + (size_t)J.dst_slot * N) = make_ulonglong2(r[0], r[E - 1]);
        } else {
            out[(size_t)J.dst_slot * N] = r[0];
        }
    }
}

// =========================================================================================================
// ROW passes
// =========================================================================================================
template <int S>
struct RowGeom {
    static constexpr int R = PPQ_ROW_PTS >> S;  // rows per CTA: PPQ_ROW_PTS points, 8 per thread
    static constexpr int TPT = 1 << (S - 3);  // threads per transform
    static constexpr int THREADS = R * TPT;
    static constexpr int SMEM = R * RowPos<S>::PITCH;
};

// Plain forward ROW pass over PAIRS of ciphertexts: the same (slot, row) of two ciphertexts uses the same twiddles, so
// a thread that owns 8 points of both loads every twiddle (and computes its address) once for 16 points.  An odd
// ciphertext at the end of a CTA's range is paired with itself (both copies are read before either is written).
// Measured on B200, C3: k_fwd_row_plain 21.0 -> 20.1 us/ct (2304 -> 2064 instructions per 16 points; 80 registers, 6
// CTAs of 128 threads per SM).
#ifndef PPQ_ROW_X2
#define PPQ_ROW_X2 1
#endif
#ifndef PPQ_ROW_X2_MINB
#define PPQ_ROW_X2_MINB 6
#endif
// The paired kernels are used from N = 2^15 on (measured: C2, N = 2^15, 85.1k -> 87.6k ct/s; C1, N = 2^14 with 27
// ciphertexts per wave, 130k -> 124k ct/s, so smaller rings keep the one-ciphertext form)
#ifndef PPQ_X2_MIN_LOGN
#define PPQ_X2_MIN_LOGN 15
#endif
#ifndef PPQ_ROW_X2_PAIRS
#define PPQ_ROW_X2_PAIRS 2  // pairs per CTA
#endif
template <int S, bool LAZY2>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, PPQ_ROW_X2_MINB)
    k_fwd_row_plain2(KCtx c, u64* __restrict__ data, size_t gs, const int* __restrict__ limb_ids, int n_g) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[2 * G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const u32 y = blockIdx.z;
    const int limb = limb_ids[y];
    if (limb < 0) return;  // block-uniform: slot not to be transformed
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const u32 gend = min((u32)n_g, (blockIdx.y + 1) * (2 * PPQ_ROW_X2_PAIRS));
    for (u32 g = blockIdx.y * (2 * PPQ_ROW_X2_PAIRS); g < gend; g += 2) {
        u64* p0 = data + g * gs + (size_t)y * N + (size_t)row * n2;
        u64* p1 = g + 1 < gend ? p0 + gs : p0;
        u64 a[8], b[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            a[k] = p0[point_of<S - 3>(u, k)];
            b[k] = p1[point_of<S - 3>(u, k)];
        }
        fwd_pass_x2<S, RowPos<S>, FWD_ROW_IN>(a, b, sm, (u32)G::SMEM, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, true, PPQ_Z(lc));
        if (FwdOut<S, FWD_ROW_IN>::B > 8 && lc.sh >= 28u) {  // block-uniform
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                a[k] = LAZY2 ? canon16_q_lazy2(a[k], lc) : canon16_q(a[k], lc);
                b[k] = LAZY2 ? canon16_q_lazy2(b[k], lc) : canon16_q(b[k], lc);
            }
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                a[k] = canon_b<FwdOut<S, FWD_ROW_IN>::B>(a[k], lc);
                b[k] = canon_b<FwdOut<S, FWD_ROW_IN>::B>(b[k], lc);
            }
        }
        st8(p0 + 8 * u, a);
        st8(p1 + 8 * u, b);
    }
}

// (S = 6, N = 2^12 / 2^13: 16 short rows per CTA; 64 registers spilled 132 bytes, so 6 CTAs/SM there)
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, S == 6 ? 6 : PPQ_ROW_MINB)
    k_fwd_row_plain(KCtx c, u64* __restrict__ data, size_t gs, const int* __restrict__ limb_ids, int n_g) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const u32 y = blockIdx.z;
    const int limb = limb_ids[y];
    if (limb < 0) return;  // block-uniform: slot not to be transformed
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const u32 gend = min((u32)n_g, (blockIdx.y + 1) * PPQ_ROW_GPC);
    for (u32 g = blockIdx.y * PPQ_ROW_GPC; g < gend; ++g) {
    u64* p = data + g * gs + (size_t)y * N + (size_t)row * n2;
    u64 a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = p[point_of<S - 3>(u, k)];
    fwd_pass<S, RowPos<S>, FWD_ROW_IN>(a, sm, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, PPQ_ROW_GPC > 1, PPQ_Z(lc));
    if (FwdOut<S, FWD_ROW_IN>::B > 8 && lc.sh >= 28u) {  // block-uniform
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = canon16_q(a[k], lc);
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = canon_b<FwdOut<S, FWD_ROW_IN>::B>(a[k], lc);
    }
    st8(p + 8 * u, a);
    }
}

// Forward ROW pass of the three small polynomials of a fresh ciphertext (v, e0 [+ m], e1: vee [g][3][ell][N], COL-pass
// output) with the Encrypt combine as its epilogue: c0 = pkb * v + e0 (+ pt), c1 = pka * v + e1, written straight into the
// ciphertext.  The three transforms of a (ciphertext, limb, row) run in one thread group with the same twiddles; v stays
// in registers.  Same canonical values as k_fwd_row_plain followed by k_encrypt_pw, without the [3][ell][N] round trip
// through HBM (C3: 63 MB per ciphertext).  grid (n1 / R, 1, n_ct * ell)
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, S == 6 ? 5 : PPQ_ROW_MINB)
    k_fwd_row_encrypt(KCtx c, const u64* __restrict__ vee, const u64* __restrict__ pkb, const u64* __restrict__ pka,
                      const u64* __restrict__ pt, u64* __restrict__ ct, int ell) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const int i = blockIdx.z % ell;
    const size_t g = blockIdx.z / ell;
    const LimbC lc = c.limbs[i];
    const ulonglong2* tw = c.twf + (size_t)i * N;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const size_t roff = (size_t)row * n2;
    u64 v[8];
#pragma unroll 1
    for (int s3 = 0; s3 < 3; ++s3) {
        const u64* p = vee + ((g * 3 + s3) * ell + i) * N + roff;
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = p[point_of<S - 3>(u, k)];
        fwd_pass<S, RowPos<S>, FWD_ROW_IN>(a, sm, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, true, PPQ_Z(lc));
        if (FwdOut<S, FWD_ROW_IN>::B > 8 && lc.sh >= 28u) {  // block-uniform
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = canon16_q(a[k], lc);
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = canon_b<FwdOut<S, FWD_ROW_IN>::B>(a[k], lc);
        }
        if (s3 == 0) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = a[k];
        } else {
            u64 key[8];
            ld8_ro(key, (s3 == 1 ? pkb : pka) + (size_t)i * N + roff + 8 * u);
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = addmod(mulmod_barrett(v[k], key[k], lc), a[k], lc.q);
            if (s3 == 1 && pt != nullptr) {
                u64 m[8];
                ld8_ro(m, pt + (g * ell + i) * N + roff + 8 * u);
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = addmod(a[k], m[k], lc.q);
            }
            st8(ct + ((g * 2 + (s3 - 1)) * ell + i) * N + roff + 8 * u, a);
        }
    }
}

// Source of an inverse ROW pass: ciphertext g of the launch is ciphertext g % n_per of buffer g / n_per (the clients of an
// aggregation pass live in separate buffers; one launch covers them all instead of one launch per client, which at
// N = 2^14 were 20 us launches of 8 us of work)
struct InvSrc {
    const u64* in[4];
    u32 n_per;
    __device__ __forceinline__ const u64* at(u32 g, size_t gs) const {
        const u32 b = g / n_per;
        const u64* base = b == 0 ? in[0] : b == 1 ? in[1] : b == 2 ? in[2] : in[3];  // selects, not an indexed local copy
        return base + (size_t)(g - b * n_per) * gs;
    }
};
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, S == 6 ? 5 : PPQ_ROW_MINB)
    k_inv_row(KCtx c, const InvSrc src, size_t in_gs, u64* out, size_t out_gs,
              const int* __restrict__ limb_map, int n_g) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const u32 y = blockIdx.z;
    const int limb = limb_map[y];
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twi + (size_t)limb * N;
    const size_t off = (size_t)y * N + (size_t)row * n2;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const u32 gend = min((u32)n_g, (blockIdx.y + 1) * PPQ_ROW_GPC);
    for (u32 g = blockIdx.y * PPQ_ROW_GPC; g < gend; ++g) {
    u64 a[8];
    ld8(a, src.at(g, in_gs) + off + 8 * u);
    inv_pass<S, RowPos<S>, INV_ROW_IN>(a, sm, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, PPQ_ROW_GPC > 1);
    u64* o = out + g * out_gs + off;
#pragma unroll
    for (int k = 0; k < 8; ++k) o[point_of<S - 3>(u, k)] = a[k];  // lazy (< 8q): consumed by k_inv_col
    }
}

// Inverse ROW pass with the Decrypt combine c0 + c1 * s as its prologue (a pointwise kernel followed by k_inv_row, without
// the [ell][N] round trip through HBM).  out [n_ct][ell][N]; grid (n1 / R, 1, n_ct * ell)
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, S == 6 ? 5 : PPQ_ROW_MINB)
    k_inv_row_decrypt(KCtx c, const u64* __restrict__ ct, const u64* __restrict__ sk, u64* __restrict__ out, int ell) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const int i = blockIdx.z % ell;
    const size_t g = blockIdx.z / ell;
    const LimbC lc = c.limbs[i];
    const ulonglong2* tw = c.twi + (size_t)i * N;
    const size_t roff = (size_t)row * n2;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    u64 a[8], c1[8], s[8];
    ld8(a, ct + ((g * 2 + 0) * ell + i) * N + roff + 8 * u);
    ld8(c1, ct + ((g * 2 + 1) * ell + i) * N + roff + 8 * u);
    ld8_ro(s, sk + (size_t)i * N + roff + 8 * u);
#pragma unroll
    for (int k = 0; k < 8; ++k)  // c0 (< q) + the product left in [0, 3q): the pass is told its input is below 4q
        a[k] += barrett128_lazy3(__umul64hi(c1[k], s[k]), c1[k] * s[k], lc);
    inv_pass<S, RowPos<S>, 4>(a, sm, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, false);
    u64* o = out + (g * ell + i) * N + roff;
#pragma unroll
    for (int k = 0; k < 8; ++k) o[point_of<S - 3>(u, k)] = a[k];  // lazy (< 8q): consumed by k_inv_col
}

// Inverse ROW pass over pairs of ciphertexts (see k_fwd_row_plain2).
// Measured on B200, C3: 6.3 + 4.5 -> 6.0 + 4.3 us/ct.
#ifndef PPQ_INV_X2
#define PPQ_INV_X2 1
#endif
#ifndef PPQ_INV_X2_MINB
#define PPQ_INV_X2_MINB 6
#endif
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, PPQ_INV_X2_MINB)
    k_inv_row2(KCtx c, const InvSrc src, size_t in_gs, u64* out, size_t out_gs, const int* __restrict__ limb_map, int n_g) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[2 * G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const u32 y = blockIdx.z;
    const int limb = limb_map[y];
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twi + (size_t)limb * N;
    const size_t off = (size_t)y * N + (size_t)row * n2;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const u32 gend = min((u32)n_g, (blockIdx.y + 1) * (2 * PPQ_ROW_X2_PAIRS));
    for (u32 g = blockIdx.y * (2 * PPQ_ROW_X2_PAIRS); g < gend; g += 2) {
        const u32 g1 = g + 1 < gend ? g + 1 : g;  // an odd tail is paired with itself (read twice, written twice)
        u64 a[8], b[8];
        ld8(a, src.at(g, in_gs) + off + 8 * u);
        ld8(b, src.at(g1, in_gs) + off + 8 * u);
        inv_pass_x2<S, RowPos<S>, INV_ROW_IN>(a, b, sm, (u32)G::SMEM, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, true);
        u64* o0 = out + g * out_gs + off;
        u64* o1 = out + g1 * out_gs + off;
#pragma unroll
        for (int k = 0; k < 8; ++k) {  // lazy (< 8q): consumed by k_inv_col
            o0[point_of<S - 3>(u, k)] = a[k];
            o1[point_of<S - 3>(u, k)] = b[k];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Pointwise inner product with the re-encryption key (EvalFastKeySwitchCoreExt), keys staged by TMA.
// One CTA owns (limb r, tile of IPT points).  Its key tile -- parts x {b,a} x IPT residues -- is brought into
// shared memory ONCE by cp.async.bulk (TMA bulk copies completing on an mbarrier) and then reused for every
// ciphertext of the wave: the rekey is read from L2/HBM once per wave instead of once per ciphertext.
// Digits arrive in EVAL form (converted limbs from k_fwd_row_plain, own-digit limbs straight from c1).
// Products use the 30-bit Karatsuba split of k_bconv (3 IMAD.WIDE per term, no carry chains).
// ---------------------------------------------------------------------------------------------------------
#ifndef PPQ_IPT
#define PPQ_IPT 512  // 24 KiB key tile (3 digits): 4 CTAs/SM; 2048 (2 CTAs/SM) left the kernel latency-bound: 14.9 -> 13.1 us/ct
#endif
constexpr int IPT = PPQ_IPT;     // points per key tile; 256 threads take 512 points (two each) per step
constexpr int IPC = IPT / 512;   // steps per ciphertext

__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, u32 bytes, u64* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

struct Kara3 {  // 30-bit split of a residue < 2^60
    u32 lo, hi, sum;
    __device__ __forceinline__ explicit Kara3(u64 v) : lo((u32)v & 0x3fffffffu), hi((u32)(v >> 30)), sum(lo + hi) {}
    struct Packed {};  // v = lo | hi << 32 (kara_pack): the halves are the two registers of v
    __device__ __forceinline__ Kara3(u64 v, Packed) : lo((u32)v), hi((u32)(v >> 32)), sum(lo + hi) {}
};
__device__ __forceinline__ u64 kara_pack(u64 v) { return (v & 0x3fffffffull) | ((v >> 30) << 32); }
struct KaraAcc {
    u64 A0 = 0, A1 = 0, A2 = 0;
    __device__ __forceinline__ void mac(const Kara3& x, const Kara3& y) {
        A0 += mulw(x.lo, y.lo);
        A2 += mulw(x.hi, y.hi);
        A1 += mulw(x.sum, y.sum);
    }
    __device__ __forceinline__ u64 reduce(const LimbC& lc) const { return kara_reduce(A0, A2, A1, lc); }
    __device__ __forceinline__ u64 reduce_lazy3(const LimbC& lc) const { return kara_reduce<true>(A0, A2, A1, lc); }
};
// Q-limb results of the inner product go to the ModDown epilogue, which tolerates [0, 3q) (x' + 8q - conv' + c0 + sum
// < 13q, canonicalised there); P-limb results feed an inverse ROW pass that wants canonical data.
#ifndef PPQ_IP_LAZY
#define PPQ_IP_LAZY 1
#endif
// Digits written by the plain forward ROW pass for the inner product may stay in [0, 2q): one conditional subtraction
// less per point.  Then a sum of `parts` products is < 2 parts q^2, inside Barrett's range 2^(64+sh) for parts <= 4.
#ifndef PPQ_DIGITS_LAZY
#define PPQ_DIGITS_LAZY 1
#endif

#ifndef PPQ_IP_MINB
#define PPQ_IP_MINB 4  // 64 registers, 4 CTAs/SM (5: 14.3, 6: 15.0 us/ct)
#endif
// AGG (deferred ModDown, ppq_agg_*): the Q-limb results are not written out per ciphertext but summed, together with
// c0 on poly 0, into the running EVAL accumulator of the slot (A.acc; A.agg == 1: first client, plain store).
#ifndef PPQ_IP_AGG_MINB
#define PPQ_IP_AGG_MINB 2
#endif
// (more than 3 digits: the per-digit source pointers no longer fit 64 registers -- 3 CTAs/SM instead of spilling)
template <int PARTS, bool AGG>
__global__ void __launch_bounds__(256, AGG ? (PARTS >= 6 ? 1 : PPQ_IP_AGG_MINB) : (PARTS <= 3 ? PPQ_IP_MINB : PARTS <= 5 ? 3 : 2)) k_ip_tma(KCtx c, IpArgs A, int n_g, int g_per_cta) {
    extern __shared__ __align__(128) unsigned char ip_smem[];
    u64* keys = reinterpret_cast<u64*>(ip_smem);              // [PARTS][2][IPT]
    u64* bar = keys + (size_t)PARTS * 2 * IPT;
    const u32 N = 1u << c.logN;
    const int r = blockIdx.y;
    const int M = A.ell + A.K;
    const int limb = r < A.ell ? r : A.L + (r - A.ell);
    const size_t off = (size_t)blockIdx.x * IPT;
    const LimbC lc = c.limbs[limb];
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, (u32)(PARTS * 2 * IPT * sizeof(u64)));
#pragma unroll
        for (int j = 0; j < PARTS; ++j) {
            const bool isq = r < A.ell;  // Q limbs come from the P^-1-scaled copy
            const size_t koff = ((size_t)j * (isq ? A.key_q_stride : A.L + A.K) + limb) * N + off;
            tma_bulk_g2s(keys + (size_t)(2 * j) * IPT, (isq ? A.key_qb : A.key_b) + koff, IPT * sizeof(u64), bar);
            tma_bulk_g2s(keys + (size_t)(2 * j + 1) * IPT, (isq ? A.key_qa : A.key_a) + koff, IPT * sizeof(u64), bar);
        }
    }
    static_assert(IPC == 1, "one 512-point step per ciphertext (PPQ_IPT = 512)");
    const int own = r < A.ell ? r / A.alpha : -1;  // the digit whose own limb r is (taken from c1 as it is)
    const int g0 = blockIdx.z * g_per_cta;
    const int g1 = min(n_g, g0 + g_per_cta);
    // digits of the next ciphertext are fetched while the current one is multiplied.  Address arithmetic is hoisted out
    // of the loop (thread-invariant bases, running slot offsets): recomputing every 64-bit address per iteration was
    // most of this kernel's instructions (profiles/r10_notes.md).
    const bool aggq = AGG && r < A.ell;  // block-uniform
    const u32 pt = threadIdx.x * 2;      // two adjacent points per 16-byte access
    const size_t in_row = (size_t)r * N + off + pt;
    const u64* dsrc[PARTS];
    size_t dstep[PARTS];
#pragma unroll
    for (int j = 0; j < PARTS; ++j) {
        const bool o = j == own;
        dsrc[j] = (o ? A.c1 + in_row : A.digits + ((size_t)j * M + r) * N + off + pt) + (size_t)g0 * (o ? A.c1_gs : A.dig_gs);
        dstep[j] = o ? A.c1_gs : A.dig_gs;
    }
    const u64* c0p = aggq ? A.c0 + (size_t)g0 * A.c0_gs + in_row : nullptr;
    u64* o0;
    size_t ostep;
    if (aggq) {
        o0 = A.acc + (size_t)g0 * A.acc_gs + in_row;
        ostep = A.acc_gs;
    } else if (r < A.ell) {
        o0 = A.extq + (size_t)g0 * A.extq_gs + in_row;
        ostep = A.extq_gs;
    } else {
        o0 = A.extp + (size_t)g0 * A.extp_gs + (size_t)(r - A.ell) * N + off + pt;
        ostep = A.extp_gs;
    }
    const size_t o1off = (size_t)(r < A.ell ? A.ell : A.K) * N;
    ulonglong2 xn[3];  // AGG, Q limbs: c0, running sum of poly 0, running sum of poly 1 (fetched with the digits)
    auto fetch = [&](size_t k, ulonglong2 (&d)[PARTS]) {  // k = slots ahead of the current output position
#pragma unroll
        for (int j = 0; j < PARTS; ++j) d[j] = __ldcs(reinterpret_cast<const ulonglong2*>(dsrc[j] + k * dstep[j]));
        if (aggq) {
            xn[0] = __ldcs(reinterpret_cast<const ulonglong2*>(c0p + k * A.c0_gs));
            if (A.agg == 2) {
                const u64* ap = o0 + k * ostep;
                xn[1] = *reinterpret_cast<const ulonglong2*>(ap);
                xn[2] = *reinterpret_cast<const ulonglong2*>(ap + o1off);
            }
        }
    };
    ulonglong2 dn[PARTS];
    if (g0 < g1) fetch(0, dn);
    mbar_wait(bar, 0);
    // the key tile is reused for every ciphertext of the wave: split it into Karatsuba halves ONCE, in place
    for (int i = threadIdx.x; i < PARTS * 2 * IPT / 2; i += 256) {
        ulonglong2* kp = reinterpret_cast<ulonglong2*>(keys) + i;
        const ulonglong2 v = *kp;
        *kp = make_ulonglong2(kara_pack(v.x), kara_pack(v.y));
    }
    __syncthreads();
    for (int g = g0; g < g1; ++g) {
        ulonglong2 d[PARTS];
#pragma unroll
        for (int j = 0; j < PARTS; ++j) d[j] = dn[j];
        ulonglong2 xc[3];
        if (AGG) {
            xc[0] = xn[0];
            xc[1] = xn[1];
            xc[2] = xn[2];
        }
        if (g + 1 < g1) fetch(1, dn);
        u64* o1 = o0 + o1off;
        KaraAcc b0, b1, a0, a1;
#pragma unroll
        for (int j = 0; j < PARTS; ++j) {
            const ulonglong2 kb = *reinterpret_cast<const ulonglong2*>(keys + (size_t)(2 * j) * IPT + pt);
            const ulonglong2 ka = *reinterpret_cast<const ulonglong2*>(keys + (size_t)(2 * j + 1) * IPT + pt);
            const Kara3 dx(d[j].x), dy(d[j].y);
            b0.mac(dx, Kara3(kb.x, Kara3::Packed{}));
            b1.mac(dy, Kara3(kb.y, Kara3::Packed{}));
            a0.mac(dx, Kara3(ka.x, Kara3::Packed{}));
            a1.mac(dy, Kara3(ka.y, Kara3::Packed{}));
        }
        if (aggq) {  // block-uniform: x' (< 3q) + c0 (+ running sum) < 5q, canonicalised once
            u64 v0 = b0.reduce_lazy3(lc) + xc[0].x, v1 = b1.reduce_lazy3(lc) + xc[0].y;
            u64 w0 = a0.reduce_lazy3(lc), w1 = a1.reduce_lazy3(lc);
            if (A.agg == 2) {
                v0 += xc[1].x;
                v1 += xc[1].y;
                w0 += xc[2].x;
                w1 += xc[2].y;
            }
            *reinterpret_cast<ulonglong2*>(o0) = make_ulonglong2(canon_any16(v0, lc), canon_any16(v1, lc));
            *reinterpret_cast<ulonglong2*>(o1) = make_ulonglong2(canon_any16(w0, lc), canon_any16(w1, lc));
        } else if (PPQ_IP_LAZY && r < A.ell) {  // block-uniform
            *reinterpret_cast<ulonglong2*>(o0) = make_ulonglong2(b0.reduce_lazy3(lc), b1.reduce_lazy3(lc));
            *reinterpret_cast<ulonglong2*>(o1) = make_ulonglong2(a0.reduce_lazy3(lc), a1.reduce_lazy3(lc));
        } else {
            *reinterpret_cast<ulonglong2*>(o0) = make_ulonglong2(b0.reduce(lc), b1.reduce(lc));
            *reinterpret_cast<ulonglong2*>(o1) = make_ulonglong2(a0.reduce(lc), a1.reduce(lc));
        }
#pragma unroll
        for (int j = 0; j < PARTS; ++j) dsrc[j] += dstep[j];
        if (aggq) c0p += A.c0_gs;
        o0 += ostep;
    }
}

// Inner product of NB clients' digits with their keys for the deferred-ModDown aggregator (ppq_agg_add_many).
// Same tiling as k_ip_tma (one CTA = limb r x IPT points, key tiles by TMA bulk copies, reused for every slot), but the
// CTA holds the tiles of NB clients and walks (slot, client): the Q-limb results of a slot's NB clients (+ their c0) are
// summed in registers and the slot's running EVAL sum is read and written ONCE per NB clients -- the read-modify-write
// of the sum was 40 % of the kernel's HBM traffic with one client per launch, and the kernel is HBM-bound.
#ifndef PPQ_IP_AGG_NB_MINB
#define PPQ_IP_AGG_NB_MINB 2
#endif
// (L2 prefetch of the next slots' lines: no effect after the address hoisting either, 93.6-94.2 vs 93.9 us/ct; removed)
template <int PARTS, int NB>
__global__ void __launch_bounds__(256, PARTS >= 7 ? 1 : PPQ_IP_AGG_NB_MINB) k_ip_agg(KCtx c, IpArgs A, IpMulti X, int n_g, int g_per_cta) {
    static_assert(IPC == 1, "k_ip_agg walks one 512-point step per (slot, client)");
    extern __shared__ __align__(128) unsigned char ip_smem[];
    u64* keys = reinterpret_cast<u64*>(ip_smem);  // [NB][PARTS][2][IPT]
    u64* bar = keys + (size_t)NB * PARTS * 2 * IPT;
    const u32 N = 1u << c.logN;
    const int r = blockIdx.y;
    const int M = A.ell + A.K;
    const int limb = r < A.ell ? r : A.L + (r - A.ell);
    const size_t off = (size_t)blockIdx.x * IPT;
    const LimbC lc = c.limbs[limb];
    const bool isq = r < A.ell;  // block-uniform
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, (u32)(NB * PARTS * 2 * IPT * sizeof(u64)));
#pragma unroll
        for (int cl = 0; cl < NB; ++cl)
#pragma unroll
            for (int j = 0; j < PARTS; ++j) {
                const size_t koff = ((size_t)j * (isq ? X.key_q_stride[cl] : A.L + A.K) + limb) * N + off;
                u64* kt = keys + ((size_t)cl * PARTS + j) * 2 * IPT;
                tma_bulk_g2s(kt, (isq ? X.key_qb[cl] : X.key_b[cl]) + koff, IPT * sizeof(u64), bar);
                tma_bulk_g2s(kt + IPT, (isq ? X.key_qa[cl] : X.key_a[cl]) + koff, IPT * sizeof(u64), bar);
            }
    }
    const int own = isq ? r / A.alpha : -1;
    const int g0 = blockIdx.z * g_per_cta;
    const int g1 = min(n_g, g0 + g_per_cta);
    const u32 pt = threadIdx.x * 2;
    // Address arithmetic is hoisted: the thread-invariant part of every stream is formed once, the slot offsets run along
    // with the loop and the client index is a compile-time constant after unrolling (ncu: 379 instructions per
    // iteration with every 64-bit address recomputed inside the loop, most of them integer multiplies for addresses).
    const size_t in_row = (size_t)r * N + off + pt;
    const u64* dig_j[PARTS];
#pragma unroll
    for (int j = 0; j < PARTS; ++j) dig_j[j] = A.digits + ((size_t)j * M + r) * N + off + pt;
    const size_t cs_dig = (size_t)n_g * A.dig_gs, own_row = (size_t)A.ell * N + in_row;
    size_t go_dig = (size_t)g0 * A.dig_gs, go_ct = (size_t)g0 * A.c1_gs, go_acc = (size_t)g0 * A.acc_gs,
           go_ext = (size_t)g0 * A.extp_gs + (size_t)(isq ? 0 : r - A.ell) * N + off + pt;
    ulonglong2 xn[3];
    auto fetch = [&](int cl, size_t gd, size_t gc, size_t ga, ulonglong2 (&d)[PARTS]) {
        const u64* cin = X.ct_in[cl] + gc;
#pragma unroll
        for (int j = 0; j < PARTS; ++j) {
            const u64* dp = (j == own) ? cin + own_row : dig_j[j] + (size_t)cl * cs_dig + gd;
            d[j] = __ldcs(reinterpret_cast<const ulonglong2*>(dp));
        }
        if (isq) {
            xn[0] = __ldcs(reinterpret_cast<const ulonglong2*>(cin + in_row));
            if (cl == 0 && A.agg == 2) {
                const u64* ap = A.acc + ga + in_row;
                xn[1] = *reinterpret_cast<const ulonglong2*>(ap);
                xn[2] = *reinterpret_cast<const ulonglong2*>(ap + (size_t)A.ell * N);
            }
        }
    };
    ulonglong2 dn[PARTS];
    if (g0 < g1) fetch(0, go_dig, go_ct, go_acc, dn);
    mbar_wait(bar, 0);
    for (int i = threadIdx.x; i < NB * PARTS * 2 * IPT / 2; i += 256) {  // Karatsuba halves, once, in place
        ulonglong2* kp = reinterpret_cast<ulonglong2*>(keys) + i;
        const ulonglong2 v = *kp;
        *kp = make_ulonglong2(kara_pack(v.x), kara_pack(v.y));
    }
    __syncthreads();
    // Q limbs: the products of TWO clients go into one set of Karatsuba sums (2 PARTS <= 7 terms of canonical operands
    // keep every partial sum and the Barrett operand in range), so a pair of clients costs four reductions per two
    // points instead of eight.  P limbs: every client's result is needed on its own (the ModDown lift is not linear).
    const bool pair = isq && NB >= 2 && 2 * PARTS <= 7;  // block-uniform
    u64 v0 = 0, v1 = 0, w0 = 0, w1 = 0;
    KaraAcc b0, b1, a0, a1;
    for (int g = g0; g < g1; ++g) {
#pragma unroll
        for (int cl = 0; cl < NB; ++cl) {
            ulonglong2 d[PARTS], xc[3];
#pragma unroll
            for (int j = 0; j < PARTS; ++j) d[j] = dn[j];
            xc[0] = xn[0];
            xc[1] = xn[1];
            xc[2] = xn[2];
            if (cl + 1 < NB) fetch(cl + 1, go_dig, go_ct, go_acc, dn);
            else if (g + 1 < g1) fetch(0, go_dig + A.dig_gs, go_ct + A.c1_gs, go_acc + A.acc_gs, dn);
            const u64* kt = keys + (size_t)cl * PARTS * 2 * IPT + pt;
            if (!pair || (cl & 1) == 0) {
                b0 = KaraAcc();
                b1 = KaraAcc();
                a0 = KaraAcc();
                a1 = KaraAcc();
            }
#pragma unroll
            for (int j = 0; j < PARTS; ++j) {
                const ulonglong2 kb = *reinterpret_cast<const ulonglong2*>(kt + (size_t)(2 * j) * IPT);
                const ulonglong2 ka = *reinterpret_cast<const ulonglong2*>(kt + (size_t)(2 * j + 1) * IPT);
                const Kara3 dx(d[j].x), dy(d[j].y);
                b0.mac(dx, Kara3(kb.x, Kara3::Packed{}));
                b1.mac(dy, Kara3(kb.y, Kara3::Packed{}));
                a0.mac(dx, Kara3(ka.x, Kara3::Packed{}));
                a1.mac(dy, Kara3(ka.y, Kara3::Packed{}));
            }
            if (isq) {
                if (cl == 0) {
                    const bool have = A.agg == 2;
                    v0 = have ? xc[1].x : 0ull;
                    v1 = have ? xc[1].y : 0ull;
                    w0 = have ? xc[2].x : 0ull;
                    w1 = have ? xc[2].y : 0ull;
                }
                v0 += xc[0].x;  // c0 of this client
                v1 += xc[0].y;
                if (!pair || (cl & 1)) {  // running sum (< q) + NB c0 + NB/2 (or NB) lazy results (< 3q each) < 16q
                    v0 += b0.reduce_lazy3(lc);
                    v1 += b1.reduce_lazy3(lc);
                    w0 += a0.reduce_lazy3(lc);
                    w1 += a1.reduce_lazy3(lc);
                }
                if (cl == NB - 1) {
                    u64* o0 = A.acc + go_acc + in_row;
                    *reinterpret_cast<ulonglong2*>(o0) = make_ulonglong2(canon_any16(v0, lc), canon_any16(v1, lc));
                    *reinterpret_cast<ulonglong2*>(o0 + (size_t)A.ell * N) = make_ulonglong2(canon_any16(w0, lc), canon_any16(w1, lc));
                }
            } else {
                u64* o0 = A.extp + (size_t)cl * X.extp_cs + go_ext;
                *reinterpret_cast<ulonglong2*>(o0) = make_ulonglong2(b0.reduce(lc), b1.reduce(lc));
                *reinterpret_cast<ulonglong2*>(o0 + (size_t)A.K * N) = make_ulonglong2(a0.reduce(lc), a1.reduce(lc));
            }
        }
        go_dig += A.dig_gs;
        go_ct += A.c1_gs;
        go_acc += A.acc_gs;
        go_ext += A.extp_gs;
    }
}

// Forward ROW pass + ModDown combine: out = (x - conv) * P^-1 (+ c0 on poly 0) (+ running sum).
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, S == 7 ? 7 : PPQ_ROW_MINB) k_fwd_row_moddown(KCtx c, MdArgs A, int n_g) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const int y = blockIdx.z;
    const int poly = y / A.ell, i = y % A.ell;
    const LimbC lc = c.limbs[i];
    const ulonglong2* tw = c.twf + (size_t)i * N;
    const size_t off = (size_t)y * N + (size_t)row * n2;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const u32 gend = min((u32)n_g, (blockIdx.y + 1) * PPQ_ROW_GPC);
    for (u32 g = blockIdx.y * PPQ_ROW_GPC; g < gend; ++g) {
    u64 a[8];
    const u64* in = A.conv + g * A.conv_gs + off;
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = ld_stream(in + point_of<S - 3>(u, k));
    u64* o = A.out + g * A.out_gs + off + 8 * u;
    prefetch_l2(A.extq + g * A.extq_gs + off + 8 * u);
    if (poly == 0 && A.ct_in) prefetch_l2(A.ct_in + g * A.in_gs + off + 8 * u);
    if (A.accumulate) prefetch_l2(o);
    fwd_pass<S, RowPos<S>, FWD_ROW_IN>(a, sm, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, PPQ_ROW_GPC > 1, PPQ_Z(lc));
    u64 x[8];
    ld8(x, A.extq + g * A.extq_gs + off + 8 * u);
    // keys and conversion constants carry P^-1 already (launch_scale_key, build_ks_plan): (x - conv) P^-1 = x' - conv'.
    // x' + 8q - conv' (+ c0) (+ running sum) < 11q, canonicalised once.
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = x[k] + (lc.q2 << 2) - fold8_b<FwdOut<S, FWD_ROW_IN>::B>(a[k], lc);
    if (poly == 0 && A.ct_in) {
        ld8(x, A.ct_in + g * A.in_gs + off + 8 * u);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] += x[k];
    }
    if (A.accumulate) {
        ld8(x, o);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] += x[k];
    }
    if (lc.sh >= 28u) {  // block-uniform
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = canon16_q(a[k], lc);
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = canon_b<16>(a[k], lc);
    }
    st8(o, a);
    }
}

// The same over pairs of ciphertexts (see k_fwd_row_plain2): twiddles shared by the two transforms.
// Measured on B200, C3: 19.1 -> 17.6 us/ct at 5 CTAs per SM (96 registers); 4 CTAs 18.2, 6 CTAs 18.7.
#ifndef PPQ_MD_X2
#define PPQ_MD_X2 1
#endif
#ifndef PPQ_MD_X2_MINB
#define PPQ_MD_X2_MINB 5
#endif
template <int S>
__device__ __forceinline__ void moddown_epilogue(u64 (&a)[8], const MdArgs& A, u32 g, size_t off, u32 u, int poly,
                                                 const LimbC& lc) {
    u64* o = A.out + g * A.out_gs + off + 8 * u;
    u64 x[8];
    ld8(x, A.extq + g * A.extq_gs + off + 8 * u);
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = x[k] + (lc.q2 << 2) - fold8_b<FwdOut<S, FWD_ROW_IN>::B>(a[k], lc);
    if (poly == 0 && A.ct_in) {
        ld8(x, A.ct_in + g * A.in_gs + off + 8 * u);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] += x[k];
    }
    if (A.accumulate) {
        ld8(x, o);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] += x[k];
    }
    if (lc.sh >= 28u) {  // block-uniform
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = canon16_q(a[k], lc);
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = canon_b<16>(a[k], lc);
    }
    st8(o, a);
}
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, PPQ_MD_X2_MINB) k_fwd_row_moddown2(KCtx c, MdArgs A, int n_g) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[2 * G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const int y = blockIdx.z;
    const int poly = y / A.ell, i = y % A.ell;
    const LimbC lc = c.limbs[i];
    const ulonglong2* tw = c.twf + (size_t)i * N;
    const size_t off = (size_t)y * N + (size_t)row * n2;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    const u32 gend = min((u32)n_g, (blockIdx.y + 1) * (2 * PPQ_ROW_X2_PAIRS));
    for (u32 g = blockIdx.y * (2 * PPQ_ROW_X2_PAIRS); g < gend; g += 2) {
        const bool two = g + 1 < gend;  // block-uniform
        const u32 g1 = two ? g + 1 : g;
        u64 a[8], b[8];
        const u64* in0 = A.conv + g * A.conv_gs + off;
        const u64* in1 = A.conv + g1 * A.conv_gs + off;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            a[k] = ld_stream(in0 + point_of<S - 3>(u, k));
            b[k] = ld_stream(in1 + point_of<S - 3>(u, k));
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const u32 ge = e ? g1 : g;
            prefetch_l2(A.extq + ge * A.extq_gs + off + 8 * u);
            if (poly == 0 && A.ct_in) prefetch_l2(A.ct_in + ge * A.in_gs + off + 8 * u);
            if (A.accumulate) prefetch_l2(A.out + ge * A.out_gs + off + 8 * u);
        }
        fwd_pass_x2<S, RowPos<S>, FWD_ROW_IN>(a, b, sm, (u32)G::SMEM, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, true, PPQ_Z(lc));
        moddown_epilogue<S>(a, A, g, off, u, poly, lc);
        if (two) moddown_epilogue<S>(b, A, g1, off, u, poly, lc);  // never twice on one ciphertext: it may accumulate
    }
}

// Forward ROW pass + rescale combine (DropLastElementAndScale, then the EvalMult constant folded in).
template <int S>
__global__ void __launch_bounds__(RowGeom<S>::THREADS, PPQ_ROW_MINB) k_fwd_row_rescale(KCtx c, RsArgs A) {
    typedef RowGeom<S> G;
    __shared__ u64 sm[G::SMEM];
    const u32 N = 1u << c.logN, n2 = 1u << S, n1 = N >> S;
    const u32 w = threadIdx.x / G::TPT, u = threadIdx.x % G::TPT;
    const u32 row = blockIdx.x * G::R + w;
    const int y = blockIdx.z;
    const u32 g = blockIdx.y;
    const int lm1 = A.ell - 1;
    const int poly = y / lm1, i = y % lm1;
    const LimbC lc = c.limbs[i];
    const ulonglong2* tw = c.twf + (size_t)i * N;
    const size_t roff = (size_t)row * n2;
    RowPos<S> pos{w * (u32)RowPos<S>::PITCH};
    u64 a[8];
    const u64* in = A.conv + g * A.conv_gs + (size_t)y * N + roff;
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = ld_stream(in + point_of<S - 3>(u, k));
    fwd_pass<S, RowPos<S>, FWD_ROW_IN>(a, sm, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, false, PPQ_Z(lc));
    u64 x[8];
    ld8(x, A.ct_in + g * A.in_gs + ((size_t)poly * A.ell + i) * N + roff + 8 * u);
    const u64 cs = A.cst[i], css = A.cst_sh[i];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        // the exact Shoup product takes any 64-bit input: fold to < 8q only, x + 8q - v < 9q
        const u64 v = fold8_b<FwdOut<S, FWD_ROW_IN>::B>(a[k], lc);
        a[k] = mul_shoup(x[k] + (lc.q2 << 2) - v, cs, css, lc.q);
    }
    st8(A.out + g * A.out_gs + (size_t)y * N + roff + 8 * u, a);
}

#ifndef PPQ_NTT_TILE
#define PPQ_NTT_TILE 0  // EXPERIMENT, not compiled by default.  Measured on B200 (profiles/r10_notes.md): bit-exact, but
                        // 8-20 % SLOWER than the two-kernel form at N = 2^12..2^14 (1.22-1.28 vs 1.33-1.58 TB/s of
                        // algorithmic bytes): the transform is bound by integer issue, not by HBM, so halving the
                        // traffic buys nothing, and one 512-thread CTA per SM (128 KiB tile) has fewer warps in flight.
#endif
#if PPQ_NTT_TILE
// =========================================================================================================
// Single-CTA transforms for N <= 2^14 (plain batched NTT API: ppq_ntt_forward / ppq_ntt_inverse, and through it the
// client path and key generation).  The limb (<= 128 KiB) lives in shared memory as an [n1][n2] tile between the COL
// and the ROW pass, so a transform crosses HBM once (16 N bytes) instead of twice; the passes are the same register
// radix-8 groups as the two-kernel form, exchanging IN PLACE in the tile (a thread always writes the words it read).
// Tile layout: row pitch n2 + n2/8 words, column c at c + c/8: conflict free for the COL pass (a warp = 32 adjacent
// columns of one row set) and for the ROW pass (RowPos' padding).
// =========================================================================================================
template <int S2>
struct TileColPos {
    static constexpr bool WARP_LOCAL = false;
    static constexpr u32 RP = (1u << S2) + (1u << (S2 - 3));
    u32 cp;  // padded column
    __device__ __forceinline__ u32 operator()(u32 p) const { return p * RP + cp; }
};
template <int S2>
struct TileRowPos {
    static constexpr bool WARP_LOCAL = (S2 - 3) <= 5;
    static constexpr u32 RP = (1u << S2) + (1u << (S2 - 3));
    u32 base;  // row * RP
    __device__ __forceinline__ u32 operator()(u32 p) const { return base + p + (p >> 3); }
};
constexpr int TILE_THREADS = 512;

template <int S1, int S2>
__global__ void __launch_bounds__(TILE_THREADS, 1) k_ntt_fwd_tile(KCtx c, u64* __restrict__ data, const int* __restrict__ limb_ids) {
    extern __shared__ __align__(16) u64 tile[];
    constexpr u32 n1 = 1u << S1, n2 = 1u << S2, N = n1 * n2;
    constexpr u32 TPC = 1u << (S1 - 3), W = TILE_THREADS / TPC;  // threads per column, columns per step
    constexpr u32 TPR = 1u << (S2 - 3), R = TILE_THREADS / TPR;  // threads per row, rows per step
    static_assert(W <= n2 && R <= n1, "tile kernel geometry");
    const int limb = limb_ids[blockIdx.x];
    if (limb < 0) return;
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    u64* p = data + (size_t)blockIdx.x * N;
    {   // COL pass: n2 / W steps of W columns
        const u32 w = threadIdx.x % W, u = threadIdx.x / W;
        for (u32 c0 = 0; c0 < n2; c0 += W) {
            const u32 col = c0 + w;
            u64 a[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = ld_stream(p + (size_t)point_of<S1 - 3>(u, k) * n2 + col);
            TileColPos<S2> pos{col + (col >> 3)};
            fwd_pass<S1, TileColPos<S2>, FWD_COL_IN>(a, tile, pos, u, tw, 1u, lc.q, lc.q2 << 1, false, PPQ_Z(lc));
#pragma unroll
            for (int k = 0; k < 8; ++k) tile[pos(8 * u + k)] = a[k];  // lazy (< 16q), read back by the ROW pass
        }
    }
    __syncthreads();
    {   // ROW pass: n1 / R steps of R rows
        const u32 wr = threadIdx.x / TPR, u = threadIdx.x % TPR;
        for (u32 r0 = 0; r0 < n1; r0 += R) {
            const u32 row = r0 + wr;
            TileRowPos<S2> pos{row * TileRowPos<S2>::RP};
            u64 a[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = tile[pos(point_of<S2 - 3>(u, k))];
            fwd_pass<S2, TileRowPos<S2>, FWD_ROW_IN>(a, tile, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, false, PPQ_Z(lc));
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = canon_any16(a[k], lc);
            st8(p + (size_t)row * n2 + 8 * u, a);
        }
    }
}

template <int S1, int S2>
__global__ void __launch_bounds__(TILE_THREADS, 1) k_ntt_inv_tile(KCtx c, u64* __restrict__ data, const int* __restrict__ limb_ids,
                                                                   const u64* __restrict__ scale) {
    extern __shared__ __align__(16) u64 tile[];
    constexpr u32 n1 = 1u << S1, n2 = 1u << S2, N = n1 * n2;
    constexpr u32 TPC = 1u << (S1 - 3), W = TILE_THREADS / TPC;
    constexpr u32 TPR = 1u << (S2 - 3), R = TILE_THREADS / TPR;
    const int limb = limb_ids[blockIdx.x];
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twi + (size_t)limb * N;
    u64* p = data + (size_t)blockIdx.x * N;
    {   // ROW pass (canonical input, < 8q out)
        const u32 wr = threadIdx.x / TPR, u = threadIdx.x % TPR;
        for (u32 r0 = 0; r0 < n1; r0 += R) {
            const u32 row = r0 + wr;
            TileRowPos<S2> pos{row * TileRowPos<S2>::RP};
            u64 a[8];
            ld8(a, p + (size_t)row * n2 + 8 * u);
            inv_pass<S2, TileRowPos<S2>, INV_ROW_IN>(a, tile, pos, u, tw, n1 + row, lc.q, lc.q2 << 1, false);
#pragma unroll
            for (int k = 0; k < 8; ++k) tile[pos(point_of<S2 - 3>(u, k))] = a[k];
        }
    }
    __syncthreads();
    {   // COL pass; the output factor (N^-1) rides on its last stage
        const u32 w = threadIdx.x % W, u = threadIdx.x / W;
        const ulonglong2 fs = __ldg(reinterpret_cast<const ulonglong2*>(scale) + 2 * blockIdx.x);
        const ulonglong2 fw = __ldg(reinterpret_cast<const ulonglong2*>(scale) + 2 * blockIdx.x + 1);
        for (u32 c0 = 0; c0 < n2; c0 += W) {
            const u32 col = c0 + w;
            TileColPos<S2> pos{col + (col >> 3)};
            u64 a[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = tile[pos(8 * u + k)];
            inv_pass<S1, TileColPos<S2>, INV_COL_IN, true>(a, tile, pos, u, tw, 1u, lc.q, lc.q2 << 1, false, fs, fw);
#pragma unroll
            for (int k = 0; k < 8; ++k) p[(size_t)point_of<S1 - 3>(u, k) * n2 + col] = canon4(a[k], lc);
        }
    }
}

#endif  // PPQ_NTT_TILE

#ifndef PPQ_FWD_CLUSTER
#define PPQ_FWD_CLUSTER 0
#endif
#if PPQ_FWD_CLUSTER
// =========================================================================================================
// EXPERIMENT (build with -DPPQ_FWD_CLUSTER=1, run with PPQ_FWD_CLUSTER=1): forward NTT of one N = 2^16 limb in ONE kernel by an 8-CTA thread-block
// cluster.  CTA r runs the COL pass of columns [32r, 32r+32) (two steps of k_fwd_col2's 16 columns), writes every
// result straight into the shared memory of the CTA that owns its row (rows [32r', 32r'+32) live in CTA r':
// st.shared::cluster through DSMEM), the cluster synchronises, and each CTA runs the ROW pass of its 32 rows out of its
// own tile.  The limb crosses HBM once (16 N bytes) instead of twice.
// MEASURED on B200 (profiles/r10_notes.md, profiles/r12_cluster_ncu.md): bit-exact, DRAM traffic halved (14.5 vs 27.6 GB
// per 216 ciphertexts), but 65.8 us/ct against 37.8 for the two kernels -- 23 % warps active at 122 registers and 106 KiB
// per CTA, barrier + long_scoreboard 37 % of the stall samples, 2x the shared-memory bank conflicts, fmaheavy 45 %.
// Not compiled by default.
// =========================================================================================================
__device__ __forceinline__ u32 cluster_ctarank() {
    u32 r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ u32 mapa_shared(u32 saddr, u32 rank) {
    u32 r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_cluster_u64(u32 addr, u64 v) {
    asm volatile("st.shared::cluster.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
struct ClusterRowPos {  // one 256-point row of the tile, RowPos' padding
    static constexpr bool WARP_LOCAL = true;
    static constexpr u32 RP = 288;
    u32 base;
    __device__ __forceinline__ u32 operator()(u32 p) const { return base + p + (p >> 3); }
};
constexpr int CL_TILE_WORDS = 32 * 288, CL_XCH_WORDS = 2 * (8 << 8);
template <bool LAZY2>
__global__ void __cluster_dims__(8, 1, 1) __launch_bounds__(256, 2)
    k_fwd_ntt_cluster(KCtx c, u64* __restrict__ data, size_t gs, const int* __restrict__ limb_ids) {
    extern __shared__ __align__(16) u64 cl_smem[];
    u64* tile = cl_smem;
    u64* sm = cl_smem + CL_TILE_WORDS;
    constexpr int S = 8;
    constexpr u32 n2 = 256, N = 65536, W2 = 8;
    const u32 rank = cluster_ctarank();
    const u32 y = blockIdx.y, g = blockIdx.z;
    const int limb = limb_ids[y];
    if (limb < 0) return;  // the same for all CTAs of the cluster
    const LimbC lc = c.limbs[limb];
    const ulonglong2* tw = c.twf + (size_t)limb * N;
    u64* p = data + g * gs + (size_t)y * N;
    {   // COL pass of my 32 columns, results scattered to the row owners
        const u32 w = threadIdx.x % W2, u = threadIdx.x / W2;
        const u32 remote = mapa_shared(smem_u32(tile), u >> 2);  // rows 8u .. 8u+7 belong to CTA u / 4
        for (u32 it = 0; it < 2; ++it) {
            const u32 col = (rank * 16 + it * 8 + w) * 2;
            u64 a[8], b[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const ulonglong2 v = __ldcs(reinterpret_cast<const ulonglong2*>(p + (size_t)point_of<S - 3>(u, k) * n2 + col));
                a[k] = v.x;
                b[k] = v.y;
            }
            fwd_pass_col2<S, W2, FWD_COL_IN>(a, b, sm, w, u, tw, lc.q, lc.q2 << 1, PPQ_Z(lc));
            const u32 cpad = col + (col >> 3);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const u32 off = (8 * (u & 3) + (u32)k) * ClusterRowPos::RP + cpad;
                st_cluster_u64(remote + off * 8u, a[k]);
                st_cluster_u64(remote + (off + 1u) * 8u, b[k]);
            }
            __syncthreads();  // the exchange buffer is reused by the next step
        }
    }
    cluster_sync_all();
    {   // ROW pass of my 32 rows, one row per warp and step
        const u32 wr = threadIdx.x / 32, u = threadIdx.x % 32;
        for (u32 it = 0; it < 4; ++it) {
            const u32 rl = it * 8 + wr, row = rank * 32 + rl;
            ClusterRowPos pos{rl * ClusterRowPos::RP};
            u64 a[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = tile[pos(point_of<S - 3>(u, k))];
            fwd_pass<S, ClusterRowPos, FWD_ROW_IN>(a, tile, pos, u, tw, 256u + row, lc.q, lc.q2 << 1, false, PPQ_Z(lc));
            if (lc.sh >= 28u) {
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = LAZY2 ? canon16_q_lazy2(a[k], lc) : canon16_q(a[k], lc);
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = canon_b<16>(a[k], lc);
            }
            st8(p + (size_t)row * n2 + 8 * u, a);
        }
    }
}

#endif  // PPQ_FWD_CLUSTER

// =========================================================================================================
// pointwise kernels: grid.x covers N/2 (two residues per thread), grid.y = flattened (ct, poly, limb)
// =========================================================================================================
__global__ void k_eval_add(KCtx c, const u64* __restrict__ a, const u64* __restrict__ b, u64* __restrict__ out,
                           int ell) {
    const u32 N = 1u << c.logN;
    const u64 q = c.limbs[blockIdx.y % ell].q;
    const size_t base = (size_t)blockIdx.y * N;
    for (u32 n = (blockIdx.x * blockDim.x + threadIdx.x) * 2; n < N; n += gridDim.x * blockDim.x * 2) {
        const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(a + base + n);
        const ulonglong2 z = *reinterpret_cast<const ulonglong2*>(b + base + n);
        *reinterpret_cast<ulonglong2*>(out + base + n) = make_ulonglong2(addmod(x.x, z.x, q), addmod(x.y, z.y, q));
    }
}
__global__ void k_finalize_mod(KCtx c, u64* __restrict__ x, int ell) {
    const u32 N = 1u << c.logN;
    const LimbC lc = c.limbs[blockIdx.y % ell];
    const size_t base = (size_t)blockIdx.y * N;
    for (u32 n = (blockIdx.x * blockDim.x + threadIdx.x) * 2; n < N; n += gridDim.x * blockDim.x * 2) {
        ulonglong2 v = *reinterpret_cast<ulonglong2*>(x + base + n);
        v.x = reduce64(v.x, lc);
        v.y = reduce64(v.y, lc);
        *reinterpret_cast<ulonglong2*>(x + base + n) = v;
    }
}
__global__ void k_mult_scalar(KCtx c, const u64* __restrict__ in, u64* __restrict__ out, const u64* __restrict__ f,
                              const u64* __restrict__ fs, int ell) {
    const u32 N = 1u << c.logN;
    const int i = blockIdx.y % ell;
    const u64 q = c.limbs[i].q, w = f[i], ws = fs[i];
    const size_t base = (size_t)blockIdx.y * N;
    for (u32 n = (blockIdx.x * blockDim.x + threadIdx.x) * 2; n < N; n += gridDim.x * blockDim.x * 2) {
        const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(in + base + n);
        *reinterpret_cast<ulonglong2*>(out + base + n) =
            make_ulonglong2(mul_shoup(x.x, w, ws, q), mul_shoup(x.y, w, ws, q));
    }
}
__global__ void k_mult_scalar_pack(KCtx c, const u64* __restrict__ in, u64* __restrict__ out, const __grid_constant__ FacPack P, int ell) {
    const u32 N = 1u << c.logN;
    const int i = blockIdx.y % ell;
    const u64 q = c.limbs[i].q, w = P.f[i], ws = P.fs[i];
    const size_t base = (size_t)blockIdx.y * N;
    for (u32 n = (blockIdx.x * blockDim.x + threadIdx.x) * 2; n < N; n += gridDim.x * blockDim.x * 2) {
        const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(in + base + n);
        *reinterpret_cast<ulonglong2*>(out + base + n) =
            make_ulonglong2(mul_shoup(x.x, w, ws, q), mul_shoup(x.y, w, ws, q));
    }
}
// ---- packed residues (the host<->device wire format of the ppq_*_host_packed calls): limb i of a polynomial is a
// little-endian bit stream of N residues of bitlen(q_i) bits each, N*bitlen(q_i)/64 words; polynomial row y of the
// batch starts at word y/ell * lay.words + lay.off[y % ell].  grid (N/512, rows)
__global__ void k_unpack(KCtx c, const u64* __restrict__ in, u64* __restrict__ out, const __grid_constant__ PackLay lay, int ell) {
    const u32 N = 1u << c.logN;
    const int i = blockIdx.y % ell;
    const u32 w = c.limbs[i].sh + 1u;
    const u64 mask = w == 64u ? ~0ull : ((1ull << w) - 1ull);
    const u64* src = in + (size_t)(blockIdx.y / ell) * lay.words + lay.off[i];
    u64* dst = out + (size_t)blockIdx.y * N;
    const u32 last = (N >> 6) * w - 1u;  // last word of the limb
    for (u32 n = (blockIdx.x * blockDim.x + threadIdx.x) * 2; n < N; n += gridDim.x * blockDim.x * 2) {
        u64 r[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const u32 bit = (n + k) * w, wi = bit >> 6, sh = bit & 63u;
            const u64 lo = src[wi];
            const u64 hi = src[wi < last ? wi + 1u : wi];
            r[k] = (sh ? (lo >> sh) | (hi << (64u - sh)) : lo) & mask;
        }
        *reinterpret_cast<ulonglong2*>(dst + n) = make_ulonglong2(r[0], r[1]);
    }
}
// one thread per packed word: the residues that overlap bits [64m, 64m + 64) of the limb's stream.  grid (ceil, rows)
__global__ void k_pack(KCtx c, const u64* __restrict__ in, u64* __restrict__ out, const __grid_constant__ PackLay lay, int ell) {
    const u32 N = 1u << c.logN;
    const int i = blockIdx.y % ell;
    const u32 w = c.limbs[i].sh + 1u;
    const u32 nw = (N >> 6) * w;
    const u64* src = in + (size_t)blockIdx.y * N;
    u64* dst = out + (size_t)(blockIdx.y / ell) * lay.words + lay.off[i];
    for (u32 m = blockIdx.x * blockDim.x + threadIdx.x; m < nw; m += gridDim.x * blockDim.x) {
        const u64 b0 = (u64)m << 6;
        u32 k = (u32)(b0 / w);
        u64 acc = 0;
        for (u64 bit = (u64)k * w; bit < b0 + 64u && k < N; bit += w, ++k) {
            const u64 v = src[k];
            acc |= bit >= b0 ? v << (bit - b0) : v >> (b0 - bit);
        }
        dst[m] = acc;
    }
}
// out[(j*out_stride + i)*N + n] = in[(j*in_stride + i)*N + n] * f[i] mod q_i ; grid (N/512, L, dnum)
__global__ void k_scale_key(KCtx c, const u64* in, int in_stride, u64* out, int out_stride,  // in may alias out
                            const u64* __restrict__ f, const u64* __restrict__ fs) {
    const u32 N = 1u << c.logN;
    const int i = blockIdx.y, j = blockIdx.z;
    const u64 q = c.limbs[i].q, w = f[i], ws = fs[i];
    const u32 n = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (n >= N) return;
    const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(in + ((size_t)j * in_stride + i) * N + n);
    *reinterpret_cast<ulonglong2*>(out + ((size_t)j * out_stride + i) * N + n) =
        make_ulonglong2(mul_shoup(x.x, w, ws, q), mul_shoup(x.y, w, ws, q));
}
// grid.y = g * ell + i
__global__ void k_encrypt_pw(KCtx c, const u64* __restrict__ pkb, const u64* __restrict__ pka,
                             const u64* __restrict__ vee, const u64* __restrict__ pt, u64* __restrict__ ct, int ell) {
    const u32 N = 1u << c.logN;
    const int i = blockIdx.y % ell;
    const size_t g = blockIdx.y / ell;
    const LimbC lc = c.limbs[i];
    const u64* v = vee + ((g * 3 + 0) * ell + i) * N;
    const u64* e0 = vee + ((g * 3 + 1) * ell + i) * N;
    const u64* e1 = vee + ((g * 3 + 2) * ell + i) * N;
    const u64* m = pt ? pt + (g * ell + i) * N : nullptr;  // nullptr: the plaintext is already inside the e0 slot
    const u64* b = pkb + (size_t)i * N;
    const u64* a = pka + (size_t)i * N;
    u64* c0 = ct + ((g * 2 + 0) * ell + i) * N;
    u64* c1 = ct + ((g * 2 + 1) * ell + i) * N;
    for (u32 n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
        const u64 vv = v[n];
        const u64 t = addmod(mulmod_barrett(vv, __ldg(b + n), lc), e0[n], lc.q);
        c0[n] = m ? addmod(t, m[n], lc.q) : t;
        c1[n] = addmod(mulmod_barrett(vv, __ldg(a + n), lc), e1[n], lc.q);
    }
}
// KeyGen pointwise step over the L+K limbs of QP: se = [2][M][N] (s, e in EVAL); pkb = e - a*s; the secret key keeps
// only its Q limbs.  grid.y = limb
__global__ void k_keygen_pw(KCtx c, const u64* __restrict__ se, const u64* __restrict__ a, u64* __restrict__ sk,
                            u64* __restrict__ pkb, int L, int M) {
    const u32 N = 1u << c.logN;
    const int r = blockIdx.y;
    const LimbC lc = c.limbs[r];
    const u64* s = se + (size_t)r * N;
    const u64* e = se + ((size_t)M + r) * N;
    const u64* ar = a + (size_t)r * N;
    u64* br = pkb + (size_t)r * N;
    for (u32 n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
        const u64 sv = s[n];
        br[n] = submod(e[n], mulmod_barrett(__ldg(ar + n), sv, lc), lc.q);
        if (r < L) sk[(size_t)r * N + n] = sv;
    }
}
// ReKeyGen pointwise step: per level-0 digit j and limb r of QP, b_j = pk.b*u + e0 (+ (P mod q_r) * s_old on the limbs
// of digit j), a_j = pk.a*u + e1.  uee = [dnum][3][M][N] (u, e0, e1 in EVAL).  grid.y = j * M + r
__global__ void k_rekeygen_pw(KCtx c, const u64* __restrict__ pkb, const u64* __restrict__ pka,
                              const u64* __restrict__ uee, const u64* __restrict__ sk_old,
                              const u64* __restrict__ pmod, u64* __restrict__ kb, u64* __restrict__ ka, int L, int M,
                              int alpha) {
    const u32 N = 1u << c.logN;
    const int r = blockIdx.y % M, j = blockIdx.y / M;
    const LimbC lc = c.limbs[r];
    const u64* u = uee + (((size_t)j * 3 + 0) * M + r) * N;
    const u64* e0 = uee + (((size_t)j * 3 + 1) * M + r) * N;
    const u64* e1 = uee + (((size_t)j * 3 + 2) * M + r) * N;
    const u64* b = pkb + (size_t)r * N;
    const u64* a = pka + (size_t)r * N;
    u64* ob = kb + (size_t)blockIdx.y * N;
    u64* oa = ka + (size_t)blockIdx.y * N;
    const bool own = r < L && r >= j * alpha && r < (j + 1) * alpha;  // block-uniform
    const u64 pm = own ? pmod[r] : 0;
    const u64* so = sk_old + (size_t)(own ? r : 0) * N;
    for (u32 n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
        const u64 uv = u[n];
        u64 bv = addmod(mulmod_barrett(__ldg(b + n), uv, lc), e0[n], lc.q);
        if (own) bv = addmod(bv, mulmod_barrett(pm, __ldg(so + n), lc), lc.q);
        ob[n] = bv;
        oa[n] = addmod(mulmod_barrett(__ldg(a + n), uv, lc), e1[n], lc.q);
    }
}

// =========================================================================================================
// launchers
// =========================================================================================================
static unsigned long long g_launches = 0;
static inline void count_launch(int n = 1) { g_launches += (unsigned long long)n; }
unsigned long long kernels_launch_count() { return g_launches; }

struct Split {
    int S1, S2;
};
static Split split_of(int logN) {
    switch (logN) {
        case 12: return {6, 6};
        case 13: return {7, 6};
        case 14: return {7, 7};
        case 15: return {8, 7};
        case 16: return {8, 8};
        case 17: return {9, 8};
        default: return {0, 0};
    }
}
template <int S>
struct ColW {  // columns per CTA: 256 threads (512 for S = 9)
    static constexpr int W = (S == 6 ? 32 : S == 7 ? 16 : 8) / PPQ_COL_WDIV;
};

#define PPQ_COL_DISPATCH(S1, CALL)        \
    switch (S1) {                         \
        case 6: { constexpr int S = 6; CALL; } break; \
        case 7: { constexpr int S = 7; CALL; } break; \
        case 8: { constexpr int S = 8; CALL; } break; \
        case 9: { constexpr int S = 9; CALL; } break; \
        default: break;                   \
    }
#define PPQ_ROW_DISPATCH(S2, CALL)        \
    switch (S2) {                         \
        case 6: { constexpr int S = 6; CALL; } break; \
        case 7: { constexpr int S = 7; CALL; } break; \
        case 8: { constexpr int S = 8; CALL; } break; \
        default: break;                   \
    }

static inline int gpc_blocks(int n_g) { return (n_g + PPQ_ROW_GPC - 1) / PPQ_ROW_GPC; }
template <int S, int MODE>
static void fwd_col_launch(const KCtx& c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs, const BcJob* jobs,
                           const u64* chat, const int* limb_ids, u64 q_src, int ny, int ng, cudaStream_t st) {
    count_launch();
    constexpr int W = ColW<S>::W;
    const int n2 = (1 << c.logN) >> S;
#if PPQ_COL2
    if constexpr (MODE == 3 && S == 8) {
        constexpr int W2 = PPQ_COL2_W;
        dim3 grid2(n2 / (2 * W2), ny, ng), block2(W2 << (S - 3));
        const u32 ell = (u32)q_src, approx = (u32)(q_src >> 32);
        const long long* sm_in = reinterpret_cast<const long long*>(src);
        const long long* co = reinterpret_cast<const long long*>(chat);
        if (n2 == (1 << S)) k_fwd_col2_small<S, W2, true><<<grid2, block2, 0, st>>>(c, sm_in, (u32)src_gs, dst, dst_gs, ell, co, approx);
        else k_fwd_col2_small<S, W2, false><<<grid2, block2, 0, st>>>(c, sm_in, (u32)src_gs, dst, dst_gs, ell, co, approx);
        return;
    }
    if constexpr (MODE == 0 && S == 8) {
        if (src == dst && src_gs == dst_gs) {
            constexpr int W2 = PPQ_COL2_W;
#if PPQ_COL2_PF
            if (ng >= 2 * PPQ_COL2_PF) {
                constexpr int SMEM = (2 * (W2 << S) + 16 * (W2 << (S - 3))) * (int)sizeof(u64);
                dim3 gridp(n2 / (2 * W2), ny, (ng + PPQ_COL2_PF - 1) / PPQ_COL2_PF), blockp(W2 << (S - 3));
                static bool attr = false;
                if (!attr) {
                    cudaFuncSetAttribute(k_fwd_col2p<S, W2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
                    cudaFuncSetAttribute(k_fwd_col2p<S, W2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
                    attr = true;
                }
                if (n2 == (1 << S)) k_fwd_col2p<S, W2, true><<<gridp, blockp, SMEM, st>>>(c, dst, dst_gs, limb_ids, ng);
                else k_fwd_col2p<S, W2, false><<<gridp, blockp, SMEM, st>>>(c, dst, dst_gs, limb_ids, ng);
                return;
            }
#endif
            dim3 grid2(n2 / (2 * W2), ny, ng), block2(W2 << (S - 3));
            if (n2 == (1 << S)) k_fwd_col2<S, W2, true><<<grid2, block2, 0, st>>>(c, dst, dst_gs, limb_ids);
            else k_fwd_col2<S, W2, false><<<grid2, block2, 0, st>>>(c, dst, dst_gs, limb_ids);
            return;
        }
    }
#endif
    dim3 grid(n2 / W, ny, ng), block(W << (S - 3));
    if (n2 == (1 << S))
        k_fwd_col<S, W, MODE, true><<<grid, block, 0, st>>>(c, src, src_gs, dst, dst_gs, jobs, chat, limb_ids, q_src);
    else
        k_fwd_col<S, W, MODE, false><<<grid, block, 0, st>>>(c, src, src_gs, dst, dst_gs, jobs, chat, limb_ids, q_src);
}

#if PPQ_NTT_TILE
template <int S1, int S2>
static void ntt_tile_launch(const KCtx& c, u64* data, const int* d_limb_ids, const u64* d_scale4, int n_polys, bool inverse,
                            cudaStream_t st) {
    const size_t smem = (size_t)(1u << S1) * TileRowPos<S2>::RP * sizeof(u64);
    count_launch();
    if (inverse) {
        cudaFuncSetAttribute(k_ntt_inv_tile<S1, S2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_ntt_inv_tile<S1, S2><<<n_polys, TILE_THREADS, smem, st>>>(c, data, d_limb_ids, d_scale4);
    } else {
        cudaFuncSetAttribute(k_ntt_fwd_tile<S1, S2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_ntt_fwd_tile<S1, S2><<<n_polys, TILE_THREADS, smem, st>>>(c, data, d_limb_ids);
    }
}
static bool ntt_tile(const KCtx& c, u64* data, const int* d_limb_ids, const u64* d_scale4, int n_polys, bool inverse,
                     cudaStream_t st) {
    static const bool off = [] { const char* e = std::getenv("PPQ_NTT_TILE"); return e && e[0] == '0'; }();
    if (off) return false;
    switch (c.logN) {
        case 12: ntt_tile_launch<6, 6>(c, data, d_limb_ids, d_scale4, n_polys, inverse, st); return true;
        case 13: ntt_tile_launch<7, 6>(c, data, d_limb_ids, d_scale4, n_polys, inverse, st); return true;
        case 14: ntt_tile_launch<7, 7>(c, data, d_limb_ids, d_scale4, n_polys, inverse, st); return true;
        default: return false;
    }
}
#else
static bool ntt_tile(const KCtx&, u64*, const int*, const u64*, int, bool, cudaStream_t) { return false; }
#endif

void launch_ntt_forward(const KCtx& c, u64* data, const int* d_limb_ids, int n_polys, cudaStream_t st) {
    if (ntt_tile(c, data, d_limb_ids, nullptr, n_polys, false, st)) return;
    count_launch();
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1;
    PPQ_COL_DISPATCH(sp.S1, (fwd_col_launch<S, 0>(c, data, 0, data, 0, nullptr, nullptr, d_limb_ids, 0, n_polys, 1, st)));
    PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_plain<S><<<dim3(n1 / RowGeom<S>::R, 1, n_polys), RowGeom<S>::THREADS, 0, st>>>(
                                c, data, 0, d_limb_ids, 1)));
}

void launch_fwd_row_plain(const KCtx& c, u64* data, size_t gs, const int* d_limb_ids, int n_slots, int n_g,
                          cudaStream_t st, bool lazy2) {
    count_launch();
    (void)lazy2;
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1;
#if PPQ_ROW_X2
    if (n_g > 1 && c.logN >= PPQ_X2_MIN_LOGN) {
        const int by = (n_g + 2 * PPQ_ROW_X2_PAIRS - 1) / (2 * PPQ_ROW_X2_PAIRS);
        if (lazy2) {
            PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_plain2<S, true><<<dim3(n1 / RowGeom<S>::R, by, n_slots), RowGeom<S>::THREADS, 0, st>>>(c, data, gs, d_limb_ids, n_g)));
        } else {
            PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_plain2<S, false><<<dim3(n1 / RowGeom<S>::R, by, n_slots), RowGeom<S>::THREADS, 0, st>>>(c, data, gs, d_limb_ids, n_g)));
        }
        return;
    }
#endif
    PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_plain<S><<<dim3(n1 / RowGeom<S>::R, gpc_blocks(n_g), n_slots), RowGeom<S>::THREADS, 0, st>>>(
                                c, data, gs, d_limb_ids, n_g)));
}

static void inv_row_launch(const KCtx& c, const InvSrc& src, size_t in_gs, u64* out, size_t out_gs, const int* d_limb_map,
                           int n_slots, int n_g, cudaStream_t st) {
    count_launch();
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1;
#if PPQ_INV_X2
    if (n_g > 1 && c.logN >= PPQ_X2_MIN_LOGN) {
        const int by = (n_g + 2 * PPQ_ROW_X2_PAIRS - 1) / (2 * PPQ_ROW_X2_PAIRS);
        PPQ_ROW_DISPATCH(sp.S2, (k_inv_row2<S><<<dim3(n1 / RowGeom<S>::R, by, n_slots), RowGeom<S>::THREADS, 0, st>>>(c, src, in_gs, out, out_gs, d_limb_map, n_g)));
        return;
    }
#endif
    PPQ_ROW_DISPATCH(sp.S2, (k_inv_row<S><<<dim3(n1 / RowGeom<S>::R, gpc_blocks(n_g), n_slots), RowGeom<S>::THREADS, 0, st>>>(
                                c, src, in_gs, out, out_gs, d_limb_map, n_g)));
}
void launch_inv_row(const KCtx& c, const u64* in, size_t in_gs, u64* out, size_t out_gs, const int* d_limb_map,
                    int n_slots, int n_g, cudaStream_t st) {
    InvSrc src{{in, in, in, in}, 0x7fffffffu};
    inv_row_launch(c, src, in_gs, out, out_gs, d_limb_map, n_slots, n_g, st);
}
void launch_inv_row_multi(const KCtx& c, const u64* const* ins, int n_in, int n_per, size_t in_gs, u64* out, size_t out_gs,
                          const int* d_limb_map, int n_slots, cudaStream_t st) {
    InvSrc src{{ins[0], ins[n_in > 1 ? 1 : 0], ins[n_in > 2 ? 2 : 0], ins[n_in > 3 ? 3 : 0]}, (u32)n_per};
    inv_row_launch(c, src, in_gs, out, out_gs, d_limb_map, n_slots, n_in * n_per, st);
}
void launch_decrypt_inv_ntt(const KCtx& c, const u64* ct, const u64* sk, u64* out, int ell, size_t n_ct,
                            const int* d_cyc_ids, const u64* d_scale4, cudaStream_t st) {
    count_launch();
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1;
    PPQ_ROW_DISPATCH(sp.S2, (k_inv_row_decrypt<S><<<dim3(n1 / RowGeom<S>::R, 1, (unsigned)(n_ct * ell)), RowGeom<S>::THREADS, 0, st>>>(
                                c, ct, sk, out, ell)));
    launch_inv_col(c, out, 0, d_cyc_ids, d_scale4, (int)(n_ct * ell), 1, st);
}
void launch_inv_col(const KCtx& c, u64* data, size_t gs, const int* d_limb_map, const u64* d_scale4, int n_slots,
                    int n_g, cudaStream_t st) {
    count_launch();
    const Split sp = split_of(c.logN);
    const int n2 = 1 << sp.S2;
    if (sp.S1 == sp.S2) {
        PPQ_COL_DISPATCH(sp.S1, (k_inv_col<S, ColW<S>::W, true><<<dim3(n2 / ColW<S>::W, n_slots, n_g), ColW<S>::W << (S - 3), 0, st>>>(
                                    c, data, gs, d_limb_map, d_scale4)));
    } else {
        PPQ_COL_DISPATCH(sp.S1, (k_inv_col<S, ColW<S>::W, false><<<dim3(n2 / ColW<S>::W, n_slots, n_g), ColW<S>::W << (S - 3), 0, st>>>(
                                    c, data, gs, d_limb_map, d_scale4)));
    }
}
void launch_ntt_inverse(const KCtx& c, u64* data, const int* d_limb_ids, const u64* d_scale4,
                        int n_polys, cudaStream_t st) {
    if (ntt_tile(c, data, d_limb_ids, d_scale4, n_polys, true, st)) return;
    launch_inv_row(c, data, 0, data, 0, d_limb_ids, n_polys, 1, st);
    launch_inv_col(c, data, 0, d_limb_ids, d_scale4, n_polys, 1, st);
}

void launch_fwd_col_plain(const KCtx& c, u64* data, size_t gs, const int* d_limb_ids, int n_slots, int n_g,
                          cudaStream_t st) {
    const Split sp = split_of(c.logN);
    PPQ_COL_DISPATCH(sp.S1, (fwd_col_launch<S, 0>(c, data, gs, data, gs, nullptr, nullptr, d_limb_ids, 0, n_slots, n_g, st)));
}

void launch_small_fwd_ntt(const KCtx& c, const int64_t* small, u64* out, int n_small, int ell, size_t n_groups,
                          const int* d_cyc_ids, cudaStream_t st, const int64_t* pt_coeff, int pt_approx) {
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1, ny = n_small * ell;
    const size_t gs = (size_t)ny << c.logN;
    const u64 mode_arg = (u64)ell | ((u64)(pt_coeff ? pt_approx : 0) << 32);
    PPQ_COL_DISPATCH(sp.S1, (fwd_col_launch<S, 3>(c, reinterpret_cast<const u64*>(small), (size_t)n_small, out, gs, nullptr,
                                                  reinterpret_cast<const u64*>(pt_coeff), nullptr, mode_arg, ny, (int)n_groups, st)));
    count_launch();
    PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_plain<S><<<dim3(n1 / RowGeom<S>::R, 1, (unsigned)(n_groups * ny)), RowGeom<S>::THREADS, 0, st>>>(
                                c, out, 0, d_cyc_ids, 1)));
}

void launch_small_fwd_encrypt(const KCtx& c, const int64_t* small, u64* vee, int ell, size_t n_ct, const u64* pkb,
                              const u64* pka, const u64* pt, u64* ct, cudaStream_t st, const int64_t* pt_coeff,
                              int pt_approx) {
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1, ny = 3 * ell;
    const size_t gs = (size_t)ny << c.logN;
    const u64 mode_arg = (u64)ell | ((u64)(pt_coeff ? pt_approx : 0) << 32);
    PPQ_COL_DISPATCH(sp.S1, (fwd_col_launch<S, 3>(c, reinterpret_cast<const u64*>(small), (size_t)3, vee, gs, nullptr,
                                                  reinterpret_cast<const u64*>(pt_coeff), nullptr, mode_arg, ny, (int)n_ct, st)));
    count_launch();
    PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_encrypt<S><<<dim3(n1 / RowGeom<S>::R, 1, (unsigned)(n_ct * ell)), RowGeom<S>::THREADS, 0, st>>>(
                                c, vee, pkb, pka, pt, ct, ell)));
}

void launch_sw_fwd_col(const KCtx& c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs, const BcJob* d_jobs,
                       int n_jobs, int n_g, u64 q_src, cudaStream_t st) {
    const Split sp = split_of(c.logN);
    PPQ_COL_DISPATCH(sp.S1, (fwd_col_launch<S, 2>(c, src, src_gs, dst, dst_gs, d_jobs, nullptr, nullptr, q_src, n_jobs, n_g, st)));
}

template <int NS>
static void bconv_launch(const KCtx& c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs, const BcJob* jobs,
                         int n_jobs, const u64* chat, const LimbC* limbs, int n_g, cudaStream_t st) {
    typedef BcPack<NS> P;
    for (int j0 = 0; j0 < n_jobs; j0 += P::MAXJ) {
        P pk;
        pk.n_jobs = n_jobs - j0 < P::MAXJ ? n_jobs - j0 : P::MAXJ;
        pk.src_first = jobs[j0].src_first;
        pk.pad[0] = pk.pad[1] = 0;
        for (int t = 0; t < pk.n_jobs; ++t) {
            const BcJob& jb = jobs[j0 + t];
            const LimbC& lc = limbs[jb.limb];
            typename P::J& J = pk.j[t];
            J.q = lc.q;
            J.mu = lc.mu;
            J.sh = lc.sh;
            J.dst_slot = jb.dst_slot;
            J.pad[0] = J.pad[1] = 0;
            for (int i = 0; i < NS; ++i) J.c[i] = chat_split(chat[jb.chat_off + i]);
        }
        count_launch();
        if (PPQ_BCONV_PAIR && NS <= 8 && (1u << c.logN) >= 256u * PPQ_BCONV_E) {
            dim3 grid((1u << c.logN) / (256 * PPQ_BCONV_E), 1, n_g);
            k_bconv2<(NS <= 8 ? NS : 1)><<<grid, 256, 0, st>>>(c.logN, src, src_gs, dst, dst_gs,
                                                               reinterpret_cast<const BcPack<(NS <= 8 ? NS : 1)>&>(pk));
        } else {
            dim3 grid((1u << c.logN) / 256, 1, n_g);
            k_bconv<NS><<<grid, 256, 0, st>>>(c.logN, src, src_gs, dst, dst_gs, pk);
        }
    }
}
bool launch_bconv(const KCtx& c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs, const BcJob* h_jobs,
                  int n_jobs, const u64* h_chat, const LimbC* h_limbs, int ns, int n_g, cudaStream_t st) {
#define PPQ_BC(NSV) case NSV: bconv_launch<NSV>(c, src, src_gs, dst, dst_gs, h_jobs, n_jobs, h_chat, h_limbs, n_g, st); return true;
    switch (ns) {
        PPQ_BC(1) PPQ_BC(2) PPQ_BC(3) PPQ_BC(4) PPQ_BC(5) PPQ_BC(6) PPQ_BC(7) PPQ_BC(8)
        PPQ_BC(9) PPQ_BC(10) PPQ_BC(11) PPQ_BC(12) PPQ_BC(13) PPQ_BC(14) PPQ_BC(15) PPQ_BC(16)
        PPQ_BC(17) PPQ_BC(18) PPQ_BC(19) PPQ_BC(20) PPQ_BC(21) PPQ_BC(22) PPQ_BC(23) PPQ_BC(24)
        PPQ_BC(25) PPQ_BC(26) PPQ_BC(27) PPQ_BC(28) PPQ_BC(29) PPQ_BC(30) PPQ_BC(31) PPQ_BC(32)
        default: return false;
    }
#undef PPQ_BC
}


template <int NS>
static void bconv_wide_launch(const KCtx& c, const u64* src, size_t src_gs, size_t in_stride, int n_in, u64* dst,
                              size_t dst_gs, int first, const BcJob* jobs, int n_jobs, const u64* chat,
                              const LimbC* limbs, int n_g, cudaStream_t st) {
    typedef BwPack<NS> P;
    constexpr int E = NS <= 8 ? 2 : 1;
    for (int j0 = 0; j0 < n_jobs; j0 += P::MAXJ) {
        P pk;
        pk.n_jobs = n_jobs - j0 < P::MAXJ ? n_jobs - j0 : P::MAXJ;
        pk.src_first = jobs[j0].src_first;
        for (int t = 0; t < pk.n_jobs; ++t) {
            const BcJob& jb = jobs[j0 + t];
            const LimbC& lc = limbs[jb.limb];
            typename P::J& J = pk.j[t];
            J.q = lc.q;
            J.mu = lc.mu;
            J.sh = lc.sh;
            J.dst_slot = jb.dst_slot;
            J.r64 = (u64)((((unsigned __int128)1) << 64) % lc.q);
            J.r64s = (u64)((((unsigned __int128)J.r64) << 64) / lc.q);
            for (int i = 0; i < NS; ++i) {
                const u64 cv = chat[jb.chat_off + i];
                J.c[i] = make_uint2((u32)(cv & 0x3fffffffull), (u32)(cv >> 30));
            }
        }
        dim3 grid((1u << c.logN) / (256 * E), 1, n_g);
        count_launch();
        k_bconv_wide<NS><<<grid, 256, 0, st>>>(c.logN, src, src_gs, in_stride, n_in, dst, dst_gs, first, pk);
    }
}
bool launch_bconv_wide(const KCtx& c, const u64* src, size_t src_gs, size_t in_stride, int n_in, u64* dst,
                       size_t dst_gs, int first, const BcJob* h_jobs, int n_jobs, const u64* h_chat,
                       const LimbC* h_limbs, int ns, int n_g, cudaStream_t st) {
    if (n_in < 1 || n_in > 8) return false;
#define PPQ_BW(NSV) case NSV: bconv_wide_launch<NSV>(c, src, src_gs, in_stride, n_in, dst, dst_gs, first, h_jobs, n_jobs, h_chat, h_limbs, n_g, st); return true;
    switch (ns) {
        PPQ_BW(1) PPQ_BW(2) PPQ_BW(3) PPQ_BW(4) PPQ_BW(5) PPQ_BW(6) PPQ_BW(7) PPQ_BW(8)
        PPQ_BW(9) PPQ_BW(10) PPQ_BW(11) PPQ_BW(12) PPQ_BW(13) PPQ_BW(14) PPQ_BW(15) PPQ_BW(16)
        default: return false;
    }
#undef PPQ_BW
}

bool launch_ip_tma(const KCtx& c, const IpArgs& a, int n_g, cudaStream_t st) {
    const int N = 1 << c.logN;
    if (N % IPT) return false;
    const int tiles = N / IPT, M = a.ell + a.K;
    // enough CTAs for >= 2 waves of 2 CTAs per SM, but as few key-tile reloads as possible
    int gz = 1;
    while (tiles * M * gz < 148 * 4 * (2048 / IPT) && gz < n_g) gz *= 2;
    const int g_per = (n_g + gz - 1) / gz;
    gz = (n_g + g_per - 1) / g_per;
    dim3 grid(tiles, M, gz);
    count_launch();
#define PPQ_IP(P)                                                                                          \
    case P: {                                                                                              \
        const size_t sm = (size_t)P * 2 * IPT * sizeof(u64) + 16;                                          \
        if (a.agg) {                                                                                       \
            cudaFuncSetAttribute(k_ip_tma<P, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm); \
            k_ip_tma<P, true><<<grid, 256, sm, st>>>(c, a, n_g, g_per);                                    \
        } else {                                                                                           \
            cudaFuncSetAttribute(k_ip_tma<P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);\
            k_ip_tma<P, false><<<grid, 256, sm, st>>>(c, a, n_g, g_per);                                   \
        }                                                                                                  \
        return true;                                                                                       \
    }
    switch (a.parts) {
        PPQ_IP(1) PPQ_IP(2) PPQ_IP(3) PPQ_IP(4) PPQ_IP(5) PPQ_IP(6) PPQ_IP(7) PPQ_IP(8)
        default: return false;
    }
#undef PPQ_IP
}
bool launch_ip_agg(const KCtx& c, const IpArgs& a, const IpMulti& m, int n_g, cudaStream_t st) {
    const int N = 1 << c.logN;
    if (N % IPT || m.nb < 1 || m.nb > 4 || m.nb == 3) return false;
    const int tiles = N / IPT, M = a.ell + a.K;
    int gz = 1;
    while (tiles * M * gz < 148 * 4 * (2048 / IPT) && gz < n_g) gz *= 2;
    const int g_per = (n_g + gz - 1) / gz;
    gz = (n_g + g_per - 1) / g_per;
    dim3 grid(tiles, M, gz);
    count_launch();
#define PPQ_IPA(P, NBV)                                                                                        \
    {                                                                                                          \
        const size_t sm = (size_t)NBV * P * 2 * IPT * sizeof(u64) + 16;                                        \
        cudaFuncSetAttribute(k_ip_agg<P, NBV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);          \
        k_ip_agg<P, NBV><<<grid, 256, sm, st>>>(c, a, m, n_g, g_per);                                          \
        return true;                                                                                           \
    }
#define PPQ_IPA_P(P)                       \
    case P:                                \
        if (m.nb == 1) PPQ_IPA(P, 1)       \
        if (m.nb == 2) PPQ_IPA(P, 2)       \
        PPQ_IPA(P, 4)
    switch (a.parts) {
        PPQ_IPA_P(1) PPQ_IPA_P(2) PPQ_IPA_P(3) PPQ_IPA_P(4) PPQ_IPA_P(5) PPQ_IPA_P(6) PPQ_IPA_P(7) PPQ_IPA_P(8)
        default: return false;
    }
#undef PPQ_IPA_P
#undef PPQ_IPA
}
#if PPQ_FWD_CLUSTER
bool launch_fwd_ntt_cluster(const KCtx& c, u64* data, size_t gs, const int* d_limb_ids, int n_slots, int n_g, cudaStream_t st,
                            bool lazy2) {
    if (c.logN != 16) return false;
    const size_t smem = (size_t)(CL_TILE_WORDS + CL_XCH_WORDS) * sizeof(u64);
    dim3 grid(8, n_slots, n_g);
    count_launch();
    if (lazy2) {
        cudaFuncSetAttribute(k_fwd_ntt_cluster<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_fwd_ntt_cluster<true><<<grid, 256, smem, st>>>(c, data, gs, d_limb_ids);
    } else {
        cudaFuncSetAttribute(k_fwd_ntt_cluster<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_fwd_ntt_cluster<false><<<grid, 256, smem, st>>>(c, data, gs, d_limb_ids);
    }
    return true;
}
#else
bool launch_fwd_ntt_cluster(const KCtx&, u64*, size_t, const int*, int, int, cudaStream_t, bool) { return false; }
#endif
void launch_fwd_row_moddown(const KCtx& c, const MdArgs& a, int n_g, cudaStream_t st) {
    count_launch();
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1;
#if PPQ_MD_X2
    if (n_g > 1 && c.logN >= PPQ_X2_MIN_LOGN) {
        const int by = (n_g + 2 * PPQ_ROW_X2_PAIRS - 1) / (2 * PPQ_ROW_X2_PAIRS);
        PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_moddown2<S><<<dim3(n1 / RowGeom<S>::R, by, 2 * a.ell), RowGeom<S>::THREADS, 0, st>>>(c, a, n_g)));
        return;
    }
#endif
    PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_moddown<S><<<dim3(n1 / RowGeom<S>::R, gpc_blocks(n_g), 2 * a.ell), RowGeom<S>::THREADS, 0, st>>>(c, a, n_g)));
}
void launch_fwd_row_rescale(const KCtx& c, const RsArgs& a, int n_g, cudaStream_t st) {
    count_launch();
    const Split sp = split_of(c.logN);
    const int n1 = 1 << sp.S1;
    PPQ_ROW_DISPATCH(sp.S2, (k_fwd_row_rescale<S><<<dim3(n1 / RowGeom<S>::R, n_g, 2 * (a.ell - 1)), RowGeom<S>::THREADS, 0, st>>>(c, a)));
}

static dim3 pw_grid(const KCtx& c, size_t ny, int per_thread) {
    const u32 N = 1u << c.logN;
    u32 gx = N / (256u * per_thread);
    if (gx == 0) gx = 1;
    return dim3(gx, (unsigned)ny, 1);
}
void launch_eval_add(const KCtx& c, const u64* a, const u64* b, u64* out, int ell, size_t n_ct, cudaStream_t st) {
    count_launch();
    k_eval_add<<<pw_grid(c, n_ct * 2 * ell, 2), 256, 0, st>>>(c, a, b, out, ell);
}
void launch_finalize_mod(const KCtx& c, u64* x, int ell, size_t n_ct, cudaStream_t st) {
    count_launch();
    k_finalize_mod<<<pw_grid(c, n_ct * 2 * ell, 2), 256, 0, st>>>(c, x, ell);
}
void launch_mult_scalar(const KCtx& c, const u64* in, u64* out, const u64* d_f, const u64* d_f_sh, int ell,
                        size_t n_ct, cudaStream_t st) {
    count_launch();
    k_mult_scalar<<<pw_grid(c, n_ct * 2 * ell, 2), 256, 0, st>>>(c, in, out, d_f, d_f_sh, ell);
}
void launch_mult_scalar_pack(const KCtx& c, const u64* in, u64* out, const FacPack& p, int ell, size_t n_rows, cudaStream_t st) {
    count_launch();
    k_mult_scalar_pack<<<pw_grid(c, n_rows, 2), 256, 0, st>>>(c, in, out, p, ell);
}
void launch_unpack(const KCtx& c, const u64* in, u64* out, const PackLay& lay, int ell, size_t n_rows, cudaStream_t st) {
    count_launch();
    k_unpack<<<pw_grid(c, n_rows, 2), 256, 0, st>>>(c, in, out, lay, ell);
}
void launch_pack(const KCtx& c, const u64* in, u64* out, const PackLay& lay, int ell, size_t n_rows, cudaStream_t st) {
    count_launch();
    k_pack<<<pw_grid(c, n_rows, 1), 256, 0, st>>>(c, in, out, lay, ell);
}
void launch_mult_scalar_polys(const KCtx& c, const u64* in, u64* out, const u64* d_f, const u64* d_f_sh, int ell,
                              size_t n_polys, cudaStream_t st) {
    count_launch();
    k_mult_scalar<<<pw_grid(c, n_polys, 2), 256, 0, st>>>(c, in, out, d_f, d_f_sh, ell);
}
void launch_scale_key(const KCtx& c, const u64* in, int in_stride, u64* out, int out_stride, const u64* d_pinv,
                      const u64* d_pinv_sh, int L, int dnum, cudaStream_t st) {
    count_launch();
    const u32 N = 1u << c.logN;
    k_scale_key<<<dim3((N / 2 + 255) / 256, L, dnum), 256, 0, st>>>(c, in, in_stride, out, out_stride, d_pinv, d_pinv_sh);
}
void launch_encrypt_pw(const KCtx& c, const u64* pkb, const u64* pka, const u64* vee, const u64* pt, u64* ct, int ell,
                       size_t n_ct, cudaStream_t st) {
    count_launch();
    k_encrypt_pw<<<pw_grid(c, n_ct * ell, 1), 256, 0, st>>>(c, pkb, pka, vee, pt, ct, ell);
}
void launch_keygen_pw(const KCtx& c, const u64* se, const u64* a, u64* sk, u64* pkb, int L, int M, cudaStream_t st) {
    count_launch();
    k_keygen_pw<<<pw_grid(c, M, 1), 256, 0, st>>>(c, se, a, sk, pkb, L, M);
}
void launch_rekeygen_pw(const KCtx& c, const u64* pkb, const u64* pka, const u64* uee, const u64* sk_old,
                        const u64* d_pmod, u64* kb, u64* ka, int L, int M, int alpha, int dnum, cudaStream_t st) {
    count_launch();
    k_rekeygen_pw<<<pw_grid(c, (size_t)dnum * M, 1), 256, 0, st>>>(c, pkb, pka, uee, sk_old, d_pmod, kb, ka, L, M, alpha);
}

const char* kernels_build_info() { return "ppq_b200 kernels: sm_100a, radix-8 register groups, 2-pass (COL,ROW) NTT"; }

}  // namespace ppq
