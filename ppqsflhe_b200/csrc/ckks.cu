// ckks.cu -- CKKS encode/decode kernels (FP64 special FFT).  See ckks.cuh; the samplers live in sampler.cu.
// Client-side path (encryptModelWeights / decryptModelWeights): the stages that fit a shared-memory tile run in one kernel.
#include "ckks.cuh"

#include <cmath>
#include <vector>

namespace ppq {
namespace ckks {

Tables::~Tables() {
    if (rot) cudaFree(rot);
    if (ksi) cudaFree(ksi);
}

bool build_tables(int logN, Tables& t, std::string& err) {
    const size_t N = (size_t)1 << logN, m = 2 * N, Nh = N / 2;
    std::vector<unsigned> rot(Nh);
    std::vector<double2> ksi(m + 1);
    size_t five = 1;
    for (size_t i = 0; i < Nh; ++i) {
        rot[i] = (unsigned)five;
        five = (five * 5) % m;
    }
    for (size_t j = 0; j < m; ++j) {
        const double ang = 2.0 * M_PI * (double)j / (double)m;
        ksi[j] = make_double2(std::cos(ang), std::sin(ang));
    }
    ksi[m] = ksi[0];
    t.logN = logN;
    cudaError_t e = cudaMalloc(&t.rot, Nh * sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMalloc(&t.ksi, (m + 1) * sizeof(double2));
    if (e == cudaSuccess) e = cudaMemcpy(t.rot, rot.data(), Nh * sizeof(unsigned), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t.ksi, ksi.data(), (m + 1) * sizeof(double2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        err = std::string("ckks tables: ") + cudaGetErrorString(e);
        return false;
    }
    return true;
}

__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

// z[g][i] = (vals[g][i], 0) zero padded
__global__ void k_load_vals(const double* __restrict__ vals, size_t n_vals, double2* __restrict__ z, unsigned Nh) {
    const size_t g = blockIdx.y;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < Nh; i += gridDim.x * blockDim.x)
        z[g * Nh + i] = make_double2(i < n_vals ? vals[g * n_vals + i] : 0.0, 0.0);
}
// one stage of FFTSpecialInv (decimation in frequency): len = current block length
__global__ void k_fft_inv_stage(double2* __restrict__ z, unsigned Nh, unsigned len, const unsigned* __restrict__ rot,
                                const double2* __restrict__ ksi, unsigned m) {
    const size_t g = blockIdx.y;
    const unsigned lenh = len >> 1, lenq = len << 2, gap = m / lenq;
    double2* a = z + g * Nh;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < Nh / 2; t += gridDim.x * blockDim.x) {
        const unsigned j = t % lenh, i = (t / lenh) * len;
        const unsigned idx = (lenq - (rot[j] % lenq)) * gap;
        const double2 x = a[i + j], y = a[i + j + lenh];
        a[i + j] = make_double2(x.x + y.x, x.y + y.y);
        a[i + j + lenh] = cmul(make_double2(x.x - y.x, x.y - y.y), ksi[idx]);
    }
}
// one stage of FFTSpecial (decimation in time)
__global__ void k_fft_fwd_stage(double2* __restrict__ z, unsigned Nh, unsigned len, const unsigned* __restrict__ rot,
                                const double2* __restrict__ ksi, unsigned m) {
    const size_t g = blockIdx.y;
    const unsigned lenh = len >> 1, lenq = len << 2, gap = m / lenq;
    double2* a = z + g * Nh;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < Nh / 2; t += gridDim.x * blockDim.x) {
        const unsigned j = t % lenh, i = (t / lenh) * len;
        const unsigned idx = (rot[j] % lenq) * gap;
        const double2 u = a[i + j], v = cmul(a[i + j + lenh], ksi[idx]);
        a[i + j] = make_double2(u.x + v.x, u.y + v.y);
        a[i + j + lenh] = make_double2(u.x - v.x, u.y - v.y);
    }
}
static dim3 grid_for(size_t work, size_t n) {
    unsigned gx = (unsigned)((work + 255) / 256);
    if (gx > 1024) gx = 1024;
    if (gx == 0) gx = 1;
    return dim3(gx, (unsigned)n, 1);
}

// ---- shared-memory FFT blocks.  The stages whose butterflies stay inside a contiguous block of 2^FB_LOG points run
// in ONE kernel on a tile held in shared memory (128 KiB of double2 for 8192 points): a whole transform for N <= 2^14,
// all but log2(Nh) - 13 stages above (N = 2^16: 13 of 15 stages in one launch instead of 13 HBM round trips).
// Round 1 launched one kernel per stage.  The arithmetic per butterfly is unchanged, so the results are the same.
constexpr int FB_LOG = 13;
constexpr int FB_THREADS = 1024;
__device__ __forceinline__ unsigned brev_n(unsigned x, int bits) { return bits ? (__brev(x) >> (32 - bits)) : 0u; }

// FFTSpecialInv stages len = B, B/2, .., 2 on block blockIdx.x of ciphertext blockIdx.y (B = min(Nh, 2^FB_LOG)).
// FROM_VALS: the tile is loaded from the real input vector (zero padded), i.e. no global stage ran before (Nh <= B).
// Then the encode epilogue: co[i] = llround(re(z[brev(i)]) / Nh * sf), co[Nh + i] = the same for the imaginary part.
template <bool FROM_VALS>
__global__ void __launch_bounds__(FB_THREADS)
    k_fft_inv_block(const double* __restrict__ vals, size_t n_vals, const double2* __restrict__ z, long long* __restrict__ co,
                    unsigned Nh, int bits, unsigned logB, const unsigned* __restrict__ rot, const double2* __restrict__ ksi,
                    unsigned m, double sf) {
    extern __shared__ double2 tile[];
    const unsigned B = 1u << logB, blk = blockIdx.x;
    const size_t g = blockIdx.y;
    for (unsigned t = threadIdx.x; t < B; t += FB_THREADS) {
        const unsigned i = blk * B + t;
        tile[t] = FROM_VALS ? make_double2(i < n_vals ? vals[g * n_vals + i] : 0.0, 0.0) : z[g * Nh + i];
    }
    __syncthreads();
    for (unsigned len = B; len >= 2; len >>= 1) {
        const unsigned lenh = len >> 1, lenq = len << 2, gap = m / lenq;
        for (unsigned t = threadIdx.x; t < B / 2; t += FB_THREADS) {
            const unsigned j = t % lenh, i = (t / lenh) * len;
            const unsigned idx = (lenq - (rot[j] % lenq)) * gap;
            const double2 x = tile[i + j], y = tile[i + j + lenh];
            tile[i + j] = make_double2(x.x + y.x, x.y + y.y);
            tile[i + j + lenh] = cmul(make_double2(x.x - y.x, x.y - y.y), ksi[idx]);
        }
        __syncthreads();
    }
    // element j = blk*B + jl goes to coefficient i = brev(j) = brev(jl, logB) << hb | brev(blk, hb); walk i upwards
    const int hb = bits - (int)logB;
    const unsigned ib = brev_n(blk, hb);
    for (unsigned t = threadIdx.x; t < B; t += FB_THREADS) {
        const double2 v = tile[brev_n(t, (int)logB)];
        const unsigned i = (t << hb) | ib;
        co[g * 2 * Nh + i] = llround(v.x / (double)Nh * sf);
        co[g * 2 * Nh + Nh + i] = llround(v.y / (double)Nh * sf);
    }
}

void launch_encode_fft(const Tables& t, const double* d_vals, size_t n_vals, double sf, double2* z, int64_t* co,
                       size_t n, cudaStream_t st) {
    const unsigned N = 1u << t.logN, Nh = N / 2, m = 2 * N;
    const int bits = t.logN - 1;
    const unsigned logB = bits < FB_LOG ? (unsigned)bits : (unsigned)FB_LOG, B = 1u << logB;
    const size_t smem = (size_t)B * sizeof(double2);
    dim3 grid(Nh / B, (unsigned)n, 1);
    if (Nh <= B) {
        cudaFuncSetAttribute(k_fft_inv_block<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_fft_inv_block<true><<<grid, FB_THREADS, smem, st>>>(d_vals, n_vals, nullptr, (long long*)co, Nh, bits, logB, t.rot, t.ksi, m, sf);
        return;
    }
    k_load_vals<<<grid_for(Nh, n), 256, 0, st>>>(d_vals, n_vals, z, Nh);
    for (unsigned len = Nh; len > B; len >>= 1)
        k_fft_inv_stage<<<grid_for(Nh / 2, n), 256, 0, st>>>(z, Nh, len, t.rot, t.ksi, m);
    cudaFuncSetAttribute(k_fft_inv_block<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_fft_inv_block<false><<<grid, FB_THREADS, smem, st>>>(nullptr, 0, z, (long long*)co, Nh, bits, logB, t.rot, t.ksi, m, sf);
}

// FFTSpecial stages len = 2, 4, .., B on a block of the BIT-REVERSED input (k_crt_to_complex stores it that way).
// TAKE_REAL: Nh <= B, the transform ends here: the real parts go straight to the output.
template <bool TAKE_REAL>
__global__ void __launch_bounds__(FB_THREADS)
    k_fft_fwd_block(double2* __restrict__ z, double* __restrict__ out, unsigned Nh, unsigned logB,
                    const unsigned* __restrict__ rot, const double2* __restrict__ ksi, unsigned m) {
    extern __shared__ double2 tile[];
    const unsigned B = 1u << logB, blk = blockIdx.x;
    const size_t g = blockIdx.y;
    for (unsigned t = threadIdx.x; t < B; t += FB_THREADS) tile[t] = z[g * Nh + blk * B + t];
    __syncthreads();
    for (unsigned len = 2; len <= B; len <<= 1) {
        const unsigned lenh = len >> 1, lenq = len << 2, gap = m / lenq;
        for (unsigned t = threadIdx.x; t < B / 2; t += FB_THREADS) {
            const unsigned j = t % lenh, i = (t / lenh) * len;
            const unsigned idx = (rot[j] % lenq) * gap;
            const double2 u = tile[i + j], v = cmul(tile[i + j + lenh], ksi[idx]);
            tile[i + j] = make_double2(u.x + v.x, u.y + v.y);
            tile[i + j + lenh] = make_double2(u.x - v.x, u.y - v.y);
        }
        __syncthreads();
    }
    for (unsigned t = threadIdx.x; t < B; t += FB_THREADS) {
        if (TAKE_REAL) out[g * Nh + blk * B + t] = tile[t].x;
        else z[g * Nh + blk * B + t] = tile[t];
    }
}

// CRT interpolation by Garner's mixed-radix digits, centred, as a double (Poly::CRTInterpolate + the
// double conversion of CKKSPackedEncoding::Decode).  Digits v_i: X = sum v_i * prod_{k<i} q_k.
// X > (Q-1)/2  <=>  digits compare above ((q_i-1)/2)_i from the top; then Q - X has digits q_i-1-v_i, plus 1.
#define PPQ_MAX_DECODE_LIMBS 64
// digits v_0..v_{n_lim-1} of the residues x_i = x[i * N] (stride N), generic limb count: O(n_lim^2) Barrett steps
__device__ __forceinline__ void garner_digits(const KCtx& c, const u64* __restrict__ x, u32 N, int n_lim, int ell,
                                              const u64* __restrict__ ginv, const u64* __restrict__ qmod, u64* v) {
    v[0] = x[0];
    for (int i = 1; i < n_lim; ++i) {
        const LimbC li = c.limbs[i];
        const u64* qm = qmod + (size_t)i * ell;  // q_k mod q_i, k < i (host table: no reduction of constants here)
        // Horner in mixed radix: acc = v_0 + q_0 (v_1 + q_1 (...)) mod q_i, one 128-bit multiply-add and ONE Barrett
        // reduction per step (acc < q_i, q_k mod q_i < q_i, v_k < 2^60: the operand stays below 2^121)
        u64 acc = 0;
        for (int k = i - 1; k >= 0; --k) {
            unsigned __int128 t = (unsigned __int128)acc * qm[k] + v[k];
            acc = barrett128((u64)(t >> 64), (u64)t, li);
        }
        v[i] = mulmod_barrett(submod(x[(size_t)i * N], acc, li.q), ginv[i], li);
    }
}
// centred value of the mixed-radix number (v_i) over the first n_lim moduli, as a double
__device__ __forceinline__ double centred_double(const KCtx& c, const u64* v, int n_lim, bool* neg_out) {
    bool neg = false;
    for (int i = n_lim - 1; i >= 0; --i) {
        const u64 h = (c.limbs[i].q - 1) >> 1;
        if (v[i] != h) {
            neg = v[i] > h;
            break;
        }
    }
    double d = 0.0;
    for (int i = n_lim - 1; i >= 0; --i) {
        const u64 q = c.limbs[i].q;
        const u64 dig = neg ? q - 1 - v[i] : v[i];
        d = d * (double)q + (double)dig;
    }
    *neg_out = neg;
    return neg ? -(d + 1.0) : d;
}
// all ell limbs (the definition); kept out of line so that its 64-word digit array costs the short path nothing
__device__ __noinline__ double crt_full(const KCtx& c, const u64* __restrict__ x, u32 N, int ell,
                                        const u64* __restrict__ ginv, const u64* __restrict__ qmod) {
    u64 v[PPQ_MAX_DECODE_LIMBS];
    garner_digits(c, x, N, ell, ell, ginv, qmod, v);
    bool neg;
    return centred_double(c, v, ell, &neg);
}
// A decrypted CKKS coefficient is small against Q (|m + noise| ~ 2^60..2^120 against Q ~ 2^(60 ell)).  If y, the centred
// value of the residues modulo P3 = q_0 q_1 q_2 alone, ALSO reproduces every other residue, then y is congruent to X
// modulo Q and |y| <= P3 / 2 < Q / 2, i.e. y IS the centred representative the full interpolation returns -- with the
// same three low digits, so the double comes out bit for bit the same (the 17 leading digits contribute exact zeros).
// 3 + 2 (ell - 3) Barrett steps per coefficient instead of ell (ell - 1) / 2; a coefficient that fails the check (it
// never does for a decryption of a valid ciphertext) takes the full path.
constexpr int CRT_FAST = 3;
static_assert(CRT_FAST == 3, "the check below is written for three digits");
__global__ void k_crt_to_complex(KCtx c, const u64* __restrict__ coeff, int ell, const u64* __restrict__ ginv,
                                 const u64* __restrict__ qmod, const u64* __restrict__ p3mod, double sf,
                                 double2* __restrict__ z) {
    const u32 N = 1u << c.logN, Nh = N / 2;
    const size_t g = blockIdx.y;
    const u64* xg = coeff + g * (size_t)ell * N;
    for (u32 n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
        const u64* x = xg + n;
        double d;
        if (ell <= CRT_FAST) {
            d = crt_full(c, x, N, ell, ginv, qmod);
        } else {
            u64 v[CRT_FAST];
            garner_digits(c, x, N, CRT_FAST, ell, ginv, qmod, v);
            bool neg;
            d = centred_double(c, v, CRT_FAST, &neg);
            bool ok = true;
            for (int i = CRT_FAST; i < ell; ++i) {
                const LimbC li = c.limbs[i];
                // Y' mod q_i = v_0 + v_1 (q_0 mod q_i) + v_2 (q_0 q_1 mod q_i): two 120-bit products and v_0 in one 128-bit
                // sum (< 2^122), ONE Barrett reduction
                const unsigned __int128 t = (unsigned __int128)v[1] * qmod[(size_t)i * ell] + (unsigned __int128)v[2] * p3mod[ell + i] + v[0];
                u64 acc = barrett128((u64)(t >> 64), (u64)t, li);
                if (neg) acc = submod(acc, p3mod[i], li.q);  // y = Y' - P3
                ok &= acc == x[(size_t)i * N];
            }
            if (!ok) d = crt_full(c, x, N, ell, ginv, qmod);
        }
        d /= sf;
        // stored at the bit-reversed position: FFTSpecial starts with that permutation (k_fft_fwd_block reads contiguously)
        if (n < Nh) z[g * Nh + (__brev(n) >> (33 - c.logN))].x = d;
        else z[g * Nh + (__brev(n - Nh) >> (33 - c.logN))].y = d;
    }
}
__global__ void k_take_real(const double2* __restrict__ z, double* __restrict__ out, unsigned Nh) {
    const size_t g = blockIdx.y;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < Nh; i += gridDim.x * blockDim.x)
        out[g * Nh + i] = z[g * Nh + i].x;
}

void launch_decode(const Tables& t, const KCtx& k, const u64* coeff, int ell, const u64* d_ginv, const u64* d_qmod,
                   const u64* d_p3mod, double sf, double2* z, double* d_vals, size_t n, cudaStream_t st) {
    const unsigned N = 1u << t.logN, Nh = N / 2, m = 2 * N;
    const int bits = t.logN - 1;
    const unsigned logB = bits < FB_LOG ? (unsigned)bits : (unsigned)FB_LOG, B = 1u << logB;
    const size_t smem = (size_t)B * sizeof(double2);
    k_crt_to_complex<<<grid_for(N, n), 256, 0, st>>>(k, coeff, ell, d_ginv, d_qmod, d_p3mod, sf, z);
    dim3 grid(Nh / B, (unsigned)n, 1);
    if (Nh <= B) {
        cudaFuncSetAttribute(k_fft_fwd_block<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_fft_fwd_block<true><<<grid, FB_THREADS, smem, st>>>(z, d_vals, Nh, logB, t.rot, t.ksi, m);
        return;
    }
    cudaFuncSetAttribute(k_fft_fwd_block<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_fft_fwd_block<false><<<grid, FB_THREADS, smem, st>>>(z, d_vals, Nh, logB, t.rot, t.ksi, m);
    for (unsigned len = 2 * B; len <= Nh; len <<= 1)
        k_fft_fwd_stage<<<grid_for(Nh / 2, n), 256, 0, st>>>(z, Nh, len, t.rot, t.ksi, m);
    k_take_real<<<grid_for(Nh, n), 256, 0, st>>>(z, d_vals, Nh);
}

}  // namespace ckks
}  // namespace ppq
