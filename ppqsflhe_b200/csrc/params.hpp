// params.hpp -- host-side parameter layer of the product library (C++): everything OpenFHE derives when the
// reference deserialises CC.json (server/src/changeCipherDomain.cpp:32-36 -> PrecomputeCRTTables) and that the
// device kernels need as constants.  This is PRODUCT code; it never touches oracle/.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace ppq {
namespace host {

typedef unsigned long long u64;

u64 mulmod(u64 a, u64 b, u64 q);
u64 powmod(u64 a, u64 e, u64 q);
u64 invmod(u64 a, u64 q);
bool is_prime(u64 n);
u64 first_prime(int bits, u64 m);      // OpenFHE nbtheory FirstPrime
u64 previous_prime(u64 q, u64 m);      // OpenFHE nbtheory PreviousPrime
u64 next_prime(u64 q, u64 m);          // OpenFHE nbtheory NextPrime
u64 min_root_of_unity(u64 m, u64 q);   // OpenFHE RootOfUnity: minimum primitive m-th root
u64 shoup(u64 w, u64 q);               // floor(w * 2^64 / q)

struct Params {
    int logN = 0, L = 0, K = 0, dnum = 0, alpha = 0, aux_bits = 60, extra_bits = 0;
    size_t N = 0;
    std::vector<u64> mod;    // q_0..q_{L-1}, p_0..p_{K-1}
    std::vector<u64> root;   // minimal primitive 2N-th roots
    std::vector<u64> ninv;   // N^-1 mod each modulus
    std::vector<double> sf;      // m_scalingFactorsReal
    std::vector<double> sfbig;   // m_scalingFactorsRealBig
    int M() const { return L + K; }
};

// Fills alpha, K, the P primes (HYBRID rule), roots, N^-1 and the FLEXIBLEAUTO(EXT) scaling factors.
bool derive(Params& p, std::string& err);
// FLEXIBLEAUTOEXT prime chain of ckksrns-parametergeneration.cpp for `depth` (L = depth + 2 with extra_bits > 0).
std::vector<u64> synthetic_chain(int logN, int depth, int first_bits, int scale_bits, int extra_bits);
// Minimal extraction of (ring dimension, Q moduli, dnum, aux/extra bits) from the cereal-JSON CryptoContext.
bool parse_ccjson(const char* text, size_t len, Params& p, std::string& err);

// twiddle tables of one limb: fwd[k] = (psi^brev(k), shoup), inv[k] = (psi^-brev(k), shoup)
// (f, shoup(f), f * psi^-brev(1), shoup(.)) per slot: the output factor of an inverse transform, folded into its last
// stage (k_inv_col); limb[i] is the modulus index of slot i
std::vector<u64> scale4(const Params& p, const std::vector<int>& limb, const std::vector<u64>& f);
void build_twiddles(const Params& p, int limb, u64* fwd_pairs /* 2N */, u64* inv_pairs /* 2N */);

struct Job {
    int src_first, ns, dst_slot, limb, chat_off;
};

// Constants of the hybrid key switch at `ell` present Q limbs (keyswitch-hybrid.cpp + ApproxSwitchCRTBasis /
// ApproxModDown tables of PrecomputeCRTTables).
struct KsPlan {
    int ell = 0, parts = 0;
    std::vector<int> c1_limb;           // [ell]
    std::vector<u64> c1_scale, c1_scale_sh;   // N^-1 * (D_j/q_i)^-1 mod q_i
    std::vector<Job> up_jobs;           // ModUp: (digit j, target slot r not in digit j)
    std::vector<u64> up_chat;           // (D_j/q_i) mod t
    std::vector<int> p_limb;            // [2K]
    std::vector<u64> p_scale, p_scale_sh;     // N^-1 * (P/p_k)^-1 mod p_k
    std::vector<Job> down_jobs;         // ModDown: (poly, i)
    std::vector<u64> down_chat;         // (P/p_k) mod q_i
    std::vector<u64> pinv, pinv_sh;     // P^-1 mod q_i
};
void build_ks_plan(const Params& p, int ell, KsPlan& out);

// Rescale by the last of `ell` limbs followed by multiplication with per-limb factors f (may be all 1).
struct RsPlan {
    int ell = 0;
    std::vector<int> last_limb;        // [2]
    std::vector<u64> last_scale, last_scale_sh;  // N^-1 mod q_last
    std::vector<Job> jobs;             // (poly, i < ell-1)
    std::vector<u64> qlinv;            // q_last^-1 mod q_i
};
void build_rs_plan(const Params& p, int ell, RsPlan& out);

// ckksrns-leveledshe.cpp GetElementForEvalMult: integer image of the real constant c at scaling factor sf.
void mult_const_factors(const Params& p, int ell, double sf, double c, u64* factors);

}  // namespace host
}  // namespace ppq
