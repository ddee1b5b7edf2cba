// ckks.cuh -- CKKS encode / decode (special FFT over the 2N-th complex roots, FP64) and the encryption noise
// sampler.  Replaces OpenFHE CKKSPackedEncoding::Encode/Decode + DiscreteFourierTransform::FFTSpecial(Inv)
// reached from cc->MakeCKKSPackedPlaintext (client/src/encryptModelWeights.cpp:82,90,109) and
// pt->GetRealPackedValue (client/src/decryptModelWeights.cpp:83,92,109).  Tolerance-checked, not bit-exact
// (values ~2^57 exceed the 53-bit mantissa).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <cuda_runtime.h>
#include "kernels.cuh"

namespace ppq {
namespace ckks {

struct Tables {
    int logN = 0;
    unsigned* rot = nullptr;   // 5^j mod 2N, j < N/2
    double2* ksi = nullptr;    // exp(2 pi i k / 2N), k <= 2N
    ~Tables();
};
bool build_tables(int logN, Tables& t, std::string& err);

// vals [n][n_vals] -> co [n][N] int64 coefficients (real parts then imaginary parts); z: [n][N/2] work
void launch_encode_fft(const Tables& t, const double* d_vals, size_t n_vals, double sf, double2* z, int64_t* co,
                       size_t n, cudaStream_t st);
// coeff [n][ell][N] canonical COEFF residues -> vals [n][N/2]
// d_ginv [ell]: (q_0..q_{i-1})^-1 mod q_i; d_qmod [ell][ell]: q_k mod q_i (Garner's mixed-radix constants)
// d_p3mod [2][ell]: q_0 q_1 q_2 mod q_i, then q_0 q_1 mod q_i (the short interpolation's check, ckks.cu)
void launch_decode(const Tables& t, const KCtx& k, const u64* coeff, int ell, const u64* d_ginv, const u64* d_qmod,
                   const u64* d_p3mod, double sf, double2* z, double* d_vals, size_t n, cudaStream_t st);
}  // namespace ckks

// plain forward ROW pass over [n_g] blocks of n_slots polys; slots whose limb id is negative are skipped
// lazy2: results may stay in [0, 2q) (digits for the key inner product; only the paired kernels honour it)
void launch_fwd_row_plain(const KCtx& c, u64* data, size_t gs, const int* d_limb_ids, int n_slots, int n_g,
                          cudaStream_t st, bool lazy2 = false);

}  // namespace ppq
