// sampler.cuh -- ChaCha20-keyed samplers (ternary, discrete Gaussian by CDT inversion, uniform residues); see sampler.cu.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include "kernels.cuh"

namespace ppq {
namespace sampler {

struct Key {
    uint32_t w[8];
};
constexpr int GAUSS_TABLE = 64;

// stream domains (first nonce word): one per kind of material, so that no two polynomials ever share a stream
enum : uint32_t { DOM_SMALL = 0x534d4c31u /* secret / noise polynomials */, DOM_UNIFORM = 0x554e4931u /* public a */ };

// sub-key for one purpose from a master key (one ChaCha20 block with a KDF nonce): secret and public material are
// generated under different keys
Key derive_key(const Key& master, uint32_t domain);
void chacha20_block_host(const uint32_t key[8], uint32_t counter, const uint32_t nonce[3], uint32_t out[16]);

// cumulative |x| table of the discrete Gaussian into constant memory (once per device)
cudaError_t upload_gauss_table(int device, double sigma);

// d_out [n_polys][N] int64: polynomial p is ternary when d_kinds[p % n_kinds] == 0, Gaussian otherwise; its stream is
// (domain, poly0 + p)
void launch_sample_small(const Key& key, uint32_t domain, uint32_t poly0, const int* d_kinds, int n_kinds, int64_t* d_out,
                         size_t N, size_t n_polys, cudaStream_t st);
// d_out [n_limbs][N]: uniform residues mod q_r, r < n_limbs
void launch_sample_uniform(const Key& key, uint32_t domain, const KCtx& k, int n_limbs, uint64_t* d_out, cudaStream_t st);

}  // namespace sampler
}  // namespace ppq
