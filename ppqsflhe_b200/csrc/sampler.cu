// sampler.cu -- secret / noise / uniform sampling for KeyGen, ReKeyGen and Encrypt on the device.
//
// Replaces what OpenFHE draws inside cc->KeyGen() (client/src/keyGen.cpp:33), cc->ReKeyGen (client/src/REkeyGen.cpp:52)
// and cc->Encrypt (client/src/encryptModelWeights.cpp:83,91,110): TernaryUniformGenerator, DiscreteGaussianGenerator
// (sigma = 3.19) and DiscreteUniformGenerator, all fed by a keyed CSPRNG.  Here:
//   * generator: ChaCha20 (RFC 8439 block function, 20 rounds) in counter mode, keyed by 256 bits.  The key comes from
//     the caller (tests) or from getrandom(2) per call (capi.cu).  Every polynomial draws from its own stream: the
//     96-bit nonce is (domain, polynomial index, chunk of 64 coefficients), the 32-bit block counter runs inside a chunk.
//     Secret material and public material never share a key: ppq_sample_keygen derives two sub-keys from the seed with
//     ChaCha20 itself (domain-separated KDF block) -- the public `a` says nothing about the key that made s and e.
//   * ternary: uniform on {-1, 0, 1} by rejection on 2-bit draws (3 is rejected); no modulo bias.
//   * discrete Gaussian: inversion of the cumulative table of |x| (64-bit fixed point, built on the host in long
//     double, tail cut at 2^-80) plus a sign bit; exact to 2^-64 per sample, no floating point on the device.
//   * uniform residues: rejection on bitlen(q)-bit draws.
// Parity tests inject the samples instead (ppq_keygen / ppq_rekeygen / ppq_encrypt take them as inputs); statistical
// tests on the device output are in tests/test_gpu_parity.py.
#include "sampler.cuh"

#include <cmath>
#include <cstring>
#include <vector>

namespace ppq {
namespace sampler {

__host__ __device__ __forceinline__ u32 rotl32(u32 x, int n) { return (x << n) | (x >> (32 - n)); }
#define PPQ_QR(a, b, c, d)     \
    a += b; d ^= a; d = rotl32(d, 16); \
    c += d; b ^= c; b = rotl32(b, 12); \
    a += b; d ^= a; d = rotl32(d, 8);  \
    c += d; b ^= c; b = rotl32(b, 7);

// RFC 8439 section 2.3: out = ChaCha20 block(key, counter, nonce)
__host__ __device__ inline void chacha20_block(const u32 (&key)[8], u32 counter, u32 n0, u32 n1, u32 n2, u32 (&out)[16]) {
    u32 s[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key[0], key[1], key[2], key[3],
                 key[4],      key[5],      key[6],      key[7],      counter, n0,     n1,     n2};
    u32 x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = s[i];
#pragma unroll 1
    for (int r = 0; r < 10; ++r) {
        PPQ_QR(x[0], x[4], x[8], x[12])
        PPQ_QR(x[1], x[5], x[9], x[13])
        PPQ_QR(x[2], x[6], x[10], x[14])
        PPQ_QR(x[3], x[7], x[11], x[15])
        PPQ_QR(x[0], x[5], x[10], x[15])
        PPQ_QR(x[1], x[6], x[11], x[12])
        PPQ_QR(x[2], x[7], x[8], x[13])
        PPQ_QR(x[3], x[4], x[9], x[14])
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) out[i] = x[i] + s[i];
}

void chacha20_block_host(const uint32_t key[8], uint32_t counter, const uint32_t nonce[3], uint32_t out[16]) {
    u32 k[8], o[16];
    for (int i = 0; i < 8; ++i) k[i] = key[i];
    chacha20_block(k, counter, nonce[0], nonce[1], nonce[2], o);
    for (int i = 0; i < 16; ++i) out[i] = o[i];
}

Key derive_key(const Key& master, uint32_t domain) {
    // KDF: the first 32 bytes of the ChaCha20 block (master, counter 0, nonce = ("ppqk", domain, 0))
    const uint32_t nonce[3] = {0x7070716bu, domain, 0u};
    uint32_t blk[16];
    chacha20_block_host(master.w, 0, nonce, blk);
    Key k;
    for (int i = 0; i < 8; ++i) k.w[i] = blk[i];
    return k;
}

// per-thread stream: words are consumed in order; the block counter advances when a block is used up
struct Stream {
    u32 key[8];
    u32 n0, n1, n2, ctr, pos;
    u32 blk[16];
    __device__ __forceinline__ void init(const Key& k, u32 a, u32 b, u32 c) {
#pragma unroll
        for (int i = 0; i < 8; ++i) key[i] = k.w[i];
        n0 = a;
        n1 = b;
        n2 = c;
        ctr = 0;
        pos = 16;
    }
    __device__ __forceinline__ u32 next() {
        if (pos == 16) {
            chacha20_block(key, ctr++, n0, n1, n2, blk);
            pos = 0;
        }
        // dynamic index into a register array would spill: select with a compile-time unrolled scan
        u32 v = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) v = (pos == (u32)i) ? blk[i] : v;
        ++pos;
        return v;
    }
    __device__ __forceinline__ u64 next64() {
        const u64 lo = next();
        return lo | ((u64)next() << 32);
    }
};

constexpr int CHUNK = 64;  // coefficients per stream / thread

// cumulative table of |X| for the discrete Gaussian, 64-bit fixed point: cdt[k] = floor(2^64 P(|X| <= k)) (saturating)
__constant__ u64 c_cdt[GAUSS_TABLE];
__constant__ int c_cdt_n;

__device__ __forceinline__ long long draw_ternary(Stream& s, u32& word, int& left) {
    for (;;) {
        if (left == 0) {
            word = s.next();
            left = 16;
        }
        const u32 v = word & 3u;
        word >>= 2;
        --left;
        if (v != 3u) return (long long)v - 1;
    }
}
__device__ __forceinline__ long long draw_gauss(Stream& s) {
    const u64 r = s.next64();
    const u32 sign = s.next() & 1u;
    int k = 0;
    const int n = c_cdt_n;
    for (int i = 0; i < n; ++i) k += (r >= c_cdt[i]) ? 1 : 0;  // constant time in r: every entry is compared
    return sign ? -(long long)k : (long long)k;                  // (k = 0 carries both signs: P(0) is not doubled in the table)
}

// out[poly][N]: kind[poly] selects ternary (0) or Gaussian (1); stream = (domain, poly0 + poly, chunk)
__global__ void k_sample_small(Key key, u32 domain, u32 poly0, const int* __restrict__ kinds, int n_kinds,
                               long long* __restrict__ out, u32 N) {
    const u32 poly = blockIdx.y;
    const u32 chunk = blockIdx.x * blockDim.x + threadIdx.x;
    if (chunk * CHUNK >= N) return;
    Stream s;
    s.init(key, domain, poly0 + poly, chunk);
    const int kind = kinds[poly % n_kinds];
    long long* o = out + (size_t)poly * N + (size_t)chunk * CHUNK;
    u32 word = 0;
    int left = 0;
    for (int i = 0; i < CHUNK && chunk * CHUNK + i < N; ++i) o[i] = kind == 0 ? draw_ternary(s, word, left) : draw_gauss(s);
}

// out[r][N] uniform in [0, q_r), r = limb0 .. limb0 + gridDim.y - 1
__global__ void k_sample_uniform(Key key, u32 domain, KCtx c, u64* __restrict__ out) {
    const u32 N = 1u << c.logN;
    const u32 r = blockIdx.y;
    const u32 chunk = blockIdx.x * blockDim.x + threadIdx.x;
    if (chunk * CHUNK >= N) return;
    const u64 q = c.limbs[r].q;
    const int drop = __clzll(q);
    Stream s;
    s.init(key, domain, r, chunk);
    u64* o = out + (size_t)r * N + (size_t)chunk * CHUNK;
    for (int i = 0; i < CHUNK && chunk * CHUNK + i < N; ++i) {
        u64 v;
        do v = s.next64() >> drop;
        while (v >= q);  // acceptance >= 1/2 per draw
        o[i] = v;
    }
}

static bool g_table_ready[64] = {};
cudaError_t upload_gauss_table(int device, double sigma) {
    if (device >= 0 && device < 64 && g_table_ready[device]) return cudaSuccess;
    // P(X = x) = rho(x) / sum_Z rho,  rho(x) = exp(-x^2 / (2 sigma^2)); two-sided mass of |x| = k > 0 is 2 rho(k)
    const int T = GAUSS_TABLE;
    std::vector<long double> rho(T + 64);
    long double tot = 0;
    for (int k = 0; k < T + 64; ++k) {
        rho[k] = expl(-(long double)k * k / (2.0L * (long double)sigma * (long double)sigma));
        tot += (k == 0 ? 1.0L : 2.0L) * rho[k];
    }
    std::vector<u64> cdt(T);
    long double cum = 0;
    int n = 0;
    for (int k = 0; k < T; ++k) {
        cum += (k == 0 ? 1.0L : 2.0L) * rho[k] / tot;
        const long double scaled = ldexpl(cum, 64);
        if (scaled >= 18446744073709551615.0L || 1.0L - cum < ldexpl(1.0L, -80)) {  // the rest of the tail is below 2^-80
            n = k;  // values above k are unreachable: r >= cdt[i] for all i < n gives at most n
            break;
        }
        cdt[k] = (u64)scaled;
        n = k + 1;
    }
    cudaError_t e = cudaMemcpyToSymbol(c_cdt, cdt.data(), sizeof(u64) * T);
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_cdt_n, &n, sizeof(int));
    if (e == cudaSuccess && device >= 0 && device < 64) g_table_ready[device] = true;
    return e;
}

void launch_sample_small(const Key& key, uint32_t domain, uint32_t poly0, const int* d_kinds, int n_kinds, int64_t* d_out,
                         size_t N, size_t n_polys, cudaStream_t st) {
    const unsigned chunks = (unsigned)((N + CHUNK - 1) / CHUNK);
    dim3 grid((chunks + 63) / 64, (unsigned)n_polys, 1);
    k_sample_small<<<grid, 64, 0, st>>>(key, domain, poly0, d_kinds, n_kinds, (long long*)d_out, (u32)N);
}
void launch_sample_uniform(const Key& key, uint32_t domain, const KCtx& k, int n_limbs, uint64_t* d_out, cudaStream_t st) {
    const unsigned N = 1u << k.logN, chunks = (N + CHUNK - 1) / CHUNK;
    dim3 grid((chunks + 63) / 64, (unsigned)n_limbs, 1);
    k_sample_uniform<<<grid, 64, 0, st>>>(key, domain, k, (u64*)d_out);
}

}  // namespace sampler
}  // namespace ppq
