/*
 * ppq_oracle.h -- CPU ORACLE (TEST INFRASTRUCTURE ONLY, never shipped, never on the product path).
 *
 * A plain-C restatement of the OpenFHE algorithms that the reference PPQSFLHE calls on its hot path
 * (cc->ReEncrypt / EvalAdd / EvalMult(.,0.5) / MakeCKKSPackedPlaintext / Encrypt / Decrypt):
 *   reference call sites: server/src/changeCipherDomain.cpp:74,89,105
 *                         server/src/aggregateEncryptedWeights.cpp:82-83,91-92,106-107
 *                         client/src/encryptModelWeights.cpp:82-83,90-91,109-110
 *                         client/src/decryptModelWeights.cpp:81-83,90-92,108-109
 * The arithmetic lives in the third-party dependency OpenFHE (openfhe-development @
 * 32362085fcfa8133f86a4e81d4440734c984b16d, >= v1.3.0), which is ABSENT from /root/reference
 * (un-vendored submodule, .SUBMODULES.json:9-15).  This file restates its published algorithms.
 *
 * PARITY PIN STATUS: "partially pinned".  Pinned against real OpenFHE artefacts that are in the tree:
 *   - server/storage/CC.json: ring dimension, the four Q primes and their 2N-th roots (prime selection
 *     rules + minimal-root rule are reproduced exactly; tests/test_oracle_params.py)
 *   - client/storage/client_{1,2}/private/client_{1,2}-private.key: a real OpenFHE secret key in
 *     EVALUATION format on all four limbs; the oracle's inverse NTT maps every limb to the SAME ternary
 *     polynomial and the forward NTT reproduces the serialized limbs bit-for-bit (pins root choice,
 *     bit-reversed output order and limb consistency; tests/golden/make_golden.py)
 *   - sample/decrypted weight pair (end-to-end CKKS tolerance).
 * NOT pinned (no OpenFHE ciphertext / rekey / executable exists here): ModUp/ModDown/rescale outputs.
 * Those follow the OpenFHE source from memory and are "parity unpinned" (see DESIGN.md).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 */
#ifndef PPQ_ORACLE_H
#define PPQ_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ppqo_ctx ppqo_ctx;

/* ---- number theory (OpenFHE core/lib/math/nbtheory: FirstPrime, PreviousPrime, NextPrime, RootOfUnity) */
uint64_t ppqo_mulmod(uint64_t a, uint64_t b, uint64_t q);
uint64_t ppqo_powmod(uint64_t a, uint64_t e, uint64_t q);
uint64_t ppqo_invmod(uint64_t a, uint64_t q);
int      ppqo_is_prime(uint64_t n);
uint64_t ppqo_first_prime(int bits, uint64_t m);
uint64_t ppqo_previous_prime(uint64_t q, uint64_t m);
uint64_t ppqo_next_prime(uint64_t q, uint64_t m);
uint64_t ppqo_min_root_of_unity(uint64_t m, uint64_t q); /* minimal primitive m-th root mod q */

/* ---- context ---------------------------------------------------------------------------------- */
/* Q moduli given explicitly (e.g. from CC.json); P primes derived by the OpenFHE HYBRID rule. */
ppqo_ctx* ppqo_ctx_create(int logN, int L, const uint64_t* q_moduli, int dnum, int aux_bits, int extra_bits);
/* CKKS FLEXIBLEAUTOEXT prime chain (ckksrns-parametergeneration.cpp) for mult depth `depth`. */
ppqo_ctx* ppqo_ctx_create_synthetic(int logN, int depth, int first_bits, int scale_bits, int extra_bits, int dnum);
void      ppqo_ctx_free(ppqo_ctx*);
int       ppqo_logN(const ppqo_ctx*);
int       ppqo_L(const ppqo_ctx*);
int       ppqo_K(const ppqo_ctx*);
int       ppqo_dnum(const ppqo_ctx*);
int       ppqo_alpha(const ppqo_ctx*);
void      ppqo_moduli(const ppqo_ctx*, uint64_t* out /* L+K */);
void      ppqo_roots(const ppqo_ctx*, uint64_t* out /* L+K */);
double    ppqo_scaling_factor(const ppqo_ctx*, int level);      /* m_scalingFactorsReal[level]    */
double    ppqo_scaling_factor_big(const ppqo_ctx*, int level);  /* m_scalingFactorsRealBig[level] */

/* ---- RNS polynomial layer: limb index r in [0, L+K) (Q limbs then P limbs), a = N residues in place */
void ppqo_ntt_fwd(const ppqo_ctx*, int limb, uint64_t* a);   /* natural -> bit-reversed EVAL order */
void ppqo_ntt_inv(const ppqo_ctx*, int limb, uint64_t* a);   /* bit-reversed EVAL -> natural COEFF */
void ppqo_ntt_fwd_naive(const ppqo_ctx*, int limb, const uint64_t* a, uint64_t* out); /* O(N^2) definition */

/* ---- hybrid key switch / PRE (keyswitch-hybrid.cpp, base-pre.cpp) ------------------------------- */
/* c1: [ell][N] EVAL.  key_b/key_a: [dnum][L+K][N] EVAL.  ks0/ks1: [ell][N]. */
void ppqo_keyswitch_core(const ppqo_ctx*, int ell, const uint64_t* c1, const uint64_t* key_b,
                         const uint64_t* key_a, uint64_t* ks0, uint64_t* ks1);
/* intermediate stages exposed for stage-by-stage parity tests */
void ppqo_modup(const ppqo_ctx*, int ell, const uint64_t* c1, uint64_t* digits /* [parts][ell+K][N] EVAL */);
void ppqo_inner_product(const ppqo_ctx*, int ell, const uint64_t* digits, const uint64_t* key_b,
                        const uint64_t* key_a, uint64_t* ct0 /* [ell+K][N] */, uint64_t* ct1);
void ppqo_moddown(const ppqo_ctx*, int ell, const uint64_t* ext /* [ell+K][N] EVAL */, uint64_t* out /* [ell][N] */);
int  ppqo_num_parts(const ppqo_ctx*, int ell);
/* ct: [2][ell][N] */
void ppqo_reencrypt(const ppqo_ctx*, int ell, const uint64_t* ct_in, const uint64_t* key_b,
                    const uint64_t* key_a, uint64_t* ct_out);
void ppqo_eval_add(const ppqo_ctx*, int ell, const uint64_t* a, const uint64_t* b, uint64_t* out);
/* DropLastElementAndScale on both polys: [2][ell][N] -> [2][ell-1][N] */
void ppqo_rescale(const ppqo_ctx*, int ell, const uint64_t* ct_in, uint64_t* ct_out);
/* GetElementForEvalMult: integer constant for real operand c at scaling factor sf, reduced per limb */
void ppqo_mult_const_factors(const ppqo_ctx*, int ell, double sf, double c, uint64_t* factors);
void ppqo_mult_scalar(const ppqo_ctx*, int ell, const uint64_t* factors, const uint64_t* ct_in, uint64_t* ct_out);

/* ---- CKKS encode / encrypt / decrypt / decode --------------------------------------------------- */
void ppqo_encode(const ppqo_ctx*, int ell, const double* vals, size_t nvals, double sf, uint64_t* pt /* [ell][N] EVAL */);
/* pk_b/pk_a: [L+K][N] EVAL (QP basis, as OpenFHE keeps it); v,e0,e1: N small signed ints (COEFF) */
void ppqo_encrypt(const ppqo_ctx*, int ell, const uint64_t* pk_b, const uint64_t* pk_a, const uint64_t* pt,
                  const int64_t* v, const int64_t* e0, const int64_t* e1, uint64_t* ct /* [2][ell][N] */);
/* sk: [>=ell][N] EVAL with limb stride N.  coeff_out: [ell][N] COEFF of c0 + c1*s */
void ppqo_decrypt_core(const ppqo_ctx*, int ell, const uint64_t* sk, const uint64_t* ct, uint64_t* coeff_out);
/* CRT-interpolate + centre + /sf + special FFT; deterministic (OpenFHE's decode noise is NOT added) */
void ppqo_decode(const ppqo_ctx*, int ell, const uint64_t* coeff, double sf, double* out /* N/2 */);
void ppqo_decrypt(const ppqo_ctx*, int ell, const uint64_t* sk, const uint64_t* ct, double sf, double* out);

/* ---- scaffolding: samplers, KeyGen (base-pke.cpp), ReKeyGen (keyswitch-hybrid.cpp KeySwitchGenInternal(sk,pk)) */
void ppqo_sample_ternary(uint64_t seed, size_t n, int64_t* out);
void ppqo_sample_gauss(uint64_t seed, size_t n, double sigma, int64_t* out);
void ppqo_sample_uniform(uint64_t seed, uint64_t q, size_t n, uint64_t* out);
/* lift small signed coefficients to limb r and NTT */
void ppqo_small_to_eval(const ppqo_ctx*, int limb, const int64_t* small, uint64_t* out);
/* sk_eval: [L+K][N]; pk_b, pk_a: [L+K][N] */
void ppqo_keygen(const ppqo_ctx*, uint64_t seed, int64_t* s_coeff /* N */, uint64_t* sk_eval, uint64_t* pk_b, uint64_t* pk_a);
void ppqo_rekeygen(const ppqo_ctx*, uint64_t seed, const uint64_t* sk_old_eval, const uint64_t* pk_new_b,
                   const uint64_t* pk_new_a, uint64_t* key_b /* [dnum][L+K][N] */, uint64_t* key_a);

/* ---- batched CPU baseline (OpenMP over ciphertexts, mirroring OpenFHE's WITH_OPENMP build) ------- */
/* acc[slot] = sum_c ReEncrypt(cts[c][slot], key[c]) for n_clients clients x n_slots cts/client.
 * cts: [n_clients][n_slots][2][ell][N]; keys: [n_clients] pairs (b,a); acc: [n_slots][2][ell][N]. */
void ppqo_reencrypt_accumulate_batch(const ppqo_ctx*, int ell, int n_clients, int n_slots,
                                     const uint64_t* cts, const uint64_t* const* key_b,
                                     const uint64_t* const* key_a, uint64_t* acc, int nthreads);
void ppqo_ntt_batch(const ppqo_ctx*, int inverse, const int* limb_ids, size_t n_polys, uint64_t* data, int nthreads);
int  ppqo_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
