/*
 * ppq_b200.h -- C ABI of libppq_b200.so: the B200-native (sm_100a) replacement for the OpenFHE calls on the
 * encrypted-aggregation hot path of CDACHPCIE25/PPQSFLHE.
 *
 * Every entry point names the reference interface it replaces (paths relative to /root/reference).  The
 * reference reaches all of them through OpenFHE's CryptoContext<DCRTPoly> (third-party, not vendored); the
 * binding a maintainer would add on the reference side is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain C: opaque handles, pointers and sizes only; no C++/torch types.
 *   - every function returns 0 on success, a negative PPQ_E* code on failure; ppq_last_error() returns a
 *     thread-local message.  There is NO CPU fallback: without a usable sm_100a device every compute entry
 *     point fails with PPQ_ECUDA.
 *   - pointers named d_* are DEVICE pointers (cudaMalloc / torch CUDA storage); h_* are HOST pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Calls are ASYNCHRONOUS on that
 *     stream (they return once the work is enqueued; ppq_sync waits) -- with these stated exceptions, which return
 *     after their work is complete: ppq_ntt_forward / ppq_ntt_inverse (the caller's host limb-id list is staged),
 *     ppq_rekey_upload / ppq_rekey_from_device, the ppq_ks_* stage views (test-facing), and the ppq_*_host calls.
 *     The first use of a limb count builds small device tables once (a blocking copy); later calls do not block.
 *     A context owns scratch memory, so use one stream at a time per context.
 *   - batch sizes are limited only by memory: calls walk large batches in chunks internally.
 *   - residues are uint64 little-endian, canonical in [0, q_i); polynomials are limb-major ([limb][N]) in
 *     EVALUATION format in OpenFHE's bit-reversed order unless stated otherwise, which is exactly the payload
 *     of a serialized OpenFHE DCRTPoly.
 *   - a ciphertext is [2][ell][N] (c0 limbs then c1 limbs); a batch is [n][2][ell][N].
 *   - limb indices are "global": 0..L-1 = Q limbs, L..L+K-1 = auxiliary P limbs.
 */
#ifndef PPQ_B200_H
#define PPQ_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define PPQ_OK 0
#define PPQ_EINVAL (-1)  /* bad argument */
#define PPQ_ECUDA (-2)   /* CUDA error or no sm_100a device */
#define PPQ_ENOMEM (-3)
#define PPQ_EPARSE (-4)  /* CC.json could not be understood */

typedef struct ppq_ctx ppq_ctx;
typedef struct ppq_rekey ppq_rekey;

const char* ppq_last_error(void);
const char* ppq_version(void);

/* ---- context: replaces Serial::DeserializeFromFile(cc_path, cc, SerType::JSON)
 *      (server/src/changeCipherDomain.cpp:32-36, aggregateEncryptedWeights.cpp:46-50,
 *       client/src/encryptModelWeights.cpp:32-36, decryptModelWeights.cpp:41-46) and the parameter side of
 *      GenCryptoContext (server/src/genCC.cpp:46-79). */
int ppq_ctx_create_from_ccjson(const char* json, size_t len, int device, ppq_ctx** out);
/* explicit Q chain (q_0..q_{L-1}); P primes, roots and tables are derived by OpenFHE's HYBRID rules */
int ppq_ctx_create(int logN, int L, const uint64_t* q_moduli, int dnum, int aux_bits, int extra_bits, int device,
                   ppq_ctx** out);
/* FLEXIBLEAUTOEXT chain for multiplicative depth `depth` (genCC.cpp:46-66 with a forced ring dimension) */
int ppq_ctx_create_synthetic(int logN, int depth, int first_bits, int scale_bits, int extra_bits, int dnum,
                             int device, ppq_ctx** out);
void ppq_ctx_destroy(ppq_ctx*);

typedef struct {
    int32_t logN, L, K, dnum, alpha, device;
} ppq_ctx_info_t;
int ppq_ctx_info(const ppq_ctx*, ppq_ctx_info_t* info);
int ppq_ctx_moduli(const ppq_ctx*, uint64_t* h_out /* L+K */);
int ppq_ctx_roots(const ppq_ctx*, uint64_t* h_out /* L+K */);
double ppq_ctx_scaling_factor(const ppq_ctx*, int level);     /* m_scalingFactorsReal[level]    */
double ppq_ctx_scaling_factor_big(const ppq_ctx*, int level); /* m_scalingFactorsRealBig[level] */
int ppq_ctx_num_parts(const ppq_ctx*, int ell);
int ppq_sync(const ppq_ctx*, void* stream);
/* ciphertexts of one internal wave (scratch is sized for it); 0 = automatic */
int ppq_ctx_set_wave(ppq_ctx*, int n_ct);

/* ---- RNS polynomial layer: replaces DCRTPoly::SetFormat(EVALUATION / COEFFICIENT), i.e. OpenFHE's
 *      ChineseRemainderTransformFTT::ForwardTransformToBitReverseInPlace / InverseTransformFromBitReverseInPlace
 *      (reached from every call below).  d_data: [n_polys][N] in place; h_limb_ids: global limb per poly. */
int ppq_ntt_forward(ppq_ctx*, uint64_t* d_data, const int32_t* h_limb_ids, size_t n_polys, void* stream);
int ppq_ntt_inverse(ppq_ctx*, uint64_t* d_data, const int32_t* h_limb_ids, size_t n_polys, void* stream);

/* ---- re-encryption key: replaces Serial::DeserializeFromFile(rekey_path, reKey, SerType::JSON)
 *      (server/src/changeCipherDomain.cpp:40-44).  b, a: [dnum][L+K][N] EVAL (EvalKey B-vector / A-vector).
 *      Both calls also prepare, once, the key's Q limbs multiplied by P^-1 mod q_i (the ModDown multiplication is
 *      folded into the key).  _from_device BORROWS d_b / d_a (never written; they must outlive the handle and be
 *      complete when the call is made -- it synchronises the default stream) and keeps that product in a private
 *      [dnum][L][N] pair of buffers. */
int ppq_rekey_upload(ppq_ctx*, const uint64_t* h_b, const uint64_t* h_a, ppq_rekey** out);
int ppq_rekey_from_device(ppq_ctx*, const uint64_t* d_b, const uint64_t* d_a, ppq_rekey** out);
void ppq_rekey_destroy(ppq_rekey*);

/* ---- cc->ReEncrypt(ct, reKey) with publicKey == nullptr (server/src/changeCipherDomain.cpp:74,89,105):
 *      d_out[g] = ReEncrypt(d_in[g]) for g < n.  d_out may not alias d_in. */
int ppq_reencrypt(ppq_ctx*, const ppq_rekey*, int ell, const uint64_t* d_in, uint64_t* d_out, size_t n,
                  void* stream);
/* fused cc->ReEncrypt + cc->EvalAdd (changeCipherDomain.cpp:105 followed by
 * aggregateEncryptedWeights.cpp:106): d_acc[g] = d_acc[g] + ReEncrypt(d_in[g]) mod q, canonical. */
int ppq_reencrypt_accumulate(ppq_ctx*, const ppq_rekey*, int ell, const uint64_t* d_in, uint64_t* d_acc, size_t n,
                             void* stream);
/* fan-out of the aggregate (the second changeCipherDomain pass, orchestration/server_fns.sh:76-80, generalised
 * to n_keys target domains): d_outs[k][g] = ReEncrypt(d_in[g], rekeys[k]) for k < n_keys.
 * rekeys and d_outs are HOST arrays of n_keys handles / device pointers ([n][2][ell][N] each).  The digit
 * decomposition + ModUp of a ciphertext is independent of the key and is computed once per ciphertext; only the key
 * inner product and the ModDown run per key.  Results are identical to n_keys ppq_reencrypt calls. */
int ppq_reencrypt_fanout(ppq_ctx*, const ppq_rekey* const* rekeys, int n_keys, int ell, const uint64_t* d_in,
                         uint64_t* const* d_outs, size_t n, void* stream);
/* stage-by-stage views of the key switch (EvalKeySwitchPrecomputeCore / EvalFastKeySwitchCoreExt /
 * ApproxModDown) for parity tests: digits [n][parts][ell+K][N] EVAL canonical, ext [n][2][ell+K][N]. */
int ppq_ks_modup(ppq_ctx*, int ell, const uint64_t* d_c1 /* [n][ell][N] */, uint64_t* d_digits, size_t n,
                 void* stream);

/* d_digits [n][parts][ell+K][N] (as ppq_ks_modup writes them) -> d_ext [n][2][ell+K][N] =
 * (sum_j digit_j * b_j, sum_j digit_j * a_j), canonical: EvalFastKeySwitchCoreExt on the key as uploaded */
int ppq_ks_inner_product(ppq_ctx*, const ppq_rekey*, int ell, const uint64_t* d_digits, uint64_t* d_ext, size_t n,
                         void* stream);
/* d_ext [n][2][ell+K][N] -> d_out [n][2][ell][N]: ApproxModDown of both polynomials */
int ppq_ks_moddown(ppq_ctx*, int ell, const uint64_t* d_ext, uint64_t* d_out, size_t n, void* stream);
/* every constant the library derives from (N, Q chain, dnum) at `ell` present limbs, as JSON text (residues as decimal
 * strings): P primes, roots, scaling factors, ModUp / ModDown / rescale tables.  Host only -- no device is touched --
 * so independent big-integer checks (tests/test_params_bigint.py) run anywhere.  *need = bytes incl. NUL; buf may be
 * NULL to query the size. */
int ppq_params_dump(int logN, int L, const uint64_t* q_moduli, int dnum, int aux_bits, int extra_bits, int ell,
                    char* buf, size_t cap, size_t* need);

/* the Q chain GenCryptoContext builds under FLEXIBLEAUTOEXT (server/src/genCC.cpp:46-68): host only */
int ppq_params_chain(int logN, int depth, int first_bits, int scale_bits, int extra_bits, uint64_t* q_out, int cap,
                     int* n_out);

/* ---- cc->EvalAdd(ct1, ct2) (server/src/aggregateEncryptedWeights.cpp:82,91,106) */
int ppq_eval_add(ppq_ctx*, int ell, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out, size_t n,
                 void* stream);
/* canonicalise lazily summed residues (after a cross-GPU uint64 sum of canonical partial sums) */
int ppq_finalize_sum(ppq_ctx*, int ell, uint64_t* d_x, size_t n, void* stream);

/* ---- cc->EvalMult(ct, c) (server/src/aggregateEncryptedWeights.cpp:83,92,107) under FLEXIBLEAUTOEXT.
 *      rescale != 0 (noiseScaleDeg == 2): DropLastElementAndScale then multiply by the integer image of c at
 *      scaling factor sf: d_in [n][2][ell][N] -> d_out [n][2][ell-1][N].
 *      rescale == 0: multiply only, d_out [n][2][ell][N].  sf is the scaling factor of the level AFTER the
 *      optional rescale (ppq_ctx_scaling_factor). */
int ppq_eval_mult_const(ppq_ctx*, int ell, const uint64_t* d_in, uint64_t* d_out, size_t n, double c, double sf,
                        int rescale, void* stream);

/* ---- cc->MakeCKKSPackedPlaintext(values) (client/src/encryptModelWeights.cpp:82,90,109):
 *      h/d_vals [n][n_vals] doubles (zero-padded to N/2 slots) -> d_pt [n][ell][N] EVAL, scale sf. */
int ppq_encode(ppq_ctx*, int ell, const double* d_vals, size_t n_vals, double sf, uint64_t* d_pt, size_t n,
               void* stream);
/* ---- cc->Encrypt(publicKey, pt) (client/src/encryptModelWeights.cpp:83,91,110) with the sampled
 *      polynomials supplied by the caller: d_small [n][3][N] int64 = (v, e0, e1) coefficients.
 *      d_pkb / d_pka: [>=ell][N] EVAL (first ell limbs of the QP public key). */
int ppq_encrypt(ppq_ctx*, int ell, const uint64_t* d_pkb, const uint64_t* d_pka, const uint64_t* d_pt,
                const int64_t* d_small, uint64_t* d_ct, size_t n, void* stream);
/* ---- ppq_encrypt without any device allocation.  NO counterpart in the reference: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 *      that forbid it).  Same arguments and bit-identical output as ppq_encrypt, but the scratch ((v, e0, e1) in EVAL form
 *      and the limb-id table) lives in one file-scope device array of 64 MiB that every GPU loading the library holds
 *      until exit, so the call never runs cudaMalloc / cudaFree or an allocating library routine.  Limit: n * 3 * ell *
 *      (8 N + 4) bytes <= 64 MiB (C1, ell 4, N 2^14: 42 ciphertexts per call -- the reference's 27-ciphertext file fits;
 *      C3, ell 20, N 2^16: 2); a larger batch is refused with PPQ_EINVAL and a message naming the maximum -- split the
 *      batch or use ppq_encrypt.  All calls on a device share the array (it keeps the last call's contents, none of which
 *      a later call reads before overwriting them): calls that overlap on the GPU -- two streams, two contexts, two
 *      threads -- RACE; serialise them on one stream.  The first call on a device resolves the array's address
 *      (cudaGetSymbolAddress); make one call before starting a graph capture.  Asynchronous on `stream`. */
int ppq_encrypt_noalloc(ppq_ctx*, int ell, const uint64_t* d_pkb, const uint64_t* d_pka, const uint64_t* d_pt,
                        const int64_t* d_small, uint64_t* d_ct, size_t n, void* stream);
/* ---- the two calls above as one (client/src/encryptModelWeights.cpp:109-110, MakeCKKSPackedPlaintext then Encrypt per
 *      chunk): the encoded coefficients are added to e0 before the transform (the NTT is linear), so a fresh ciphertext
 *      costs 3*ell forward limb transforms instead of 4*ell.  d_ct is bit-identical to ppq_encode + ppq_encrypt. */
int ppq_encode_encrypt(ppq_ctx*, int ell, const uint64_t* d_pkb, const uint64_t* d_pka, const double* d_vals,
                       size_t n_vals, double sf, const int64_t* d_small, uint64_t* d_ct, size_t n, void* stream);
/* fills d_small [n][3][N] with (ternary, discrete Gaussian sigma 3.19, Gaussian) polynomials: ChaCha20 streams keyed
 * by seed32 (32 bytes), one stream per polynomial; ternary by rejection on 2-bit draws, Gaussian by inversion of a
 * 64-bit cumulative table (csrc/sampler.cu).  seed32 == NULL draws a fresh 256-bit key from getrandom(2) for this
 * call -- the only mode for real keys and ciphertexts; a caller-supplied seed is for tests and reproducible runs. */
int ppq_sample_encrypt_noise(ppq_ctx*, const uint8_t* seed32, int64_t* d_small, size_t n, void* stream);
/* the samplers' generator, exposed for known-answer tests: one ChaCha20 block (RFC 8439 2.3), computed on the host
 * by the same inline function the device kernels use */
int ppq_chacha20_block(const uint32_t* key8, uint32_t counter, const uint32_t* nonce3, uint32_t* out16);
/* ---- cc->Decrypt(privateKey, ct, &pt) core (client/src/decryptModelWeights.cpp:81,90,108):
 *      d_coeff [n][ell][N] = COEFF residues of c0 + c1*s.  d_sk: [>=ell][N] EVAL. */
int ppq_decrypt_core(ppq_ctx*, int ell, const uint64_t* d_sk, const uint64_t* d_ct, uint64_t* d_coeff, size_t n,
                     void* stream);
/* ---- pt->GetRealPackedValue() (client/src/decryptModelWeights.cpp:83,92,109): CRT-compose, centre, divide
 *      by sf, special FFT; d_vals [n][N/2].  OpenFHE's randomised decode noise is not added. */
int ppq_decode(ppq_ctx*, int ell, const uint64_t* d_coeff, double sf, double* d_vals, size_t n, void* stream);

/* ---- cc->KeyGen() (client/src/keyGen.cpp:33; OpenFHE base-pke.cpp KeyGenInternal) with the sampled
 *      polynomials supplied by the caller: d_s, d_e [N] int64 (ternary secret, Gaussian error sigma 3.19),
 *      d_a [L+K][N] uniform residues (EVAL).  Out: d_sk [L][N] (the secret on the Q limbs, EVAL),
 *      d_pkb [L+K][N] = e - a*s.  The public key is the pair (d_pkb, d_a) over QP. */
int ppq_keygen(ppq_ctx*, const int64_t* d_s, const int64_t* d_e, const uint64_t* d_a, uint64_t* d_sk,
               uint64_t* d_pkb, void* stream);
/* draws ppq_keygen's inputs on the device.  The secret half (s, e) and the public half (a, uniform residues by
 * rejection) come from DIFFERENT ChaCha20 keys -- two independent getrandom(2) draws when seed32 == NULL, two
 * KDF sub-keys of seed32 otherwise (tests) -- so the published `a` carries no information about s. */
int ppq_sample_keygen(ppq_ctx*, const uint8_t* seed32, int64_t* d_s, int64_t* d_e, uint64_t* d_a, void* stream);
/* ---- cc->ReKeyGen(oldPrivateKey, newPublicKey) (client/src/REkeyGen.cpp:52; OpenFHE keyswitch-hybrid.cpp
 *      KeySwitchGenInternal(oldSk, newPk)): per level-0 digit j, b_j = pk.b*u + e0 + (P mod q_r)*s_old on the
 *      limbs of digit j, a_j = pk.a*u + e1.  d_sk_old [L][N], d_pkb / d_pka [L+K][N] (EVAL);
 *      d_small [dnum][3][N] int64 = (u, e0, e1) per digit (ppq_sample_encrypt_noise(seed, d_small, dnum)
 *      draws them).  Out: d_kb, d_ka [dnum][L+K][N], the layout ppq_rekey_from_device takes. */
int ppq_rekeygen(ppq_ctx*, const uint64_t* d_sk_old, const uint64_t* d_pkb, const uint64_t* d_pka,
                 const int64_t* d_small, uint64_t* d_kb, uint64_t* d_ka, void* stream);

/* ---- C-client aggregation: sum over clients of cc->ReEncrypt(ct_c, reKey_c) per ciphertext slot, i.e.
 *      changeCipherDomain.cpp:105 applied to every client's file followed by the EvalAdd chain of
 *      aggregateEncryptedWeights.cpp:106 (the reference hard-codes two clients; SURVEY.md 8 rows a1+a2, e).
 *      The result is BIT-IDENTICAL to ppq_reencrypt + ppq_eval_add per client, but the linear tail of ApproxModDown
 *      (base conversion P->Q of the integer lifts, forward NTT, subtraction, +c0) is applied to sums: the conversion
 *      once per group of up to 8 clients, the 2*ell forward NTTs once per slot in ppq_agg_finish.
 *      d_in: one client's [n_slots][2][ell][N]; d_out: [n_slots][2][ell][N], canonical.  finish resets the handle. */
typedef struct ppq_agg ppq_agg;
int ppq_agg_create(ppq_ctx*, int ell, size_t n_slots, ppq_agg** out);
int ppq_agg_add(ppq_agg*, const ppq_rekey*, const uint64_t* d_in, void* stream);
/* several clients at once (HOST arrays of n_clients handles / device pointers): identical result; the library takes
 * them a few per pass so that the slot sums are read and written once per pass instead of once per client */
int ppq_agg_add_many(ppq_agg*, const ppq_rekey* const* rekeys, const uint64_t* const* d_ins, size_t n_clients,
                     void* stream);
int ppq_agg_finish(ppq_agg*, uint64_t* d_out, void* stream);
size_t ppq_agg_count(const ppq_agg*); /* clients added since the last finish */
void ppq_agg_destroy(ppq_agg*);

/* ---- cross-GPU sum of per-rank partial aggregates (SURVEY.md 8 rows b, e): one process (or thread) per GPU, clients
 *      sharded over ranks, ONE ncclReduce(ncclUint64, ncclSum) of canonical residues (< 2^60, so up to 15 ranks
 *      cannot wrap) + a mod-q kernel on the root.  NCCL is bound at first use (dlopen of libnccl.so.2).
 *      ppq_comm_unique_id: rank 0 makes the 128-byte id and hands it to the other ranks by any means;
 *      ppq_comm_from_nccl wraps a communicator the host already owns (ncclComm_t passed as void*, not destroyed). */
typedef struct ppq_comm ppq_comm;
int ppq_comm_unique_id(uint8_t* id128);
int ppq_comm_init(ppq_ctx*, int world, int rank, const uint8_t* id128, ppq_comm** out);
int ppq_comm_from_nccl(void* nccl_comm, int world, int rank, int device, ppq_comm** out);
void ppq_comm_destroy(ppq_comm*);
/* d_acc [n][2][ell][N] in place: the sum lands on `root` (root < 0: on every rank), canonical */
int ppq_reduce_across_ranks(ppq_ctx*, ppq_comm*, int ell, uint64_t* d_acc, size_t n, int root, void* stream);

/* ---- shard planning for ranks that do not all move data at the same rate (host only, no device is touched): cost[r] =
 *      seconds rank r needs per client (e.g. one timed ppq_copy_ceiling_host_packed of an equal shard / its clients);
 *      counts[r] = clients for rank r, minimising max_r counts[r] * cost[r] with sum counts = n_clients.  Deterministic:
 *      every rank holding the same cost vector computes the same plan.  The reference aggregates on one host
 *      (server/src/aggregateEncryptedWeights.cpp:68-110); this belongs to SURVEY.md 8 row e. */
int ppq_plan_shards(size_t n_clients, int world, const double* cost, size_t* counts);

/* ---- host-buffer convenience (the call a drop-in server makes once every client's upload has arrived):
 *      h_cts[client] -> [n_slots][2][ell][N] in host memory (pinned for full PCIe rate); rekeys[n_clients];
 *      h_acc [n_slots][2][ell][N] = sum over clients of ReEncrypt(ct, rekey[client]) (ppq_agg_* inside).  Uploads
 *      are double buffered against the kernels; device buffers, streams and events live in the context and are
 *      reused from call to call.  Synchronous.
 *      _ranks: this rank's clients only; the per-rank sums are combined with ppq_reduce_across_ranks and h_acc is
 *      written on `root` (may be NULL elsewhere); n_clients may be 0 on a rank.
 *      ppq_copy_ceiling_host: the same uploads and download on the same streams without any kernel -- what the
 *      box's host->device path allows (bench.py reports e2e against it). */
int ppq_reencrypt_aggregate_host(ppq_ctx*, const ppq_rekey* const* rekeys, int ell, const uint64_t* const* h_cts,
                                 size_t n_clients, size_t n_slots, uint64_t* h_acc);
int ppq_reencrypt_aggregate_host_ranks(ppq_ctx*, ppq_comm*, int root, const ppq_rekey* const* rekeys, int ell,
                                       const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                       uint64_t* h_acc);
int ppq_copy_ceiling_host(ppq_ctx*, int ell, const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                          uint64_t* h_acc);

/* ---- packed residues: the wire format of the *_host_packed calls.  A residue of limb i needs bitlen(q_i) bits, not 64
 *      (C3: 1142 of 1280 bits per coefficient), and the host->device copy is what bounds the end-to-end path.  A packed
 *      ciphertext is [2][ell] limbs; limb i = N residues of bitlen(q_i) bits each as a little-endian bit stream
 *      (N * bitlen(q_i) / 64 words).  ppq_packed_words = u64 words of one packed ciphertext (0 on a bad argument).
 *      The deserialiser of a cereal blob (8-byte residues, changeCipherDomain.cpp:68-72, aggregateEncryptedWeights.cpp:
 *      18-30) writes this layout instead of [2][ell][N] u64: ppq_pack_host / ppq_unpack_host are that conversion on the
 *      host (format only, no arithmetic); ppq_pack / ppq_unpack are the device kernels, asynchronous on `stream`. */
size_t ppq_packed_words(const ppq_ctx*, int ell);
int ppq_pack_host(const ppq_ctx*, int ell, const uint64_t* h_in, uint64_t* h_packed, size_t n);
int ppq_unpack_host(const ppq_ctx*, int ell, const uint64_t* h_packed, uint64_t* h_out, size_t n);
int ppq_pack(ppq_ctx*, int ell, const uint64_t* d_in, uint64_t* d_packed, size_t n, void* stream);
int ppq_unpack(ppq_ctx*, int ell, const uint64_t* d_packed, uint64_t* d_out, size_t n, void* stream);
/* the host-buffer calls above with packed host buffers on both sides: h_cts[client] -> [n_slots] packed ciphertexts,
 * h_acc -> [n_slots] packed ciphertexts.  The upload is unpacked on the device (one HBM-bound kernel per client, hidden
 * under the next client's copy), the aggregate packed before its download; the arithmetic and the result are those of
 * ppq_reencrypt_aggregate_host.  Residues must be canonical, as everywhere. */
int ppq_reencrypt_aggregate_host_packed(ppq_ctx*, const ppq_rekey* const* rekeys, int ell, const uint64_t* const* h_cts,
                                        size_t n_clients, size_t n_slots, uint64_t* h_acc);
int ppq_reencrypt_aggregate_host_packed_ranks(ppq_ctx*, ppq_comm*, int root, const ppq_rekey* const* rekeys, int ell,
                                              const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                              uint64_t* h_acc);
int ppq_copy_ceiling_host_packed(ppq_ctx*, int ell, const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                 uint64_t* h_acc);

/* ---- per-stage device timing of the key-switch pipeline (CUDA events on the caller's stream).
 *      stage s < PPQ_PROFILE_STAGES: see ppq_profile_stage_name.  ms / launches: arrays of that length. */
#define PPQ_PROFILE_STAGES 11
int ppq_profile_enable(ppq_ctx*, int on);
int ppq_profile_read(ppq_ctx*, double* ms, uint64_t* launches, int reset);
const char* ppq_profile_stage_name(int stage);
/* kernels launched by the library since it was loaded (bench.py's gpu_launches) */
uint64_t ppq_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
