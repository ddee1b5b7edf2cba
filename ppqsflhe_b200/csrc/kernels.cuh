// kernels.cuh -- launcher interface between the C-ABI layer (capi.cu) and the sm_100a kernels (kernels.cu).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include "modarith.cuh"

namespace ppq {

// Device-resident view of a context: per-limb constants and the twiddle tables (w, w_shoup) for every
// modulus of Q u P, forward and inverse, each [M][N] ulonglong2.
struct KCtx {
    const LimbC* limbs;
    const ulonglong2* twf;
    const ulonglong2* twi;
    int logN;
};

// One fused "base conversion -> forward COL pass" job (ModUp of a digit to one target limb, or ModDown of the
// P part to one Q limb): out[dst_slot] = COLpass( sum_i src[src_first+i] * chat[chat_off+i]  mod q_limb ).
struct BcJob {
    int src_first;  // first source poly (index inside one ciphertext's source block)
    int ns;         // number of source limbs
    int dst_slot;   // destination poly inside one ciphertext's destination block
    int limb;       // global limb index of the target modulus
    int chat_off;   // offset into the chat[] constant array
    int pad[3];
};

// ---- plain batched NTT (C-ABI ppq_ntt_forward / ppq_ntt_inverse).  data: [n_polys][N] in place,
// d_limb_ids: device int[n_polys].  d_scale / d_scale_sh: per-POLY constant folded into the inverse (N^-1).
void launch_ntt_forward(const KCtx& c, u64* data, const int* d_limb_ids, int n_polys, cudaStream_t st);
// d_scale4: per polynomial (f, shoup(f), f * psi^-brev(1) mod q, shoup(.)) with f the output factor (here N^-1): the
// last inverse stage applies it (host::scale4 builds the array)
void launch_ntt_inverse(const KCtx& c, u64* data, const int* d_limb_ids, const u64* d_scale4, int n_polys,
                        cudaStream_t st);

// ---- generic strided inverse:  for g < n_g, y < n_slots:
//   out[g*out_gs + y*N ..] = iNTT(in[g*in_gs + y*N ..]) * scale[y]   (limb = limb_map[y]), canonical.
// row pass and column pass are separate launches; `tmp` may alias `out`.
void launch_inv_row(const KCtx& c, const u64* in, size_t in_gs, u64* out, size_t out_gs, const int* d_limb_map,
                    int n_slots, int n_g, cudaStream_t st);
// the same over n_in <= 4 source buffers of n_per ciphertexts each in ONE launch: ciphertext g of the launch is
// ciphertext g % n_per of ins[g / n_per]; out is contiguous over all n_in * n_per ciphertexts
void launch_inv_row_multi(const KCtx& c, const u64* const* ins, int n_in, int n_per, size_t in_gs, u64* out, size_t out_gs,
                          const int* d_limb_map, int n_slots, cudaStream_t st);
void launch_inv_col(const KCtx& c, u64* data, size_t gs, const int* d_limb_map, const u64* d_scale4, int n_slots,
                    int n_g, cudaStream_t st);

// ---- approximate CRT basis switch, pointwise (ApproxSwitchCRTBasis).  All `n_jobs` jobs share src_first and
// ns (= `ns`); dst[g][job.dst_slot] = sum_i src[g][src_first+i] * chat[job.chat_off+i] mod q_{job.limb}.
// h_jobs / h_chat / h_limbs are HOST arrays: the constants of a launch are packed into the kernel parameter
// (constant bank) by the launcher.  Returns false if ns is outside 1..32.
bool launch_bconv(const KCtx& c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs, const BcJob* h_jobs,
                  int n_jobs, const u64* h_chat, const LimbC* h_limbs, int ns, int n_g, cudaStream_t st);
// ---- the same conversion over the INTEGER SUM of n_in <= 8 source sets (deferred ModDown, ppq_agg_*): the conversion
// is linear in the integer lifts, so sum_c conv(u_c) = conv(sum_c u_c) mod q exactly; the sum of up to 8 canonical
// residues (< 2^63) is formed on the fly from src + c*in_stride, c < n_in, and reduced with a 128-bit-wide reduction.
// dst (canonical) is overwritten when `first`, else incremented mod q.  ns <= 16.
bool launch_bconv_wide(const KCtx& c, const u64* src, size_t src_gs, size_t in_stride, int n_in, u64* dst,
                       size_t dst_gs, int first, const BcJob* h_jobs, int n_jobs, const u64* h_chat,
                       const LimbC* h_limbs, int ns, int n_g, cudaStream_t st);
// (c0, c1, c0 + c1, 0) of a conversion constant c < 2^60 (30-bit Karatsuba split, see k_bconv)
inline uint4 chat_split(u64 c) {
    const u32 c0 = (u32)(c & 0x3fffffffull), c1 = (u32)(c >> 30);
    return make_uint4(c0, c1, c0 + c1, 0u);
}
// ---- plain forward COL pass in place over [n_g] blocks of n_slots polys (negative limb id = skip the slot);
// output is lazy [0,8q) and must be consumed by a forward ROW pass.
void launch_fwd_col_plain(const KCtx& c, u64* data, size_t gs, const int* d_limb_ids, int n_slots, int n_g,
                          cudaStream_t st);
// ---- EXPERIMENT: COL + ROW of an N = 2^16 limb in one kernel (8-CTA cluster, transpose through DSMEM); same result as
// launch_fwd_col_plain followed by launch_fwd_row_plain.  Returns false for other ring sizes.
bool launch_fwd_ntt_cluster(const KCtx& c, u64* data, size_t gs, const int* d_limb_ids, int n_slots, int n_g, cudaStream_t st,
                            bool lazy2);
// ---- centred SwitchModulus of ONE source limb with modulus q_src (rescale, DropLastElementAndScale) fused
// into the forward COL pass.
void launch_sw_fwd_col(const KCtx& c, const u64* src, size_t src_gs, u64* dst, size_t dst_gs, const BcJob* d_jobs,
                       int n_jobs, int n_g, u64 q_src, cudaStream_t st);

// ---- key inner product (hybrid key switch, EvalFastKeySwitchCoreExt) on digits that are already in EVAL form.
struct IpArgs {
    const u64* c1;       // EVAL c1 limbs of the batch: c1 + g*c1_gs + r*N (own-digit limbs are used as they are)
    size_t c1_gs;
    const u64* digits;   // COL-pass output of bc_fwd_col: digits + g*dig_gs + (j*(ell+K) + r)*N
    size_t dig_gs;
    const u64* key_b;    // [dnum][L+K][N]  (P limbs are read from here)
    const u64* key_a;
    const u64* key_qb;   // Q limbs pre-multiplied by P^-1 mod q_i: key_qb + (j*key_q_stride + i)*N  (ppq_rekey)
    const u64* key_qa;
    int key_q_stride;
    u64* extq;           // out, Q limbs:  extq + g*extq_gs + (poly*ell + r)*N, EVAL canonical
    size_t extq_gs;
    u64* extp;           // out, P limbs after the inverse ROW pass: extp + g*extp_gs + (poly*K + k)*N
    size_t extp_gs;
    int ell, K, L, parts, alpha;
    // deferred ModDown (ppq_agg_*): agg != 0 sums the Q-limb results (+ c0 on poly 0) into the slot's running EVAL
    // accumulator acc + g*acc_gs + (poly*ell + r)*N instead of writing extq (1: first client, store; 2: add)
    int agg = 0;
    u64* acc = nullptr;
    size_t acc_gs = 0;
    const u64* c0 = nullptr;  // c0 + g*c0_gs + r*N
    size_t c0_gs = 0;
};
// ---- several clients per inner-product launch (ppq_agg_add_many): client cl < nb has its own input ciphertexts, key
// and P-limb output block; digits of client cl, slot g sit at a.digits + (cl*n_g + g)*a.dig_gs.  A CTA sums the Q-limb
// results of the nb clients of a slot in registers, so the slot's running sum is read and written once per nb clients.
struct IpMulti {
    int nb = 1;
    const u64* ct_in[4] = {};   // [n_g][2][ell][N] of each client (c0 then c1 limbs), slot stride a.c1_gs
    const u64* key_b[4] = {};
    const u64* key_a[4] = {};
    const u64* key_qb[4] = {};
    const u64* key_qa[4] = {};
    int key_q_stride[4] = {};
    size_t extp_cs = 0;         // client stride of the P-limb output: a.extp + cl*extp_cs + g*a.extp_gs
};
bool launch_ip_agg(const KCtx& c, const IpArgs& a, const IpMulti& m, int n_g, cudaStream_t st);
// ---- split form: digits already in EVAL (k_fwd_row_plain on a.digits); pointwise inner product with the key tile
// staged in shared memory by TMA bulk copies and reused across the wave.  extp receives EVAL P limbs (run
// launch_inv_row on it afterwards).  Returns false if the shape is unsupported (parts > 8 or N < 512).
bool launch_ip_tma(const KCtx& c, const IpArgs& a, int n_g, cudaStream_t st);

// ---- forward ROW pass fused with the ModDown combine (+ c0, + optional accumulate = EvalAdd).
struct MdArgs {
    const u64* conv;   // COL-pass output: conv + g*conv_gs + (poly*ell + i)*N
    size_t conv_gs;
    const u64* extq;   // extq + g*extq_gs + (poly*ell + i)*N
    size_t extq_gs;
    const u64* ct_in;  // original ciphertexts [g][2][ell][N] (c0 is added to poly 0)
    size_t in_gs;
    u64* out;          // [g][2][ell][N]
    size_t out_gs;
    int ell;
    int accumulate;    // 0: out = result ; 1: out = out + result mod q (EvalAdd into the running sum)
};
// ---- key_q = key * P^-1 mod q_i on the Q limbs of a re-encryption key (once per key): with the ModDown conversion
// constants scaled the same way, (x - conv) * P^-1 becomes x' - conv' and the ModDown epilogue has no multiplication
void launch_scale_key(const KCtx& c, const u64* in, int in_stride, u64* out, int out_stride, const u64* d_pinv,
                      const u64* d_pinv_sh, int L, int dnum, cudaStream_t st);
void launch_fwd_row_moddown(const KCtx& c, const MdArgs& a, int n_g, cudaStream_t st);

// ---- forward ROW pass fused with the rescale combine: out = (x - v) * cst[i] mod q_i
struct RsArgs {
    const u64* conv;   // conv + g*conv_gs + (poly*(ell-1) + i)*N
    size_t conv_gs;
    const u64* ct_in;  // [g][2][ell][N]
    size_t in_gs;
    u64* out;          // [g][2][ell-1][N]
    size_t out_gs;
    u64 cst[64];       // [ell-1] per-limb factor q_last^-1 * EvalMult constant, BY VALUE: the call stays asynchronous
    u64 cst_sh[64];
    int ell;
};
void launch_fwd_row_rescale(const KCtx& c, const RsArgs& a, int n_g, cudaStream_t st);

// ---- pointwise helpers
// out = a + b mod q   (layout [n][2][ell][N], limb i -> modulus i)
void launch_eval_add(const KCtx& c, const u64* a, const u64* b, u64* out, int ell, size_t n_ct, cudaStream_t st);
// x = x mod q for lazily summed residues (after a cross-rank u64 sum)
void launch_finalize_mod(const KCtx& c, u64* x, int ell, size_t n_ct, cudaStream_t st);
// out = in * f[i] mod q_i
void launch_mult_scalar(const KCtx& c, const u64* in, u64* out, const u64* d_f, const u64* d_f_sh, int ell,
                        size_t n_ct, cudaStream_t st);
// per-limb factors passed by value (kernel parameter): nothing to stage, nothing to wait for
struct FacPack {
    u64 f[64], fs[64];
};
void launch_mult_scalar_pack(const KCtx& c, const u64* in, u64* out, const FacPack& p, int ell, size_t n_rows, cudaStream_t st);
// packed residues: limb i of a polynomial = N residues of bitlen(q_i) bits as a little-endian bit stream; row y
// ([ct][poly][limb] flattened) of a batch starts at word (y / ell) * words + off[y % ell]
struct PackLay {
    u32 off[64];  // word offset of limb i inside a packed polynomial
    u32 words;    // words of one packed polynomial (ell limbs)
};
void launch_unpack(const KCtx& c, const u64* in, u64* out, const PackLay& lay, int ell, size_t n_rows, cudaStream_t st);
void launch_pack(const KCtx& c, const u64* in, u64* out, const PackLay& lay, int ell, size_t n_rows, cudaStream_t st);
// same over a flat list of n_polys polynomials whose limb index is (poly index % ell)
void launch_mult_scalar_polys(const KCtx& c, const u64* in, u64* out, const u64* d_f, const u64* d_f_sh, int ell,
                              size_t n_polys, cudaStream_t st);
// forward NTT of small signed polynomials on the first `ell` limbs: out[g][s][i] = NTT_i(small[g][s]), g < n_groups,
// s < n_small (the residues are formed inside the COL pass); d_cyc_ids: limb ids 0..ell-1 repeated (>= n_groups*n_small*ell)
// pt_coeff != nullptr ([n_groups][N] int64, needs n_small >= 2): slot 1 becomes NTT(small[g][1] + pt_coeff[g] * 2^pt_approx)
void launch_small_fwd_ntt(const KCtx& c, const int64_t* small, u64* out, int n_small, int ell, size_t n_groups,
                          const int* d_cyc_ids, cudaStream_t st, const int64_t* pt_coeff = nullptr, int pt_approx = 0);
// Encrypt from the sampled polynomials in one pair of launches: COL pass of small [n_ct][3][N] (v, e0, e1; the encoded
// plaintext coefficients pt_coeff * 2^pt_approx join e0 when given) into vee [n_ct][3][ell][N], then the ROW pass with the
// combine ct[g][0] = pkb*v + e0 (+ pt), ct[g][1] = pka*v + e1 as its epilogue (vee is never read back as EVAL data).
// pt: optional EVAL plaintext [n_ct][ell][N]
void launch_small_fwd_encrypt(const KCtx& c, const int64_t* small, u64* vee, int ell, size_t n_ct, const u64* pkb,
                              const u64* pka, const u64* pt, u64* ct, cudaStream_t st, const int64_t* pt_coeff = nullptr,
                              int pt_approx = 0);
// encrypt combine: ct[g][0][i] = pkb[i]*v + e0 + m ; ct[g][1][i] = pka[i]*v + e1   (all EVAL); pt may be nullptr (m = 0)
void launch_encrypt_pw(const KCtx& c, const u64* pkb, const u64* pka, const u64* vee /*[g][3][ell][N]*/,
                       const u64* pt /*[g][ell][N]*/, u64* ct, int ell, size_t n_ct, cudaStream_t st);
// Decrypt core in two launches: out[g][i] = iNTT_i(c0 + c1 * s), canonical COEFF residues; the combine is the prologue
// of the inverse ROW pass.  d_cyc_ids / d_scale4: limb ids 0..ell-1 repeated and the N^-1 factors in that order
void launch_decrypt_inv_ntt(const KCtx& c, const u64* ct, const u64* sk, u64* out, int ell, size_t n_ct,
                            const int* d_cyc_ids, const u64* d_scale4, cudaStream_t st);
// key generation: se [2][M][N] EVAL (s, e), a [M][N] -> sk [L][N], pkb [M][N] = e - a*s
void launch_keygen_pw(const KCtx& c, const u64* se, const u64* a, u64* sk, u64* pkb, int L, int M, cudaStream_t st);
// uee [dnum][3][M][N] EVAL (u, e0, e1), sk_old [L][N], d_pmod [L] = P mod q_r -> kb, ka [dnum][M][N]
void launch_rekeygen_pw(const KCtx& c, const u64* pkb, const u64* pka, const u64* uee, const u64* sk_old,
                        const u64* d_pmod, u64* kb, u64* ka, int L, int M, int alpha, int dnum, cudaStream_t st);

const char* kernels_build_info();
// kernels launched by this library since load (bench.py's gpu_launches)
unsigned long long kernels_launch_count();

}  // namespace ppq
