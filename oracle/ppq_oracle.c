/*
 * ppq_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE ONLY).  See ppq_oracle.h for scope and pin status.
 *
 * Every function names the OpenFHE source it restates (OpenFHE @ 32362085, absent from /root/reference)
 * and the reference call site that reaches it.  All outputs are canonical residues in [0, q), so
 * bit-exactness depends only on the mathematical definitions (which root, which ordering, approximate
 * vs exact base conversion, centred vs plain lift), never on the reduction strategy used here.
 */
#include "ppq_oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;
typedef __int128 i128;
typedef uint64_t u64;

#define MAXLIMBS 128

struct ppqo_ctx {
    int logN, L, K, dnum, alpha, extra_bits;
    size_t N;
    u64 mod[MAXLIMBS];  /* q_0..q_{L-1}, p_0..p_{K-1} */
    u64 root[MAXLIMBS]; /* minimal primitive 2N-th roots */
    u64* psi;           /* [M][N] psi^{brev(k)} */
    u64* psi_sh;        /* Shoup companions floor(w*2^64/q) */
    u64* ipsi;          /* [M][N] psi^{-brev(k)} */
    u64* ipsi_sh;
    u64 ninv[MAXLIMBS], ninv_sh[MAXLIMBS];
    double sf[MAXLIMBS], sfbig[MAXLIMBS];
    /* CKKS special FFT tables (core/lib/math/dftransform.cpp) */
    size_t* rot_group; /* 5^i mod 2N, i < N/2 */
    double* ksi_re;    /* cos(2 pi j / 2N), j <= 2N */
    double* ksi_im;
};

/* ------------------------------------------------------------------ number theory (nbtheory) */
u64 ppqo_mulmod(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }

u64 ppqo_powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    while (e) {
        if (e & 1) r = ppqo_mulmod(r, a, q);
        a = ppqo_mulmod(a, a, q);
        e >>= 1;
    }
    return r;
}

u64 ppqo_invmod(u64 a, u64 q) { return ppqo_powmod(a, q - 2, q); /* q prime */ }

int ppqo_is_prime(u64 n) {
    /* deterministic Miller-Rabin for 64-bit n */
    static const u64 bases[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return 0;
    for (size_t i = 0; i < sizeof(bases) / sizeof(bases[0]); i++) {
        if (n == bases[i]) return 1;
        if (n % bases[i] == 0) return 0;
    }
    u64 d = n - 1;
    int s = 0;
    while ((d & 1) == 0) { d >>= 1; s++; }
    for (size_t i = 0; i < sizeof(bases) / sizeof(bases[0]); i++) {
        u64 x = ppqo_powmod(bases[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) {
            x = ppqo_mulmod(x, x, n);
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* nbtheory FirstPrime(nBits, m): smallest prime >= 2^nBits + 1 that is 1 mod m */
u64 ppqo_first_prime(int bits, u64 m) {
    u64 q = (u64)1 << bits;
    u64 r = q % m;
    u64 qn = q + 1 - r;
    if (r > 0) qn += m;
    while (!ppqo_is_prime(qn)) qn += m;
    return qn;
}
/* nbtheory PreviousPrime(q, m): next smaller prime congruent to q mod m (strictly below q) */
u64 ppqo_previous_prime(u64 q, u64 m) {
    do { q -= m; } while (!ppqo_is_prime(q));
    return q;
}
u64 ppqo_next_prime(u64 q, u64 m) {
    do { q += m; } while (!ppqo_is_prime(q));
    return q;
}

/* nbtheory RootOfUnity(m, q): OpenFHE returns the MINIMUM over all primitive m-th roots (m = 2N, power of 2) */
u64 ppqo_min_root_of_unity(u64 m, u64 q) {
    /* find any primitive m-th root: g^((q-1)/m) with g a non-residue => order exactly m (m power of two) */
    u64 r = 0;
    for (u64 g = 2;; g++) {
        r = ppqo_powmod(g, (q - 1) / m, q);
        if (ppqo_powmod(r, m / 2, q) == q - 1) break;
    }
    /* all primitive roots are r^k, k odd */
    u64 r2 = ppqo_mulmod(r, r, q);
    u64 x = r, best = r;
    for (u64 k = 1; k < m; k += 2) {
        if (x < best) best = x;
        x = ppqo_mulmod(x, r2, q);
    }
    return best;
}

/* ------------------------------------------------------------------ small helpers */
static inline u64 shoup_of(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
static inline u64 mulmod_sh(u64 x, u64 w, u64 wsh, u64 q) {
    u64 hi = (u64)(((u128)x * wsh) >> 64);
    u64 r = x * w - hi * q;
    return r - (q & (0 - (u64)(r >= q))); /* branch-free: the comparison is a coin flip on random residues */
}
/* 128-bit value mod q by Barrett reduction, as OpenFHE reduces the 128-bit accumulators of ApproxSwitchCRTBasis
 * (BarrettUint128ModUint64) -- same residues as `%`, without the libgcc division (__umodti3) that made the first
 * version of this port spend most of its time dividing.  mu = floor(2^(64+sh) / q), sh = bitlen(q) - 1; valid for
 * x < 2^(63 + bitlen(q)), e.g. a sum of 8 products of a 60-bit value and a residue. */
typedef struct { u64 q, mu; int sh; } bar_t;
static inline bar_t bar_of(u64 q) {
    bar_t b;
    b.q = q;
    b.sh = 63 - __builtin_clzll(q);
    b.mu = (u64)((((u128)1) << (64 + b.sh)) / q);
    return b;
}
static inline u64 bar_red(u128 x, bar_t b) {
    u64 x1 = (u64)(x >> b.sh);
    u64 qh = (u64)(((u128)x1 * b.mu) >> 64);
    u64 r = (u64)x - qh * b.q;
    while (r >= b.q) r -= b.q;
    return r;
}
static inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return s - (q & (0 - (u64)(s >= q))); }
static inline u64 submod(u64 a, u64 b, u64 q) { return a - b + (q & (0 - (u64)(a < b))); }
static inline size_t brev(size_t x, int bits) {
    size_t r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}
static inline u64 signed_mod(i128 v, u64 q) {
    i128 r = v % (i128)q;
    if (r < 0) r += q;
    return (u64)r;
}

/* multiword helpers (little-endian u64 words) */
static int mw_bitlen(const u64* w, int n) {
    for (int i = n - 1; i >= 0; i--)
        if (w[i]) return i * 64 + (64 - __builtin_clzll(w[i]));
    return 0;
}
static void mw_mul_word(u64* w, int n, u64 m) { /* w *= m, in place, n words (caller guarantees room) */
    u64 carry = 0;
    for (int i = 0; i < n; i++) {
        u128 t = (u128)w[i] * m + carry;
        w[i] = (u64)t;
        carry = (u64)(t >> 64);
    }
}
static void mw_add_word(u64* w, int n, u64 a) {
    for (int i = 0; i < n && a; i++) {
        u64 s = w[i] + a;
        a = s < a;
        w[i] = s;
    }
}
static int mw_cmp(const u64* a, const u64* b, int n) {
    for (int i = n - 1; i >= 0; i--) {
        if (a[i] != b[i]) return a[i] > b[i] ? 1 : -1;
    }
    return 0;
}
static void mw_sub(u64* r, const u64* a, const u64* b, int n) { /* r = a - b, a >= b */
    u64 borrow = 0;
    for (int i = 0; i < n; i++) {
        u64 bi = b[i] + borrow;
        u64 nb = (bi < borrow) || (a[i] < bi);
        r[i] = a[i] - bi;
        borrow = nb;
    }
}
static double mw_to_double(const u64* w, int n) {
    double r = 0.0;
    for (int i = n - 1; i >= 0; i--) r = r * 18446744073709551616.0 + (double)w[i];
    return r;
}

/* ------------------------------------------------------------------ context */
static void build_tables(ppqo_ctx* c) {
    int M = c->L + c->K;
    size_t N = c->N;
    c->psi = malloc(sizeof(u64) * M * N);
    c->psi_sh = malloc(sizeof(u64) * M * N);
    c->ipsi = malloc(sizeof(u64) * M * N);
    c->ipsi_sh = malloc(sizeof(u64) * M * N);
    for (int r = 0; r < M; r++) {
        u64 q = c->mod[r];
        c->root[r] = ppqo_min_root_of_unity(2 * N, q);
        u64 psi = c->root[r], ipsi = ppqo_invmod(psi, q);
        u64 *t = c->psi + (size_t)r * N, *ts = c->psi_sh + (size_t)r * N;
        u64 *it = c->ipsi + (size_t)r * N, *its = c->ipsi_sh + (size_t)r * N;
        u64 pw = 1, ipw = 1;
        for (size_t k = 0; k < N; k++) {
            size_t b = brev(k, c->logN);
            t[b] = pw;
            it[b] = ipw;
            pw = ppqo_mulmod(pw, psi, q);
            ipw = ppqo_mulmod(ipw, ipsi, q);
        }
        for (size_t k = 0; k < N; k++) {
            ts[k] = shoup_of(t[k], q);
            its[k] = shoup_of(it[k], q);
        }
        c->ninv[r] = ppqo_invmod((u64)N % q, q);
        c->ninv_sh[r] = shoup_of(c->ninv[r], q);
    }
    /* dftransform.cpp: m_rotGroup, m_ksiPows for cyclotomic order m = 2N */
    size_t m = 2 * N, Nh = N / 2;
    c->rot_group = malloc(sizeof(size_t) * Nh);
    c->ksi_re = malloc(sizeof(double) * (m + 1));
    c->ksi_im = malloc(sizeof(double) * (m + 1));
    size_t five = 1;
    for (size_t i = 0; i < Nh; i++) {
        c->rot_group[i] = five;
        five = (five * 5) % m;
    }
    for (size_t j = 0; j < m; j++) {
        double ang = 2.0 * M_PI * (double)j / (double)m;
        c->ksi_re[j] = cos(ang);
        c->ksi_im[j] = sin(ang);
    }
    c->ksi_re[m] = c->ksi_re[0];
    c->ksi_im[m] = c->ksi_im[0];
}

/* rns-cryptoparameters.cpp PrecomputeCRTTables (HYBRID): alpha, K and the P primes;
 * ckksrns-cryptoparameters.cpp: m_scalingFactorsReal / m_scalingFactorsRealBig (FLEXIBLEAUTO*) */
static void derive_aux(ppqo_ctx* c, int aux_bits) {
    int L = c->L;
    u64 m = 2 * (u64)c->N;
    c->alpha = (L + c->dnum - 1) / c->dnum;
    if (L - c->alpha * (c->dnum - 1) <= 0) {
        fprintf(stderr, "ppqo: bad dnum %d for L=%d\n", c->dnum, L);
        abort();
    }
    int maxbits = 0;
    for (int j = 0; j < c->dnum; j++) {
        u64 w[MAXLIMBS + 1];
        memset(w, 0, sizeof(w));
        w[0] = 1;
        int s = j * c->alpha, e = s + c->alpha > L ? L : s + c->alpha;
        for (int i = s; i < e; i++) mw_mul_word(w, MAXLIMBS + 1, c->mod[i]);
        int b = mw_bitlen(w, MAXLIMBS + 1);
        if (b > maxbits) maxbits = b;
    }
    c->K = (maxbits + aux_bits - 1) / aux_bits;
    u64 pprev = ppqo_first_prime(aux_bits, m);
    for (int k = 0; k < c->K; k++) {
        int found;
        u64 p;
        do {
            p = ppqo_previous_prime(pprev, m);
            found = 0;
            for (int i = 0; i < L; i++)
                if (c->mod[i] == p) found = 1;
            pprev = p;
        } while (found);
        c->mod[L + k] = p;
    }
    /* scaling factors */
    memset(c->sf, 0, sizeof(c->sf));
    memset(c->sfbig, 0, sizeof(c->sfbig));
    c->sf[0] = (double)c->mod[L - 1];
    int preset = 1;
    if (c->extra_bits > 0 && L >= 2) {
        c->sf[1] = (double)c->mod[L - 2];
        preset = 2;
    }
    for (int k = preset; k < L; k++) {
        double prev = c->sf[k - 1];
        c->sf[k] = prev * prev / (double)c->mod[L - k];
    }
    if (L >= 2) {
        c->sfbig[0] = (c->extra_bits == 0) ? c->sf[0] * c->sf[0] : c->sf[0] * c->sf[1];
        for (int k = 1; k < L - 1; k++) c->sfbig[k] = c->sf[k] * c->sf[k];
    }
}

ppqo_ctx* ppqo_ctx_create(int logN, int L, const u64* q_moduli, int dnum, int aux_bits, int extra_bits) {
    ppqo_ctx* c = calloc(1, sizeof(*c));
    c->logN = logN;
    c->N = (size_t)1 << logN;
    c->L = L;
    c->dnum = dnum;
    c->extra_bits = extra_bits;
    for (int i = 0; i < L; i++) c->mod[i] = q_moduli[i];
    derive_aux(c, aux_bits);
    build_tables(c);
    return c;
}

/* ckksrns-parametergeneration.cpp (FLEXIBLEAUTOEXT branch): numPrimes = depth+1 scaling-chain primes plus
 * one extra prime of (extra_bits-1) bits (verified against CC.json for depth 2: SURVEY 0.3). */
ppqo_ctx* ppqo_ctx_create_synthetic(int logN, int depth, int first_bits, int scale_bits, int extra_bits, int dnum) {
    u64 m = (u64)2 << logN;
    int np = depth + 1;
    u64 q[MAXLIMBS];
    q[np - 1] = ppqo_first_prime(scale_bits, m);
    double sf = (double)q[np - 1];
    int cnt = 0;
    for (int i = np - 2; i >= 1; i--) {
        sf = sf * sf / (double)q[i + 1];
        u64 sfi = (u64)llround(sf), rem = sfi % m, cand;
        int same;
        if ((cnt & 1) == 0) {
            cand = sfi - m - rem + 1;
            do {
                cand = ppqo_previous_prime(cand, m);
                same = 0;
                for (int j = i + 1; j < np; j++) if (q[j] == cand) same = 1;
            } while (same);
        } else {
            cand = sfi + m - rem + 1;
            do {
                cand = ppqo_next_prime(cand, m);
                same = 0;
                for (int j = i + 1; j < np; j++) if (q[j] == cand) same = 1;
            } while (same);
        }
        q[i] = cand;
        cnt++;
    }
    if (first_bits == scale_bits) {
        u64 mn = q[np - 1];
        for (int i = 1; i < np; i++) if (q[i] < mn) mn = q[i];
        q[0] = ppqo_previous_prime(mn, m);
    } else {
        q[0] = ppqo_previous_prime(ppqo_first_prime(first_bits, m), m);
    }
    int L = np;
    if (extra_bits > 0) {
        u64 e = ppqo_first_prime(extra_bits - 1, m);
        int same;
        do {
            same = 0;
            for (int j = 0; j < np; j++) if (q[j] == e) same = 1;
            if (same) e = ppqo_next_prime(e, m);
        } while (same);
        q[np] = e;
        L = np + 1;
    }
    return ppqo_ctx_create(logN, L, q, dnum, 60, extra_bits);
}

void ppqo_ctx_free(ppqo_ctx* c) {
    if (!c) return;
    free(c->psi); free(c->psi_sh); free(c->ipsi); free(c->ipsi_sh);
    free(c->rot_group); free(c->ksi_re); free(c->ksi_im);
    free(c);
}
int ppqo_logN(const ppqo_ctx* c) { return c->logN; }
int ppqo_L(const ppqo_ctx* c) { return c->L; }
int ppqo_K(const ppqo_ctx* c) { return c->K; }
int ppqo_dnum(const ppqo_ctx* c) { return c->dnum; }
int ppqo_alpha(const ppqo_ctx* c) { return c->alpha; }
void ppqo_moduli(const ppqo_ctx* c, u64* out) { memcpy(out, c->mod, sizeof(u64) * (c->L + c->K)); }
void ppqo_roots(const ppqo_ctx* c, u64* out) { memcpy(out, c->root, sizeof(u64) * (c->L + c->K)); }
double ppqo_scaling_factor(const ppqo_ctx* c, int level) { return c->sf[level]; }
double ppqo_scaling_factor_big(const ppqo_ctx* c, int level) { return c->sfbig[level]; }

/* ------------------------------------------------------------------ NTT
 * transformnat-impl.h ChineseRemainderTransformFTT: ForwardTransformToBitReverseInPlace (Cooley-Tukey,
 * natural in, bit-reversed out, twiddles psi^{brev(k)}) and InverseTransformFromBitReverseInPlace
 * (Gentleman-Sande, times N^-1).  EVAL[i] = sum_k a_k psi^{(2 brev(i)+1) k}. */
void ppqo_ntt_fwd(const ppqo_ctx* c, int limb, u64* a) {
    size_t N = c->N;
    u64 q = c->mod[limb];
    const u64 *w = c->psi + (size_t)limb * N, *ws = c->psi_sh + (size_t)limb * N;
    size_t t = N;
    for (size_t m = 1; m < N; m <<= 1) {
        t >>= 1;
        for (size_t i = 0; i < m; i++) {
            size_t j1 = 2 * i * t;
            u64 W = w[m + i], Ws = ws[m + i];
            for (size_t j = j1; j < j1 + t; j++) {
                u64 U = a[j], V = mulmod_sh(a[j + t], W, Ws, q);
                a[j] = addmod(U, V, q);
                a[j + t] = submod(U, V, q);
            }
        }
    }
}

void ppqo_ntt_inv(const ppqo_ctx* c, int limb, u64* a) {
    size_t N = c->N;
    u64 q = c->mod[limb];
    const u64 *w = c->ipsi + (size_t)limb * N, *ws = c->ipsi_sh + (size_t)limb * N;
    size_t t = 1;
    for (size_t m = N; m > 1; m >>= 1) {
        size_t j1 = 0, h = m >> 1;
        for (size_t i = 0; i < h; i++) {
            u64 W = w[h + i], Ws = ws[h + i];
            for (size_t j = j1; j < j1 + t; j++) {
                u64 U = a[j], V = a[j + t];
                a[j] = addmod(U, V, q);
                a[j + t] = mulmod_sh(submod(U, V, q), W, Ws, q);
            }
            j1 += 2 * t;
        }
        t <<= 1;
    }
    for (size_t j = 0; j < N; j++) a[j] = mulmod_sh(a[j], c->ninv[limb], c->ninv_sh[limb], q);
}

void ppqo_ntt_fwd_naive(const ppqo_ctx* c, int limb, const u64* a, u64* out) {
    size_t N = c->N;
    u64 q = c->mod[limb], psi = c->root[limb];
    for (size_t i = 0; i < N; i++) {
        u64 x = ppqo_powmod(psi, 2 * brev(i, c->logN) + 1, q), pw = 1, acc = 0;
        for (size_t k = 0; k < N; k++) {
            acc = addmod(acc, ppqo_mulmod(a[k] % q, pw, q), q);
            pw = ppqo_mulmod(pw, x, q);
        }
        out[i] = acc;
    }
}

/* ------------------------------------------------------------------ hybrid key switching
 * keyswitch-hybrid.cpp: EvalKeySwitchPrecomputeCore / EvalFastKeySwitchCoreExt / EvalFastKeySwitchCore;
 * dcrtpoly-impl.h: ApproxSwitchCRTBasis, ApproxModDown.  Reached from cc->ReEncrypt
 * (server/src/changeCipherDomain.cpp:74,89,105). */
int ppqo_num_parts(const ppqo_ctx* c, int ell) {
    int p = (ell + c->alpha - 1) / c->alpha;
    return p > c->dnum ? c->dnum : p;
}
static inline int ext_limb(const ppqo_ctx* c, int ell, int r) { return r < ell ? r : c->L + (r - ell); }

/* ApproxSwitchCRTBasis: src limbs (COEFF, canonical) with moduli sm[0..ns) -> one target modulus t.
 * y_i = x_i * (D/q_i)^-1 mod q_i ;  out = (sum_i y_i * ((D/q_i) mod t)) mod t  -- no overflow correction. */
static void approx_switch(size_t N, int ns, const u64* sm, u64* const* y /* already scaled */, u64 t, u64* out) {
    u64 chat[MAXLIMBS];
    for (int i = 0; i < ns; i++) {
        u64 p = 1 % t;
        for (int k = 0; k < ns; k++)
            if (k != i) p = ppqo_mulmod(p, sm[k] % t, t);
        chat[i] = p;
    }
    const bar_t bt = bar_of(t);
    if (ns <= 8) { /* the common case: one reduction per output, no test inside the product loop */
        for (size_t n = 0; n < N; n++) {
            u128 acc = 0;
            for (int i = 0; i < ns; i++) acc += (u128)y[i][n] * chat[i];
            out[n] = bar_red(acc, bt);
        }
        return;
    }
    for (size_t n = 0; n < N; n++) {
        u128 acc = 0;
        for (int i = 0; i < ns; i++) {
            acc += (u128)y[i][n] * chat[i];
            if ((i & 7) == 7 && i + 1 < ns) acc = bar_red(acc, bt); /* at most 8 products (+ one residue) per reduction */
        }
        out[n] = bar_red(acc, bt);
    }
}
static void scale_by_hat_inv(size_t N, int ns, const u64* sm, u64** x) {
    for (int i = 0; i < ns; i++) {
        u64 q = sm[i], p = 1;
        for (int k = 0; k < ns; k++)
            if (k != i) p = ppqo_mulmod(p, sm[k] % q, q);
        u64 inv = ppqo_invmod(p, q), invs = shoup_of(inv, q);
        for (size_t n = 0; n < N; n++) x[i][n] = mulmod_sh(x[i][n], inv, invs, q);
    }
}

void ppqo_modup(const ppqo_ctx* c, int ell, const u64* c1, u64* digits) {
    size_t N = c->N;
    int M = ell + c->K, parts = ppqo_num_parts(c, ell);
    for (int j = 0; j < parts; j++) {
        int s = j * c->alpha, e = s + c->alpha > ell ? ell : s + c->alpha, ns = e - s;
        u64* buf = malloc(sizeof(u64) * N * ns);
        u64* y[MAXLIMBS];
        u64 sm[MAXLIMBS];
        for (int i = 0; i < ns; i++) {
            y[i] = buf + (size_t)i * N;
            memcpy(y[i], c1 + (size_t)(s + i) * N, sizeof(u64) * N);
            ppqo_ntt_inv(c, s + i, y[i]);
            sm[i] = c->mod[s + i];
        }
        scale_by_hat_inv(N, ns, sm, y);
        u64* dj = digits + (size_t)j * M * N;
        for (int r = 0; r < M; r++) {
            u64* dst = dj + (size_t)r * N;
            if (r >= s && r < e) {
                memcpy(dst, c1 + (size_t)r * N, sizeof(u64) * N);
            } else {
                int g = ext_limb(c, ell, r);
                approx_switch(N, ns, sm, y, c->mod[g], dst);
                ppqo_ntt_fwd(c, g, dst);
            }
        }
        free(buf);
    }
}

void ppqo_inner_product(const ppqo_ctx* c, int ell, const u64* digits, const u64* kb, const u64* ka, u64* ct0, u64* ct1) {
    size_t N = c->N;
    int M = ell + c->K, MK = c->L + c->K, parts = ppqo_num_parts(c, ell);
    for (int r = 0; r < M; r++) {
        int g = ext_limb(c, ell, r); /* key limb index: r for Q limbs, L + k for P limbs */
        const bar_t bq = bar_of(c->mod[g]);
        u64 *o0 = ct0 + (size_t)r * N, *o1 = ct1 + (size_t)r * N;
        for (size_t n = 0; n < N; n++) { /* sum of <= 8 products of residues, reduced once (same residue as term by term) */
            u128 a0 = 0, a1 = 0;
            for (int j = 0; j < parts; j++) {
                u64 d = digits[((size_t)j * M + r) * N + n];
                a0 += (u128)d * kb[((size_t)j * MK + g) * N + n];
                a1 += (u128)d * ka[((size_t)j * MK + g) * N + n];
            }
            o0[n] = bar_red(a0, bq);
            o1[n] = bar_red(a1, bq);
        }
    }
}

void ppqo_moddown(const ppqo_ctx* c, int ell, const u64* ext, u64* out) {
    size_t N = c->N;
    int K = c->K, L = c->L;
    u64* buf = malloc(sizeof(u64) * N * K);
    u64* y[MAXLIMBS];
    const u64* pm = c->mod + L;
    for (int k = 0; k < K; k++) {
        y[k] = buf + (size_t)k * N;
        memcpy(y[k], ext + (size_t)(ell + k) * N, sizeof(u64) * N);
        ppqo_ntt_inv(c, L + k, y[k]);
    }
    scale_by_hat_inv(N, K, pm, y);
    u64* w = malloc(sizeof(u64) * N);
    for (int i = 0; i < ell; i++) {
        u64 q = c->mod[i];
        approx_switch(N, K, pm, y, q, w);
        ppqo_ntt_fwd(c, i, w);
        u64 P = 1;
        for (int k = 0; k < K; k++) P = ppqo_mulmod(P, pm[k] % q, q);
        u64 pinv = ppqo_invmod(P, q), pinvs = shoup_of(pinv, q);
        const u64* x = ext + (size_t)i * N;
        u64* o = out + (size_t)i * N;
        for (size_t n = 0; n < N; n++) o[n] = mulmod_sh(submod(x[n], w[n], q), pinv, pinvs, q);
    }
    free(w);
    free(buf);
}

void ppqo_keyswitch_core(const ppqo_ctx* c, int ell, const u64* c1, const u64* kb, const u64* ka, u64* ks0, u64* ks1) {
    size_t N = c->N;
    int M = ell + c->K, parts = ppqo_num_parts(c, ell);
    u64* digits = malloc(sizeof(u64) * N * M * parts);
    u64* t0 = malloc(sizeof(u64) * N * M);
    u64* t1 = malloc(sizeof(u64) * N * M);
    ppqo_modup(c, ell, c1, digits);
    ppqo_inner_product(c, ell, digits, kb, ka, t0, t1);
    ppqo_moddown(c, ell, t0, ks0);
    ppqo_moddown(c, ell, t1, ks1);
    free(digits); free(t0); free(t1);
}

/* base-pre.cpp PREBase::ReEncrypt with publicKey == nullptr (INDCPA, as the reference calls it):
 * result = clone; KeySwitchInPlace: c0 += ks0 ; c1 = ks1. */
void ppqo_reencrypt(const ppqo_ctx* c, int ell, const u64* ct_in, const u64* kb, const u64* ka, u64* ct_out) {
    size_t N = c->N, PS = (size_t)ell * N;
    u64* ks0 = malloc(sizeof(u64) * PS);
    ppqo_keyswitch_core(c, ell, ct_in + PS, kb, ka, ks0, ct_out + PS);
    for (int i = 0; i < ell; i++) {
        u64 q = c->mod[i];
        for (size_t n = 0; n < N; n++) ct_out[i * N + n] = addmod(ct_in[i * N + n], ks0[i * N + n], q);
    }
    free(ks0);
}

/* rns-leveledshe.cpp EvalAddCore on equal level / noiseScaleDeg (aggregateEncryptedWeights.cpp:82,91,106) */
void ppqo_eval_add(const ppqo_ctx* c, int ell, const u64* a, const u64* b, u64* out) {
    size_t N = c->N;
    for (int p = 0; p < 2; p++)
        for (int i = 0; i < ell; i++) {
            u64 q = c->mod[i];
            size_t o = ((size_t)p * ell + i) * N;
            for (size_t n = 0; n < N; n++) out[o + n] = addmod(a[o + n], b[o + n], q);
        }
}

/* dcrtpoly-impl.h DropLastElementAndScale (via ModReduceInternalInPlace, first half of EvalMult(ct, 0.5),
 * aggregateEncryptedWeights.cpp:83,92,107): out_i = (x_i - NTT_i(centred(r) mod q_i)) * q_last^-1 mod q_i,
 * centred lift threshold r > floor(q_last/2) (NativeVector::SwitchModulus). */
void ppqo_rescale(const ppqo_ctx* c, int ell, const u64* ct_in, u64* ct_out) {
    size_t N = c->N;
    int last = ell - 1;
    u64 ql = c->mod[last], half = ql >> 1;
    u64* r = malloc(sizeof(u64) * N);
    u64* t = malloc(sizeof(u64) * N);
    for (int p = 0; p < 2; p++) {
        memcpy(r, ct_in + ((size_t)p * ell + last) * N, sizeof(u64) * N);
        ppqo_ntt_inv(c, last, r);
        for (int i = 0; i < last; i++) {
            u64 q = c->mod[i], qlm = ql % q;
            for (size_t n = 0; n < N; n++) {
                u64 v = r[n] % q;
                t[n] = r[n] > half ? submod(v, qlm, q) : v;
            }
            ppqo_ntt_fwd(c, i, t);
            u64 inv = ppqo_invmod(qlm, q), invs = shoup_of(inv, q);
            const u64* x = ct_in + ((size_t)p * ell + i) * N;
            u64* o = ct_out + ((size_t)p * last + i) * N;
            for (size_t n = 0; n < N; n++) o[n] = mulmod_sh(submod(x[n], t[n], q), inv, invs, q);
        }
    }
    free(r); free(t);
}

/* ckksrns-leveledshe.cpp GetElementForEvalMult (HAVE_INT128 build): large = (int128)(c * sf + 0.5) */
void ppqo_mult_const_factors(const ppqo_ctx* c, int ell, double sf, double cst, u64* factors) {
    int logSF = (int)ceil(log2(fabs(sf)));
    int logValid = logSF <= 125 ? logSF : 125;
    double approx = pow(2.0, logSF - logValid);
    i128 large = (i128)(cst / approx * sf + 0.5);
    for (int i = 0; i < ell; i++) factors[i] = signed_mod(large, c->mod[i]);
}
void ppqo_mult_scalar(const ppqo_ctx* c, int ell, const u64* factors, const u64* ct_in, u64* ct_out) {
    size_t N = c->N;
    for (int p = 0; p < 2; p++)
        for (int i = 0; i < ell; i++) {
            u64 q = c->mod[i], f = factors[i], fs = shoup_of(f, q);
            size_t o = ((size_t)p * ell + i) * N;
            for (size_t n = 0; n < N; n++) ct_out[o + n] = mulmod_sh(ct_in[o + n], f, fs, q);
        }
}

/* ------------------------------------------------------------------ CKKS encoding
 * dftransform.cpp FFTSpecialInv / FFTSpecial; ckkspackedencoding.cpp Encode (NATIVEINT 64 branch) / Decode.
 * Reached from MakeCKKSPackedPlaintext (encryptModelWeights.cpp:82,90,109) and GetRealPackedValue
 * (decryptModelWeights.cpp:83,92,109).  Full packing: slots = N/2. */
static void bit_reverse_cplx(double* re, double* im, size_t n) {
    for (size_t i = 1, j = 0; i < n; i++) {
        size_t bit = n >> 1;
        for (; j >= bit; bit >>= 1) j -= bit;
        j += bit;
        if (i < j) {
            double t = re[i]; re[i] = re[j]; re[j] = t;
            t = im[i]; im[i] = im[j]; im[j] = t;
        }
    }
}
static void fft_special_inv(const ppqo_ctx* c, double* re, double* im, size_t size) {
    size_t m = 2 * c->N;
    for (size_t len = size; len >= 1; len >>= 1) {
        size_t lenh = len >> 1, lenq = len << 2, gap = m / lenq;
        for (size_t i = 0; i < size; i += len)
            for (size_t j = 0; j < lenh; j++) {
                size_t idx = (lenq - (c->rot_group[j] % lenq)) * gap;
                double ur = re[i + j] + re[i + j + lenh], ui = im[i + j] + im[i + j + lenh];
                double vr = re[i + j] - re[i + j + lenh], vi = im[i + j] - im[i + j + lenh];
                double wr = c->ksi_re[idx], wi = c->ksi_im[idx];
                re[i + j] = ur; im[i + j] = ui;
                re[i + j + lenh] = vr * wr - vi * wi;
                im[i + j + lenh] = vr * wi + vi * wr;
            }
    }
    bit_reverse_cplx(re, im, size);
    for (size_t i = 0; i < size; i++) { re[i] /= (double)size; im[i] /= (double)size; }
}
static void fft_special(const ppqo_ctx* c, double* re, double* im, size_t size) {
    size_t m = 2 * c->N;
    bit_reverse_cplx(re, im, size);
    for (size_t len = 2; len <= size; len <<= 1) {
        size_t lenh = len >> 1, lenq = len << 2, gap = m / lenq;
        for (size_t i = 0; i < size; i += len)
            for (size_t j = 0; j < lenh; j++) {
                size_t idx = (c->rot_group[j] % lenq) * gap;
                double wr = c->ksi_re[idx], wi = c->ksi_im[idx];
                double xr = re[i + j + lenh], xi = im[i + j + lenh];
                double vr = xr * wr - xi * wi, vi = xr * wi + xi * wr;
                double ur = re[i + j], ui = im[i + j];
                re[i + j] = ur + vr; im[i + j] = ui + vi;
                re[i + j + lenh] = ur - vr; im[i + j + lenh] = ui - vi;
            }
    }
}

void ppqo_encode(const ppqo_ctx* c, int ell, const double* vals, size_t nvals, double sf, u64* pt) {
    size_t N = c->N, Nh = N / 2;
    double* re = calloc(Nh, sizeof(double));
    double* im = calloc(Nh, sizeof(double));
    for (size_t i = 0; i < nvals && i < Nh; i++) re[i] = vals[i];
    fft_special_inv(c, re, im, Nh);
    int64_t* co = malloc(sizeof(int64_t) * N);
    for (size_t i = 0; i < Nh; i++) {
        co[i] = llround(re[i] * sf);
        co[i + Nh] = llround(im[i] * sf);
    }
    for (int i = 0; i < ell; i++) {
        u64 q = c->mod[i];
        u64* o = pt + (size_t)i * N;
        for (size_t n = 0; n < N; n++) o[n] = signed_mod((i128)co[n], q);
        ppqo_ntt_fwd(c, i, o);
    }
    free(re); free(im); free(co);
}

/* rns-pke.cpp EncryptZeroCore + Encrypt: c0 = b*v + e0 + m ; c1 = a*v + e1 (noise scale 1 for CKKS);
 * public key towers beyond the plaintext's are dropped.  Samples are INJECTED (encryptModelWeights.cpp:83). */
void ppqo_small_to_eval(const ppqo_ctx* c, int limb, const int64_t* small, u64* out) {
    u64 q = c->mod[limb];
    for (size_t n = 0; n < c->N; n++) out[n] = signed_mod((i128)small[n], q);
    ppqo_ntt_fwd(c, limb, out);
}
void ppqo_encrypt(const ppqo_ctx* c, int ell, const u64* pkb, const u64* pka, const u64* pt, const int64_t* v,
                  const int64_t* e0, const int64_t* e1, u64* ct) {
    size_t N = c->N;
    u64* tv = malloc(sizeof(u64) * N);
    u64* te = malloc(sizeof(u64) * N);
    for (int i = 0; i < ell; i++) {
        u64 q = c->mod[i];
        ppqo_small_to_eval(c, i, v, tv);
        ppqo_small_to_eval(c, i, e0, te);
        u64* c0 = ct + (size_t)i * N;
        u64* c1 = ct + ((size_t)ell + i) * N;
        for (size_t n = 0; n < N; n++)
            c0[n] = addmod(addmod(ppqo_mulmod(pkb[(size_t)i * N + n], tv[n], q), te[n], q), pt[(size_t)i * N + n], q);
        ppqo_small_to_eval(c, i, e1, te);
        for (size_t n = 0; n < N; n++) c1[n] = addmod(ppqo_mulmod(pka[(size_t)i * N + n], tv[n], q), te[n], q);
    }
    free(tv); free(te);
}

/* rns-pke.cpp DecryptCore: b = c0 + c1*s (EVAL) then COEFFICIENT (decryptModelWeights.cpp:81,90,108) */
void ppqo_decrypt_core(const ppqo_ctx* c, int ell, const u64* sk, const u64* ct, u64* coeff) {
    size_t N = c->N;
    for (int i = 0; i < ell; i++) {
        u64 q = c->mod[i];
        u64* o = coeff + (size_t)i * N;
        for (size_t n = 0; n < N; n++)
            o[n] = addmod(ct[(size_t)i * N + n], ppqo_mulmod(ct[((size_t)ell + i) * N + n], sk[(size_t)i * N + n], q), q);
        ppqo_ntt_inv(c, i, o);
    }
}

/* Poly CRTInterpolate + CKKSPackedEncoding::Decode, WITHOUT OpenFHE's randomised decode noise */
void ppqo_decode(const ppqo_ctx* c, int ell, const u64* coeff, double sf, double* out) {
    size_t N = c->N, Nh = N / 2;
    int W = ell + 1;
    /* Garner constants: inv[i] = (q_0..q_{i-1})^-1 mod q_i */
    u64 ginv[MAXLIMBS];
    for (int i = 1; i < ell; i++) {
        u64 p = 1;
        for (int k = 0; k < i; k++) p = ppqo_mulmod(p, c->mod[k] % c->mod[i], c->mod[i]);
        ginv[i] = ppqo_invmod(p, c->mod[i]);
    }
    u64 Q[MAXLIMBS + 1], Qh[MAXLIMBS + 1];
    memset(Q, 0, sizeof(Q));
    Q[0] = 1;
    for (int i = 0; i < ell; i++) mw_mul_word(Q, W, c->mod[i]);
    for (int i = 0; i < W; i++) Qh[i] = (Q[i] >> 1) | (i + 1 < W ? Q[i + 1] << 63 : 0);
    double* re = malloc(sizeof(double) * Nh);
    double* im = malloc(sizeof(double) * Nh);
    for (size_t n = 0; n < N; n++) {
        u64 v[MAXLIMBS];
        v[0] = coeff[n];
        for (int i = 1; i < ell; i++) {
            u64 qi = c->mod[i];
            /* evaluate current mixed-radix number mod q_i */
            u64 acc = 0;
            for (int k = i - 1; k >= 0; k--) acc = addmod(ppqo_mulmod(acc, c->mod[k] % qi, qi), v[k] % qi, qi);
            v[i] = ppqo_mulmod(submod(coeff[(size_t)i * N + n], acc, qi), ginv[i], qi);
        }
        u64 X[MAXLIMBS + 1];
        memset(X, 0, sizeof(u64) * W);
        X[0] = v[ell - 1];
        for (int i = ell - 2; i >= 0; i--) {
            mw_mul_word(X, W, c->mod[i]);
            mw_add_word(X, W, v[i]);
        }
        double d;
        if (mw_cmp(X, Qh, W) > 0) {
            u64 T[MAXLIMBS + 1];
            mw_sub(T, Q, X, W);
            d = -mw_to_double(T, W);
        } else {
            d = mw_to_double(X, W);
        }
        d /= sf;
        if (n < Nh) re[n] = d; else im[n - Nh] = d;
    }
    fft_special(c, re, im, Nh);
    for (size_t i = 0; i < Nh; i++) out[i] = re[i];
    free(re); free(im);
}

void ppqo_decrypt(const ppqo_ctx* c, int ell, const u64* sk, const u64* ct, double sf, double* out) {
    u64* co = malloc(sizeof(u64) * c->N * ell);
    ppqo_decrypt_core(c, ell, sk, ct, co);
    ppqo_decode(c, ell, co, sf, out);
    free(co);
}

/* ------------------------------------------------------------------ samplers (test scaffolding; our own convention) */
static inline u64 splitmix64(u64* s) {
    u64 z = (*s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
void ppqo_sample_ternary(u64 seed, size_t n, int64_t* out) {
    u64 s = seed;
    for (size_t i = 0; i < n; i++) out[i] = (int64_t)(splitmix64(&s) % 3) - 1;
}
void ppqo_sample_gauss(u64 seed, size_t n, double sigma, int64_t* out) {
    u64 s = seed;
    for (size_t i = 0; i < n; i++) {
        double u1 = ((double)(splitmix64(&s) >> 11) + 1.0) / 9007199254740993.0;
        double u2 = (double)(splitmix64(&s) >> 11) / 9007199254740992.0;
        out[i] = llround(sigma * sqrt(-2.0 * log(u1)) * cos(2.0 * M_PI * u2));
    }
}
void ppqo_sample_uniform(u64 seed, u64 q, size_t n, u64* out) {
    u64 s = seed;
    for (size_t i = 0; i < n; i++) out[i] = splitmix64(&s) % q;
}

/* base-pke.cpp KeyGenInternal: s ternary, a uniform, e Gauss over QP (EVAL); b = e - a*s */
void ppqo_keygen(const ppqo_ctx* c, u64 seed, int64_t* s_coeff, u64* sk, u64* pkb, u64* pka) {
    size_t N = c->N;
    int M = c->L + c->K;
    int64_t* e = malloc(sizeof(int64_t) * N);
    u64* te = malloc(sizeof(u64) * N);
    ppqo_sample_ternary(seed * 1000003ULL + 1, N, s_coeff);
    ppqo_sample_gauss(seed * 1000003ULL + 2, N, 3.19, e);
    for (int r = 0; r < M; r++) {
        u64 q = c->mod[r];
        ppqo_small_to_eval(c, r, s_coeff, sk + (size_t)r * N);
        ppqo_small_to_eval(c, r, e, te);
        ppqo_sample_uniform(seed * 1000003ULL + 100 + r, q, N, pka + (size_t)r * N);
        for (size_t n = 0; n < N; n++)
            pkb[(size_t)r * N + n] = submod(te[n], ppqo_mulmod(pka[(size_t)r * N + n], sk[(size_t)r * N + n], q), q);
    }
    free(e); free(te);
}

/* keyswitch-hybrid.cpp KeySwitchGenInternal(oldSk, newPk), reached from cc->ReKeyGen (client/src/REkeyGen.cpp:52):
 * per level-0 digit j: b_j = pk.b*u + e0 (+ (P mod q_r) * s_old on limbs of digit j), a_j = pk.a*u + e1 */
void ppqo_rekeygen(const ppqo_ctx* c, u64 seed, const u64* sk_old, const u64* pkb, const u64* pka, u64* kb, u64* ka) {
    size_t N = c->N;
    int L = c->L, M = c->L + c->K;
    int64_t* u = malloc(sizeof(int64_t) * N);
    int64_t* e0 = malloc(sizeof(int64_t) * N);
    int64_t* e1 = malloc(sizeof(int64_t) * N);
    u64* tu = malloc(sizeof(u64) * N);
    u64* t0 = malloc(sizeof(u64) * N);
    u64* t1 = malloc(sizeof(u64) * N);
    for (int j = 0; j < c->dnum; j++) {
        int s = j * c->alpha, e = s + c->alpha > L ? L : s + c->alpha;
        ppqo_sample_ternary(seed * 7919ULL + 10 * j + 1, N, u);
        ppqo_sample_gauss(seed * 7919ULL + 10 * j + 2, N, 3.19, e0);
        ppqo_sample_gauss(seed * 7919ULL + 10 * j + 3, N, 3.19, e1);
        for (int r = 0; r < M; r++) {
            u64 q = c->mod[r];
            ppqo_small_to_eval(c, r, u, tu);
            ppqo_small_to_eval(c, r, e0, t0);
            ppqo_small_to_eval(c, r, e1, t1);
            u64 Pm = 1;
            for (int k = 0; k < c->K; k++) Pm = ppqo_mulmod(Pm, c->mod[L + k] % q, q);
            u64* ob = kb + ((size_t)j * M + r) * N;
            u64* oa = ka + ((size_t)j * M + r) * N;
            for (size_t n = 0; n < N; n++) {
                u64 b = addmod(ppqo_mulmod(pkb[(size_t)r * N + n], tu[n], q), t0[n], q);
                if (r >= s && r < e) b = addmod(b, ppqo_mulmod(Pm, sk_old[(size_t)r * N + n], q), q);
                ob[n] = b;
                oa[n] = addmod(ppqo_mulmod(pka[(size_t)r * N + n], tu[n], q), t1[n], q);
            }
        }
    }
    free(u); free(e0); free(e1); free(tu); free(t0); free(t1);
}

/* ------------------------------------------------------------------ batched CPU baseline */
int ppqo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ppqo_reencrypt_accumulate_batch(const ppqo_ctx* c, int ell, int n_clients, int n_slots, const u64* cts,
                                     const u64* const* kb, const u64* const* ka, u64* acc, int nthreads) {
    size_t N = c->N, CT = 2 * (size_t)ell * N;
    memset(acc, 0, sizeof(u64) * CT * n_slots);
    (void)nthreads;
    int nt = 1;
#ifdef _OPENMP
    nt = nthreads > 0 ? nthreads : omp_get_max_threads();
#endif
    if (n_slots >= nt) {
        /* enough slots for every thread: a thread owns whole slots and adds its clients in order -- no lock */
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic) num_threads(nt)
#endif
        for (int s = 0; s < n_slots; s++) {
            u64* tmp = malloc(sizeof(u64) * CT);
            for (int cl = 0; cl < n_clients; cl++) {
                ppqo_reencrypt(c, ell, cts + ((size_t)cl * n_slots + s) * CT, kb[cl], ka[cl], tmp);
                ppqo_eval_add(c, ell, acc + (size_t)s * CT, tmp, acc + (size_t)s * CT);
            }
            free(tmp);
        }
        return;
    }
    int total = n_clients * n_slots;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic) num_threads(nt)
#endif
    for (int w = 0; w < total; w++) {
        int cl = w / n_slots, s = w % n_slots;
        u64* tmp = malloc(sizeof(u64) * CT);
        ppqo_reencrypt(c, ell, cts + ((size_t)cl * n_slots + s) * CT, kb[cl], ka[cl], tmp);
#ifdef _OPENMP
#pragma omp critical(ppqo_acc)
#endif
        ppqo_eval_add(c, ell, acc + (size_t)s * CT, tmp, acc + (size_t)s * CT);
        free(tmp);
    }
}

void ppqo_ntt_batch(const ppqo_ctx* c, int inverse, const int* limb_ids, size_t n_polys, u64* data, int nthreads) {
    (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : omp_get_max_threads())
#endif
    for (size_t p = 0; p < n_polys; p++) {
        if (inverse) ppqo_ntt_inv(c, limb_ids[p], data + p * c->N);
        else ppqo_ntt_fwd(c, limb_ids[p], data + p * c->N);
    }
}
