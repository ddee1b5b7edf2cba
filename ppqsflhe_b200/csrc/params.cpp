// params.cpp -- host-side parameter derivation (PRODUCT code; see params.hpp).
//
// Restates, for the constants the kernels need, what OpenFHE computes at context load time:
//   core/lib/math/nbtheory*.cpp               FirstPrime / PreviousPrime / NextPrime / RootOfUnity (minimal root)
//   pke/lib/schemerns/rns-cryptoparameters.cpp PrecomputeCRTTables (HYBRID: alpha, P primes, QHat tables)
//   pke/lib/scheme/ckksrns/ckksrns-cryptoparameters.cpp  scaling factors (FLEXIBLEAUTO / FLEXIBLEAUTOEXT)
//   pke/lib/scheme/ckksrns/ckksrns-parametergeneration.cpp  the prime chain (synthetic contexts)
// OpenFHE itself is absent from /root/reference (un-vendored submodule, .SUBMODULES.json:9-15); the pins are
// server/storage/CC.json (primes + roots, reproduced exactly) -- see DESIGN.md.
#include "params.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace ppq {
namespace host {

typedef unsigned __int128 u128;
typedef __int128 i128;

u64 mulmod(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }
u64 powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
    }
    return r;
}
u64 invmod(u64 a, u64 q) { return powmod(a % q, q - 2, q); }
u64 shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }

bool is_prime(u64 n) {
    if (n < 2) return false;
    static const u64 sp[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (u64 p : sp) {
        if (n == p) return true;
        if (n % p == 0) return false;
    }
    u64 d = n - 1;
    int s = 0;
    while (!(d & 1)) { d >>= 1; ++s; }
    for (u64 a : sp) {  // deterministic for n < 2^64
        u64 x = powmod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool composite = true;
        for (int r = 1; r < s && composite; ++r) {
            x = mulmod(x, x, n);
            if (x == n - 1) composite = false;
        }
        if (composite) return false;
    }
    return true;
}

u64 first_prime(int bits, u64 m) {
    u64 q = (u64)1 << bits, r = q % m;
    u64 c = q + 1 - r;
    if (r > 0) c += m;
    while (!is_prime(c)) c += m;
    return c;
}
u64 previous_prime(u64 q, u64 m) {
    do q -= m; while (!is_prime(q));
    return q;
}
u64 next_prime(u64 q, u64 m) {
    do q += m; while (!is_prime(q));
    return q;
}

u64 min_root_of_unity(u64 m, u64 q) {
    u64 r = 0;
    for (u64 g = 2;; ++g) {  // m is a power of two: order exactly m  <=>  r^(m/2) = -1
        r = powmod(g, (q - 1) / m, q);
        if (powmod(r, m / 2, q) == q - 1) break;
    }
    const u64 r2 = mulmod(r, r, q);
    u64 best = r, x = r;
    for (u64 k = 1; k < m; k += 2) {  // all primitive roots are the odd powers
        best = std::min(best, x);
        x = mulmod(x, r2, q);
    }
    return best;
}

static int bitlen_product(const std::vector<u64>& f) {
    std::vector<u64> w(f.size() + 1, 0);
    w[0] = 1;
    for (u64 m : f) {
        u64 carry = 0;
        for (auto& x : w) {
            u128 t = (u128)x * m + carry;
            x = (u64)t;
            carry = (u64)(t >> 64);
        }
    }
    for (int i = (int)w.size() - 1; i >= 0; --i)
        if (w[i]) return i * 64 + 64 - __builtin_clzll(w[i]);
    return 0;
}

bool derive(Params& p, std::string& err) {
    if (p.logN < 12 || p.logN > 17) { err = "ring dimension must be 2^12..2^17"; return false; }
    p.N = (size_t)1 << p.logN;
    const u64 m = 2 * (u64)p.N;
    if (p.L < 1 || (int)p.mod.size() < p.L) { err = "missing Q moduli"; return false; }
    p.mod.resize(p.L);
    for (u64 q : p.mod)
        if (q >= ((u64)1 << 60) || q % m != 1 || !is_prime(q)) { err = "Q modulus must be a prime < 2^60 with q = 1 mod 2N"; return false; }
    if (p.dnum < 1 || p.dnum > 8) { err = "dnum must be in 1..8"; return false; }
    p.alpha = (p.L + p.dnum - 1) / p.dnum;
    if (p.L - p.alpha * (p.dnum - 1) <= 0) { err = "dnum too large for the number of limbs"; return false; }
    if (p.alpha > 32) { err = "more than 32 limbs per digit is not supported"; return false; }
    int maxbits = 0;
    for (int j = 0; j < p.dnum; ++j) {
        std::vector<u64> f(p.mod.begin() + j * p.alpha, p.mod.begin() + std::min(p.L, (j + 1) * p.alpha));
        maxbits = std::max(maxbits, bitlen_product(f));
    }
    p.K = (maxbits + p.aux_bits - 1) / p.aux_bits;
    if (p.K > 32) { err = "more than 32 auxiliary limbs is not supported"; return false; }
    u64 prev = first_prime(p.aux_bits, m);
    for (int k = 0; k < p.K; ++k) {
        u64 c;
        do {
            c = previous_prime(prev, m);
            prev = c;
        } while (std::find(p.mod.begin(), p.mod.begin() + p.L, c) != p.mod.begin() + p.L);
        p.mod.push_back(c);
    }
    const int M = p.M();
    p.root.resize(M);
    p.ninv.resize(M);
    for (int r = 0; r < M; ++r) {
        p.root[r] = min_root_of_unity(m, p.mod[r]);
        p.ninv[r] = invmod((u64)p.N % p.mod[r], p.mod[r]);
    }
    p.sf.assign(p.L, 0.0);
    p.sfbig.assign(p.L, 0.0);
    p.sf[0] = (double)p.mod[p.L - 1];
    int preset = 1;
    if (p.extra_bits > 0 && p.L >= 2) {
        p.sf[1] = (double)p.mod[p.L - 2];
        preset = 2;
    }
    for (int k = preset; k < p.L; ++k) p.sf[k] = p.sf[k - 1] * p.sf[k - 1] / (double)p.mod[p.L - k];
    if (p.L >= 2) {
        p.sfbig[0] = p.extra_bits == 0 ? p.sf[0] * p.sf[0] : p.sf[0] * p.sf[1];
        for (int k = 1; k < p.L - 1; ++k) p.sfbig[k] = p.sf[k] * p.sf[k];
    }
    return true;
}

std::vector<u64> synthetic_chain(int logN, int depth, int first_bits, int scale_bits, int extra_bits) {
    const u64 m = (u64)2 << logN;
    const int np = depth + 1;
    std::vector<u64> q(np + (extra_bits > 0 ? 1 : 0), 0);
    q[np - 1] = first_prime(scale_bits, m);
    double sf = (double)q[np - 1];
    int cnt = 0;
    auto used = [&](u64 c, int from, int to) {
        for (int j = from; j < to; ++j)
            if (q[j] == c) return true;
        return false;
    };
    for (int i = np - 2; i >= 1; --i) {
        sf = sf * sf / (double)q[i + 1];
        const u64 sfi = (u64)std::llround(sf), rem = sfi % m;
        u64 c;
        if ((cnt & 1) == 0) {
            c = sfi - m - rem + 1;
            do c = previous_prime(c, m); while (used(c, i + 1, np));
        } else {
            c = sfi + m - rem + 1;
            do c = next_prime(c, m); while (used(c, i + 1, np));
        }
        q[i] = c;
        ++cnt;
    }
    if (first_bits == scale_bits) {
        u64 mn = q[np - 1];
        for (int i = 1; i < np; ++i) mn = std::min(mn, q[i]);
        q[0] = previous_prime(mn, m);
    } else {
        q[0] = previous_prime(first_prime(first_bits, m), m);
    }
    if (extra_bits > 0) {
        u64 e = first_prime(extra_bits - 1, m);
        while (used(e, 0, np)) e = next_prime(e, m);
        q[np] = e;
    }
    return q;
}

// ---- minimal scanner for the cereal JSON CryptoContext (server/storage/CC.json)
static const char* find_key(const char* s, const char* end, const char* key) {
    const size_t n = std::strlen(key);
    for (const char* p = s; p + n + 2 <= end; ++p)
        if (*p == '"' && std::memcmp(p + 1, key, n) == 0 && p[n + 1] == '"') return p + n + 2;
    return nullptr;
}
static const char* skip_to_value(const char* p, const char* end) {
    while (p < end && (*p == ' ' || *p == ':' || *p == '\n' || *p == '\r' || *p == '\t')) ++p;
    return p;
}
static bool scalar_after(const char* s, const char* end, const char* key, u64& out) {
    const char* p = find_key(s, end, key);
    if (!p) return false;
    p = skip_to_value(p, end);
    if (p >= end || *p < '0' || *p > '9') return false;
    out = std::strtoull(p, nullptr, 10);
    return true;
}

bool parse_ccjson(const char* text, size_t len, Params& p, std::string& err) {
    const char* end = text + len;
    u64 rd = 0, v = 0;
    if (!scalar_after(text, end, "rd", rd) || rd == 0 || (rd & (rd - 1))) { err = "CC.json: ring dimension (rd) not found"; return false; }
    p.logN = 63 - __builtin_clzll(rd);
    p.mod.clear();
    std::vector<u64> roots;
    // every limb record is  "cm": {"cereal_class_version": 1, "v": <u64>}, "ru": {"v": <u64>} ; the composite
    // modulus has "v": [ ... ] and is skipped.
    const char* s = text;
    while (const char* c = find_key(s, end, "cm")) {
        const char* vp = find_key(c, end, "v");
        if (!vp) break;
        vp = skip_to_value(vp, end);
        s = vp;
        if (vp < end && *vp >= '0' && *vp <= '9') {
            p.mod.push_back(std::strtoull(vp, nullptr, 10));
            u64 ru = 0;
            const char* r = find_key(vp, end, "ru");
            if (r && scalar_after(r, end, "v", ru)) roots.push_back(ru);
        }
    }
    if (p.mod.empty()) { err = "CC.json: no limb moduli found"; return false; }
    p.L = (int)p.mod.size();
    p.dnum = scalar_after(text, end, "dnum", v) ? (int)v : 0;
    p.aux_bits = scalar_after(text, end, "ab", v) ? (int)v : 60;
    p.extra_bits = scalar_after(text, end, "eb", v) ? (int)v : 0;
    if (scalar_after(text, end, "ks", v) && v != 2) { err = "CC.json: only HYBRID key switching (ks = 2) is supported"; return false; }
    if (p.dnum == 0) { err = "CC.json: dnum not found"; return false; }
    if (!derive(p, err)) return false;
    for (size_t i = 0; i < roots.size() && i < (size_t)p.L; ++i)
        if (roots[i] != p.root[i]) { err = "CC.json: serialized root of unity differs from the derived minimal root"; return false; }
    return true;
}

static inline size_t brev(size_t x, int bits) {
    size_t r = 0;
    for (int i = 0; i < bits; ++i) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

std::vector<u64> scale4(const Params& p, const std::vector<int>& limb, const std::vector<u64>& f) {
    std::vector<u64> o(4 * limb.size());
    for (size_t i = 0; i < limb.size(); ++i) {
        const u64 q = p.mod[limb[i]];
        // inv[1] of build_twiddles: psi^-brev(1) = psi^-(N/2)
        const u64 w1 = powmod(invmod(p.root[limb[i]], q), (u64)p.N / 2, q);
        o[4 * i] = f[i];
        o[4 * i + 1] = shoup(f[i], q);
        o[4 * i + 2] = mulmod(f[i], w1, q);
        o[4 * i + 3] = shoup(o[4 * i + 2], q);
    }
    return o;
}

void build_twiddles(const Params& p, int limb, u64* fwd, u64* inv) {
    const u64 q = p.mod[limb], psi = p.root[limb], ipsi = invmod(psi, q);
    u64 pw = 1, ipw = 1;
    for (size_t k = 0; k < p.N; ++k) {
        const size_t b = brev(k, p.logN);
        fwd[2 * b] = pw;
        fwd[2 * b + 1] = shoup(pw, q);
        inv[2 * b] = ipw;
        inv[2 * b + 1] = shoup(ipw, q);
        pw = mulmod(pw, psi, q);
        ipw = mulmod(ipw, ipsi, q);
    }
}

// product of mods[first..last) except index skip, reduced mod t
static u64 hat_mod(const std::vector<u64>& mods, int first, int last, int skip, u64 t) {
    u64 r = 1 % t;
    for (int i = first; i < last; ++i)
        if (i != skip) r = mulmod(r, mods[i] % t, t);
    return r;
}

void build_ks_plan(const Params& p, int ell, KsPlan& o) {
    const int L = p.L, K = p.K, M = ell + K;
    o = KsPlan();
    o.ell = ell;
    o.parts = std::min((ell + p.alpha - 1) / p.alpha, p.dnum);
    o.c1_limb.resize(ell);
    o.c1_scale.resize(ell);
    o.c1_scale_sh.resize(ell);
    for (int j = 0; j < o.parts; ++j) {
        const int s = j * p.alpha, e = std::min(s + p.alpha, ell);
        for (int i = s; i < e; ++i) {
            const u64 q = p.mod[i];
            const u64 hatinv = invmod(hat_mod(p.mod, s, e, i, q), q);
            o.c1_limb[i] = i;
            o.c1_scale[i] = mulmod(hatinv, p.ninv[i], q);
            o.c1_scale_sh[i] = shoup(o.c1_scale[i], q);
        }
        for (int r = 0; r < M; ++r) {
            if (r >= s && r < e) continue;
            const int g = r < ell ? r : L + (r - ell);
            Job jb{s, e - s, j * M + r, g, (int)o.up_chat.size()};
            for (int i = s; i < e; ++i) o.up_chat.push_back(hat_mod(p.mod, s, e, i, p.mod[g]));
            o.up_jobs.push_back(jb);
        }
    }
    o.p_limb.resize(2 * K);
    o.p_scale.resize(2 * K);
    o.p_scale_sh.resize(2 * K);
    for (int poly = 0; poly < 2; ++poly)
        for (int k = 0; k < K; ++k) {
            const u64 q = p.mod[L + k];
            const u64 hatinv = invmod(hat_mod(p.mod, L, L + K, L + k, q), q);
            o.p_limb[poly * K + k] = L + k;
            o.p_scale[poly * K + k] = mulmod(hatinv, p.ninv[L + k], q);
            o.p_scale_sh[poly * K + k] = shoup(o.p_scale[poly * K + k], q);
        }
    o.pinv.resize(ell);
    o.pinv_sh.resize(ell);
    for (int i = 0; i < ell; ++i) {
        const u64 q = p.mod[i];
        o.pinv[i] = invmod(hat_mod(p.mod, L, L + K, -1, q), q);
        o.pinv_sh[i] = shoup(o.pinv[i], q);
    }
    for (int poly = 0; poly < 2; ++poly)
        for (int i = 0; i < ell; ++i) {
            Job jb{poly * K, K, poly * ell + i, i, (int)o.down_chat.size()};
            // [P/p_k]_{q_i} * P^-1 mod q_i: the ModDown multiplication by P^-1 is folded into the conversion (and, on
            // the other operand, into the key: launch_scale_key), both exact modulo q_i
            for (int k = 0; k < K; ++k)
                o.down_chat.push_back(mulmod(hat_mod(p.mod, L, L + K, L + k, p.mod[i]), o.pinv[i], p.mod[i]));
            o.down_jobs.push_back(jb);
        }
}

void build_rs_plan(const Params& p, int ell, RsPlan& o) {
    o = RsPlan();
    o.ell = ell;
    const int last = ell - 1;
    const u64 ql = p.mod[last];
    o.last_limb = {last, last};
    o.last_scale = {p.ninv[last], p.ninv[last]};
    o.last_scale_sh = {shoup(p.ninv[last], ql), shoup(p.ninv[last], ql)};
    o.qlinv.resize(last);
    for (int i = 0; i < last; ++i) o.qlinv[i] = invmod(ql % p.mod[i], p.mod[i]);
    for (int poly = 0; poly < 2; ++poly)
        for (int i = 0; i < last; ++i) o.jobs.push_back(Job{poly, 1, poly * last + i, i, 0});
}

void mult_const_factors(const Params& p, int ell, double sf, double c, u64* factors) {
    const int logSF = (int)std::ceil(std::log2(std::fabs(sf)));
    const int logValid = logSF <= 125 ? logSF : 125;
    const double approx = std::pow(2.0, logSF - logValid);
    const i128 large = (i128)(c / approx * sf + 0.5);
    for (int i = 0; i < ell; ++i) {
        i128 r = large % (i128)p.mod[i];
        if (r < 0) r += p.mod[i];
        factors[i] = (u64)r;
    }
}

}  // namespace host
}  // namespace ppq
