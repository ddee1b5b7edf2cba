// ppq_wire.hpp -- file formats of the four PPQSFLHE tools, host side (C++17, header only).
//
//   * container:  {"weights_summary":[{"layer","shape","mean","std_dev","values":[...]}]}  -- the schema written
//     by client/src/encryptModelWeights.cpp:64-129 and rewritten by server/src/changeCipherDomain.cpp:57-124,
//     server/src/aggregateEncryptedWeights.cpp:64-120; read by client/src/decryptModelWeights.cpp:68-121.
//   * base64:     standard alphabet, padding, NO newlines (lib/base64_utils.h:10-49, OpenSSL BIO_FLAGS_BASE64_NO_NL).
//   * keys / CC:  cereal JSON archives as OpenFHE writes them (client/src/keyGen.cpp:41-48, REkeyGen.cpp:60,
//     server/src/genCC.cpp:89).  The DCRTPoly element layout is taken from a REAL file in the reference tree,
//     client/storage/client_1/private/client_1-private.key:  {"v":[{"v":{"ptr_wrapper":{"data":{"v":[residues],
//     "m":{"v":modulus}}}},"f":format,"p":...}, ...],"f":format,"p":...}.
//   * ciphertext blob: the reference stores base64(cereal PortableBinary(Ciphertext<DCRTPoly>)), which embeds the
//     whole crypto context per blob.  ppq_cereal.hpp writes and reads that layout (field order from
//     the two real cereal-JSON files in the reference tree, total size pinned by the recorded file sizes; byte identity
//     with a real OpenFHE blob UNVERIFIED -- none exists in the tree).  It is the default on write; PPQ_BLOB=ppqct1
//     selects the older self-describing private blob ("PPQCT1\0\0" magic, the CiphertextImpl scalars, then per
//     polynomial and limb: modulus, length, raw residues).  Codec::decode detects either.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#endif

#include "json_min.hpp"
#include "ppq_cereal.hpp"

namespace wire {

// ------------------------------------------------------------------------------------------------ files
// whole-file reads and writes with one read()/write() loop: a container is tens of megabytes and an iostream copy of it
// costs as much as decoding it
inline std::string slurp(const std::string& path) {
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    std::string out;
    if (std::fseek(f, 0, SEEK_END) == 0) {
        const long n = std::ftell(f);
        if (n > 0) out.resize((size_t)n);
        std::rewind(f);
    }
    size_t got = 0;
    if (!out.empty()) got = std::fread(&out[0], 1, out.size(), f);
    out.resize(got);
    char buf[65536];  // not seekable (a pipe) or still growing: take the rest
    for (size_t n; (n = std::fread(buf, 1, sizeof buf, f)) > 0;) out.append(buf, n);
    const bool bad = std::ferror(f) != 0;
    std::fclose(f);
    if (bad) throw std::runtime_error("read error on " + path);
    return out;
}
inline void spit(const std::string& path, const std::string& data) {
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    const size_t put = std::fwrite(data.data(), 1, data.size(), f);
    const bool bad = put != data.size() || std::fclose(f) != 0;
    if (bad) throw std::runtime_error("short write to " + path);
}

// ------------------------------------------------------------------------------------------------ base64
// standard alphabet, '=' padding, no line breaks on output (lib/base64_utils.h:10-49: BIO_FLAGS_BASE64_NO_NL); line
// breaks are tolerated on input.  Three bytes per step through a 4096-entry pair table / four characters per step
// through a 256-entry table: a blob is ~1.4 MB of text and a round decodes and encodes hundreds of them.
namespace b64detail {
struct Tables {
    char pair[4096][2];  // 12 bits -> two characters
    int8_t dec[256];     // character -> 6 bits, -1 invalid
    Tables() {
        static const char T[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
        for (int i = 0; i < 4096; ++i) {
            pair[i][0] = T[i >> 6];
            pair[i][1] = T[i & 63];
        }
        std::memset(dec, -1, sizeof dec);
        for (int i = 0; i < 64; ++i) dec[(unsigned char)T[i]] = (int8_t)i;
    }
};
inline const Tables& tables() {
    static const Tables t;  // initialised once, thread-safe (blobs are coded from OpenMP loops)
    return t;
}
#if defined(__x86_64__) && defined(__GNUC__)
#define PPQ_B64_AVX2 1
// 32 characters -> 24 bytes / 24 bytes -> 32 characters per step with byte shuffles (the published vector algorithm of
// Mula & Lemire, "Faster Base64 Encoding and Decoding Using AVX2 Instructions"): classify each character by its two
// nibbles (invalid ones stop the loop and are left to the scalar code), add the per-range offset, then pack 4 x 6 bits
// into 3 bytes with two multiply-adds.  Returns the number of characters / bytes consumed.
__attribute__((target("avx2"))) inline size_t decode_avx2(const unsigned char* p, size_t n, unsigned char* o) {
    const __m256i lut_lo = _mm256_setr_epi8(0x15, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x13, 0x1A, 0x1B, 0x1B,
                                            0x1B, 0x1A, 0x15, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x11, 0x13, 0x1A,
                                            0x1B, 0x1B, 0x1B, 0x1A);
    const __m256i lut_hi = _mm256_setr_epi8(0x10, 0x10, 0x01, 0x02, 0x04, 0x08, 0x04, 0x08, 0x10, 0x10, 0x10, 0x10, 0x10, 0x10,
                                            0x10, 0x10, 0x10, 0x10, 0x01, 0x02, 0x04, 0x08, 0x04, 0x08, 0x10, 0x10, 0x10, 0x10,
                                            0x10, 0x10, 0x10, 0x10);
    const __m256i lut_roll = _mm256_setr_epi8(0, 16, 19, 4, -65, -65, -71, -71, 0, 0, 0, 0, 0, 0, 0, 0, 0, 16, 19, 4, -65, -65,
                                              -71, -71, 0, 0, 0, 0, 0, 0, 0, 0);
    const __m256i mask_2f = _mm256_set1_epi8(0x2f);
    const __m256i pack = _mm256_setr_epi8(2, 1, 0, 6, 5, 4, 10, 9, 8, 14, 13, 12, -1, -1, -1, -1, 2, 1, 0, 6, 5, 4, 10, 9, 8, 14,
                                          13, 12, -1, -1, -1, -1);
    const __m256i lanes = _mm256_setr_epi32(0, 1, 2, 4, 5, 6, -1, -1);
    size_t i = 0;
    // the 32-byte store writes 8 bytes past the 24 it produces: stop while 45 characters (>= 33 output bytes) remain
    for (; i + 45 <= n; i += 32, o += 24) {
        __m256i str = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(p + i));
        const __m256i hi_nib = _mm256_and_si256(_mm256_srli_epi32(str, 4), mask_2f);
        const __m256i lo_nib = _mm256_and_si256(str, mask_2f);
        const __m256i hi = _mm256_shuffle_epi8(lut_hi, hi_nib);
        const __m256i lo = _mm256_shuffle_epi8(lut_lo, lo_nib);
        if (!_mm256_testz_si256(lo, hi)) break;
        const __m256i eq_2f = _mm256_cmpeq_epi8(str, mask_2f);
        const __m256i roll = _mm256_shuffle_epi8(lut_roll, _mm256_add_epi8(eq_2f, hi_nib));
        str = _mm256_add_epi8(str, roll);
        const __m256i ab_bc = _mm256_maddubs_epi16(str, _mm256_set1_epi32(0x01400140));
        __m256i out = _mm256_madd_epi16(ab_bc, _mm256_set1_epi32(0x00011000));
        out = _mm256_shuffle_epi8(out, pack);
        out = _mm256_permutevar8x32_epi32(out, lanes);
        _mm256_storeu_si256(reinterpret_cast<__m256i*>(o), out);
    }
    return i;
}
__attribute__((target("avx2"))) inline size_t encode_avx2(const unsigned char* p, size_t n, char* o) {
    const __m256i spread = _mm256_setr_epi8(1, 0, 2, 1, 4, 3, 5, 4, 7, 6, 8, 7, 10, 9, 11, 10, 1, 0, 2, 1, 4, 3, 5, 4, 7, 6, 8, 7,
                                            10, 9, 11, 10);
    const __m256i lut = _mm256_setr_epi8(65, 71, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -19, -16, 0, 0, 65, 71, -4, -4, -4, -4,
                                         -4, -4, -4, -4, -4, -4, -19, -16, 0, 0);
    size_t i = 0;
    for (; i + 28 <= n; i += 24, o += 32) {  // the second 16-byte load reads bytes 12..27
        const __m128i lo = _mm_loadu_si128(reinterpret_cast<const __m128i*>(p + i));
        const __m128i hi = _mm_loadu_si128(reinterpret_cast<const __m128i*>(p + i + 12));
        __m256i in = _mm256_shuffle_epi8(_mm256_set_m128i(hi, lo), spread);
        const __m256i t0 = _mm256_and_si256(in, _mm256_set1_epi32(0x0fc0fc00));
        const __m256i t1 = _mm256_mulhi_epu16(t0, _mm256_set1_epi32(0x04000040));
        const __m256i t2 = _mm256_and_si256(in, _mm256_set1_epi32(0x003f03f0));
        const __m256i t3 = _mm256_mullo_epi16(t2, _mm256_set1_epi32(0x01000010));
        in = _mm256_or_si256(t1, t3);  // one 6-bit value per byte
        __m256i idx = _mm256_subs_epu8(in, _mm256_set1_epi8(51));
        idx = _mm256_sub_epi8(idx, _mm256_cmpgt_epi8(in, _mm256_set1_epi8(25)));
        _mm256_storeu_si256(reinterpret_cast<__m256i*>(o), _mm256_add_epi8(in, _mm256_shuffle_epi8(lut, idx)));
    }
    return i;
}
inline bool have_avx2() {
    static const bool yes = __builtin_cpu_supports("avx2");
    return yes;
}
#endif
}  // namespace b64detail
// `out` is resized (its capacity reused) and overwritten
inline void base64_encode_to(std::string& out, const unsigned char* p, size_t n) {
    const b64detail::Tables& t = b64detail::tables();
    out.resize((n + 2) / 3 * 4);
    char* o = &out[0];
    size_t i = 0;
#ifdef PPQ_B64_AVX2
    if (b64detail::have_avx2()) {
        i = b64detail::encode_avx2(p, n, o);
        o += i / 3 * 4;
    }
#endif
    for (; i + 2 < n; i += 3, o += 4) {
        const uint32_t v = ((uint32_t)p[i] << 16) | ((uint32_t)p[i + 1] << 8) | p[i + 2];
        std::memcpy(o, t.pair[v >> 12], 2);
        std::memcpy(o + 2, t.pair[v & 4095], 2);
    }
    if (i < n) {
        uint32_t v = (uint32_t)p[i] << 16;
        if (i + 1 < n) v |= (uint32_t)p[i + 1] << 8;
        std::memcpy(o, t.pair[v >> 12], 2);
        o[2] = (i + 1 < n) ? t.pair[v & 4095][0] : '=';
        o[3] = '=';
    }
}
inline std::string base64_encode(const unsigned char* p, size_t n) {
    std::string out;
    base64_encode_to(out, p, n);
    return out;
}
inline void base64_decode_to(std::vector<unsigned char>& out, const std::string& s) {
    const b64detail::Tables& t = b64detail::tables();
    out.resize(s.size() / 4 * 3 + 3);
    unsigned char* o = out.data();
    const unsigned char* p = reinterpret_cast<const unsigned char*>(s.data());
    const unsigned char* e = p + s.size();
    uint32_t acc = 0;
    int bits = 0;
    while (p < e) {
#ifdef PPQ_B64_AVX2
        if (bits == 0 && b64detail::have_avx2()) {
            const size_t used = b64detail::decode_avx2(p, (size_t)(e - p), o);
            p += used;
            o += used / 4 * 3;
        }
#endif
        if (bits == 0) {  // aligned: whole quads of valid characters, three bytes per step
            while (e - p >= 4) {
                const int a = t.dec[p[0]], b = t.dec[p[1]], c = t.dec[p[2]], d = t.dec[p[3]];
                if ((a | b | c | d) < 0) break;  // padding, a line break or an invalid character: one careful step
                const uint32_t v = ((uint32_t)a << 18) | ((uint32_t)b << 12) | ((uint32_t)c << 6) | (uint32_t)d;
                o[0] = (unsigned char)(v >> 16);
                o[1] = (unsigned char)(v >> 8);
                o[2] = (unsigned char)v;
                o += 3;
                p += 4;
            }
            if (p >= e) break;
        }
        const unsigned char ch = *p++;
        if (ch == '=') break;
        const int8_t d = t.dec[ch];
        if (d < 0) {
            if (ch == '\n' || ch == '\r') continue;
            throw std::runtime_error("base64: invalid character");
        }
        acc = (acc << 6) | (uint32_t)d;
        bits += 6;
        if (bits >= 8) {
            bits -= 8;
            *o++ = (unsigned char)(acc >> bits);
        }
    }
    out.resize((size_t)(o - out.data()));
}
inline std::vector<unsigned char> base64_decode(const std::string& s) {
    std::vector<unsigned char> out;
    base64_decode_to(out, s);
    return out;
}

// ------------------------------------------------------------------------------------------------ ciphertext blob
struct CtMeta {  // the scalar members of lbcrypto::CiphertextImpl ("d","l","h","e","s","si","sl")
    uint32_t noise_scale_deg = 2, level = 0, hops_level = 0, encoding = 3 /* CKKS_PACKED_ENCODING */, slots = 0;
    double scaling_factor = 1.0;
    uint64_t scaling_factor_int = 1;
    std::string key_tag;  // CryptoObject::keyTag (cereal blobs only; the private blob does not carry it)
};
struct Ct {
    CtMeta meta;
    uint32_t ell = 0, N = 0;
    std::vector<uint64_t> moduli;  // [ell]
    std::vector<uint64_t> data;    // [2][ell][N] EVAL, bit-reversed order
};
static const char CT_MAGIC[8] = {'P', 'P', 'Q', 'C', 'T', '1', 0, 0};

template <class T>
inline void put(std::string& o, T v) { o.append(reinterpret_cast<const char*>(&v), sizeof v); }
template <class T>
inline T get(const unsigned char*& p, const unsigned char* e) {
    if ((size_t)(e - p) < sizeof(T)) throw std::runtime_error("ciphertext blob truncated");
    T v;
    std::memcpy(&v, p, sizeof v);
    p += sizeof v;
    return v;
}

// data: [2][ell][N]
inline std::string encode_ct(const CtMeta& m, uint32_t ell, uint32_t N, const uint64_t* moduli, const uint64_t* data) {
    std::string o;
    o.reserve(96 + (size_t)2 * ell * (16 + (size_t)N * 8));
    o.append(CT_MAGIC, 8);
    put<uint32_t>(o, m.noise_scale_deg);
    put<uint32_t>(o, m.level);
    put<uint32_t>(o, m.hops_level);
    put<uint32_t>(o, m.encoding);
    put<double>(o, m.scaling_factor);
    put<uint64_t>(o, m.scaling_factor_int);
    put<uint32_t>(o, m.slots);
    put<uint32_t>(o, 0u);               // format: 0 = EVALUATION
    put<uint64_t>(o, 2);                // number of polynomials
    for (int poly = 0; poly < 2; ++poly) {
        put<uint64_t>(o, ell);
        for (uint32_t i = 0; i < ell; ++i) {
            put<uint64_t>(o, moduli[i]);
            put<uint64_t>(o, N);
            o.append(reinterpret_cast<const char*>(data + ((size_t)poly * ell + i) * N), (size_t)N * 8);
        }
    }
    return base64_encode(reinterpret_cast<const unsigned char*>(o.data()), o.size());
}
inline Ct decode_ct(const std::string& b64) {
    const std::vector<unsigned char> raw = base64_decode(b64);
    const unsigned char* p = raw.data();
    const unsigned char* e = p + raw.size();
    if (raw.size() < 8 || std::memcmp(p, CT_MAGIC, 8) != 0)
        throw std::runtime_error(
            "ciphertext blob is not in PPQCT1 format (an OpenFHE cereal-binary blob cannot be read: its layout is "
            "unverifiable without OpenFHE, see host/ppq_wire.hpp)");
    p += 8;
    Ct c;
    c.meta.noise_scale_deg = get<uint32_t>(p, e);
    c.meta.level = get<uint32_t>(p, e);
    c.meta.hops_level = get<uint32_t>(p, e);
    c.meta.encoding = get<uint32_t>(p, e);
    c.meta.scaling_factor = get<double>(p, e);
    c.meta.scaling_factor_int = get<uint64_t>(p, e);
    c.meta.slots = get<uint32_t>(p, e);
    if (get<uint32_t>(p, e) != 0) throw std::runtime_error("ciphertext blob: not in EVALUATION format");
    if (get<uint64_t>(p, e) != 2) throw std::runtime_error("ciphertext blob: expected 2 polynomials");
    for (int poly = 0; poly < 2; ++poly) {
        const uint64_t ell = get<uint64_t>(p, e);
        if (ell == 0 || ell > 4096) throw std::runtime_error("ciphertext blob: bad limb count");
        if (poly == 0) {
            c.ell = (uint32_t)ell;
            c.moduli.resize(ell);
        } else if (ell != c.ell) {
            throw std::runtime_error("ciphertext blob: polynomials differ in limb count");
        }
        for (uint32_t i = 0; i < c.ell; ++i) {
            const uint64_t q = get<uint64_t>(p, e), n = get<uint64_t>(p, e);
            if (poly == 0 && i == 0) {
                if (n == 0 || (n & (n - 1)) || n > (1u << 20)) throw std::runtime_error("ciphertext blob: bad ring dimension");
                c.N = (uint32_t)n;
                c.data.resize((size_t)2 * c.ell * c.N);
            }
            if (n != c.N) throw std::runtime_error("ciphertext blob: limbs differ in length");
            if (poly == 0) c.moduli[i] = q;
            else if (c.moduli[i] != q) throw std::runtime_error("ciphertext blob: polynomials differ in moduli");
            if ((size_t)(e - p) < (size_t)n * 8) throw std::runtime_error("ciphertext blob truncated");
            std::memcpy(c.data.data() + ((size_t)poly * c.ell + i) * c.N, p, (size_t)n * 8);
            p += (size_t)n * 8;
        }
    }
    if (p != e) throw std::runtime_error("ciphertext blob: trailing bytes");
    return c;
}

// ------------------------------------------------------------------------------------------------ blob codec
// One per session: knows the crypto context the blobs embed.  encode: cereal PortableBinary unless PPQ_BLOB=ppqct1;
// decode: whichever of the two the blob is.
struct Codec {
    cerealpb::Context cx;
    bool write_cereal = true;
    mutable bool warned = false;
    void init(const std::string& ccjson_text, uint32_t N, const uint64_t* moduli, const uint64_t* roots, size_t L) {
        cx.init(jmin::parse(ccjson_text), N, moduli, roots, L);
        const char* e = std::getenv("PPQ_BLOB");
        write_cereal = !(e && std::string(e) == "ppqct1");
    }
    void warn_once() const {
        bool first;
#pragma omp critical(ppq_codec_warn)
        {
            first = !warned;
            warned = true;
        }
        if (first && !std::getenv("PPQ_QUIET"))
            std::fprintf(stderr, "[ppq] note: cereal PortableBinary ciphertext blobs are written/read to a layout derived from the "
                                 "reference's own cereal-JSON files and file sizes; byte identity is unverified against OpenFHE\n");
    }
    std::string encode(const CtMeta& m, uint32_t ell, const uint64_t* data) const {
        std::string out;
        encode_to(out, m, ell, data);
        return out;
    }
    // the blob replaces the contents of `out`, reusing its capacity (a re-encoded blob is never longer than the one it
    // replaces) and a per-thread scratch archive: no megabyte allocation per blob inside the callers' OpenMP loops
    void encode_to(std::string& out, const CtMeta& m, uint32_t ell, const uint64_t* data) const {
        if (!write_cereal) {
            out = encode_ct(m, ell, cx.N, cx.moduli.data(), data);
            return;
        }
        warn_once();
        cerealpb::Meta cm;
        cm.noise_scale_deg = m.noise_scale_deg;
        cm.level = m.level;
        cm.hops_level = m.hops_level;
        cm.encoding = m.encoding;
        cm.slots = m.slots;
        cm.scaling_factor = m.scaling_factor;
        cm.scaling_factor_int = m.scaling_factor_int;
        cm.key_tag = m.key_tag.empty() ? std::string(32, '0') : m.key_tag;
        static thread_local std::string raw;
        cerealpb::encode_to(raw, cx, cm, ell, data);
        base64_encode_to(out, reinterpret_cast<const unsigned char*>(raw.data()), raw.size());
    }
    // dst != nullptr: residues straight into the caller's batch ([2][ell][N], at most dst_cap words), Ct::data left empty
    Ct decode(const std::string& b64, uint64_t* dst = nullptr, size_t dst_cap = 0) const {
        if (b64.compare(0, 8, "UFBRQ1Qx") == 0) {  // base64 of "PPQCT1"
            Ct o = decode_ct(b64);
            if (dst) {
                if (o.data.size() > dst_cap) throw std::runtime_error("ciphertext blob: more limbs than the batch it belongs to");
                std::memcpy(dst, o.data.data(), o.data.size() * sizeof(uint64_t));
                o.data.clear();
            }
            return o;
        }
        static thread_local std::vector<unsigned char> raw;
        base64_decode_to(raw, b64);
        if (!cerealpb::looks_like_cereal(raw.data(), raw.size()))
            throw std::runtime_error("ciphertext blob is neither a cereal PortableBinary Ciphertext archive nor PPQCT1");
        warn_once();
        cerealpb::Ct c = cerealpb::decode(cx, raw.data(), raw.size(), dst, dst_cap);
        Ct o;
        o.meta.noise_scale_deg = c.meta.noise_scale_deg;
        o.meta.level = c.meta.level;
        o.meta.hops_level = c.meta.hops_level;
        o.meta.encoding = c.meta.encoding;
        o.meta.slots = c.meta.slots;
        o.meta.scaling_factor = c.meta.scaling_factor;
        o.meta.scaling_factor_int = c.meta.scaling_factor_int;
        o.meta.key_tag = c.meta.key_tag;
        o.ell = c.ell;
        o.N = c.N;
        o.moduli = std::move(c.moduli);
        o.data = std::move(c.data);
        return o;
    }
};

// ------------------------------------------------------------------------------------------------ cereal JSON keys
struct Poly {  // one DCRTPoly: limbs[i] has N residues mod moduli[i]
    std::vector<uint64_t> moduli;
    std::vector<std::vector<uint64_t>> limbs;
    int format = 0;
};
inline bool looks_like_dcrtpoly(const jmin::Value& v) {
    if (!v.is_object()) return false;
    const jmin::Value* a = v.find("v");
    if (!a || !a->is_array() || a->size() == 0) return false;
    const jmin::Value& first = a->at(0);
    return first.is_object() && first.find("v") && first.at("v").is_object() && first.at("v").find("ptr_wrapper");
}
inline Poly read_dcrtpoly(const jmin::Value& v) {
    if (!looks_like_dcrtpoly(v)) throw std::runtime_error("key file: element is not a serialized DCRTPoly");
    Poly p;
    if (const jmin::Value* f = v.find("f")) p.format = (int)f->as_i64();
    for (const auto& np : v.at("v").arr) {
        const jmin::Value& d = np->at("v").at("ptr_wrapper").at("data");
        const jmin::Value& vals = d.at("v");
        std::vector<uint64_t> limb(vals.size());
        vals.numbers_to(limb.data());
        p.moduli.push_back(d.at("m").at("v").as_u64());
        p.limbs.push_back(std::move(limb));
    }
    return p;
}
// depth-first search for the first member called `key` anywhere below v
inline const jmin::Value* find_member(const jmin::Value& v, const std::string& key) {
    if (v.is_object()) {
        for (const auto& kv : v.obj) {
            if (kv.first == key) return kv.second.get();
        }
        for (const auto& kv : v.obj) {
            if (kv.first == "cc") continue;  // the embedded crypto context never holds key material
            if (const jmin::Value* r = find_member(*kv.second, key)) return r;
        }
    } else if (v.is_array()) {
        for (const auto& x : v.arr)
            if (const jmin::Value* r = find_member(*x, key)) return r;
    }
    return nullptr;
}
// the same without skipping the embedded context (for CC.json itself)
inline const jmin::Value* find_member_any(const jmin::Value& v, const std::string& key) {
    if (v.is_object()) {
        for (const auto& kv : v.obj)
            if (kv.first == key) return kv.second.get();
        for (const auto& kv : v.obj)
            if (const jmin::Value* r = find_member_any(*kv.second, key)) return r;
    } else if (v.is_array()) {
        for (const auto& x : v.arr)
            if (const jmin::Value* r = find_member_any(*x, key)) return r;
    }
    return nullptr;
}
// PrivateKeyImpl: member "s" (client/src/keyGen.cpp:46 writes it; decryptModelWeights.cpp:49-54 reads it)
inline Poly read_private_key(const std::string& text) {
    jmin::Ptr root = jmin::parse_lazy(text);
    const jmin::Value* s = find_member(*root, "s");
    if (!s) throw std::runtime_error("private key file: no member \"s\"");
    return read_dcrtpoly(*s);
}
// PublicKeyImpl: member "h" = [b, a] (keyGen.cpp:41; encryptModelWeights.cpp:44-48)
inline std::vector<Poly> read_public_key(const std::string& text) {
    jmin::Ptr root = jmin::parse_lazy(text);
    const jmin::Value* h = find_member(*root, "h");
    if (!h || !h->is_array() || h->size() != 2) throw std::runtime_error("public key file: no member \"h\" with two elements");
    return {read_dcrtpoly(h->at(0)), read_dcrtpoly(h->at(1))};
}
// EvalKeyRelinImpl: member "k" = [A-vector, B-vector], each a list of dnum DCRTPolys (REkeyGen.cpp:60;
// changeCipherDomain.cpp:40-44)
struct EvalKeyPolys {
    std::vector<Poly> a, b;
};
inline EvalKeyPolys read_eval_key(const std::string& text) {
    jmin::Ptr root = jmin::parse_lazy(text);
    const jmin::Value* k = find_member(*root, "k");
    if (!k || !k->is_array() || k->size() != 2) throw std::runtime_error("re-encryption key file: no member \"k\" with two vectors");
    EvalKeyPolys out;
    for (const auto& x : k->at(0).arr) out.a.push_back(read_dcrtpoly(*x));
    for (const auto& x : k->at(1).arr) out.b.push_back(read_dcrtpoly(*x));
    if (out.a.empty() || out.a.size() != out.b.size()) throw std::runtime_error("re-encryption key file: A and B vectors differ in length");
    return out;
}

// writers of the same element layout (used by the test scaffolding that fabricates key files)
inline jmin::Ptr dcrtpoly_to_json(const Poly& p) {
    jmin::Ptr o = jmin::make(jmin::Value::Object);
    jmin::Ptr arr = jmin::make(jmin::Value::Array);
    for (size_t i = 0; i < p.limbs.size(); ++i) {
        jmin::Ptr vals = jmin::make(jmin::Value::Array);
        for (uint64_t x : p.limbs[i]) vals->arr.push_back(jmin::make_u64(x));
        jmin::Ptr m = jmin::make(jmin::Value::Object);
        jmin::set(m, "v", jmin::make_u64(p.moduli[i]));
        jmin::Ptr data = jmin::make(jmin::Value::Object);
        jmin::set(data, "cereal_class_version", jmin::make_u64(1));
        jmin::set(data, "v", vals);
        jmin::set(data, "m", m);
        jmin::Ptr pw = jmin::make(jmin::Value::Object);
        jmin::set(pw, "valid", jmin::make_u64(1));
        jmin::set(pw, "data", data);
        jmin::Ptr vv = jmin::make(jmin::Value::Object);
        jmin::set(vv, "polymorphic_id", jmin::make_u64(1073741824));
        jmin::set(vv, "ptr_wrapper", pw);
        jmin::Ptr np = jmin::make(jmin::Value::Object);
        jmin::set(np, "cereal_class_version", jmin::make_u64(1));
        jmin::set(np, "v", vv);
        jmin::set(np, "f", jmin::make_u64((uint64_t)p.format));
        arr->arr.push_back(np);
    }
    jmin::set(o, "cereal_class_version", jmin::make_u64(1));
    jmin::set(o, "v", arr);
    jmin::set(o, "f", jmin::make_u64((uint64_t)p.format));
    return o;
}

// CryptoObject::keyTag of a key file ("kt" as a STRING; the "kt" inside the embedded context is the scheme object)
inline std::string read_key_tag(const std::string& text) {
    jmin::Ptr root = jmin::parse_lazy(text);
    std::string tag;
    std::vector<const jmin::Value*> stack{root.get()};
    while (!stack.empty() && tag.empty()) {
        const jmin::Value* v = stack.back();
        stack.pop_back();
        if (v->is_object()) {
            for (const auto& kv : v->obj) {
                if (kv.first == "kt" && kv.second->is_string()) tag = kv.second->s;
                else if (kv.first != "cc" && !(kv.second->is_array() && kv.second->size() > 64)) stack.push_back(kv.second.get());
            }
        } else if (v->is_array()) {
            for (const auto& x : v->arr) stack.push_back(x.get());
        }
    }
    return tag;
}
// the cereal shared_ptr envelope around a key object: value0 -> ptr_wrapper -> data -> {kt, member}
inline jmin::Ptr wrap_key_member(const char* member, jmin::Ptr payload, const std::string& key_tag = "") {
    jmin::Ptr data = jmin::make(jmin::Value::Object);
    jmin::set(data, "cereal_class_version", jmin::make_u64(0));
    if (!key_tag.empty()) jmin::set(data, "kt", jmin::make_string(key_tag));
    jmin::set(data, member, payload);
    jmin::Ptr pw = jmin::make(jmin::Value::Object);
    jmin::set(pw, "id", jmin::make_u64(2147483649ull));
    jmin::set(pw, "data", data);
    jmin::Ptr v0 = jmin::make(jmin::Value::Object);
    jmin::set(v0, "polymorphic_id", jmin::make_u64(1073741824));
    jmin::set(v0, "ptr_wrapper", pw);
    jmin::Ptr root = jmin::make(jmin::Value::Object);
    jmin::set(root, "value0", v0);
    return root;
}
// [limbs][N] contiguous residues -> Poly (EVALUATION format)
inline Poly poly_from_flat(const uint64_t* data, size_t limbs, size_t N, const uint64_t* moduli) {
    Poly p;
    for (size_t i = 0; i < limbs; ++i) {
        p.limbs.emplace_back(data + i * N, data + (i + 1) * N);
        p.moduli.push_back(moduli[i]);
    }
    return p;
}
// key files as client/src/keyGen.cpp:41-48 and client/src/REkeyGen.cpp:60 write them (SerType::JSON)
inline void write_private_key(const std::string& path, const Poly& s, const std::string& key_tag = "") {
    spit(path, jmin::dump(*wrap_key_member("s", dcrtpoly_to_json(s), key_tag), 4, false));
}
inline void write_public_key(const std::string& path, const Poly& b, const Poly& a, const std::string& key_tag = "") {
    jmin::Ptr h = jmin::make(jmin::Value::Array);
    h->arr.push_back(dcrtpoly_to_json(b));
    h->arr.push_back(dcrtpoly_to_json(a));
    spit(path, jmin::dump(*wrap_key_member("h", h, key_tag), 4, false));
}
inline void write_eval_key(const std::string& path, const std::vector<Poly>& a, const std::vector<Poly>& b,
                           const std::string& key_tag = "") {
    jmin::Ptr k = jmin::make(jmin::Value::Array), av = jmin::make(jmin::Value::Array), bv = jmin::make(jmin::Value::Array);
    for (const Poly& p : a) av->arr.push_back(dcrtpoly_to_json(p));
    for (const Poly& p : b) bv->arr.push_back(dcrtpoly_to_json(p));
    k->arr.push_back(av);
    k->arr.push_back(bv);
    spit(path, jmin::dump(*wrap_key_member("k", k, key_tag), 4, false));
}

// ------------------------------------------------------------------------------------------------ container
struct Layer {
    jmin::Ptr layer, shape;          // copied through verbatim
    std::string mean, std_dev;       // base64 blobs (encrypted files)
    std::vector<std::string> values;
};
inline std::vector<Layer> read_container(const std::string& path) {
    jmin::Ptr root = jmin::parse(slurp(path));
    const jmin::Value* ws = root->find("weights_summary");
    if (!ws || !ws->is_array()) throw std::runtime_error(path + ": no \"weights_summary\" array");
    std::vector<Layer> out;
    for (const auto& e : ws->arr) {
        Layer l;
        for (const auto& kv : e->obj) {
            if (kv.first == "layer") l.layer = kv.second;
            else if (kv.first == "shape") l.shape = kv.second;
        }
        if (!l.layer || !l.shape) throw std::runtime_error(path + ": entry without layer/shape");
        // the blobs are megabytes each: move them out of the parsed tree (it is dropped right after) instead of copying
        auto take = [&](const jmin::Value& v) {
            v.as_string();  // type check
            return std::move(const_cast<jmin::Value&>(v).s);
        };
        l.mean = take(e->at("mean"));
        l.std_dev = take(e->at("std_dev"));
        for (const auto& v : e->at("values").arr) l.values.push_back(take(*v));
        out.push_back(std::move(l));
    }
    return out;
}
inline void write_container(const std::string& path, const std::vector<Layer>& layers) {
    // same text as dump(root, 2) of {"weights_summary": [{layer, mean, shape, std_dev, values}]} with sorted keys (the
    // reference prints with `std::setw(2) << nlohmann::json`), written straight into one buffer: the blobs are not copied
    // into a tree first
    size_t need = 64;
    for (const Layer& l : layers) {
        need += 256 + l.mean.size() + l.std_dev.size();
        for (const std::string& v : l.values) need += v.size() + 16;
    }
    std::string out;
    out.reserve(need + need / 64);
    out += "{\n  \"weights_summary\": [";
    if (layers.empty()) out += "]";
    for (size_t k = 0; k < layers.size(); ++k) {
        const Layer& l = layers[k];
        out += k ? ",\n    {\n" : "\n    {\n";
        out += "      \"layer\": ";
        jmin::dump_to(out, *l.layer, 2, 6, true);
        out += ",\n      \"mean\": ";
        jmin::escape_to(out, l.mean);
        out += ",\n      \"shape\": ";
        jmin::dump_to(out, *l.shape, 2, 6, true);
        out += ",\n      \"std_dev\": ";
        jmin::escape_to(out, l.std_dev);
        out += ",\n      \"values\": ";
        if (l.values.empty()) {
            out += "[]";
        } else {
            out += "[\n";
            for (size_t i = 0; i < l.values.size(); ++i) {
                out += "        ";
                jmin::escape_to(out, l.values[i]);
                out += i + 1 < l.values.size() ? ",\n" : "\n";
            }
            out += "      ]";
        }
        out += "\n    }";
        if (k + 1 == layers.size()) out += "\n  ]";
    }
    out += "\n}\n";
    spit(path, out);
}
inline bool same_json(const jmin::Value& a, const jmin::Value& b) { return jmin::dump(a, 0) == jmin::dump(b, 0); }

}  // namespace wire
