// genCC -- drop-in for the reference's server/src/genCC.cpp: reads config_cc.json (MultiplicativeDepth,
// ScalingModSize, BatchSize, PREMode; :36-66), derives the CKKS-RNS context GenCryptoContext would build (ring dimension
// from the HEStd_128_classic table, FLEXIBLEAUTOEXT prime chain, HYBRID key switching, all six features enabled; :68-76)
// and writes it as the cereal-JSON CryptoContext the other tools load (:89).  Host only: no device is touched.
//
// The reference fixes both paths at compile time (-DCONFIG_PATH / -DOUTPUT_PATH, Makefile:23-24, genCC.cpp:10-16); here
// they are optional argv with the same defaults.  Pin: with the reference's own config this program reproduces
// server/storage/CC.json BYTE FOR BYTE (tests/test_host_tools.py) -- primes, roots, composite modulus, field order,
// number formatting and pointer ids.
#include <cmath>
#include <iostream>

#include "../include/ppq_b200.h"
#include "ppq_wire.hpp"

#ifndef CONFIG_PATH
#define CONFIG_PATH "server/config/config_cc.json"
#endif
#ifndef OUTPUT_PATH
#define OUTPUT_PATH "server/storage/CC.json"
#endif

using jmin::Ptr;
using jmin::make;
using jmin::make_u64;
using jmin::set;

static Ptr obj() { return make(jmin::Value::Object); }
static Ptr num(const char* text) { return jmin::make_number_text(text); }

// smallest ring dimension whose HEStd_128_classic bound (uniform ternary secret) admits log2(QP) bits
// (OpenFHE core/lib/lattice/stdlatticeparms.cpp [OFHE-mem]; CC.json's N = 16384 at logQP = 280 agrees)
static int ring_log_for(int log_qp) {
    static const int tbl[][2] = {{10, 27}, {11, 54}, {12, 109}, {13, 218}, {14, 438}, {15, 881}, {16, 1747}, {17, 3523}};
    for (const auto& r : tbl)
        if (log_qp <= r[1]) return r[0];
    return -1;
}

static Ptr int_object(uint64_t v, bool with_version) {
    Ptr o = obj();
    if (with_version) set(o, "cereal_class_version", make_u64(1));
    set(o, "v", make_u64(v));
    return o;
}
static Ptr big_object(const std::vector<uint64_t>& limbs, uint32_t msb, bool with_version) {
    Ptr o = obj();
    if (with_version) set(o, "cereal_class_version", make_u64(1));
    Ptr a = make(jmin::Value::Array);
    for (uint64_t x : limbs) a->arr.push_back(make_u64(x));
    set(o, "v", a);
    set(o, "m", make_u64(msb));
    return o;
}

int main(int argc, char* argv[]) {
    const std::string configPath = argc > 1 ? argv[1] : CONFIG_PATH;
    const std::string outputPath = argc > 2 ? argv[2] : OUTPUT_PATH;
    if (argc > 3) {
        std::cerr << "Usage: " << argv[0] << " [config_cc.json [CC.json]]" << std::endl;
        return 1;
    }
    int depth = 1, scaling = 50, batch = 0;  // CCParams<CryptoContextCKKSRNS> defaults [OFHE-mem]
    const int first = 60, extra = 20, aux = 60;
    try {
        jmin::Ptr cfg;
        try {
            cfg = jmin::parse(wire::slurp(configPath));
        } catch (const std::exception&) {
            std::cerr << "Failed to open " << configPath << std::endl;
            return 1;
        }
        if (const jmin::Value* v = cfg->find("MultiplicativeDepth")) depth = (int)v->as_i64();
        if (const jmin::Value* v = cfg->find("ScalingModSize")) scaling = (int)v->as_i64();
        if (const jmin::Value* v = cfg->find("BatchSize")) batch = (int)v->as_i64();
        if (const jmin::Value* v = cfg->find("PREMode")) {
            const std::string mode = v->as_string();
            if (mode == "FIXED_NOISE") {  // the reference maps this to FIXEDMANUAL scaling (:59-60); not supported here
                std::cerr << "PREMode FIXED_NOISE selects FIXEDMANUAL scaling, which ppq_b200 does not implement" << std::endl;
                return 1;
            } else if (mode != "INDCPA") {
                std::cerr << "Unknown PREMode in config: " << mode << std::endl;
                return 1;
            }
        }
        if (depth < 1 || depth > 60 || scaling < 20 || scaling > 59) {
            std::cerr << "MultiplicativeDepth / ScalingModSize out of range" << std::endl;
            return 1;
        }
        // ---- parameters GenCryptoContext derives
        const int L = depth + 2;                                     // depth + 1 scaling levels + the extra limb
        const int dnum = depth > 3 ? 3 : 2;                          // ComputeNumLargeDigits(0, depth)
        const int q_bits = first + depth * scaling + extra;
        const int alpha_bits = (q_bits + dnum - 1) / dnum;
        const int log_qp = q_bits + aux * ((alpha_bits + aux - 1) / aux);
        const int logN = ring_log_for(log_qp);
        if (logN < 12) {
            std::cerr << "parameters need a ring dimension outside 2^12..2^17 (log QP = " << log_qp << ")" << std::endl;
            return 1;
        }
        const uint64_t N = (uint64_t)1 << logN;
        if (batch == 0) batch = (int)(N / 2);
        if (batch < 1 || (uint64_t)batch > N / 2) {
            std::cerr << "BatchSize must be in 1..N/2" << std::endl;
            return 1;
        }
        std::vector<uint64_t> q(64);
        int nq = 0;
        if (ppq_params_chain(logN, depth, first, scaling, extra, q.data(), (int)q.size(), &nq) != PPQ_OK || nq != L) {
            std::cerr << "prime chain: " << ppq_last_error() << std::endl;
            return 1;
        }
        q.resize(L);
        size_t need = 0;
        if (ppq_params_dump(logN, L, q.data(), dnum, aux, extra, L, nullptr, 0, &need) != PPQ_OK) {
            std::cerr << "parameter derivation: " << ppq_last_error() << std::endl;
            return 1;
        }
        std::string buf(need, '\0');
        ppq_params_dump(logN, L, q.data(), dnum, aux, extra, L, &buf[0], need, &need);
        jmin::Ptr derived = jmin::parse(std::string(buf.c_str()));
        std::vector<uint64_t> roots(L);
        for (int i = 0; i < L; ++i) roots[i] = derived->at("roots").at(i).as_u64();
        std::vector<uint64_t> big;
        uint32_t msb = 0;
        cerealpb::big_product(q.data(), (uint32_t)L, big, msb);

        // ---- the cereal-JSON tree (field order and pointer ids as OpenFHE writes them: server/storage/CC.json)
        const uint64_t MSB = 2147483648ull, SAME = 1073741824ull;
        uint64_t next_id = 1;
        auto new_id = [&]() { return make_u64(MSB | next_id++); };
        Ptr root = obj(), v0 = obj(), pw0 = obj(), d0 = obj();
        set(v0, "polymorphic_id", make_u64(SAME));
        set(pw0, "id", new_id());
        set(d0, "cereal_class_version", make_u64(0));
        // "cc": CryptoParametersCKKSRNS
        Ptr cc = obj(), ccpw = obj(), ccd = obj(), rns = obj(), rlwe = obj(), base = obj();
        set(cc, "polymorphic_id", make_u64(MSB | 1));
        set(cc, "polymorphic_name", jmin::make_string("lbcrypto::CryptoParametersCKKSRNS"));
        set(ccpw, "id", new_id());
        set(ccd, "cereal_class_version", make_u64(0));
        set(rns, "cereal_class_version", make_u64(0));
        set(rlwe, "cereal_class_version", make_u64(0));
        set(base, "cereal_class_version", make_u64(0));
        // elp: ILDCRTParams
        Ptr elp = obj(), elpw = obj(), eld = obj(), el0 = obj();
        set(elp, "polymorphic_id", make_u64(SAME));
        set(elpw, "id", new_id());
        set(eld, "cereal_class_version", make_u64(1));
        set(el0, "cereal_class_version", make_u64(1));
        set(el0, "co", make_u64(2 * N));
        set(el0, "rd", make_u64(N));
        set(el0, "cm", big_object(big, msb, true));
        set(el0, "ru", big_object({0}, 0, false));
        set(el0, "bm", big_object({0}, 0, false));
        set(el0, "br", big_object({0}, 0, false));
        set(eld, "value0", el0);
        Ptr plist = make(jmin::Value::Array);
        for (int i = 0; i < L; ++i) {
            Ptr np = obj(), npw = obj(), nd = obj(), n0 = obj();
            set(np, "polymorphic_id", make_u64(SAME));
            set(npw, "id", new_id());
            if (i == 0) {
                set(nd, "cereal_class_version", make_u64(1));
                set(n0, "cereal_class_version", make_u64(1));
            }
            set(n0, "co", make_u64(2 * N));
            set(n0, "rd", make_u64(N));
            set(n0, "cm", int_object(q[i], i == 0));
            set(n0, "ru", int_object(roots[i], false));
            set(n0, "bm", int_object(0, false));
            set(n0, "br", int_object(0, false));
            set(nd, "value0", n0);
            set(npw, "data", nd);
            set(np, "ptr_wrapper", npw);
            plist->arr.push_back(np);
        }
        set(eld, "p", plist);
        set(elpw, "data", eld);
        set(elp, "ptr_wrapper", elpw);
        set(base, "elp", elp);
        // enp: EncodingParams
        Ptr enp = obj(), enpw = obj(), end = obj();
        set(enp, "polymorphic_id", make_u64(SAME));
        set(enpw, "id", new_id());
        set(end, "cereal_class_version", make_u64(1));
        set(end, "m", make_u64((uint64_t)scaling));
        set(end, "ru", int_object(0, false));
        set(end, "bm", int_object(0, false));
        set(end, "bru", int_object(0, false));
        set(end, "g", make_u64(0));
        set(end, "bs", make_u64((uint64_t)batch));
        set(enpw, "data", end);
        set(enp, "ptr_wrapper", enpw);
        set(base, "enp", enp);
        set(rlwe, "value0", base);
        set(rlwe, "dp", num("3.190000057220459"));  // float(3.19) as cereal prints it
        set(rlwe, "am", num("36.0"));
        set(rlwe, "ns", make_u64(1));
        set(rlwe, "rw", make_u64(0));
        set(rlwe, "nest", num("0.0"));
        set(rlwe, "muld", make_u64(2));
        set(rlwe, "addc", make_u64(0));
        set(rlwe, "kswc", make_u64(0));
        set(rlwe, "phops", make_u64(0));
        set(rlwe, "md", make_u64((uint64_t)depth));
        set(rlwe, "mo", make_u64(1));    // UNIFORM_TERNARY
        set(rlwe, "pmo", make_u64(1));   // INDCPA
        set(rlwe, "mmo", make_u64(1));
        set(rlwe, "exm", make_u64(0));
        set(rlwe, "dnm", make_u64(0));
        set(rlwe, "slv", make_u64(0));   // HEStd_128_classic
        set(rlwe, "fdp", num("0.0"));
        set(rlwe, "ss", num("30.0"));
        set(rlwe, "aq", num("1.0"));
        set(rlwe, "tp", make_u64(1));
        set(rns, "value0", rlwe);
        set(rns, "ks", make_u64(2));     // HYBRID
        set(rns, "rs", make_u64(3));     // FLEXIBLEAUTOEXT
        set(rns, "encs", make_u64(0));
        set(rns, "muls", make_u64(1));
        set(rns, "dnum", make_u64((uint64_t)dnum));
        set(rns, "ab", make_u64((uint64_t)aux));
        set(rns, "eb", make_u64((uint64_t)extra));
        set(rns, "ccl", make_u64(3));
        set(rns, "cd", make_u64(1));
        set(rns, "rws", make_u64(64));
        set(rns, "cdt", make_u64(0));
        set(ccd, "value0", rns);
        set(ccpw, "data", ccd);
        set(cc, "ptr_wrapper", ccpw);
        set(d0, "cc", cc);
        // "kt": SchemeCKKSRNS with PKE | KEYSWITCH | PRE | LEVELEDSHE | ADVANCEDSHE | MULTIPARTY enabled (genCC.cpp:71-76)
        Ptr kt = obj(), ktpw = obj(), ktd = obj(), kt1 = obj(), kt2 = obj(), nullp = obj(), nullq = obj();
        set(kt, "polymorphic_id", make_u64(MSB | 2));
        set(kt, "polymorphic_name", jmin::make_string("lbcrypto::SchemeCKKSRNS"));
        set(ktpw, "id", new_id());
        set(ktd, "cereal_class_version", make_u64(0));
        set(kt1, "cereal_class_version", make_u64(0));
        set(kt2, "cereal_class_version", make_u64(0));
        set(nullp, "polymorphic_id", make_u64(0));
        set(nullq, "polymorphic_id", make_u64(0));
        set(kt2, "fhe", nullp);
        set(kt2, "schswitch", nullq);
        set(kt2, "enabled", make_u64(63));
        set(kt1, "value0", kt2);
        set(ktd, "value0", kt1);
        set(ktpw, "data", ktd);
        set(kt, "ptr_wrapper", ktpw);
        set(d0, "kt", kt);
        set(d0, "si", make_u64(1));      // CKKSRNS_SCHEME
        set(pw0, "data", d0);
        set(v0, "ptr_wrapper", pw0);
        set(root, "value0", v0);
        try {
            wire::spit(outputPath, jmin::dump(*root, 4, false));
        } catch (const std::exception&) {
            std::cerr << "Failed to serialize CryptoContext to CC.json" << std::endl;
            return 1;
        }
        std::cout << "CryptoContext Generated and saved to : " << outputPath << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "genCC: " << e.what() << std::endl;
        return 1;
    }
}
