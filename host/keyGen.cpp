// keyGen -- drop-in for the reference's client/src/keyGen.cpp (same argv, exit codes, progress lines, cereal-JSON key
// files), with cc->KeyGen() (:33) replaced by ppq_sample_keygen + ppq_keygen through the C ABI of libppq_b200.so.
// Sampler key: 256 bits from getrandom(2) inside the library per call; PPQ_SEED is a test-only override.  No CPU fallback.
#include <random>

#include "ppq_session.hpp"

using namespace ppqhost;

int main(int argc, char* argv[]) {
    if (argc != 4) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> <pubkey_out> <privkey_out>" << std::endl;
        return 1;
    }
    const std::string cc_path = argv[1], pubkey_out = argv[2], privkey_out = argv[3];
    try {
        // Step 1: CryptoContext (reference :26-31)
        std::unique_ptr<Session> s;
        try {
            s.reset(new Session(cc_path));
        } catch (const std::exception& e) {
            std::cerr << "[keyGen] ERROR: cannot load CryptoContext from " << cc_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[keyGen] CryptoContext loaded from " << cc_path << std::endl;

        // Step 2: key pair (reference :33-38), sampled and computed on the device
        const ppqhost::SamplerSeed seed;  // NULL (fresh OS entropy per call) unless the test-only PPQ_SEED is set
        const size_t N = s->N, L = (size_t)s->L, M = L + (size_t)s->K;
        Dev<int64_t> d_s(N), d_e(N);
        Dev<uint64_t> d_a(M * N), d_sk(L * N), d_pkb(M * N);
        ck(ppq_sample_keygen(s->cc, seed.ptr(), d_s.p, d_e.p, d_a.p, nullptr), "ppq_sample_keygen");
        ck(ppq_keygen(s->cc, d_s.p, d_e.p, d_a.p, d_sk.p, d_pkb.p, nullptr), "ppq_keygen");
        ck(ppq_sync(s->cc, nullptr), "ppq_sync");
        std::vector<uint64_t> sk(L * N), pkb(M * N), pka(M * N);
        d_sk.download(sk.data(), sk.size());
        d_pkb.download(pkb.data(), pkb.size());
        d_a.download(pka.data(), pka.size());
        std::cout << "[keyGen] Public and Private keys generated" << std::endl;

        // Step 3: key files (reference :41-48)
        const std::string tag = new_key_tag();  // both halves of the pair carry the same CryptoObject key tag
        try {
            wire::write_private_key(privkey_out, wire::poly_from_flat(sk.data(), L, N, s->moduli.data()), tag);
        } catch (const std::exception&) {
            std::cerr << "[keyGen] ERROR: Failed to save private key to " << privkey_out << std::endl;
            return 1;
        }
        try {
            wire::write_public_key(pubkey_out, wire::poly_from_flat(pkb.data(), M, N, s->moduli.data()),
                                   wire::poly_from_flat(pka.data(), M, N, s->moduli.data()), tag);
        } catch (const std::exception&) {
            std::cerr << "[keyGen] ERROR: Failed to save public key to " << pubkey_out << std::endl;
            return 1;
        }
        std::cout << "[keyGen] Keys saved: priv=" << privkey_out << " pub=" << pubkey_out << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "[keyGen] ERROR: " << e.what() << std::endl;
        return 1;
    }
}
