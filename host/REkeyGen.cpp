// REkeyGen -- drop-in for the reference's client/src/REkeyGen.cpp (same argv, exit codes, progress lines, cereal-JSON
// key files), with cc->ReKeyGen(privKey, pubKey) (:52) replaced by ppq_sample_encrypt_noise + ppq_rekeygen through the
// C ABI of libppq_b200.so.  Sampler key: 256 bits from getrandom(2) inside the library per call; PPQ_SEED is a test-only override.  No CPU fallback.
#include <random>

#include "ppq_session.hpp"

using namespace ppqhost;

int main(int argc, char* argv[]) {
    if (argc != 5) {
        std::cerr << "Usage: " << argv[0] << " <cc.json> <client_privkey.json> <peer_pubkey.json> <rekey_out.json>" << std::endl;
        return 1;
    }
    const std::string cc_path = argv[1], client_sk = argv[2], client_pk = argv[3], rekey_path = argv[4];
    try {
        std::unique_ptr<Session> s;
        try {
            s.reset(new Session(cc_path));
        } catch (const std::exception& e) {
            std::cerr << "Error loading CryptoContext from " << cc_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[ReKeyGen] CryptoContext loaded from " << cc_path << std::endl;
        const size_t N = s->N, L = (size_t)s->L, M = L + (size_t)s->K, dnum = (size_t)s->dnum;

        std::vector<uint64_t> sk;
        try {
            sk = s->flatten(wire::read_private_key(wire::slurp(client_sk)), L, "private key");
        } catch (const std::exception& e) {
            std::cerr << "Error loading Client private key from " << client_sk << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[ReKeyGen] Client Private Key loaded from " << client_sk << std::endl;

        std::vector<uint64_t> pkb, pka;
        std::string peer_tag;  // the re-encryption key carries the tag of the key it switches TO
        try {
            const std::string pk_text = wire::slurp(client_pk);
            peer_tag = wire::read_key_tag(pk_text);
            const std::vector<wire::Poly> h = wire::read_public_key(pk_text);
            pkb = s->flatten(h[0], M, "public key b");
            pka = s->flatten(h[1], M, "public key a");
        } catch (const std::exception& e) {
            std::cerr << "Error loading Peer public key from " << client_pk << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[ReKeyGen] Peer Public Key loaded from " << client_pk << std::endl;

        // reference :52 -- dnum digits of (u, e0, e1) drawn on the device, then the key itself
        const ppqhost::SamplerSeed seed;  // NULL (fresh OS entropy per call) unless the test-only PPQ_SEED is set
        Dev<uint64_t> d_sk(L * N), d_pkb(M * N), d_pka(M * N), d_kb(dnum * M * N), d_ka(dnum * M * N);
        Dev<int64_t> d_small(dnum * 3 * N);
        d_sk.upload(sk.data(), sk.size());
        d_pkb.upload(pkb.data(), pkb.size());
        d_pka.upload(pka.data(), pka.size());
        try {
            ck(ppq_sample_encrypt_noise(s->cc, seed.ptr(), d_small.p, dnum, nullptr), "ppq_sample_encrypt_noise");
            ck(ppq_rekeygen(s->cc, d_sk.p, d_pkb.p, d_pka.p, d_small.p, d_kb.p, d_ka.p, nullptr), "ppq_rekeygen");
            ck(ppq_sync(s->cc, nullptr), "ppq_sync");
        } catch (const std::exception& e) {
            std::cerr << "[ReKeyGen] Re-encryption key generation failed (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::vector<uint64_t> kb(dnum * M * N), ka(dnum * M * N);
        d_kb.download(kb.data(), kb.size());
        d_ka.download(ka.data(), ka.size());
        std::cout << "[ReKeyGen] Re-encryption key generated successfully" << std::endl;

        try {
            std::vector<wire::Poly> av, bv;
            for (size_t j = 0; j < dnum; ++j) {
                av.push_back(wire::poly_from_flat(ka.data() + j * M * N, M, N, s->moduli.data()));
                bv.push_back(wire::poly_from_flat(kb.data() + j * M * N, M, N, s->moduli.data()));
            }
            wire::write_eval_key(rekey_path, av, bv, peer_tag);
        } catch (const std::exception&) {
            std::cerr << "[ReKeyGen] Failed to save re-encryption key to " << rekey_path << std::endl;
            return 1;
        }
        std::cout << "[ReKeyGen] Re-encryption key saved to " << rekey_path << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "[ReKeyGen] ERROR: " << e.what() << std::endl;
        return 1;
    }
}
