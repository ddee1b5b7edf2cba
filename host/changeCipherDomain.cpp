// changeCipherDomain -- drop-in for the reference's server/src/changeCipherDomain.cpp (same argv, exit codes,
// progress lines and file schema), with cc->ReEncrypt (:74,89,105) replaced by ONE batched ppq_reencrypt call
// through the C ABI of libppq_b200.so.  No CPU fallback: without a B200 the context load fails and the tool exits 1.
#include "ppq_ops.hpp"

using namespace ppqhost;

int main(int argc, char* argv[]) {
    if (argc != 5) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> <rekey_path> <input_encfile> <output_encfile>" << std::endl;
        return 1;
    }
    const std::string cc_path = argv[1], rekey_path = argv[2], input_encfile = argv[3], output_encfile = argv[4];
    try {
        // Step 1: CryptoContext (reference :32-36)
        std::unique_ptr<Session> s;
        try {
            s.reset(new Session(cc_path));
        } catch (const std::exception& e) {
            std::cerr << "[recrypt] ERROR: Failed to load CryptoContext: " << cc_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[recrypt] CryptoContext loaded\n";

        // Step 2: re-encryption key (reference :40-44) -> device-resident [dnum][L+K][N] B and A vectors
        ppq_rekey* rk = nullptr;
        std::string rekey_tag;  // a re-encrypted ciphertext carries the tag of the key domain it now lives in
        try {
            rk = load_rekey(*s, rekey_path, &rekey_tag);
        } catch (const std::exception& e) {
            std::cerr << "[recrypt] ERROR: Failed to load ReKey from " << rekey_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[recrypt] ReKey loaded\n";

        // Step 3: encrypted weights (reference :48-55)
        std::vector<wire::Layer> layers;
        try {
            layers = wire::read_container(input_encfile);
        } catch (const std::exception& e) {
            std::cerr << "[recrypt] ERROR: Could not open input encrypted weights file (" << e.what() << ")\n";
            return 1;
        }

        // Step 4: ReEncrypt every field (reference :61-115), all ciphertexts of the file in one call per limb count
        std::vector<BlobRef> refs = blob_refs(layers);
        std::map<uint32_t, Batch> groups = decode_all(*s, refs);
        for (auto& kv : groups) {
            Batch& b = kv.second;
            Dev<uint64_t> d_in(b.data.size()), d_out(b.data.size());
            d_in.upload(b.data.data(), b.data.size());
            ck(ppq_reencrypt(s->cc, rk, (int)b.ell, d_in.p, d_out.p, b.n(), nullptr), "ppq_reencrypt");
            ck(ppq_sync(s->cc, nullptr), "ppq_sync");
            d_out.download(b.data.data(), b.data.size());
            if (!rekey_tag.empty())
                for (auto& m : b.meta) m.key_tag = rekey_tag;
            encode_back(*s, b, b.ell, b.meta, b.data.data(), refs);  // level / scale / slots unchanged by a key switch
        }
        ppq_rekey_destroy(rk);

        // Step 5: output (reference :118-124)
        try {
            wire::write_container(output_encfile, layers);
        } catch (const std::exception&) {
            std::cerr << "[recrypt] ERROR: Failed to open output file\n";
            return 1;
        }
        std::cout << "[recrypt] Re-encryption completed successfully. Output: " << output_encfile << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "[recrypt] ERROR: " << e.what() << std::endl;
        return 1;
    }
}
