// decryptModelWeights -- drop-in for the reference's client/src/decryptModelWeights.cpp: cc->Decrypt +
// GetRealPackedValue for every blob (:81-83,90-92,108-109) as one ppq_decrypt_core + ppq_decode batch; scalars are
// SetLength(1) (:82,91), value chunks are concatenated and trimmed to the product of the shape (:100-116).
// OpenFHE's randomised decode noise (IND-CPA-D countermeasure) is not added: the output is deterministic.
#include "ppq_session.hpp"

using namespace ppqhost;

int main(int argc, char* argv[]) {
    if (argc != 5) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> <privkey_path> <input_encfile> <output_file>" << std::endl;
        return 1;
    }
    const std::string cc_path = argv[1], privkey_path = argv[2], input_encfile = argv[3], output_file = argv[4];
    try {
        std::unique_ptr<Session> s;
        try {
            s.reset(new Session(cc_path));
        } catch (const std::exception& e) {
            std::cerr << "[decrypt] ERROR: Failed to load CryptoContext from " << cc_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[decrypt] CryptoContext loaded\n";

        std::vector<uint64_t> sk;
        try {
            const wire::Poly p = wire::read_private_key(wire::slurp(privkey_path));
            sk = s->flatten(p, p.limbs.size(), "private key");
            if ((int)p.limbs.size() < 1) throw std::runtime_error("empty key");
        } catch (const std::exception& e) {
            std::cerr << "[decrypt] ERROR: Failed to load private key from " << privkey_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        const size_t sk_limbs = sk.size() / s->N;
        std::cout << "[decrypt] Private key loaded\n";

        std::vector<wire::Layer> layers;
        try {
            layers = wire::read_container(input_encfile);
        } catch (const std::exception& e) {
            std::cerr << "[decrypt] ERROR: Could not open input file: " << input_encfile << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[decrypt] Encrypted weights loaded\n";

        std::vector<BlobRef> refs = blob_refs(layers);
        std::map<uint32_t, Batch> groups = decode_all(*s, refs);
        const size_t slots = s->batch_size;  // CC.json "bs" (full packing is enforced by Session)
        std::vector<std::vector<double>> plain(refs.size());
        Dev<uint64_t> d_sk(sk.size());
        d_sk.upload(sk.data(), sk.size());
        for (auto& kv : groups) {
            Batch& b = kv.second;
            if (b.ell > sk_limbs) throw std::runtime_error("ciphertext has more limbs than the private key");
            const size_t n = b.n();
            Dev<uint64_t> d_ct(b.data.size()), d_co(n * b.ell * s->N);
            Dev<double> d_vals(n * slots);
            d_ct.upload(b.data.data(), b.data.size());
            // ciphertexts of one file share their scaling factor in the reference flow; decode in runs of equal scale
            size_t k0 = 0;
            while (k0 < n) {
                size_t k1 = k0 + 1;
                while (k1 < n && b.meta[k1].scaling_factor == b.meta[k0].scaling_factor && k1 - k0 < 8192 &&
                       (k1 - k0 + 1) * b.ell <= 65535)
                    ++k1;
                ck(ppq_decrypt_core(s->cc, (int)b.ell, d_sk.p, d_ct.p + k0 * 2 * b.ell * s->N, d_co.p + k0 * b.ell * s->N, k1 - k0, nullptr), "ppq_decrypt_core");
                ck(ppq_decode(s->cc, (int)b.ell, d_co.p + k0 * b.ell * s->N, b.meta[k0].scaling_factor, d_vals.p + k0 * slots, k1 - k0, nullptr), "ppq_decode");
                k0 = k1;
            }
            ck(ppq_sync(s->cc, nullptr), "ppq_sync");
            std::vector<double> h(n * slots);
            d_vals.download(h.data(), h.size());
            for (size_t k = 0; k < n; ++k) plain[b.index[k]].assign(h.begin() + k * slots, h.begin() + (k + 1) * slots);
        }

        // Step 4: plaintext JSON (reference :68-121)
        jmin::Ptr root = jmin::make(jmin::Value::Object);
        jmin::Ptr arr = jmin::make(jmin::Value::Array);
        size_t r = 0;
        for (const auto& l : layers) {
            jmin::Ptr o = jmin::make(jmin::Value::Object);
            jmin::set(o, "layer", l.layer);
            jmin::set(o, "shape", l.shape);
            jmin::set(o, "mean", jmin::make_double(plain[r++][0]));
            jmin::set(o, "std_dev", jmin::make_double(plain[r++][0]));
            size_t expected = 1;
            for (const auto& d : l.shape->arr) expected *= (size_t)d->as_u64();
            jmin::Ptr vals = jmin::make(jmin::Value::Array);
            size_t count = 0;
            for (size_t j = 0; j < l.values.size(); ++j, ++r)
                for (double v : plain[r]) {
                    if (count++ >= expected) break;  // trim the zero padding of the last chunk
                    vals->arr.push_back(jmin::make_double(v));
                }
            jmin::set(o, "values", vals);
            arr->arr.push_back(o);
        }
        jmin::set(root, "weights_summary", arr);
        try {
            wire::spit(output_file, jmin::dump(*root, 2) + "\n");
        } catch (const std::exception&) {
            std::cerr << "[decrypt] ERROR: Failed to open output file: " << output_file << std::endl;
            return 1;
        }
        std::cout << "[decrypt] Decryption completed successfully. Output: " << output_file << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "[decrypt] ERROR: " << e.what() << std::endl;
        return 1;
    }
}
