// encryptModelWeights -- drop-in for the reference's client/src/encryptModelWeights.cpp: skips "optimizer/" tensors
// (:71-74), encrypts mean and std_dev as one-value packed plaintexts (:81-94) and the values in zero-padded chunks of
// the batch size (:100-115).  cc->MakeCKKSPackedPlaintext + cc->Encrypt for EVERY plaintext of the file run as one
// ppq_encode + ppq_sample_encrypt_noise + ppq_encrypt batch through the C ABI.
#include <random>

#include "ppq_session.hpp"

using namespace ppqhost;

int main(int argc, char* argv[]) {
    if (argc != 5) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> <pubkey_path> <input_weights> <output_encfile>" << std::endl;
        return 1;
    }
    const std::string cc_path = argv[1], pubkey_path = argv[2], input_weights = argv[3], output_encfile = argv[4];
    try {
        std::unique_ptr<Session> s;
        try {
            s.reset(new Session(cc_path));
        } catch (const std::exception& e) {
            std::cerr << "[encrypt] ERROR: Failed to deserialize CryptoContext from " << cc_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[encrypt] CryptoContext loaded from " << cc_path << std::endl;
        const size_t batchSize = s->batch_size;  // cc->GetEncodingParams()->GetBatchSize() (:39-40): CC.json "bs"
        std::cout << "[encrypt] Batch size from CryptoContext = " << batchSize << std::endl;

        std::vector<uint64_t> pkb, pka;
        std::string key_tag;
        try {
            const std::string pk_text = wire::slurp(pubkey_path);
            key_tag = wire::read_key_tag(pk_text);  // a ciphertext carries the tag of the key it was made under
            const std::vector<wire::Poly> h = wire::read_public_key(pk_text);
            const size_t M = (size_t)s->L + s->K;
            pkb = s->flatten(h[0], M, "public key b");
            pka = s->flatten(h[1], M, "public key a");
        } catch (const std::exception& e) {
            std::cerr << "[encrypt] ERROR: Failed to deserialize public key from " << pubkey_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[encrypt] Public key loaded from " << pubkey_path << std::endl;

        jmin::Ptr root;
        try {
            root = jmin::parse(wire::slurp(input_weights));
        } catch (const std::exception& e) {
            std::cerr << "[encrypt] ERROR: Could not open input weights file: " << input_weights << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[encrypt] Weights loaded from " << input_weights << std::endl;

        // gather every plaintext vector of the file: [n][batchSize], zero padded
        std::vector<wire::Layer> out;
        std::vector<double> vals;
        std::vector<std::string*> dst;
        const jmin::Value& ws = root->at("weights_summary");
        out.reserve(ws.size());
        for (const auto& w : ws.arr) {
            const std::string layerName = w->at("layer").as_string();
            if (layerName.rfind("optimizer/", 0) == 0) {
                std::cout << "[encrypt] Skipping optimizer layer: " << layerName << std::endl;
                continue;
            }
            out.emplace_back();
            wire::Layer& l = out.back();
            for (const auto& kv : w->obj) {
                if (kv.first == "layer") l.layer = kv.second;
                if (kv.first == "shape") l.shape = kv.second;
            }
            const jmin::Value& samples = w->at("values");
            const size_t n_chunks = (samples.size() + batchSize - 1) / batchSize;
            l.values.resize(n_chunks);
            const size_t base = vals.size();
            vals.resize(base + (2 + n_chunks) * batchSize, 0.0);
            vals[base] = w->at("mean").as_double();
            vals[base + batchSize] = w->at("std_dev").as_double();
            for (size_t i = 0; i < samples.size(); ++i) vals[base + 2 * batchSize + i] = samples.arr[i]->as_double();
            dst.push_back(&l.mean);
            dst.push_back(&l.std_dev);
            for (auto& v : l.values) dst.push_back(&v);
        }
        const size_t n = dst.size(), ell = (size_t)s->L, ct = 2 * ell * s->N;
        if (n) {
            const double sf = ppq_ctx_scaling_factor_big(s->cc, 0);
            Dev<double> d_vals(vals.size());
            Dev<uint64_t> d_ct(n * ct), d_b(pkb.size()), d_a(pka.size());
            Dev<int64_t> d_small(n * 3 * s->N);
            d_vals.upload(vals.data(), vals.size());
            d_b.upload(pkb.data(), pkb.size());
            d_a.upload(pka.data(), pka.size());
            const ppqhost::SamplerSeed seed;  // NULL (fresh OS entropy per call) unless the test-only PPQ_SEED is set
            ck(ppq_sample_encrypt_noise(s->cc, seed.ptr(), d_small.p, n, nullptr), "ppq_sample_encrypt_noise");
            // MakeCKKSPackedPlaintext + Encrypt of every chunk of the file in one call (3*ell limb transforms per ciphertext)
            ck(ppq_encode_encrypt(s->cc, (int)ell, d_b.p, d_a.p, d_vals.p, batchSize, sf, d_small.p, d_ct.p, n, nullptr),
               "ppq_encode_encrypt");
            ck(ppq_sync(s->cc, nullptr), "ppq_sync");
            std::vector<uint64_t> h_ct(n * ct);
            d_ct.download(h_ct.data(), h_ct.size());
            wire::CtMeta m;  // fresh CKKS ciphertext under FLEXIBLEAUTOEXT: level 0, noiseScaleDeg 2, scale = sfBig(0)
            m.noise_scale_deg = 2;
            m.level = 0;
            m.scaling_factor = sf;
            m.slots = (uint32_t)batchSize;
            m.key_tag = key_tag;
#pragma omp parallel for schedule(dynamic)
            for (size_t k = 0; k < n; ++k) s->codec.encode_to(*dst[k], m, (uint32_t)ell, h_ct.data() + k * ct);
        }
        try {
            wire::write_container(output_encfile, out);
        } catch (const std::exception&) {
            std::cerr << "[encrypt] ERROR: Failed to write to output file: " << output_encfile << std::endl;
            return 1;
        }
        std::cout << "[encrypt] Encryption completed successfully and saved in " << output_encfile << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "[encrypt] ERROR: " << e.what() << std::endl;
        return 1;
    }
}
