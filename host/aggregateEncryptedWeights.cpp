// aggregateEncryptedWeights -- drop-in for the reference's server/src/aggregateEncryptedWeights.cpp: tensors are
// matched by layer + shape (:68-72), values pairwise up to the shorter list (:98-101); cc->EvalAdd (:82,91,106) and
// cc->EvalMult(ct_sum, 0.5) (:83,92,107) run as ONE ppq_eval_add + ONE ppq_eval_mult_const call over all pairs.
#include "ppq_session.hpp"

using namespace ppqhost;

int main(int argc, char* argv[]) {
    if (argc != 5) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> <client2_encfile> <client1to2_encfile> <output_aggfile>" << std::endl;
        return 1;
    }
    const std::string cc_path = argv[1], client2_file = argv[2], client1to2_file = argv[3], output_file = argv[4];
    try {
        std::unique_ptr<Session> s;
        try {
            s.reset(new Session(cc_path));
        } catch (const std::exception& e) {
            std::cerr << "[agg] ERROR: Failed to load CryptoContext from " << cc_path << " (" << e.what() << ")" << std::endl;
            return 1;
        }
        std::cout << "[agg] CryptoContext loaded\n";

        std::vector<wire::Layer> c2 = wire::read_container(client2_file), c1 = wire::read_container(client1to2_file);

        // Step 3: pair up blobs exactly as the reference's double loop does
        std::vector<wire::Layer> out;
        std::vector<const std::string*> lhs, rhs;  // ct1 (client 1 -> 2), ct2 (client 2)
        std::vector<std::string*> dst;
        size_t n_pairs = 0;
        for (const auto& w2 : c2)
            for (const auto& w1 : c1) {
                if (!wire::same_json(*w1.layer, *w2.layer)) continue;
                if (!wire::same_json(*w1.shape, *w2.shape)) continue;
                n_pairs += 2 + std::min(w1.values.size(), w2.values.size());
            }
        out.reserve(c2.size() * c1.size());
        for (const auto& w2 : c2)
            for (const auto& w1 : c1) {
                if (!wire::same_json(*w1.layer, *w2.layer)) continue;
                if (!wire::same_json(*w1.shape, *w2.shape)) continue;
                out.emplace_back();
                wire::Layer& a = out.back();
                a.layer = w2.layer;
                a.shape = w2.shape;
                a.values.resize(std::min(w1.values.size(), w2.values.size()));
                lhs.push_back(&w1.mean);    rhs.push_back(&w2.mean);    dst.push_back(&a.mean);
                lhs.push_back(&w1.std_dev); rhs.push_back(&w2.std_dev); dst.push_back(&a.std_dev);
                for (size_t j = 0; j < a.values.size(); ++j) {
                    lhs.push_back(&w1.values[j]); rhs.push_back(&w2.values[j]); dst.push_back(&a.values[j]);
                }
            }
        (void)n_pairs;

        // decode, grouped by limb count (operands of a pair must agree in level and noise scale degree, as in OpenFHE
        // without automatic level adjustment at equal depth)
        struct Group { std::vector<size_t> idx; std::vector<wire::CtMeta> meta; std::vector<uint64_t> a, b; };
        std::map<uint32_t, Group> groups;
        for (size_t i = 0; i < lhs.size(); ++i) {
            wire::Ct x = s->codec.decode(*lhs[i]), y = s->codec.decode(*rhs[i]);
            if (x.N != s->N || y.N != s->N) throw std::runtime_error("ciphertext ring dimension differs from CC.json");
            if (x.ell != y.ell || x.meta.level != y.meta.level || x.meta.noise_scale_deg != y.meta.noise_scale_deg)
                throw std::runtime_error("EvalAdd operands differ in level / noise scale degree");
            for (uint32_t k = 0; k < x.ell; ++k)
                if (x.moduli[k] != s->moduli[k] || y.moduli[k] != s->moduli[k]) throw std::runtime_error("ciphertext limb modulus does not match CC.json's basis");
            Group& g = groups[x.ell];
            g.idx.push_back(i);
            g.meta.push_back(x.meta);
            g.a.insert(g.a.end(), x.data.begin(), x.data.end());
            g.b.insert(g.b.end(), y.data.begin(), y.data.end());
        }
        for (auto& kv : groups) {
            const uint32_t ell = kv.first;
            Group& g = kv.second;
            const size_t n = g.idx.size();
            // EvalMult(ct, 0.5) under FLEXIBLEAUTOEXT: rescale first iff noiseScaleDeg == 2 (all members of a group
            // produced by the reference flow share it; split if a file ever mixes them)
            std::map<uint32_t, std::vector<size_t>> by_deg;
            for (size_t k = 0; k < n; ++k) by_deg[g.meta[k].noise_scale_deg].push_back(k);
            for (auto& dv : by_deg) {
                const bool rescale = dv.first == 2;
                if (rescale && ell < 2) throw std::runtime_error("EvalMult: no limb left to rescale");
                const std::vector<size_t>& sel = dv.second;
                const uint32_t lo = rescale ? ell - 1 : ell;
                const size_t ct_in = (size_t)2 * ell * s->N, ct_out = (size_t)2 * lo * s->N, m = sel.size();
                std::vector<uint64_t> ha(m * ct_in), hb(m * ct_in), ho(m * ct_out);
                for (size_t k = 0; k < m; ++k) {
                    std::copy(g.a.begin() + sel[k] * ct_in, g.a.begin() + (sel[k] + 1) * ct_in, ha.begin() + k * ct_in);
                    std::copy(g.b.begin() + sel[k] * ct_in, g.b.begin() + (sel[k] + 1) * ct_in, hb.begin() + k * ct_in);
                }
                Dev<uint64_t> da(ha.size()), db(hb.size()), dsum(ha.size()), dout(ho.size());
                da.upload(ha.data(), ha.size());
                db.upload(hb.data(), hb.size());
                const size_t step_add = std::max<size_t>(1, 65535 / (2 * ell)), step_mul = 8192;
                for (size_t k0 = 0; k0 < m; k0 += step_add) {
                    const size_t k1 = std::min(m, k0 + step_add);
                    ck(ppq_eval_add(s->cc, (int)ell, da.p + k0 * ct_in, db.p + k0 * ct_in, dsum.p + k0 * ct_in, k1 - k0, nullptr), "ppq_eval_add");
                }
                const wire::CtMeta& m0 = g.meta[sel[0]];
                const uint32_t level_out = m0.level + (rescale ? 1 : 0);
                const double sf = ppq_ctx_scaling_factor(s->cc, (int)level_out);
                for (size_t k0 = 0; k0 < m; k0 += step_mul) {
                    const size_t k1 = std::min(m, k0 + step_mul);
                    ck(ppq_eval_mult_const(s->cc, (int)ell, dsum.p + k0 * ct_in, dout.p + k0 * ct_out, k1 - k0, 0.5, sf, rescale ? 1 : 0, nullptr), "ppq_eval_mult_const");
                }
                ck(ppq_sync(s->cc, nullptr), "ppq_sync");
                dout.download(ho.data(), ho.size());
                for (size_t k = 0; k < m; ++k) {
                    wire::CtMeta mo = g.meta[sel[k]];
                    if (rescale) {
                        mo.scaling_factor /= (double)s->moduli[ell - 1];
                        mo.level += 1;
                        mo.noise_scale_deg -= 1;
                    }
                    mo.noise_scale_deg += 1;
                    mo.scaling_factor *= sf;
                    *dst[g.idx[sel[k]]] = s->codec.encode(mo, lo, ho.data() + k * ct_out);
                }
            }
        }

        wire::write_container(output_file, out);
        std::cout << "[agg] Aggregation completed successfully. Output: " << output_file << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "[agg] ERROR: " << e.what() << std::endl;
        return 1;
    }
}
