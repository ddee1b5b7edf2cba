// ppq_ops.hpp -- the file-level operations the tools and the resident server share (C++17 over the C ABI).
#pragma once
#include <chrono>
#include <memory>

#include "ppq_session.hpp"

namespace ppqhost {

inline double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Serial::DeserializeFromFile(rekey_path, reKey, SerType::JSON) (server/src/changeCipherDomain.cpp:40-44) -> a
// device-resident handle.  `tag_out`: CryptoObject key tag of the file ("" when it carries none).
inline ppq_rekey* load_rekey(const Session& s, const std::string& rekey_path, std::string* tag_out = nullptr) {
    const std::string text = wire::slurp(rekey_path);
    if (tag_out) *tag_out = wire::read_key_tag(text);
    const wire::EvalKeyPolys ek = wire::read_eval_key(text);
    if ((int)ek.a.size() != s.dnum)
        throw std::runtime_error("key has " + std::to_string(ek.a.size()) + " digits, context dnum is " + std::to_string(s.dnum));
    const size_t M = (size_t)s.L + s.K, per = M * s.N;
    std::vector<uint64_t> b(per * s.dnum), a(per * s.dnum);
    for (int j = 0; j < s.dnum; ++j) {
        const std::vector<uint64_t> bj = s.flatten(ek.b[j], M, "rekey B"), aj = s.flatten(ek.a[j], M, "rekey A");
        std::copy(bj.begin(), bj.end(), b.begin() + j * per);
        std::copy(aj.begin(), aj.end(), a.begin() + j * per);
    }
    ppq_rekey* rk = nullptr;
    ck(ppq_rekey_upload(s.cc, b.data(), a.data(), &rk), "ppq_rekey_upload");
    return rk;
}

// One encrypted-weights container with every blob decoded into ONE host batch [n][2][ell][N] (file order).
struct EncFile {
    std::vector<wire::Layer> layers;
    std::vector<BlobRef> refs;
    Batch batch;  // all blobs; a file of the reference flow holds one limb count
};
inline void read_enc_file(const Session& s, const std::string& path, EncFile& f) {
    f.layers = wire::read_container(path);
    f.refs = blob_refs(f.layers);
    std::map<uint32_t, Batch> groups = decode_all(s, f.refs);
    if (groups.size() != 1) throw std::runtime_error(path + ": ciphertexts with different limb counts in one file");
    f.batch = std::move(groups.begin()->second);
}
// same tensor list (layer + shape, the reference's match rule: aggregateEncryptedWeights.cpp:68-72) and chunk counts?
inline bool same_structure(const EncFile& a, const EncFile& b) {
    if (a.layers.size() != b.layers.size() || a.batch.ell != b.batch.ell) return false;
    for (size_t i = 0; i < a.layers.size(); ++i) {
        if (!wire::same_json(*a.layers[i].layer, *b.layers[i].layer)) return false;
        if (!wire::same_json(*a.layers[i].shape, *b.layers[i].shape)) return false;
        if (a.layers[i].values.size() != b.layers[i].values.size()) return false;
    }
    return true;
}

}  // namespace ppqhost
