// ppq_served -- resident aggregation server.
//
// The reference runs a round as three short-lived processes (orchestration/server_fns.sh:62-80:
// changeCipherDomain c1->c2, aggregateEncryptedWeights, changeCipherDomain c2->c1), each of which re-parses CC.json and
// a 26.8 MB JSON re-encryption key (server/src/changeCipherDomain.cpp:32-44) before touching a ciphertext, while the
// Mongoose drop-box (server/src/runMserver.cpp:237-285) only moves files.  This process keeps the context and every
// re-encryption key RESIDENT on the GPU and runs a whole round per request, for C clients instead of two:
//
//     agg_T   = ( sum_{i != T} ReEncrypt(enc_i, rk_{i->T}) + enc_T ) * (1/C)        ppq_reencrypt_aggregate_host,
//                                                                                   ppq_eval_add, ppq_eval_mult_const
//     out_i   = ReEncrypt(agg_T, rk_{T->i})  for every i != T                        ppq_reencrypt_fanout
//
// Files keep the reference's schema (weights_summary containers of base64 blobs), so the HTTP drop-box, the clients and
// the orchestrator are untouched: it is started once where the orchestrator used to exec the tools.
//
// Protocol: one request per line on stdin (or on a UNIX socket with --socket PATH), one reply line per request.
//     LOAD_REKEY <name> <rekey.json>                              -> OK rekey <name> parse_ms=..
//     ROUND target=<enc_T.json> out=<agg_T.json> up=<name>:<enc_i.json>... down=<name>:<out_i.json>...
//                                                                 -> OK clients=C cts=n decode_ms=.. gpu_ms=.. encode_ms=.. total_ms=..
//     STATS                                                       -> OK rounds=.. rekeys=.. key_parses=..
//     QUIT                                                        -> OK bye
// Errors: "ERR <message>"; the process stays up.
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <sstream>

#include "ppq_ops.hpp"

using namespace ppqhost;

struct Server {
    std::unique_ptr<Session> s;
    std::map<std::string, ppq_rekey*> rekeys;
    std::map<std::string, std::string> rekey_tag;
    size_t rounds = 0, key_parses = 0;

    ~Server() {
        for (auto& kv : rekeys) ppq_rekey_destroy(kv.second);
    }

    std::string load_rekey_cmd(const std::string& name, const std::string& path) {
        const double t0 = now_ms();
        std::string tag;
        ppq_rekey* rk = load_rekey(*s, path, &tag);
        ++key_parses;
        auto it = rekeys.find(name);
        if (it != rekeys.end()) ppq_rekey_destroy(it->second);
        rekeys[name] = rk;
        rekey_tag[name] = tag;
        std::ostringstream o;
        o << "OK rekey " << name << " parse_ms=" << (now_ms() - t0);
        return o.str();
    }

    ppq_rekey* key(const std::string& name) const {
        auto it = rekeys.find(name);
        if (it == rekeys.end()) throw std::runtime_error("unknown rekey '" + name + "' (LOAD_REKEY it first)");
        return it->second;
    }

    std::string round(const std::string& target, const std::string& out, const std::vector<std::pair<std::string, std::string>>& up,
                      const std::vector<std::pair<std::string, std::string>>& down) {
        const double t0 = now_ms();
        // ---- decode every upload (host): one batch per client
        EncFile tf;
        read_enc_file(*s, target, tf);
        std::vector<EncFile> uf(up.size());
        for (size_t i = 0; i < up.size(); ++i) {
            read_enc_file(*s, up[i].second, uf[i]);
            if (!same_structure(uf[i], tf)) throw std::runtime_error(up[i].second + ": tensors differ from the target client's file");
        }
        const uint32_t ell = tf.batch.ell;
        const size_t n = tf.batch.n(), N = s->N, ct = (size_t)2 * ell * N;
        const wire::CtMeta m0 = tf.batch.meta[0];
        for (const auto& f : uf)
            for (const auto& m : f.batch.meta)
                if (m.level != m0.level || m.noise_scale_deg != m0.noise_scale_deg)
                    throw std::runtime_error("EvalAdd operands differ in level / noise scale degree");
        const double t1 = now_ms();
        // ---- device: re-encrypt + sum the uploads, add the target client's own ciphertexts, scale by 1/C
        const size_t C = up.size() + 1;
        Dev<uint64_t> d_sum(n * ct), d_t(n * ct);
        d_t.upload(tf.batch.data.data(), n * ct);
        if (!up.empty()) {
            std::vector<const ppq_rekey*> rks;
            std::vector<const uint64_t*> ptrs;
            for (size_t i = 0; i < up.size(); ++i) {
                rks.push_back(key(up[i].first));
                ptrs.push_back(uf[i].batch.data.data());
            }
            std::vector<uint64_t> h_acc(n * ct);
            ck(ppq_reencrypt_aggregate_host(s->cc, rks.data(), (int)ell, ptrs.data(), up.size(), n, h_acc.data()), "ppq_reencrypt_aggregate_host");
            d_sum.upload(h_acc.data(), n * ct);
            const size_t step = std::max<size_t>(1, 65535 / (2 * ell));
            for (size_t k0 = 0; k0 < n; k0 += step) {
                const size_t k1 = std::min(n, k0 + step);
                ck(ppq_eval_add(s->cc, (int)ell, d_sum.p + k0 * ct, d_t.p + k0 * ct, d_sum.p + k0 * ct, k1 - k0, nullptr), "ppq_eval_add");
            }
        } else {
            cu(cudaMemcpy(d_sum.p, d_t.p, n * ct * sizeof(uint64_t), cudaMemcpyDeviceToDevice), "copy");
        }
        // cc->EvalMult(sum, 1/C) under FLEXIBLEAUTOEXT: rescale first iff noiseScaleDeg == 2 (aggregateEncryptedWeights.cpp:83)
        const bool rescale = m0.noise_scale_deg == 2;
        if (rescale && ell < 2) throw std::runtime_error("EvalMult: no limb left to rescale");
        const uint32_t lo = rescale ? ell - 1 : ell;
        const size_t ct_out = (size_t)2 * lo * N;
        const double sf = ppq_ctx_scaling_factor(s->cc, (int)(m0.level + (rescale ? 1 : 0)));
        Dev<uint64_t> d_agg(n * ct_out);
        for (size_t k0 = 0; k0 < n; k0 += 8192) {
            const size_t k1 = std::min(n, k0 + 8192);
            ck(ppq_eval_mult_const(s->cc, (int)ell, d_sum.p + k0 * ct, d_agg.p + k0 * ct_out, k1 - k0, 1.0 / (double)C, sf, rescale ? 1 : 0, nullptr),
               "ppq_eval_mult_const");
        }
        wire::CtMeta mo = m0;
        if (rescale) {
            mo.scaling_factor /= (double)s->moduli[ell - 1];
            mo.level += 1;
            mo.noise_scale_deg -= 1;
        }
        mo.noise_scale_deg += 1;
        mo.scaling_factor *= sf;
        // ---- fan-out of the aggregate to the other key domains (second changeCipherDomain pass, server_fns.sh:76-80)
        std::vector<Dev<uint64_t>> d_out(down.size());
        if (!down.empty()) {
            std::vector<const ppq_rekey*> rks;
            std::vector<uint64_t*> outs;
            for (size_t i = 0; i < down.size(); ++i) {
                rks.push_back(key(down[i].first));
                d_out[i].alloc(n * ct_out);
                outs.push_back(d_out[i].p);
            }
            ck(ppq_reencrypt_fanout(s->cc, rks.data(), (int)down.size(), (int)lo, d_agg.p, outs.data(), n, nullptr), "ppq_reencrypt_fanout");
        }
        ck(ppq_sync(s->cc, nullptr), "ppq_sync");
        std::vector<uint64_t> h(n * ct_out);
        d_agg.download(h.data(), h.size());
        const double t2 = now_ms();
        // ---- encode: same containers, new blobs
        std::vector<wire::CtMeta> metas(n, mo);
        for (size_t k = 0; k < n; ++k) metas[k].slots = tf.batch.meta[k].slots;
        encode_back(*s, tf.batch, lo, metas, h, tf.refs);
        wire::write_container(out, tf.layers);
        for (size_t i = 0; i < down.size(); ++i) {
            d_out[i].download(h.data(), h.size());
            std::vector<wire::CtMeta> md = metas;
            const std::string& tag = rekey_tag[down[i].first];
            if (!tag.empty())
                for (auto& m : md) m.key_tag = tag;
            encode_back(*s, tf.batch, lo, md, h, tf.refs);
            wire::write_container(down[i].second, tf.layers);
        }
        const double t3 = now_ms();
        ++rounds;
        std::ostringstream o;
        o << "OK clients=" << C << " cts=" << n << " limbs_out=" << lo << " decode_ms=" << (t1 - t0) << " gpu_ms=" << (t2 - t1)
          << " encode_ms=" << (t3 - t2) << " total_ms=" << (t3 - t0);
        return o.str();
    }

    std::string handle(const std::string& line, bool& quit) {
        std::istringstream in(line);
        std::string cmd;
        in >> cmd;
        try {
            if (cmd == "LOAD_REKEY") {
                std::string name, path;
                in >> name >> path;
                if (name.empty() || path.empty()) return "ERR usage: LOAD_REKEY <name> <path>";
                return load_rekey_cmd(name, path);
            }
            if (cmd == "ROUND") {
                std::string target, out, tok;
                std::vector<std::pair<std::string, std::string>> up, down;
                while (in >> tok) {
                    const size_t eq = tok.find('=');
                    if (eq == std::string::npos) return "ERR bad token '" + tok + "'";
                    const std::string k = tok.substr(0, eq), v = tok.substr(eq + 1);
                    if (k == "target") target = v;
                    else if (k == "out") out = v;
                    else if (k == "up" || k == "down") {
                        const size_t c = v.find(':');
                        if (c == std::string::npos) return "ERR " + k + " needs <rekey>:<file>";
                        (k == "up" ? up : down).emplace_back(v.substr(0, c), v.substr(c + 1));
                    } else return "ERR unknown key '" + k + "'";
                }
                if (target.empty() || out.empty()) return "ERR ROUND needs target= and out=";
                return round(target, out, up, down);
            }
            if (cmd == "STATS") {
                std::ostringstream o;
                o << "OK rounds=" << rounds << " rekeys=" << rekeys.size() << " key_parses=" << key_parses;
                return o.str();
            }
            if (cmd == "QUIT") {
                quit = true;
                return "OK bye";
            }
            return "ERR unknown command '" + cmd + "'";
        } catch (const std::exception& e) {
            return std::string("ERR ") + e.what();
        }
    }
};

static int serve_socket(Server& sv, const std::string& path) {
    const int fd = socket(AF_UNIX, SOCK_STREAM, 0);
    if (fd < 0) return 1;
    sockaddr_un a{};
    a.sun_family = AF_UNIX;
    std::strncpy(a.sun_path, path.c_str(), sizeof(a.sun_path) - 1);
    unlink(path.c_str());
    if (bind(fd, reinterpret_cast<sockaddr*>(&a), sizeof a) != 0 || listen(fd, 4) != 0) {
        std::cerr << "[served] ERROR: cannot listen on " << path << std::endl;
        return 1;
    }
    std::cout << "[served] listening on " << path << std::endl;
    bool quit = false;
    while (!quit) {
        const int c = accept(fd, nullptr, nullptr);
        if (c < 0) break;
        std::string buf;
        char tmp[4096];
        ssize_t r;
        while (!quit && (r = read(c, tmp, sizeof tmp)) > 0) {
            buf.append(tmp, (size_t)r);
            size_t nl;
            while (!quit && (nl = buf.find('\n')) != std::string::npos) {
                const std::string reply = sv.handle(buf.substr(0, nl), quit) + "\n";
                buf.erase(0, nl + 1);
                if (write(c, reply.data(), reply.size()) < 0) break;
            }
        }
        close(c);
    }
    close(fd);
    unlink(path.c_str());
    return 0;
}

int main(int argc, char* argv[]) {
    if (argc < 2) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> [--socket PATH] [--rekey NAME=PATH]..." << std::endl;
        return 1;
    }
    Server sv;
    try {
        sv.s.reset(new Session(argv[1]));
    } catch (const std::exception& e) {
        std::cerr << "[served] ERROR: Failed to load CryptoContext from " << argv[1] << " (" << e.what() << ")" << std::endl;
        return 1;
    }
    std::cout << "[served] CryptoContext loaded" << std::endl;
    std::string sock;
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--socket" && i + 1 < argc) sock = argv[++i];
        else if (a == "--rekey" && i + 1 < argc) {
            const std::string v = argv[++i];
            const size_t eq = v.find('=');
            if (eq == std::string::npos) {
                std::cerr << "[served] ERROR: --rekey NAME=PATH" << std::endl;
                return 1;
            }
            bool q = false;
            const std::string r = sv.handle("LOAD_REKEY " + v.substr(0, eq) + " " + v.substr(eq + 1), q);
            std::cout << "[served] " << r << std::endl;
            if (r.compare(0, 2, "OK") != 0) return 1;
        } else {
            std::cerr << "[served] ERROR: unknown argument " << a << std::endl;
            return 1;
        }
    }
    if (!sock.empty()) return serve_socket(sv, sock);
    std::cout << "[served] ready" << std::endl;
    std::string line;
    bool quit = false;
    while (!quit && std::getline(std::cin, line)) {
        if (line.empty()) continue;
        std::cout << sv.handle(line, quit) << std::endl;
    }
    return 0;
}
