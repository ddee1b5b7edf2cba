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
//
// --http <sConfig.json>: the same process is ALSO the file drop-box of the reference (server/src/runMserver.cpp:237-285,
// same endpoints, same sConfig.json keys; HTTP/1.1, multipart uploads with the part name "file", `Expect: 100-continue`
// answered), generalised from the hard-coded C1/C2 to CLIENT_<i>_* keys for i = 1..C:
//     GET  /getCC, /sendPbKeyC<i>, /download/<rel>             POST /uploadPubKeyC<i>, /uploadReKeyC<i>, /uploadEncWeightsC<i>,
//                                                                   /uploadDomainChange, /uploadAggregated, /uploadDomainChangeAgg
// and it CALLS THE GPU PATH: an uploaded re-encryption key is loaded onto the device at once, and when the encrypted
// weights of all C clients of a round have arrived the round runs in-process (target domain = client C, as in the
// reference's flow for C = 2) and writes AGGREGATED_ENCRYPTED_WEIGHTS_PATH and OUTPUT_AGGREGATED_DOMAIN_CHANGED_PATH[_<i>],
// which the clients then GET exactly as before (orchestration/comm_fns.sh:134-145).  The orchestrator's three server
// steps per round (server_fns.sh:62-80) become no-ops.  Re-encryption keys: CLIENT_<i>_REKEY is client i -> client C for
// i < C, and CLIENT_<C>_REKEY (C = 2) or CLIENT_<C>_REKEY_TO_<i> is client C -> client i.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <omp.h>

#include <cctype>
#include <filesystem>
#include <fstream>
#include <sstream>

#include "ppq_ops.hpp"

using namespace ppqhost;

struct Server {
    std::unique_ptr<Session> s;
    std::map<std::string, ppq_rekey*> rekeys;
    std::map<std::string, std::string> rekey_tag;
    size_t rounds = 0, key_parses = 0;
    // device buffers of a round, kept from round to round (grow-only): no cudaMalloc / cudaFree in the steady state
    DevGrow<uint64_t> dv_sum, dv_t, dv_agg;
    std::vector<std::unique_ptr<DevGrow<uint64_t>>> dv_out;

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
        std::vector<EncFile> uf(up.size());
        {   // the files are independent: one reader per file, each decoding its blobs with its own inner team
            std::vector<std::string> errs(up.size() + 1);
            omp_set_max_active_levels(2);
            const int inner = std::max(1, omp_get_num_procs() / (int)(up.size() + 1));
#pragma omp parallel for schedule(static, 1) num_threads((int)up.size() + 1)
            for (size_t i = 0; i <= up.size(); ++i) {
                omp_set_num_threads(inner);
                try {
                    if (i == 0) read_enc_file(*s, target, tf);
                    else read_enc_file(*s, up[i - 1].second, uf[i - 1]);
                } catch (const std::exception& e) {
                    errs[i] = e.what()[0] ? e.what() : "read error";
                }
            }
            for (const std::string& e : errs)
                if (!e.empty()) throw std::runtime_error(e);
        }
        for (size_t i = 0; i < up.size(); ++i)
            if (!same_structure(uf[i], tf)) throw std::runtime_error(up[i].second + ": tensors differ from the target client's file");
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
        struct View {  // the round's device buffers live in the server
            uint64_t* p;
            void upload(const uint64_t* h, size_t count) { cu(cudaMemcpy(p, h, count * sizeof(uint64_t), cudaMemcpyHostToDevice), "H2D copy"); }
            void download(uint64_t* h, size_t count) const { cu(cudaMemcpy(h, p, count * sizeof(uint64_t), cudaMemcpyDeviceToHost), "D2H copy"); }
        };
        View d_sum{dv_sum.ensure(n * ct)}, d_t{dv_t.ensure(n * ct)};
        d_t.upload(tf.batch.data.data(), n * ct);
        if (!up.empty()) {
            std::vector<const ppq_rekey*> rks;
            std::vector<const uint64_t*> ptrs;
            for (size_t i = 0; i < up.size(); ++i) {
                rks.push_back(key(up[i].first));
                ptrs.push_back(uf[i].batch.data.data());
            }
            HVec h_acc(n * ct);
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
        View d_agg{dv_agg.ensure(n * ct_out)};
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
        std::vector<View> d_out(down.size());
        if (!down.empty()) {
            std::vector<const ppq_rekey*> rks;
            std::vector<uint64_t*> outs;
            while (dv_out.size() < down.size()) dv_out.emplace_back(new DevGrow<uint64_t>);
            for (size_t i = 0; i < down.size(); ++i) {
                rks.push_back(key(down[i].first));
                d_out[i].p = dv_out[i]->ensure(n * ct_out);
                outs.push_back(d_out[i].p);
            }
            ck(ppq_reencrypt_fanout(s->cc, rks.data(), (int)down.size(), (int)lo, d_agg.p, outs.data(), n, nullptr), "ppq_reencrypt_fanout");
        }
        ck(ppq_sync(s->cc, nullptr), "ppq_sync");
        HVec h(n * ct_out);
        d_agg.download(h.data(), h.size());
        const double t2 = now_ms();
        // ---- encode: same containers, new blobs; the outputs are independent, one writer per file (the aggregate re-uses
        // the target file's container in place, every fan-out file gets its own copy of it)
        std::vector<wire::CtMeta> metas(n, mo);
        for (size_t k = 0; k < n; ++k) metas[k].slots = tf.batch.meta[k].slots;
        {
            const size_t n_out = down.size() + 1;
            std::vector<std::string> errs(n_out), tags(n_out);
            for (size_t i = 1; i < n_out; ++i) {
                auto it = rekey_tag.find(down[i - 1].first);
                if (it != rekey_tag.end()) tags[i] = it->second;
            }
            std::vector<std::vector<wire::Layer>> copies(n_out);
            omp_set_max_active_levels(2);
            const int inner = std::max(1, omp_get_num_procs() / (int)n_out);
#pragma omp parallel num_threads((int)n_out)
            {
                const size_t tid = (size_t)omp_get_thread_num(), nth = (size_t)omp_get_num_threads();
                omp_set_num_threads(inner);
                // first every fan-out writer copies the container (blob texts of the right size to overwrite) ...
                for (size_t i = tid; i < n_out; i += nth)
                    if (i > 0) {
                        try {
                            copies[i] = tf.layers;
                        } catch (const std::exception& e) {
                            errs[i] = "out of memory";
                        }
                    }
#pragma omp barrier
                // ... then the aggregate overwrites the original in place
                for (size_t i = tid; i < n_out; i += nth) {
                    if (!errs[i].empty()) continue;
                    try {
                        if (i == 0) {
                            encode_back(*s, tf.batch, lo, metas, h.data(), tf.refs);
                            wire::write_container(out, tf.layers);
                        } else {
                            HVec hi(n * ct_out);
                            d_out[i - 1].download(hi.data(), hi.size());
                            std::vector<wire::CtMeta> md = metas;
                            if (!tags[i].empty())
                                for (auto& m : md) m.key_tag = tags[i];
                            const std::vector<BlobRef> refs = blob_refs(copies[i]);
                            encode_back(*s, tf.batch, lo, md, hi.data(), refs);
                            wire::write_container(down[i - 1].second, copies[i]);
                        }
                    } catch (const std::exception& e) {
                        errs[i] = e.what()[0] ? e.what() : "write error";
                    }
                }
            }
            for (const std::string& e : errs)
                if (!e.empty()) throw std::runtime_error(e);
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

// ------------------------------------------------------------------------------------------------ HTTP drop-box
struct HttpCfg {
    std::string ip = "0.0.0.0", cc_path;
    int port = 8000, C = 0;
    std::map<std::string, std::string> clients;  // the "CLIENTS" object of sConfig.json, verbatim
    const std::string* get(const std::string& k) const {
        auto it = clients.find(k);
        return it == clients.end() ? nullptr : &it->second;
    }
};
static HttpCfg load_http_cfg(const std::string& path) {
    jmin::Ptr j = jmin::parse(wire::slurp(path));
    HttpCfg c;
    if (const jmin::Value* m = j->find("mSConfig")) {
        if (const jmin::Value* v = m->find("SERVER_IP")) c.ip = v->as_string();
        if (const jmin::Value* v = m->find("SERVER_PORT")) c.port = (int)v->as_i64();
    }
    c.cc_path = j->at("CC").at("path").as_string();
    for (const auto& kv : j->at("CLIENTS").obj) c.clients[kv.first] = kv.second->as_string();
    while (c.get("CLIENT_" + std::to_string(c.C + 1) + "_ENCRYPTED_WEIGHTS_PATH")) ++c.C;
    if (c.C < 1) throw std::runtime_error("sConfig.json: no CLIENT_<i>_ENCRYPTED_WEIGHTS_PATH");
    return c;
}
static bool read_n(int fd, std::string& buf, size_t want) {
    char tmp[65536];
    while (buf.size() < want) {
        const ssize_t r = read(fd, tmp, std::min(sizeof tmp, want - buf.size()));
        if (r <= 0) return false;
        buf.append(tmp, (size_t)r);
    }
    return true;
}
static void http_reply(int fd, int code, const std::string& ctype, const std::string& body) {
    const char* txt = code == 200 ? "OK" : code == 404 ? "Not Found" : code == 405 ? "Method Not Allowed" : "Internal Server Error";
    std::ostringstream h;
    h << "HTTP/1.1 " << code << " " << txt << "\r\n";
    if (!ctype.empty()) h << "Content-Type: " << ctype << "\r\n";
    h << "Content-Length: " << body.size() << "\r\nConnection: close\r\n\r\n";
    const std::string head = h.str();
    (void)!write(fd, head.data(), head.size());
    size_t off = 0;
    while (off < body.size()) {
        const ssize_t w = write(fd, body.data() + off, body.size() - off);
        if (w <= 0) break;
        off += (size_t)w;
    }
}
// body of the multipart part called "file" (the reference's handle_upload takes exactly that part, :152-160)
static bool multipart_file(const std::string& ctype, const std::string& body, std::string& out) {
    const size_t b = ctype.find("boundary=");
    if (b == std::string::npos) return false;
    std::string boundary = ctype.substr(b + 9);
    if (!boundary.empty() && boundary[0] == '"') boundary = boundary.substr(1, boundary.find('"', 1) - 1);
    const std::string delim = "--" + boundary;
    size_t pos = body.find(delim);
    while (pos != std::string::npos) {
        pos += delim.size();
        if (body.compare(pos, 2, "--") == 0) break;
        const size_t hend = body.find("\r\n\r\n", pos);
        if (hend == std::string::npos) break;
        const std::string headers = body.substr(pos, hend - pos);
        const size_t next = body.find("\r\n" + delim, hend + 4);
        if (next == std::string::npos) break;
        if (headers.find("name=\"file\"") != std::string::npos) {
            out = body.substr(hend + 4, next - (hend + 4));
            return true;
        }
        pos = next + 2;
    }
    return false;
}

struct HttpServer {
    Server& sv;
    HttpCfg cfg;
    std::vector<bool> have;  // encrypted weights of client i received this round
    size_t rounds_run = 0;
    HttpServer(Server& s, const HttpCfg& c) : sv(s), cfg(c), have((size_t)c.C, false) {}

    std::string rekey_name_up(int i) const { return "up" + std::to_string(i); }
    std::string rekey_name_down(int i) const { return "down" + std::to_string(i); }
    // which resident key a just-uploaded CLIENT_<i>_REKEY[_TO_<k>] file is
    void maybe_load_rekey(const std::string& key, const std::string& path) {
        const int C = cfg.C;
        for (int i = 1; i <= C; ++i) {
            if (i < C && key == "CLIENT_" + std::to_string(i) + "_REKEY") {
                bool q = false;
                std::cout << "[served] " << sv.handle("LOAD_REKEY " + rekey_name_up(i) + " " + path, q) << std::endl;
            }
            if (key == "CLIENT_" + std::to_string(C) + "_REKEY_TO_" + std::to_string(i) || (C == 2 && i == 1 && key == "CLIENT_2_REKEY")) {
                bool q = false;
                std::cout << "[served] " << sv.handle("LOAD_REKEY " + rekey_name_down(i) + " " + path, q) << std::endl;
            }
        }
    }
    void maybe_run_round() {
        for (bool h : have)
            if (!h) return;
        const int C = cfg.C;
        std::ostringstream cmd;
        cmd << "ROUND target=" << *cfg.get("CLIENT_" + std::to_string(C) + "_ENCRYPTED_WEIGHTS_PATH")
            << " out=" << *cfg.get("AGGREGATED_ENCRYPTED_WEIGHTS_PATH");
        for (int i = 1; i < C; ++i) {
            cmd << " up=" << rekey_name_up(i) << ":" << *cfg.get("CLIENT_" + std::to_string(i) + "_ENCRYPTED_WEIGHTS_PATH");
            const std::string* o = cfg.get(i == 1 ? "OUTPUT_AGGREGATED_DOMAIN_CHANGED_PATH" : "OUTPUT_AGGREGATED_DOMAIN_CHANGED_PATH_" + std::to_string(i));
            if (o && sv.rekeys.count(rekey_name_down(i))) cmd << " down=" << rekey_name_down(i) << ":" << *o;
        }
        bool q = false;
        const std::string r = sv.handle(cmd.str(), q);
        std::cout << "[served] round " << ++rounds_run << ": " << r << std::endl;
        std::fill(have.begin(), have.end(), false);
    }

    void serve_file(int fd, const std::string& path) {
        std::string data;
        try {
            data = wire::slurp(path);
        } catch (const std::exception&) {
            http_reply(fd, 404, "", "Not found\n");
            return;
        }
        std::cout << "[SERVER] Serving file " << path << " (" << data.size() << " bytes)" << std::endl;
        http_reply(fd, 200, "application/octet-stream", data);
    }
    void upload(int fd, const std::string& method, const std::string& ctype, const std::string& body, const std::string& key) {
        if (method != "POST") return http_reply(fd, 405, "", "Method not allowed\n");
        const std::string* dest = cfg.get(key);
        std::string file;
        if (!dest || !multipart_file(ctype, body, file)) return http_reply(fd, 500, "", "Error: cannot open file for writing\n");
        try {
            const std::filesystem::path dp(*dest);
            if (dp.has_parent_path()) std::filesystem::create_directories(dp.parent_path());
            wire::spit(*dest, file);
        } catch (const std::exception&) {
            return http_reply(fd, 500, "", "Error: cannot open file for writing\n");
        }
        std::cout << "[SERVER] Received " << file.size() << " bytes, saved to " << *dest << std::endl;
        http_reply(fd, 200, "application/json", "{\"status\":\"received\"}");
        if (key.find("_REKEY") != std::string::npos) maybe_load_rekey(key, *dest);
        for (int i = 1; i <= cfg.C; ++i)
            if (key == "CLIENT_" + std::to_string(i) + "_ENCRYPTED_WEIGHTS_PATH") {
                have[(size_t)i - 1] = true;
                maybe_run_round();
            }
    }
    void handle(int fd) {
        std::string buf;
        char tmp[8192];
        size_t hend;
        while ((hend = buf.find("\r\n\r\n")) == std::string::npos) {
            const ssize_t r = read(fd, tmp, sizeof tmp);
            if (r <= 0) return;
            buf.append(tmp, (size_t)r);
            if (buf.size() > (1u << 20)) return;
        }
        std::istringstream rl(buf.substr(0, buf.find("\r\n")));
        std::string method, uri;
        rl >> method >> uri;
        std::string headers = buf.substr(0, hend + 4), lower = headers;
        for (auto& ch : lower) ch = (char)std::tolower((unsigned char)ch);
        auto header = [&](const std::string& name) {
            const size_t p = lower.find("\r\n" + name + ":");
            if (p == std::string::npos) return std::string();
            const size_t s0 = p + name.size() + 3, e0 = headers.find("\r\n", s0);
            std::string v = headers.substr(s0, e0 - s0);
            while (!v.empty() && v[0] == ' ') v.erase(0, 1);
            return v;
        };
        const size_t clen = (size_t)std::strtoull(header("content-length").c_str(), nullptr, 10);
        if (lower.find("expect: 100-continue") != std::string::npos) (void)!write(fd, "HTTP/1.1 100 Continue\r\n\r\n", 25);
        std::string body = buf.substr(hend + 4);
        if (clen > ((size_t)1 << 32) || !read_n(fd, body, clen)) return;
        const std::string ctype = header("content-type");
        auto client_of = [&](const std::string& prefix) -> int {  // "/uploadReKeyC7" -> 7
            if (uri.compare(0, prefix.size(), prefix) != 0) return 0;
            const int i = std::atoi(uri.c_str() + prefix.size());
            return (i >= 1 && i <= cfg.C && uri == prefix + std::to_string(i)) ? i : 0;
        };
        int i;
        if (uri == "/getCC" && method == "GET") {
            std::cout << "[SERVER] Serving " << cfg.cc_path << std::endl;
            serve_file(fd, cfg.cc_path);
        } else if ((i = client_of("/sendPbKeyC")) && method == "GET") {
            const std::string* p = cfg.get("CLIENT_" + std::to_string(i) + "_PUBLIC");
            if (p) serve_file(fd, *p);
            else http_reply(fd, 500, "", "Error: cannot open pubkey file\n");
        } else if (uri.size() > 10 && uri.compare(0, 10, "/download/") == 0) {
            const std::string rel = uri.substr(10);
            const size_t slash = cfg.cc_path.rfind('/');
            const std::string base = slash == std::string::npos ? "." : cfg.cc_path.substr(0, slash);
            if (rel.find("..") != std::string::npos) http_reply(fd, 404, "", "Not found\n");
            else serve_file(fd, base + "/" + rel);
        } else if ((i = client_of("/uploadPubKeyC"))) {
            upload(fd, method, ctype, body, "CLIENT_" + std::to_string(i) + "_PUBLIC");
        } else if ((i = client_of("/uploadReKeyC"))) {
            upload(fd, method, ctype, body, "CLIENT_" + std::to_string(i) + "_REKEY");
        } else if (uri.compare(0, 13, "/uploadReKeyC") == 0 && uri.find("To") != std::string::npos) {  // /uploadReKeyC<C>To<i>
            const int to = std::atoi(uri.c_str() + uri.find("To") + 2);
            upload(fd, method, ctype, body, "CLIENT_" + std::to_string(cfg.C) + "_REKEY_TO_" + std::to_string(to));
        } else if ((i = client_of("/uploadEncWeightsC"))) {
            upload(fd, method, ctype, body, "CLIENT_" + std::to_string(i) + "_ENCRYPTED_WEIGHTS_PATH");
        } else if (uri == "/uploadDomainChange") {
            upload(fd, method, ctype, body, "OUTPUT_DOMAIN_CHANGED_PATH");
        } else if (uri == "/uploadAggregated") {
            upload(fd, method, ctype, body, "AGGREGATED_ENCRYPTED_WEIGHTS_PATH");
        } else if (uri == "/uploadDomainChangeAgg") {
            upload(fd, method, ctype, body, "OUTPUT_AGGREGATED_DOMAIN_CHANGED_PATH");
        } else if (uri == "/shutdown" && method == "POST") {  // not in the reference: lets a test stop the listener
            http_reply(fd, 200, "application/json", "{\"status\":\"bye\"}");
            stop = true;
        } else {
            http_reply(fd, 404, "", "Not found\n");
        }
    }
    bool stop = false;
    int run() {
        const int fd = socket(AF_INET, SOCK_STREAM, 0);
        if (fd < 0) return 1;
        int one = 1;
        setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
        sockaddr_in a{};
        a.sin_family = AF_INET;
        a.sin_port = htons((uint16_t)cfg.port);
        if (inet_pton(AF_INET, cfg.ip.c_str(), &a.sin_addr) != 1) a.sin_addr.s_addr = INADDR_ANY;
        if (bind(fd, reinterpret_cast<sockaddr*>(&a), sizeof a) != 0 || listen(fd, 16) != 0) {
            std::cerr << "[SERVER] Failed to start server on http://" << cfg.ip << ":" << cfg.port << std::endl;
            return 1;
        }
        std::cout << "[SERVER] ppq_served HTTP server running on http://" << cfg.ip << ":" << cfg.port << " (" << cfg.C << " clients)" << std::endl;
        while (!stop) {
            const int c = accept(fd, nullptr, nullptr);
            if (c < 0) break;
            try {
                handle(c);
            } catch (const std::exception& e) {
                http_reply(c, 500, "", std::string("Error: ") + e.what() + "\n");
            }
            close(c);
        }
        close(fd);
        return 0;
    }
};

int main(int argc, char* argv[]) {
    if (argc < 2) {
        std::cerr << "Usage: " << argv[0] << " <cc_path> [--socket PATH] [--rekey NAME=PATH]...\n       " << argv[0]
                  << " --http <sConfig.json>" << std::endl;
        return 1;
    }
    Server sv;
    if (std::string(argv[1]) == "--http") {
        if (argc != 3) {
            std::cerr << "Usage: " << argv[0] << " --http <sConfig.json>" << std::endl;
            return 1;
        }
        try {
            const HttpCfg cfg = load_http_cfg(argv[2]);
            sv.s.reset(new Session(cfg.cc_path));
            HostPool::get().pinned = true;  // resident process: ciphertext batches in pooled pinned memory
            std::cout << "[served] CryptoContext loaded" << std::endl;
            HttpServer hs(sv, cfg);
            for (const auto& kv : cfg.clients)  // keys already on disk from an earlier run stay usable
                if (kv.first.find("_REKEY") != std::string::npos && std::ifstream(kv.second).good()) hs.maybe_load_rekey(kv.first, kv.second);
            return hs.run();
        } catch (const std::exception& e) {
            std::cerr << "[SERVER] ERROR: " << e.what() << std::endl;
            return 1;
        }
    }
    try {
        sv.s.reset(new Session(argv[1]));
        HostPool::get().pinned = true;  // resident process: ciphertext batches in pooled pinned memory
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
