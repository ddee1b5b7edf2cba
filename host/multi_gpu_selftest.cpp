// multi_gpu_selftest -- a plain C++ host (no Python, no torch) aggregating over several GPUs through the C ABI alone:
// one thread per device, clients sharded round-robin (SURVEY.md 8 row e), every rank calls
// ppq_reencrypt_aggregate_host_ranks (local ppq_agg_* + ONE ncclReduce + mod q on the root), and the root's result is
// compared bit for bit with the single-GPU aggregate of all clients; a second pass does the same over PACKED host buffers
// (ppq_reencrypt_aggregate_host_packed_ranks) with the shards ppq_plan_shards gives for unequal per-rank costs.  NCCL comes from the system libnccl.so.2 (dlopen).
//
//     multi_gpu_selftest [n_gpus] [n_clients] [n_slots] [logN] [depth]        exit 0 = identical
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../include/ppq_b200.h"

#define CK(call)                                                                          \
    do {                                                                                  \
        if ((call) != PPQ_OK) {                                                           \
            std::fprintf(stderr, "%s failed: %s\n", #call, ppq_last_error());             \
            std::exit(1);                                                                 \
        }                                                                                 \
    } while (0)

int main(int argc, char** argv) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        std::fprintf(stderr, "no CUDA device (libppq_b200 has no CPU fallback)\n");
        return 1;
    }
    const int G = argc > 1 ? std::atoi(argv[1]) : ndev;
    const int C = argc > 2 ? std::atoi(argv[2]) : 6;
    const size_t S = argc > 3 ? (size_t)std::atoi(argv[3]) : 3;
    const int logN = argc > 4 ? std::atoi(argv[4]) : 13, depth = argc > 5 ? std::atoi(argv[5]) : 4;
    if (G < 1 || G > ndev || C < 1) {
        std::fprintf(stderr, "need 1 <= n_gpus <= %d and at least one client\n", ndev);
        return 1;
    }
    // one context per device, same parameters
    std::vector<ppq_ctx*> ctx(G);
    for (int r = 0; r < G; ++r) CK(ppq_ctx_create_synthetic(logN, depth, 60, 50, 20, 2, r, &ctx[r]));
    ppq_ctx_info_t info;
    CK(ppq_ctx_info(ctx[0], &info));
    const size_t N = (size_t)1 << info.logN, M = (size_t)info.L + info.K, ell = (size_t)info.L;
    std::vector<uint64_t> mod(M);
    CK(ppq_ctx_moduli(ctx[0], mod.data()));
    const size_t ct = 2 * ell * N, blk = ct * S, keyn = (size_t)info.dnum * M * N;
    // synthetic canonical residues: C clients' ciphertexts and re-encryption keys
    std::mt19937_64 rng(2024);
    std::vector<std::vector<uint64_t>> cts(C, std::vector<uint64_t>(blk)), kb(C, std::vector<uint64_t>(keyn)), ka(C, std::vector<uint64_t>(keyn));
    for (int c = 0; c < C; ++c) {
        for (size_t i = 0; i < blk; ++i) cts[c][i] = rng() % mod[(i / N) % ell];
        for (size_t i = 0; i < keyn; ++i) {
            kb[c][i] = rng() % mod[(i / N) % M];
            ka[c][i] = rng() % mod[(i / N) % M];
        }
    }
    // reference: everything on device 0
    std::vector<uint64_t> want(blk), got(blk);
    {
        std::vector<ppq_rekey*> rk(C);
        std::vector<const uint64_t*> ptr(C);
        for (int c = 0; c < C; ++c) {
            CK(ppq_rekey_upload(ctx[0], kb[c].data(), ka[c].data(), &rk[c]));
            ptr[c] = cts[c].data();
        }
        CK(ppq_reencrypt_aggregate_host(ctx[0], rk.data(), (int)ell, ptr.data(), (size_t)C, S, want.data()));
        for (auto* k : rk) ppq_rekey_destroy(k);
    }
    // pass 0: u64 host buffers, rank r owns clients {c : c % G == r}
    // pass 1: PACKED host buffers (bitlen(q_i) bits per residue) and shards from ppq_plan_shards for ranks that copy at
    //         different rates (contiguous blocks of counts[r] clients)
    const size_t pw = ppq_packed_words(ctx[0], (int)ell);
    std::vector<std::vector<uint64_t>> pk(C, std::vector<uint64_t>(pw * S));
    for (int c = 0; c < C; ++c) CK(ppq_pack_host(ctx[0], (int)ell, cts[c].data(), pk[c].data(), S));
    std::vector<double> cost(G);
    for (int r = 0; r < G; ++r) cost[r] = 1.0 + 0.6 * (r % 2);
    std::vector<size_t> counts(G);
    CK(ppq_plan_shards((size_t)C, G, cost.data(), counts.data()));
    bool all_same = true;
    for (int pass = 0; pass < 2; ++pass) {
        uint8_t id[128];
        CK(ppq_comm_unique_id(id));
        std::vector<std::thread> th;
        std::vector<int> rc(G, 0);
        std::vector<uint64_t> got_p(pw * S);
        std::fill(got.begin(), got.end(), 0);
        for (int r = 0; r < G; ++r)
            th.emplace_back([&, r]() {
                ppq_comm* comm = nullptr;
                if (ppq_comm_init(ctx[r], G, r, id, &comm) != PPQ_OK) {
                    std::fprintf(stderr, "rank %d: ppq_comm_init: %s\n", r, ppq_last_error());
                    rc[r] = 1;
                    return;
                }
                std::vector<int> mine;
                if (pass == 0) {
                    for (int c = r; c < C; c += G) mine.push_back(c);
                } else {
                    size_t first = 0;
                    for (int q = 0; q < r; ++q) first += counts[q];
                    for (size_t c = first; c < first + counts[r]; ++c) mine.push_back((int)c);
                }
                std::vector<ppq_rekey*> rk;
                std::vector<const uint64_t*> ptr;
                for (int c : mine) {
                    ppq_rekey* k = nullptr;
                    if (ppq_rekey_upload(ctx[r], kb[c].data(), ka[c].data(), &k) != PPQ_OK) rc[r] = 1;
                    rk.push_back(k);
                    ptr.push_back(pass == 0 ? cts[c].data() : pk[c].data());
                }
                const int e = pass == 0
                                  ? ppq_reencrypt_aggregate_host_ranks(ctx[r], comm, 0, rk.data(), (int)ell, ptr.data(), rk.size(), S,
                                                                       r == 0 ? got.data() : nullptr)
                                  : ppq_reencrypt_aggregate_host_packed_ranks(ctx[r], comm, 0, rk.data(), (int)ell, ptr.data(),
                                                                              rk.size(), S, r == 0 ? got_p.data() : nullptr);
                if (e != PPQ_OK) {
                    std::fprintf(stderr, "rank %d: %s\n", r, ppq_last_error());
                    rc[r] = 1;
                }
                for (auto* k : rk) ppq_rekey_destroy(k);
                ppq_comm_destroy(comm);
            });
        for (auto& t : th) t.join();
        for (int r = 0; r < G; ++r)
            if (rc[r]) return 1;
        if (pass == 1) CK(ppq_unpack_host(ctx[0], (int)ell, got_p.data(), got.data(), S));
        const bool ok = std::memcmp(want.data(), got.data(), blk * sizeof(uint64_t)) == 0;
        std::printf("multi_gpu_selftest: pass %d (%s): %s\n", pass,
                    pass == 0 ? "u64 buffers, round-robin shards" : "packed buffers, planned shards", ok ? "identical" : "MISMATCH");
        all_same = all_same && ok;
    }
    const bool same = all_same;
    std::printf("multi_gpu_selftest: %d GPUs, %d clients x %zu cts, N=2^%d, %zu limbs: %s\n", G, C, S, info.logN, ell,
                same ? "IDENTICAL to the single-GPU aggregate" : "MISMATCH");
    for (auto* c : ctx) ppq_ctx_destroy(c);
    return same ? 0 : 1;
}
