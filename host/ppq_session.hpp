// ppq_session.hpp -- what the four tools share: the context loaded from CC.json through the C ABI, device/pinned
// buffers, and the batching of every ciphertext blob of a container file into ONE device batch (the reference loops
// ciphertext by ciphertext, e.g. server/src/changeCipherDomain.cpp:61-115; the B200 path takes them all per call).
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <iostream>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <utility>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/ppq_b200.h"
#include "ppq_wire.hpp"

namespace ppqhost {

inline void cu(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}
inline void ck(int rc, const char* what) {
    if (rc != PPQ_OK) throw std::runtime_error(std::string(what) + ": " + ppq_last_error());
}

template <class T>
struct Dev {
    T* p = nullptr;
    size_t n = 0;
    Dev() {}
    explicit Dev(size_t count) { alloc(count); }
    Dev(const Dev&) = delete;
    Dev& operator=(const Dev&) = delete;
    ~Dev() { if (p) cudaFree(p); }
    void alloc(size_t count) {
        if (p) cudaFree(p);
        p = nullptr;
        n = count;
        if (count) cu(cudaMalloc(&p, count * sizeof(T)), "cudaMalloc");
    }
    void upload(const T* h, size_t count) { cu(cudaMemcpy(p, h, count * sizeof(T), cudaMemcpyHostToDevice), "H2D copy"); }
    void download(T* h, size_t count) const { cu(cudaMemcpy(h, p, count * sizeof(T), cudaMemcpyDeviceToHost), "D2H copy"); }
};

// Host memory of the ciphertext batches.  In the one-shot tools it is plain malloc.  A resident process
// (ppq_served) turns the pool on: blocks come from cudaHostAlloc (pinned: host<->device copies run at the PCIe rate, not
// through the driver's pageable staging) and go back to a free list instead of being released, so a steady-state round
// allocates nothing.
struct HostPool {
    bool pinned = false;
    std::mutex mu;
    std::multimap<size_t, void*> free_blocks;  // by capacity
    std::map<void*, size_t> live;              // pinned blocks handed out -> capacity
    static HostPool& get() {
        static HostPool p;
        return p;
    }
    void* take(size_t bytes) {
        if (!pinned) {
            void* q = std::malloc(bytes ? bytes : 1);
            if (!q) throw std::bad_alloc();
            return q;
        }
        std::lock_guard<std::mutex> lk(mu);
        auto it = free_blocks.lower_bound(bytes);
        if (it != free_blocks.end() && it->first <= 2 * bytes + (1u << 20)) {
            void* q = it->second;
            live[q] = it->first;
            free_blocks.erase(it);
            return q;
        }
        void* q = nullptr;
        const size_t cap = bytes ? bytes : 1;
        if (cudaHostAlloc(&q, cap, cudaHostAllocDefault) != cudaSuccess) {
            (void)cudaGetLastError();  // no pinned memory left: a pageable block works too, only slower to copy
            q = std::malloc(cap);
            if (!q) throw std::bad_alloc();
            return q;                  // not in `live`: give() frees it
        }
        live[q] = cap;
        return q;
    }
    void give(void* q) {
        if (!q) return;
        std::unique_lock<std::mutex> lk(mu);
        auto it = live.find(q);
        if (it == live.end()) {  // a malloc block (the pool was off when it was taken)
            lk.unlock();
            std::free(q);
            return;
        }
        free_blocks.emplace(it->second, q);
        live.erase(it);
    }
};
template <class T>
struct PoolAlloc {
    typedef T value_type;
    PoolAlloc() {}
    template <class U>
    PoolAlloc(const PoolAlloc<U>&) {}
    T* allocate(size_t n) { return static_cast<T*>(HostPool::get().take(n * sizeof(T))); }
    void deallocate(T* p, size_t) { HostPool::get().give(p); }
    // resize() leaves new residues uninitialised (every word is written by the decoder or a device->host copy)
    template <class U>
    void construct(U* p) { ::new (static_cast<void*>(p)) U; }
    template <class U, class... A>
    void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
    template <class U>
    bool operator==(const PoolAlloc<U>&) const { return true; }
    template <class U>
    bool operator!=(const PoolAlloc<U>&) const { return false; }
};
typedef std::vector<uint64_t, PoolAlloc<uint64_t>> HVec;

// grow-only device buffer for a resident process (Dev<T> allocates and frees per use)
template <class T>
struct DevGrow {
    T* p = nullptr;
    size_t cap = 0;
    DevGrow() {}
    DevGrow(const DevGrow&) = delete;
    DevGrow& operator=(const DevGrow&) = delete;
    ~DevGrow() { if (p) cudaFree(p); }
    T* ensure(size_t count) {
        if (count > cap) {
            if (p) cudaFree(p);
            p = nullptr;
            cap = 0;
            cu(cudaMalloc(&p, count * sizeof(T)), "cudaMalloc");
            cap = count;
        }
        return p;
    }
};

struct Session {
    ppq_ctx* cc = nullptr;
    ppq_ctx_info_t info{};
    size_t N = 0;
    int L = 0, K = 0, dnum = 0;
    std::vector<uint64_t> moduli;  // L + K
    std::vector<uint64_t> roots;   // L + K
    wire::Codec codec;             // ciphertext blobs <-> [2][ell][N]
    uint32_t batch_size = 0;       // EncodingParams "bs" of CC.json

    explicit Session(const std::string& cc_path) {
        const std::string text = wire::slurp(cc_path);
        int device = 0;
        if (const char* e = std::getenv("PPQ_DEVICE")) device = std::atoi(e);
        ck(ppq_ctx_create_from_ccjson(text.data(), text.size(), device, &cc), "ppq_ctx_create_from_ccjson");
        ck(ppq_ctx_info(cc, &info), "ppq_ctx_info");
        N = (size_t)1 << info.logN;
        L = info.L;
        K = info.K;
        dnum = info.dnum;
        moduli.resize((size_t)L + K);
        ck(ppq_ctx_moduli(cc, moduli.data()), "ppq_ctx_moduli");
        roots.resize((size_t)L + K);
        ck(ppq_ctx_roots(cc, roots.data()), "ppq_ctx_roots");
        codec.init(text, (uint32_t)N, moduli.data(), roots.data(), (size_t)L);
        // CC.json's EncodingParams batch size (client/src/encryptModelWeights.cpp:40 GetBatchSize): the chunking of
        // `values`.  Sparse packing (bs < N/2) is not implemented -- fail loudly instead of diverging silently.
        if (const jmin::Value* enp = wire::find_member_any(*codec.cx.cc_root, "enp"))
            if (const jmin::Value* bs = wire::find_member_any(*enp, "bs")) batch_size = (uint32_t)bs->as_u64();
        if (batch_size == 0) batch_size = (uint32_t)(N / 2);
        if (batch_size != N / 2)
            throw std::runtime_error("CC.json batch size " + std::to_string(batch_size) + " != N/2: sparse packing is not supported");
    }
    Session(const Session&) = delete;
    ~Session() { if (cc) ppq_ctx_destroy(cc); }

    // a DCRTPoly read from a key file -> [limbs][N] contiguous, checked against this context's basis
    std::vector<uint64_t> flatten(const wire::Poly& p, size_t want_limbs, const char* what) const {
        if (p.limbs.size() != want_limbs)
            throw std::runtime_error(std::string(what) + ": expected " + std::to_string(want_limbs) + " limbs, file has " +
                                     std::to_string(p.limbs.size()));
        if (p.format != 0) throw std::runtime_error(std::string(what) + ": element is not in EVALUATION format");
        std::vector<uint64_t> out(want_limbs * N);
        for (size_t i = 0; i < want_limbs; ++i) {
            if (p.moduli[i] != moduli[i]) throw std::runtime_error(std::string(what) + ": limb modulus does not match CC.json's basis");
            if (p.limbs[i].size() != N) throw std::runtime_error(std::string(what) + ": limb length is not the ring dimension");
            std::copy(p.limbs[i].begin(), p.limbs[i].end(), out.begin() + i * N);
        }
        return out;
    }
};

// every blob of a container, in file order, with a pointer back to where its replacement goes
struct BlobRef {
    std::string* slot;
};
inline std::vector<BlobRef> blob_refs(std::vector<wire::Layer>& layers) {
    std::vector<BlobRef> r;
    for (auto& l : layers) {
        r.push_back({&l.mean});
        r.push_back({&l.std_dev});
        for (auto& v : l.values) r.push_back({&v});
    }
    return r;
}

// A batch of ciphertexts with the same limb count, resident on the host as [n][2][ell][N].
struct Batch {
    uint32_t ell = 0;
    std::vector<size_t> index;         // position in the file-order blob list
    std::vector<wire::CtMeta> meta;
    HVec data;
    size_t n() const { return index.size(); }
};
// decode all blobs, grouped by limb count (one group in every file the reference flow produces)
inline std::map<uint32_t, Batch> decode_all(const Session& s, const std::vector<BlobRef>& refs) {
    // base64 + blob decoding is the host-side cost of a file (a 38 MB container holds 27 blobs): one blob per thread,
    // residues written straight into the batch.  Blobs of equal text length have equal limb counts, so the first blob of
    // each length is decoded alone to learn the shape, the batches are allocated (and touched) once by this thread, and
    // the parallel loop allocates nothing; a blob that turns out to have another shape sends the file down the
    // general path below.
    std::map<size_t, uint32_t> ell_of_len;
    std::vector<wire::CtMeta> metas(refs.size());
    std::vector<uint32_t> ells(refs.size(), 0);
    std::map<uint32_t, Batch> groups;
    bool fast = true;
    try {
        for (size_t i = 0; i < refs.size(); ++i) {
            const size_t len = refs[i].slot->size();
            if (!ell_of_len.count(len)) {
                uint64_t probe_cap = (size_t)2 * s.L * s.N;
                std::vector<uint64_t> probe(probe_cap);
                ell_of_len[len] = s.codec.decode(*refs[i].slot, probe.data(), probe_cap).ell;
            }
            ells[i] = ell_of_len[len];
            Batch& b = groups[ells[i]];
            b.ell = ells[i];
            b.index.push_back(i);
        }
    } catch (const std::exception&) {
        fast = false;  // the general path reports the error with its own message
    }
    if (fast) {
        std::vector<size_t> pos(refs.size());
        for (auto& kv : groups) {
            Batch& b = kv.second;
            b.data.resize(b.n() * (size_t)2 * b.ell * s.N);
            b.meta.resize(b.n());
            for (size_t k = 0; k < b.n(); ++k) pos[b.index[k]] = k;
        }
        int bad = 0;
#pragma omp parallel for schedule(dynamic) reduction(+ : bad)
        for (size_t i = 0; i < refs.size(); ++i) {
            try {
                Batch& b = groups[ells[i]];
                const size_t ct = (size_t)2 * b.ell * s.N;
                wire::Ct c = s.codec.decode(*refs[i].slot, b.data.data() + pos[i] * ct, ct);
                bool ok = c.ell == b.ell && c.N == s.N;
                for (uint32_t k = 0; ok && k < c.ell; ++k) ok = c.moduli[k] == s.moduli[k];
                if (!ok) ++bad;
                b.meta[pos[i]] = std::move(c.meta);
            } catch (const std::exception&) {
                ++bad;
            }
        }
        if (!bad) return groups;
    }
    groups.clear();
    std::vector<wire::Ct> cts(refs.size());
    std::vector<std::string> errs(refs.size());
#pragma omp parallel for schedule(dynamic)
    for (size_t i = 0; i < refs.size(); ++i) {
        try {
            cts[i] = s.codec.decode(*refs[i].slot);
        } catch (const std::exception& e) {
            errs[i] = e.what()[0] ? e.what() : "decode error";
        }
    }
    for (size_t i = 0; i < refs.size(); ++i) {
        if (!errs[i].empty()) throw std::runtime_error(errs[i]);
        wire::Ct& c = cts[i];
        if (c.N != s.N) throw std::runtime_error("ciphertext ring dimension differs from CC.json");
        if ((int)c.ell > s.L) throw std::runtime_error("ciphertext has more limbs than the context");
        for (uint32_t k = 0; k < c.ell; ++k)
            if (c.moduli[k] != s.moduli[k]) throw std::runtime_error("ciphertext limb modulus does not match CC.json's basis");
        Batch& b = groups[c.ell];
        b.ell = c.ell;
        b.index.push_back(i);
        b.meta.push_back(c.meta);
    }
    for (auto& kv : groups) {  // gather: [n][2][ell][N]
        Batch& b = kv.second;
        const size_t ct = (size_t)2 * b.ell * s.N;
        b.data.resize(b.n() * ct);
#pragma omp parallel for schedule(static)
        for (size_t k = 0; k < b.n(); ++k) std::memcpy(b.data.data() + k * ct, cts[b.index[k]].data.data(), ct * sizeof(uint64_t));
    }
    return groups;
}
inline void encode_back(const Session& s, const Batch& b, uint32_t ell_out, const std::vector<wire::CtMeta>& meta,
                        const uint64_t* data, const std::vector<BlobRef>& refs) {
    const size_t ct = (size_t)2 * ell_out * s.N;
    s.codec.warn_once();
    // each new blob overwrites the text of the blob it replaces (same or shorter), so nothing is allocated in the loop
#pragma omp parallel for schedule(dynamic)
    for (size_t k = 0; k < b.n(); ++k) s.codec.encode_to(*refs[b.index[k]].slot, meta[k], ell_out, data + k * ct);
}

// CryptoObject::keyTag of a new key pair: 32 hex digits (OpenFHE's GenerateUniqueKeyID writes 128 random bits this way)
inline std::string new_key_tag() {
    unsigned char b[16];
    if (const char* e = std::getenv("PPQ_SEED")) {
        const uint64_t v = std::strtoull(e, nullptr, 0) * 0x9E3779B97F4A7C15ull + 1;
        std::memcpy(b, &v, 8);
        std::memcpy(b + 8, &v, 8);
    } else {
        FILE* f = std::fopen("/dev/urandom", "rb");
        if (!f || std::fread(b, 1, 16, f) != 16) throw std::runtime_error("cannot read /dev/urandom");
        std::fclose(f);
    }
    static const char* hx = "0123456789abcdef";
    std::string t;
    for (unsigned char c : b) {
        t += hx[c >> 4];
        t += hx[c & 15];
    }
    return t;
}
// Key material for the device samplers.  Production: NULL, i.e. the library draws a fresh 256-bit ChaCha20 key from
// getrandom(2) for every call.  PPQ_SEED (an integer) is a TEST-ONLY switch for reproducible runs: it is widened to
// 32 bytes, `salt` separating the calls of one process.  Never set it when generating real keys or ciphertexts.
struct SamplerSeed {
    uint8_t bytes[32];
    bool fixed = false;
    explicit SamplerSeed(uint64_t salt = 0) {
        std::memset(bytes, 0, sizeof(bytes));
        if (const char* e = std::getenv("PPQ_SEED")) {
            const uint64_t v = std::strtoull(e, nullptr, 0);
            std::memcpy(bytes, &v, 8);
            std::memcpy(bytes + 8, &salt, 8);
            fixed = true;
        }
    }
    const uint8_t* ptr() const { return fixed ? bytes : nullptr; }
};

}  // namespace ppqhost
