// capi.cu -- the C ABI of libppq_b200.so (include/ppq_b200.h): context/key/scratch management and the
// orchestration of kernel launches for each reference-facing operation.  No CPU fallback: every compute entry
// point needs a CUDA device and fails loudly otherwise.
#include <dlfcn.h>
#include <sys/random.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ppq_b200.h"
#include "ckks.cuh"
#include "kernels.cuh"
#include "params.hpp"
#include "sampler.cuh"

using namespace ppq;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return fail(PPQ_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));               \
    } while (0)

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    cudaError_t upload(const std::vector<T>& h) {
        release();
        n = h.size();
        if (!n) return cudaSuccess;
        cudaError_t e = cudaMalloc(&p, n * sizeof(T));
        if (e != cudaSuccess) return e;
        return cudaMemcpy(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice);
    }
    cudaError_t ensure(size_t want) {
        if (want <= n) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) n = want;
        return e;
    }
};

static std::vector<BcJob> to_dev_jobs(const std::vector<host::Job>& j) {
    std::vector<BcJob> o(j.size());
    for (size_t i = 0; i < j.size(); ++i) o[i] = BcJob{j[i].src_first, j[i].ns, j[i].dst_slot, j[i].limb, j[i].chat_off, {0, 0, 0}};
    return o;
}

struct KsPlanDev {
    host::KsPlan h;
    struct Digit { int job_first, n_jobs, ns; };
    std::vector<Digit> digits;  // ModUp jobs grouped by source digit
    DevBuf<int> dig_ids;        // [parts*(ell+K)] limb of every digit slot, -1 for the digit's own limbs
    DevBuf<int> conv_ids;       // [2*ell] limb of every ModDown slot
    DevBuf<int> c1_limb, p_limb;
    DevBuf<u64> c1_scale, c1_scale_sh, p_scale, p_scale_sh, pinv, pinv_sh;
    std::vector<BcJob> up_jobs, down_jobs;  // host: packed into kernel parameters per launch (launch_bconv)
};
struct RsPlanDev {
    host::RsPlan h;
    DevBuf<int> last_limb;
    DevBuf<u64> last_scale, last_scale_sh, cst, cst_sh;
    DevBuf<BcJob> jobs;
};

struct ppq_ctx {
    host::Params hp;
    int device = 0;
    KCtx k{};
    DevBuf<LimbC> limbs;
    std::vector<LimbC> h_limbs;
    DevBuf<ulonglong2> twf, twi;
    std::map<int, std::unique_ptr<KsPlanDev>> ks;
    std::map<int, std::unique_ptr<RsPlanDev>> rs;
    DevBuf<u64> key_pinv, key_pinv_sh;  // [L] P^-1 mod q_i (+ Shoup), used once per re-encryption key
    DevBuf<u64> scratch;   // key-switch wave scratch
    DevBuf<u64> scratch2;  // rescale / encrypt / decrypt scratch
    DevBuf<int> ids;       // limb ids of the plain NTT API
    DevBuf<u64> ids_scale; // [n][4] output factor N^-1 per poly (host::scale4)
    std::vector<u64> ntt_scale4;  // [L+K][4], built on first use
    std::map<int, std::unique_ptr<DevBuf<int>>> cyc;    // limb ids 0..ell-1 repeated (65536 entries) per ell
    std::map<int, std::unique_ptr<DevBuf<u64>>> cyc_scale;  // [65536][4] inverse-NTT output factors in the same order
    std::map<int, std::unique_ptr<DevBuf<u64>>> garner;  // per ell: [ell] Garner inverses then [ell] partial products as doubles
    DevBuf<u64> pmod;      // [L] P mod q_r (ReKeyGen)
    DevBuf<int> kinds3;    // {ternary, gauss, gauss}: the polynomial kinds of one (v, e0, e1) / (u, e0, e1) triple
    ckks::Tables fft;
    int wave = 0;
    // per-stage CUDA-event profiling of the key-switch pipeline (ppq_profile_*): a chain of events on the launching
    // stream; the time between two consecutive marks is attributed to the stage named by the later one
    static const int NSTAGE = 11;
    bool prof = false;
    struct Mark { int stage; cudaEvent_t ev; };  // stage -1 starts a chain
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    std::vector<Mark> marks;
    double prof_ms[NSTAGE] = {};
    unsigned long long prof_launches[NSTAGE] = {};
    struct HostPath;
    HostPath* hostpath = nullptr;  // persistent buffers / streams of ppq_reencrypt_aggregate_host
    ~ppq_ctx();
};
struct ppq_rekey {
    ppq_ctx* ctx = nullptr;
    int device = 0;  // kept here so that a key may be destroyed after its context
    const u64* b = nullptr;
    const u64* a = nullptr;
    // Q limbs multiplied by P^-1 mod q_i (see launch_scale_key): in place for an uploaded key (stride L+K), a private
    // [dnum][L][N] copy for a borrowed one (the caller's buffers are never written)
    const u64* qb = nullptr;
    const u64* qa = nullptr;
    int q_stride = 0;
    DevBuf<u64> own_b, own_a, own_qb, own_qa;
};

extern "C" const char* ppq_last_error(void) { return g_err.c_str(); }
extern "C" const char* ppq_version(void) { return "ppq_b200 0.1 (sm_100a)"; }

static int ctx_finish(std::unique_ptr<ppq_ctx>& c, int device, ppq_ctx** out) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(PPQ_ECUDA, "no CUDA device: libppq_b200 has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(PPQ_EINVAL, "bad device ordinal");
    CU(cudaSetDevice(device));
    c->device = device;
    const host::Params& hp = c->hp;
    const int M = hp.M();
    const size_t N = hp.N;
    std::vector<LimbC> lc(M);
    for (int r = 0; r < M; ++r) {
        const u64 q = hp.mod[r];
        const int bl = 64 - __builtin_clzll(q);
        lc[r].q = q;
        lc[r].q2 = 2 * q;
        lc[r].sh = (u32)(bl - 1);
        lc[r].mu = (u64)((((unsigned __int128)1) << (64 + bl - 1)) / q);
        lc[r].pad = 0;
    }
    CU(c->limbs.upload(lc));
    c->h_limbs = lc;
    {
        std::vector<ulonglong2> f((size_t)M * N), i((size_t)M * N);
#pragma omp parallel for
        for (int r = 0; r < M; ++r)
            host::build_twiddles(hp, r, reinterpret_cast<host::u64*>(f.data() + (size_t)r * N),
                                 reinterpret_cast<host::u64*>(i.data() + (size_t)r * N));
        CU(c->twf.upload(f));
        CU(c->twi.upload(i));
    }
    c->k.limbs = c->limbs.p;
    c->k.twf = c->twf.p;
    c->k.twi = c->twi.p;
    c->k.logN = hp.logN;
    std::string err;
    if (!ckks::build_tables(hp.logN, c->fft, err)) return fail(PPQ_ECUDA, err);
    CU(c->kinds3.upload(std::vector<int>{0, 1, 1}));
    CU(sampler::upload_gauss_table(device, 3.19));
    *out = c.release();
    return PPQ_OK;
}

extern "C" int ppq_ctx_create(int logN, int L, const uint64_t* q, int dnum, int aux_bits, int extra_bits, int device,
                              ppq_ctx** out) {
    if (!q || !out || L < 1) return fail(PPQ_EINVAL, "ppq_ctx_create: bad arguments");
    std::unique_ptr<ppq_ctx> c(new ppq_ctx);
    c->hp.logN = logN;
    c->hp.L = L;
    c->hp.dnum = dnum;
    c->hp.aux_bits = aux_bits;
    c->hp.extra_bits = extra_bits;
    c->hp.mod.assign(q, q + L);
    std::string err;
    if (!host::derive(c->hp, err)) return fail(PPQ_EINVAL, err);
    return ctx_finish(c, device, out);
}
extern "C" int ppq_ctx_create_from_ccjson(const char* json, size_t len, int device, ppq_ctx** out) {
    if (!json || !out) return fail(PPQ_EINVAL, "ppq_ctx_create_from_ccjson: bad arguments");
    std::unique_ptr<ppq_ctx> c(new ppq_ctx);
    std::string err;
    if (!host::parse_ccjson(json, len, c->hp, err)) return fail(PPQ_EPARSE, err);
    return ctx_finish(c, device, out);
}
extern "C" int ppq_ctx_create_synthetic(int logN, int depth, int first_bits, int scale_bits, int extra_bits, int dnum,
                                        int device, ppq_ctx** out) {
    if (!out || depth < 1 || logN < 12 || logN > 17) return fail(PPQ_EINVAL, "ppq_ctx_create_synthetic: bad arguments");
    std::vector<host::u64> q = host::synthetic_chain(logN, depth, first_bits, scale_bits, extra_bits);
    return ppq_ctx_create(logN, (int)q.size(), reinterpret_cast<const uint64_t*>(q.data()), dnum, 60, extra_bits, device, out);
}
extern "C" void ppq_ctx_destroy(ppq_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    delete c;
}
extern "C" int ppq_ctx_info(const ppq_ctx* c, ppq_ctx_info_t* o) {
    if (!c || !o) return fail(PPQ_EINVAL, "ppq_ctx_info: null");
    o->logN = c->hp.logN;
    o->L = c->hp.L;
    o->K = c->hp.K;
    o->dnum = c->hp.dnum;
    o->alpha = c->hp.alpha;
    o->device = c->device;
    return PPQ_OK;
}
extern "C" int ppq_ctx_moduli(const ppq_ctx* c, uint64_t* o) {
    if (!c || !o) return fail(PPQ_EINVAL, "null");
    std::memcpy(o, c->hp.mod.data(), sizeof(u64) * c->hp.M());
    return PPQ_OK;
}
extern "C" int ppq_ctx_roots(const ppq_ctx* c, uint64_t* o) {
    if (!c || !o) return fail(PPQ_EINVAL, "null");
    std::memcpy(o, c->hp.root.data(), sizeof(u64) * c->hp.M());
    return PPQ_OK;
}
extern "C" double ppq_ctx_scaling_factor(const ppq_ctx* c, int level) {
    return (c && level >= 0 && level < c->hp.L) ? c->hp.sf[level] : 0.0;
}
extern "C" double ppq_ctx_scaling_factor_big(const ppq_ctx* c, int level) {
    return (c && level >= 0 && level < c->hp.L) ? c->hp.sfbig[level] : 0.0;
}
extern "C" int ppq_ctx_num_parts(const ppq_ctx* c, int ell) {
    if (!c || ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "bad ell");
    int p = (ell + c->hp.alpha - 1) / c->hp.alpha;
    return p > c->hp.dnum ? c->hp.dnum : p;
}
extern "C" int ppq_sync(const ppq_ctx* c, void* stream) {
    if (!c) return fail(PPQ_EINVAL, "null");
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    return PPQ_OK;
}
extern "C" int ppq_ctx_set_wave(ppq_ctx* c, int n) {
    if (!c || n < 0) return fail(PPQ_EINVAL, "bad wave");
    c->wave = n;
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- plain NTT
static int ntt_common(ppq_ctx* c, u64* d, const int32_t* ids, size_t n, void* stream, bool inverse) {
    if (!c || !d || !ids) return fail(PPQ_EINVAL, "ppq_ntt: null argument");
    if (n == 0) return PPQ_OK;
    if (n > 65535) {  // walked in chunks: the grid limit is not the caller's problem
        for (size_t s0 = 0; s0 < n; s0 += 32768)
            if (int rc = ntt_common(c, d + s0 * c->hp.N, ids + s0, std::min<size_t>(32768, n - s0), stream, inverse)) return rc;
        return PPQ_OK;
    }
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int M = c->hp.M();
    std::vector<int> h(ids, ids + n);
    for (size_t i = 0; i < n; ++i)
        if (h[i] < 0 || h[i] >= M) return fail(PPQ_EINVAL, "ppq_ntt: limb id out of range");
    std::vector<u64> sc;
    if (inverse) {
        if (c->ntt_scale4.empty()) {  // per limb, once: (N^-1, shoup, N^-1 psi^-brev(1), shoup)
            std::vector<int> all(M);
            for (int r = 0; r < M; ++r) all[r] = r;
            c->ntt_scale4 = host::scale4(c->hp, all, c->hp.ninv);
        }
        sc.resize(4 * n);
        for (size_t i = 0; i < n; ++i)
            for (int k = 0; k < 4; ++k) sc[4 * i + k] = c->ntt_scale4[4 * h[i] + k];
    }
    CU(c->ids.ensure(n));
    CU(cudaMemcpyAsync(c->ids.p, h.data(), n * sizeof(int), cudaMemcpyHostToDevice, st));
    if (inverse) {
        CU(c->ids_scale.ensure(4 * n));
        CU(cudaMemcpyAsync(c->ids_scale.p, sc.data(), 4 * n * sizeof(u64), cudaMemcpyHostToDevice, st));
        launch_ntt_inverse(c->k, d, c->ids.p, c->ids_scale.p, (int)n, st);
    } else {
        launch_ntt_forward(c->k, d, c->ids.p, (int)n, st);
    }
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));  // h / sc are stack-owned staging for the async copies
    return PPQ_OK;
}
extern "C" int ppq_ntt_forward(ppq_ctx* c, uint64_t* d, const int32_t* ids, size_t n, void* stream) {
    return ntt_common(c, (u64*)d, ids, n, stream, false);
}
extern "C" int ppq_ntt_inverse(ppq_ctx* c, uint64_t* d, const int32_t* ids, size_t n, void* stream) {
    return ntt_common(c, (u64*)d, ids, n, stream, true);
}

// ------------------------------------------------------------------------------------------- rekeys
// P^-1 mod q_i for every Q limb, and the scaled Q part of a key (ModDown without a multiplication, kernels.cuh)
static int rekey_scale_q(ppq_ctx* c, ppq_rekey* k, bool in_place) {
    const host::Params& hp = c->hp;
    const int L = hp.L, M = hp.M();
    if (!c->key_pinv.p) {
        host::KsPlan plan;
        host::build_ks_plan(hp, L, plan);
        CU(c->key_pinv.upload(plan.pinv));
        CU(c->key_pinv_sh.upload(plan.pinv_sh));
    }
    u64 *qb, *qa;
    if (in_place) {
        qb = k->own_b.p;
        qa = k->own_a.p;
        k->q_stride = M;
    } else {
        const size_t n = (size_t)hp.dnum * L * hp.N;
        CU(k->own_qb.ensure(n));
        CU(k->own_qa.ensure(n));
        qb = k->own_qb.p;
        qa = k->own_qa.p;
        k->q_stride = L;
    }
    // legacy default stream: ordered after whatever produced a borrowed key on a blocking stream
    launch_scale_key(c->k, k->b, M, qb, k->q_stride, c->key_pinv.p, c->key_pinv_sh.p, L, hp.dnum, 0);
    launch_scale_key(c->k, k->a, M, qa, k->q_stride, c->key_pinv.p, c->key_pinv_sh.p, L, hp.dnum, 0);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(0));
    k->qb = qb;
    k->qa = qa;
    return PPQ_OK;
}

extern "C" int ppq_rekey_upload(ppq_ctx* c, const uint64_t* hb, const uint64_t* ha, ppq_rekey** out) {
    if (!c || !hb || !ha || !out) return fail(PPQ_EINVAL, "ppq_rekey_upload: null argument");
    CU(cudaSetDevice(c->device));
    std::unique_ptr<ppq_rekey> k(new ppq_rekey);
    const size_t n = (size_t)c->hp.dnum * c->hp.M() * c->hp.N;
    CU(k->own_b.ensure(n));
    CU(k->own_a.ensure(n));
    CU(cudaMemcpy(k->own_b.p, hb, n * sizeof(u64), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(k->own_a.p, ha, n * sizeof(u64), cudaMemcpyHostToDevice));
    k->ctx = c;
    k->device = c->device;
    k->b = k->own_b.p;
    k->a = k->own_a.p;
    int rc = rekey_scale_q(c, k.get(), true);
    if (rc != PPQ_OK) return rc;
    *out = k.release();
    return PPQ_OK;
}
extern "C" int ppq_rekey_from_device(ppq_ctx* c, const uint64_t* db, const uint64_t* da, ppq_rekey** out) {
    if (!c || !db || !da || !out) return fail(PPQ_EINVAL, "ppq_rekey_from_device: null argument");
    CU(cudaSetDevice(c->device));
    std::unique_ptr<ppq_rekey> k(new ppq_rekey);
    k->ctx = c;
    k->device = c->device;
    k->b = (const u64*)db;
    k->a = (const u64*)da;
    int rc = rekey_scale_q(c, k.get(), false);
    if (rc != PPQ_OK) return rc;
    *out = k.release();
    return PPQ_OK;
}
extern "C" void ppq_rekey_destroy(ppq_rekey* k) {
    if (!k) return;
    cudaSetDevice(k->device);
    delete k;
}

// ------------------------------------------------------------------------------------------- key switch
static int get_ks_plan(ppq_ctx* c, int ell, KsPlanDev** out) {
    auto it = c->ks.find(ell);
    if (it == c->ks.end()) {
        std::unique_ptr<KsPlanDev> p(new KsPlanDev);
        host::build_ks_plan(c->hp, ell, p->h);
        CU(p->c1_limb.upload(p->h.c1_limb));
        CU(p->c1_scale.upload(host::scale4(c->hp, p->h.c1_limb, p->h.c1_scale)));
        CU(p->p_limb.upload(p->h.p_limb));
        CU(p->p_scale.upload(host::scale4(c->hp, p->h.p_limb, p->h.p_scale)));
        CU(p->pinv.upload(p->h.pinv));
        CU(p->pinv_sh.upload(p->h.pinv_sh));
        p->up_jobs = to_dev_jobs(p->h.up_jobs);
        p->down_jobs = to_dev_jobs(p->h.down_jobs);
        {
            const int M = ell + c->hp.K;
            std::vector<int> ids((size_t)p->h.parts * M, -1), cids(2 * (size_t)ell);
            for (size_t i = 0; i < p->h.up_jobs.size(); ++i) {
                const host::Job& jb = p->h.up_jobs[i];
                ids[jb.dst_slot] = jb.limb;
                if (p->digits.empty() || p->h.up_jobs[p->digits.back().job_first].src_first != jb.src_first)
                    p->digits.push_back({(int)i, 0, jb.ns});
                p->digits.back().n_jobs++;
            }
            for (int i = 0; i < 2 * ell; ++i) cids[i] = i % ell;
            CU(p->dig_ids.upload(ids));
            CU(p->conv_ids.upload(cids));
        }
        it = c->ks.emplace(ell, std::move(p)).first;
    }
    *out = it->second.get();
    return PPQ_OK;
}

// limb ids (0, 1, .., ell-1) repeated: the id table of every batched transform whose polynomials cycle through the first
// `ell` limbs.  Built once per ell (the only synchronising step), then every call that needs it is asynchronous.
static const int CYC_LEN = 65536;
static int cyc_ids(ppq_ctx* c, int ell, const int** out) {
    auto it = c->cyc.find(ell);
    if (it == c->cyc.end()) {
        std::unique_ptr<DevBuf<int>> b(new DevBuf<int>);
        std::vector<int> h(CYC_LEN);
        for (int i = 0; i < CYC_LEN; ++i) h[i] = i % ell;
        CU(b->upload(h));
        it = c->cyc.emplace(ell, std::move(b)).first;
    }
    *out = it->second->p;
    return PPQ_OK;
}

struct WaveLayout {
    size_t y, dig, extq, extp, conv, per_ct;  // sizes in u64 per ciphertext
};
// keep_digits: the digits outlive the ModDown of one key (fan-out to several keys), so the ModDown conversion gets
// its own buffer instead of reusing theirs
static WaveLayout wave_layout(const ppq_ctx* c, int ell, int parts, bool keep_digits) {
    const size_t N = c->hp.N, K = c->hp.K, M = ell + K;
    WaveLayout w;
    w.y = (size_t)ell * N;
    w.dig = std::max((size_t)parts * M, (size_t)2 * ell) * N;  // digits, later reused for the ModDown conversion
    w.extq = (size_t)2 * ell * N;
    w.extp = (size_t)2 * K * N;
    w.conv = keep_digits ? (size_t)2 * ell * N : 0;
    w.per_ct = w.y + w.dig + w.extq + w.extp + w.conv;
    return w;
}
static int pick_wave(const ppq_ctx* c, const WaveLayout& w, size_t n) {
    size_t g;
    if (c->wave > 0) {
        g = (size_t)c->wave;
    } else {
        const char* e = std::getenv("PPQ_WAVE");
        if (e && std::atoi(e) > 0) {
            g = (size_t)std::atoi(e);
        } else {
            // Measured on B200 (profiles/r01_notes.md): the pipeline is integer-issue bound, not HBM bound, so large
            // waves (fewer, fuller launches, key tiles reused over more ciphertexts) beat L2-resident small ones:
            // C3 13 / 26 / 54 ciphertexts per wave = 5676 / 5862 / 6038 ct/s.  6 GiB of scratch (of 180 GB).
            const size_t budget = (size_t)6 << 30;
            g = budget / (w.per_ct * sizeof(u64));
            if (g < 1) g = 1;
        }
    }
    if (g > n) g = n;
    if (g > 32768) g = 32768;
    return (int)g;
}

static void prof_collect(ppq_ctx* c) {
    if (c->marks.empty()) return;
    cudaEventSynchronize(c->marks.back().ev);
    for (size_t i = 1; i < c->marks.size(); ++i) {
        const int s = c->marks[i].stage;
        if (s < 0) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c->marks[i - 1].ev, c->marks[i].ev) == cudaSuccess) {
            c->prof_ms[s] += ms;
            c->prof_launches[s] += 1;
        }
    }
    c->marks.clear();
    c->ev_used = 0;
}
// stage < 0: start of a chain (before the first launch of a call); stage >= 0: the launches since the previous mark
static void prof_mark(ppq_ctx* c, cudaStream_t st, int stage) {
    if (!c->prof) return;
    if (c->ev_used >= 8192) prof_collect(c);
    if (c->ev_used == c->ev_pool.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        c->ev_pool.push_back(e);
    }
    cudaEvent_t e = c->ev_pool[c->ev_used++];
    cudaEventRecord(e, st);
    c->marks.push_back({stage, e});
}
extern "C" int ppq_profile_enable(ppq_ctx* c, int on) {
    if (!c) return fail(PPQ_EINVAL, "null");
    CU(cudaSetDevice(c->device));
    prof_collect(c);
    c->prof = on != 0;
    return PPQ_OK;
}
extern "C" int ppq_profile_read(ppq_ctx* c, double* ms, uint64_t* launches, int reset) {
    if (!c || !ms || !launches) return fail(PPQ_EINVAL, "null");
    CU(cudaSetDevice(c->device));
    prof_collect(c);
    for (int s = 0; s < ppq_ctx::NSTAGE; ++s) {
        ms[s] = c->prof_ms[s];
        launches[s] = c->prof_launches[s];
        if (reset) {
            c->prof_ms[s] = 0;
            c->prof_launches[s] = 0;
        }
    }
    return PPQ_OK;
}
extern "C" uint64_t ppq_launch_count(void) { return (uint64_t)kernels_launch_count(); }
extern "C" const char* ppq_profile_stage_name(int s) {
    static const char* names[ppq_ctx::NSTAGE] = {"k_inv_row(c1)",      "k_inv_col(c1,scale)", "k_bconv(modup)",
                                                 "k_fwd_col(modup)",   "k_fwd_row(modup)",    "k_ip_tma|k_ip_agg",
                                                 "k_inv_row(P)",       "k_inv_col(P,scale)",  "k_bconv(moddown)",
                                                 "k_fwd_col(moddown)", "k_fwd_row_moddown"};
    // (ppq_agg_*: stage 8 is k_bconv_wide once per group of clients, stages 9-10 run once per ppq_agg_finish)
    return (s >= 0 && s < ppq_ctx::NSTAGE) ? names[s] : "";
}

// shared driver: out[g] = (accumulate ? out[g] : 0) + ReEncrypt(in[g])
// digits may reach k_ip_tma in [0, 2q) when a sum of `parts` products of a 2q-bounded digit and a key residue stays
// inside the Barrett range (kernels.cu, PPQ_DIGITS_LAZY)
#ifndef PPQ_DIGITS_LAZY
#define PPQ_DIGITS_LAZY 1
#endif
#define PPQ_DIGITS_LAZY_OK(parts) (PPQ_DIGITS_LAZY && (parts) <= 4)
// ReEncrypt of n ciphertexts under n_keys re-encryption keys: outs[k] = ReEncrypt(in, rks[k]).  The digit
// decomposition + ModUp of a ciphertext (steps 1-3b and the forward ROW pass of the digits) does not depend on the
// key, so with several keys (the fan-out of the aggregate) it is done once per wave and only the
// inner product and the ModDown are repeated per key.
static int reencrypt_multi(ppq_ctx* c, const ppq_rekey* const* rks, int n_keys, int ell, const u64* in, u64* const* outs,
                           size_t n, void* stream, int accumulate) {
    if (!c || !rks || !in || !outs || n_keys < 1) return fail(PPQ_EINVAL, "ppq_reencrypt: null argument");
    for (int k = 0; k < n_keys; ++k) {
        if (!rks[k] || !outs[k]) return fail(PPQ_EINVAL, "ppq_reencrypt: null argument");
        if (rks[k]->ctx != c) return fail(PPQ_EINVAL, "ppq_reencrypt: rekey belongs to another context");
    }
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_reencrypt: bad limb count");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    KsPlanDev* p;
    int rc = get_ks_plan(c, ell, &p);
    if (rc) return rc;
    const size_t N = c->hp.N;
    const int K = c->hp.K, L = c->hp.L, parts = p->h.parts;
    const bool keep = n_keys > 1;
    const WaveLayout w = wave_layout(c, ell, parts, keep);
    const int G = pick_wave(c, w, n);
    CU(c->scratch.ensure(w.per_ct * G));
    u64* y = c->scratch.p;
    u64* dig = y + w.y * G;
    u64* extq = dig + w.dig * G;
    u64* extp = extq + w.extq * G;
    u64* conv = keep ? extp + w.extp * G : dig;
    const size_t conv_gs = keep ? w.conv : w.dig;
    const size_t ct = (size_t)2 * ell * N;
    for (size_t g0 = 0; g0 < n; g0 += G) {
        const int ng = (int)std::min((size_t)G, n - g0);
        const u64* in_g = in + g0 * ct;
        prof_mark(c, st, -1);
        // 1-2. c1 -> COEFF, scaled by N^-1 (D_j/q_i)^-1
        launch_inv_row(c->k, in_g + (size_t)ell * N, ct, y, w.y, p->c1_limb.p, ell, ng, st);
        prof_mark(c, st, 0);
        launch_inv_col(c->k, y, w.y, p->c1_limb.p, p->c1_scale.p, ell, ng, st);
        prof_mark(c, st, 1);
        // 3. ModUp: approximate basis switch of every digit to its complementary limbs (pointwise)
        for (const auto& d : p->digits)
            if (!launch_bconv(c->k, y, w.y, dig, w.dig, p->up_jobs.data() + d.job_first, d.n_jobs, p->h.up_chat.data(), c->h_limbs.data(), d.ns, ng, st))
                return fail(PPQ_EINVAL, "digit wider than 32 limbs");
        prof_mark(c, st, 2);
        // 3b. forward COL pass of the converted limbs (the digit's own limbs stay as they are in c1)
        launch_fwd_col_plain(c->k, dig, w.dig, p->dig_ids.p, parts * (ell + K), ng, st);
        prof_mark(c, st, 3);
        // 4. forward ROW pass + inner product with the rekey (+ inverse ROW pass on the P limbs)
        launch_fwd_row_plain(c->k, dig, w.dig, p->dig_ids.p, parts * (ell + K), ng, st, PPQ_DIGITS_LAZY_OK(parts));
        prof_mark(c, st, 4);
        for (int key = 0; key < n_keys; ++key) {
        const ppq_rekey* rk = rks[key];
        u64* out_g = outs[key] + g0 * ct;
        if (key > 0) prof_mark(c, st, -1);
        IpArgs ia;
        ia.c1 = in_g + (size_t)ell * N;
        ia.c1_gs = ct;
        ia.digits = dig;
        ia.dig_gs = w.dig;
        ia.key_b = rk->b;
        ia.key_a = rk->a;
        ia.key_qb = rk->qb;
        ia.key_qa = rk->qa;
        ia.key_q_stride = rk->q_stride;
        ia.extq = extq;
        ia.extq_gs = w.extq;
        ia.extp = extp;
        ia.extp_gs = w.extp;
        ia.ell = ell;
        ia.K = K;
        ia.L = L;
        ia.parts = parts;
        ia.alpha = c->hp.alpha;
        if (!launch_ip_tma(c->k, ia, ng, st)) return fail(PPQ_EINVAL, "ppq_reencrypt: the inner product supports at most 8 digits");
        prof_mark(c, st, 5);
        launch_inv_row(c->k, extp, w.extp, extp, w.extp, p->p_limb.p, 2 * K, ng, st);
        prof_mark(c, st, 6);
        // 5. P limbs: inverse COL pass, scaled by N^-1 (P/p_k)^-1
        launch_inv_col(c->k, extp, w.extp, p->p_limb.p, p->p_scale.p, 2 * K, ng, st);
        prof_mark(c, st, 7);
        // 6. ModDown conversion P -> Q (pointwise; with one key the digits buffer is dead and reused), then its COL pass
        if (!launch_bconv(c->k, extp, w.extp, conv, conv_gs, p->down_jobs.data(), ell, p->h.down_chat.data(), c->h_limbs.data(), K, ng, st))
            return fail(PPQ_EINVAL, "more than 32 auxiliary limbs");
        if (!launch_bconv(c->k, extp, w.extp, conv, conv_gs, p->down_jobs.data() + ell, ell, p->h.down_chat.data(), c->h_limbs.data(), K, ng, st))
            return fail(PPQ_EINVAL, "more than 32 auxiliary limbs");
        prof_mark(c, st, 8);
        launch_fwd_col_plain(c->k, conv, conv_gs, p->conv_ids.p, 2 * ell, ng, st);
        prof_mark(c, st, 9);
        // 7. forward ROW pass + (x - conv) P^-1 + c0 (+ running sum)
        MdArgs ma;
        ma.conv = conv;
        ma.conv_gs = conv_gs;
        ma.extq = extq;
        ma.extq_gs = w.extq;
        ma.ct_in = in_g;
        ma.in_gs = ct;
        ma.out = out_g;
        ma.out_gs = ct;
        ma.ell = ell;
        ma.accumulate = accumulate;
        launch_fwd_row_moddown(c->k, ma, ng, st);
        prof_mark(c, st, 10);
        }  // keys
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}
static int reencrypt_common(ppq_ctx* c, const ppq_rekey* rk, int ell, const u64* in, u64* out, size_t n, void* stream,
                            int accumulate) {
    if (!rk || !out) return fail(PPQ_EINVAL, "ppq_reencrypt: null argument");
    return reencrypt_multi(c, &rk, 1, ell, in, &out, n, stream, accumulate);
}
extern "C" int ppq_reencrypt(ppq_ctx* c, const ppq_rekey* rk, int ell, const uint64_t* in, uint64_t* out, size_t n,
                             void* stream) {
    if (in == out) return fail(PPQ_EINVAL, "ppq_reencrypt: d_out may not alias d_in");
    return reencrypt_common(c, rk, ell, (const u64*)in, (u64*)out, n, stream, 0);
}
extern "C" int ppq_reencrypt_fanout(ppq_ctx* c, const ppq_rekey* const* rekeys, int n_keys, int ell, const uint64_t* in,
                                    uint64_t* const* outs, size_t n, void* stream) {
    if (!rekeys || !outs || n_keys < 1) return fail(PPQ_EINVAL, "ppq_reencrypt_fanout: null argument");
    for (int k = 0; k < n_keys; ++k) {
        if (outs[k] == in) return fail(PPQ_EINVAL, "ppq_reencrypt_fanout: an output may not alias d_in");
        for (int j = 0; j < k; ++j)
            if (outs[k] == outs[j]) return fail(PPQ_EINVAL, "ppq_reencrypt_fanout: outputs may not alias each other");
    }
    return reencrypt_multi(c, rekeys, n_keys, ell, (const u64*)in, (u64* const*)outs, n, stream, 0);
}
extern "C" int ppq_reencrypt_accumulate(ppq_ctx* c, const ppq_rekey* rk, int ell, const uint64_t* in, uint64_t* acc,
                                        size_t n, void* stream) {
    if (in == acc) return fail(PPQ_EINVAL, "ppq_reencrypt_accumulate: d_acc may not alias d_in");
    return reencrypt_common(c, rk, ell, (const u64*)in, (u64*)acc, n, stream, 1);
}

extern "C" int ppq_ks_modup(ppq_ctx* c, int ell, const uint64_t* d_c1, uint64_t* d_digits, size_t n, void* stream) {
    if (!c || !d_c1 || !d_digits) return fail(PPQ_EINVAL, "ppq_ks_modup: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_ks_modup: bad limb count");
    if (n == 0) return PPQ_OK;
    if (n > 32768) return fail(PPQ_EINVAL, "ppq_ks_modup: batch too large");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    KsPlanDev* p;
    int rc = get_ks_plan(c, ell, &p);
    if (rc) return rc;
    const size_t N = c->hp.N;
    const int K = c->hp.K, M = ell + K, parts = p->h.parts;
    const size_t ys = (size_t)ell * N, ds = (size_t)parts * M * N;
    CU(c->scratch.ensure(ys * n));
    u64* y = c->scratch.p;
    launch_inv_row(c->k, (const u64*)d_c1, ys, y, ys, p->c1_limb.p, ell, (int)n, st);
    launch_inv_col(c->k, y, ys, p->c1_limb.p, p->c1_scale.p, ell, (int)n, st);
    for (const auto& d : p->digits)
        if (!launch_bconv(c->k, y, ys, (u64*)d_digits, ds, p->up_jobs.data() + d.job_first, d.n_jobs, p->h.up_chat.data(), c->h_limbs.data(), d.ns, (int)n, st))
            return fail(PPQ_EINVAL, "digit wider than 32 limbs");
    launch_fwd_col_plain(c->k, (u64*)d_digits, ds, p->dig_ids.p, parts * M, (int)n, st);
    // finish the converted limbs with a plain ROW pass; own-digit limbs are copies of c1
    std::vector<int> ids((size_t)parts * M);
    for (int j = 0; j < parts; ++j)
        for (int r = 0; r < M; ++r) {
            const bool own = r >= j * c->hp.alpha && r < std::min((j + 1) * c->hp.alpha, ell);
            ids[(size_t)j * M + r] = own ? -1 : (r < ell ? r : c->hp.L + (r - ell));
        }
    CU(c->ids.ensure(ids.size()));
    CU(cudaMemcpyAsync(c->ids.p, ids.data(), ids.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    for (size_t g = 0; g < n; ++g) {
        for (int j = 0; j < parts; ++j) {
            const int s = j * c->hp.alpha, e = std::min(s + c->hp.alpha, ell);
            CU(cudaMemcpyAsync((u64*)d_digits + g * ds + ((size_t)j * M + s) * N, (const u64*)d_c1 + g * ys + (size_t)s * N,
                               (size_t)(e - s) * N * sizeof(u64), cudaMemcpyDeviceToDevice, st));
        }
    }
    launch_fwd_row_plain(c->k, (u64*)d_digits, ds, c->ids.p, parts * M, (int)n, st);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    return PPQ_OK;
}

// stage view of EvalFastKeySwitchCoreExt: ext[g][poly][r] = sum_j digits[g][j][r] * key_j[r] mod q_r, canonical, on the
// plain (unscaled) key.  d_digits as written by ppq_ks_modup.
extern "C" int ppq_ks_inner_product(ppq_ctx* c, const ppq_rekey* rk, int ell, const uint64_t* d_digits, uint64_t* d_ext,
                                    size_t n, void* stream) {
    if (!c || !rk || !d_digits || !d_ext) return fail(PPQ_EINVAL, "ppq_ks_inner_product: null argument");
    if (rk->ctx != c) return fail(PPQ_EINVAL, "ppq_ks_inner_product: rekey belongs to another context");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_ks_inner_product: bad limb count");
    if (n == 0) return PPQ_OK;
    if (n * 2 * (size_t)ell > 65535) return fail(PPQ_EINVAL, "ppq_ks_inner_product: batch too large");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    KsPlanDev* p;
    if (int rc = get_ks_plan(c, ell, &p)) return rc;
    const size_t N = c->hp.N;
    const int K = c->hp.K, L = c->hp.L, M = ell + K, parts = p->h.parts, alpha = c->hp.alpha;
    const size_t ds = (size_t)parts * M * N, c1s = (size_t)ell * N, qs = (size_t)2 * ell * N, ps = (size_t)2 * K * N;
    CU(c->scratch2.ensure((c1s + qs + ps) * n));
    u64* c1 = c->scratch2.p;
    u64* extq = c1 + c1s * n;
    u64* extp = extq + qs * n;
    for (int j = 0; j < parts; ++j) {  // the digit's own limbs are read from a c1-shaped block by the kernel
        const int s0 = j * alpha, e0 = std::min(s0 + alpha, ell);
        CU(cudaMemcpy2DAsync(c1 + (size_t)s0 * N, c1s * sizeof(u64), (const u64*)d_digits + ((size_t)j * M + s0) * N,
                             ds * sizeof(u64), (size_t)(e0 - s0) * N * sizeof(u64), n, cudaMemcpyDeviceToDevice, st));
    }
    IpArgs ia;
    ia.c1 = c1;
    ia.c1_gs = c1s;
    ia.digits = (const u64*)d_digits;
    ia.dig_gs = ds;
    ia.key_b = rk->b;
    ia.key_a = rk->a;
    ia.key_qb = rk->b;  // unscaled Q limbs: this view returns the textbook inner product
    ia.key_qa = rk->a;
    ia.key_q_stride = L + K;
    ia.extq = extq;
    ia.extq_gs = qs;
    ia.extp = extp;
    ia.extp_gs = ps;
    ia.ell = ell;
    ia.K = K;
    ia.L = L;
    ia.parts = parts;
    ia.alpha = alpha;
    if (!launch_ip_tma(c->k, ia, (int)n, st)) return fail(PPQ_EINVAL, "ppq_ks_inner_product: unsupported shape");
    launch_finalize_mod(c->k, extq, ell, n, st);  // Q-limb results leave the kernel in [0, 3q)
    u64* ext = (u64*)d_ext;
    const size_t es = (size_t)2 * M * N;
    for (int poly = 0; poly < 2; ++poly) {
        CU(cudaMemcpy2DAsync(ext + (size_t)poly * M * N, es * sizeof(u64), extq + (size_t)poly * ell * N, qs * sizeof(u64),
                             (size_t)ell * N * sizeof(u64), n, cudaMemcpyDeviceToDevice, st));
        CU(cudaMemcpy2DAsync(ext + ((size_t)poly * M + ell) * N, es * sizeof(u64), extp + (size_t)poly * K * N, ps * sizeof(u64),
                             (size_t)K * N * sizeof(u64), n, cudaMemcpyDeviceToDevice, st));
    }
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    return PPQ_OK;
}

// stage view of ApproxModDown: out[g][poly][i] = (ext[g][poly][i] - NTT(conv_{P->q_i}(iNTT(P limbs)))) * P^-1 mod q_i
extern "C" int ppq_ks_moddown(ppq_ctx* c, int ell, const uint64_t* d_ext, uint64_t* d_out, size_t n, void* stream) {
    if (!c || !d_ext || !d_out) return fail(PPQ_EINVAL, "ppq_ks_moddown: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_ks_moddown: bad limb count");
    if (n == 0) return PPQ_OK;
    if (n * 2 * (size_t)ell > 65535) return fail(PPQ_EINVAL, "ppq_ks_moddown: batch too large");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    KsPlanDev* p;
    if (int rc = get_ks_plan(c, ell, &p)) return rc;
    const size_t N = c->hp.N;
    const int K = c->hp.K, M = ell + K;
    const size_t qs = (size_t)2 * ell * N, ps = (size_t)2 * K * N, es = (size_t)2 * M * N;
    CU(c->scratch2.ensure((2 * qs + ps) * n));
    u64* extq = c->scratch2.p;
    u64* conv = extq + qs * n;
    u64* extp = conv + qs * n;
    const u64* ext = (const u64*)d_ext;
    for (int poly = 0; poly < 2; ++poly) {
        CU(cudaMemcpy2DAsync(extq + (size_t)poly * ell * N, qs * sizeof(u64), ext + (size_t)poly * M * N, es * sizeof(u64),
                             (size_t)ell * N * sizeof(u64), n, cudaMemcpyDeviceToDevice, st));
        CU(cudaMemcpy2DAsync(extp + (size_t)poly * K * N, ps * sizeof(u64), ext + ((size_t)poly * M + ell) * N, es * sizeof(u64),
                             (size_t)K * N * sizeof(u64), n, cudaMemcpyDeviceToDevice, st));
    }
    // the product path folds P^-1 into the key and into the conversion constants; here the Q limbs get it explicitly
    launch_mult_scalar(c->k, extq, extq, p->pinv.p, p->pinv_sh.p, ell, n, st);
    launch_inv_row(c->k, extp, ps, extp, ps, p->p_limb.p, 2 * K, (int)n, st);
    launch_inv_col(c->k, extp, ps, p->p_limb.p, p->p_scale.p, 2 * K, (int)n, st);
    for (int poly = 0; poly < 2; ++poly)
        if (!launch_bconv(c->k, extp, ps, conv, qs, p->down_jobs.data() + (size_t)poly * ell, ell, p->h.down_chat.data(),
                          c->h_limbs.data(), K, (int)n, st))
            return fail(PPQ_EINVAL, "more than 32 auxiliary limbs");
    launch_fwd_col_plain(c->k, conv, qs, p->conv_ids.p, 2 * ell, (int)n, st);
    MdArgs ma;
    ma.conv = conv;
    ma.conv_gs = qs;
    ma.extq = extq;
    ma.extq_gs = qs;
    ma.ct_in = nullptr;
    ma.in_gs = 0;
    ma.out = (u64*)d_out;
    ma.out_gs = qs;
    ma.ell = ell;
    ma.accumulate = 0;
    launch_fwd_row_moddown(c->k, ma, (int)n, st);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    return PPQ_OK;
}

// FLEXIBLEAUTOEXT prime chain of GenCryptoContext for (depth, firstModSize, scalingModSize, extraBits): host only
extern "C" int ppq_params_chain(int logN, int depth, int first_bits, int scale_bits, int extra_bits, uint64_t* q_out,
                                int cap, int* n_out) {
    if (!q_out || !n_out || depth < 1 || logN < 10 || logN > 17 || first_bits > 60 || scale_bits > 60 || scale_bits < 16)
        return fail(PPQ_EINVAL, "ppq_params_chain: bad arguments");
    const std::vector<host::u64> q = host::synthetic_chain(logN, depth, first_bits, scale_bits, extra_bits);
    *n_out = (int)q.size();
    if ((int)q.size() > cap) return fail(PPQ_EINVAL, "ppq_params_chain: output array too small");
    for (size_t i = 0; i < q.size(); ++i) q_out[i] = q[i];
    return PPQ_OK;
}

// Host-only view of everything the library derives from (N, Q chain, dnum): no device needed.  JSON text.
static void js_arr(std::string& o, const char* name, const std::vector<u64>& v) {
    o += "\"";
    o += name;
    o += "\": [";
    for (size_t i = 0; i < v.size(); ++i) {
        if (i) o += ", ";
        o += "\"" + std::to_string(v[i]) + "\"";
    }
    o += "]";
}
extern "C" int ppq_params_dump(int logN, int L, const uint64_t* q, int dnum, int aux_bits, int extra_bits, int ell,
                               char* buf, size_t cap, size_t* need) {
    if (!q || L < 1 || !need) return fail(PPQ_EINVAL, "ppq_params_dump: bad arguments");
    host::Params hp;
    hp.logN = logN;
    hp.L = L;
    hp.dnum = dnum;
    hp.aux_bits = aux_bits;
    hp.extra_bits = extra_bits;
    hp.mod.assign(q, q + L);
    std::string err;
    if (!host::derive(hp, err)) return fail(PPQ_EINVAL, err);
    if (ell < 1 || ell > L) return fail(PPQ_EINVAL, "ppq_params_dump: bad ell");
    host::KsPlan ks;
    host::build_ks_plan(hp, ell, ks);
    std::string o = "{";
    o += "\"logN\": " + std::to_string(hp.logN) + ", \"L\": " + std::to_string(hp.L) + ", \"K\": " + std::to_string(hp.K) +
         ", \"dnum\": " + std::to_string(hp.dnum) + ", \"alpha\": " + std::to_string(hp.alpha) + ", \"ell\": " +
         std::to_string(ell) + ", \"parts\": " + std::to_string(ks.parts) + ", ";
    js_arr(o, "moduli", hp.mod);
    o += ", ";
    js_arr(o, "roots", hp.root);
    o += ", ";
    js_arr(o, "ninv", hp.ninv);
    o += ", \"sf\": [";
    char tmp[64];
    for (size_t i = 0; i < hp.sf.size(); ++i) {
        std::snprintf(tmp, sizeof(tmp), "%s%.17g", i ? ", " : "", hp.sf[i]);
        o += tmp;
    }
    o += "], \"sfbig\": [";
    for (size_t i = 0; i < hp.sfbig.size(); ++i) {
        std::snprintf(tmp, sizeof(tmp), "%s%.17g", i ? ", " : "", hp.sfbig[i]);
        o += tmp;
    }
    o += "], ";
    js_arr(o, "c1_scale", ks.c1_scale);   // N^-1 (D_j/q_i)^-1 mod q_i, i < ell
    o += ", ";
    js_arr(o, "p_scale", ks.p_scale);     // N^-1 (P/p_k)^-1 mod p_k, twice (one per polynomial)
    o += ", ";
    js_arr(o, "pinv", ks.pinv);           // P^-1 mod q_i
    o += ", ";
    js_arr(o, "up_chat", ks.up_chat);     // per ModUp job, the digit's (D_j/q_i) mod t
    o += ", ";
    js_arr(o, "down_chat", ks.down_chat); // per ModDown job, (P/p_k) P^-1 mod q_i
    o += ", \"up_jobs\": [";
    for (size_t i = 0; i < ks.up_jobs.size(); ++i) {
        const host::Job& j = ks.up_jobs[i];
        o += (i ? ", [" : "[") + std::to_string(j.src_first) + ", " + std::to_string(j.ns) + ", " + std::to_string(j.dst_slot) +
             ", " + std::to_string(j.limb) + ", " + std::to_string(j.chat_off) + "]";
    }
    o += "]";
    if (ell >= 2) {
        host::RsPlan rs;
        host::build_rs_plan(hp, ell, rs);
        o += ", ";
        js_arr(o, "rescale_qlinv", rs.qlinv);
    }
    o += "}";
    *need = o.size() + 1;
    if (buf && cap >= o.size() + 1) std::memcpy(buf, o.c_str(), o.size() + 1);
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- EvalAdd / EvalMult
// Batches larger than one launch's grid are walked in chunks inside the library: no "batch too large" at the ABI.
static size_t rows_chunk(int rows_per_ct) { return std::max<size_t>(1, 65535 / (size_t)rows_per_ct); }

extern "C" int ppq_eval_add(ppq_ctx* c, int ell, const uint64_t* a, const uint64_t* b, uint64_t* o, size_t n,
                            void* stream) {
    if (!c || !a || !b || !o) return fail(PPQ_EINVAL, "ppq_eval_add: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_eval_add: bad limb count");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    const size_t ct = (size_t)2 * ell * c->hp.N, step = rows_chunk(2 * ell);
    for (size_t s0 = 0; s0 < n; s0 += step)
        launch_eval_add(c->k, (const u64*)a + s0 * ct, (const u64*)b + s0 * ct, (u64*)o + s0 * ct, ell, std::min(step, n - s0),
                        (cudaStream_t)stream);
    CU(cudaGetLastError());
    return PPQ_OK;
}
extern "C" int ppq_finalize_sum(ppq_ctx* c, int ell, uint64_t* x, size_t n, void* stream) {
    if (!c || !x) return fail(PPQ_EINVAL, "ppq_finalize_sum: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_finalize_sum: bad limb count");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    const size_t ct = (size_t)2 * ell * c->hp.N, step = rows_chunk(2 * ell);
    for (size_t s0 = 0; s0 < n; s0 += step)
        launch_finalize_mod(c->k, (u64*)x + s0 * ct, ell, std::min(step, n - s0), (cudaStream_t)stream);
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_eval_mult_const(ppq_ctx* c, int ell, const uint64_t* in, uint64_t* out, size_t n, double cst,
                                   double sf, int rescale, void* stream) {
    if (!c || !in || !out) return fail(PPQ_EINVAL, "ppq_eval_mult_const: null argument");
    if (ell < (rescale ? 2 : 1) || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_eval_mult_const: bad limb count");
    if (ell > 64) return fail(PPQ_EINVAL, "ppq_eval_mult_const: more than 64 limbs");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    const int lo = rescale ? ell - 1 : ell;
    std::vector<u64> f(2 * (size_t)lo);
    host::mult_const_factors(c->hp, lo, sf, cst, f.data());
    if (!rescale) {
        FacPack pk{};
        for (int i = 0; i < lo; ++i) {
            pk.f[i] = f[i];
            pk.fs[i] = host::shoup(f[i], c->hp.mod[i]);
        }
        const size_t ct = (size_t)2 * ell * N, step = rows_chunk(2 * ell);
        for (size_t s0 = 0; s0 < n; s0 += step)
            launch_mult_scalar_pack(c->k, (const u64*)in + s0 * ct, (u64*)out + s0 * ct, pk, ell, std::min(step, n - s0) * 2 * ell, st);
        CU(cudaGetLastError());
        return PPQ_OK;
    }
    auto it = c->rs.find(ell);
    if (it == c->rs.end()) {
        std::unique_ptr<RsPlanDev> p(new RsPlanDev);
        host::build_rs_plan(c->hp, ell, p->h);
        CU(p->last_limb.upload(p->h.last_limb));
        CU(p->last_scale.upload(host::scale4(c->hp, p->h.last_limb, p->h.last_scale)));
        CU(p->jobs.upload(to_dev_jobs(p->h.jobs)));
        it = c->rs.emplace(ell, std::move(p)).first;
    }
    RsPlanDev* p = it->second.get();
    RsArgs ra{};
    for (int i = 0; i < lo; ++i) {  // (x - t) * q_last^-1 * f  ==  (x - t) * (q_last^-1 f mod q)
        ra.cst[i] = host::mulmod(f[i], p->h.qlinv[i], c->hp.mod[i]);
        ra.cst_sh[i] = host::shoup(ra.cst[i], c->hp.mod[i]);
    }
    // scratch: last limbs in COEFF [n][2][N] + converted limbs [n][2][ell-1][N]
    const size_t rs_ = (size_t)2 * N, cs_ = (size_t)2 * lo * N, ct = (size_t)2 * ell * N;
    const size_t step = 8192;
    CU(c->scratch2.ensure((rs_ + cs_) * std::min(n, step)));
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t m = std::min(step, n - s0);
        const u64* in_c = (const u64*)in + s0 * ct;
        u64* r = c->scratch2.p;
        u64* conv = r + rs_ * m;
        // gather the two last limbs of every ciphertext (poly stride ell*N) via the strided inverse ROW pass
        for (int poly = 0; poly < 2; ++poly)
            launch_inv_row(c->k, in_c + ((size_t)poly * ell + lo) * N, ct, r + (size_t)poly * N, rs_, p->last_limb.p, 1, (int)m, st);
        launch_inv_col(c->k, r, rs_, p->last_limb.p, p->last_scale.p, 2, (int)m, st);
        launch_sw_fwd_col(c->k, r, rs_, conv, cs_, p->jobs.p, (int)p->h.jobs.size(), (int)m, c->hp.mod[lo], st);
        ra.conv = conv;
        ra.conv_gs = cs_;
        ra.ct_in = in_c;
        ra.in_gs = ct;
        ra.out = (u64*)out + s0 * cs_;
        ra.out_gs = cs_;
        ra.ell = ell;
        launch_fwd_row_rescale(c->k, ra, (int)m, st);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- CKKS client side
extern "C" int ppq_encode(ppq_ctx* c, int ell, const double* d_vals, size_t n_vals, double sf, uint64_t* d_pt, size_t n,
                          void* stream) {
    if (!c || !d_vals || !d_pt) return fail(PPQ_EINVAL, "ppq_encode: null argument");
    if (ell < 1 || ell > c->hp.L || ell > 64) return fail(PPQ_EINVAL, "ppq_encode: bad limb count");
    if (n_vals > c->hp.N / 2) return fail(PPQ_EINVAL, "ppq_encode: more values than slots");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    const int* ids;
    if (int rc = cyc_ids(c, ell, &ids)) return rc;
    // Scaling factors beyond what an int64 coefficient can hold (FLEXIBLEAUTOEXT at level 0 with 59-bit scaling
    // primes: sfBig = 2^79) are applied in two steps, as OpenFHE's CKKSPackedEncoding::Encode does with its
    // approxFactor: round at sf / 2^a, then multiply the residues by 2^a mod q_i.  a keeps |coefficient| < 2^62 for
    // slot values up to 16 in magnitude; encode is tolerance-checked only.
    int sf_exp = 0;
    std::frexp(sf, &sf_exp);
    const int approx = sf_exp > 58 ? sf_exp - 58 : 0;
    FacPack pk{};
    if (approx > 0)
        for (int i = 0; i < ell; ++i) {
            const u64 q = c->hp.mod[i];
            u64 t = 1, b = 2 % q;
            for (int e = approx; e; e >>= 1, b = host::mulmod(b, b, q))
                if (e & 1) t = host::mulmod(t, b, q);
            pk.f[i] = t;
            pk.fs[i] = host::shoup(t, q);
        }
    const size_t step = rows_chunk(ell);
    // scratch2: complex work [m][N/2] double2  (= m*N u64)  +  int64 coefficients [m][N]
    CU(c->scratch2.ensure(2 * std::min(n, step) * N));
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t m = std::min(step, n - s0);
        double2* z = reinterpret_cast<double2*>(c->scratch2.p);
        int64_t* co = reinterpret_cast<int64_t*>(c->scratch2.p + m * N);
        u64* pt = (u64*)d_pt + s0 * ell * N;
        ckks::launch_encode_fft(c->fft, d_vals + s0 * n_vals, n_vals, std::ldexp(sf, -approx), z, co, m, st);
        launch_small_fwd_ntt(c->k, co, pt, 1, ell, m, ids, st);
        if (approx > 0) launch_mult_scalar_pack(c->k, pt, pt, pk, ell, m * ell, st);  // the NTT is linear: 2^a applied after it
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

// MakeCKKSPackedPlaintext + Encrypt in one call: the encoded coefficients join e0 BEFORE the transform (the NTT is
// linear), so a fresh ciphertext costs 3*ell forward limb transforms instead of 4*ell; output identical to ppq_encode
// followed by ppq_encrypt.
extern "C" int ppq_encode_encrypt(ppq_ctx* c, int ell, const uint64_t* pkb, const uint64_t* pka, const double* d_vals,
                                  size_t n_vals, double sf, const int64_t* small, uint64_t* ct, size_t n, void* stream) {
    if (!c || !pkb || !pka || !d_vals || !small || !ct) return fail(PPQ_EINVAL, "ppq_encode_encrypt: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_encode_encrypt: bad limb count");
    if (n_vals > c->hp.N / 2) return fail(PPQ_EINVAL, "ppq_encode_encrypt: more values than slots");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    int sf_exp = 0;
    std::frexp(sf, &sf_exp);
    const int approx = sf_exp > 58 ? sf_exp - 58 : 0;  // as ppq_encode: round at sf / 2^approx, then times 2^approx
    if (approx > 62) return fail(PPQ_EINVAL, "ppq_encode_encrypt: scaling factor out of range");
    const size_t step = rows_chunk(3 * ell);
    const size_t m_max = std::min(n, step);
    // scratch2: (v, m + e0, e1) in EVAL [m][3][ell][N], then the FFT work [m][N/2] double2 and the coefficients [m][N]
    CU(c->scratch2.ensure(m_max * (3 * (size_t)ell + 2) * N));
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t m = std::min(step, n - s0);
        u64* vee = c->scratch2.p;
        double2* z = reinterpret_cast<double2*>(vee + m * 3 * ell * N);
        int64_t* co = reinterpret_cast<int64_t*>(vee + m * (3 * (size_t)ell + 1) * N);
        ckks::launch_encode_fft(c->fft, d_vals + s0 * n_vals, n_vals, std::ldexp(sf, -approx), z, co, m, st);
        launch_small_fwd_encrypt(c->k, small + s0 * 3 * N, vee, ell, m, (const u64*)pkb, (const u64*)pka, nullptr,
                                 (u64*)ct + s0 * 2 * ell * N, st, co, approx);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_encrypt(ppq_ctx* c, int ell, const uint64_t* pkb, const uint64_t* pka, const uint64_t* pt,
                           const int64_t* small, uint64_t* ct, size_t n, void* stream) {
    if (!c || !pkb || !pka || !pt || !small || !ct) return fail(PPQ_EINVAL, "ppq_encrypt: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_encrypt: bad limb count");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    const size_t step = rows_chunk(3 * ell);
    CU(c->scratch2.ensure(std::min(n, step) * 3 * ell * N));
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t m = std::min(step, n - s0);
        u64* vee = c->scratch2.p;  // [m][3][ell][N]
        launch_small_fwd_encrypt(c->k, small + s0 * 3 * N, vee, ell, m, (const u64*)pkb, (const u64*)pka,
                                 (const u64*)pt + s0 * ell * N, (u64*)ct + s0 * 2 * ell * N, st);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

// known-answer hook for the samplers' generator (RFC 8439 section 2.3.2 vector in tests/test_capi_exports.py); host only
extern "C" int ppq_chacha20_block(const uint32_t* key8, uint32_t counter, const uint32_t* nonce3, uint32_t* out16) {
    if (!key8 || !nonce3 || !out16) return fail(PPQ_EINVAL, "ppq_chacha20_block: null argument");
    sampler::chacha20_block_host(key8, counter, nonce3, out16);
    return PPQ_OK;
}
// 32 bytes of key material for the samplers: the caller's (tests, reproducible runs) or fresh OS entropy per call
static int sampler_key(const uint8_t* seed32, sampler::Key& k) {
    if (seed32) {
        std::memcpy(k.w, seed32, 32);
        return PPQ_OK;
    }
    size_t got = 0;
    while (got < 32) {
        const ssize_t r = getrandom(reinterpret_cast<uint8_t*>(k.w) + got, 32 - got, 0);
        if (r < 0) return fail(PPQ_EINVAL, "getrandom failed: no entropy source for key material");
        got += (size_t)r;
    }
    return PPQ_OK;
}
extern "C" int ppq_sample_encrypt_noise(ppq_ctx* c, const uint8_t* seed32, int64_t* d_small, size_t n, void* stream) {
    if (!c || !d_small) return fail(PPQ_EINVAL, "ppq_sample_encrypt_noise: null argument");
    if (n == 0) return PPQ_OK;
    if (n * 3 > 65535) return fail(PPQ_EINVAL, "ppq_sample_encrypt_noise: at most 21845 triples per call");
    CU(cudaSetDevice(c->device));
    sampler::Key master;
    if (int rc = sampler_key(seed32, master)) return rc;
    const sampler::Key k = sampler::derive_key(master, 0x454e4352u /* "ENCR" */);
    sampler::launch_sample_small(k, sampler::DOM_SMALL, 0, c->kinds3.p, 3, d_small, c->hp.N, n * 3, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return PPQ_OK;
}

// ---- key generation.  Samples come in as small integers so that the arithmetic is bit-exact
// against the oracle; ppq_sample_keygen / ppq_sample_encrypt_noise draw them on the device.
static int small_to_eval(ppq_ctx* c, const int64_t* small, u64* out, int n_small, size_t n_groups, cudaStream_t st) {
    const int M = c->hp.M();
    const int* ids;
    if (int rc = cyc_ids(c, M, &ids)) return rc;
    launch_small_fwd_ntt(c->k, small, out, n_small, M, n_groups, ids, st);
    return PPQ_OK;
}

extern "C" int ppq_keygen(ppq_ctx* c, const int64_t* d_s, const int64_t* d_e, const uint64_t* d_a, uint64_t* d_sk,
                          uint64_t* d_pkb, void* stream) {
    if (!c || !d_s || !d_e || !d_a || !d_sk || !d_pkb) return fail(PPQ_EINVAL, "ppq_keygen: null argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    const int M = c->hp.M();
    CU(c->scratch2.ensure(2 * (size_t)M * N + 2 * N));
    u64* se = c->scratch2.p;                                            // [2][M][N]
    int64_t* small = reinterpret_cast<int64_t*>(se + 2 * (size_t)M * N);  // [2][N]
    CU(cudaMemcpyAsync(small, d_s, N * sizeof(int64_t), cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(small + N, d_e, N * sizeof(int64_t), cudaMemcpyDeviceToDevice, st));
    if (int rc = small_to_eval(c, small, se, 2, 1, st)) return rc;
    launch_keygen_pw(c->k, se, (const u64*)d_a, (u64*)d_sk, (u64*)d_pkb, c->hp.L, M, st);
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_sample_keygen(ppq_ctx* c, const uint8_t* seed32, int64_t* d_s, int64_t* d_e, uint64_t* d_a, void* stream) {
    if (!c || !d_s || !d_e || !d_a) return fail(PPQ_EINVAL, "ppq_sample_keygen: null argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    // the secret half (s, e) and the public half (a) are drawn under DIFFERENT keys: sub-keys of the seed by a
    // domain-separated KDF, or two independent 256-bit draws from the OS when no seed is given
    sampler::Key ksec, kpub;
    if (seed32) {
        sampler::Key master;
        if (int rc = sampler_key(seed32, master)) return rc;
        ksec = sampler::derive_key(master, 0x53454352u /* "SECR" */);
        kpub = sampler::derive_key(master, 0x5055424cu /* "PUBL" */);
    } else {
        if (int rc = sampler_key(nullptr, ksec)) return rc;
        if (int rc = sampler_key(nullptr, kpub)) return rc;
    }
    sampler::launch_sample_small(ksec, sampler::DOM_SMALL, 0, c->kinds3.p, 1, d_s, c->hp.N, 1, st);      // ternary
    sampler::launch_sample_small(ksec, sampler::DOM_SMALL, 1, c->kinds3.p + 1, 1, d_e, c->hp.N, 1, st);  // Gaussian
    sampler::launch_sample_uniform(kpub, sampler::DOM_UNIFORM, c->k, c->hp.M(), d_a, st);
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_rekeygen(ppq_ctx* c, const uint64_t* d_sk_old, const uint64_t* d_pkb, const uint64_t* d_pka,
                            const int64_t* d_small, uint64_t* d_kb, uint64_t* d_ka, void* stream) {
    if (!c || !d_sk_old || !d_pkb || !d_pka || !d_small || !d_kb || !d_ka)
        return fail(PPQ_EINVAL, "ppq_rekeygen: null argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    const int L = c->hp.L, M = c->hp.M(), dnum = c->hp.dnum;
    CU(c->scratch2.ensure((size_t)dnum * 3 * M * N));
    u64* uee = c->scratch2.p;  // [dnum][3][M][N]
    if (int rc = small_to_eval(c, d_small, uee, 3, (size_t)dnum, st)) return rc;
    if (!c->pmod.p) {  // P mod q_r, once per context
        std::vector<u64> pm(L);
        for (int r = 0; r < L; ++r) {
            u64 v = 1;
            for (int k = 0; k < c->hp.K; ++k) v = host::mulmod(v, c->hp.mod[L + k] % c->hp.mod[r], c->hp.mod[r]);
            pm[r] = v;
        }
        CU(c->pmod.upload(pm));
    }
    launch_rekeygen_pw(c->k, (const u64*)d_pkb, (const u64*)d_pka, uee, (const u64*)d_sk_old, c->pmod.p, (u64*)d_kb,
                       (u64*)d_ka, L, M, c->hp.alpha, dnum, st);
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_decrypt_core(ppq_ctx* c, int ell, const uint64_t* sk, const uint64_t* ct, uint64_t* coeff, size_t n,
                                void* stream) {
    if (!c || !sk || !ct || !coeff) return fail(PPQ_EINVAL, "ppq_decrypt_core: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_decrypt_core: bad limb count");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    const int* ids;
    if (int rc = cyc_ids(c, ell, &ids)) return rc;
    if (c->ntt_scale4.empty()) {  // per limb, once: (N^-1, shoup, N^-1 psi^-brev(1), shoup)
        std::vector<int> all(c->hp.M());
        for (int r = 0; r < c->hp.M(); ++r) all[r] = r;
        c->ntt_scale4 = host::scale4(c->hp, all, c->hp.ninv);
    }
    auto it = c->cyc_scale.find(ell);
    if (it == c->cyc_scale.end()) {  // the output factors of the inverse transform in the same cyclic order as the ids
        std::unique_ptr<DevBuf<u64>> b(new DevBuf<u64>);
        std::vector<u64> h((size_t)4 * CYC_LEN);
        for (int i = 0; i < CYC_LEN; ++i)
            for (int k = 0; k < 4; ++k) h[4 * (size_t)i + k] = c->ntt_scale4[4 * (size_t)(i % ell) + k];
        CU(b->upload(h));
        it = c->cyc_scale.emplace(ell, std::move(b)).first;
    }
    const size_t step = rows_chunk(ell);
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t m = std::min(step, n - s0);
        u64* co = (u64*)coeff + s0 * ell * N;
        launch_decrypt_inv_ntt(c->k, (const u64*)ct + s0 * 2 * ell * N, (const u64*)sk, co, ell, m, ids, it->second->p, st);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_decode(ppq_ctx* c, int ell, const uint64_t* coeff, double sf, double* d_vals, size_t n,
                          void* stream) {
    if (!c || !coeff || !d_vals) return fail(PPQ_EINVAL, "ppq_decode: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_decode: bad limb count");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = c->hp.N;
    auto it = c->garner.find(ell);
    if (it == c->garner.end()) {
        // Garner constants, once per ell: ginv[i] = (q_0..q_{i-1})^-1 mod q_i, then the table q_k mod q_i
        // ... and q_0 q_1 q_2 mod q_i for the short interpolation's check
        std::vector<u64> g((size_t)ell + (size_t)ell * ell + 2 * (size_t)ell, 0);
        for (int i = 1; i < ell; ++i) {
            u64 pr = 1;
            for (int k = 0; k < i; ++k) pr = host::mulmod(pr, c->hp.mod[k] % c->hp.mod[i], c->hp.mod[i]);
            g[i] = host::invmod(pr, c->hp.mod[i]);
        }
        for (int i = 0; i < ell; ++i)
            for (int k = 0; k < ell; ++k) g[(size_t)ell + (size_t)i * ell + k] = c->hp.mod[k] % c->hp.mod[i];
        for (int i = 3; i < ell; ++i) {
            u64 pr = 1;
            for (int k = 0; k < 3; ++k) pr = host::mulmod(pr, c->hp.mod[k] % c->hp.mod[i], c->hp.mod[i]);
            g[(size_t)ell + (size_t)ell * ell + i] = pr;
            g[(size_t)ell + (size_t)ell * ell + ell + i] =
                host::mulmod(c->hp.mod[0] % c->hp.mod[i], c->hp.mod[1] % c->hp.mod[i], c->hp.mod[i]);
        }
        std::unique_ptr<DevBuf<u64>> b(new DevBuf<u64>);
        CU(b->upload(g));
        it = c->garner.emplace(ell, std::move(b)).first;
    }
    const u64* gin = it->second->p;
    const size_t step = 8192;
    CU(c->scratch2.ensure(std::min(n, step) * N));
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t m = std::min(step, n - s0);
        double2* z = reinterpret_cast<double2*>(c->scratch2.p);
        ckks::launch_decode(c->fft, c->k, (const u64*)coeff + s0 * ell * N, ell, gin, gin + ell, gin + ell + (size_t)ell * ell, sf, z, d_vals + s0 * (N / 2), m, st);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- aggregation (deferred ModDown)
// sum_c ReEncrypt(ct_c, rk_c) for the ciphertexts of one slot, bit-identical to re-encrypting every client's
// ciphertext and adding the results (changeCipherDomain.cpp:105 followed by aggregateEncryptedWeights.cpp:106), with
// the LINEAR tail of ApproxModDown taken out of the per-client loop:
//   ReEncrypt(ct)_poly = c0[poly==0] + (x - NTT(conv(u))) * P^-1,   x = <digits, key> on the Q limbs,
//   u_k = canonical iNTT of the P limbs times (P/p_k)^-1,  conv(u)_i = sum_k u_k [P/p_k]_{q_i}  (no correction).
// Modulo q_i both the NTT and conv are linear (conv in the integer lifts u_k), so
//   sum_c ReEncrypt(ct_c) = sum_c (c0_c + x_c P^-1)  -  NTT( sum_k (sum_c u_{k,c}) [P/p_k]_{q_i} P^-1 ).
// Per client only the non-linear part runs (ModUp, inner product, inverse NTT of the P limbs); the inner product
// adds its Q-limb results and c0 into accE; the conversion runs once per `group` <= 8 clients on the integer sum of
// their u (k_bconv_wide) into accW; ppq_agg_finish runs the 2*ell forward NTTs once per slot.
struct ppq_agg {
    ppq_ctx* ctx = nullptr;
    int ell = 0;
    size_t n_slots = 0;
    DevBuf<u64> accE, accW, ring;  // [n_slots][2*ell][N] each; ring: [group][n_slots][2K][N] COEFF u of the open group
    int group = 8, in_group = 0;
    size_t count = 0;     // clients added since the last finish
    bool w_valid = false; // accW holds a sum
    bool legacy = false;  // shapes the split inner product does not cover: plain ppq_reencrypt_accumulate into accE
};

extern "C" int ppq_agg_create(ppq_ctx* c, int ell, size_t n_slots, ppq_agg** out) {
    if (!c || !out) return fail(PPQ_EINVAL, "ppq_agg_create: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_agg_create: bad limb count");
    if (n_slots < 1 || n_slots > 32768) return fail(PPQ_EINVAL, "ppq_agg_create: 1..32768 slots");
    CU(cudaSetDevice(c->device));
    std::unique_ptr<ppq_agg> a(new ppq_agg);
    a->ctx = c;
    a->ell = ell;
    a->n_slots = n_slots;
    const size_t N = c->hp.N, K = c->hp.K;
    a->legacy = K > 16;  // the wide conversion holds <= 16 source limbs
    int grp = 8;
    if (const char* e = std::getenv("PPQ_AGG_GROUP")) grp = std::atoi(e);
    a->group = grp < 1 ? 1 : grp > 8 ? 8 : grp;
    CU(a->accE.ensure(n_slots * 2 * ell * N));
    if (!a->legacy) {
        CU(a->accW.ensure(n_slots * 2 * ell * N));
        CU(a->ring.ensure((size_t)a->group * n_slots * 2 * K * N));
    }
    *out = a.release();
    return PPQ_OK;
}
extern "C" void ppq_agg_destroy(ppq_agg* a) {
    if (!a) return;
    cudaSetDevice(a->ctx->device);
    delete a;
}
extern "C" size_t ppq_agg_count(const ppq_agg* a) { return a ? a->count : 0; }

static int agg_flush(ppq_agg* a, KsPlanDev* p, cudaStream_t st) {
    ppq_ctx* c = a->ctx;
    if (a->in_group == 0) return PPQ_OK;
    const size_t N = c->hp.N;
    const int K = c->hp.K, ell = a->ell;
    const size_t ext_gs = (size_t)2 * K * N, ring_stride = a->n_slots * ext_gs, ct = (size_t)2 * ell * N;
    prof_mark(c, st, -1);
    for (int poly = 0; poly < 2; ++poly)
        if (!launch_bconv_wide(c->k, a->ring.p, ext_gs, ring_stride, a->in_group, a->accW.p, ct, a->w_valid ? 0 : 1,
                               p->down_jobs.data() + (size_t)poly * ell, ell, p->h.down_chat.data(), c->h_limbs.data(), K,
                               (int)a->n_slots, st))
            return fail(PPQ_EINVAL, "ppq_agg: more than 16 auxiliary limbs");
    prof_mark(c, st, 8);
    a->w_valid = true;
    a->in_group = 0;
    return PPQ_OK;
}

extern "C" int ppq_agg_add(ppq_agg* a, const ppq_rekey* rk, const uint64_t* d_in, void* stream) {
    if (!a || !rk || !d_in) return fail(PPQ_EINVAL, "ppq_agg_add: null argument");
    ppq_ctx* c = a->ctx;
    if (rk->ctx != c) return fail(PPQ_EINVAL, "ppq_agg_add: rekey belongs to another context");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int ell = a->ell;
    const size_t N = c->hp.N, n = a->n_slots;
    const size_t ct = (size_t)2 * ell * N;
    if (a->legacy) {
        if (a->count == 0) CU(cudaMemsetAsync(a->accE.p, 0, n * ct * sizeof(u64), st));
        int rc = reencrypt_common(c, rk, ell, (const u64*)d_in, a->accE.p, n, stream, 1);
        if (rc == PPQ_OK) a->count++;
        return rc;
    }
    KsPlanDev* p;
    int rc = get_ks_plan(c, ell, &p);
    if (rc) return rc;
    const int K = c->hp.K, L = c->hp.L, parts = p->h.parts, M = ell + K;
    WaveLayout w;  // only the ModUp half needs wave scratch: COEFF c1 and the digits
    w.y = (size_t)ell * N;
    w.dig = (size_t)parts * M * N;
    w.extq = w.extp = w.conv = 0;
    w.per_ct = w.y + w.dig;
    const int G = pick_wave(c, w, n);
    CU(c->scratch.ensure(w.per_ct * G));
    u64* y = c->scratch.p;
    u64* dig = y + w.y * G;
    const size_t ext_gs = (size_t)2 * K * N;
    u64* ring_e = a->ring.p + (size_t)a->in_group * n * ext_gs;
    const u64* in = (const u64*)d_in;
    for (size_t g0 = 0; g0 < n; g0 += G) {
        const int ng = (int)std::min((size_t)G, n - g0);
        const u64* in_g = in + g0 * ct;
        u64* extp = ring_e + g0 * ext_gs;
        prof_mark(c, st, -1);
        launch_inv_row(c->k, in_g + (size_t)ell * N, ct, y, w.y, p->c1_limb.p, ell, ng, st);
        prof_mark(c, st, 0);
        launch_inv_col(c->k, y, w.y, p->c1_limb.p, p->c1_scale.p, ell, ng, st);
        prof_mark(c, st, 1);
        for (const auto& d : p->digits)
            if (!launch_bconv(c->k, y, w.y, dig, w.dig, p->up_jobs.data() + d.job_first, d.n_jobs, p->h.up_chat.data(), c->h_limbs.data(), d.ns, ng, st))
                return fail(PPQ_EINVAL, "digit wider than 32 limbs");
        prof_mark(c, st, 2);
        launch_fwd_col_plain(c->k, dig, w.dig, p->dig_ids.p, parts * M, ng, st);
        prof_mark(c, st, 3);
        launch_fwd_row_plain(c->k, dig, w.dig, p->dig_ids.p, parts * M, ng, st, PPQ_DIGITS_LAZY_OK(parts));
        prof_mark(c, st, 4);
        IpArgs ia;
        ia.c1 = in_g + (size_t)ell * N;
        ia.c1_gs = ct;
        ia.digits = dig;
        ia.dig_gs = w.dig;
        ia.key_b = rk->b;
        ia.key_a = rk->a;
        ia.key_qb = rk->qb;
        ia.key_qa = rk->qa;
        ia.key_q_stride = rk->q_stride;
        ia.extq = nullptr;
        ia.extq_gs = 0;
        ia.extp = extp;
        ia.extp_gs = ext_gs;
        ia.ell = ell;
        ia.K = K;
        ia.L = L;
        ia.parts = parts;
        ia.alpha = c->hp.alpha;
        ia.agg = a->count == 0 ? 1 : 2;
        ia.acc = a->accE.p + g0 * ct;
        ia.acc_gs = ct;
        ia.c0 = in_g;
        ia.c0_gs = ct;
        if (!launch_ip_tma(c->k, ia, ng, st)) return fail(PPQ_EINVAL, "ppq_agg_add: unsupported shape");
        prof_mark(c, st, 5);
        launch_inv_row(c->k, extp, ext_gs, extp, ext_gs, p->p_limb.p, 2 * K, ng, st);
        prof_mark(c, st, 6);
        launch_inv_col(c->k, extp, ext_gs, p->p_limb.p, p->p_scale.p, 2 * K, ng, st);
        prof_mark(c, st, 7);
    }
    a->in_group++;
    a->count++;
    if (a->in_group == a->group) rc = agg_flush(a, p, st);
    if (rc == PPQ_OK) CU(cudaGetLastError());
    return rc;
}

// several clients per pass (device-resident inputs): the ModUp half runs over nb x n_slots ciphertexts per launch and the
// inner product sums the nb clients of a slot in registers before it touches the running sum (k_ip_agg)
static int agg_add_batch(ppq_agg* a, const ppq_rekey* const* rks, const uint64_t* const* ins, int nb, cudaStream_t st) {
    ppq_ctx* c = a->ctx;
    const int ell = a->ell;
    const size_t N = c->hp.N, n = a->n_slots, ct = (size_t)2 * ell * N;
    KsPlanDev* p;
    if (int rc = get_ks_plan(c, ell, &p)) return rc;
    const int K = c->hp.K, L = c->hp.L, parts = p->h.parts, M = ell + K;
    const size_t ys = (size_t)ell * N, ds = (size_t)parts * M * N, ng = n * nb;
    CU(c->scratch.ensure((ys + ds) * ng));
    u64* y = c->scratch.p;
    u64* dig = y + ys * ng;
    const size_t ext_gs = (size_t)2 * K * N, ring_stride = n * ext_gs;
    u64* extp = a->ring.p + (size_t)a->in_group * ring_stride;
    prof_mark(c, st, -1);
    {   // c1 of the nb clients of this pass (separate input buffers) in one launch
        const u64* c1s[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int cl = 0; cl < nb; ++cl) c1s[cl] = (const u64*)ins[cl] + (size_t)ell * N;
        launch_inv_row_multi(c->k, c1s, nb, (int)n, ct, y, ys, p->c1_limb.p, ell, st);
    }
    prof_mark(c, st, 0);
    launch_inv_col(c->k, y, ys, p->c1_limb.p, p->c1_scale.p, ell, (int)ng, st);
    prof_mark(c, st, 1);
    for (const auto& d : p->digits)
        if (!launch_bconv(c->k, y, ys, dig, ds, p->up_jobs.data() + d.job_first, d.n_jobs, p->h.up_chat.data(), c->h_limbs.data(), d.ns, (int)ng, st))
            return fail(PPQ_EINVAL, "digit wider than 32 limbs");
    prof_mark(c, st, 2);
    // digits canonical here: k_ip_agg puts the products of two clients into one Barrett operand (2 parts terms)
    static const bool use_cluster = [] { const char* e = std::getenv("PPQ_FWD_CLUSTER"); return e && e[0] == '1'; }();
    if (use_cluster && launch_fwd_ntt_cluster(c->k, dig, ds, p->dig_ids.p, parts * M, (int)ng, st, false)) {
        prof_mark(c, st, 3);  // experiment: both passes in one cluster kernel, booked under stage 3
        prof_mark(c, st, 4);
    } else {
        launch_fwd_col_plain(c->k, dig, ds, p->dig_ids.p, parts * M, (int)ng, st);
        prof_mark(c, st, 3);
        launch_fwd_row_plain(c->k, dig, ds, p->dig_ids.p, parts * M, (int)ng, st, false);
        prof_mark(c, st, 4);
    }
    IpArgs ia;
    ia.c1 = nullptr;
    ia.c1_gs = ct;
    ia.digits = dig;
    ia.dig_gs = ds;
    ia.key_b = ia.key_a = ia.key_qb = ia.key_qa = nullptr;
    ia.key_q_stride = 0;
    ia.extq = nullptr;
    ia.extq_gs = 0;
    ia.extp = extp;
    ia.extp_gs = ext_gs;
    ia.ell = ell;
    ia.K = K;
    ia.L = L;
    ia.parts = parts;
    ia.alpha = c->hp.alpha;
    ia.agg = a->count == 0 ? 1 : 2;
    ia.acc = a->accE.p;
    ia.acc_gs = ct;
    IpMulti X;
    X.nb = nb;
    X.extp_cs = ring_stride;
    for (int cl = 0; cl < nb; ++cl) {
        X.ct_in[cl] = (const u64*)ins[cl];
        X.key_b[cl] = rks[cl]->b;
        X.key_a[cl] = rks[cl]->a;
        X.key_qb[cl] = rks[cl]->qb;
        X.key_qa[cl] = rks[cl]->qa;
        X.key_q_stride[cl] = rks[cl]->q_stride;
    }
    if (!launch_ip_agg(c->k, ia, X, (int)n, st)) return fail(PPQ_EINVAL, "ppq_agg_add_many: unsupported shape");
    prof_mark(c, st, 5);
    launch_inv_row(c->k, extp, ext_gs, extp, ext_gs, p->p_limb.p, 2 * K, (int)ng, st);
    prof_mark(c, st, 6);
    launch_inv_col(c->k, extp, ext_gs, p->p_limb.p, p->p_scale.p, 2 * K, (int)ng, st);
    prof_mark(c, st, 7);
    a->in_group += nb;
    a->count += (size_t)nb;
    if (a->in_group == a->group) return agg_flush(a, p, st);
    return PPQ_OK;
}

extern "C" int ppq_agg_add_many(ppq_agg* a, const ppq_rekey* const* rks, const uint64_t* const* d_ins, size_t n_clients,
                                void* stream) {
    if (!a || !rks || !d_ins) return fail(PPQ_EINVAL, "ppq_agg_add_many: null argument");
    ppq_ctx* c = a->ctx;
    for (size_t i = 0; i < n_clients; ++i) {
        if (!rks[i] || !d_ins[i]) return fail(PPQ_EINVAL, "ppq_agg_add_many: null client entry");
        if (rks[i]->ctx != c) return fail(PPQ_EINVAL, "ppq_agg_add_many: rekey belongs to another context");
    }
    CU(cudaSetDevice(c->device));
    int want = 4;  // clients per pass: 1 / 2 / 4 = 96.7 / 94.4 / 93.8 us per ciphertext at C3 (profiles/r10_notes.md)
    if (const char* e = std::getenv("PPQ_AGG_NB")) want = std::atoi(e);
    const size_t per_client = ((size_t)a->ell + (size_t)ppq_ctx_num_parts(c, a->ell) * (a->ell + c->hp.K)) * c->hp.N * a->n_slots * sizeof(u64);
    size_t i = 0;
    while (i < n_clients) {
        int nb = 1;
        if (!a->legacy && c->wave == 0 && !std::getenv("PPQ_WAVE")) {
            for (int cand : {4, 2}) {
                if (cand > want || (size_t)cand > n_clients - i || a->in_group + cand > a->group) continue;
                if (per_client * cand > ((size_t)24 << 30)) continue;  // scratch budget
                nb = cand;
                break;
            }
        }
        int rc;
        if (nb == 1) rc = ppq_agg_add(a, rks[i], d_ins[i], stream);
        else rc = agg_add_batch(a, rks + i, d_ins + i, nb, (cudaStream_t)stream);
        if (rc) return rc;
        i += (size_t)nb;
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

extern "C" int ppq_agg_finish(ppq_agg* a, uint64_t* d_out, void* stream) {
    if (!a || !d_out) return fail(PPQ_EINVAL, "ppq_agg_finish: null argument");
    if (a->count == 0) return fail(PPQ_EINVAL, "ppq_agg_finish: nothing was added");
    ppq_ctx* c = a->ctx;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int ell = a->ell;
    const size_t N = c->hp.N, n = a->n_slots, ct = (size_t)2 * ell * N;
    if (a->legacy) {
        CU(cudaMemcpyAsync(d_out, a->accE.p, n * ct * sizeof(u64), cudaMemcpyDeviceToDevice, st));
        a->count = 0;
        return PPQ_OK;
    }
    KsPlanDev* p;
    int rc = get_ks_plan(c, ell, &p);
    if (rc) return rc;
    rc = agg_flush(a, p, st);
    if (rc) return rc;
    prof_mark(c, st, -1);
    launch_fwd_col_plain(c->k, a->accW.p, ct, p->conv_ids.p, 2 * ell, (int)n, st);
    prof_mark(c, st, 9);
    MdArgs ma;
    ma.conv = a->accW.p;
    ma.conv_gs = ct;
    ma.extq = a->accE.p;
    ma.extq_gs = ct;
    ma.ct_in = nullptr;  // c0 is already inside accE
    ma.in_gs = 0;
    ma.out = (u64*)d_out;
    ma.out_gs = ct;
    ma.ell = ell;
    ma.accumulate = 0;
    launch_fwd_row_moddown(c->k, ma, (int)n, st);
    prof_mark(c, st, 10);
    a->count = 0;
    a->w_valid = false;
    CU(cudaGetLastError());
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- cross-GPU sum (NCCL)
// NCCL is bound at first use with dlopen: inside a process that already carries a libnccl.so.2 (torch's) that copy is
// the one used; a plain C++ host gets the system library.  No link-time dependency, no second copy.
namespace {
struct Nccl {
    void* h = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, char[128], int) = nullptr;  // ncclUniqueId is passed BY VALUE (128-byte struct)
    int (*CommDestroy)(void*) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, void*, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
struct NcclId { char internal[128]; };
typedef int (*InitRankFn)(void**, int, NcclId, int);
Nccl g_nccl;
int nccl_load() {
    if (g_nccl.h) return PPQ_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(PPQ_ECUDA, std::string("cannot load libnccl.so.2: ") + dlerror());
    Nccl n;
    n.h = h;
    n.GetUniqueId = (int (*)(void*))dlsym(h, "ncclGetUniqueId");
    n.CommInitRank = (int (*)(void**, int, char[128], int))dlsym(h, "ncclCommInitRank");
    n.CommDestroy = (int (*)(void*))dlsym(h, "ncclCommDestroy");
    n.Reduce = (int (*)(const void*, void*, size_t, int, int, int, void*, cudaStream_t))dlsym(h, "ncclReduce");
    n.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(h, "ncclAllReduce");
    n.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
    if (!n.GetUniqueId || !n.CommInitRank || !n.CommDestroy || !n.Reduce || !n.AllReduce || !n.GetErrorString)
        return fail(PPQ_ECUDA, "libnccl.so.2 lacks an expected symbol");
    g_nccl = n;
    return PPQ_OK;
}
const int kNcclUint64 = 5, kNcclSum = 0;  // ncclDataType_t / ncclRedOp_t values of nccl.h (stable ABI since 2.0)
}  // namespace
#define NC(call)                                                                                          \
    do {                                                                                                  \
        int e_ = (call);                                                                                  \
        if (e_ != 0) return fail(PPQ_ECUDA, std::string(#call) + ": " + g_nccl.GetErrorString(e_));       \
    } while (0)

struct ppq_comm {
    void* comm = nullptr;  // ncclComm_t
    int world = 1, rank = 0, device = 0;
    bool owned = false;
};
extern "C" int ppq_comm_unique_id(uint8_t* id128) {
    if (!id128) return fail(PPQ_EINVAL, "ppq_comm_unique_id: null");
    if (int rc = nccl_load()) return rc;
    NC(g_nccl.GetUniqueId(id128));
    return PPQ_OK;
}
extern "C" int ppq_comm_init(ppq_ctx* c, int world, int rank, const uint8_t* id128, ppq_comm** out) {
    if (!c || !id128 || !out || world < 1 || rank < 0 || rank >= world) return fail(PPQ_EINVAL, "ppq_comm_init: bad arguments");
    if (world > 15) return fail(PPQ_EINVAL, "ppq_comm_init: a u64 sum of canonical residues (< 2^60) holds for at most 15 ranks");
    if (int rc = nccl_load()) return rc;
    CU(cudaSetDevice(c->device));
    std::unique_ptr<ppq_comm> m(new ppq_comm);
    NcclId id;
    std::memcpy(id.internal, id128, 128);
    NC(((InitRankFn)g_nccl.CommInitRank)(&m->comm, world, id, rank));
    m->world = world;
    m->rank = rank;
    m->device = c->device;
    m->owned = true;
    *out = m.release();
    return PPQ_OK;
}
extern "C" int ppq_comm_from_nccl(void* nccl_comm, int world, int rank, int device, ppq_comm** out) {
    if (!nccl_comm || !out || world < 1 || world > 15 || rank < 0 || rank >= world) return fail(PPQ_EINVAL, "ppq_comm_from_nccl: bad arguments");
    if (int rc = nccl_load()) return rc;
    std::unique_ptr<ppq_comm> m(new ppq_comm);
    m->comm = nccl_comm;
    m->world = world;
    m->rank = rank;
    m->device = device;
    *out = m.release();
    return PPQ_OK;
}
extern "C" void ppq_comm_destroy(ppq_comm* m) {
    if (!m) return;
    if (m->owned && m->comm && g_nccl.CommDestroy) {
        cudaSetDevice(m->device);
        g_nccl.CommDestroy(m->comm);
    }
    delete m;
}
// sum of the ranks' canonical partial aggregates, in place on `root` (all ranks when root < 0), then mod q
extern "C" int ppq_reduce_across_ranks(ppq_ctx* c, ppq_comm* m, int ell, uint64_t* d_acc, size_t n, int root,
                                       void* stream) {
    if (!c || !m || !d_acc) return fail(PPQ_EINVAL, "ppq_reduce_across_ranks: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_reduce_across_ranks: bad limb count");
    if (root >= m->world) return fail(PPQ_EINVAL, "ppq_reduce_across_ranks: bad root");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cnt = n * 2 * (size_t)ell * c->hp.N;
    if (m->world > 1) {
        if (root < 0) NC(g_nccl.AllReduce(d_acc, d_acc, cnt, kNcclUint64, kNcclSum, m->comm, st));
        else NC(g_nccl.Reduce(d_acc, d_acc, cnt, kNcclUint64, kNcclSum, root, m->comm, st));
    }
    if (m->world > 1 && (root < 0 || root == m->rank)) {
        const size_t step = 65535 / (2 * (size_t)ell) ? 65535 / (2 * (size_t)ell) : 1;
        for (size_t s0 = 0; s0 < n; s0 += step)
            launch_finalize_mod(c->k, (u64*)d_acc + s0 * 2 * ell * c->hp.N, ell, std::min(step, n - s0), st);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- shard planning
// Clients per rank for a job whose ranks do not all move data at the same rate (measured on the 8-GPU B200 box with all
// ranks copying: 23 GB/s host->device on GPUs 0-3, 35 GB/s on GPUs 4-7, profiles/r16_notes.md): cost[r] is the time rank r
// needs per client (e.g. one timed ppq_copy_ceiling_host_packed of an equal shard, divided by its clients); counts[r]
// minimises max_r counts[r] * cost[r] with sum counts = n_clients (greedy on the next finishing time, exact for this
// objective; ties go to the lower rank, so every rank that holds the same cost vector computes the same plan).
extern "C" int ppq_plan_shards(size_t n_clients, int world, const double* cost, size_t* counts) {
    if (world < 1 || !cost || !counts) return fail(PPQ_EINVAL, "ppq_plan_shards: bad argument");
    for (int r = 0; r < world; ++r)
        if (!(cost[r] > 0.0) || !std::isfinite(cost[r])) return fail(PPQ_EINVAL, "ppq_plan_shards: costs must be positive and finite");
    for (int r = 0; r < world; ++r) counts[r] = 0;
    for (size_t c = 0; c < n_clients; ++c) {
        int best = 0;
        for (int r = 1; r < world; ++r)
            if ((double)(counts[r] + 1) * cost[r] < (double)(counts[best] + 1) * cost[best]) best = r;
        ++counts[best];
    }
    return PPQ_OK;
}

// ------------------------------------------------------------------------------------------- packed residues
// Wire format of the *_host_packed calls: a residue of limb i needs bitlen(q_i) bits, not 64 (C3: 1142 of 1280 bits per
// coefficient), and the host->device copy is what bounds the end-to-end path.  Packed ciphertext = [2][ell] limbs, limb
// i = N residues as a little-endian bit stream of bitlen(q_i) bits each (N * bitlen / 64 words, N is a multiple of 64).
static bool pack_layout(const ppq_ctx* c, int ell, PackLay& lay) {
    if (!c || ell < 1 || ell > c->hp.L || ell > 64) return false;
    u32 off = 0;
    for (int i = 0; i < ell; ++i) {
        lay.off[i] = off;
        off += (u32)(c->hp.N / 64) * (c->h_limbs[i].sh + 1u);
    }
    for (int i = ell; i < 64; ++i) lay.off[i] = 0;
    lay.words = off;
    return true;
}
extern "C" size_t ppq_packed_words(const ppq_ctx* c, int ell) {
    PackLay lay;
    return pack_layout(c, ell, lay) ? (size_t)2 * lay.words : 0;
}
extern "C" int ppq_pack_host(const ppq_ctx* c, int ell, const uint64_t* h_in, uint64_t* h_packed, size_t n) {
    PackLay lay;
    if (!h_in || !h_packed || !pack_layout(c, ell, lay)) return fail(PPQ_EINVAL, "ppq_pack_host: bad argument");
    const size_t N = c->hp.N;
    const long rows = (long)(n * 2 * (size_t)ell);
#pragma omp parallel for schedule(static)
    for (long y = 0; y < rows; ++y) {
        const int i = (int)(y % ell);
        const unsigned w = c->h_limbs[i].sh + 1u;
        const uint64_t* src = h_in + (size_t)y * N;
        uint64_t* dst = h_packed + (size_t)(y / ell) * lay.words + lay.off[i];
        unsigned __int128 acc = 0;  // bits not yet written, low end first
        unsigned have = 0;
        for (size_t k = 0; k < N; ++k) {
            acc |= (unsigned __int128)src[k] << have;
            have += w;
            if (have >= 64) {
                *dst++ = (u64)acc;
                acc >>= 64;
                have -= 64;
            }
        }
    }
    return PPQ_OK;
}
extern "C" int ppq_unpack_host(const ppq_ctx* c, int ell, const uint64_t* h_packed, uint64_t* h_out, size_t n) {
    PackLay lay;
    if (!h_packed || !h_out || !pack_layout(c, ell, lay)) return fail(PPQ_EINVAL, "ppq_unpack_host: bad argument");
    const size_t N = c->hp.N;
    const long rows = (long)(n * 2 * (size_t)ell);
#pragma omp parallel for schedule(static)
    for (long y = 0; y < rows; ++y) {
        const int i = (int)(y % ell);
        const unsigned w = c->h_limbs[i].sh + 1u;
        const u64 mask = w == 64 ? ~0ull : ((1ull << w) - 1ull);
        const uint64_t* src = h_packed + (size_t)(y / ell) * lay.words + lay.off[i];
        uint64_t* dst = h_out + (size_t)y * N;
        unsigned __int128 acc = 0;
        unsigned have = 0;
        for (size_t k = 0; k < N; ++k) {
            if (have < w) {
                acc |= (unsigned __int128)(*src++) << have;
                have += 64;
            }
            dst[k] = (u64)acc & mask;
            acc >>= w;
            have -= w;
        }
    }
    return PPQ_OK;
}
static int pack_dev(ppq_ctx* c, int ell, const uint64_t* in, uint64_t* out, size_t n, void* stream, bool unpack,
                    const char* who) {
    PackLay lay;
    if (!in || !out || !pack_layout(c, ell, lay)) return fail(PPQ_EINVAL, std::string(who) + ": bad argument");
    if (n == 0) return PPQ_OK;
    CU(cudaSetDevice(c->device));
    const size_t ct = (size_t)2 * ell * c->hp.N, pk = (size_t)2 * lay.words, step = rows_chunk(2 * ell);
    for (size_t s0 = 0; s0 < n; s0 += step) {
        const size_t rows = std::min(step, n - s0) * 2 * ell;
        if (unpack) launch_unpack(c->k, (const u64*)in + s0 * pk, (u64*)out + s0 * ct, lay, ell, rows, (cudaStream_t)stream);
        else launch_pack(c->k, (const u64*)in + s0 * ct, (u64*)out + s0 * pk, lay, ell, rows, (cudaStream_t)stream);
    }
    CU(cudaGetLastError());
    return PPQ_OK;
}
extern "C" int ppq_unpack(ppq_ctx* c, int ell, const uint64_t* d_packed, uint64_t* d_out, size_t n, void* stream) {
    return pack_dev(c, ell, d_packed, d_out, n, stream, true, "ppq_unpack");
}
extern "C" int ppq_pack(ppq_ctx* c, int ell, const uint64_t* d_in, uint64_t* d_packed, size_t n, void* stream) {
    return pack_dev(c, ell, d_in, d_packed, n, stream, false, "ppq_pack");
}

// ------------------------------------------------------------------------------------------- host-buffer path
// Persistent per-context resources of the host-buffer calls: two device upload buffers, the device aggregate, two
// streams and the hand-over events.  Created on first use, reused while the shape fits, freed with the context:
// nothing is allocated inside a steady-state call and no early return can leak.
struct ppq_ctx::HostPath {
    DevBuf<u64> in[2], out;
    DevBuf<u64> pk_in[2], pk_out;  // packed staging of the *_host_packed calls
    cudaStream_t sc = nullptr, sx = nullptr;
    cudaEvent_t up[2] = {}, done[2] = {};
    bool made = false;
    ppq_agg* agg = nullptr;
    ~HostPath() {
        if (agg) ppq_agg_destroy(agg);
        if (made) {
            for (int i = 0; i < 2; ++i) {
                cudaEventDestroy(up[i]);
                cudaEventDestroy(done[i]);
            }
            cudaStreamDestroy(sc);
            cudaStreamDestroy(sx);
        }
    }
};
ppq_ctx::~ppq_ctx() {
    delete hostpath;
    for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
}

static int hostpath_get(ppq_ctx* c, int ell, size_t n_slots, size_t packed_blk, ppq_ctx::HostPath** out) {
    if (!c->hostpath) c->hostpath = new ppq_ctx::HostPath;
    ppq_ctx::HostPath* h = c->hostpath;
    if (!h->made) {
        CU(cudaStreamCreateWithFlags(&h->sc, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&h->sx, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            CU(cudaEventCreateWithFlags(&h->up[i], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->done[i], cudaEventDisableTiming));
        }
        h->made = true;
    }
    const size_t blk = (size_t)2 * ell * c->hp.N * n_slots;
    CU(h->in[0].ensure(blk));
    CU(h->in[1].ensure(blk));
    CU(h->out.ensure(blk));
    if (packed_blk) {
        CU(h->pk_in[0].ensure(packed_blk));
        CU(h->pk_in[1].ensure(packed_blk));
        CU(h->pk_out.ensure(packed_blk));
    }
    if (h->agg && (h->agg->ell != ell || h->agg->n_slots != n_slots || h->agg->count != 0)) {
        ppq_agg_destroy(h->agg);
        h->agg = nullptr;
    }
    if (!h->agg)
        if (int rc = ppq_agg_create(c, ell, n_slots, &h->agg)) return rc;
    *out = h;
    return PPQ_OK;
}

// copy_only: the same uploads/downloads on the same streams without any kernel (the box's copy ceiling for e2e)
// packed: h_cts / h_acc hold packed residues (ppq_packed_words per ciphertext); the upload is unpacked on the device
static int aggregate_host_impl(ppq_ctx* c, ppq_comm* comm, int root, const ppq_rekey* const* rekeys, int ell,
                               const uint64_t* const* h_cts, size_t n_clients, size_t n_slots, uint64_t* h_acc,
                               bool copy_only, bool packed = false) {
    if (!c || !h_cts || (!rekeys && !copy_only)) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host: bad limb count");
    if (!n_slots) return PPQ_OK;
    if (!n_clients && !comm) return PPQ_OK;
    for (size_t cl = 0; cl < n_clients; ++cl)
        if (!h_cts[cl] || (!copy_only && !rekeys[cl])) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host: null client entry");
    const bool is_root = !comm || root < 0 || comm->rank == root;
    if (is_root && !h_acc) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host: null output");
    CU(cudaSetDevice(c->device));
    ppq_ctx::HostPath* h;
    PackLay lay;
    if (packed && !pack_layout(c, ell, lay)) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host_packed: more than 64 limbs");
    const size_t pblk = packed ? (size_t)2 * lay.words * n_slots : 0;
    if (int rc = hostpath_get(c, ell, n_slots, pblk, &h)) return rc;
    const size_t blk = (size_t)2 * ell * c->hp.N * n_slots;
    int rc = PPQ_OK;
    cudaError_t ce = cudaSuccess;
#define HP(call) do { if (rc == PPQ_OK && ce == cudaSuccess) ce = (call); } while (0)
    // double-buffered upload: client c+1 is copied while client c is re-encrypted
    for (size_t cl = 0; cl < n_clients && rc == PPQ_OK && ce == cudaSuccess; ++cl) {
        const int b = (int)(cl & 1);
        if (cl >= 2) HP(cudaStreamWaitEvent(h->sc, h->done[b], 0));
        if (packed) HP(cudaMemcpyAsync(h->pk_in[b].p, h_cts[cl], pblk * sizeof(u64), cudaMemcpyHostToDevice, h->sc));
        else HP(cudaMemcpyAsync(h->in[b].p, h_cts[cl], blk * sizeof(u64), cudaMemcpyHostToDevice, h->sc));
        HP(cudaEventRecord(h->up[b], h->sc));
        HP(cudaStreamWaitEvent(h->sx, h->up[b], 0));
        if (packed && !copy_only && rc == PPQ_OK && ce == cudaSuccess)
            rc = ppq_unpack(c, ell, (const uint64_t*)h->pk_in[b].p, (uint64_t*)h->in[b].p, n_slots, h->sx);
        if (!copy_only && rc == PPQ_OK && ce == cudaSuccess) rc = ppq_agg_add(h->agg, rekeys[cl], (const uint64_t*)h->in[b].p, h->sx);
        HP(cudaEventRecord(h->done[b], h->sx));
    }
    if (!copy_only && rc == PPQ_OK && ce == cudaSuccess) {
        if (n_clients) rc = ppq_agg_finish(h->agg, (uint64_t*)h->out.p, h->sx);
        else HP(cudaMemsetAsync(h->out.p, 0, blk * sizeof(u64), h->sx));  // a rank without clients contributes zero
        if (rc == PPQ_OK && comm) rc = ppq_reduce_across_ranks(c, comm, ell, (uint64_t*)h->out.p, n_slots, root, h->sx);
    }
    if (is_root && packed) {
        if (!copy_only && rc == PPQ_OK && ce == cudaSuccess)
            rc = ppq_pack(c, ell, (const uint64_t*)h->out.p, (uint64_t*)h->pk_out.p, n_slots, h->sx);
        HP(cudaMemcpyAsync(h_acc, h->pk_out.p, pblk * sizeof(u64), cudaMemcpyDeviceToHost, h->sx));
    } else if (is_root) {
        HP(cudaMemcpyAsync(h_acc, h->out.p, blk * sizeof(u64), cudaMemcpyDeviceToHost, h->sx));
    }
#undef HP
    // always drain both streams: the caller's host buffers must not be in flight when this returns, error or not
    const cudaError_t e1 = cudaStreamSynchronize(h->sc), e2 = cudaStreamSynchronize(h->sx);
    if (rc != PPQ_OK) {
        if (h->agg) {  // leave no half-built sum behind
            ppq_agg_destroy(h->agg);
            h->agg = nullptr;
        }
        return rc;
    }
    if (ce != cudaSuccess) return fail(PPQ_ECUDA, std::string("ppq_reencrypt_aggregate_host: ") + cudaGetErrorString(ce));
    if (e1 != cudaSuccess || e2 != cudaSuccess)
        return fail(PPQ_ECUDA, std::string("ppq_reencrypt_aggregate_host: ") + cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
    return PPQ_OK;
}
extern "C" int ppq_reencrypt_aggregate_host(ppq_ctx* c, const ppq_rekey* const* rekeys, int ell,
                                            const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                            uint64_t* h_acc) {
    if (!rekeys || !h_acc) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host: null argument");
    return aggregate_host_impl(c, nullptr, 0, rekeys, ell, h_cts, n_clients, n_slots, h_acc, false);
}
extern "C" int ppq_reencrypt_aggregate_host_ranks(ppq_ctx* c, ppq_comm* comm, int root, const ppq_rekey* const* rekeys,
                                                  int ell, const uint64_t* const* h_cts, size_t n_clients,
                                                  size_t n_slots, uint64_t* h_acc) {
    if (!comm) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host_ranks: null communicator");
    return aggregate_host_impl(c, comm, root, rekeys, ell, h_cts, n_clients, n_slots, h_acc, false);
}
extern "C" int ppq_copy_ceiling_host(ppq_ctx* c, int ell, const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                     uint64_t* h_acc) {
    return aggregate_host_impl(c, nullptr, 0, nullptr, ell, h_cts, n_clients, n_slots, h_acc, true);
}
extern "C" int ppq_reencrypt_aggregate_host_packed(ppq_ctx* c, const ppq_rekey* const* rekeys, int ell,
                                                   const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                                   uint64_t* h_acc) {
    if (!rekeys || !h_acc) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host_packed: null argument");
    return aggregate_host_impl(c, nullptr, 0, rekeys, ell, h_cts, n_clients, n_slots, h_acc, false, true);
}
extern "C" int ppq_reencrypt_aggregate_host_packed_ranks(ppq_ctx* c, ppq_comm* comm, int root,
                                                         const ppq_rekey* const* rekeys, int ell,
                                                         const uint64_t* const* h_cts, size_t n_clients, size_t n_slots,
                                                         uint64_t* h_acc) {
    if (!comm) return fail(PPQ_EINVAL, "ppq_reencrypt_aggregate_host_packed_ranks: null communicator");
    return aggregate_host_impl(c, comm, root, rekeys, ell, h_cts, n_clients, n_slots, h_acc, false, true);
}
extern "C" int ppq_copy_ceiling_host_packed(ppq_ctx* c, int ell, const uint64_t* const* h_cts, size_t n_clients,
                                            size_t n_slots, uint64_t* h_acc) {
    return aggregate_host_impl(c, nullptr, 0, nullptr, ell, h_cts, n_clients, n_slots, h_acc, true, true);
}

// ------------------------------------------------------------------------------------------- allocation-free Encrypt
// ppq_encrypt for callers that may not allocate device memory (CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.  The scratch of
// ppq_encrypt -- (v, e0, e1) in EVAL form, [n][3][ell][N] words, and the limb-id table of its ROW pass -- lives in ONE
// file-scope device array of PPQ_NOALLOC_SCRATCH_BYTES instead of context-owned cudaMalloc buffers.  Every GPU that loads
// the library holds the array until exit.  In host code `g_noalloc_scratch` is only a stand-in for the symbol: its
// device address is resolved per device with cudaGetSymbolAddress (below) and that pointer is what the kernels get.
#define PPQ_NOALLOC_SCRATCH_BYTES (64ull << 20)
#define PPQ_NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) unsigned char g_noalloc_scratch[PPQ_NOALLOC_SCRATCH_BYTES];
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[PPQ_NOALLOC_MAX_DEVICES] = {};

// ids[i] = i % ell: the limb of polynomial row i of [n][3][ell] (what cyc_ids uploads for the allocating path)
static __global__ void k_noalloc_fill_ids(int* __restrict__ ids, int n, int ell) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ids[i] = i % ell;
}

// once per device, with that device current
static int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= PPQ_NOALLOC_MAX_DEVICES) return fail(PPQ_EINVAL, "ppq_encrypt_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(PPQ_ECUDA, std::string("ppq_encrypt_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[device];
    return PPQ_OK;
}

extern "C" int ppq_encrypt_noalloc(ppq_ctx* c, int ell, const uint64_t* pkb, const uint64_t* pka, const uint64_t* pt,
                                   const int64_t* small, uint64_t* ct, size_t n, void* stream) {
    if (!c || !pkb || !pka || !pt || !small || !ct) return fail(PPQ_EINVAL, "ppq_encrypt_noalloc: null argument");
    if (ell < 1 || ell > c->hp.L) return fail(PPQ_EINVAL, "ppq_encrypt_noalloc: bad limb count");
    if (n == 0) return PPQ_OK;
    const size_t N = c->hp.N;
    const size_t rows = n * 3 * (size_t)ell;
    // layout: [rows][N] u64 (v, e0, e1 in EVAL), then [rows] int limb ids (both offsets are multiples of 256 bytes)
    const size_t need = rows * N * sizeof(u64) + rows * sizeof(int);
    if (n > 65535 / (3 * (size_t)ell) || need > PPQ_NOALLOC_SCRATCH_BYTES)
        return fail(PPQ_EINVAL, "ppq_encrypt_noalloc: batch needs " + std::to_string(need) + " bytes of scratch, the maximum is " +
                                    std::to_string((unsigned long long)PPQ_NOALLOC_SCRATCH_BYTES) +
                                    " (n * 3 * ell * (8 N + 4)); split the batch or use ppq_encrypt");
    CU(cudaSetDevice(c->device));
    unsigned char* base = nullptr;
    if (int rc = noalloc_scratch(c->device, &base)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    u64* vee = reinterpret_cast<u64*>(base);
    int* ids = reinterpret_cast<int*>(base + rows * N * sizeof(u64));
    // nothing in the array has to start at zero: the ids are rewritten here and every word of vee is written by the COL
    // pass before the ROW pass reads it, so the previous call's contents are never seen
    k_noalloc_fill_ids<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(ids, (int)rows, ell);
    launch_small_fwd_ntt(c->k, small, vee, 3, ell, n, ids, st);
    launch_encrypt_pw(c->k, (const u64*)pkb, (const u64*)pka, vee, (const u64*)pt, (u64*)ct, ell, n, st);
    CU(cudaGetLastError());
    return PPQ_OK;
}
