// engine_ctx.cu — context lifetime: parameter tables in HBM, level plans, stream-ordered allocation.
#include <cstring>
#include <random>
#include <stdexcept>

#include <algorithm>

#include <sys/random.h>

#include <cstdio>
#include <cstring>

#include "engine.h"

namespace fheb200 {

static thread_local const Context *tls_ctx = nullptr;
static thread_local int tls_idx = 0;
void Context::bind_thread(int i) { tls_ctx = this, tls_idx = i; }
void Context::unbind_thread() {
    if (tls_ctx == this) tls_ctx = nullptr;
}
void Context::bind_thread_stream() {
    if (!forked_) return;
    if (tls_ctx == this) stream = tls_idx < 0 ? main_stream_ : side_[tls_idx % fork_k_];
    else stream = fork_current_;
}

DevMem::~DevMem() {
    if (!p || !ctx) return;
    // handles are dropped by whichever host thread holds the last reference: take the context's lock (and that thread's stream)
    // unless the context itself is going away
    std::unique_ptr<CtxLock> lock;
    if (keep) lock.reset(new CtxLock(*ctx));
    cudaStream_t s = stream ? stream : ctx->stream;
    // Inside a forked region a buffer that existed before the fork is readable by EVERY side stream, so no single event
    // orders its release: park it; join() frees it on the main stream once all side streams have been joined.
    if (ctx->park_if_shared(p, epoch)) return;
    if (s != ctx->stream && ctx->free_ev_) { // last used on another stream than the one it lives on: free after that use
        cudaEventRecord(ctx->free_ev_, ctx->stream);
        cudaStreamWaitEvent(s, ctx->free_ev_, 0);
    }
    cudaFreeAsync(p, s);
}

Mem Context::alloc(size_t words) {
    auto m = std::make_shared<DevMem>();
    m->ctx = this;
    m->keep = weak_from_this().lock(); // null only while the context is being constructed or destroyed
    m->words = words;
    m->stream = stream, m->epoch = epoch;
    FHE_CUDA(cudaMallocAsync((void **)&m->p, words * sizeof(u64), stream));
    return m;
}
void Context::sync() {
    FHE_CUDA(cudaStreamSynchronize(stream));
    if (copy_stream_) {
        FHE_CUDA(cudaStreamSynchronize(copy_stream_));
        FHE_CUDA(cudaStreamSynchronize(out_stream_));
        CtxLock g(*this);
        drain_copies(true);
    }
    if (forked_)
        for (int i = 0; i < fork_k_; i++) FHE_CUDA(cudaStreamSynchronize(side_[i]));
}

cudaEvent_t Context::take_event() {
    if (!free_events_.empty()) {
        cudaEvent_t e = free_events_.back();
        free_events_.pop_back();
        return e;
    }
    cudaEvent_t e;
    FHE_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    return e;
}
void Context::drain_copies(bool wait) {
    while (!pending_.empty()) {
        PendingCopy &c = pending_.front();
        if (wait) FHE_CUDA(cudaEventSynchronize(c.done));
        else if (cudaEventQuery(c.done) != cudaSuccess) break;
        auto it = inbound_.find(c.keep.get());
        if (it != inbound_.end() && it->second == c.done) inbound_.erase(it); // complete: nothing left to wait for
        free_events_.push_back(c.done);
        pending_.pop_front(); // releases the buffer (stream-ordered free on the stream it lives on)
    }
}
void Context::copy_in(void *dst, const void *src, size_t bytes, const Mem &owner) {
    CtxLock g(*this);
    if (!copy_stream_) {
        FHE_CUDA(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
        FHE_CUDA(cudaStreamCreateWithFlags(&out_stream_, cudaStreamNonBlocking));
    }
    drain_copies(false);
    cudaEvent_t ready = take_event(), done = take_event();
    FHE_CUDA(cudaEventRecord(ready, stream)); // the allocation (and whatever used the memory before) is ordered on `stream`
    FHE_CUDA(cudaStreamWaitEvent(copy_stream_, ready, 0));
    FHE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, copy_stream_));
    FHE_CUDA(cudaEventRecord(done, copy_stream_));
    free_events_.push_back(ready);
    pending_.push_back(PendingCopy{done, owner});
    inbound_[owner.get()] = done;
}
void Context::wait_copy_in(const Mem &owner) {
    CtxLock g(*this);
    auto it = inbound_.find(owner.get());
    if (it == inbound_.end()) return;
    FHE_CUDA(cudaStreamWaitEvent(stream, it->second, 0));
    inbound_.erase(it);
}
void Context::copy_out(void *dst, const void *src, size_t bytes, const Mem &owner) {
    CtxLock g(*this);
    if (!copy_stream_) {
        FHE_CUDA(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
        FHE_CUDA(cudaStreamCreateWithFlags(&out_stream_, cudaStreamNonBlocking));
    }
    drain_copies(false);
    cudaEvent_t ready = take_event(), done = take_event();
    FHE_CUDA(cudaEventRecord(ready, stream));
    FHE_CUDA(cudaStreamWaitEvent(out_stream_, ready, 0));
    FHE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, out_stream_));
    FHE_CUDA(cudaEventRecord(done, out_stream_));
    free_events_.push_back(ready);
    pending_.push_back(PendingCopy{done, owner}); // the source stays allocated until the copy has run
}

void Context::fork(int k) {
    CtxLock g(*this);
    if (forked_) throw std::runtime_error("fork: already forked");
    if (k < 1 || k > 16) throw std::invalid_argument("fork: 1..16 streams");
    FHE_CUDA(cudaSetDevice(device));
    if (!fork_ev_) {
        FHE_CUDA(cudaEventCreateWithFlags(&fork_ev_, cudaEventDisableTiming));
        FHE_CUDA(cudaEventCreateWithFlags(&free_ev_, cudaEventDisableTiming));
    }
    while ((int)side_.size() < k) {
        cudaStream_t s;
        cudaEvent_t e;
        FHE_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        FHE_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        side_.push_back(s), side_done_.push_back(e);
    }
    main_stream_ = fork_current_ = stream;
    FHE_CUDA(cudaEventRecord(fork_ev_, main_stream_));
    for (int i = 0; i < k; i++) FHE_CUDA(cudaStreamWaitEvent(side_[i], fork_ev_, 0));
    fork_k_ = k, forked_ = true, fork_epoch_ = ++epoch;
}
void Context::use_stream(int i) {
    CtxLock g(*this);
    if (!forked_) throw std::runtime_error("use_stream: not forked");
    stream = fork_current_ = i < 0 ? main_stream_ : side_[i % fork_k_];
}
void Context::join() {
    CtxLock g(*this);
    if (!forked_) return;
    for (int i = 0; i < fork_k_; i++) {
        FHE_CUDA(cudaEventRecord(side_done_[i], side_[i]));
        FHE_CUDA(cudaStreamWaitEvent(main_stream_, side_done_[i], 0));
    }
    stream = main_stream_;
    forked_ = false, ++epoch;
    std::vector<u64 *> parked;
    {
        std::lock_guard<std::mutex> pg(park_mu_);
        parked.swap(parked_);
    }
    for (u64 *q : parked) cudaFreeAsync(q, main_stream_); // ordered after every side stream's last kernel
}
bool Context::park_if_shared(u64 *q, unsigned buf_epoch) {
    std::lock_guard<std::mutex> pg(park_mu_);
    if (!forked_ || buf_epoch >= fork_epoch_) return false;
    parked_.push_back(q);
    return true;
}

void *Context::owned_alloc(size_t bytes, const void *init) {
    void *d = nullptr;
    FHE_CUDA(cudaMalloc(&d, bytes));
    if (init) FHE_CUDA(cudaMemcpy(d, init, bytes, cudaMemcpyHostToDevice));
    owned_.push_back(d);
    return d;
}

std::shared_ptr<Context> Context::create(int logn, int depth, int scale_bits, int first_bits, int dnum, int batch,
                                         int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        throw std::runtime_error("fheb200: no CUDA device available (this engine has no CPU fallback)");
    if (device < 0 || device >= count) throw std::invalid_argument("fheb200: bad device ordinal");
    auto C = std::make_shared<Context>();
    C->P = Params::create(logn, depth, scale_bits, first_bits, dnum);
    const Params &P = *C->P;
    if (P.L + P.K > MAX_LIMBS) throw std::invalid_argument("fheb200: L+K exceeds MAX_LIMBS");
    if (batch <= 0) batch = P.n / 2;
    if (batch > P.n / 2 || (batch & (batch - 1))) throw std::invalid_argument("fheb200: batch size must be a power of two <= n/2");
    C->batch = batch;
    C->device = device;
    FHE_CUDA(cudaSetDevice(device));
    FHE_CUDA(cudaStreamCreateWithFlags(&C->stream, cudaStreamNonBlocking));
    cudaMemPool_t pool;
    FHE_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t keep = UINT64_MAX;
    FHE_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    C->reseed_from_os();

    const int F = P.L + P.K, n = P.n;
    DeviceTables &T = C->T;
    T.logn = logn, T.n = n, T.F = F;
    T.mc = (ModConst *)C->owned_alloc(sizeof(ModConst) * F, P.mc.data());
    std::vector<u64> ninv(2 * F);
    for (int i = 0; i < F; i++) ninv[i] = P.n_inv[i], ninv[F + i] = P.n_inv_s[i];
    T.ninv = (u64 *)C->owned_alloc(sizeof(u64) * 2 * F, ninv.data());
    size_t tb = sizeof(ulonglong2) * (size_t)F * n;
    T.tw = (ulonglong2 *)C->owned_alloc(tb, nullptr);
    T.twi = (ulonglong2 *)C->owned_alloc(tb, nullptr);
    T.twf = (double *)C->owned_alloc(sizeof(double) * (size_t)F * n, nullptr);
    T.twif = (double *)C->owned_alloc(sizeof(double) * (size_t)F * n, nullptr);
    std::vector<double> qd(2 * (size_t)F);
    for (int i = 0; i < F; i++) qd[2 * i] = (double)P.moduli[i], qd[2 * i + 1] = (double)(1.0L / (long double)P.moduli[i]);
    T.qd = (double2 *)C->owned_alloc(sizeof(double) * qd.size(), qd.data());
    ntt_register_moduli(T, P.mc.data());
    std::vector<std::vector<double>> fwd(F), iwd(F);
    std::vector<std::vector<u64>> fw(F), iw(F);
#pragma omp parallel for schedule(dynamic)
    for (int i = 0; i < F; i++) {
        std::vector<u64> w, ws, wi, wis;
        P.twiddles(i, w, ws, wi, wis);
        fw[i].resize(2 * (size_t)n), iw[i].resize(2 * (size_t)n);
        fwd[i].resize(n), iwd[i].resize(n);
        for (int k = 0; k < n; k++) fwd[i][k] = (double)w[k], iwd[i][k] = (double)wi[k];
        for (int k = 0; k < n; k++) {
            fw[i][2 * k] = w[k], fw[i][2 * k + 1] = ws[k];
            iw[i][2 * k] = wi[k], iw[i][2 * k + 1] = wis[k];
        }
    }
    // factored twiddles of the last three stages (n = 2^17, FP64-path moduli): twa[s - 14][T] = tw[2^s + T 2^(s-6)] / tw[2^s];
    // kept only if every table entry is reproduced exactly (FHEB200_NTT_TWFACTOR=0 disables)
    {
        const char *e = getenv("FHEB200_NTT_TWFACTOR");
        if (logn == 17 && !(e && *e == '0')) {
            std::vector<double> fa((size_t)F * 192, 0.0), ia((size_t)F * 192, 0.0);
            bool ok = true;
            for (int i = 0; i < F && ok; i++) {
                if (P.mc[i].bits > 42) continue; // integer path: tables unused
                const u64 q = P.moduli[i];
                for (int dir = 0; dir < 2 && ok; dir++) {
                    const std::vector<u64> &tb = dir ? iw[i] : fw[i]; // (w, Shoup) pairs
                    for (int s = 14; s <= 16 && ok; s++) {
                        const u64 inv0 = invmod_h(tb[2 * ((size_t)1 << s)], q);
                        for (int t = 0; t < 64 && ok; t++) {
                            const u64 a = mulmod_h(tb[2 * (((size_t)1 << s) + ((size_t)t << (s - 6)))], inv0, q);
                            (dir ? ia : fa)[(size_t)i * 192 + (s - 14) * 64 + t] = (double)a;
                            for (size_t r = 0; r < ((size_t)1 << (s - 6)); r++)
                                if (mulmod_h(tb[2 * (((size_t)1 << s) + r)], a, q) != tb[2 * (((size_t)1 << s) + ((size_t)t << (s - 6)) + r)]) {
                                    ok = false;
                                    break;
                                }
                        }
                    }
                }
            }
            if (ok) {
                T.twa = (double *)C->owned_alloc(sizeof(double) * fa.size(), fa.data());
                T.twia = (double *)C->owned_alloc(sizeof(double) * ia.size(), ia.data());
            }
        }
    }
    for (int i = 0; i < F; i++) {
        size_t off = (size_t)i * n, b = sizeof(ulonglong2) * n;
        FHE_CUDA(cudaMemcpy(T.tw + off, fw[i].data(), b, cudaMemcpyHostToDevice));
        FHE_CUDA(cudaMemcpy(T.twi + off, iw[i].data(), b, cudaMemcpyHostToDevice));
        FHE_CUDA(cudaMemcpy(T.twf + off, fwd[i].data(), sizeof(double) * n, cudaMemcpyHostToDevice));
        FHE_CUDA(cudaMemcpy(T.twif + off, iwd[i].data(), sizeof(double) * n, cudaMemcpyHostToDevice));
    }
    return C;
}

Context::~Context() {
    if (forked_) stream = main_stream_;
    try {
        comm_destroy();
    } catch (...) {
    }
    if (copy_stream_) cudaStreamSynchronize(copy_stream_), cudaStreamSynchronize(out_stream_);
    for (auto &c : pending_) cudaEventDestroy(c.done);
    pending_.clear();
    inbound_.clear();
    for (auto e : free_events_) cudaEventDestroy(e);
    if (copy_stream_) cudaStreamDestroy(copy_stream_), cudaStreamDestroy(out_stream_);
    for (auto s : side_) cudaStreamSynchronize(s);
    for (u64 *q : parked_) cudaFree(q); // a forked region that was never joined
    parked_.clear();
    if (stream) cudaStreamSynchronize(stream);
    for (auto s : side_) cudaStreamDestroy(s);
    for (auto e : side_done_) cudaEventDestroy(e);
    if (fork_ev_) cudaEventDestroy(fork_ev_);
    if (free_ev_) cudaEventDestroy(free_ev_), free_ev_ = nullptr;
    if (evk_mult) cudaFree(evk_mult);
    for (auto &kv : evk_rot) cudaFree(kv.second);
    for (void *p : owned_) cudaFree(p);
    for (auto &s : stage_) {
        if (s.host) cudaFreeHost(s.host);
        if (s.done) cudaEventDestroy(s.done);
    }
    if (stream && own_stream) cudaStreamDestroy(stream);
}

// 64 bytes of OS entropy: getrandom(2), else /dev/urandom; never a weaker source
static void os_entropy(unsigned char *out, size_t len) {
    size_t got = 0;
    while (got < len) {
        ssize_t r = getrandom(out + got, len - got, 0);
        if (r <= 0) break;
        got += (size_t)r;
    }
    if (got < len) {
        FILE *f = fopen("/dev/urandom", "rb");
        if (f) {
            got += fread(out + got, 1, len - got, f);
            fclose(f);
        }
    }
    if (got < len) throw std::runtime_error("fheb200: no OS entropy source (getrandom and /dev/urandom both failed)");
}
void Context::reseed_from_os() {
    unsigned char k[64];
    os_entropy(k, sizeof(k));
    set_rng_keys(k);
    rng_pinned = false;
    memset(k, 0, sizeof(k));
}
void Context::set_rng_keys(const unsigned char *k64) {
    CtxLock g(*this);
    rng_secret = RngState(), rng_public = RngState();
    memcpy(rng_secret.key, k64, 32), memcpy(rng_public.key, k64 + 32, 32);
    rng_secret.secure = rng_public.secure = 1;
    rng_pinned = true;
    nonce = 1;
}
void Context::get_rng_keys(unsigned char *k64) const {
    if (!rng_secret.secure) throw std::runtime_error("fheb200: the context is in the deterministic test mode; it has no key material to share");
    memcpy(k64, rng_secret.key, 32), memcpy(k64 + 32, rng_public.key, 32);
}

size_t Context::key_bytes() const { return (evk_rot.size() + (evk_mult ? 1 : 0)) * evk_words() * sizeof(u64); }

const uint32_t *Context::desc(const std::vector<uint32_t> &d) {
    CtxLock g(*this);
    auto it = descs_.find(d);
    if (it != descs_.end()) return it->second;
    uint32_t *p = (uint32_t *)owned_alloc(sizeof(uint32_t) * d.size(), d.data());
    ntt_register_desc(p, d.data(), d.size());
    descs_[d] = p;
    return p;
}

// C[src * cstride + tgt_pos]; targets = (output slot, modulus index) pairs (tensor-core base conversion tables, baseconv.cu)
ConvTable Context::make_conv_table(const std::vector<u64> &C, int cstride, int nsrc, std::vector<std::pair<int, int>> tg,
                               const std::vector<int> &tpos) {
typedef unsigned __int128 u128;
auto bits_of = [&](int qp) { return P->mc[qp].bits; };
    // targets needing 8 byte planes first, so that as few column groups as possible carry them
    std::vector<int> order(tg.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return (bits_of(tg[a].second) > 48) > (bits_of(tg[b].second) > 48); });
    ConvTable t;
    t.nsrc = nsrc, t.KS = (nsrc + 3) / 4, t.NT = ((int)tg.size() + 7) / 8;
    std::vector<int> tslot(t.NT * 8, -1), tqp(t.NT * 8, 0);
    std::vector<unsigned char> planes(t.NT, 6);
    std::vector<uint32_t> B((size_t)t.KS * t.NT * 8 * 32 * 2, 0);
    for (size_t c = 0; c < order.size(); c++) {
        tslot[c] = tg[order[c]].first, tqp[c] = tg[order[c]].second;
        if (bits_of(tqp[c]) > 48) planes[c / 8] = 8;
    }
    for (int ks = 0; ks < t.KS; ks++)
        for (int nt = 0; nt < t.NT; nt++)
            for (int lane = 0; lane < 32; lane++) {
                const int g = lane >> 2, tq = lane & 3, col = 8 * nt + g;
                if (col >= (int)order.size()) continue;
                const u64 p = P->moduli[tqp[col]];
                for (int reg = 0; reg < 2; reg++) {
                    const int j = 4 * ks + 2 * reg + (tq >> 1);
                    if (j >= nsrc) continue;
                    const u64 c0 = C[(size_t)j * cstride + tpos[order[col]]] % p;
                    for (int i = 0; i < 4; i++) {
                        const int a = 4 * (tq & 1) + i;
                        const u64 cp = (u64)(((u128)c0 << (8 * a)) % p);
                        for (int b = 0; b < 8; b++)
                            B[(((size_t)(ks * t.NT + nt) * 8 + b) * 32 + lane) * 2 + reg] |= (uint32_t)((cp >> (8 * b)) & 0xff) << (8 * i);
                    }
                }
            }
    t.B = (const uint2 *)owned_alloc(sizeof(uint32_t) * B.size(), B.data());
    t.tslot = (const int *)owned_alloc(sizeof(int) * tslot.size(), tslot.data());
    std::vector<uint32_t> tconst(tqp.size() * 4, 0);
    for (size_t c = 0; c < order.size(); c++) {
        const u64 p = P->moduli[tqp[c]];
        const int bits = bits_of(tqp[c]);
        if (bits < 17 || bits > 62) throw std::runtime_error("fheb200: modulus size outside the base conversion's range");
        tconst[4 * c] = (uint32_t)p, tconst[4 * c + 1] = (uint32_t)(p >> 32);
        tconst[4 * c + 2] = (uint32_t)(((u128)1 << (bits + 31)) / p), tconst[4 * c + 3] = (uint32_t)bits;
    }
    t.tconst = (const uint4 *)owned_alloc(sizeof(uint32_t) * tconst.size(), tconst.data());
    t.planes = (const unsigned char *)owned_alloc(planes.size(), planes.data());
    if (unsigned char *b5 = bc5_make_table(t.B, t.KS, t.NT)) owned_.push_back(b5), t.B5 = b5;
    return t;
}

const DevicePlan &Context::plan(int ell) {
    CtxLock g(*this);
    auto it = plans_.find(ell);
    if (it != plans_.end()) return it->second;
    const LevelPlan &H = P->plan(ell);
    if (H.beta > 8) throw std::runtime_error("fheb200: more than 8 digits unsupported");
    DevicePlan D;
    D.ell = ell, D.beta = H.beta, D.width = H.width, D.K = P->K;
    auto up64 = [&](const std::vector<u64> &v) { return (u64 *)owned_alloc(sizeof(u64) * v.size(), v.data()); };
    // [src][tgt] u64 table -> [tgt][src padded to even] (lo30, hi30) pairs for the split dot products
    auto split = [&](const std::vector<u64> &v, int nsrc, int ntgt) {
        const int sp = nsrc + (nsrc & 1);
        std::vector<uint32_t> o((size_t)ntgt * sp * 2, 0);
        for (int i = 0; i < nsrc; i++)
            for (int t = 0; t < ntgt; t++) {
                u64 h = v[(size_t)i * ntgt + t];
                o[((size_t)t * sp + i) * 2] = (uint32_t)(h & 0x3fffffffu), o[((size_t)t * sp + i) * 2 + 1] = (uint32_t)(h >> 30);
            }
        return (uint2 *)owned_alloc(sizeof(uint32_t) * o.size(), o.data());
    };
    D.qp = (int *)owned_alloc(sizeof(int) * H.qp.size(), H.qp.data());
    for (int j = 0; j < H.beta; j++) {
        D.dstart[j] = H.dstart[j], D.dsize[j] = H.dsize[j];
        D.hat_inv[j] = up64(H.hat_inv[j]), D.hat_inv_s[j] = up64(H.hat_inv_s[j]);
        D.hat2[j] = split(H.hat[j], H.dsize[j], H.width);
    }
    D.p_hat_inv = up64(H.p_hat_inv), D.p_hat_inv_s = up64(H.p_hat_inv_s);
    D.p_hat2 = split(H.p_hat, P->K, ell);
    D.p_inv = up64(H.p_inv), D.p_inv_s = up64(H.p_inv_s);
    std::vector<uint32_t> dq, dm, dp2, dq2;
    for (int i = 0; i < ell; i++) dq.push_back(((uint32_t)i << 8) | (uint32_t)i);
    // batch order = launch order of the CTAs: limbs of equal modulus are adjacent so that their twiddles are read from
    // HBM once and then hit in L2
    // ... and the 60-bit P limbs (integer path, 1.6x the time of an FP64 limb) go first so that they do not form the tail
    for (int tt = 0; tt < H.width; tt++) {
        const int t = tt < P->K ? ell + tt : tt - P->K;
        for (int j = 0; j < H.beta; j++)
            if (t < H.dstart[j] || t >= H.dstart[j] + H.dsize[j])
                dm.push_back(((uint32_t)(j * H.width + t) << 8) | (uint32_t)H.qp[t]);
    }
    for (int k = 0; k < P->K; k++) // (same order as DevicePlan::down_scale)
        for (int b = 0; b < 2; b++) dp2.push_back(((uint32_t)(b * H.width + ell + k) << 8) | (uint32_t)(P->L + k));
    for (int i = 0; i < ell; i++)
        for (int b = 0; b < 2; b++) dq2.push_back(((uint32_t)(b * ell + i) << 8) | (uint32_t)i);
    typedef unsigned __int128 u128;
    auto shoup_of = [&](u64 v, u64 q) { return (u64)(((u128)v << 64) / q); };
    auto mulmod = [&](u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); };
    {
        std::vector<u64> up(2 * (size_t)ell, 0), down(4 * (size_t)P->K, 0);
        for (int j = 0; j < H.beta; j++) {
            std::vector<std::pair<int, int>> tg;
            std::vector<int> tpos;
            for (int t = 0; t < H.width; t++)
                if (t < H.dstart[j] || t >= H.dstart[j] + H.dsize[j]) tg.push_back({t, H.qp[t]}), tpos.push_back(t);
            D.up[j] = make_conv_table(H.hat[j], H.width, H.dsize[j], tg, tpos);
            for (int i = 0; i < H.dsize[j]; i++) {
                const int limb = H.dstart[j] + i;
                const u64 q = P->moduli[limb];
                up[limb] = mulmod(P->n_inv[limb], H.hat_inv[j][i], q), up[ell + limb] = shoup_of(up[limb], q);
            }
        }
        std::vector<std::pair<int, int>> tg;
        std::vector<int> tpos;
        for (int i = 0; i < ell; i++) tg.push_back({i, i}), tpos.push_back(i);
        D.down = make_conv_table(H.p_hat, ell, P->K, tg, tpos);
        for (int k = 0; k < P->K; k++)
            for (int b = 0; b < 2; b++) { // entry order of desc_p2
                const u64 q = P->moduli[P->L + k];
                const u64 v = mulmod(P->n_inv[P->L + k], H.p_hat_inv[k], q);
                down[2 * k + b] = v, down[2 * P->K + 2 * k + b] = shoup_of(v, q);
            }
        D.up_scale = (u64 *)owned_alloc(sizeof(u64) * up.size(), up.data());
        D.down_scale = (u64 *)owned_alloc(sizeof(u64) * down.size(), down.data());
    }
    D.desc_q = (uint32_t *)desc(dq);
    D.desc_modup = (uint32_t *)desc(dm);
    D.n_modup = (int)dm.size();
    D.desc_p2 = (uint32_t *)desc(dp2);
    D.desc_q2 = (uint32_t *)desc(dq2);
    return plans_.emplace(ell, D).first->second;
}

// ModDown by R = P * q_{ell-1} (relinearise-and-rescale in one division, engine_arith.cu: relinearize_reduced): the same ModDown
// routines run on a plan whose "special" limbs are q_{ell-1} followed by the K primes of P (positions ell-1 .. ell+K-1 of an
// accumulator, contiguous) and whose ordinary limbs are q_0 .. q_{ell-2}.  Only the fields ModDown reads are filled.
const DevicePlan &Context::plan_merged(int ell) {
    CtxLock g(*this);
    auto it = plans_m_.find(ell);
    if (it != plans_m_.end()) return it->second;
    if (ell < 2) throw std::out_of_range("merged ModDown plan: needs at least two limbs");
    const LevelPlan &H = P->plan(ell);
    const int K1 = P->K + 1, le = ell - 1;
    std::vector<int> sp; // modulus index of special limb j
    sp.push_back(ell - 1);
    for (int k = 0; k < P->K; k++) sp.push_back(P->L + k);
    auto prod_except = [&](int skip, u64 mod) { // product of the special moduli except sp[skip] (skip < 0: all), modulo mod
        u64 r = 1;
        for (int x = 0; x < K1; x++)
            if (x != skip) r = mulmod_h(r, P->moduli[sp[x]] % mod, mod);
        return r;
    };
    DevicePlan D;
    D.ell = le, D.beta = 0, D.width = H.width, D.K = K1;
    D.qp = (int *)owned_alloc(sizeof(int) * H.qp.size(), H.qp.data());
    // conversion table with one more source: u = round(sum_j y_j / r_j) times (-R mod q_i) makes the lift exact (centred)
    std::vector<u64> r_hat((size_t)(K1 + 1) * le), r_inv(le), r_inv_s(le), down(4 * (size_t)K1, 0);
    std::vector<double> recip(K1);
    for (int j = 0; j < K1; j++) {
        const u64 rj = P->moduli[sp[j]];
        const u64 hat_inv = invmod_h(prod_except(j, rj), rj);
        const u64 v = mulmod_h(P->n_inv[sp[j]], hat_inv, rj);
        for (int b = 0; b < 2; b++) down[2 * j + b] = v, down[2 * K1 + 2 * j + b] = shoup_h(v, rj); // entry order of desc_p2
        for (int i = 0; i < le; i++) r_hat[(size_t)j * le + i] = prod_except(j, P->moduli[i]);
        recip[j] = (double)(1.0L / (long double)rj);
    }
    for (int i = 0; i < le; i++) {
        const u64 qi = P->moduli[i], R = prod_except(-1, qi);
        r_hat[(size_t)K1 * le + i] = R ? qi - R : 0;
    }
    D.exact_lift = 1;
    D.r_recip = (double *)owned_alloc(sizeof(double) * K1, recip.data());
    for (int i = 0; i < le; i++) {
        const u64 qi = P->moduli[i];
        r_inv[i] = invmod_h(prod_except(-1, qi), qi), r_inv_s[i] = shoup_h(r_inv[i], qi);
    }
    D.p_inv = (u64 *)owned_alloc(sizeof(u64) * le, r_inv.data());
    D.p_inv_s = (u64 *)owned_alloc(sizeof(u64) * le, r_inv_s.data());
    D.down_scale = (u64 *)owned_alloc(sizeof(u64) * down.size(), down.data());
    std::vector<std::pair<int, int>> tg;
    std::vector<int> tpos;
    for (int i = 0; i < le; i++) tg.push_back({i, i}), tpos.push_back(i);
    D.down = make_conv_table(r_hat, le, K1 + 1, tg, tpos);
    std::vector<uint32_t> dp2, dq2;
    for (int j = 0; j < K1; j++)
        for (int b = 0; b < 2; b++) dp2.push_back(((uint32_t)(b * H.width + le + j) << 8) | (uint32_t)sp[j]);
    for (int i = 0; i < le; i++)
        for (int b = 0; b < 2; b++) dq2.push_back(((uint32_t)(b * le + i) << 8) | (uint32_t)i);
    D.desc_p2 = (uint32_t *)desc(dp2);
    D.desc_q2 = (uint32_t *)desc(dq2);
    return plans_m_.emplace(ell, D).first->second;
}

LimbScalars Context::scalars_for(long double value, int ell, bool shoup) const {
    // round(value) as a signed 128-bit integer, reduced modulo each active modulus
    LimbScalars s;
    memset(&s, 0, sizeof(s));
    bool neg = value < 0;
    long double a = neg ? -value : value;
    u128 mag;
    if (a < 9.0e18L) {
        mag = (u128)(u64)llroundl(a);
    } else { // >= 2^62.9: the 64-bit mantissa already holds an integer
        int ex;
        long double mant = frexpl(a, &ex);
        u64 mi = (u64)ldexpl(mant, 64);
        if (ex > 126) throw std::overflow_error("fheb200: scaled constant exceeds 126 bits");
        mag = ex >= 64 ? ((u128)mi << (ex - 64)) : ((u128)mi >> (64 - ex));
    }
    for (int i = 0; i < ell; i++) {
        u64 q = P->moduli[i];
        u64 r = (u64)(mag % q);
        if (neg && r) r = q - r;
        s.v[i] = r;
        s.s[i] = shoup ? shoup_h(r, q) : 0;
    }
    return s;
}

} // namespace fheb200
