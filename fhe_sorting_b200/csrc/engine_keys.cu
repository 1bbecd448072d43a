// engine_keys.cu — key generation (device resident), CKKS encode/decode, encrypt, decrypt.
// Reference call sites: KeyGen / EvalMultKeyGen / EvalRotateKeyGen (tests/DirectSortTest.cpp:43-47),
// MakeCKKSPackedPlaintext + Encrypt (src/encryption.cpp:5-12), Decrypt + GetRealPackedValue (src/encryption.cpp:14-27).
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "engine.h"

namespace fheb200 {

enum { ST_SK = 1, ST_PK_A = 2, ST_PK_E = 3, ST_ENC_V = 4, ST_ENC_E0 = 5, ST_ENC_E1 = 6, ST_EVK_A = 7, ST_EVK_E = 8 };
static inline u64 stream_id(u64 tag, u64 digit, u64 id) { return tag | (digit << 8) | (id << 16); }

static std::vector<uint32_t> iota_desc(int count, int slot0 = 0, int qp0 = 0) {
    std::vector<uint32_t> d;
    for (int i = 0; i < count; i++) d.push_back(((uint32_t)(slot0 + i) << 8) | (uint32_t)(qp0 + i));
    return d;
}

Keys Context::keygen(u64 sd) {
    CtxLock g(*this);
    FHE_CUDA(cudaSetDevice(device));
    if (sd != 0) { // TEST-ONLY: 64-bit seeded, non-cryptographic generator shared with the CPU oracle
        if (!allow_deterministic)
            throw std::runtime_error("KeyGen: a non-zero seed selects the deterministic test generator, which is refused unless "
                                     "fheb200_ctx_allow_deterministic_keys(ctx, 1) was called; pass seed 0 for ChaCha20 keyed from OS entropy");
        rng_secret = RngState(), rng_public = RngState();
        rng_secret.seed = rng_public.seed = sd;
        nonce = 1;
    } else if (!rng_pinned) {
        reseed_from_os(); // every key pair gets its own 2 x 256 bits
    }
    const int n = P->n, L = P->L, F = L + P->K;
    auto K = std::make_shared<KeySet>();
    K->s = alloc((size_t)F * n);
    K->pk = alloc((size_t)2 * L * n);
    Mem small = alloc((size_t)n / 2 + 1); // n ints
    int *sm = (int *)small->p;
    const uint32_t *dF = desc(iota_desc(F)), *dL = desc(iota_desc(L));
    k_sample_small(T, SMALL_TERNARY, rng_secret, stream_id(ST_SK, 0, 0), sm, stream);
    k_expand_small(T, sm, K->s->p, dF, F, stream);
    k_ntt_fwd(T, K->s->p, dF, F, stream);
    // public key: b = e - a s over the Q limbs
    u64 *b = K->pk->p, *a = K->pk->p + (size_t)L * n;
    k_sample_uniform(T, rng_public, stream_id(ST_PK_A, 0, 0), 0, L, a, stream);
    k_sample_small(T, SMALL_GAUSS, rng_secret, stream_id(ST_PK_E, 0, 0), sm, stream);
    k_expand_small(T, sm, b, dL, L, stream);
    k_ntt_fwd(T, b, dL, L, stream);
    LimbScalars zero;
    memset(&zero, 0, sizeof(zero));
    k_evk_finish(T, L, a, b, K->s->p, K->s->p, zero, b, stream);
    return K;
}

// one switching key old -> new: [digit][b|a][F][n]; b_j = e_j - a_j s_new + [limb in digit j] (P mod q_i) s_old
u64 *Context::gen_switch_key(u64 key_id, const u64 *s_old, const u64 *s_new) {
    const int n = P->n, L = P->L, K = P->K, F = L + K, beta = P->beta(L);
    u64 *evk = nullptr;
    FHE_CUDA(cudaMalloc((void **)&evk, evk_words() * sizeof(u64)));
    Mem small = alloc((size_t)n / 2 + 1);
    const uint32_t *dF = desc(iota_desc(F));
    for (int j = 0; j < beta; j++) {
        u64 *b = evk + ((size_t)j * 2 + 0) * F * n, *a = evk + ((size_t)j * 2 + 1) * F * n;
        k_sample_uniform(T, rng_public, stream_id(ST_EVK_A, j, key_id), 0, F, a, stream);
        k_sample_small(T, SMALL_GAUSS, rng_secret, stream_id(ST_EVK_E, j, key_id), (int *)small->p, stream);
        k_expand_small(T, (int *)small->p, b, dF, F, stream);
        k_ntt_fwd(T, b, dF, F, stream);
        LimbScalars pm;
        memset(&pm, 0, sizeof(pm));
        int s = j * P->alpha, d = std::min(P->alpha, L - s);
        for (int i = s; i < s + d; i++) {
            u64 q = P->moduli[i], v = 1;
            for (int x = 0; x < K; x++) v = mulmod_h(v, P->moduli[L + x] % q, q);
            pm.v[i] = v, pm.s[i] = shoup_h(v, q);
        }
        k_evk_finish(T, F, a, b, s_new, s_old, pm, b, stream);
    }
    return evk;
}

void Context::gen_mult_key(const Keys &k) {
    CtxLock g(*this);
    FHE_CUDA(cudaSetDevice(device));
    if (!k || !k->s) throw std::runtime_error("EvalMultKeyGen: secret key required");
    const int n = P->n, F = P->L + P->K;
    Mem s2 = alloc((size_t)F * n);
    // s^2 limb-wise over all F moduli: ew kernels index moduli by limb, and QP order == limb order here
    k_binary(T, OP_MUL, F, k->s->p, k->s->p, s2->p, stream);
    if (evk_mult) cudaFree(evk_mult);
    evk_mult = gen_switch_key(0, s2->p, k->s->p);
    sync();
}

void Context::gen_rot_keys(const Keys &k, const std::vector<int> &indices) {
    CtxLock g(*this);
    FHE_CUDA(cudaSetDevice(device));
    if (!k || !k->s) throw std::runtime_error("EvalRotateKeyGen: secret key required");
    const int n = P->n, F = P->L + P->K;
    Mem sp = alloc((size_t)F * n);
    for (int idx : indices) {
        uint32_t gal = P->galois_for_rotation(idx);
        if (gal == 1 || evk_rot.count(gal)) continue;
        // key switches s -> sigma_{g^-1}(s); the rotation applies sigma_g afterwards
        k_automorphism(T, P->galois_inverse(gal), F, k->s->p, sp->p, stream);
        evk_rot[gal] = gen_switch_key(gal, k->s->p, sp->p);
    }
    sync();
}

// ---------------------------------------------------------------- encode / decode (host special FFT)
static void bitrev_perm(std::vector<std::complex<double>> &v) {
    size_t sz = v.size();
    int lg = 0;
    while (((size_t)1 << lg) < sz) lg++;
    for (size_t i = 0; i < sz; i++) {
        size_t r = bitrev_h((uint32_t)i, lg);
        if (i < r) std::swap(v[i], v[r]);
    }
}
struct FftTables {
    std::vector<std::complex<double>> ksi; // exp(2 pi i k / M), M = 4*size
    std::vector<int> rot;                  // 5^j mod M
};
static const FftTables &fft_tables(int size) {
    static std::mutex m;
    static std::map<int, FftTables> cache;
    std::lock_guard<std::mutex> g(m);
    auto it = cache.find(size);
    if (it != cache.end()) return it->second;
    FftTables t;
    int M = 4 * size;
    t.ksi.resize(M + 1);
    for (int k = 0; k <= M; k++) {
        double ang = 2.0 * M_PI * (double)k / (double)M;
        t.ksi[k] = std::complex<double>(std::cos(ang), std::sin(ang));
    }
    t.rot.resize(size);
    long long p = 1;
    for (int j = 0; j < size; j++, p = p * 5 % M) t.rot[j] = (int)p;
    return cache.emplace(size, std::move(t)).first->second;
}
void fft_special(std::vector<std::complex<double>> &v) {
    int size = (int)v.size(), M = 4 * size;
    if (size < 2) return;
    const FftTables &t = fft_tables(size);
    bitrev_perm(v);
    for (int len = 2; len <= size; len <<= 1) {
        int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (int i = 0; i < size; i += len)
            for (int j = 0; j < lenh; j++) {
                std::complex<double> w = t.ksi[(t.rot[j] % lenq) * gap];
                std::complex<double> u = v[i + j], x = v[i + j + lenh] * w;
                v[i + j] = u + x;
                v[i + j + lenh] = u - x;
            }
    }
}
void fft_special_inv(std::vector<std::complex<double>> &v) {
    int size = (int)v.size(), M = 4 * size;
    if (size < 2) return;
    const FftTables &t = fft_tables(size);
    for (int len = size; len >= 2; len >>= 1) {
        int lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (int i = 0; i < size; i += len)
            for (int j = 0; j < lenh; j++) {
                std::complex<double> w = t.ksi[(lenq - (t.rot[j] % lenq)) * gap];
                std::complex<double> u = v[i + j] + v[i + j + lenh], x = (v[i + j] - v[i + j + lenh]) * w;
                v[i + j] = u;
                v[i + j + lenh] = x;
            }
    }
    bitrev_perm(v);
    for (auto &x : v) x /= (double)size;
}

static void to_i128_words(long double x, u64 &lo, u64 &hi) {
    bool neg = x < 0;
    long double a = neg ? -x : x;
    u128 mag;
    if (a < 9.0e18L) {
        mag = (u128)(u64)llroundl(a);
    } else {
        int ex;
        long double mant = frexpl(a, &ex);
        if (ex > 126) throw std::overflow_error("fheb200: encoded coefficient exceeds 126 bits");
        u64 mi = (u64)ldexpl(mant, 64);
        mag = ex >= 64 ? ((u128)mi << (ex - 64)) : ((u128)mi >> (64 - ex));
    }
    i128 v = neg ? -(i128)mag : (i128)mag;
    lo = (u64)(u128)v, hi = (u64)((u128)v >> 64);
}

std::pair<double2 *, int *> Context::fft_device_tables(int slots) {
    auto it = fft_tabs_.find(slots);
    if (it != fft_tabs_.end()) return it->second;
    const FftTables &t = fft_tables(slots);
    static_assert(sizeof(std::complex<double>) == sizeof(double2), "complex layout");
    double2 *ksi = (double2 *)owned_alloc(sizeof(double2) * t.ksi.size(), t.ksi.data());
    int *rot = (int *)owned_alloc(sizeof(int) * t.rot.size(), t.rot.data());
    return fft_tabs_[slots] = std::make_pair(ksi, rot);
}

// slots values -> special inverse FFT -> * scale, round -> residues on `ell` limbs -> NTT; everything after the upload of
// the slot values runs on the device (encode.cu)
void Context::upload(void *dst, const void *src, size_t bytes) {
    if (stage_.empty()) {
        stage_bytes_ = sizeof(double2) * (size_t)(P->n / 2);
        stage_.resize(8);
        for (auto &s : stage_) {
            FHE_CUDA(cudaHostAlloc(&s.host, stage_bytes_, cudaHostAllocDefault));
            FHE_CUDA(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        }
    }
    while (bytes) {
        Stage &s = stage_[stage_next_++ % stage_.size()];
        const size_t part = std::min(bytes, stage_bytes_);
        FHE_CUDA(cudaEventSynchronize(s.done)); // returns at once unless the copy that last used this slot is still queued
        std::memcpy(s.host, src, part);
        FHE_CUDA(cudaMemcpyAsync(dst, s.host, part, cudaMemcpyHostToDevice, stream));
        FHE_CUDA(cudaEventRecord(s.done, stream));
        dst = (char *)dst + part, src = (const char *)src + part, bytes -= part;
    }
}

Mem Context::encode_at(const Plaintext &pt, int ell, double scale) {
    CtxLock g(*this);
    FHE_CUDA(cudaSetDevice(device));
    const int n = P->n, slots = pt.slots;
    if (pt.base) { // rotation of an encoded plaintext = the automorphism X -> X^(5^rot) of its polynomial: one permutation
        Mem src = encoding_for(pt.base, ell, scale), out = alloc((size_t)ell * n);
        const uint32_t gal = P->galois_for_rotation(pt.rot);
        if (gal == 1) return src;
        k_automorphism(T, gal, ell, src->p, out->p, stream);
        return out;
    }
    Mem vbuf = alloc((size_t)2 * slots), rbuf;
    size_t nv;
    if (pt.real_only) {
        nv = std::min((size_t)slots, pt.re.size());
        rbuf = alloc(std::max<size_t>(nv, 1));
        if (nv) upload(rbuf->p, pt.re.data(), sizeof(double) * nv);
    } else {
        nv = std::min((size_t)slots, pt.vals.size());
        if (nv < (size_t)slots) FHE_CUDA(cudaMemsetAsync(vbuf->p, 0, sizeof(double2) * slots, stream));
        if (nv) upload(vbuf->p, pt.vals.data(), sizeof(double2) * nv);
    }
    auto tabs = fft_device_tables(slots);
    Mem out = alloc((size_t)ell * n);
    k_encode(T, (double2 *)vbuf->p, pt.real_only ? (const double *)rbuf->p : nullptr, (int)nv, slots, scale, tabs.first,
             tabs.second, ell, out->p, stream);
    k_ntt_fwd(T, out->p, plan(ell).desc_q, ell, stream);
    return out;
}

Pt Context::new_plaintext(size_t count, int nsd, int level, int slots) {
    if (slots <= 0) slots = batch;
    if ((slots & (slots - 1)) || slots > P->n / 2) throw std::invalid_argument("MakeCKKSPackedPlaintext: slots must be a power of two <= n/2");
    if ((int)count > slots) throw std::invalid_argument("MakeCKKSPackedPlaintext: more values than slots");
    if (level < 0 || level >= P->L) throw std::invalid_argument("MakeCKKSPackedPlaintext: bad level");
    auto p = std::make_shared<Plaintext>();
    p->slots = slots;
    p->length = count;
    p->level = level;
    // FLEXIBLEAUTOEXT: level-0 plaintexts live at the big scale with degree 2; deeper ones at sf[level]^nsd
    if (level == 0) {
        p->nsd = 2;
        p->scale = P->sf_big[0];
    } else {
        p->nsd = nsd < 1 ? 1 : nsd;
        p->scale = std::pow(P->sf[level], (double)p->nsd);
    }
    p->ell = P->L - level;
    // encoded lazily (encoding_for): a mask built at level 0 is only ever used at the level of the ciphertext it multiplies
    return p;
}
Pt Context::make_plaintext(const std::vector<std::complex<double>> &vals, int nsd, int level, int slots) {
    Pt p = new_plaintext(vals.size(), nsd, level, slots);
    p->vals = vals;
    return p;
}
Pt Context::rotate_plaintext(const Pt &base, int index) {
    if (!base) throw std::invalid_argument("rotate_plaintext: null plaintext");
    if (base->base) return rotate_plaintext(base->base, base->rot + index);
    auto p = std::make_shared<Plaintext>();
    p->slots = base->slots, p->level = base->level, p->nsd = base->nsd, p->scale = base->scale, p->ell = base->ell;
    p->length = base->slots, p->log_precision = base->log_precision;
    p->base = base, p->rot = index;
    return p;
}
Pt Context::make_plaintext_real(const double *vals, size_t count, int nsd, int level, int slots) {
    Pt p = new_plaintext(count, nsd, level, slots);
    p->re.assign(vals, vals + count);
    p->real_only = true;
    return p;
}

Ct Context::encrypt(const Keys &k, const Pt &pt) {
    CtxLock g(*this);
    FHE_CUDA(cudaSetDevice(device));
    if (!k || !k->pk) throw std::runtime_error("Encrypt: public key required");
    if (!pt) throw std::runtime_error("Encrypt: null plaintext");
    Mem ptmem = encoding_for(pt, pt->ell, pt->scale);
    const int n = P->n, L = P->L, ell = pt->ell;
    u64 nc = nonce.fetch_add(1);
    auto c = std::make_shared<Ciphertext>();
    c->npoly = 2, c->ell = ell, c->level = pt->level, c->nsd = pt->nsd, c->scale = pt->scale, c->slots = pt->slots;
    c->mem = alloc((size_t)2 * ell * n);
    Mem tmp = alloc((size_t)3 * ell * n), small = alloc((size_t)n / 2 + 1);
    const uint32_t *dq = plan(ell).desc_q;
    const u64 tags[3] = {ST_ENC_V, ST_ENC_E0, ST_ENC_E1};
    for (int x = 0; x < 3; x++) {
        u64 *dst = tmp->p + (size_t)x * ell * n;
        k_sample_small(T, x == 0 ? SMALL_TERNARY : SMALL_GAUSS, rng_secret, stream_id(tags[x], 0, nc), (int *)small->p, stream);
        k_expand_small(T, (int *)small->p, dst, dq, ell, stream);
        k_ntt_fwd(T, dst, dq, ell, stream);
    }
    k_encrypt_finish(T, ell, k->pk->p, k->pk->p + (size_t)L * n, tmp->p, tmp->p + (size_t)ell * n,
                     tmp->p + (size_t)2 * ell * n, ptmem->p, c->poly(0, n), c->poly(1, n), stream);
    return c;
}

Pt Context::decrypt(const Keys &k, const Ct &ct) {
    CtxLock g(*this);
    FHE_CUDA(cudaSetDevice(device));
    if (!k || !k->s) throw std::runtime_error("Decrypt: secret key required");
    if (!ct || ct->npoly != 2) throw std::runtime_error("Decrypt: expected a 2-element ciphertext");
    const int n = P->n, slots = ct->slots;
    // the message is small against Q: keep just enough limbs for |m| < Q_k / 2 (dropping limbs = mod reduction)
    int kk = 0;
    double bits = 0, need = std::log2(ct->scale) + 30.0;
    while (kk < ct->ell && (bits < need || kk < 1)) bits += std::log2((double)P->moduli[kk]), kk++;
    if (kk > 8) throw std::runtime_error("Decrypt: too many limbs needed");
    Mem m = alloc((size_t)kk * n);
    k_decrypt_core(T, kk, ct->poly(0, n), ct->poly(1, n), k->s->p, m->p, stream);
    k_ntt_inv(T, m->p, plan(ct->ell).desc_q, kk, stream);
    std::vector<u64> h((size_t)kk * n);
    FHE_CUDA(cudaMemcpyAsync(h.data(), m->p, h.size() * sizeof(u64), cudaMemcpyDeviceToHost, stream));
    FHE_CUDA(cudaStreamSynchronize(stream));
    // Garner mixed radix with exact sign decision
    std::vector<std::vector<u64>> inv(kk, std::vector<u64>(kk, 0));
    for (int i = 0; i < kk; i++)
        for (int j = 0; j < i; j++) inv[i][j] = invmod_h(P->moduli[j] % P->moduli[i], P->moduli[i]);
    const int gap = n / (2 * slots);
    auto centred = [&](size_t j) -> long double {
        u64 d[8];
        for (int i = 0; i < kk; i++) {
            u64 qi = P->moduli[i], t = h[(size_t)i * n + j];
            for (int x = 0; x < i; x++) {
                u64 dx = d[x] % qi;
                t = mulmod_h(t >= dx ? t - dx : t + qi - dx, inv[i][x], qi);
            }
            d[i] = t;
        }
        bool neg = d[kk - 1] > (P->moduli[kk - 1] >> 1);
        if (neg)
            for (int i = 0; i < kk; i++) d[i] = P->moduli[i] - 1 - d[i];
        long double r = 0;
        for (int i = kk - 1; i >= 0; i--) r = r * (long double)P->moduli[i] + (long double)d[i];
        return neg ? -(r + 1) : r;
    };
    auto p = std::make_shared<Plaintext>();
    p->slots = slots, p->level = ct->level, p->nsd = ct->nsd, p->scale = ct->scale, p->ell = ct->ell;
    p->vals.resize(slots);
    for (int i = 0; i < slots; i++) {
        long double re = centred((size_t)i * gap) / (long double)ct->scale;
        long double im = centred((size_t)n / 2 + (size_t)i * gap) / (long double)ct->scale;
        p->vals[i] = std::complex<double>((double)re, (double)im);
    }
    if (slots > 1) fft_special(p->vals);
    p->length = slots;
    // crude precision estimate (OpenFHE reports log2 of the imaginary-part noise)
    double mx = 0;
    for (auto &z : p->vals) mx = std::max(mx, std::fabs(z.imag()));
    p->log_precision = mx > 0 ? -std::log2(mx) : 64.0;
    return p;
}

} // namespace fheb200
