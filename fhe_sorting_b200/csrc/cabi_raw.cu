// cabi_raw.cu — raw RNS import/export, single-primitive entry points (parity tests) and device timing loops.
#include <cstring>
#include <stdexcept>
#include <string>

#include "../../include/fhe_b200.h"
#include "engine.h"
#include "modarith.cuh"

using namespace fheb200;

struct fheb200_ctx { std::shared_ptr<Context> c; };
struct fheb200_keys { Keys k; };
struct fheb200_ct { Ct c; };
struct fheb200_pt { Pt p; };

extern "C" const char *fheb200_last_error(void);
namespace fheb200 { void set_last_error(const char *); }
static thread_local std::string g_err_raw;
#define GUARD(...)                                   \
    try {                                            \
        __VA_ARGS__;                                     \
        return 0;                                    \
    } catch (const std::exception &e) {              \
        fheb200::set_last_error(e.what());           \
        return 1;                                    \
    } catch (...) {                                  \
        fheb200::set_last_error("unknown error");    \
        return 1;                                    \
    }
static void need(const void *p, const char *what) {
    if (!p) throw std::invalid_argument(std::string("null argument: ") + what);
}
static void d2h(Context &C, void *dst, const void *src, size_t bytes) {
    FHE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, C.stream));
    C.sync();
}
static void h2d(Context &C, void *dst, const void *src, size_t bytes) {
    FHE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, C.stream));
    C.sync();
}
#define LOCKCTX(ctx)                                              \
    need(ctx, "ctx");                                             \
    Context &C = *ctx->c;                                         \
    CtxLock lock__(C);                                            \
    FHE_CUDA(cudaSetDevice(C.device));                            \
    const int n = C.P->n;                                         \
    (void)n

extern "C" {

unsigned long long fheb200_launch_count(void) { return g_launches.load(); }
unsigned long long fheb200_algo_bytes(void) { return g_algo_bytes.load(); }
unsigned long long fheb200_op_count(int which) { return which >= 0 && which < 4 ? g_op_count[which].load() : 0ull; }

int fheb200_ct_export(fheb200_ctx *ctx, fheb200_ct *ct, uint64_t *c0, uint64_t *c1, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        need(ct, "ct");
        size_t w = (size_t)ct->c->ell * n;
        if (words < w) throw std::invalid_argument("ct_export: buffer too small");
        if (c0) d2h(C, c0, ct->c->poly(0, n), w * 8);
        if (c1) d2h(C, c1, ct->c->poly(1, n), w * 8);
    })
}
int fheb200_ct_import(fheb200_ctx *ctx, const uint64_t *c0, const uint64_t *c1, int ell, int nsd, double scale, int slots, fheb200_ct **out) {
    GUARD({
        LOCKCTX(ctx);
        need(c0, "c0"); need(c1, "c1"); need(out, "out");
        if (ell < 1 || ell > C.P->L) throw std::invalid_argument("ct_import: bad limb count");
        auto c = std::make_shared<Ciphertext>();
        c->npoly = 2, c->ell = ell, c->level = C.P->L - ell, c->nsd = nsd, c->scale = scale, c->slots = slots > 0 ? slots : C.batch;
        c->mem = C.alloc((size_t)2 * ell * n);
        h2d(C, c->poly(0, n), c0, (size_t)ell * n * 8);
        h2d(C, c->poly(1, n), c1, (size_t)ell * n * 8);
        *out = new fheb200_ct{c};
    })
}
// Asynchronous variants for pinned host buffers: the copies run on the context's copy stream and overlap the kernels of
// neighbouring work items.  The host buffers must stay untouched (import) / are valid (export) only after fheb200_ctx_sync.
int fheb200_ct_import_async(fheb200_ctx *ctx, const uint64_t *c0, const uint64_t *c1, int ell, int nsd, double scale, int slots,
                            fheb200_ct **out) {
    GUARD({
        LOCKCTX(ctx);
        need(c0, "c0"); need(c1, "c1"); need(out, "out");
        if (ell < 1 || ell > C.P->L) throw std::invalid_argument("ct_import: bad limb count");
        auto c = std::make_shared<Ciphertext>();
        c->npoly = 2, c->ell = ell, c->level = C.P->L - ell, c->nsd = nsd, c->scale = scale, c->slots = slots > 0 ? slots : C.batch;
        c->mem = C.alloc((size_t)2 * ell * n);
        C.copy_in(c->poly(0, n), c0, (size_t)ell * n * 8, c->mem);
        C.copy_in(c->poly(1, n), c1, (size_t)ell * n * 8, c->mem);
        *out = new fheb200_ct{c};
    })
}
int fheb200_ct_ready(fheb200_ctx *ctx, fheb200_ct *ct) {
    GUARD({ LOCKCTX(ctx); need(ct, "ct"); C.wait_copy_in(ct->c->mem); })
}
int fheb200_ct_export_async(fheb200_ctx *ctx, fheb200_ct *ct, uint64_t *c0, uint64_t *c1, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        need(ct, "ct");
        size_t w = (size_t)ct->c->ell * n;
        if (words < w) throw std::invalid_argument("ct_export: buffer too small");
        if (c0) C.copy_out(c0, ct->c->poly(0, n), w * 8, ct->c->mem);
        if (c1) C.copy_out(c1, ct->c->poly(1, n), w * 8, ct->c->mem);
    })
}
int fheb200_pt_export(fheb200_ctx *ctx, fheb200_pt *pt, uint64_t *out, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        need(pt, "pt");
        Mem ptmem = C.encoding_for(pt->p, pt->p->ell, pt->p->scale);
        size_t w = (size_t)pt->p->ell * n;
        if (words < w) throw std::invalid_argument("pt_export: buffer too small");
        d2h(C, out, ptmem->p, w * 8);
    })
}
int fheb200_keys_export(fheb200_ctx *ctx, fheb200_keys *k, int which, uint64_t *out, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        need(k, "keys");
        const int L = C.P->L, F = L + C.P->K;
        size_t w = (size_t)(which == 0 ? F : L) * n;
        if (words < w) throw std::invalid_argument("keys_export: buffer too small");
        if (which == 0) {
            if (!k->k->s) throw std::runtime_error("keys_export: no secret key");
            d2h(C, out, k->k->s->p, w * 8);
        } else {
            d2h(C, out, k->k->pk->p + (which == 2 ? w : 0), w * 8);
        }
    })
}
int fheb200_evk_export(fheb200_ctx *ctx, uint32_t galois, uint64_t *out, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        if (words < C.evk_words()) throw std::invalid_argument("evk_export: buffer too small");
        const u64 *src = nullptr;
        if (galois == 0) src = C.evk_mult;
        else if (C.evk_rot.count(galois)) src = C.evk_rot[galois];
        if (!src) throw std::runtime_error("evk_export: key not present");
        d2h(C, out, src, C.evk_words() * 8);
    })
}
// ---- raw import of key material (the other half of keys_export / evk_export: serialisation, src/sort.h:31-73)
// public key (b, a), each [L][n], evaluation domain; the resulting key set has no secret key (evaluation-only party)
int fheb200_keys_import_public(fheb200_ctx *ctx, const uint64_t *b, const uint64_t *a, size_t words_each, fheb200_keys **out) {
    GUARD({
        LOCKCTX(ctx);
        need(b, "b"); need(a, "a"); need(out, "out");
        const size_t w = (size_t)C.P->L * n;
        if (words_each != w) throw std::invalid_argument("keys_import_public: expected L * ring_dim words per polynomial");
        auto K = std::make_shared<KeySet>();
        K->pk = C.alloc(2 * w);
        h2d(C, K->pk->p, b, w * 8);
        h2d(C, K->pk->p + w, a, w * 8);
        *out = new fheb200_keys{K};
    })
}
// evaluation key [digit][2][L+K][n] for Galois element `galois` (0 = the relinearisation key); replaces a key already present
int fheb200_evk_import(fheb200_ctx *ctx, uint32_t galois, const uint64_t *data, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        need(data, "data");
        if (words != C.evk_words()) throw std::invalid_argument("evk_import: wrong size for this context");
        if (galois != 0 && (galois & 1u) == 0) throw std::invalid_argument("evk_import: Galois elements are odd");
        u64 *dst = nullptr;
        FHE_CUDA(cudaMalloc((void **)&dst, words * 8));
        h2d(C, dst, data, words * 8);
        u64 *&slot = galois == 0 ? C.evk_mult : C.evk_rot[galois];
        if (slot) cudaFree(slot);
        slot = dst;
    })
}
// Galois elements of the rotation keys present, ascending
int fheb200_evk_list(fheb200_ctx *ctx, uint32_t *out, size_t capacity, size_t *count) {
    GUARD({
        LOCKCTX(ctx);
        size_t k = 0;
        for (auto &kv : C.evk_rot) {
            if (out && k < capacity) out[k] = kv.first;
            k++;
        }
        if (count) *count = k;
    })
}
int fheb200_has_mult_key(fheb200_ctx *ctx, int *present) {
    GUARD({ LOCKCTX(ctx); need(present, "present"); *present = C.evk_mult != nullptr; })
}
int fheb200_ctx_evk_words(fheb200_ctx *ctx, size_t *words) {
    GUARD({ LOCKCTX(ctx); need(words, "words"); *words = C.evk_words(); })
}
int fheb200_ctx_params(fheb200_ctx *ctx, fheb200_params *p) {
    GUARD({
        LOCKCTX(ctx);
        need(p, "params");
        p->log_ring_dim = C.P->logn, p->mult_depth = C.P->depth, p->scaling_mod_size = C.P->scale_bits;
        p->first_mod_size = C.P->first_bits, p->batch_size = C.batch, p->num_large_digits = C.P->dnum, p->device = C.device;
    })
}
int fheb200_galois_for_rotation(fheb200_ctx *ctx, int index, uint32_t *g) { GUARD({ need(ctx, "ctx"); *g = ctx->c->P->galois_for_rotation(index); }) }

int fheb200_raw_ntt(fheb200_ctx *ctx, int inverse, int qp, uint64_t *poly) {
    GUARD({
        LOCKCTX(ctx);
        if (qp < 0 || qp >= C.T.F) throw std::out_of_range("raw_ntt: modulus index");
        Mem m = C.alloc(n);
        h2d(C, m->p, poly, (size_t)n * 8);
        const uint32_t *d = C.desc({(uint32_t)qp});
        if (inverse) k_ntt_inv(C.T, m->p, d, 1, C.stream);
        else k_ntt_fwd(C.T, m->p, d, 1, C.stream);
        d2h(C, poly, m->p, (size_t)n * 8);
    })
}
int fheb200_raw_automorphism(fheb200_ctx *ctx, uint32_t g, int limbs, const uint64_t *in, uint64_t *out) {
    GUARD({
        LOCKCTX(ctx);
        Mem a = C.alloc((size_t)limbs * n), b = C.alloc((size_t)limbs * n);
        h2d(C, a->p, in, (size_t)limbs * n * 8);
        k_automorphism(C.T, g, limbs, a->p, b->p, C.stream);
        d2h(C, out, b->p, (size_t)limbs * n * 8);
    })
}
int fheb200_raw_modup(fheb200_ctx *ctx, int ell, const uint64_t *c, uint64_t *digits, size_t words) {
    GUARD({
        LOCKCTX(ctx);
        const DevicePlan &D = C.plan(ell);
        size_t w = (size_t)D.beta * D.width * n;
        if (words < w) throw std::invalid_argument("raw_modup: buffer too small");
        Mem a = C.alloc((size_t)ell * n), dg = C.alloc(w);
        h2d(C, a->p, c, (size_t)ell * n * 8);
        C.modup(ell, a->p, dg->p);
        d2h(C, digits, dg->p, w * 8);
    })
}
int fheb200_raw_rescale(fheb200_ctx *ctx, int ell, const uint64_t *in, uint64_t *out) {
    GUARD({
        LOCKCTX(ctx);
        Mem a = C.alloc((size_t)ell * n), b = C.alloc((size_t)(ell - 1) * n);
        h2d(C, a->p, in, (size_t)ell * n * 8);
        C.rescale_poly(ell, 1, a->p, b->p);
        d2h(C, out, b->p, (size_t)(ell - 1) * n * 8);
    })
}

// ---- single-kernel entry points for the limb-level parity tests (tests/test_gpu_kernels_limb.py)
// out[g][limbs2][n] = sum_t in[t][limbs2][n] * w[g][t][limb % ell]  (the Paterson-Stockmeyer leaf sums);
// variant 0: integer tensor cores through the cp.async rings (lc_tc_ring), 1: tensor cores, operands from global memory
// (lc_tc), 2: CUDA cores (ew_lincomb, G <= 4)
int fheb200_raw_lincomb(fheb200_ctx *ctx, int ell, int limbs2, int count, int G, const uint64_t *in, const uint64_t *w, uint64_t *out,
                        int variant) {
    GUARD({
        LOCKCTX(ctx);
        need(in, "in"); need(w, "w"); need(out, "out");
        if (ell < 1 || ell > C.P->L || limbs2 % ell || count < 1 || G < 1 || G > (variant == 2 ? 4 : 8))
            throw std::invalid_argument("raw_lincomb: bad shape");
        const size_t poly = (size_t)limbs2 * n;
        Mem din = C.alloc(poly * count), dout = C.alloc(poly * G);
        h2d(C, din->p, in, poly * count * 8);
        std::vector<u64> blob((size_t)count + G + (size_t)G * count * ell);
        for (int t = 0; t < count; t++) blob[t] = (u64)(uintptr_t)(din->p + poly * t);
        for (int g = 0; g < G; g++) blob[count + g] = (u64)(uintptr_t)(dout->p + poly * g);
        for (size_t i = 0; i < (size_t)G * count * ell; i++) blob[(size_t)count + G + i] = (w[i] & 0x3fffffffull) | ((w[i] >> 30) << 32);
        Mem d = C.alloc(blob.size());
        h2d(C, d->p, blob.data(), blob.size() * 8);
        if (variant == 2) {
            k_lincomb(C.T, ell, limbs2, count, G, (const u64 *const *)d->p, (const uint2 *)(d->p + count + G), (u64 *const *)(d->p + count), C.stream);
        } else if (variant == 3) { // tcgen05 / TMEM form
            Mem tables = C.alloc(lincomb_tc_scratch_words(ell, count));
            if (!k_lincomb_tcgen05(C.T, ell, limbs2, count, G, (const u64 *const *)d->p, (const uint2 *)(d->p + count + G), (u64 *const *)(d->p + count),
                                   tables->p, C.stream))
                throw std::invalid_argument("raw_lincomb: shape not covered by the tcgen05 kernel");
        } else {
            Mem tables = C.alloc(lincomb_tc_scratch_words(ell, count));
            k_lincomb_tc(C.T, ell, limbs2, count, G, (const u64 *const *)d->p, (const uint2 *)(d->p + count + G), (u64 *const *)(d->p + count),
                         tables->p, C.stream, variant == 0 ? 1 : 0);
        }
        d2h(C, out, dout->p, poly * G * 8);
    })
}
// o_b = (accumulate ? o_b : 0) + a_b * pt on ell limbs (ew_mul_pt)
int fheb200_raw_mul_pt(fheb200_ctx *ctx, int ell, const uint64_t *a0, const uint64_t *a1, const uint64_t *pt, uint64_t *o0, uint64_t *o1,
                       int accumulate) {
    GUARD({
        LOCKCTX(ctx);
        const size_t w = (size_t)ell * n;
        Mem d = C.alloc(5 * w);
        h2d(C, d->p, a0, w * 8); h2d(C, d->p + w, a1, w * 8); h2d(C, d->p + 2 * w, pt, w * 8);
        h2d(C, d->p + 3 * w, o0, w * 8); h2d(C, d->p + 4 * w, o1, w * 8);
        k_mul_pt(C.T, ell, d->p, d->p + w, d->p + 2 * w, d->p + 3 * w, d->p + 4 * w, accumulate, C.stream);
        d2h(C, o0, d->p + 3 * w, w * 8); d2h(C, o1, d->p + 4 * w, w * 8);
    })
}
// (d0, d1, d2) = (a0 b0, a0 b1 + a1 b0, a1 b1) on ell limbs (ew_tensor; b == a: ew_tensor_square)
int fheb200_raw_tensor(fheb200_ctx *ctx, int ell, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0, const uint64_t *b1,
                       uint64_t *d0, uint64_t *d1, uint64_t *d2) {
    GUARD({
        LOCKCTX(ctx);
        const size_t w = (size_t)ell * n;
        Mem d = C.alloc(7 * w);
        h2d(C, d->p, a0, w * 8); h2d(C, d->p + w, a1, w * 8);
        if (b0) { h2d(C, d->p + 2 * w, b0, w * 8); h2d(C, d->p + 3 * w, b1, w * 8); }
        if (b0) k_tensor(C.T, ell, d->p, d->p + w, d->p + 2 * w, d->p + 3 * w, d->p + 4 * w, d->p + 5 * w, d->p + 6 * w, C.stream);
        else k_tensor_square(C.T, ell, d->p, d->p + w, d->p + 4 * w, d->p + 5 * w, d->p + 6 * w, C.stream);
        d2h(C, d0, d->p + 4 * w, w * 8); d2h(C, d1, d->p + 5 * w, w * 8); d2h(C, d2, d->p + 6 * w, w * 8);
    })
}
// S[2][ell + K][n] += pt (.) sigma_g(v) in the extended basis (ks_dh_accumulate; acc may be null: the unrotated term)
int fheb200_raw_dh_accumulate(fheb200_ctx *ctx, int ell, const uint64_t *acc, const uint64_t *cP, const uint64_t *pt, uint64_t *S,
                              uint32_t galois) {
    GUARD({
        LOCKCTX(ctx);
        const DevicePlan &D = C.plan(ell);
        const size_t W = (size_t)D.width * n, Q = (size_t)ell * n;
        Mem dacc = C.alloc(2 * W), dcp = C.alloc(2 * Q), dpt = C.alloc(W), dS = C.alloc(2 * W);
        if (acc) h2d(C, dacc->p, acc, 2 * W * 8);
        h2d(C, dcp->p, cP, 2 * Q * 8); h2d(C, dpt->p, pt, W * 8); h2d(C, dS->p, S, 2 * W * 8);
        k_dh_accumulate(C.T, D, acc ? dacc->p : nullptr, dcp->p, dpt->p, dS->p, galois, C.stream);
        d2h(C, S, dS->p, 2 * W * 8);
    })
}
// evk inner product alone: acc[2][ell + K][n] = sum_j digits[j] (.) evk[j] with the key of Galois element g (0: relin key)
int fheb200_raw_inner_product(fheb200_ctx *ctx, int ell, uint32_t galois, const uint64_t *digits, uint64_t *acc) {
    GUARD({
        LOCKCTX(ctx);
        const DevicePlan &D = C.plan(ell);
        const u64 *evk = galois == 0 ? C.evk_mult : (C.evk_rot.count(galois) ? C.evk_rot[galois] : nullptr);
        if (!evk) throw std::runtime_error("raw_inner_product: key not present");
        const size_t W = (size_t)D.width * n;
        Mem dd = C.alloc(D.beta * W), da = C.alloc(2 * W);
        h2d(C, dd->p, digits, D.beta * W * 8);
        k_inner_product(C.T, D, dd->p, evk, da->p, C.stream);
        d2h(C, acc, da->p, 2 * W * 8);
    })
}
// ModDown alone: acc[2][ell + K][n] (evaluation domain) -> out_b = round(acc_b / P) on ell limbs
int fheb200_raw_moddown(fheb200_ctx *ctx, int ell, const uint64_t *acc, uint64_t *out0, uint64_t *out1) {
    GUARD({
        LOCKCTX(ctx);
        const DevicePlan &D = C.plan(ell);
        const size_t W = (size_t)D.width * n, Q = (size_t)ell * n;
        Mem da = C.alloc(2 * W), o = C.alloc(2 * Q);
        h2d(C, da->p, acc, 2 * W * 8);
        C.moddown(ell, da->p, nullptr, nullptr, o->p, o->p + Q, 1u);
        d2h(C, out0, o->p, Q * 8); d2h(C, out1, o->p + Q, Q * 8);
    })
}

// Forward NTT of `count` limbs that already lie in DEVICE memory ([count][n] u64, limb i reduced modulo modulus qp_indices[i]),
// written to d_out ([count][n], may equal d_in), on the context's stream.  Allocates nothing: see k_ntt_fwd_static.  Returns
// after the launch; fheb200_ctx_sync (or stream order) makes the result visible.
int fheb200_ntt_forward_noalloc(fheb200_ctx *ctx, const int *qp_indices, int count, const uint64_t *d_in, uint64_t *d_out) {
    GUARD({
        LOCKCTX(ctx);
        need(qp_indices, "qp_indices"); need(d_in, "d_in"); need(d_out, "d_out");
        if (count < 1 || count > ntt_static_max_limbs())
            throw std::invalid_argument("ntt_forward_noalloc: count must be 1.." + std::to_string(ntt_static_max_limbs()) +
                                        " limbs (size of the static scratch array)");
        uint32_t desc[256];
        static_assert(sizeof(desc) / sizeof(desc[0]) >= 256, "descriptor buffer");
        if (ntt_static_max_limbs() > 256) throw std::logic_error("ntt_forward_noalloc: descriptor buffer too small");
        for (int i = 0; i < count; i++) {
            if (qp_indices[i] < 0 || qp_indices[i] >= C.T.F) throw std::out_of_range("ntt_forward_noalloc: modulus index");
            desc[i] = ((uint32_t)i << 8) | (uint32_t)qp_indices[i];
        }
        k_ntt_fwd_static(C.T, d_in, d_out, desc, count, C.stream);
    })
}
int fheb200_ntt_forward_noalloc_max_limbs(void) { return ntt_static_max_limbs(); }

// host-side evaluation of the generator the sampling kernels use (known-answer test of the ChaCha20 block function)
int fheb200_rng_block(const uint32_t *key8, uint64_t stream, uint64_t index, uint64_t *out2) {
    GUARD({
        need(key8, "key"); need(out2, "out");
        uint32_t k[8];
        for (int i = 0; i < 8; i++) k[i] = key8[i];
        u64 w0, w1;
        chacha20_128(k, stream, index, w0, w1);
        out2[0] = w0, out2[1] = w1;
    })
}

int fheb200_ct_device_view(fheb200_ct *ct, void **device_ptr, size_t *words) {
    GUARD({
        need(ct, "ct");
        need(device_ptr, "device_ptr");
        *device_ptr = ct->c->mem->p;
        if (words) *words = (size_t)ct->c->npoly * ct->c->ell * ct->c->mem->ctx->P->n;
    })
}
int fheb200_ct_reduce_partial_sums(fheb200_ctx *ctx, fheb200_ct *ct) {
    GUARD({
        LOCKCTX(ctx);
        need(ct, "ct");
        ct->c->reduced_cache.reset();
        for (int p = 0; p < ct->c->npoly; p++) k_reduce_words(C.T, ct->c->ell, ct->c->poly(p, n), C.stream);
    })
}

// `reps` rotations of the same ciphertext back to back; device time by CUDA events on the context stream
int fheb200_bench_rotate(fheb200_ctx *ctx, fheb200_ct *ct, int index, int reps, float *ms) {
    GUARD({
        LOCKCTX(ctx);
        need(ct, "ct");
        cudaEvent_t e0, e1;
        FHE_CUDA(cudaEventCreate(&e0));
        FHE_CUDA(cudaEventCreate(&e1));
        FHE_CUDA(cudaEventRecord(e0, C.stream));
        for (int r = 0; r < reps; r++) { Ct o = C.rotate(ct->c, index); }
        FHE_CUDA(cudaEventRecord(e1, C.stream));
        FHE_CUDA(cudaEventSynchronize(e1));
        FHE_CUDA(cudaEventElapsedTime(ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    })
}

// `reps` passes over `count` rotation indices of the same ciphertext (each index has its own evk, so the key stream is
// never L2 resident); hoisted != 0 shares one EvalFastRotationPrecompute per pass. Device time by CUDA events.
int fheb200_bench_rotations(fheb200_ctx *ctx, fheb200_ct *ct, const int *indices, size_t count, int reps, int hoisted,
                            float *ms) {
    GUARD({
        LOCKCTX(ctx);
        need(ct, "ct");
        need(indices, "indices");
        cudaEvent_t e0, e1;
        FHE_CUDA(cudaEventCreate(&e0));
        FHE_CUDA(cudaEventCreate(&e1));
        FHE_CUDA(cudaEventRecord(e0, C.stream));
        for (int r = 0; r < reps; r++) {
            if (hoisted) {
                auto pre = C.fast_rotation_precompute(ct->c);
                for (size_t i = 0; i < count; i++) { Ct o = C.fast_rotation(ct->c, indices[i], pre); }
            } else {
                for (size_t i = 0; i < count; i++) { Ct o = C.rotate(ct->c, indices[i]); }
            }
        }
        FHE_CUDA(cudaEventRecord(e1, C.stream));
        FHE_CUDA(cudaEventSynchronize(e1));
        FHE_CUDA(cudaEventElapsedTime(ms, e0, e1));
        FHE_CUDA(cudaGetLastError());
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    })
}

// isolated kernels on scratch data at `ell` active limbs: "ntt_fwd", "ntt_inv", "inner_product", "modup_conv",
// "moddown_conv", "automorphism", "tensor", "add", and the CUDA-core conversions "modup_conv_scalar", "moddown_conv_scalar"
int fheb200_bench_kernel(fheb200_ctx *ctx, const char *name, int ell, int reps, float *ms, double *algo_bytes) {
    GUARD({
        LOCKCTX(ctx);
        need(name, "name");
        const DevicePlan &D = C.plan(ell);
        const std::string k(name);
        const size_t limb = (size_t)n * 8;
        Mem a = C.alloc((size_t)D.beta * D.width * n), b = C.alloc((size_t)3 * D.width * n);
        FHE_CUDA(cudaMemsetAsync(a->p, 0, a->words * 8, C.stream));
        FHE_CUDA(cudaMemsetAsync(b->p, 0, b->words * 8, C.stream));
        if (!C.evk_mult && k == "inner_product") throw std::runtime_error("bench_kernel: needs the relinearisation key");
        double bytes = 0;
        // "lincomb": 4 weighted sums of 63 polynomials of 2 * ell limbs each (the Paterson-Stockmeyer leaf pass of a
        // degree-6510 evaluation, k = 64): pointer table and split weights as Context::lincomb lays them out
        const bool lc_tc = [] { const char *e = getenv("FHEB200_LINCOMB"); return e ? atoi(e) == 6 : true; }();
        const int lc_count = 63, lc_G = lc_tc ? 8 : 4;
        Mem lc_in, lc_out, lc_tab, lc_scratch;
        if (k == "lincomb") {
            lc_in = C.alloc((size_t)lc_count * 2 * ell * n), lc_out = C.alloc((size_t)lc_G * 2 * ell * n);
            {
                RngState bench_rng; // fixed test pattern (one limb of noise, replicated below)
                bench_rng.seed = 7;
                k_sample_uniform(C.T, bench_rng, 99, 0, 1, lc_in->p, C.stream);
            }
            for (size_t off = (size_t)n; off < lc_in->words; off *= 2)
                FHE_CUDA(cudaMemcpyAsync(lc_in->p + off, lc_in->p, std::min(off, lc_in->words - off) * 8, cudaMemcpyDeviceToDevice, C.stream));
            std::vector<u64> blob((size_t)lc_count + lc_G + (size_t)lc_G * lc_count * ell);
            for (int t = 0; t < lc_count; t++) blob[t] = (u64)(uintptr_t)(lc_in->p + (size_t)t * 2 * ell * n);
            for (int g = 0; g < lc_G; g++) blob[lc_count + g] = (u64)(uintptr_t)(lc_out->p + (size_t)g * 2 * ell * n);
            for (size_t i = (size_t)lc_count + lc_G; i < blob.size(); i++) {
                const size_t rel = i - lc_count - lc_G;
                u64 v = (0x9e3779b97f4a7c15ull * (rel + 1)) % C.P->moduli[rel % ell];
                blob[i] = (v & 0x3fffffffull) | ((v >> 30) << 32);
            }
            lc_tab = C.alloc(blob.size());
            lc_scratch = C.alloc(lincomb_tc_scratch_words(ell, lc_count));
            FHE_CUDA(cudaMemcpyAsync(lc_tab->p, blob.data(), blob.size() * 8, cudaMemcpyHostToDevice, C.stream));
            FHE_CUDA(cudaStreamSynchronize(C.stream));
        }
        // the NTTs walk over six disjoint regions of ell limbs (6 x 43 MB at 41 limbs: more than the 126 MB L2), so that no
        // launch finds its input in L2 from the launch before
        int iter = 0;
        auto region = [&]() {
            const int r = iter++ % 6;
            return (r < 3 ? a->p : b->p) + (size_t)(r % 3) * D.width * n;
        };
        auto run = [&]() {
            if (k == "ntt_fwd") { k_ntt_fwd(C.T, region(), D.desc_q, ell, C.stream); bytes = 2.0 * ell * limb; }
            else if (k == "ntt_inv") { k_ntt_inv(C.T, region(), D.desc_q, ell, C.stream); bytes = 2.0 * ell * limb; }
            else if (k == "inner_product") { k_inner_product(C.T, D, a->p, C.evk_mult, b->p, C.stream); bytes = (3.0 * D.beta + 2.0) * D.width * limb; }
            else if (k == "modup_conv_scalar") { k_modup(C.T, D, b->p, b->p, a->p, C.stream); bytes = (double)(ell + D.beta * D.width) * limb; }
            else if (k == "moddown_conv_scalar") { k_moddown_conv(C.T, D, b->p, a->p, C.stream); bytes = 2.0 * (D.K + ell) * limb; }
            else if (k == "modup_conv") { k_modup_tc(C.T, D, b->p, b->p, a->p, C.stream); bytes = (double)(ell + D.beta * D.width) * limb; }
            else if (k == "moddown_conv") { k_moddown_conv_tc(C.T, D, b->p, a->p, C.stream); bytes = 2.0 * (D.K + ell) * limb; }
            else if (k == "automorphism") { k_automorphism(C.T, 5, ell, a->p, b->p, C.stream); bytes = 2.0 * ell * limb; }
            else if (k == "tensor") { k_tensor(C.T, ell, a->p, a->p + (size_t)ell * n, a->p, a->p + (size_t)ell * n, b->p, b->p + (size_t)ell * n, b->p + (size_t)2 * ell * n, C.stream); bytes = 7.0 * ell * limb; }
            else if (k == "add") { k_binary(C.T, OP_ADD, ell, a->p, a->p + (size_t)ell * n, b->p, C.stream); bytes = 3.0 * ell * limb; }
            else if (k == "lincomb") {
                if (lc_tc)
                    k_lincomb_tc(C.T, ell, 2 * ell, lc_count, lc_G, (const u64 *const *)lc_tab->p, (const uint2 *)(lc_tab->p + lc_count + lc_G),
                                 (u64 *const *)(lc_tab->p + lc_count), lc_scratch->p, C.stream);
                else
                    k_lincomb(C.T, ell, 2 * ell, lc_count, lc_G, (const u64 *const *)lc_tab->p, (const uint2 *)(lc_tab->p + lc_count + lc_G),
                              (u64 *const *)(lc_tab->p + lc_count), C.stream);
                bytes = (double)(lc_count + lc_G) * 2 * ell * limb;
            }
            else throw std::invalid_argument("bench_kernel: unknown kernel " + k);
        };
        run(); // warm-up
        cudaEvent_t e0, e1;
        FHE_CUDA(cudaEventCreate(&e0));
        FHE_CUDA(cudaEventCreate(&e1));
        FHE_CUDA(cudaEventRecord(e0, C.stream));
        for (int r = 0; r < reps; r++) run();
        FHE_CUDA(cudaEventRecord(e1, C.stream));
        FHE_CUDA(cudaEventSynchronize(e1));
        FHE_CUDA(cudaEventElapsedTime(ms, e0, e1));
        FHE_CUDA(cudaGetLastError());
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (algo_bytes) *algo_bytes = bytes;
        if (k == "lincomb") { // checksum of the outputs on stderr: identical across kernel variants (FHEB200_LINCOMB)
            std::vector<u64> h((size_t)4 * 2 * ell * n); // the first four outputs: the same sums whatever the outputs per pass
            FHE_CUDA(cudaMemcpy(h.data(), lc_out->p, h.size() * 8, cudaMemcpyDeviceToHost));
            u64 x = 0;
            for (size_t i = 0; i < h.size(); i++) x = x * 1099511628211ull + h[i];
            fprintf(stderr, "lincomb checksum %016llx\n", (unsigned long long)x);
        }
    })
}

} // extern "C"
