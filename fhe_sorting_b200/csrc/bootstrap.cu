// bootstrap.cu — CKKS bootstrapping on the engine's own operations (plan: bootstrap_plan.cpp).
//
//   1. bring the ciphertext to its last two limbs, multiply by round(msg_scale q_1 / scale) and rescale: one limb q_0,
//      message polynomial msg_scale * m with msg_scale = q_0 2^-correction (so that sin(2 pi t / q_0) is linear in it)
//   2. ModRaise: the two polynomials as centred integers on the limbs of level 1; decrypts to msg_scale m + q_0 I, |I| < K
//   3. SubSum: log2(n / 2s) rotate-and-adds project I onto the subring of the s-slot packing (and multiply by n / 2s)
//   4. CoeffsToSlots (grouped FFT factors, hoisted rotations), + conjugate: 2s real slots t_k / (q_0 K)
//   5. EvalMod: Chebyshev series of cos(2 pi (K u - 1/4) / 2^r) by Paterson-Stockmeyer, then r double-angle steps
//   6. SlotsToCoeffs, + rotate by s
// Reference call sites: src/k-way/EvalUtils.cpp:57-86 (checkLevelAndBoot), src/sign.cpp:163-170 (lazyBootstrap).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#include "engine.h"

namespace fheb200 {

#define LOCK                                           \
    CtxLock lock__(*this);                             \
    FHE_CUDA(cudaSetDevice(device))

static const int RAISE_LEVEL = 1; // the raised ciphertext lives at level 1 (the FLEXIBLEAUTOEXT extra limb stays unused)

static int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}

void Context::bootstrap_setup(int budget_enc, int budget_dec, int slots) {
    LOCK;
    if (slots <= 0) slots = batch;
    if (slots > P->n / 4 || (slots & (slots - 1)))
        throw std::invalid_argument("EvalBootstrapSetup: fully packed bootstrapping is unsupported: slots must be a power of two <= ring_dim / 4 (got " +
                                    std::to_string(slots) + "; the reference's k-way fixture packs <= 1024)");
    auto B = std::make_shared<BootState>();
    B->plan = make_boot_plan(P->n, slots, budget_enc, budget_dec, (double)P->moduli[0], P->sf[RAISE_LEVEL],
                             env_int("FHEB200_BOOT_CORRECTION", 10));
    if (RAISE_LEVEL + B->plan.depth() > P->depth)
        throw std::invalid_argument("EvalBootstrapSetup: multiplicative depth " + std::to_string(P->depth) +
                                    " is below the bootstrapping depth " + std::to_string(RAISE_LEVEL + B->plan.depth()));
    int level = RAISE_LEVEL;
    for (auto &st : B->plan.cts) {
        std::vector<Pt> v;
        for (auto &d : st.diag) v.push_back(make_plaintext(d, 1, level, 2 * slots));
        B->cts_pt.push_back(std::move(v));
        level++;
    }
    level += B->plan.cheb_depth() + B->plan.r;
    for (auto &st : B->plan.stc) {
        std::vector<Pt> v;
        for (auto &d : st.diag) v.push_back(make_plaintext(d, 1, std::min(level, P->L - 1), 2 * slots));
        B->stc_pt.push_back(std::move(v));
        level++;
    }
    B->cts_qp.resize(B->plan.cts.size()), B->stc_qp.resize(B->plan.stc.size());
    boot[slots] = B;
}
int Context::bootstrap_depth(int slots) const {
    auto it = boot.find(slots > 0 ? slots : batch);
    if (it == boot.end()) throw std::runtime_error("EvalBootstrap: call EvalBootstrapSetup for this slot count first");
    return RAISE_LEVEL + it->second->plan.depth();
}

void Context::gen_galois_keys(const Keys &k, const std::vector<uint32_t> &galois) {
    LOCK;
    if (!k || !k->s) throw std::runtime_error("key generation: secret key required");
    const int n = P->n, F = P->L + P->K;
    Mem sp = alloc((size_t)F * n);
    for (uint32_t gal : galois) {
        if (gal == 1 || evk_rot.count(gal)) continue;
        k_automorphism(T, P->galois_inverse(gal), F, k->s->p, sp->p, stream);
        evk_rot[gal] = gen_switch_key(gal, k->s->p, sp->p);
    }
    sync();
}
void Context::bootstrap_keygen(const Keys &k, int slots) {
    LOCK;
    if (slots <= 0) slots = batch;
    auto it = boot.find(slots);
    if (it == boot.end()) throw std::runtime_error("EvalBootstrapKeyGen: call EvalBootstrapSetup for this slot count first");
    gen_rot_keys(k, it->second->plan.rotations());
    gen_galois_keys(k, {(uint32_t)(2 * P->n - 1)});
}

Ct Context::rotate_galois(const Ct &a, uint32_t g) {
    LOCK;
    if (a->npoly != 2) throw std::runtime_error("automorphism: relinearise first");
    if (g == 1) return clone(a);
    auto it = evk_rot.find(g);
    if (it == evk_rot.end()) throw std::runtime_error("automorphism: no evaluation key for Galois element " + std::to_string(g));
    const int n = P->n, ell = a->ell;
    const DevicePlan &D = plan(ell);
    Mem digits = alloc((size_t)D.beta * D.width * n);
    modup(ell, a->poly(1, n), digits->p);
    auto r = std::make_shared<Ciphertext>(*a);
    r->mem = alloc((size_t)2 * ell * n);
    keyswitch_digits(ell, digits->p, it->second, a->poly(0, n), nullptr, r->poly(0, n), r->poly(1, n), g);
    return r;
}
Ct Context::conjugate(const Ct &a) { return rotate_galois(a, (uint32_t)(2 * P->n - 1)); }

Ct Context::mod_raise(const Ct &a, int ell_out, double scale_label) {
    LOCK;
    if (a->ell != 1 || a->npoly != 2) throw std::runtime_error("ModRaise: expected a two-element ciphertext on one limb");
    const int n = P->n;
    Mem coef = alloc((size_t)2 * n);
    std::vector<uint32_t> d1 = {(0u << 8) | 0u, (1u << 8) | 0u}, dall;
    k_ntt_inv(T, a->mem->p, coef->p, desc(d1), 2, stream);
    auto r = std::make_shared<Ciphertext>();
    r->npoly = 2, r->ell = ell_out, r->level = P->L - ell_out, r->nsd = 1, r->scale = scale_label, r->slots = a->slots;
    r->mem = alloc((size_t)2 * ell_out * n);
    k_modraise_lift(T, ell_out, 2, coef->p, (size_t)n, r->mem->p, stream);
    for (int i = 0; i < ell_out; i++)
        for (int p = 0; p < 2; p++) dall.push_back(((uint32_t)(p * ell_out + i) << 8) | (uint32_t)i);
    k_ntt_fwd(T, r->mem->p, desc(dall), 2 * ell_out, stream);
    return r;
}

Mem Context::encode_qp(const Plaintext &pt, int ell, double scale) {
    LOCK;
    if (pt.base || pt.real_only) throw std::logic_error("encode_qp: complex slot values expected");
    const int n = P->n, slots = pt.slots, W = ell + P->K;
    Mem vbuf = alloc((size_t)2 * slots);
    const size_t nv = std::min((size_t)slots, pt.vals.size());
    if (nv < (size_t)slots) FHE_CUDA(cudaMemsetAsync(vbuf->p, 0, sizeof(double2) * slots, stream));
    if (nv) upload(vbuf->p, pt.vals.data(), sizeof(double2) * nv);
    auto tabs = fft_device_tables(slots);
    Mem out = alloc((size_t)W * n);
    k_encode(T, (double2 *)vbuf->p, nullptr, (int)nv, slots, scale, tabs.first, tabs.second, ell, out->p, stream, P->L, P->K);
    std::vector<uint32_t> d;
    for (int t = 0; t < W; t++) d.push_back(((uint32_t)t << 8) | (uint32_t)(t < ell ? t : P->L + t - ell));
    k_ntt_fwd(T, out->p, desc(d), W, stream);
    return out;
}

// out = sum_i diag_i (.) rotate(x, rot_i).  The rotations of one stage share a hoisted ModUp.  With a QP cache (double
// hoisting) the products are formed and summed in the extended basis - diag_i encoded on the P limbs too, the unrotated
// c_b entering as P * c_b - and the stage pays for ONE ModDown instead of one per rotation.  The result stays at degree
// 2 and is rescaled once by whoever multiplies next.
Ct Context::linear_stage(const Ct &x0, const BootStage &st, const std::vector<Pt> &diag,
                         std::vector<std::tuple<int, double, Mem>> *qp_cache) {
    LOCK;
    Ct x = x0->nsd == 2 ? reduced(x0) : x0;
    static const bool single = [] { const char *e = getenv("FHEB200_BOOT_HOIST"); return e && std::string(e) == "single"; }();
    if (single || !qp_cache || forked()) {
        std::shared_ptr<Precomp> pre;
        Ct acc;
        for (size_t i = 0; i < st.rot.size(); i++) {
            Ct r = x;
            if (st.rot[i] != 0) {
                if (!pre) pre = fast_rotation_precompute(x);
                r = fast_rotation(x, st.rot[i], pre);
            }
            Ct t = mult_pt(r, diag[i]);
            acc = acc ? add(acc, t) : t;
        }
        return acc;
    }
    const int n = P->n, ell = x->ell, L = P->L, K = P->K;
    const DevicePlan &D = plan(ell);
    const int W = D.width;
    const double ps = P->sf[x->level];
    if (qp_cache->size() != diag.size()) qp_cache->assign(diag.size(), std::make_tuple(0, 0.0, Mem()));
    // P * c_b on the Q limbs
    LimbScalars pm;
    memset(&pm, 0, sizeof(pm));
    for (int i = 0; i < ell; i++) {
        u64 q = P->moduli[i], v = 1;
        for (int k = 0; k < K; k++) v = mulmod_h(v, P->moduli[L + k] % q, q);
        pm.v[i] = v, pm.s[i] = shoup_h(v, q);
    }
    Mem cP = alloc((size_t)2 * ell * n), S = alloc((size_t)2 * W * n), acc = alloc((size_t)2 * W * n);
    for (int b = 0; b < 2; b++) k_mul_scalar(T, ell, x->poly(b, n), pm, cP->p + (size_t)b * ell * n, stream);
    FHE_CUDA(cudaMemsetAsync(S->p, 0, (size_t)2 * W * n * sizeof(u64), stream));
    std::shared_ptr<Precomp> pre;
    for (size_t i = 0; i < st.rot.size(); i++) {
        auto &slot = (*qp_cache)[i];
        if (!std::get<2>(slot) || std::get<0>(slot) != ell || std::get<1>(slot) != ps)
            slot = std::make_tuple(ell, ps, encode_qp(*diag[i], ell, ps));
        const u64 *pt = std::get<2>(slot)->p;
        const uint32_t g = P->galois_for_rotation(st.rot[i]);
        if (g == 1u) {
            k_dh_accumulate(T, D, nullptr, cP->p, pt, S->p, 1u, stream);
            continue;
        }
        auto it = evk_rot.find(g);
        if (it == evk_rot.end()) throw std::runtime_error("EvalBootstrap: no evaluation key for rotation index " + std::to_string(st.rot[i]));
        if (!pre) pre = fast_rotation_precompute(x);
        k_inner_product(T, D, pre->digits->p, it->second, acc->p, stream);
        k_dh_accumulate(T, D, acc->p, cP->p, pt, S->p, g, stream);
    }
    auto r = std::make_shared<Ciphertext>(*x);
    r->mem = alloc((size_t)2 * ell * n);
    moddown(ell, S->p, nullptr, nullptr, r->poly(0, n), r->poly(1, n), 1u);
    r->nsd = x->nsd + 1;
    r->scale = x->scale * ps;
    return r;
}

Ct Context::bootstrap(const Ct &a, int stop_after) {
    LOCK;
    if (!a || a->npoly != 2) throw std::runtime_error("EvalBootstrap: expected a relinearised ciphertext");
    const int slots = a->slots > 0 ? a->slots : batch;
    auto it = boot.find(slots);
    if (it == boot.end())
        throw std::runtime_error("EvalBootstrap: no EvalBootstrapSetup for " + std::to_string(slots) + " slots");
    BootState &B = *it->second;
    const BootPlan &bp = B.plan;
    const int n = P->n;
    static const bool verbose = getenv("FHEB200_BOOT_VERBOSE") != nullptr;
    // FHEB200_BOOT_VERBOSE: synchronised wall time and kernel launches of every phase
    auto t_last = std::chrono::steady_clock::now();
    unsigned long long l_last = g_launches.load();
    auto phase = [&](const char *name, const Ct &c) {
        if (!verbose) return;
        sync();
        auto now = std::chrono::steady_clock::now();
        unsigned long long l = g_launches.load();
        fprintf(stderr, "bootstrap %-14s %8.2f ms %6llu launches -> level %d nsd %d limbs %d\n", name,
                std::chrono::duration<double, std::milli>(now - t_last).count(), l - l_last, c->level, c->nsd, c->ell);
        t_last = now, l_last = l;
    };

    // 1. last two limbs -> one limb carrying msg_scale * m
    Ct x = a->nsd == 2 ? reduced(a) : a;
    if (x->ell < 2) throw std::runtime_error("EvalBootstrap: the ciphertext has no level left (bootstrap one level earlier)");
    if (x->ell > 2) x = level_reduce(x, x->ell - 2);
    {
        const long double mult = (long double)bp.msg_scale * (long double)P->moduli[1] / (long double)x->scale;
        if (mult < 2.0L) throw std::runtime_error("EvalBootstrap: scale too large for the correction factor");
        LimbScalars s = scalars_for(mult, 2, true);
        auto y = std::make_shared<Ciphertext>(*x);
        y->mem = alloc((size_t)2 * 2 * n);
        for (int p = 0; p < 2; p++) k_mul_scalar(T, 2, x->poly(p, n), s, y->poly(p, n), stream);
        y->nsd = 2;
        y->scale = x->scale * (double)roundl(mult);
        x = mod_reduce(y); // scale == msg_scale up to 2^-40 relative
    }
    phase("scale-down", x);
    if (stop_after == 1) return x;
    // 2. ModRaise
    Ct r = mod_raise(x, P->L - RAISE_LEVEL, bp.raise_scale);
    r->slots = 2 * slots;
    phase("ModRaise", r);
    if (stop_after == 5) return r;
    // 3. SubSum
    for (int step = slots; step < n / 2; step <<= 1) r = add(r, rotate(r, step));
    phase("SubSum", r);
    if (stop_after == 2) return r;
    // 4. CoeffsToSlots + conjugate
    for (size_t i = 0; i < bp.cts.size(); i++) r = linear_stage(r, bp.cts[i], B.cts_pt[i], &B.cts_qp[i]);
    r = reduced(r);
    r = add(r, conjugate(r));
    phase("CoeffsToSlots", r);
    if (stop_after == 3) return r;
    // 5. EvalMod
    if (bp.even) {
        r = add_const(r, -bp.shift);                  // u = (x - delta) / (1 + delta)
        Ct sq = mult(r, r);
        r = add_const(add(sq, sq), -1.0);             // y = T_2(u)
    }
    r = eval_chebyshev(r, bp.cheb, -1.0, 1.0);        // sum_j c_2j T_j(y) = cos(w u)
    phase("Chebyshev", r);
    for (int i = 0; i < bp.r; i++) {
        Ct sq = mult(r, r);
        r = add_const(add(sq, sq), -1.0);
    }
    phase("double-angle", r);
    if (stop_after == 4) return r;
    // 6. SlotsToCoeffs + fold of the two halves
    for (size_t i = 0; i < bp.stc.size(); i++) r = linear_stage(r, bp.stc[i], B.stc_pt[i], &B.stc_qp[i]);
    r = add(r, rotate(r, slots));
    if (r.use_count() > 1) r = std::make_shared<Ciphertext>(*r);
    r->slots = slots;
    phase("SlotsToCoeffs", r);
    return r;
}

} // namespace fheb200
