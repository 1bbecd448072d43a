// engine_arith.cu — EvalAdd / EvalSub / EvalMult (ct, plaintext, constant) with FLEXIBLEAUTOEXT alignment.
// Reference call sites: src/comparison.cpp:13-38, src/sign.cpp:23-56, src/sort_algo.h:319-364,454-504,706-742.
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <stdexcept>

#include "engine.h"

namespace fheb200 {

#define LOCK                                           \
    CtxLock lock__(*this);                             \
    FHE_CUDA(cudaSetDevice(device))

// bring x (shallower) to the level, degree and scale of t — restates OpenFHE's AdjustLevelsAndDepthInPlace
static Ct raise_to(Context *C, Ct x, const Ct &t) {
    const Params &P = *C->P;
    const double scf1 = x->scale, scf = P.sf[x->level], q1 = (double)P.moduli[x->ell - 1];
    if (x->nsd == 2) {
        if (t->nsd == 2) {
            x = C->mult_core_const(x, t->scale / scf1 * q1 / scf);
            x = C->mod_reduce(x);
            if (x->level < t->level) x = C->level_reduce(x, t->level - x->level);
        } else if (x->level + 1 == t->level) {
            x = C->mod_reduce(x);
        } else {
            x = C->mult_core_const(x, P.sf_big[t->level - 1] / scf1 * q1 / scf);
            x = C->mod_reduce(x);
            if (x->level + 1 < t->level) x = C->level_reduce(x, t->level - x->level - 1);
            x = C->mod_reduce(x);
        }
    } else {
        if (t->nsd == 2) {
            x = C->mult_core_const(x, t->scale / scf1 / scf);
            x = C->level_reduce(x, t->level - x->level);
        } else {
            x = C->mult_core_const(x, P.sf_big[t->level - 1] / scf1 / scf);
            if (x->level + 1 < t->level) x = C->level_reduce(x, t->level - x->level - 1);
            x = C->mod_reduce(x);
        }
    }
    if (x.get() == nullptr || x->level != t->level || x->nsd != t->nsd)
        throw std::logic_error("fheb200: level alignment failed");
    if (x.use_count() > 1) x = std::make_shared<Ciphertext>(*x); // never mutate a shared object
    x->scale = t->scale;
    return x;
}

void Context::adjust_pair(Ct &a, Ct &b) {
    if (a->level < b->level)
        a = raise_to(this, a, b);
    else if (b->level < a->level)
        b = raise_to(this, b, a);
    else if (a->nsd < b->nsd)
        a = mult_core_const(a, 1.0);
    else if (b->nsd < a->nsd)
        b = mult_core_const(b, 1.0);
}
void Context::adjust_pair_to_one(Ct &a, Ct &b) {
    adjust_pair(a, b);
    if (a->nsd == 2) {
        bool same = a.get() == b.get();
        a = reduced(a);
        b = same ? a : reduced(b);
    }
}

static void check_pair(const Ct &a, const Ct &b) {
    if (!a || !b) throw std::invalid_argument("fheb200: null ciphertext");
    if (a->npoly != 2 || b->npoly != 2) throw std::runtime_error("fheb200: expected relinearised ciphertexts");
}

Ct Context::add(const Ct &a0, const Ct &b0) {
    LOCK;
    check_pair(a0, b0);
    Ct a = a0, b = b0;
    adjust_pair(a, b);
    const int n = P->n;
    auto r = std::make_shared<Ciphertext>(*a);
    r->mem = alloc((size_t)2 * a->ell * n);
    r->slots = std::max(a->slots, b->slots);
    for (int p = 0; p < 2; p++) k_binary(T, OP_ADD, a->ell, a->poly(p, n), b->poly(p, n), r->poly(p, n), stream);
    count_algo(8.0 * n * 6.0 * a->ell);
    return r;
}
Ct Context::sub(const Ct &a0, const Ct &b0) {
    LOCK;
    check_pair(a0, b0);
    Ct a = a0, b = b0;
    adjust_pair(a, b);
    const int n = P->n;
    auto r = std::make_shared<Ciphertext>(*a);
    r->mem = alloc((size_t)2 * a->ell * n);
    r->slots = std::max(a->slots, b->slots);
    for (int p = 0; p < 2; p++) k_binary(T, OP_SUB, a->ell, a->poly(p, n), b->poly(p, n), r->poly(p, n), stream);
    count_algo(8.0 * n * 6.0 * a->ell);
    return r;
}
Ct Context::add_many(const std::vector<Ct> &v) {
    LOCK;
    if (v.empty()) throw std::invalid_argument("EvalAddMany: empty input");
    Ct acc = v[0];
    for (size_t i = 1; i < v.size(); i++) acc = add(acc, v[i]);
    return v.size() == 1 ? clone(acc) : acc;
}

Mem Context::encoding_for(const Pt &pt, int ell, double scale) {
    if (pt->mem && pt->ell == ell && pt->scale == scale && usable(pt->mem)) return pt->mem;
    for (auto &e : pt->alt)
        if (std::get<0>(e) == ell && std::get<1>(e) == scale && usable(std::get<2>(e))) return std::get<2>(e);
    Mem m = encode_at(*pt, ell, scale);
    // inside a forked region an encoding made by one stream is unusable by its siblings (usable()), so every stream keeps
    // its own entries: room for 4 per stream, or they evict each other on every call
    const size_t cap = forked_ ? 4 * (size_t)(fork_k_ + 1) : 4;
    while (pt->alt.size() >= cap) pt->alt.erase(pt->alt.begin());
    pt->alt.emplace_back(ell, scale, m);
    return m;
}

// ct + plaintext: the plaintext is (re-)encoded at exactly the ciphertext's limbs and scale
Ct Context::add_pt(const Ct &a, const Pt &p, bool negate_pt) {
    LOCK;
    if (!a || !p) throw std::invalid_argument("EvalAdd: null operand");
    const int n = P->n;
    Mem enc = encoding_for(p, a->ell, a->scale);
    Ct r = clone(a);
    r->slots = std::max(a->slots, p->slots);
    k_binary(T, negate_pt ? OP_SUB : OP_ADD, a->ell, r->poly(0, n), enc->p, r->poly(0, n), stream);
    return r;
}
// ct * plaintext: ciphertext brought to degree 1, plaintext encoded at sf[level]; result degree 2
Ct Context::mult_pt(const Ct &a0, const Pt &p) {
    LOCK;
    if (!a0 || !p) throw std::invalid_argument("EvalMult: null operand");
    Ct a = a0->nsd == 2 ? reduced(a0) : a0;
    const int n = P->n;
    const double ps = P->sf[a->level];
    Mem enc = encoding_for(p, a->ell, ps);
    auto r = std::make_shared<Ciphertext>(*a);
    r->mem = alloc((size_t)2 * a->ell * n);
    r->slots = std::max(a->slots, p->slots);
    k_mul_pt(T, a->ell, a->poly(0, n), a->poly(1, n), enc->p, r->poly(0, n), r->poly(1, n), 0, stream);
    count_algo(8.0 * n * 5.0 * a->ell);
    r->nsd = a->nsd + 1;
    r->scale = a->scale * ps;
    return r;
}
// acc + ct * plaintext in one pass over the operands (the mask-and-accumulate loops of the sort drivers: 7 limb streams
// instead of the 11 of EvalMult followed by EvalAdd).  Bit-identical to add(acc, mult_pt(a, p)); falls back to exactly that
// when the accumulator does not already sit at the product's limbs, degree and scale.  FHEB200_MULT_ADD=0 forces the fallback.
Ct Context::mult_pt_add(const Ct &acc, const Ct &a0, const Pt &p) {
    static const bool fused = [] { const char *e = getenv("FHEB200_MULT_ADD"); return !(e && *e == '0'); }();
    if (!acc) return mult_pt(a0, p);
    {
        LOCK;
        if (!a0 || !p) throw std::invalid_argument("EvalMult: null operand");
        Ct a = a0->nsd == 2 ? reduced(a0) : a0;
        const double ps = P->sf[a->level];
        if (fused && acc->npoly == 2 && a->npoly == 2 && acc->ell == a->ell && acc->level == a->level && acc->nsd == a->nsd + 1 &&
            acc->scale == a->scale * ps && usable(acc->mem)) {
            const int n = P->n;
            Mem enc = encoding_for(p, a->ell, ps);
            Ct r = (acc.use_count() == 1 && acc->mem.use_count() == 1) ? acc : clone(acc);
            r->reduced_cache.reset();
            r->slots = std::max(std::max(acc->slots, a->slots), p->slots);
            k_mul_pt(T, a->ell, a->poly(0, n), a->poly(1, n), enc->p, r->poly(0, n), r->poly(1, n), 1, stream);
            count_algo(8.0 * n * 7.0 * a->ell);
            return r;
        }
    }
    return add(acc, mult_pt(a0, p));
}
// ct * ct with relinearisation (EvalMult, EvalMultAndRelinearize, EvalSquare)
Ct Context::mult(const Ct &a0, const Ct &b0) { return relinearize(mult_lazy(a0, b0)); }

// The tensor product alone.  Paterson-Stockmeyer adds half of its products to another product before anything is
// multiplied by them again, so those can share one relinearisation (poly_eval.cu: eval_inner).
Ct Context::mult_lazy(const Ct &a0, const Ct &b0) {
    LOCK;
    check_pair(a0, b0);
    Ct a = a0, b = b0;
    adjust_pair_to_one(a, b);
    const int n = P->n, ell = a->ell;
    auto r = std::make_shared<Ciphertext>(*a);
    r->npoly = 3;
    r->mem = alloc((size_t)3 * ell * n);
    r->slots = std::max(a->slots, b->slots);
    k_tensor(T, ell, a->poly(0, n), a->poly(1, n), b->poly(0, n), b->poly(1, n), r->poly(0, n), r->poly(1, n), r->poly(2, n), stream);
    count_algo(8.0 * n * 7.0 * ell); // tensor: 4 polynomials in, 3 out
    r->nsd = a->nsd + b->nsd;
    r->scale = a->scale * b->scale;
    return r;
}
Ct Context::relinearize(const Ct &a) {
    LOCK;
    if (!a) throw std::invalid_argument("fheb200: null ciphertext");
    if (a->npoly == 2) return a;
    if (a->npoly != 3) throw std::runtime_error("relinearize: expected 3 polynomials");
    if (!evk_mult) throw std::runtime_error("EvalMult: relinearisation key missing (call EvalMultKeyGen)");
    const int n = P->n, ell = a->ell;
    const DevicePlan &D = plan(ell);
    Mem digits = alloc((size_t)D.beta * D.width * n);
    modup(ell, a->poly(2, n), digits->p);
    auto r = std::make_shared<Ciphertext>(*a);
    r->npoly = 2;
    r->mem = alloc((size_t)2 * ell * n);
    keyswitch_digits(ell, digits->p, evk_mult, a->poly(0, n), a->poly(1, n), r->poly(0, n), r->poly(1, n), 1u);
    count_algo(8.0 * n * (5.0 * ell + 2.0 * D.beta * D.width)); // 3 polynomials in, the evk stream, 2 out
    return r;
}
// relinearise and rescale in one division (VERDICT r1 "next lever", DESIGN.md round-2 results): ModDown divides the key-switched
// accumulator by P and rounds, rescale divides the sum by q_last and rounds; here d0, d1 are lifted into the accumulator (times P)
// and one ModDown by P * q_last (Context::plan_merged) does both: one forward NTT batch over 2 (ell - 1) limbs instead of two
// batches over 2 ell and 2 (ell - 1), and no separate pass over the dropped limb.  Same level and scale bookkeeping as
// reduced(relinearize(a)); the integers differ by the rounding of one division instead of two (at most a few units per
// coefficient: tests/test_gpu_scheme.py).  FHEB200_MERGED_RESCALE=0 restores the two steps.
Ct Context::relinearize_reduced(const Ct &a) {
    LOCK;
    static const bool merged = [] { const char *e = getenv("FHEB200_MERGED_RESCALE"); return !(e && *e == '0'); }();
    if (!a) throw std::invalid_argument("fheb200: null ciphertext");
    if (!merged || a->npoly != 3 || a->nsd != 2 || a->ell < 2) { // what the callers did before: relinearise, rescale only a degree-2 result
        Ct r = relinearize(a);
        return r->nsd == 2 ? reduced(r) : r;
    }
    if (!evk_mult) throw std::runtime_error("EvalMult: relinearisation key missing (call EvalMultKeyGen)");
    const int n = P->n, ell = a->ell;
    const DevicePlan &D = plan(ell);
    const DevicePlan &Dm = plan_merged(ell);
    g_op_count[2]++;
    Mem digits = alloc((size_t)D.beta * D.width * n);
    modup(ell, a->poly(2, n), digits->p);
    Mem acc = alloc((size_t)2 * D.width * n);
    k_inner_product(T, D, digits->p, evk_mult, acc->p, stream);
    LimbScalars pm; // P mod q_i
    memset(&pm, 0, sizeof(pm));
    for (int i = 0; i < ell; i++) {
        const u64 qi = P->moduli[i];
        u64 r = 1;
        for (int k = 0; k < P->K; k++) r = mulmod_h(r, P->moduli[P->L + k] % qi, qi);
        pm.v[i] = r, pm.s[i] = shoup_h(r, qi);
    }
    for (int b = 0; b < 2; b++) k_scalar_mad(T, ell, a->poly(b, n), pm, acc->p + (size_t)b * D.width * n, stream);
    auto r = std::make_shared<Ciphertext>(*a);
    r->npoly = 2;
    r->ell = ell - 1;
    r->level = P->L - r->ell;
    r->reduced_cache.reset();
    r->mem = alloc((size_t)2 * (ell - 1) * n);
    moddown_plan(Dm, acc->p, nullptr, nullptr, r->poly(0, n), r->poly(1, n), 1u);
    count_algo(8.0 * n * (5.0 * ell + 2.0 * D.beta * D.width + 2.0 * (ell - 1)));
    r->nsd = a->nsd - 1;
    r->scale = a->scale / (double)P->moduli[ell - 1];
    return r;
}
Ct Context::add_lazy(const Ct &a0, const Ct &b0) {
    LOCK;
    if (!a0 || !b0) throw std::invalid_argument("fheb200: null ciphertext");
    if (a0->npoly == 2 && b0->npoly == 2) return add(a0, b0);
    Ct a = a0, b = b0;
    adjust_pair(a, b); // the alignment helpers work on any number of polynomials
    if (a->npoly < b->npoly) std::swap(a, b);
    const int n = P->n;
    auto r = std::make_shared<Ciphertext>(*a);
    r->mem = alloc((size_t)a->npoly * a->ell * n);
    r->slots = std::max(a->slots, b->slots);
    for (int p = 0; p < a->npoly; p++) {
        if (p < b->npoly) k_binary(T, OP_ADD, a->ell, a->poly(p, n), b->poly(p, n), r->poly(p, n), stream);
        else FHE_CUDA(cudaMemcpyAsync(r->poly(p, n), a->poly(p, n), (size_t)a->ell * n * sizeof(u64), cudaMemcpyDeviceToDevice, stream));
    }
    return r;
}

} // namespace fheb200
