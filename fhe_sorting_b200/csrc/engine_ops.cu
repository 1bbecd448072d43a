// engine_ops.cu — scheme operations: key switching pipeline, rotation (plain and hoisted), multiplication with
// relinearisation, rescale, and the FLEXIBLEAUTOEXT level/scale alignment used by add/sub/mult.
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "engine.h"

namespace fheb200 {

#define LOCK                                           \
    CtxLock lock__(*this);                             \
    FHE_CUDA(cudaSetDevice(device))

static Ct new_like(Context *C, const Ct &a, int npoly, int ell) {
    auto c = std::make_shared<Ciphertext>(*a);
    c->npoly = npoly;
    c->ell = ell;
    c->level = C->P->L - ell;
    c->mem = C->alloc((size_t)npoly * ell * C->P->n);
    return c;
}

Ct Context::clone(const Ct &a) {
    LOCK;
    Ct c = new_like(this, a, a->npoly, a->ell);
    FHE_CUDA(cudaMemcpyAsync(c->mem->p, a->mem->p, (size_t)a->npoly * a->ell * P->n * sizeof(u64), cudaMemcpyDeviceToDevice, stream));
    return c;
}

// ------------------------------------------------------------------ key switching pipeline
// FHEB200_CONV=scalar keeps the CUDA-core base conversion (keyswitch.cu) for A/B measurements; both are exact
static bool tensor_core_conv() {
    static const bool on = [] {
        const char *e = getenv("FHEB200_CONV");
        return !(e && std::string(e) == "scalar");
    }();
    return on;
}
void Context::modup(int ell, const u64 *c, u64 *digits) {
    const DevicePlan &D = plan(ell);
    const int n = P->n;
    Mem coef = alloc((size_t)ell * n);
    if (tensor_core_conv()) { // (Q_j/q_i)^-1 rides on the inverse NTT's final scaling; contraction on the integer tensor cores
        k_ntt_inv(T, c, coef->p, D.desc_q, ell, stream, D.up_scale);
        k_modup_tc(T, D, coef->p, c, digits, stream);
    } else {
        k_ntt_inv(T, c, coef->p, D.desc_q, ell, stream);
        k_modup(T, D, coef->p, c, digits, stream);
    }
    k_ntt_fwd(T, digits, D.desc_modup, D.n_modup, stream);
}
void Context::keyswitch_digits(int ell, const u64 *digits, const u64 *evk, const u64 *add0, const u64 *add1, u64 *out0,
                               u64 *out1, uint32_t galois) {
    const DevicePlan &D = plan(ell);
    Mem acc = alloc((size_t)2 * D.width * P->n);
    k_inner_product(T, D, digits, evk, acc->p, stream);
    g_op_count[galois == 1u ? 1 : 3]++;
    moddown(ell, acc->p, add0, add1, out0, out1, galois);
}
// acc[2][ell + K][n] (evaluation domain, overwritten) -> out_b = sigma_g(round(acc_b / P) + add_b) on the ell Q limbs
void Context::moddown(int ell, u64 *acc, const u64 *add0, const u64 *add1, u64 *out0, u64 *out1, uint32_t galois) {
    moddown_plan(plan(ell), acc, add0, add1, out0, out1, galois);
}
// D: the level's own plan (special limbs = P) or a merged one (special limbs = q_last and P, Context::plan_merged)
void Context::moddown_plan(const DevicePlan &D, u64 *acc, const u64 *add0, const u64 *add1, u64 *out0, u64 *out1, uint32_t galois) {
    const int n = P->n, ell = D.ell;
    Mem conv = alloc((size_t)2 * ell * n);
    if (tensor_core_conv() || D.K != P->K) { // (the CUDA-core conversion is instantiated for K special limbs only)
        k_ntt_inv(T, acc, acc, D.desc_p2, 2 * D.K, stream, D.down_scale);
        Mem over;
        if (D.exact_lift) { // the quotient is final here (no rescale follows to absorb the lift's overshoot): make the lift exact
            over = alloc((size_t)2 * n);
            k_lift_overflow(T, D, acc, over->p, stream);
        }
        k_moddown_conv_tc(T, D, acc, conv->p, stream, over ? over->p : nullptr);
    } else {
        k_ntt_inv(T, acc, D.desc_p2, 2 * D.K, stream);
        k_moddown_conv(T, D, acc, conv->p, stream);
    }
    if (k_moddown_ntt_fused(T, D, conv->p, acc, add0, add1, out0, out1, galois == 1u ? 1u : P->galois_inverse(galois), stream)) return;
    k_ntt_fwd(T, conv->p, D.desc_q2, 2 * ell, stream);
    k_moddown_finish(T, D, acc, conv->p, add0, add1, out0, out1, galois, stream);
}

Ct Context::rotate(const Ct &a, int index) {
    LOCK;
    if (a->npoly != 2) throw std::runtime_error("EvalRotate: relinearise first");
    uint32_t g = P->galois_for_rotation(index);
    if (g == 1) return clone(a);
    auto it = evk_rot.find(g);
    if (it == evk_rot.end()) throw std::runtime_error("EvalRotate: no evaluation key for rotation index " + std::to_string(index));
    const int n = P->n, ell = a->ell;
    const DevicePlan &D = plan(ell);
    Mem digits = alloc((size_t)D.beta * D.width * n);
    modup(ell, a->poly(1, n), digits->p);
    count_algo(8.0 * n * (4.0 * ell + 2.0 * D.beta * D.width));
    Ct r = new_like(this, a, 2, ell);
    keyswitch_digits(ell, digits->p, it->second, a->poly(0, n), nullptr, r->poly(0, n), r->poly(1, n), g);
    return r;
}
std::shared_ptr<Precomp> Context::fast_rotation_precompute(const Ct &a) {
    LOCK;
    if (a->npoly != 2) throw std::runtime_error("EvalFastRotationPrecompute: relinearise first");
    const DevicePlan &D = plan(a->ell);
    auto pre = std::make_shared<Precomp>();
    pre->ell = a->ell;
    pre->digits = alloc((size_t)D.beta * D.width * P->n);
    modup(a->ell, a->poly(1, P->n), pre->digits->p);
    return pre;
}
Ct Context::fast_rotation(const Ct &a, int index, const std::shared_ptr<Precomp> &pre) {
    LOCK;
    uint32_t g = P->galois_for_rotation(index);
    if (g == 1) return clone(a);
    if (!pre || pre->ell != a->ell) throw std::runtime_error("EvalFastRotation: precomputation does not match the ciphertext");
    auto it = evk_rot.find(g);
    if (it == evk_rot.end()) throw std::runtime_error("EvalFastRotation: no evaluation key for rotation index " + std::to_string(index));
    const int n = P->n;
    {
        const DevicePlan &D = plan(a->ell);
        count_algo(8.0 * n * (4.0 * a->ell + 2.0 * D.beta * D.width));
    }
    Ct r = new_like(this, a, 2, a->ell);
    keyswitch_digits(a->ell, pre->digits->p, it->second, a->poly(0, n), nullptr, r->poly(0, n), r->poly(1, n), g);
    return r;
}

// ------------------------------------------------------------------ rescale / level reduce
void Context::rescale_poly(int ell, int npoly, const u64 *in, u64 *out) {
    const int n = P->n;
    const LevelPlan &H = P->plan(ell);
    LimbScalars qi;
    memset(&qi, 0, sizeof(qi));
    for (int i = 0; i < ell - 1; i++) qi.v[i] = H.ql_inv[i], qi.s[i] = H.ql_inv_s[i];
    // the dropped limb of poly p goes through the inverse NTT straight out of `in` (slot p * ell of both views)
    Mem last = alloc(((size_t)(npoly - 1) * ell + 1) * n), tmp = alloc((size_t)npoly * (ell - 1) * n);
    std::vector<uint32_t> dl, dt;
    for (int p = 0; p < npoly; p++) dl.push_back(((uint32_t)(p * ell) << 8) | (uint32_t)(ell - 1));
    for (int i = 0; i < ell - 1; i++) // limb-major: the polys' limbs of one modulus share their twiddles in L2
        for (int p = 0; p < npoly; p++) dt.push_back(((uint32_t)(p * (ell - 1) + i) << 8) | (uint32_t)i);
    k_ntt_inv(T, in + (size_t)(ell - 1) * n, last->p, desc(dl), npoly, stream);
    if (k_rescale_ntt_fused(T, ell, npoly, last->p, in, tmp->p, out, desc(dt), qi, stream)) return;
    k_rescale_lift(T, ell, npoly, last->p, (size_t)ell * n, tmp->p, stream);
    k_ntt_fwd(T, tmp->p, desc(dt), npoly * (ell - 1), stream);
    k_rescale_finish(T, ell - 1, npoly, in, tmp->p, qi, out, stream);
}
Ct Context::mod_reduce(const Ct &a) {
    LOCK;
    if (a->ell < 2) throw std::runtime_error("rescale: no limbs left (multiplicative depth exhausted)");
    Ct r = new_like(this, a, a->npoly, a->ell - 1);
    g_op_count[0]++;
    count_algo(8.0 * P->n * a->npoly * (2.0 * a->ell - 1.0));
    rescale_poly(a->ell, a->npoly, a->mem->p, r->mem->p);
    r->nsd = a->nsd - 1;
    r->scale = a->scale / (double)P->moduli[a->ell - 1];
    return r;
}
Ct Context::rescale(const Ct &a) { return mod_reduce(a); }
Ct Context::reduced(const Ct &a) {
    LOCK;
    Ct r = a->reduced_cache;
    if (!r || !usable(r->mem)) { // (inside a forked region a copy made by another stream may not be complete yet)
        r = mod_reduce(a);
        // kept for later calls unless `a` is shared with the other streams of a forked region
        if (!forked() || a->mem->stream == stream) a->reduced_cache = r;
    }
    r->slots = a->slots; // SetSlots may have been called on the handle since
    return r;
}
Ct Context::level_reduce(const Ct &a, int levels) {
    LOCK;
    if (levels <= 0) return a;
    if (a->ell - levels < 1) throw std::runtime_error("LevelReduce: not enough limbs");
    const int n = P->n, ne = a->ell - levels;
    Ct r = new_like(this, a, a->npoly, ne);
    for (int p = 0; p < a->npoly; p++)
        FHE_CUDA(cudaMemcpyAsync(r->poly(p, n), a->poly(p, n), (size_t)ne * n * sizeof(u64), cudaMemcpyDeviceToDevice, stream));
    return r;
}

// ------------------------------------------------------------------ scalar helpers
Ct Context::mult_core_const(const Ct &a, double c) { // EvalMultCoreInPlace: times round(c * sf[level])
    LOCK;
    const int n = P->n;
    double sfl = P->sf[a->level];
    LimbScalars s = scalars_for((long double)c * (long double)sfl, a->ell, true);
    Ct r = new_like(this, a, a->npoly, a->ell);
    for (int p = 0; p < a->npoly; p++) k_mul_scalar(T, a->ell, a->poly(p, n), s, r->poly(p, n), stream);
    count_algo(8.0 * n * a->npoly * 2.0 * a->ell);
    r->nsd = a->nsd + 1;
    r->scale = a->scale * sfl;
    return r;
}
std::vector<Ct> Context::lincomb(const std::vector<Ct> &in, const std::vector<std::vector<double>> &w) {
    LOCK;
    std::vector<Ct> out(w.size());
    if (in.empty() || w.empty()) return out;
    const int n = P->n, ell = in[0]->ell, npoly = in[0]->npoly, count = (int)in.size();
    for (auto &c : in)
        if (c->ell != ell || c->npoly != npoly || c->nsd != in[0]->nsd)
            throw std::runtime_error("lincomb: inputs must share level, degree and size");
    const double sfl = P->sf[in[0]->level];
    std::vector<int> live; // outputs with at least one non-zero weight
    for (size_t r = 0; r < w.size(); r++) {
        bool any = false;
        for (double x : w[r]) any = any || x != 0.0;
        if (any) live.push_back((int)r);
    }
    // FHEB200_LINCOMB=6: the sums on the integer tensor cores, 8 outputs per pass over the inputs (baseconv.cu: lc_tc)
    static const bool tc = [] { const char *e = getenv("FHEB200_LINCOMB"); return e ? atoi(e) == 6 : true; }();
    const bool use_tc = tc && count >= 8;
    count_algo(8.0 * n * npoly * ell * ((double)count + (double)live.size()));
    const size_t per_pass = use_tc ? 8 : 4;
    for (size_t g0 = 0; g0 < live.size(); g0 += per_pass) {
        const int G = (int)std::min<size_t>(per_pass, live.size() - g0);
        // one blob: count input pointers, G output pointers, G*count*ell split weights
        std::vector<u64> blob((size_t)count + G + (size_t)G * count * ell);
        for (int t = 0; t < count; t++) blob[t] = (u64)(uintptr_t)in[t]->mem->p;
        for (int g = 0; g < G; g++) {
            Ct o = new_like(this, in[0], npoly, ell);
            o->nsd = in[0]->nsd + 1;
            o->scale = in[0]->scale * sfl;
            out[live[g0 + g]] = o;
            blob[count + g] = (u64)(uintptr_t)o->mem->p;
            for (int t = 0; t < count; t++) {
                LimbScalars sc = scalars_for((long double)w[live[g0 + g]][t] * (long double)sfl, ell, false);
                for (int i = 0; i < ell; i++)
                    blob[(size_t)count + G + ((size_t)g * count + t) * ell + i] = (sc.v[i] & 0x3fffffffull) | ((sc.v[i] >> 30) << 32);
            }
        }
        Mem d = alloc(blob.size());
        FHE_CUDA(cudaMemcpyAsync(d->p, blob.data(), blob.size() * sizeof(u64), cudaMemcpyHostToDevice, stream));
        if (use_tc) {
            Mem tables = alloc(lincomb_tc_scratch_words(ell, count));
            k_lincomb_tc(T, ell, npoly * ell, count, G, (const u64 *const *)d->p, (const uint2 *)(d->p + count + G),
                         (u64 *const *)(d->p + count), tables->p, stream);
        } else {
            k_lincomb(T, ell, npoly * ell, count, G, (const u64 *const *)d->p, (const uint2 *)(d->p + count + G),
                      (u64 *const *)(d->p + count), stream);
        }
    }
    return out;
}
Ct Context::mult_const(const Ct &a, double c) {
    LOCK;
    Ct x = a->nsd == 2 ? reduced(a) : a;
    return mult_core_const(x, c);
}
Ct Context::add_const(const Ct &a, double c) {
    LOCK;
    const int n = P->n;
    LimbScalars s = scalars_for((long double)c * (long double)a->scale, a->ell, false);
    Ct r = clone(a);
    k_add_scalar(T, a->ell, r->poly(0, n), s, r->poly(0, n), stream);
    return r;
}
Ct Context::negate(const Ct &a) {
    LOCK;
    Ct r = new_like(this, a, a->npoly, a->ell);
    for (int p = 0; p < a->npoly; p++) k_negate(T, a->ell, a->poly(p, P->n), r->poly(p, P->n), stream);
    return r;
}

} // namespace fheb200
