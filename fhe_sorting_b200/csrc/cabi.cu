// cabi.cu — extern "C" boundary (include/fhe_b200.h) over the engine.
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>

#include "../../include/fhe_b200.h"
#include "engine.h"

using namespace fheb200;

struct fheb200_ctx { std::shared_ptr<Context> c; };
struct fheb200_keys { Keys k; };
struct fheb200_ct { Ct c; };
struct fheb200_pt { Pt p; };
struct fheb200_pre { std::shared_ptr<Precomp> p; };

static thread_local std::string g_err;
namespace fheb200 { void set_last_error(const char *m) { g_err = m; } }
// FHEB200_TRACE=1: per entry point, number of calls, kernels launched and synchronised wall time (printed at exit).
// Tracing synchronises the device after every call, so it measures GPU time per operation, not pipelined latency.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
namespace {
struct TraceRow { unsigned long long calls = 0, launches = 0; double ms = 0; };
std::map<std::string, TraceRow> g_trace;
std::mutex g_trace_mu;
bool trace_on() {
    static const bool on = [] {
        const char *e = getenv("FHEB200_TRACE");
        bool v = e && *e && *e != '0';
        if (v) atexit([] {
            double tot = 0;
            for (auto &kv : g_trace) tot += kv.second.ms;
            fprintf(stderr, "%-36s %9s %10s %11s %7s\n", "fheb200 entry point", "calls", "launches", "ms (sync)", "share");
            for (auto &kv : g_trace)
                fprintf(stderr, "%-36s %9llu %10llu %11.2f %7.3f\n", kv.first.c_str(), kv.second.calls, kv.second.launches,
                        kv.second.ms, kv.second.ms / (tot > 0 ? tot : 1));
        });
        return v;
    }();
    return on;
}
struct OpTrace {
    const char *name;
    unsigned long long l0 = 0;
    std::chrono::steady_clock::time_point t0;
    bool on;
    explicit OpTrace(const char *n) : name(n), on(trace_on()) {
        if (on) l0 = g_launches.load(), t0 = std::chrono::steady_clock::now();
    }
    ~OpTrace() {
        if (!on) return;
        static const bool host_only = [] { const char *e = getenv("FHEB200_TRACE"); return e && *e == '2'; }();
        if (!host_only) cudaDeviceSynchronize(); // FHEB200_TRACE=2: time spent in the call by the launching thread only
        double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::lock_guard<std::mutex> g(g_trace_mu);
        TraceRow &r = g_trace[name];
        r.calls++, r.launches += g_launches.load() - l0, r.ms += ms;
    }
};
} // namespace
#define GUARD(...)                          \
    try {                                   \
        OpTrace trace__(__func__);          \
        __VA_ARGS__;                            \
        return 0;                           \
    } catch (const std::exception &e) {     \
        g_err = e.what();                   \
        return 1;                           \
    } catch (...) {                         \
        g_err = "unknown error";            \
        return 1;                           \
    }
static void need(const void *p, const char *what) {
    if (!p) throw std::invalid_argument(std::string("null argument: ") + what);
}
static fheb200_ct *wrap(const Ct &c) { return new fheb200_ct{c}; }

extern "C" {

const char *fheb200_last_error(void) { return g_err.c_str(); }
const char *fheb200_version(void) { return "fheb200 0.1 (sm_100a)"; }

int fheb200_ctx_create(const fheb200_params *p, fheb200_ctx **out) {
    GUARD({
        need(p, "params");
        need(out, "out");
        auto c = Context::create(p->log_ring_dim, p->mult_depth, p->scaling_mod_size, p->first_mod_size ? p->first_mod_size : 60,
                                 p->num_large_digits ? p->num_large_digits : 3, p->batch_size, p->device);
        *out = new fheb200_ctx{c};
    })
}
int fheb200_ctx_release(fheb200_ctx *ctx) { GUARD({ delete ctx; }) }
int fheb200_ctx_set_stream(fheb200_ctx *ctx, void *s) {
    GUARD({
        need(ctx, "ctx");
        CtxLock g(*ctx->c);
        ctx->c->sync();
        if (ctx->c->own_stream && ctx->c->stream) cudaStreamDestroy(ctx->c->stream);
        ctx->c->stream = (cudaStream_t)s;
        ctx->c->own_stream = false;
    })
}
int fheb200_ctx_sync(fheb200_ctx *ctx) { GUARD({ need(ctx, "ctx"); ctx->c->sync(); }) }
int fheb200_ctx_fork(fheb200_ctx *ctx, int k) { GUARD({ need(ctx, "ctx"); ctx->c->fork(k); }) }
int fheb200_ctx_use_stream(fheb200_ctx *ctx, int i) { GUARD({ need(ctx, "ctx"); ctx->c->use_stream(i); }) }
int fheb200_ctx_bind_thread(fheb200_ctx *ctx, int i) { GUARD({ need(ctx, "ctx"); ctx->c->bind_thread(i); }) }
int fheb200_ctx_unbind_thread(fheb200_ctx *ctx) { GUARD({ need(ctx, "ctx"); ctx->c->unbind_thread(); }) }
int fheb200_ctx_join(fheb200_ctx *ctx) { GUARD({ need(ctx, "ctx"); ctx->c->join(); }) }
int fheb200_ctx_info(fheb200_ctx *ctx, int *n, int *nq, int *np, int *alpha, int *batch) {
    GUARD({
        need(ctx, "ctx");
        const Params &P = *ctx->c->P;
        if (n) *n = P.n;
        if (nq) *nq = P.L;
        if (np) *np = P.K;
        if (alpha) *alpha = P.alpha;
        if (batch) *batch = ctx->c->batch;
    })
}
int fheb200_ctx_modulus(fheb200_ctx *ctx, int i, uint64_t *q, uint64_t *psi) {
    GUARD({
        need(ctx, "ctx");
        const Params &P = *ctx->c->P;
        if (i < 0 || i >= P.L + P.K) throw std::out_of_range("modulus index");
        if (q) *q = P.moduli[i];
        if (psi) *psi = P.psi[i];
    })
}
int fheb200_ctx_scaling_factor(fheb200_ctx *ctx, int level, double *sf, double *sfb) {
    GUARD({
        need(ctx, "ctx");
        const Params &P = *ctx->c->P;
        if (level < 0 || level >= P.L) throw std::out_of_range("level");
        if (sf) *sf = P.sf[level];
        if (sfb) *sfb = level < P.L - 1 ? P.sf_big[level] : 0.0;
    })
}
int fheb200_ctx_key_bytes(fheb200_ctx *ctx, uint64_t *bytes) { GUARD({ need(ctx, "ctx"); *bytes = ctx->c->key_bytes(); }) }
int fheb200_params_chain(const fheb200_params *p, uint64_t *moduli, uint64_t *psi, int cap, int *nq, int *np, int *alpha) {
    GUARD({
        need(p, "params");
        auto P = Params::create(p->log_ring_dim, p->mult_depth, p->scaling_mod_size, p->first_mod_size ? p->first_mod_size : 60,
                                p->num_large_digits ? p->num_large_digits : 3);
        if (nq) *nq = P->L;
        if (np) *np = P->K;
        if (alpha) *alpha = P->alpha;
        for (int i = 0; i < P->L + P->K && i < cap; i++) {
            if (moduli) moduli[i] = P->moduli[i];
            if (psi) psi[i] = P->psi[i];
        }
    })
}

int fheb200_ctx_allow_deterministic_keys(fheb200_ctx *ctx, int allow) {
    GUARD({ need(ctx, "ctx"); ctx->c->allow_deterministic = allow != 0; })
}
int fheb200_ctx_get_rng_keys(fheb200_ctx *ctx, unsigned char *keys64) {
    GUARD({ need(ctx, "ctx"); need(keys64, "keys64"); ctx->c->get_rng_keys(keys64); })
}
int fheb200_ctx_set_rng_keys(fheb200_ctx *ctx, const unsigned char *keys64) {
    GUARD({ need(ctx, "ctx"); need(keys64, "keys64"); ctx->c->set_rng_keys(keys64); })
}
int fheb200_keygen(fheb200_ctx *ctx, uint64_t seed, fheb200_keys **out) {
    GUARD({ need(ctx, "ctx"); need(out, "out"); *out = new fheb200_keys{ctx->c->keygen(seed)}; })
}
int fheb200_keys_release(fheb200_keys *k) { GUARD({ delete k; }) }
int fheb200_eval_mult_keygen(fheb200_ctx *ctx, fheb200_keys *k) { GUARD({ need(ctx, "ctx"); need(k, "keys"); ctx->c->gen_mult_key(k->k); }) }
int fheb200_eval_rotate_keygen(fheb200_ctx *ctx, fheb200_keys *k, const int *idx, size_t count) {
    GUARD({ need(ctx, "ctx"); need(k, "keys"); ctx->c->gen_rot_keys(k->k, std::vector<int>(idx, idx + count)); })
}

int fheb200_make_packed_plaintext(fheb200_ctx *ctx, const double *v, size_t count, int nsd, int level, int slots, fheb200_pt **out) {
    GUARD({
        need(ctx, "ctx"); need(out, "out");
        *out = new fheb200_pt{ctx->c->make_plaintext_real(v, count, nsd, level, slots)};
    })
}
int fheb200_pt_rotate(fheb200_ctx *ctx, fheb200_pt *pt, int index, fheb200_pt **out) {
    GUARD({ need(ctx, "ctx"); need(pt, "pt"); need(out, "out"); *out = new fheb200_pt{ctx->c->rotate_plaintext(pt->p, index)}; })
}
int fheb200_pt_release(fheb200_pt *pt) { GUARD({ delete pt; }) }
int fheb200_pt_values(fheb200_pt *pt, double *out, size_t cap, size_t *count) {
    GUARD({
        need(pt, "pt");
        const Plaintext &P = pt->p->base ? *pt->p->base : *pt->p;
        const size_t have = P.real_only ? P.re.size() : P.vals.size();
        if (pt->p->base) { // rotated view: slot i = slot (i + rot) mod slots of the base (missing entries are zero)
            const long S = P.slots;
            if (count) *count = (size_t)S;
            for (long i = 0; i < S && (size_t)i < cap; i++) {
                const size_t j = (size_t)((((i + pt->p->rot) % S) + S) % S);
                out[i] = j < have ? (P.real_only ? P.re[j] : P.vals[j].real()) : 0.0;
            }
            return 0;
        }
        size_t nv = std::min(P.length, have);
        if (count) *count = nv;
        for (size_t i = 0; i < nv && i < cap; i++) out[i] = P.real_only ? P.re[i] : P.vals[i].real();
    })
}
int fheb200_pt_set_length(fheb200_pt *pt, size_t len) {
    GUARD({ need(pt, "pt"); pt->p->length = len;
            if (pt->p->real_only) { if (pt->p->re.size() < len) pt->p->re.resize(len); }
            else if (pt->p->vals.size() < len) pt->p->vals.resize(len); })
}
int fheb200_pt_log_precision(fheb200_pt *pt, double *bits) { GUARD({ need(pt, "pt"); *bits = pt->p->log_precision; }) }
int fheb200_encrypt(fheb200_ctx *ctx, fheb200_keys *k, fheb200_pt *pt, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(k, "keys"); need(pt, "pt"); *out = wrap(ctx->c->encrypt(k->k, pt->p)); })
}
int fheb200_decrypt(fheb200_ctx *ctx, fheb200_keys *k, fheb200_ct *ct, fheb200_pt **out) {
    GUARD({ need(ctx, "ctx"); need(k, "keys"); need(ct, "ct"); *out = new fheb200_pt{ctx->c->decrypt(k->k, ct->c)}; })
}

int fheb200_ct_release(fheb200_ct *ct) { GUARD({ delete ct; }) }
int fheb200_ct_clone(fheb200_ctx *ctx, fheb200_ct *ct, fheb200_ct **out) { GUARD({ need(ctx, "ctx"); need(ct, "ct"); *out = wrap(ctx->c->clone(ct->c)); }) }
int fheb200_ct_info(fheb200_ct *ct, int *level, int *nsd, int *slots, int *limbs, double *scale) {
    GUARD({
        need(ct, "ct");
        if (level) *level = ct->c->level;
        if (nsd) *nsd = ct->c->nsd;
        if (slots) *slots = ct->c->slots;
        if (limbs) *limbs = ct->c->ell;
        if (scale) *scale = ct->c->scale;
    })
}
int fheb200_ct_set_slots(fheb200_ct *ct, int slots) {
    GUARD({
        need(ct, "ct");
        if (slots <= 0 || (slots & (slots - 1))) throw std::invalid_argument("SetSlots: power of two expected");
        ct->c->slots = slots;
    })
}

#define BIN(name, expr)                                                                       \
    int name(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct *b, fheb200_ct **out) {              \
        GUARD({ need(ctx, "ctx"); need(a, "a"); need(b, "b"); need(out, "out"); *out = wrap(expr); }) \
    }
BIN(fheb200_eval_add, ctx->c->add(a->c, b->c))
BIN(fheb200_eval_sub, ctx->c->sub(a->c, b->c))
BIN(fheb200_eval_mult, ctx->c->mult(a->c, b->c))
BIN(fheb200_eval_mult_rescaled, ctx->c->relinearize_reduced(ctx->c->mult_lazy(a->c, b->c)))

int fheb200_eval_add_plain(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); need(p, "p"); *out = wrap(ctx->c->add_pt(a->c, p->p, false)); })
}
int fheb200_eval_sub_plain(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); need(p, "p"); *out = wrap(ctx->c->add_pt(a->c, p->p, true)); })
}
int fheb200_eval_add_const(fheb200_ctx *ctx, fheb200_ct *a, double c, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); *out = wrap(ctx->c->add_const(a->c, c)); })
}
int fheb200_eval_negate(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out) { GUARD({ need(ctx, "ctx"); need(a, "a"); *out = wrap(ctx->c->negate(a->c)); }) }
int fheb200_eval_add_many(fheb200_ctx *ctx, fheb200_ct *const *cts, size_t count, fheb200_ct **out) {
    GUARD({
        need(ctx, "ctx"); need(cts, "cts");
        std::vector<Ct> v;
        for (size_t i = 0; i < count; i++) { need(cts[i], "cts[i]"); v.push_back(cts[i]->c); }
        *out = wrap(ctx->c->add_many(v));
    })
}
int fheb200_eval_mult_plain(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); need(p, "p"); *out = wrap(ctx->c->mult_pt(a->c, p->p)); })
}
int fheb200_eval_mult_plain_add(fheb200_ctx *ctx, fheb200_ct *acc, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(acc, "acc"); need(a, "a"); need(p, "p"); *out = wrap(ctx->c->mult_pt_add(acc->c, a->c, p->p)); })
}
int fheb200_eval_mult_const(fheb200_ctx *ctx, fheb200_ct *a, double c, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); *out = wrap(ctx->c->mult_const(a->c, c)); })
}
int fheb200_eval_rescale(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out) { GUARD({ need(ctx, "ctx"); need(a, "a"); *out = wrap(ctx->c->rescale(a->c)); }) }

int fheb200_eval_rotate(fheb200_ctx *ctx, fheb200_ct *a, int index, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); *out = wrap(ctx->c->rotate(a->c, index)); })
}
int fheb200_fast_rotation_precompute(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pre **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); *out = new fheb200_pre{ctx->c->fast_rotation_precompute(a->c)}; })
}
int fheb200_pre_release(fheb200_pre *pre) { GUARD({ delete pre; }) }
int fheb200_eval_fast_rotation(fheb200_ctx *ctx, fheb200_ct *a, int index, fheb200_pre *pre, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "a"); need(pre, "pre"); *out = wrap(ctx->c->fast_rotation(a->c, index, pre->p)); })
}
int fheb200_has_rotation_key(fheb200_ctx *ctx, int index, int *present) {
    GUARD({ need(ctx, "ctx"); uint32_t g = ctx->c->P->galois_for_rotation(index); *present = (g == 1 || ctx->c->has_rot_key(g)) ? 1 : 0; })
}

int fheb200_eval_chebyshev_series(fheb200_ctx *ctx, fheb200_ct *x, const double *co, size_t count, double a, double b, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(x, "x"); need(co, "coeffs"); *out = wrap(ctx->c->eval_chebyshev(x->c, std::vector<double>(co, co + count), a, b)); })
}
int fheb200_eval_poly(fheb200_ctx *ctx, fheb200_ct *x, const double *co, size_t count, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(x, "x"); need(co, "coeffs"); *out = wrap(ctx->c->eval_poly_linear(x->c, std::vector<double>(co, co + count))); })
}
int fheb200_chebyshev_coefficients(double (*f)(double, void *), void *user, double a, double b, int degree, double *out) {
    GUARD({
        need((void *)f, "f"); need(out, "out");
        auto v = chebyshev_coefficients([&](double x) { return f(x, user); }, a, b, degree);
        memcpy(out, v.data(), v.size() * sizeof(double));
    })
}

// ---- bootstrapping
int fheb200_eval_bootstrap_setup(fheb200_ctx *ctx, int budget_enc, int budget_dec, int slots) {
    GUARD({ need(ctx, "ctx"); ctx->c->bootstrap_setup(budget_enc, budget_dec, slots); })
}
int fheb200_eval_bootstrap_keygen(fheb200_ctx *ctx, fheb200_keys *k, int slots) {
    GUARD({ need(ctx, "ctx"); need(k, "keys"); ctx->c->bootstrap_keygen(k->k, slots); })
}
int fheb200_eval_bootstrap(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "ct"); need(out, "out"); *out = wrap(ctx->c->bootstrap(a->c)); })
}
// Meta-BTS (Bae, Cheon, Cho, Kim, Kim 2022; what OpenFHE's EvalBootstrap(ct, numIterations, precision) runs, reference
// call site src/sort_algo.h:1110-1111): bootstrap, measure the error of that pass against the input at the input's level,
// scale the error up by 2^p, bootstrap it, scale it back and subtract.  One more pass per extra iteration; each removes
// ~p bits of the previous pass's error.  p is capped at 12: a single pass here is accurate to ~2^-16 for |m| <= 1 and the
// scaled error must stay inside the bootstrapping's input range.
static Ct bootstrap_iterated(Context &C, const Ct &a, int iterations, int precision) {
    if (iterations <= 1) return C.bootstrap(a);
    const int p = std::max(1, std::min(precision > 0 ? precision : 12, 12));
    Ct first = bootstrap_iterated(C, a, iterations - 1, precision);
    Ct err = C.sub(first, a);                       // aligned to the input's level: the error of the pass(es) so far
    for (int i = 0; i < p; i++) err = C.add(err, err); // times 2^p without spending a level
    Ct second = C.bootstrap(err);
    return C.sub(first, C.mult_const(second, std::ldexp(1.0, -p)));
}
int fheb200_eval_bootstrap_iterated(fheb200_ctx *ctx, fheb200_ct *a, int iterations, int precision, fheb200_ct **out) {
    GUARD({
        need(ctx, "ctx"); need(a, "ct"); need(out, "out");
        if (iterations < 1 || iterations > 4) throw std::invalid_argument("EvalBootstrap: numIterations must be 1..4");
        *out = wrap(bootstrap_iterated(*ctx->c, a->c, iterations, precision));
    })
}
int fheb200_eval_bootstrap_partial(fheb200_ctx *ctx, fheb200_ct *a, int stop_after, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "ct"); need(out, "out"); *out = wrap(ctx->c->bootstrap(a->c, stop_after)); })
}
int fheb200_bootstrap_depth(fheb200_ctx *ctx, int slots, int *levels) {
    GUARD({ need(ctx, "ctx"); need(levels, "levels"); *levels = ctx->c->bootstrap_depth(slots); })
}
int fheb200_eval_conjugate(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out) {
    GUARD({ need(ctx, "ctx"); need(a, "ct"); need(out, "out"); *out = wrap(ctx->c->conjugate(a->c)); })
}
int fheb200_bootstrap_plan_info(int ring_dim, int slots, int budget_enc, int budget_dec, double q0, double raise_scale,
                                int correction, int *depth, int *k_range, int *double_angles, int *rotations, int capacity,
                                int *num_rotations, double *cheb, int cheb_capacity, int *num_cheb, double *shift) {
    GUARD({
        BootPlan P = make_boot_plan(ring_dim, slots, budget_enc, budget_dec, q0, raise_scale, correction);
        if (depth) *depth = P.depth();
        if (k_range) *k_range = P.K;
        if (double_angles) *double_angles = P.r;
        auto R = P.rotations();
        if (num_rotations) *num_rotations = (int)R.size();
        if (rotations) for (int i = 0; i < (int)R.size() && i < capacity; i++) rotations[i] = R[i];
        if (shift) *shift = P.shift;
        if (num_cheb) *num_cheb = (int)P.cheb.size();
        if (cheb) for (int i = 0; i < (int)P.cheb.size() && i < cheb_capacity; i++) cheb[i] = P.cheb[i];
    })
}
int fheb200_bootstrap_plan_apply(int ring_dim, int slots, int budget_enc, int budget_dec, double q0, double raise_scale,
                                 int correction, int which, const double *in_re_im, double *out_re_im) {
    GUARD({
        need(in_re_im, "in"); need(out_re_im, "out");
        BootPlan P = make_boot_plan(ring_dim, slots, budget_enc, budget_dec, q0, raise_scale, correction);
        const int m = 2 * slots;
        std::vector<std::complex<double>> v(m), w(m);
        for (int j = 0; j < m; j++) v[j] = std::complex<double>(in_re_im[2 * j], in_re_im[2 * j + 1]);
        for (auto &st : which == 0 ? P.cts : P.stc) {
            std::fill(w.begin(), w.end(), std::complex<double>(0, 0));
            for (size_t t = 0; t < st.rot.size(); t++)
                for (int j = 0; j < m; j++) w[j] += st.diag[t][j] * v[(j + st.rot[t]) % m];
            v = w;
        }
        for (int j = 0; j < m; j++) out_re_im[2 * j] = v[j].real(), out_re_im[2 * j + 1] = v[j].imag();
    })
}

} // extern "C"
