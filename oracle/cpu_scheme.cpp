// oracle/cpu_scheme.cpp — a CPU implementation of the C ABI of include/fhe_b200.h (the subset the sort drivers call) on top
// of the CPU oracle's RNS primitives (ckks_oracle.h).
//
// TEST INFRASTRUCTURE ONLY.  Built into oracle/libfheb200_cpu.so and, together with the reference's unmodified drivers, into
// oracle/_ref/libref_sort_cpu.so, so that an encrypted sort runs end to end on the host cores with no GPU and no product code:
//   * an independent decrypted-value cross-check of the GPU sorts (tests/test_gpu_cpu_crosscheck.py);
//   * the CPU arm of bench.py's sort latency (BASELINE config 1: "DirectSort N = 16 ... runs on CPU").
// What it restates: the scheme layer that OpenFHE puts between the reference's calls and the RNS arithmetic - FLEXIBLEAUTOEXT
// level / scale bookkeeping (Kim-Papadimitriou-Polyakov 2020: lazy rescale before the next multiplication, exact real scaling
// factor per level, one extra level for fresh ciphertexts), operand alignment, EvalChebyshevSeries by Paterson-Stockmeyer
// (depth ceil(log2 k) + m, as OpenFHE), EvalPoly by binary powers.  One relinearisation per product, no fused kernels, no
// hoisting tricks: plain loops over the oracle's primitives.  No bootstrapping (the k-way network stays GPU-only).
// Parity status: as the oracle's - unpinned against OpenFHE (absent here); anchored on the reference's decrypted known answers.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include <omp.h>

#include "../include/fhe_b200.h"
#include "ckks_oracle.h"

typedef uint64_t u64;
typedef unsigned __int128 u128;

struct CtData {
    std::vector<u64> c0, c1; // [ell][n], evaluation domain
    std::vector<u64> view;   // c0 || c1 while a caller's all-reduce works on the words (fheb200_ct_device_view)
    int ell = 0, level = 0, nsd = 1, slots = 0;
    double scale = 1.0;
};
typedef std::shared_ptr<CtData> CtP;

struct fheb200_ctx {
    orc_ctx *o = nullptr;
    int logn = 0, n = 0, L = 0, K = 0, alpha = 0, depth = 0, batch = 0;
    std::vector<u64> q;           // QP order
    std::vector<double> sf, sfb;  // scaling factors per level
    u64 seed = 0, nonce = 1;
    std::vector<u64> evk_mult;
    std::map<uint32_t, std::vector<u64>> evk_rot;
    std::shared_ptr<struct KeyMaterial> keys;
};
struct KeyMaterial {
    std::vector<u64> s, pkb, pka;
};
struct fheb200_keys {
    std::shared_ptr<KeyMaterial> k;
};
struct fheb200_ct {
    CtP c;
};
struct fheb200_pt {
    std::vector<double> v;
    int slots = 0, level = 0, nsd = 1;
    size_t length = 0;
    double scale = 1.0;
    double log_precision = 0;
};
struct fheb200_pre {
    std::vector<u64> digits;
    int ell = 0;
};

static thread_local std::string g_err;
#define GUARD(...)                  \
    try {                           \
        __VA_ARGS__;                \
        return 0;                   \
    } catch (const std::exception &e) { \
        g_err = e.what();           \
        return 1;                   \
    }
static void need(const void *p, const char *what) {
    if (!p) throw std::invalid_argument(std::string("null argument: ") + what);
}

// ---------------------------------------------------------------- scalars and limb helpers
// round(value) as a signed integer of up to 126 bits, reduced modulo each of the first ell moduli
static std::vector<u64> scalars_for(const fheb200_ctx &C, long double value, int ell) {
    const bool neg = value < 0;
    const long double a = neg ? -value : value;
    u128 mag;
    if (a < 9.0e18L) {
        mag = (u128)(u64)llroundl(a);
    } else {
        int ex;
        const long double mant = frexpl(a, &ex);
        const u64 mi = (u64)ldexpl(mant, 64);
        if (ex > 126) throw std::overflow_error("cpu scheme: scaled constant exceeds 126 bits");
        mag = ex >= 64 ? ((u128)mi << (ex - 64)) : ((u128)mi >> (64 - ex));
    }
    std::vector<u64> s(ell);
    for (int i = 0; i < ell; i++) {
        u64 r = (u64)(mag % C.q[i]);
        s[i] = (neg && r) ? C.q[i] - r : r;
    }
    return s;
}
static CtP like(const CtP &a) {
    auto r = std::make_shared<CtData>();
    r->ell = a->ell, r->level = a->level, r->nsd = a->nsd, r->slots = a->slots, r->scale = a->scale;
    r->c0.resize(a->c0.size()), r->c1.resize(a->c1.size());
    return r;
}
// EvalMultCore by a real constant: times round(c * sf[level]); degree + 1
static CtP mult_core_const(const fheb200_ctx &C, const CtP &a, long double c) {
    const double sfl = C.sf[a->level];
    std::vector<u64> s = scalars_for(C, c * (long double)sfl, a->ell);
    CtP r = like(a);
    orc_mul_scalar(C.o, a->ell, a->c0.data(), s.data(), r->c0.data());
    orc_mul_scalar(C.o, a->ell, a->c1.data(), s.data(), r->c1.data());
    r->nsd = a->nsd + 1, r->scale = a->scale * sfl;
    return r;
}
static CtP mod_reduce(const fheb200_ctx &C, const CtP &a) { // rescale: drop the last limb
    if (a->ell < 2) throw std::runtime_error("rescale: no limbs left (multiplicative depth exhausted)");
    auto r = std::make_shared<CtData>();
    r->ell = a->ell - 1, r->level = a->level + 1, r->nsd = a->nsd - 1, r->slots = a->slots;
    r->scale = a->scale / (double)C.q[a->ell - 1];
    r->c0.resize((size_t)r->ell * C.n), r->c1.resize((size_t)r->ell * C.n);
    orc_rescale(C.o, a->ell, a->c0.data(), r->c0.data());
    orc_rescale(C.o, a->ell, a->c1.data(), r->c1.data());
    return r;
}
static CtP level_reduce(const fheb200_ctx &C, const CtP &a, int levels) {
    if (levels <= 0) return a;
    if (a->ell - levels < 1) throw std::runtime_error("LevelReduce: not enough limbs");
    auto r = std::make_shared<CtData>(*a);
    r->ell = a->ell - levels, r->level = a->level + levels;
    r->c0.resize((size_t)r->ell * C.n), r->c1.resize((size_t)r->ell * C.n);
    return r;
}
static CtP reduced(const fheb200_ctx &C, const CtP &a) { return a->nsd == 2 ? mod_reduce(C, a) : a; }

// Bring the shallower operand x to the level, degree and scale of t (OpenFHE's AdjustLevelsAndDepth under FLEXIBLEAUTOEXT):
// one multiplication by the ratio of the scaling factors, placed so that the following rescale lands exactly on t's scale.
static CtP raise_to(const fheb200_ctx &C, CtP x, const CtP &t) {
    const long double have = x->scale, step = C.sf[x->level], qlast = (long double)C.q[x->ell - 1];
    if (x->nsd == 2) {
        if (t->nsd == 2) {
            x = mod_reduce(C, mult_core_const(C, x, (long double)t->scale / have * qlast / step));
            x = level_reduce(C, x, t->level - x->level);
        } else if (x->level + 1 == t->level) {
            x = mod_reduce(C, x);
        } else {
            x = mod_reduce(C, mult_core_const(C, x, (long double)C.sfb[t->level - 1] / have * qlast / step));
            x = level_reduce(C, x, t->level - x->level - 1);
            x = mod_reduce(C, x);
        }
    } else {
        if (t->nsd == 2) {
            x = level_reduce(C, mult_core_const(C, x, (long double)t->scale / have / step), t->level - x->level);
        } else {
            x = mult_core_const(C, x, (long double)C.sfb[t->level - 1] / have / step);
            x = level_reduce(C, x, t->level - x->level - 1);
            x = mod_reduce(C, x);
        }
    }
    if (x->level != t->level || x->nsd != t->nsd) throw std::logic_error("cpu scheme: level alignment failed");
    x = std::make_shared<CtData>(*x);
    x->scale = t->scale;
    return x;
}
static void align(const fheb200_ctx &C, CtP &a, CtP &b) {
    if (a->level < b->level) a = raise_to(C, a, b);
    else if (b->level < a->level) b = raise_to(C, b, a);
    else if (a->nsd < b->nsd) a = mult_core_const(C, a, 1.0L);
    else if (b->nsd < a->nsd) b = mult_core_const(C, b, 1.0L);
}
static CtP add_sub(const fheb200_ctx &C, CtP a, CtP b, bool sub) {
    align(C, a, b);
    CtP r = like(a);
    r->slots = std::max(a->slots, b->slots);
    (sub ? orc_sub : orc_add)(C.o, a->ell, a->c0.data(), b->c0.data(), r->c0.data());
    (sub ? orc_sub : orc_add)(C.o, a->ell, a->c1.data(), b->c1.data(), r->c1.data());
    return r;
}
static CtP mult_ct(const fheb200_ctx &C, CtP a, CtP b) { // tensor product + relinearisation, no rescale
    if (C.evk_mult.empty()) throw std::runtime_error("EvalMult: relinearisation key missing (call EvalMultKeyGen)");
    align(C, a, b);
    if (a->nsd == 2) {
        const bool same = a.get() == b.get();
        a = mod_reduce(C, a);
        b = same ? a : mod_reduce(C, b);
    }
    CtP r = like(a);
    r->slots = std::max(a->slots, b->slots);
    orc_mult_relin(C.o, a->ell, a->c0.data(), a->c1.data(), b->c0.data(), b->c1.data(), C.evk_mult.data(), r->c0.data(), r->c1.data());
    r->nsd = a->nsd + b->nsd, r->scale = a->scale * b->scale;
    return r;
}
static CtP mult_const(const fheb200_ctx &C, const CtP &a, double c) { return mult_core_const(C, reduced(C, a), c); }
static CtP add_const(const fheb200_ctx &C, const CtP &a, double c) {
    std::vector<u64> s = scalars_for(C, (long double)c * (long double)a->scale, a->ell);
    CtP r = std::make_shared<CtData>(*a);
    orc_add_scalar(C.o, a->ell, a->c0.data(), s.data(), r->c0.data());
    return r;
}
static std::vector<u64> encode(const fheb200_ctx &C, const fheb200_pt &p, int ell, double scale) {
    std::vector<u64> out((size_t)ell * C.n);
    orc_encode(C.o, p.v.data(), (int)p.v.size(), p.slots, scale, ell, out.data());
    return out;
}
static CtP mult_pt(const fheb200_ctx &C, const CtP &a0, const fheb200_pt &p) {
    CtP a = reduced(C, a0);
    const double ps = C.sf[a->level];
    std::vector<u64> e = encode(C, p, a->ell, ps);
    CtP r = like(a);
    r->slots = std::max(a->slots, p.slots);
    orc_mul(C.o, a->ell, a->c0.data(), e.data(), r->c0.data());
    orc_mul(C.o, a->ell, a->c1.data(), e.data(), r->c1.data());
    r->nsd = a->nsd + 1, r->scale = a->scale * ps;
    return r;
}
static CtP add_pt(const fheb200_ctx &C, const CtP &a, const fheb200_pt &p, bool negate) {
    std::vector<u64> e = encode(C, p, a->ell, a->scale);
    CtP r = std::make_shared<CtData>(*a);
    r->slots = std::max(a->slots, p.slots);
    (negate ? orc_sub : orc_add)(C.o, a->ell, a->c0.data(), e.data(), r->c0.data());
    return r;
}
static uint32_t galois_of(const fheb200_ctx &C, int index) { return orc_rot_to_galois(C.o, index); }
static CtP rotate(const fheb200_ctx &C, const CtP &a, int index) {
    const uint32_t g = galois_of(C, index);
    if (g == 1) return std::make_shared<CtData>(*a);
    auto it = C.evk_rot.find(g);
    if (it == C.evk_rot.end()) throw std::runtime_error("EvalRotate: no evaluation key for rotation index " + std::to_string(index));
    CtP r = like(a);
    orc_rotate(C.o, a->ell, g, a->c0.data(), a->c1.data(), it->second.data(), r->c0.data(), r->c1.data());
    return r;
}

// ---------------------------------------------------------------- Chebyshev series (Paterson-Stockmeyer) and power-basis polynomials
typedef std::vector<double> Poly; // p = sum p[i] T_i
static int degree_of(const Poly &p) {
    int d = (int)p.size() - 1;
    while (d > 0 && p[d] == 0.0) d--;
    return d;
}
// f = quo * g + rem in the Chebyshev basis (T_a T_b = (T_{a+b} + T_|a-b|) / 2), deg rem < deg g; dense long division
static void cheb_divmod(const Poly &f, const Poly &g, Poly &quo, Poly &rem) {
    const int dg = degree_of(g);
    int df = degree_of(f);
    rem = f;
    rem.resize(std::max(df, dg) + 1, 0.0);
    quo.assign(df >= dg ? df - dg + 1 : 1, 0.0);
    for (int top = df; top >= dg && dg > 0; top--) {
        if (rem[top] == 0.0) continue;
        const int sh = top - dg;
        const double c = sh > 0 ? 2.0 * rem[top] / g[dg] : rem[top] / g[dg]; // T_sh T_dg contributes T_top with weight 1/2
        quo[sh] = c;
        for (int i = 0; i <= dg; i++) {
            if (g[i] == 0.0) continue;
            if (sh == 0) rem[i] -= c * g[i];
            else rem[i + sh] -= 0.5 * c * g[i], rem[std::abs(i - sh)] -= 0.5 * c * g[i];
        }
        rem[top] = 0.0;
    }
    rem.resize(std::max(dg, 1));
}
struct PsRun {
    const fheb200_ctx &C;
    int k = 0, m = 0;
    std::vector<CtP> baby, giant; // baby[i] = T_{i+1}(u) aligned to T_k ; giant[i] = T_{k 2^i}(u)
    std::vector<CtP> baby1;       // baby steps at degree 1 (operands of the scalar products)
    explicit PsRun(const fheb200_ctx &c) : C(c) {}
    const CtP &deg1(int i) {
        if (baby1.empty()) baby1.resize(baby.size());
        if (!baby1[i]) baby1[i] = reduced(C, baby[i]);
        return baby1[i];
    }
};
static CtP twice(const fheb200_ctx &C, const CtP &a) { return add_sub(C, a, a, false); }
// sum_{1 <= i < k} p_i T_i (null when every weight vanishes)
static CtP scalar_sum(PsRun &R, const Poly &p) {
    CtP acc;
    for (int i = 1; i < R.k && i < (int)p.size(); i++) {
        if (p[i] == 0.0) continue;
        CtP t = mult_core_const(R.C, R.deg1(i - 1), p[i]);
        acc = acc ? add_sub(R.C, acc, t, false) : t;
    }
    return acc;
}
// degree-k(2^m - 1) polynomial whose leading coefficient is a power of two
static CtP ps_node(PsRun &R, const Poly &f, int m) {
    const fheb200_ctx &C = R.C;
    const int k = R.k;
    if (m == 1) {
        CtP r = scalar_sum(R, f);
        const double lead = (int)f.size() > k ? f[k] : 0.0;
        if (lead != 0.0) {
            const long j = std::lround(std::log2(std::fabs(lead)));
            CtP top;
            if (lead > 0 && j >= 0 && std::fabs(std::ldexp(1.0, (int)j) - lead) < 1e-9 * lead) {
                top = R.baby[k - 1];
                for (long x = 0; x < j; x++) top = twice(C, top);
            } else {
                top = mult_const(C, R.baby[k - 1], lead);
            }
            r = r ? add_sub(C, r, top, false) : top;
        }
        if (!r) throw std::logic_error("cpu scheme: empty Paterson-Stockmeyer leaf");
        return f[0] != 0.0 ? add_const(C, r, f[0]) : r;
    }
    // f = (T_{k 2^(m-1)} + c) q + s with deg q = deg s = k(2^(m-1) - 1); s is made monic so that the recursion stays in shape
    const int half = k * (1 << (m - 1)), low = half - k;
    Poly Tg(half + 1, 0.0), q, r, c, s;
    Tg[half] = 1.0;
    cheb_divmod(f, Tg, q, r);
    r.resize(std::max((int)r.size(), low + 1), 0.0);
    r[low] -= 1.0;
    cheb_divmod(r, q, c, s);
    s.resize(low + 1, 0.0);
    s[low] += 1.0;
    CtP left = R.giant[m - 1];
    if (CtP cu = scalar_sum(R, c)) left = add_sub(C, left, cu, false);
    if (c[0] != 0.0) left = add_const(C, left, c[0]);
    CtP prod = mult_ct(C, left, ps_node(R, q, m - 1));
    return add_sub(C, prod, ps_node(R, s, m - 1), false);
}
static CtP to_unit(const fheb200_ctx &C, const CtP &x, double a, double b) {
    if (std::fabs(a + 1.0) < 1e-12 && std::fabs(b - 1.0) < 1e-12) return x;
    return add_const(C, mult_const(C, x, 2.0 / (b - a)), -(a + b) / (b - a));
}
static std::vector<CtP> cheb_basis(const fheb200_ctx &C, const CtP &u, int count) { // T_2i = 2 T_i^2 - 1, T_2i+1 = 2 T_i T_i+1 - T_1
    std::vector<CtP> T(count);
    T[0] = u;
    for (int i = 2; i <= count; i++) {
        if (i & 1) T[i - 1] = add_sub(C, twice(C, mult_ct(C, T[i / 2 - 1], T[i / 2])), T[0], true);
        else T[i - 1] = add_const(C, twice(C, mult_ct(C, T[i / 2 - 1], T[i / 2 - 1])), -1.0);
    }
    return T;
}
static CtP eval_chebyshev(const fheb200_ctx &C, const CtP &x, const std::vector<double> &coeffs, double a, double b) {
    const int n = degree_of(coeffs);
    if (n < 1) throw std::invalid_argument("EvalChebyshevSeries: degree must be >= 1");
    CtP u = to_unit(C, x, a, b);
    if (n < 5) { // linear evaluation, as OpenFHE below degree 5
        std::vector<CtP> T = cheb_basis(C, u, n);
        CtP acc;
        for (int i = 1; i <= n; i++) {
            if (coeffs[i] == 0.0) continue;
            CtP t = mult_const(C, T[i - 1], coeffs[i]);
            acc = acc ? add_sub(C, acc, t, false) : t;
        }
        if (!acc) acc = mult_const(C, T[0], 0.0);
        return add_const(C, acc, coeffs[0] / 2.0);
    }
    PsRun R(C);
    int best_depth = 1 << 30;
    long best_cost = 1L << 60;
    for (int k = 2; k <= n + 1; k++) { // smallest depth ceil(log2 k) + m with k (2^m - 1) > n, then fewest products
        int lg = 0, m = 1;
        while ((1 << lg) < k) lg++;
        while ((long)k * ((1L << m) - 1) <= n) m++;
        const long cost = k + 2 * m + (1L << (m - 1)) - 4;
        if (lg + m < best_depth || (lg + m == best_depth && cost < best_cost)) best_depth = lg + m, best_cost = cost, R.k = k, R.m = m;
    }
    const int k = R.k, m = R.m;
    R.baby = cheb_basis(C, u, k);
    for (int i = 0; i + 1 < k; i++) {
        CtP t = R.baby[i], d = R.baby[k - 1];
        align(C, t, d);
        R.baby[i] = t;
    }
    R.giant.resize(m);
    R.giant[0] = R.baby[k - 1];
    for (int i = 1; i < m; i++) R.giant[i] = add_const(C, twice(C, mult_ct(C, R.giant[i - 1], R.giant[i - 1])), -1.0);
    CtP Ttop = R.giant[0]; // T_{k(2^m - 1)} by T_{a+b} = 2 T_a T_b - T_{a-b}, a = k 2^i, b = k(2^i - 1)
    for (int i = 1; i < m; i++) Ttop = add_sub(C, twice(C, mult_ct(C, Ttop, R.giant[i])), R.giant[0], true);
    Poly f(coeffs.begin(), coeffs.begin() + n + 1);
    f[0] /= 2.0; // series convention: c_0 / 2 + sum c_i T_i
    const int top = k * ((1 << m) - 1);
    f.resize(top + 1, 0.0);
    f[top] += 1.0;
    return add_sub(C, ps_node(R, f, m), Ttop, true);
}
// EvalPolyLinear: power basis, binary powers
static CtP eval_poly(const fheb200_ctx &C, const CtP &x, const std::vector<double> &coeffs) {
    int n = (int)coeffs.size() - 1;
    while (n > 0 && coeffs[n] == 0.0) n--;
    if (n < 1) throw std::invalid_argument("EvalPoly: degree must be >= 1");
    std::vector<CtP> pw(n + 1);
    pw[1] = x;
    for (int i = 2; i <= n; i++) {
        int hi = 1;
        while (hi * 2 <= i) hi *= 2;
        pw[i] = (i == hi) ? mult_ct(C, pw[i / 2], pw[i / 2]) : mult_ct(C, pw[hi], pw[i - hi]);
    }
    CtP acc;
    for (int i = 1; i <= n; i++) {
        if (coeffs[i] == 0.0) continue;
        CtP t = mult_const(C, pw[i], coeffs[i]);
        acc = acc ? add_sub(C, acc, t, false) : t;
    }
    if (!acc) acc = mult_const(C, pw[1], 0.0);
    return add_const(C, acc, coeffs[0]);
}

// ---------------------------------------------------------------- the C ABI (same names and meaning as include/fhe_b200.h)
static fheb200_ct *wrap(const CtP &c) { return new fheb200_ct{c}; }
static fheb200_ctx &ctx_of(fheb200_ctx *c) {
    need(c, "ctx");
    return *c;
}

extern "C" {
const char *fheb200_last_error(void) { return g_err.c_str(); }
const char *fheb200_version(void) { return "fheb200-cpu-oracle (test infrastructure)"; }
unsigned long long fheb200_launch_count(void) { return 0; }
unsigned long long fheb200_algo_bytes(void) { return 0; }
unsigned long long fheb200_op_count(int) { return 0; }
int fheb200_cpu_threads(void) { const char *th = std::getenv("FHEB200_CPU_THREADS"); return th && std::atoi(th) > 0 ? std::atoi(th) : omp_get_max_threads(); }

int fheb200_ctx_create(const fheb200_params *p, fheb200_ctx **out) {
    GUARD({
        need(p, "params"); need(out, "out");
        auto C = std::make_unique<fheb200_ctx>();
        const char *th = std::getenv("FHEB200_CPU_THREADS"); // limb loops of the oracle: all host cores unless told otherwise
        orc_set_threads(th && std::atoi(th) > 0 ? std::atoi(th) : omp_get_max_threads());
        C->o = orc_ctx_create(p->log_ring_dim, p->mult_depth, p->scaling_mod_size, p->first_mod_size, p->num_large_digits > 0 ? p->num_large_digits : 3);
        if (!C->o) throw std::invalid_argument("cpu scheme: bad parameters");
        C->logn = p->log_ring_dim, C->n = orc_n(C->o), C->L = orc_L(C->o), C->K = orc_K(C->o), C->alpha = orc_alpha(C->o);
        C->depth = p->mult_depth, C->batch = p->batch_size > 0 ? p->batch_size : C->n / 2;
        for (int i = 0; i < C->L + C->K; i++) C->q.push_back(orc_modulus(C->o, i));
        for (int l = 0; l < C->L; l++) C->sf.push_back(orc_scale_real(C->o, l)), C->sfb.push_back(orc_scale_big(C->o, l));
        *out = C.release();
    })
}
int fheb200_ctx_release(fheb200_ctx *c) {
    if (c) {
        orc_ctx_free(c->o);
        delete c;
    }
    return 0;
}
int fheb200_ctx_set_stream(fheb200_ctx *, void *) { return 0; }
int fheb200_ctx_sync(fheb200_ctx *) { return 0; }
int fheb200_ctx_fork(fheb200_ctx *, int) { return 0; }
int fheb200_ctx_use_stream(fheb200_ctx *, int) { return 0; }
int fheb200_ctx_bind_thread(fheb200_ctx *, int) { return 0; }
int fheb200_ctx_unbind_thread(fheb200_ctx *) { return 0; }
int fheb200_ctx_join(fheb200_ctx *) { return 0; }
int fheb200_ctx_info(fheb200_ctx *c, int *n, int *nq, int *np, int *digit, int *batch) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        if (n) *n = C.n;
        if (nq) *nq = C.L;
        if (np) *np = C.K;
        if (digit) *digit = C.alpha;
        if (batch) *batch = C.batch;
    })
}
int fheb200_ctx_params(fheb200_ctx *c, fheb200_params *p) {
    GUARD({ (void)ctx_of(c); need(p, "params"); throw std::runtime_error("cpu scheme: serialisation is not provided"); })
}
int fheb200_ctx_allow_deterministic_keys(fheb200_ctx *, int) { return 0; }
int fheb200_keygen(fheb200_ctx *c, uint64_t seed, fheb200_keys **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(out, "out");
        if (seed == 0) throw std::invalid_argument("cpu scheme: KeyGen needs a non-zero (test) seed - the oracle has no entropy source");
        C.seed = seed, C.nonce = 1;
        auto K = std::make_shared<KeyMaterial>();
        const size_t n = C.n;
        K->s.resize((size_t)(C.L + C.K) * n), K->pkb.resize((size_t)C.L * n), K->pka.resize((size_t)C.L * n);
        orc_keygen_secret(C.o, seed, K->s.data());
        orc_keygen_public(C.o, seed, K->s.data(), K->pkb.data(), K->pka.data());
        C.keys = K;
        *out = new fheb200_keys{K};
    })
}
int fheb200_keys_release(fheb200_keys *k) {
    delete k;
    return 0;
}
int fheb200_eval_mult_keygen(fheb200_ctx *c, fheb200_keys *k) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(k, "keys");
        C.evk_mult.resize(orc_evk_words(C.o));
        orc_keygen_mult(C.o, C.seed, k->k->s.data(), C.evk_mult.data());
    })
}
int fheb200_eval_rotate_keygen(fheb200_ctx *c, fheb200_keys *k, const int *idx, size_t count) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(k, "keys");
        for (size_t i = 0; i < count; i++) {
            const uint32_t g = galois_of(C, idx[i]);
            if (g == 1 || C.evk_rot.count(g)) continue;
            std::vector<u64> &e = C.evk_rot[g];
            e.resize(orc_evk_words(C.o));
            orc_keygen_rot(C.o, C.seed, g, k->k->s.data(), e.data());
        }
    })
}
int fheb200_has_rotation_key(fheb200_ctx *c, int index, int *present) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(present, "present"); const uint32_t g = galois_of(C, index); *present = g == 1 || C.evk_rot.count(g); })
}
int fheb200_make_packed_plaintext(fheb200_ctx *c, const double *values, size_t count, int nsd, int level, int slots, fheb200_pt **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(out, "out");
        if (slots <= 0) slots = C.batch;
        if ((slots & (slots - 1)) || slots > C.n / 2) throw std::invalid_argument("MakeCKKSPackedPlaintext: slots must be a power of two <= n/2");
        if ((int)count > slots) throw std::invalid_argument("MakeCKKSPackedPlaintext: more values than slots");
        if (level < 0 || level >= C.L) throw std::invalid_argument("MakeCKKSPackedPlaintext: bad level");
        auto p = std::make_unique<fheb200_pt>();
        p->v.assign(values, values + count);
        p->slots = slots, p->length = count, p->level = level;
        if (level == 0) p->nsd = 2, p->scale = C.sfb[0]; // FLEXIBLEAUTOEXT: fresh data lives at the big scale, degree 2
        else p->nsd = nsd < 1 ? 1 : nsd, p->scale = std::pow(C.sf[level], (double)p->nsd);
        *out = p.release();
    })
}
int fheb200_pt_rotate(fheb200_ctx *c, fheb200_pt *pt, int index, fheb200_pt **out) {
    GUARD({
        (void)ctx_of(c);
        need(pt, "pt"); need(out, "out");
        auto p = std::make_unique<fheb200_pt>(*pt);
        p->v.assign((size_t)pt->slots, 0.0);
        for (int i = 0; i < pt->slots; i++) {
            const int src = ((i + index) % pt->slots + pt->slots) % pt->slots;
            p->v[i] = src < (int)pt->v.size() ? pt->v[src] : 0.0;
        }
        p->length = pt->slots;
        *out = p.release();
    })
}
int fheb200_pt_release(fheb200_pt *p) {
    delete p;
    return 0;
}
int fheb200_pt_values(fheb200_pt *p, double *out, size_t capacity, size_t *count) {
    GUARD({
        need(p, "pt");
        const size_t len = p->length ? p->length : p->v.size();
        if (count) *count = len;
        for (size_t i = 0; out && i < len && i < capacity; i++) out[i] = i < p->v.size() ? p->v[i] : 0.0;
    })
}
int fheb200_pt_set_length(fheb200_pt *p, size_t len) { GUARD({ need(p, "pt"); p->length = len; }) }
int fheb200_pt_log_precision(fheb200_pt *p, double *bits) { GUARD({ need(p, "pt"); need(bits, "bits"); *bits = p->log_precision; }) }

int fheb200_encrypt(fheb200_ctx *c, fheb200_keys *k, fheb200_pt *pt, fheb200_ct **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(k, "keys"); need(pt, "pt"); need(out, "out");
        const int ell = C.L - pt->level;
        std::vector<u64> e = encode(C, *pt, ell, pt->scale);
        auto r = std::make_shared<CtData>();
        r->ell = ell, r->level = pt->level, r->nsd = pt->nsd, r->scale = pt->scale, r->slots = pt->slots;
        r->c0.resize((size_t)ell * C.n), r->c1.resize((size_t)ell * C.n);
        orc_encrypt(C.o, C.seed, C.nonce++, ell, k->k->pkb.data(), k->k->pka.data(), e.data(), r->c0.data(), r->c1.data());
        *out = wrap(r);
    })
}
int fheb200_decrypt(fheb200_ctx *c, fheb200_keys *k, fheb200_ct *ct, fheb200_pt **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(k, "keys"); need(ct, "ct"); need(out, "out");
        const CtData &a = *ct->c;
        auto p = std::make_unique<fheb200_pt>();
        p->slots = a.slots, p->level = a.level, p->nsd = a.nsd, p->scale = a.scale, p->length = a.slots;
        p->v.resize(a.slots);
        std::vector<double> im(a.slots);
        orc_decrypt(C.o, a.ell, a.c0.data(), a.c1.data(), k->k->s.data(), a.slots, a.scale, p->v.data(), im.data());
        double worst = 0;
        for (double v : im) worst = std::max(worst, std::fabs(v));
        p->log_precision = worst > 0 ? -std::log2(worst) : 60.0; // the imaginary parts estimate the noise, as OpenFHE's decoder
        *out = p.release();
    })
}
int fheb200_ct_release(fheb200_ct *ct) {
    delete ct;
    return 0;
}
int fheb200_ct_clone(fheb200_ctx *, fheb200_ct *ct, fheb200_ct **out) { GUARD({ need(ct, "ct"); need(out, "out"); *out = wrap(std::make_shared<CtData>(*ct->c)); }) }
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
int fheb200_ct_set_slots(fheb200_ct *ct, int slots) { GUARD({ need(ct, "ct"); ct->c->slots = slots; }) }

#define BIN(name, expr)                                                                          \
    int name(fheb200_ctx *c, fheb200_ct *a, fheb200_ct *b, fheb200_ct **out) {                   \
        GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); need(b, "b"); need(out, "out"); *out = wrap(expr); }) \
    }
BIN(fheb200_eval_add, add_sub(C, a->c, b->c, false))
BIN(fheb200_eval_sub, add_sub(C, a->c, b->c, true))
BIN(fheb200_eval_mult, mult_ct(C, a->c, b->c))
BIN(fheb200_eval_mult_rescaled, mod_reduce(C, mult_ct(C, a->c, b->c))) // the two-step form the merged division is compared with
int fheb200_eval_add_plain(fheb200_ctx *c, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); need(p, "p"); *out = wrap(add_pt(C, a->c, *p, false)); })
}
int fheb200_eval_sub_plain(fheb200_ctx *c, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); need(p, "p"); *out = wrap(add_pt(C, a->c, *p, true)); })
}
int fheb200_eval_add_const(fheb200_ctx *c, fheb200_ct *a, double v, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); *out = wrap(add_const(C, a->c, v)); })
}
int fheb200_eval_negate(fheb200_ctx *c, fheb200_ct *a, fheb200_ct **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(a, "a");
        CtP z = like(a->c);
        std::fill(z->c0.begin(), z->c0.end(), 0), std::fill(z->c1.begin(), z->c1.end(), 0);
        *out = wrap(add_sub(C, z, a->c, true));
    })
}
int fheb200_eval_add_many(fheb200_ctx *c, fheb200_ct *const *cts, size_t count, fheb200_ct **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        if (!count) throw std::invalid_argument("EvalAddMany: empty input");
        CtP acc = cts[0]->c;
        for (size_t i = 1; i < count; i++) acc = add_sub(C, acc, cts[i]->c, false);
        *out = wrap(count == 1 ? std::make_shared<CtData>(*acc) : acc);
    })
}
int fheb200_eval_mult_plain(fheb200_ctx *c, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); need(p, "p"); *out = wrap(mult_pt(C, a->c, *p)); })
}
int fheb200_eval_mult_plain_add(fheb200_ctx *c, fheb200_ct *acc, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(a, "a"); need(p, "p");
        CtP prod = mult_pt(C, a->c, *p);
        *out = wrap(acc ? add_sub(C, acc->c, prod, false) : prod);
    })
}
int fheb200_eval_mult_const(fheb200_ctx *c, fheb200_ct *a, double v, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); *out = wrap(mult_const(C, a->c, v)); })
}
int fheb200_eval_rescale(fheb200_ctx *c, fheb200_ct *a, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); *out = wrap(mod_reduce(C, a->c)); })
}
int fheb200_eval_rotate(fheb200_ctx *c, fheb200_ct *a, int index, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(a, "a"); *out = wrap(rotate(C, a->c, index)); })
}
int fheb200_fast_rotation_precompute(fheb200_ctx *c, fheb200_ct *a, fheb200_pre **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(a, "a"); need(out, "out");
        auto p = std::make_unique<fheb200_pre>();
        p->ell = a->c->ell;
        p->digits.resize((size_t)orc_beta(C.o, p->ell) * (p->ell + C.K) * C.n);
        orc_modup(C.o, p->ell, a->c->c1.data(), p->digits.data());
        *out = p.release();
    })
}
int fheb200_pre_release(fheb200_pre *p) {
    delete p;
    return 0;
}
int fheb200_eval_fast_rotation(fheb200_ctx *c, fheb200_ct *a, int index, fheb200_pre *pre, fheb200_ct **out) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(a, "a"); need(pre, "pre");
        const uint32_t g = galois_of(C, index);
        if (g == 1) {
            *out = wrap(std::make_shared<CtData>(*a->c));
        } else {
            auto it = C.evk_rot.find(g);
            if (it == C.evk_rot.end()) throw std::runtime_error("EvalFastRotation: no evaluation key for rotation index " + std::to_string(index));
            if (pre->ell != a->c->ell) throw std::runtime_error("EvalFastRotation: precomputation does not match the ciphertext");
            CtP r = like(a->c);
            orc_rotate_hoisted(C.o, a->c->ell, g, a->c->c0.data(), pre->digits.data(), it->second.data(), r->c0.data(), r->c1.data());
            *out = wrap(r);
        }
    })
}
int fheb200_eval_chebyshev_series(fheb200_ctx *c, fheb200_ct *x, const double *coeffs, size_t count, double a, double b, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(x, "x"); need(coeffs, "coeffs"); *out = wrap(eval_chebyshev(C, x->c, std::vector<double>(coeffs, coeffs + count), a, b)); })
}
int fheb200_eval_poly(fheb200_ctx *c, fheb200_ct *x, const double *coeffs, size_t count, fheb200_ct **out) {
    GUARD({ fheb200_ctx &C = ctx_of(c); need(x, "x"); need(coeffs, "coeffs"); *out = wrap(eval_poly(C, x->c, std::vector<double>(coeffs, coeffs + count))); })
}
int fheb200_chebyshev_coefficients(double (*f)(double, void *), void *user, double a, double b, int degree, double *out) {
    GUARD({ need((const void *)f, "f"); need(out, "out"); orc_cheb_coefficients(f, user, a, b, degree, out); })
}
// not provided on the CPU: bootstrapping, serialisation, collectives (the sorts that need them stay GPU-only)
#define ABSENT(sig, what) \
    int sig { g_err = std::string("cpu scheme: ") + what + " is not provided"; return 1; }
ABSENT(fheb200_eval_bootstrap_setup(fheb200_ctx *, int, int, int), "EvalBootstrapSetup")
ABSENT(fheb200_eval_bootstrap_keygen(fheb200_ctx *, fheb200_keys *, int), "EvalBootstrapKeyGen")
ABSENT(fheb200_eval_bootstrap(fheb200_ctx *, fheb200_ct *, fheb200_ct **), "EvalBootstrap")
ABSENT(fheb200_eval_bootstrap_iterated(fheb200_ctx *, fheb200_ct *, int, int, fheb200_ct **), "EvalBootstrap")
// The callback form of the multi-rank combine (include/fhe_b200.h): the caller sums the words of the partial ciphertexts of all
// ranks as 64-bit integers (gloo all-reduce in tests/test_multi_rank.py), then every word is reduced modulo its limb's modulus
int fheb200_ct_device_view(fheb200_ct *ct, void **ptr, size_t *words) {
    GUARD({
        need(ct, "ct"); need(ptr, "ptr"); need(words, "words");
        CtData &a = *ct->c;
        a.view = a.c0;
        a.view.insert(a.view.end(), a.c1.begin(), a.c1.end());
        *ptr = a.view.data(), *words = a.view.size();
    })
}
int fheb200_ct_reduce_partial_sums(fheb200_ctx *c, fheb200_ct *ct) {
    GUARD({
        fheb200_ctx &C = ctx_of(c);
        need(ct, "ct");
        CtData &a = *ct->c;
        const size_t n = C.n, half = (size_t)a.ell * n;
        if (a.view.size() != 2 * half) throw std::runtime_error("reduce_partial_sums: no view taken of this ciphertext");
        for (int i = 0; i < a.ell; i++)
            for (size_t k = 0; k < n; k++) {
                a.c0[(size_t)i * n + k] = a.view[(size_t)i * n + k] % C.q[i];
                a.c1[(size_t)i * n + k] = a.view[half + (size_t)i * n + k] % C.q[i];
            }
        a.view.clear(), a.view.shrink_to_fit();
    })
}
ABSENT(fheb200_ct_all_reduce_sum(fheb200_ctx *, fheb200_ct *), "all-reduce")
ABSENT(fheb200_comm_unique_id(unsigned char *), "NCCL")
ABSENT(fheb200_ctx_comm_init(fheb200_ctx *, const unsigned char *, int, int), "NCCL")
ABSENT(fheb200_keys_export(fheb200_ctx *, fheb200_keys *, int, uint64_t *, size_t), "key export")
ABSENT(fheb200_keys_import_public(fheb200_ctx *, const uint64_t *, const uint64_t *, size_t, fheb200_keys **), "key import")
ABSENT(fheb200_evk_export(fheb200_ctx *, uint32_t, uint64_t *, size_t), "key export")
ABSENT(fheb200_evk_import(fheb200_ctx *, uint32_t, const uint64_t *, size_t), "key import")
ABSENT(fheb200_evk_list(fheb200_ctx *, uint32_t *, size_t, size_t *), "key list")
ABSENT(fheb200_has_mult_key(fheb200_ctx *, int *), "key query")
ABSENT(fheb200_ctx_evk_words(fheb200_ctx *, size_t *), "key size")
ABSENT(fheb200_ct_export(fheb200_ctx *, fheb200_ct *, uint64_t *, uint64_t *, size_t), "ciphertext export")
ABSENT(fheb200_ct_import(fheb200_ctx *, const uint64_t *, const uint64_t *, int, int, double, int, fheb200_ct **), "ciphertext import")
}
