// poly_eval.cu — polynomial evaluation on ciphertexts: Chebyshev series by Paterson-Stockmeyer
// (EvalChebyshevSeriesPS: src/sort_algo.h:629,728,967; src/sign.cpp:76), the linear variant
// (EvalChebyshevSeries for degree < 5), EvalChebyshevFunction / EvalChebyshevCoefficients
// (src/sign.cpp:616,630; utils/generate_cheb_coeffs.cpp:17) and EvalPolyLinear (src/mehp24/mehp24_utils.cpp:256-259).
//
// PS layout (Chen-Chillotti-Song style, as OpenFHE): baby steps T_1..T_k, giants T_{k 2^i}; the series is made
// monic of degree k(2^m - 1) so that every leaf carries a power-of-two leading coefficient and no scalar
// multiplication sits on the critical path.  Levels consumed: ceil(log2 k) + m.
#include <cmath>
#include <stdexcept>

#include "engine.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>

namespace fheb200 {

typedef std::vector<double> Poly; // true Chebyshev coefficients: p = sum p[i] T_i

namespace {
struct HostTimer { // FHEB200_PSTIME=1: where the launching thread spends its time inside one Paterson-Stockmeyer evaluation
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double lap() {
        auto t = std::chrono::steady_clock::now();
        double ms = std::chrono::duration<double, std::milli>(t - t0).count();
        t0 = t;
        return ms;
    }
};
double g_split_ms = 0;
} // namespace


static int degree_of(const Poly &p) {
    int d = (int)p.size() - 1;
    while (d > 0 && p[d] == 0.0) d--;
    return d;
}
// r += w * T_j * g  in the Chebyshev basis: T_j T_i = (T_{i+j} + T_{|i-j|}) / 2 ; nz = indices of g's non-zero entries
static void add_shifted(Poly &r, double w, int j, const Poly &g, const std::vector<int> &nz) {
    const int dg = nz.empty() ? 0 : nz.back();
    if ((int)r.size() < dg + j + 1) r.resize(dg + j + 1, 0.0);
    for (int i : nz) {
        if (j == 0) {
            r[i] += w * g[i];
        } else {
            r[i + j] += 0.5 * w * g[i];
            r[std::abs(i - j)] += 0.5 * w * g[i];
        }
    }
}
// f = q g + r, deg r < deg g.  The divisors of the Paterson-Stockmeyer recursion are either a single T_j or met once per
// node, so the work is linear in the number of non-zero terms (a dense scan of g per quotient term made the host side
// of a degree-6510 evaluation cost more than its 5700 kernel launches).
static void cheb_divide(const Poly &f, const Poly &g, Poly &q, Poly &r) {
    int n = degree_of(f), k = degree_of(g);
    std::vector<int> nz;
    for (int i = 0; i <= k; i++)
        if (g[i] != 0.0) nz.push_back(i);
    r = f;
    r.resize(n + 1);
    q.assign(n >= k ? n - k + 1 : 1, 0.0);
    while (n >= k) {
        if (r[n] == 0.0) {
            n--;
            if (n < 0) break;
            continue;
        }
        double lead = (n > k) ? 2.0 * r[n] / g[k] : r[n] / g[k];
        q[n - k] = lead;
        add_shifted(r, -lead, n - k, g, nz);
        r[n] = 0.0; // exact cancellation of the leading term
        n--;
        if (n < 0) break;
    }
    r.resize(std::max(k, 1));
    if (k == 0) r.assign(1, 0.0);
}

struct PSState {
    Context *C;
    int k, m;
    std::vector<Ct> T;  // T[i] = T_{i+1}(u), i < k   (all aligned to T[k-1])
    std::vector<Ct> T2; // T2[i] = T_{k 2^i}(u)
    std::vector<Ct> Tr; // T[i] brought to degree 1 once (the scalar products of every leaf start from these)
    std::vector<Ct> ready; // weighted sums of the baby steps, precomputed in traversal order by one fused kernel family
    size_t next = 0;
    const Ct &reduced(int i) {
        if (Tr.empty()) Tr.resize(T.size());
        if (!Tr[i]) Tr[i] = T[i]->nsd == 2 ? C->mod_reduce(T[i]) : T[i];
        return Tr[i];
    }
};

static Ct double_ct(Context *C, const Ct &a) { return C->add(a, a); }

// sum_{i=1..deg} w_i T_i + c0 ; returns null if everything vanishes
static Ct weighted_sum(PSState &S, const Poly &p, int upto) {
    if (!S.ready.empty()) {
        if (S.next >= S.ready.size()) throw std::logic_error("chebyshev PS: weighted-sum queue exhausted");
        return S.ready[S.next++];
    }
    Ct acc;
    for (int i = 1; i <= upto && i < (int)p.size(); i++) {
        if (p[i] == 0.0) continue;
        Ct t = S.C->mult_core_const(S.reduced(i - 1), p[i]);
        acc = acc ? S.C->add(acc, t) : t;
    }
    return acc;
}
// polynomial of degree exactly k whose T_k coefficient is a power of two
static Ct eval_leaf(PSState &S, const Poly &p) {
    const int k = S.k;
    double lead = p.size() > (size_t)k ? p[k] : 0.0;
    Ct r = weighted_sum(S, p, k - 1);
    if (lead != 0.0) {
        double l2 = std::log2(std::fabs(lead));
        long j = std::lround(l2);
        Ct top;
        if (lead > 0 && std::fabs(l2 - (double)j) < 1e-9 && j >= 0) {
            top = S.T[k - 1];
            for (long x = 0; x < j; x++) top = double_ct(S.C, top);
        } else {
            top = S.C->mult_const(S.T[k - 1], lead);
        }
        r = r ? S.C->add(r, top) : top;
    }
    if (!r) throw std::logic_error("chebyshev PS: empty leaf");
    if (p[0] != 0.0) r = S.C->add_const(r, p[0]);
    return r;
}
// f = (T_{k 2^(m-1)} + c) q + s2 with deg q = deg s2 = k(2^(m-1) - 1), both with power-of-two leading coefficient
static void ps_split(int k, int m, const Poly &f, Poly &q, Poly &c, Poly &s2) {
    const int k2m2k = k * (1 << (m - 1)) - k;
    Poly Tkm(k2m2k + k + 1, 0.0);
    Tkm.back() = 1.0;
    Poly r, s;
    cheb_divide(f, Tkm, q, r);
    Poly r2 = r;
    if ((int)r2.size() < k2m2k + 1) r2.resize(k2m2k + 1, 0.0);
    r2[k2m2k] -= 1.0;
    cheb_divide(r2, q, c, s);
    s2 = s;
    s2.resize(k2m2k + 1, 0.0);
    s2[k2m2k] += 1.0;
}
// the weight vectors eval_inner will ask weighted_sum for, in the order it asks
static void ps_collect(int k, int m, const Poly &f, std::vector<Poly> &req) {
    if (m == 1) {
        req.push_back(f);
        return;
    }
    Poly q, c, s2;
    ps_split(k, m, f, q, c, s2);
    req.push_back(c);
    ps_collect(k, m - 1, q, req);
    ps_collect(k, m - 1, s2, req);
}
// f has degree exactly k(2^m - 1) with power-of-two leading coefficient
static Ct eval_inner(PSState &S, const Poly &f, int m) {
    const int k = S.k;
    if (m == 1) return eval_leaf(S, f);
    Poly q, c, s2;
    {
        HostTimer ts;
        ps_split(k, m, f, q, c, s2);
        g_split_ms += ts.lap();
    }
    Ct cu = weighted_sum(S, c, k - 1);
    // a multiplicand has to be relinearised and brought to degree 1: one merged division where the recursion returned the
    // unrelinearised degree-2 sum (Context::relinearize_reduced)
    Ct qu = S.C->relinearize_reduced(eval_inner(S, q, m - 1));
    Ct su = eval_inner(S, s2, m - 1);                   // ... a summand not: it may keep its s^2 polynomial
    Ct left = cu ? S.C->add(S.T2[m - 1], cu) : S.T2[m - 1];
    if (c[0] != 0.0) left = S.C->add_const(left, c[0]);
    // product and remainder share one key switch, done by whoever multiplies by (or returns) the sum: half of the
    // products of the recursion never pay for a relinearisation of their own
    static const bool lazy = [] { const char *e = getenv("FHEB200_PS_LAZY_RELIN"); return !(e && *e == '0'); }();
    if (!lazy) return S.C->add(S.C->mult(left, qu), S.C->relinearize(su));
    return S.C->add_lazy(S.C->mult_lazy(left, qu), su);
}

static void choose_km(int n, int &k, int &m) {
    int bestD = 1 << 30;
    long bestCost = 1L << 60;
    for (int kk = 2; kk <= std::max(2, n + 1); kk++) {
        int lg = 0;
        while ((1 << lg) < kk) lg++;
        for (int mm = 1; mm <= 20; mm++) {
            if ((long)kk * ((1L << mm) - 1) <= n) continue;
            int D = lg + mm;
            long cost = kk + 2 * mm + (1L << (mm - 1)) - 4;
            if (D < bestD || (D == bestD && cost < bestCost)) bestD = D, bestCost = cost, k = kk, m = mm;
            break; // larger m only costs more for this k
        }
    }
}

// u = (2x - (a+b)) / (b-a); identity when [a,b] = [-1,1] (no level consumed)
static Ct to_unit_interval(Context *C, const Ct &x, double a, double b) {
    if (std::fabs(a + 1.0) < 1e-12 && std::fabs(b - 1.0) < 1e-12) return x;
    Ct u = C->mult_const(x, 2.0 / (b - a));
    return C->add_const(u, -(a + b) / (b - a));
}
// T_1..T_count by T_2i = 2 T_i^2 - 1, T_2i+1 = 2 T_i T_{i+1} - T_1
static std::vector<Ct> chebyshev_basis(Context *C, const Ct &u, int count) {
    std::vector<Ct> T(count);
    T[0] = u;
    // T_1 brought to the level and degree of a product is the same ciphertext for every product of that level: align it once per
    // level instead of once per odd index (each alignment is a scalar product and a rescale)
    std::map<std::pair<int, int>, Ct> t1_at;
    for (int i = 2; i <= count; i++) {
        if (i & 1) {
            Ct prod = C->mult(T[i / 2 - 1], T[i / 2]);
            Ct &t1 = t1_at[{prod->level, prod->nsd}];
            if (!t1) {
                Ct t = T[0], d = prod;
                C->adjust_pair(t, d);
                if (d.get() == prod.get()) t1 = t; // (the product itself was not touched: t is T_1 in its shape)
            }
            T[i - 1] = C->sub(double_ct(C, prod), t1 ? t1 : T[0]);
        } else {
            Ct sq = C->mult(T[i / 2 - 1], T[i / 2 - 1]);
            T[i - 1] = C->add_const(double_ct(C, sq), -1.0);
        }
    }
    return T;
}

Ct Context::eval_chebyshev_linear(const Ct &x, const std::vector<double> &coeffs, double a, double b) {
    CtxLock g(*this);
    int n = degree_of(coeffs);
    if (n < 1) throw std::invalid_argument("EvalChebyshevSeries: degree must be >= 1");
    Ct u = to_unit_interval(this, x, a, b);
    std::vector<Ct> T = chebyshev_basis(this, u, n);
    Ct acc;
    for (int i = 1; i <= n; i++) {
        if (coeffs[i] == 0.0) continue;
        Ct t = mult_const(T[i - 1], coeffs[i]);
        acc = acc ? add(acc, t) : t;
    }
    if (!acc) acc = mult_const(T[0], 0.0);
    return add_const(acc, coeffs[0] / 2.0);
}

Ct Context::eval_chebyshev(const Ct &x, const std::vector<double> &coeffs, double a, double b) {
    CtxLock g(*this);
    static const bool pstime = getenv("FHEB200_PSTIME") != nullptr;
    HostTimer tm;
    double t_basis = 0, t_giant = 0, t_collect = 0, t_lincomb = 0;
    int n = degree_of(coeffs);
    if (n < 5) return eval_chebyshev_linear(x, coeffs, a, b);
    PSState S;
    S.C = this;
    choose_km(n, S.k, S.m);
    const int k = S.k, m = S.m;
    Ct u = to_unit_interval(this, x, a, b);
    S.T = chebyshev_basis(this, u, k);
    // One alignment of all baby steps to the deepest one.  Only their degree-1 forms are ever used below (the scalar products of
    // the leaf sums), so the alignment is done on those: one scalar product + one rescale per baby step instead of aligning the
    // degree-2 forms and rescaling them again (FHEB200_PS_ALIGN_REDUCED=0 restores that).
    static const bool align_reduced = [] { const char *e = getenv("FHEB200_PS_ALIGN_REDUCED"); return !(e && *e == '0'); }();
    if (align_reduced) {
        S.Tr.resize(k);
        const Ct d1 = S.T[k - 1]->nsd == 2 ? reduced(S.T[k - 1]) : S.T[k - 1];
        for (int i = 0; i + 1 < k; i++) {
            Ct t = S.T[i]->nsd == 2 ? reduced(S.T[i]) : S.T[i], d = d1;
            adjust_pair(t, d);
            if (t->nsd != 1 || t->ell != d1->ell) { // (not the shape the shortcut assumes: fall back for this one)
                Ct t2 = S.T[i], d2 = S.T[k - 1];
                adjust_pair(t2, d2);
                S.T[i] = t2;
                continue;
            }
            S.Tr[i] = t;
        }
    } else {
        for (int i = 0; i + 1 < k; i++) {
            Ct t = S.T[i], d = S.T[k - 1];
            adjust_pair(t, d);
            S.T[i] = t;
        }
    }
    t_basis = tm.lap();
    S.T2.resize(m);
    S.T2[0] = S.T[k - 1];
    for (int i = 1; i < m; i++) S.T2[i] = add_const(double_ct(this, mult(S.T2[i - 1], S.T2[i - 1])), -1.0);
    // T_{k(2^m - 1)} by T_{a+b} = 2 T_a T_b - T_{a-b} with a = k 2^i, b = k(2^i - 1), a - b = k
    Ct T2km1 = S.T2[0];
    for (int i = 1; i < m; i++) T2km1 = sub(double_ct(this, mult(T2km1, S.T2[i])), S.T2[0]);
    t_giant = tm.lap();
    Poly f(coeffs.begin(), coeffs.begin() + n + 1);
    f[0] /= 2.0; // the series convention is c_0/2 + sum c_i T_i
    const int top = k * ((1 << m) - 1);
    f.resize(top + 1, 0.0);
    f[top] += 1.0;
    if (k > 2) { // every weighted sum of T_1..T_{k-1} the recursion needs, four outputs per pass over the baby steps
        std::vector<Poly> req;
        ps_collect(k, m, f, req);
        t_collect = tm.lap();
        std::vector<Ct> in(k - 1);
        for (int i = 0; i + 1 < k; i++) in[i] = S.reduced(i);
        std::vector<std::vector<double>> w(req.size(), std::vector<double>(k - 1, 0.0));
        for (size_t r = 0; r < req.size(); r++)
            for (int i = 1; i < k && i < (int)req[r].size(); i++) w[r][i - 1] = req[r][i];
        S.ready = lincomb(in, w);
        t_lincomb = tm.lap();
    }
    g_split_ms = 0;
    Ct r = eval_inner(S, f, m);
    if (pstime)
        fprintf(stderr, "PS deg %d k %d m %d ell %d: basis %.2f giant %.2f collect %.2f lincomb %.2f inner %.2f (splits %.2f) ms host, ends at launch %llu\n",
                n, k, m, x->ell, t_basis, t_giant, t_collect, t_lincomb, tm.lap(), g_split_ms, (unsigned long long)g_launches.load());
    return sub(relinearize(r), T2km1);
}

// power-basis evaluation with binary powers, depth ceil(log2 deg) + 1
Ct Context::eval_poly_linear(const Ct &x, const std::vector<double> &coeffs) {
    CtxLock g(*this);
    int n = degree_of(coeffs);
    if (n < 1) throw std::invalid_argument("EvalPoly: degree must be >= 1");
    std::vector<char> need(n + 1, 0);
    for (int i = 1; i <= n; i++) need[i] = coeffs[i] != 0.0;
    for (int i = n; i >= 2; i--)
        if (need[i]) need[i / 2] = 1, need[i - i / 2] = 1;
    std::vector<Ct> pw(n + 1);
    pw[1] = x;
    for (int i = 2; i <= n; i++)
        if (need[i]) pw[i] = mult(pw[i / 2], pw[i - i / 2]);
    Ct acc;
    for (int i = 1; i <= n; i++) {
        if (coeffs[i] == 0.0) continue;
        Ct t = mult_const(pw[i], coeffs[i]);
        acc = acc ? add(acc, t) : t;
    }
    return coeffs[0] != 0.0 ? add_const(acc, coeffs[0]) : acc;
}

std::vector<double> chebyshev_coefficients(const std::function<double(double)> &f, double a, double b, int degree) {
    if (degree < 0) throw std::invalid_argument("EvalChebyshevCoefficients: negative degree");
    const int N = degree + 1;
    const double bma = 0.5 * (b - a), bpa = 0.5 * (b + a);
    std::vector<double> fv(N), out(N);
    for (int i = 0; i < N; i++) fv[i] = f(std::cos(M_PI * (i + 0.5) / N) * bma + bpa);
    for (int kx = 0; kx < N; kx++) {
        double s = 0;
        for (int i = 0; i < N; i++) s += fv[i] * std::cos(M_PI * kx * (i + 0.5) / N);
        out[kx] = s * 2.0 / N;
    }
    return out;
}

} // namespace fheb200
