// bootstrap_plan.cpp — see bootstrap_plan.h.  Everything here is double-precision host arithmetic.
//
// Slot vectors.  A ciphertext with `s` slots holds a polynomial t(Y), Y = X^(n/2s), of 2s real coefficients; with
// w_k = t_k + i t_{k+s} its slots are z = U w, U[j][k] = xi^(5^j k), xi = exp(2 pi i / 4s).  The special FFT factors U as
// S_s ... S_4 S_2 R (R = bit reversal, S_len = radix-2 butterflies over blocks of len).  CoeffsToSlots applies
// S_2^-1 ... S_s^-1 (s first), leaving w in bit-reversed order, which the slot-wise EvalMod does not mind; SlotsToCoeffs
// applies S_2 first ... S_s last.  Consecutive factors are multiplied together into `budget` groups.
//
// Both transforms run on 2s slots so that one EvalMod serves the real and the imaginary coefficients: the last
// CoeffsToSlots group carries -i on its upper half, so that v + conj(v) = [2 Re w ; 2 Im w]; the first SlotsToCoeffs
// group carries rotate(c, r) with c = [1 ; i] on its diagonals, and out + rotate(out, s) = [U(a + i b) ; same].
#include "bootstrap_plan.h"

#include <cmath>
#include <cstdlib>
#include <functional>
#include <map>
#include <set>
#include <stdexcept>

namespace fheb200 {

std::vector<double> chebyshev_coefficients(const std::function<double(double)> &f, double a, double b, int degree);

typedef std::complex<double> cd;
typedef std::map<int, std::vector<cd>> Diags; // rotation (mod s) -> diagonal of length s

static std::vector<int> rot_group(int s) {
    std::vector<int> r(s);
    long long p = 1;
    for (int j = 0; j < s; j++, p = p * 5 % (4LL * s)) r[j] = (int)p;
    return r;
}
static void put(Diags &D, int r, int s, int pos, cd v) {
    auto &d = D[r];
    if (d.empty()) d.assign(s, cd(0, 0));
    d[pos] += v;
}
// S_len (inverse = false) or S_len^-1 (inverse = true, with its factor 1/2) as diagonals over Z_s
static Diags butterfly(int s, int len, bool inverse) {
    const int M = 4 * s, lenh = len / 2, lenq = len * 4, gap = M / lenq;
    const std::vector<int> rg = rot_group(s);
    Diags D;
    const int up = lenh % s, down = (s - lenh) % s;
    for (int i = 0; i < s; i += len)
        for (int j = 0; j < lenh; j++) {
            const double ang = 2.0 * M_PI * (double)((rg[j] % lenq) * gap) / (double)M;
            const int lo = i + j, hi = i + j + lenh;
            if (!inverse) { // lo: v[lo] + w v[hi] ; hi: v[lo] - w v[hi]
                const cd w(std::cos(ang), std::sin(ang));
                put(D, 0, s, lo, 1.0), put(D, up, s, lo, w);
                put(D, down, s, hi, 1.0), put(D, 0, s, hi, -w);
            } else {        // lo: (v[lo] + v[hi]) / 2 ; hi: conj(w) (v[lo] - v[hi]) / 2
                const cd w(0.5 * std::cos(ang), -0.5 * std::sin(ang));
                put(D, 0, s, lo, 0.5), put(D, up, s, lo, 0.5);
                put(D, down, s, hi, w), put(D, 0, s, hi, -w);
            }
        }
    return D;
}
// A after B: (AB)_{r1+r2}[j] += A_{r1}[j] B_{r2}[j + r1]
static Diags compose(const Diags &A, const Diags &B, int s) {
    Diags R;
    for (auto &a : A)
        for (auto &b : B) {
            auto &d = R[(a.first + b.first) % s];
            if (d.empty()) d.assign(s, cd(0, 0));
            for (int j = 0; j < s; j++) d[j] += a.second[j] * b.second[(j + a.first) % s];
        }
    return R;
}
static std::vector<std::vector<int>> split(const std::vector<int> &lens, int budget) {
    std::vector<std::vector<int>> g;
    const int total = (int)lens.size();
    int pos = 0;
    for (int i = 0; i < budget; i++) {
        int cnt = total / budget + (i < total % budget ? 1 : 0);
        g.emplace_back(lens.begin() + pos, lens.begin() + pos + cnt);
        pos += cnt;
    }
    return g;
}
static BootStage to_stage(const Diags &D, int s, double factor) {
    BootStage st;
    for (auto &e : D) {
        double mx = 0;
        for (auto &v : e.second) mx = std::max(mx, std::abs(v));
        if (mx < 1e-13) continue; // cancelled diagonal
        st.rot.push_back(e.first);
        std::vector<cd> d(2 * s);
        for (int j = 0; j < 2 * s; j++) d[j] = e.second[j % s] * factor;
        st.diag.push_back(std::move(d));
    }
    return st;
}

BootPlan make_boot_plan(int n, int slots, int b_enc, int b_dec, double q0, double raise_scale, int correction) {
    if (slots < 2 || (slots & (slots - 1))) throw std::invalid_argument("EvalBootstrapSetup: slots must be a power of two >= 2");
    if (4 * slots > n) throw std::invalid_argument("EvalBootstrapSetup: fully packed bootstrapping (slots = ring_dim / 2) is not built; use slots <= ring_dim / 4");
    BootPlan P;
    P.n = n, P.slots = slots, P.q0 = q0, P.raise_scale = raise_scale, P.correction = correction;
    P.msg_scale = std::ldexp(q0, -correction);
    const int s = slots;
    int logs = 0;
    while ((1 << logs) < s) logs++;
    b_enc = std::max(1, std::min(b_enc, logs));
    b_dec = std::max(1, std::min(b_dec, logs));
    const double g = (double)n / (2.0 * s); // SubSum multiplies the polynomial by n / 2s
    std::vector<int> down, up;
    for (int len = s; len >= 2; len >>= 1) down.push_back(len);
    for (int len = 2; len <= s; len <<= 1) up.push_back(len);

    // CoeffsToSlots: slots of the raised ciphertext = U w g / raise_scale; wanted: x / (1 + delta) with x = [Re w ; Im w] /
    // (q0 K) in [-1, 1] and delta = 1 / 4K (see EvalMod below); conj-add doubles
    {
        const char *e = getenv("FHEB200_BOOT_EVEN");
        P.even = !(e && *e == '0');
    }
    const double delta = P.even ? 1.0 / (4.0 * (double)P.K) : 0.0;
    const double k_enc = std::pow(raise_scale / (2.0 * g * q0 * (double)P.K * (1.0 + delta)), 1.0 / b_enc);
    auto ge = split(down, b_enc);
    for (size_t i = 0; i < ge.size(); i++) {
        Diags D;
        for (int len : ge[i]) D = D.empty() ? butterfly(s, len, true) : compose(butterfly(s, len, true), D, s);
        BootStage st = to_stage(D, s, k_enc);
        if (i + 1 == ge.size())
            for (auto &d : st.diag)
                for (int j = s; j < 2 * s; j++) d[j] *= cd(0, -1);
        P.cts.push_back(std::move(st));
    }
    // SlotsToCoeffs: EvalMod leaves sin(2 pi t / q0) ~ 2 pi msg_scale m / q0 for message coefficient m
    const double k_dec = std::pow(q0 / (2.0 * M_PI * P.msg_scale), 1.0 / b_dec);
    auto gd = split(up, b_dec);
    for (size_t i = 0; i < gd.size(); i++) {
        Diags D;
        for (int len : gd[i]) D = D.empty() ? butterfly(s, len, false) : compose(butterfly(s, len, false), D, s);
        BootStage st = to_stage(D, s, k_dec);
        if (i == 0)
            for (size_t t = 0; t < st.rot.size(); t++)
                for (int j = 0; j < 2 * s; j++)
                    if ((j + st.rot[t]) % (2 * s) >= s) st.diag[t][j] *= cd(0, 1);
        P.stc.push_back(std::move(st));
    }
    // EvalMod: cos(2 pi (K x - 1/4) / 2^r) = cos(w u) with u = (x - delta) / (1 + delta) in [-1, 1) and
    // w = 2 pi K (1 + delta) / 2^r.  Even in u, so it is a series in T_2j(u) = T_j(y), y = 2 u^2 - 1: half the degree for one
    // extra square.  cos(w sqrt((y + 1) / 2)) is entire in y, so plain Chebyshev interpolation in y converges as fast.
    // The conditioning in y is worse near u = 0 (|dH/dy| up to w^2 / 4 against |dh/du| <= w): about one bit of precision
    // for 5 of 21 multiplications; FHEB200_BOOT_EVEN=0 evaluates the degree-88 series in u = x instead.
    P.shift = delta / (1.0 + delta);
    const double w = 2.0 * M_PI * (double)P.K * (1.0 + delta) / (double)(1 << P.r);
    if (P.even) {
        P.degree = 44;
        P.cheb = chebyshev_coefficients([w](double y) { return std::cos(w * std::sqrt(std::max(0.0, (y + 1.0) / 2.0))); }, -1.0, 1.0, P.degree);
    } else {
        P.degree = 88;
        const double quarter = M_PI / 2.0 / (double)(1 << P.r);
        P.cheb = chebyshev_coefficients([w, quarter](double u) { return std::cos(w * u - quarter); }, -1.0, 1.0, P.degree);
    }
    return P;
}

std::vector<int> BootPlan::rotations() const {
    std::set<int> R;
    for (int step = slots; step < n / 2; step <<= 1) R.insert(step); // SubSum; `slots` is also the final fold
    R.insert(slots);
    for (auto *v : {&cts, &stc})
        for (auto &st : *v)
            for (int x : st.rot)
                if (x) R.insert(x);
    return std::vector<int>(R.begin(), R.end());
}

} // namespace fheb200
