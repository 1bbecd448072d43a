// params.cpp — modulus chain, roots, scaling factors and base-conversion plans (host only).
#include "params.h"

#include <algorithm>
#include <cmath>
#include <stdexcept>

namespace fheb200 {

u64 mulmod_h(u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); }
u64 powmod_h(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod_h(r, a, q);
        a = mulmod_h(a, a, q);
    }
    return r;
}
u64 invmod_h(u64 a, u64 q) { return powmod_h(a % q, q - 2, q); }

static int bitlen(u64 x) { return x ? 64 - __builtin_clzll(x) : 0; }

ModConst make_modconst(u64 q) {
    ModConst m{};
    m.q = q;
    m.bits = (uint32_t)bitlen(q);
    m.mu = (u64)(((u128)1 << (m.bits + 63)) / q);
    m.r64 = (u64)((((u128)1) << 64) / q);
    m.c64 = (u64)((((u128)1) << 64) % q);
    m.c64s = shoup_h(m.c64, q);
    return m;
}

static bool miller_rabin(u64 n) {
    if (n < 2) return false;
    for (u64 p : {2ULL, 3ULL, 5ULL, 7ULL, 11ULL, 13ULL, 17ULL, 19ULL, 23ULL, 29ULL, 31ULL, 37ULL})
        if (n % p == 0) return n == p;
    u64 d = n - 1;
    int s = 0;
    while (!(d & 1)) d >>= 1, s++;
    for (u64 a : {2ULL, 3ULL, 5ULL, 7ULL, 11ULL, 13ULL, 17ULL, 19ULL, 23ULL, 29ULL, 31ULL, 37ULL}) {
        u64 x = powmod_h(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool witness = true;
        for (int r = 1; r < s && witness; r++) {
            x = mulmod_h(x, x, n);
            if (x == n - 1) witness = false;
        }
        if (witness) return false;
    }
    return true;
}
// NTT-friendly primes (= 1 mod 2n) around a power of two
static u64 prime_at_or_above(int bits, u64 m) {
    u64 c = (1ULL << bits) + 1;
    while (!miller_rabin(c)) c += m;
    return c;
}
static u64 prime_above(u64 c, u64 m) {
    for (c += m; !miller_rabin(c); c += m) {}
    return c;
}
static u64 prime_below(u64 c, u64 m) {
    for (c -= m; !miller_rabin(c); c -= m) {}
    return c;
}
static inline uint32_t brev(uint32_t x, int bits) { return bitrev_h(x, bits); }

// smallest primitive 2n-th root of unity
static u64 smallest_root(u64 m, u64 q) {
    u64 g = 0;
    for (u64 c = 2;; c++) {
        g = powmod_h(c, (q - 1) / m, q);
        if (powmod_h(g, m >> 1, q) == q - 1) break;
    }
    u64 g2 = mulmod_h(g, g, q), cur = g, best = g;
    for (u64 e = 1; e < m; e += 2, cur = mulmod_h(cur, g2, q)) best = std::min(best, cur);
    return best;
}

std::shared_ptr<Params> Params::create(int logn, int depth, int scale_bits, int first_bits, int dnum) {
    if (logn < 10 || logn > 17) throw std::invalid_argument("ring dimension must be 2^10 .. 2^17");
    if (depth < 1 || depth > 60) throw std::invalid_argument("multiplicative depth out of range");
    if (scale_bits < 20 || scale_bits > 59 || first_bits < scale_bits || first_bits > 60)
        throw std::invalid_argument("scaling/first modulus size out of range");
    if (dnum < 1) throw std::invalid_argument("dnum must be >= 1");
    auto P = std::make_shared<Params>();
    P->logn = logn;
    P->n = 1 << logn;
    P->depth = depth;
    P->scale_bits = scale_bits;
    P->first_bits = first_bits;
    P->dnum = dnum;
    P->L = depth + 2;
    const u64 m = 2ULL << logn;
    const int np = depth + 1;
    std::vector<u64> Q(P->L, 0);
    auto taken = [&](u64 v) { return std::find(Q.begin(), Q.end(), v) != Q.end(); };

    // scaling primes: keep sf_{k+1} = sf_k^2 / q close to 2^scale by alternating below/above the estimate
    Q[np - 1] = prime_at_or_above(scale_bits, m);
    double est = (double)Q[np - 1];
    for (int i = np - 2, flip = 0; i >= 1; i--, flip ^= 1) {
        est = est * est / (double)Q[i + 1];
        u64 r = (u64)std::llround(est);
        u64 c = flip ? r + m - r % m + 1 : r - m - r % m + 1;
        do c = flip ? prime_above(c, m) : prime_below(c, m);
        while (taken(c));
        Q[i] = c;
    }
    {   // first modulus
        u64 c;
        if (first_bits == scale_bits) {
            c = *std::min_element(Q.begin() + 1, Q.begin() + np);
        } else {
            c = prime_at_or_above(first_bits, m);
        }
        do c = prime_below(c, m);
        while (taken(c));
        Q[0] = c;
    }
    {   // FLEXIBLEAUTOEXT extra modulus (absorbs the fresh-encryption noise at the first rescale)
        int eb = std::min(scale_bits + 20, 60);
        u64 c;
        if (eb == first_bits) {
            c = Q[0];
            do c = prime_below(c, m);
            while (taken(c));
        } else {
            c = prime_at_or_above(eb, m);
            while (taken(c)) c = prime_above(c, m);
        }
        Q[P->L - 1] = c;
    }
    P->alpha = (P->L + dnum - 1) / dnum;
    int widest = 0;
    for (int s = 0; s < P->L; s += P->alpha) {
        int b = 0;
        for (int i = s; i < std::min(P->L, s + P->alpha); i++) b += bitlen(Q[i]);
        widest = std::max(widest, b);
    }
    P->K = (widest + 59) / 60;
    P->moduli = Q;
    for (u64 c = prime_at_or_above(60, m); (int)P->moduli.size() < P->L + P->K;) {
        c = prime_below(c, m);
        if (!taken(c)) P->moduli.push_back(c);
    }
    int F = P->L + P->K;
    P->psi.resize(F);
    P->mc.resize(F);
    P->n_inv.resize(F);
    P->n_inv_s.resize(F);
#pragma omp parallel for schedule(dynamic)
    for (int i = 0; i < F; i++) {
        u64 q = P->moduli[i];
        P->psi[i] = smallest_root(m, q);
        P->mc[i] = make_modconst(q);
        P->n_inv[i] = invmod_h((u64)P->n, q);
        P->n_inv_s[i] = shoup_h(P->n_inv[i], q);
    }
    // real scaling factors: sf[0] is the fresh "big" pseudo level, sf[l] the scale of a level-l, degree-1 ciphertext
    int S = P->L;
    P->sf.assign(S, 0.0);
    P->sf[0] = (double)Q[S - 1];
    P->sf[1] = (double)Q[S - 2];
    for (int k = 2; k < S; k++) P->sf[k] = P->sf[k - 1] * P->sf[k - 1] / (double)Q[S - k];
    P->sf_big.assign(S - 1, 0.0);
    P->sf_big[0] = P->sf[0] * P->sf[1];
    for (int k = 1; k < S - 1; k++) P->sf_big[k] = P->sf[k] * P->sf[k];
    return P;
}

void Params::twiddles(int qp, std::vector<u64> &w, std::vector<u64> &ws, std::vector<u64> &wi, std::vector<u64> &wis) const {
    u64 q = moduli[qp], g = psi[qp], gi = invmod_h(g, q);
    w.resize(n), ws.resize(n), wi.resize(n), wis.resize(n);
    u64 a = 1, b = 1;
    for (int i = 0; i < n; i++) {
        uint32_t r = brev((uint32_t)i, logn);
        w[r] = a, wi[r] = b;
        a = mulmod_h(a, g, q), b = mulmod_h(b, gi, q);
    }
    for (int i = 0; i < n; i++) ws[i] = shoup_h(w[i], q), wis[i] = shoup_h(wi[i], q);
}

uint32_t Params::galois_inverse(uint32_t k) const {
    u64 m = 2ULL << logn, r = 1, b = k;
    for (u64 e = (u64)n - 1; e; e >>= 1) { // |Z_2n^*| = n
        if (e & 1) r = r * b % m;
        b = b * b % m;
    }
    return (uint32_t)r;
}
uint32_t Params::galois_for_rotation(int rot) const {
    u64 m = 2ULL << logn, g = 5, k = 1;
    long slots = n / 2;
    long r = ((long)rot % slots + slots) % slots; // left rotation by r slots; the orbit of 5 has order n/2
    for (u64 e = (u64)r; e; e >>= 1) {
        if (e & 1) k = k * g % m;
        g = g * g % m;
    }
    return (uint32_t)k;
}

const LevelPlan &Params::plan(int ell) const {
    std::lock_guard<std::mutex> lock(mu_);
    auto it = plans_.find(ell);
    if (it != plans_.end()) return it->second;
    if (ell < 1 || ell > L) throw std::out_of_range("level plan: bad limb count");
    LevelPlan T;
    T.ell = ell;
    T.beta = beta(ell);
    T.width = ell + K;
    for (int t = 0; t < T.width; t++) T.qp.push_back(t < ell ? t : L + (t - ell));
    auto prod_except = [&](int first, int count, int skip, u64 mod) {
        u64 r = 1;
        for (int x = first; x < first + count; x++)
            if (x != skip) r = mulmod_h(r, moduli[x] % mod, mod);
        return r;
    };
    for (int j = 0; j < T.beta; j++) {
        int s = j * alpha, d = std::min(alpha, ell - s);
        T.dstart.push_back(s);
        T.dsize.push_back(d);
        std::vector<u64> hi(d), his(d), h((size_t)d * T.width);
        for (int i = 0; i < d; i++) {
            u64 qi = moduli[s + i];
            hi[i] = invmod_h(prod_except(s, d, s + i, qi), qi);
            his[i] = shoup_h(hi[i], qi);
            for (int t = 0; t < T.width; t++) h[(size_t)i * T.width + t] = prod_except(s, d, s + i, moduli[T.qp[t]]);
        }
        T.hat_inv.push_back(hi);
        T.hat_inv_s.push_back(his);
        T.hat.push_back(h);
    }
    T.p_hat_inv.resize(K), T.p_hat_inv_s.resize(K), T.p_hat.resize((size_t)K * ell);
    T.p_inv.resize(ell), T.p_inv_s.resize(ell), T.ql_inv.resize(ell), T.ql_inv_s.resize(ell);
    for (int k = 0; k < K; k++) {
        u64 pk = moduli[L + k];
        T.p_hat_inv[k] = invmod_h(prod_except(L, K, L + k, pk), pk);
        T.p_hat_inv_s[k] = shoup_h(T.p_hat_inv[k], pk);
        for (int i = 0; i < ell; i++) T.p_hat[(size_t)k * ell + i] = prod_except(L, K, L + k, moduli[i]);
    }
    for (int i = 0; i < ell; i++) {
        u64 qi = moduli[i];
        T.p_inv[i] = invmod_h(prod_except(L, K, -1, qi), qi);
        T.p_inv_s[i] = shoup_h(T.p_inv[i], qi);
        if (i < ell - 1) {
            T.ql_inv[i] = invmod_h(moduli[ell - 1] % qi, qi);
            T.ql_inv_s[i] = shoup_h(T.ql_inv[i], qi);
        }
    }
    return plans_.emplace(ell, std::move(T)).first->second;
}

} // namespace fheb200
