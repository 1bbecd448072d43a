/*
 * oracle/ckks_oracle.cpp — CPU oracle for the RNS-CKKS hot path.  TEST INFRASTRUCTURE ONLY.
 * See ckks_oracle.h for the scope statement ("PARITY UNPINNED" against OpenFHE: the library is absent).
 *
 * Reference call sites this restates (paths relative to /root/reference):
 *   EvalRotate / EvalFastRotation        src/rotation.h:224,231,286,344-350
 *   EvalMult / EvalSquare / Relinearize  src/sign.cpp:23-33,46-56 ; src/comparison.cpp:19,38
 *   rescale (implicit, FLEXIBLEAUTOEXT)  every EvalMult above
 *   MakeCKKSPackedPlaintext / Encrypt    src/encryption.cpp:5-12 ; src/sort_algo.h:317,341,356
 *   Decrypt / GetRealPackedValue         src/encryption.cpp:14-27
 *   KeyGen / EvalMultKeyGen / EvalRotateKeyGen   tests/DirectSortTest.cpp:43-47
 * Plain, loop-level C++ with unsigned __int128 arithmetic; deliberately independent of the CUDA engine.
 */
#include "ckks_oracle.h"

#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef uint64_t u64;
typedef int64_t i64;
typedef unsigned __int128 u128;
typedef __int128 i128;

static int g_threads = 1;

/* ------------------------------------------------------------------ modular arithmetic */
/* a*b mod q.  With reduced operands the 128-bit product has a high word below q, so one hardware 128/64 division gives
 * the remainder (x86-64 divq); any other case goes through the compiler's 128-bit remainder.  Hot loops with a constant
 * factor (NTT butterflies) use the Shoup form below instead, as OpenFHE's native backend does (ModMulFastConst). */
static inline u64 mulmod(u64 a, u64 b, u64 q) {
    u128 z = (u128)a * b;
#if defined(__x86_64__)
    u64 hi = (u64)(z >> 64), lo = (u64)z;
    if (hi < q) {
        u64 quo, rem;
        __asm__("divq %4" : "=a"(quo), "=d"(rem) : "a"(lo), "d"(hi), "r"(q) : "cc");
        (void)quo;
        return rem;
    }
#endif
    return (u64)(z % q);
}
/* 128-bit value mod q (lazy sums of products are reduced once) */
static inline u64 reduce128(u128 z, u64 q) {
#if defined(__x86_64__)
    u64 hi = (u64)(z >> 64) % q, lo = (u64)z, quo, rem;
    __asm__("divq %4" : "=a"(quo), "=d"(rem) : "a"(lo), "d"(hi), "r"(q) : "cc");
    (void)quo;
    return rem;
#else
    return (u64)(z % q);
#endif
}
/* Shoup multiplication by a constant w < q with ws = floor(w 2^64 / q): exact result in [0, q) for any a < 2^64 */
static inline u64 shoup_of(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
static inline u64 mulmod_shoup(u64 a, u64 w, u64 ws, u64 q) {
    u64 h = (u64)(((u128)a * ws) >> 64);
    u64 r = a * w - h * q;
    return r >= q ? r - q : r;
}
static inline u64 addmod(u64 a, u64 b, u64 q) {
    u64 s = a + b;
    return s >= q ? s - q : s;
}
static inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
static u64 powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    while (e) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
        e >>= 1;
    }
    return r;
}
static u64 invmod(u64 a, u64 q) { return powmod(a, q - 2, q); } /* q prime */

static bool is_prime(u64 n) {
    if (n < 2) return false;
    static const u64 small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (u64 p : small) {
        if (n % p == 0) return n == p;
    }
    u64 d = n - 1;
    int r = 0;
    while ((d & 1) == 0) {
        d >>= 1;
        r++;
    }
    for (u64 a : small) { /* deterministic for n < 3.3e24 */
        u64 x = powmod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool comp = true;
        for (int i = 1; i < r; i++) {
            x = mulmod(x, x, n);
            if (x == n - 1) {
                comp = false;
                break;
            }
        }
        if (comp) return false;
    }
    return true;
}
/* primes congruent to 1 mod m (m = 2n): smallest >= 2^bits, next above, previous below */
static u64 first_prime(int bits, u64 m) {
    u64 q = ((u64)1 << bits) + 1; /* m | 2^bits for bits > log2 m */
    while (!is_prime(q)) q += m;
    return q;
}
static u64 next_prime(u64 q, u64 m) {
    do q += m;
    while (!is_prime(q));
    return q;
}
static u64 prev_prime(u64 q, u64 m) {
    do q -= m;
    while (!is_prime(q));
    return q;
}
static inline uint32_t bitrev(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
/* minimal primitive m-th root of unity mod q (m a power of two), as OpenFHE's RootOfUnity does */
static u64 min_primitive_root(u64 m, u64 q) {
    u64 r = 0;
    for (u64 c = 2;; c++) {
        r = powmod(c, (q - 1) / m, q);
        if (powmod(r, m / 2, q) == q - 1) break;
    }
    u64 best = r, r2 = mulmod(r, r, q), cur = r;
    for (u64 k = 1; k < m; k += 2) { /* all odd powers are the primitive roots */
        if (cur < best) best = cur;
        cur = mulmod(cur, r2, q);
    }
    return best;
}

/* ------------------------------------------------------------------ context */
struct ModTables {
    u64 q, psi, psi_inv, n_inv;
    std::vector<u64> psi_br, psi_inv_br; /* psi^bitrev(i), psi^-bitrev(i) */
    std::vector<u64> psi_br_s, psi_inv_br_s; /* their Shoup companions */
    u64 n_inv_s;
};
struct LevelTables { /* per ell: partial-digit base-conversion constants */
    int beta;
    std::vector<int> dstart, dsize;                 /* digit j covers Q limbs [dstart, dstart+dsize) */
    std::vector<std::vector<u64>> hat_inv;          /* [j][i]   (Q_j/q_i)^-1 mod q_i            */
    std::vector<std::vector<std::vector<u64>>> hat; /* [j][i][t] (Q_j/q_i) mod modulus(t), t over ell+K */
    std::vector<u64> p_hat_inv;                     /* [k]   (P/p_k)^-1 mod p_k                 */
    std::vector<std::vector<u64>> p_hat;            /* [k][i] (P/p_k) mod q_i, i<ell            */
    std::vector<u64> p_inv;                         /* [i]   P^-1 mod q_i                       */
};
struct orc_ctx {
    int logn, n, depth, scale_bits, first_bits, dnum;
    int L, K, alpha;
    std::vector<ModTables> mods; /* QP order */
    std::vector<double> sf, sf_big;
    mutable std::map<int, LevelTables> lvl;
    mutable std::mutex mu;
};

static void build_mod(ModTables &t, u64 q, int logn) {
    int n = 1 << logn;
    t.q = q;
    t.psi = min_primitive_root(2 * (u64)n, q);
    t.psi_inv = invmod(t.psi, q);
    t.n_inv = invmod((u64)n, q);
    t.psi_br.resize(n);
    t.psi_inv_br.resize(n);
    t.psi_br_s.resize(n);
    t.psi_inv_br_s.resize(n);
    t.n_inv_s = shoup_of(t.n_inv, q);
    u64 p = 1, pi = 1;
    for (int i = 0; i < n; i++) {
        uint32_t r = bitrev((uint32_t)i, logn);
        t.psi_br[r] = p;
        t.psi_inv_br[r] = pi;
        t.psi_br_s[r] = shoup_of(p, q);
        t.psi_inv_br_s[r] = shoup_of(pi, q);
        p = mulmod(p, t.psi, q);
        pi = mulmod(pi, t.psi_inv, q);
    }
}

static int bits_of(u64 q) {
    int b = 0;
    while (q) {
        b++;
        q >>= 1;
    }
    return b;
}

/* Modulus chain as OpenFHE's CKKS parameter generation builds it for FLEXIBLEAUTOEXT + HYBRID
 * (restated from memory of the published source; see header: parity unpinned):
 *   q_depth = first prime >= 2^scale ; q_{depth-1..1} alternate below/above the running scaling-factor
 *   estimate sf <- sf^2 / q_{i+1} so that consecutive real scaling factors stay close to 2^scale ;
 *   q_0 = largest prime below the first prime >= 2^first ; extra prime of min(scale+20, 60) bits ;
 *   P = K 60-bit primes not in Q, K = ceil(max digit bits / 60). */
extern "C" orc_ctx *orc_ctx_create(int logn, int depth, int scale_bits, int first_bits, int dnum) {
    if (logn < 3 || logn > 17 || depth < 1 || scale_bits < 20 || scale_bits > 60 || first_bits > 60 ||
        first_bits < scale_bits || dnum < 1)
        return nullptr;
    orc_ctx *c = new orc_ctx();
    c->logn = logn;
    c->n = 1 << logn;
    c->depth = depth;
    c->scale_bits = scale_bits;
    c->first_bits = first_bits;
    c->dnum = dnum;
    const u64 m = 2 * (u64)c->n;
    const int numPrimes = depth + 1;
    c->L = depth + 2;
    std::vector<u64> q(c->L, 0);
    auto used = [&](u64 v) { return std::find(q.begin(), q.end(), v) != q.end(); };

    q[numPrimes - 1] = first_prime(scale_bits, m);
    {
        double sf = (double)q[numPrimes - 1];
        unsigned cnt = 0;
        for (int i = numPrimes - 2; i >= 1; i--) {
            sf = sf * sf / (double)q[i + 1];
            u64 sfInt = (u64)std::llround(sf);
            u64 rem = sfInt % m;
            if ((cnt & 1) == 0) {
                u64 cand = sfInt - m - rem + 1;
                do cand = prev_prime(cand, m);
                while (used(cand));
                q[i] = cand;
            } else {
                u64 cand = sfInt + m - rem + 1;
                do cand = next_prime(cand, m);
                while (used(cand));
                q[i] = cand;
            }
            cnt++;
        }
    }
    if (first_bits == scale_bits) {
        u64 lo = q[numPrimes - 1];
        for (int i = 1; i < numPrimes; i++) lo = std::min(lo, q[i]);
        u64 cand = lo;
        do cand = prev_prime(cand, m);
        while (used(cand));
        q[0] = cand;
    } else {
        u64 cand = first_prime(first_bits, m);
        do cand = prev_prime(cand, m);
        while (used(cand));
        q[0] = cand;
    }
    {
        int ebits = std::min(scale_bits + 20, 60);
        u64 cand;
        if (ebits == first_bits) {
            cand = q[0];
            do cand = prev_prime(cand, m);
            while (used(cand));
        } else {
            cand = first_prime(ebits, m);
            while (used(cand)) cand = next_prime(cand, m);
        }
        q[c->L - 1] = cand;
    }
    /* digits */
    c->alpha = (c->L + dnum - 1) / dnum;
    int maxbits = 0;
    for (int j = 0; j * c->alpha < c->L; j++) {
        int b = 0;
        for (int i = j * c->alpha; i < std::min(c->L, (j + 1) * c->alpha); i++) b += bits_of(q[i]);
        maxbits = std::max(maxbits, b);
    }
    c->K = (maxbits + 59) / 60;
    std::vector<u64> p;
    {
        u64 cand = first_prime(60, m);
        while ((int)p.size() < c->K) {
            cand = prev_prime(cand, m);
            if (!used(cand)) p.push_back(cand);
        }
    }
    c->mods.resize(c->L + c->K);
#pragma omp parallel for schedule(dynamic) num_threads(g_threads > 1 ? g_threads : 8)
    for (int i = 0; i < c->L + c->K; i++) build_mod(c->mods[i], i < c->L ? q[i] : p[i - c->L], logn);
    /* real scaling factors, FLEXIBLEAUTOEXT: level 0 is the fresh "big" level */
    int sizeQ = c->L;
    c->sf.assign(sizeQ, 0.0);
    c->sf[0] = (double)q[sizeQ - 1];
    c->sf[1] = (double)q[sizeQ - 2];
    for (int k = 2; k < sizeQ; k++) c->sf[k] = c->sf[k - 1] * c->sf[k - 1] / (double)q[sizeQ - k];
    c->sf_big.assign(sizeQ - 1, 0.0);
    c->sf_big[0] = c->sf[0] * c->sf[1];
    for (int k = 1; k < sizeQ - 1; k++) c->sf_big[k] = c->sf[k] * c->sf[k];
    return c;
}
extern "C" void orc_ctx_free(orc_ctx *c) { delete c; }
extern "C" int orc_n(const orc_ctx *c) { return c->n; }
extern "C" int orc_L(const orc_ctx *c) { return c->L; }
extern "C" int orc_K(const orc_ctx *c) { return c->K; }
extern "C" int orc_alpha(const orc_ctx *c) { return c->alpha; }
extern "C" uint64_t orc_modulus(const orc_ctx *c, int i) { return c->mods[i].q; }
extern "C" uint64_t orc_psi(const orc_ctx *c, int i) { return c->mods[i].psi; }
extern "C" double orc_scale_real(const orc_ctx *c, int l) { return c->sf[l]; }
extern "C" double orc_scale_big(const orc_ctx *c, int l) { return c->sf_big[l]; }
extern "C" void orc_set_threads(int t) { g_threads = t < 1 ? 1 : t; }

/* ------------------------------------------------------------------ NTT (negacyclic, bit-reversed twiddles) */
/* forward: Cooley-Tukey, natural order in -> bit-reversed order out */
extern "C" void orc_ntt_fwd(const orc_ctx *c, int idx, u64 *a) {
    const ModTables &T = c->mods[idx];
    const u64 q = T.q;
    int n = c->n, t = n;
    for (int m = 1; m < n; m <<= 1) {
        t >>= 1;
        for (int i = 0; i < m; i++) {
            const u64 w = T.psi_br[m + i], ws = T.psi_br_s[m + i];
            int j1 = 2 * i * t;
            for (int j = j1; j < j1 + t; j++) {
                u64 u = a[j], v = mulmod_shoup(a[j + t], w, ws, q);
                a[j] = addmod(u, v, q);
                a[j + t] = submod(u, v, q);
            }
        }
    }
}
/* inverse: Gentleman-Sande, bit-reversed in -> natural out, then * n^-1 */
extern "C" void orc_ntt_inv(const orc_ctx *c, int idx, u64 *a) {
    const ModTables &T = c->mods[idx];
    const u64 q = T.q;
    int n = c->n, t = 1;
    for (int m = n; m > 1; m >>= 1) {
        int h = m >> 1, j1 = 0;
        for (int i = 0; i < h; i++) {
            const u64 w = T.psi_inv_br[h + i], ws = T.psi_inv_br_s[h + i];
            for (int j = j1; j < j1 + t; j++) {
                u64 u = a[j], v = a[j + t];
                a[j] = addmod(u, v, q);
                a[j + t] = mulmod_shoup(submod(u, v, q), w, ws, q);
            }
            j1 += 2 * t;
        }
        t <<= 1;
    }
    for (int j = 0; j < n; j++) a[j] = mulmod_shoup(a[j], T.n_inv, T.n_inv_s, q);
}

/* ------------------------------------------------------------------ automorphisms */
extern "C" uint32_t orc_rot_to_galois(const orc_ctx *c, int r) {
    u64 m = 2 * (u64)c->n;
    u64 g = 5;
    if (r < 0) { /* 5^-1 mod 2n */
        u64 inv = 1, base = 5; /* 5^(m/4 - 1) since 5 has order n/2 = m/4 */
        u64 e = m / 4 - 1;
        while (e) {
            if (e & 1) inv = inv * base % m;
            base = base * base % m;
            e >>= 1;
        }
        g = inv;
        r = -r;
    }
    u64 k = 1;
    for (int i = 0; i < r; i++) k = k * g % m;
    return (uint32_t)k;
}
/* out[rev(j)] = in[rev(idx(j))], idx(j) = ((2j+1)k mod 2n - 1)/2 : b(psi^(2j+1)) = a(psi^((2j+1)k)) */
extern "C" void orc_automorphism_eval(const orc_ctx *c, uint32_t k, int limbs, const u64 *in, u64 *out) {
    int n = c->n, logn = c->logn;
    u64 m = 2 * (u64)n;
    std::vector<uint32_t> map(n);
    for (int j = 0; j < n; j++) {
        u64 idx = (((2 * (u64)j + 1) * k) % m) >> 1;
        map[bitrev(j, logn)] = bitrev((uint32_t)idx, logn);
    }
    for (int l = 0; l < limbs; l++)
        for (int j = 0; j < n; j++) out[(size_t)l * n + j] = in[(size_t)l * n + map[j]];
}
extern "C" void orc_automorphism_coeff(const orc_ctx *c, uint32_t k, int idx, const u64 *in, u64 *out) {
    int n = c->n;
    u64 q = c->mods[idx].q, m = 2 * (u64)n;
    for (int i = 0; i < n; i++) {
        u64 e = ((u64)i * k) % m;
        if (e < (u64)n)
            out[e] = in[i];
        else
            out[e - n] = in[i] ? q - in[i] : 0;
    }
}

/* ------------------------------------------------------------------ elementwise (a5, a6), tensor (a3), rescale (a4) */
#define PAR_LIMBS _Pragma("omp parallel for schedule(static) num_threads(g_threads) if (g_threads > 1)")

extern "C" void orc_add(const orc_ctx *c, int ell, const u64 *a, const u64 *b, u64 *o) {
    int n = c->n;
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        for (int j = 0; j < n; j++) o[(size_t)l * n + j] = addmod(a[(size_t)l * n + j], b[(size_t)l * n + j], q);
    }
}
extern "C" void orc_sub(const orc_ctx *c, int ell, const u64 *a, const u64 *b, u64 *o) {
    int n = c->n;
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        for (int j = 0; j < n; j++) o[(size_t)l * n + j] = submod(a[(size_t)l * n + j], b[(size_t)l * n + j], q);
    }
}
extern "C" void orc_mul(const orc_ctx *c, int ell, const u64 *a, const u64 *b, u64 *o) {
    int n = c->n;
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        for (int j = 0; j < n; j++) o[(size_t)l * n + j] = mulmod(a[(size_t)l * n + j], b[(size_t)l * n + j], q);
    }
}
extern "C" void orc_mul_scalar(const orc_ctx *c, int ell, const u64 *a, const u64 *s, u64 *o) {
    int n = c->n;
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        for (int j = 0; j < n; j++) o[(size_t)l * n + j] = mulmod(a[(size_t)l * n + j], s[l], q);
    }
}
extern "C" void orc_add_scalar(const orc_ctx *c, int ell, const u64 *a, const u64 *s, u64 *o) {
    int n = c->n;
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        for (int j = 0; j < n; j++) o[(size_t)l * n + j] = addmod(a[(size_t)l * n + j], s[l], q);
    }
}
/* d0 = a0 b0, d1 = a0 b1 + a1 b0, d2 = a1 b1 */
extern "C" void orc_tensor(const orc_ctx *c, int ell, const u64 *a0, const u64 *a1, const u64 *b0, const u64 *b1,
                           u64 *d0, u64 *d1, u64 *d2) {
    int n = c->n;
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        for (int j = 0; j < n; j++) {
            size_t x = (size_t)l * n + j;
            u64 A0 = a0[x], A1 = a1[x], B0 = b0[x], B1 = b1[x];
            d0[x] = mulmod(A0, B0, q);
            d1[x] = addmod(mulmod(A0, B1, q), mulmod(A1, B0, q), q);
            d2[x] = mulmod(A1, B1, q);
        }
    }
}
/* Drop the last limb and divide by q_last with rounding to nearest: the last limb goes to the coefficient
 * domain, is lifted to its centred representative, reduced mod each remaining q_i, transformed back and
 * out_i = (in_i - lift_i) * q_last^-1 mod q_i.  (OpenFHE: DCRTPoly::DropLastElementAndScale.) */
extern "C" void orc_rescale(const orc_ctx *c, int ell, const u64 *in, u64 *out) {
    int n = c->n;
    u64 ql = c->mods[ell - 1].q, half = ql >> 1;
    std::vector<u64> last(in + (size_t)(ell - 1) * n, in + (size_t)ell * n);
    orc_ntt_inv(c, ell - 1, last.data());
    PAR_LIMBS for (int l = 0; l < ell - 1; l++) {
        u64 q = c->mods[l].q, qlinv = invmod(ql % q, q);
        std::vector<u64> t(n);
        for (int j = 0; j < n; j++) {
            u64 v = last[j];
            if (v > half) { /* negative representative v - ql */
                u64 d = (ql - v) % q;
                t[j] = d ? q - d : 0;
            } else
                t[j] = v % q;
        }
        orc_ntt_fwd(c, l, t.data());
        for (int j = 0; j < n; j++)
            out[(size_t)l * n + j] = mulmod(submod(in[(size_t)l * n + j], t[j], q), qlinv, q);
    }
}

/* ------------------------------------------------------------------ hybrid key switching (a1, a2) */
static const LevelTables &level_tables(const orc_ctx *c, int ell) {
    std::lock_guard<std::mutex> g(c->mu);
    auto it = c->lvl.find(ell);
    if (it != c->lvl.end()) return it->second;
    LevelTables T;
    int K = c->K, L = c->L, a = c->alpha;
    T.beta = (ell + a - 1) / a;
    auto modulus_t = [&](int t) { return t < ell ? c->mods[t].q : c->mods[L + (t - ell)].q; };
    for (int j = 0; j < T.beta; j++) {
        int s = j * a, d = std::min(a, ell - s);
        T.dstart.push_back(s);
        T.dsize.push_back(d);
        std::vector<u64> hi(d);
        std::vector<std::vector<u64>> h(d, std::vector<u64>(ell + K));
        for (int i = 0; i < d; i++) {
            u64 qi = c->mods[s + i].q, prod = 1;
            for (int x = 0; x < d; x++)
                if (x != i) prod = mulmod(prod, c->mods[s + x].q % qi, qi);
            hi[i] = invmod(prod, qi);
            for (int t = 0; t < ell + K; t++) {
                u64 mt = modulus_t(t), pr = 1;
                for (int x = 0; x < d; x++)
                    if (x != i) pr = mulmod(pr, c->mods[s + x].q % mt, mt);
                h[i][t] = pr;
            }
        }
        T.hat_inv.push_back(hi);
        T.hat.push_back(h);
    }
    T.p_hat_inv.resize(K);
    T.p_hat.assign(K, std::vector<u64>(ell));
    T.p_inv.resize(ell);
    for (int k = 0; k < K; k++) {
        u64 pk = c->mods[L + k].q, prod = 1;
        for (int x = 0; x < K; x++)
            if (x != k) prod = mulmod(prod, c->mods[L + x].q % pk, pk);
        T.p_hat_inv[k] = invmod(prod, pk);
        for (int i = 0; i < ell; i++) {
            u64 qi = c->mods[i].q, pr = 1;
            for (int x = 0; x < K; x++)
                if (x != k) pr = mulmod(pr, c->mods[L + x].q % qi, qi);
            T.p_hat[k][i] = pr;
        }
    }
    for (int i = 0; i < ell; i++) {
        u64 qi = c->mods[i].q, pr = 1;
        for (int x = 0; x < K; x++) pr = mulmod(pr, c->mods[L + x].q % qi, qi);
        T.p_inv[i] = invmod(pr, qi);
    }
    return c->lvl.emplace(ell, std::move(T)).first->second;
}
extern "C" int orc_beta(const orc_ctx *c, int ell) { return (ell + c->alpha - 1) / c->alpha; }
extern "C" uint64_t orc_evk_words(const orc_ctx *c) {
    return (uint64_t)orc_beta(c, c->L) * 2 * (c->L + c->K) * c->n;
}
/* QP index of "level limb" t at level ell: t<ell -> Q limb t ; else P limb t-ell */
static inline int qp_of(const orc_ctx *c, int ell, int t) { return t < ell ? t : c->L + (t - ell); }

/* ModUp (approximate fast basis extension, no correction term): digit j of c, in coefficient form, is
 * scaled by (Q_j/q_i)^-1 mod q_i and re-expanded as sum_i x_i (Q_j/q_i) mod every modulus outside the
 * digit; limbs inside the digit keep c's own evaluation-domain values. */
extern "C" void orc_modup(const orc_ctx *c, int ell, const u64 *cin, u64 *digits) {
    const LevelTables &T = level_tables(c, ell);
    int n = c->n, K = c->K, W = ell + K;
    std::vector<u64> coef((size_t)ell * n);
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        memcpy(&coef[(size_t)l * n], cin + (size_t)l * n, sizeof(u64) * n);
        orc_ntt_inv(c, l, &coef[(size_t)l * n]);
    }
    for (int j = 0; j < T.beta; j++) {
        int s = T.dstart[j], d = T.dsize[j];
        u64 *D = digits + (size_t)j * W * n;
        std::vector<u64> x((size_t)d * n);
        PAR_LIMBS for (int i = 0; i < d; i++) {
            const u64 qi = c->mods[s + i].q, h = T.hat_inv[j][i], hs = shoup_of(h, qi);
            for (int k = 0; k < n; k++) x[(size_t)i * n + k] = mulmod_shoup(coef[(size_t)(s + i) * n + k], h, hs, qi);
        }
        PAR_LIMBS for (int t = 0; t < W; t++) {
            u64 *dst = D + (size_t)t * n;
            if (t >= s && t < s + d) {
                memcpy(dst, cin + (size_t)t * n, sizeof(u64) * n);
                continue;
            }
            int qp = qp_of(c, ell, t);
            u64 mt = c->mods[qp].q;
            /* sum_i x_i * hat_i accumulated in 128 bits (d <= 16 terms below 2^120) and reduced once: the same residue
             * as reducing every term */
            for (int k = 0; k < n; k++) {
                u128 acc = 0;
                for (int i = 0; i < d; i++) acc += (u128)x[(size_t)i * n + k] * T.hat[j][i][t];
                dst[k] = reduce128(acc, mt);
            }
            orc_ntt_fwd(c, qp, dst);
        }
    }
}

/* acc_b[t] = sum_j digit_j[t] * evk[j][b][qp(t)] over the beta digits present at this level */
extern "C" void orc_inner_product(const orc_ctx *c, int ell, const u64 *digits, const u64 *evk, u64 *acc0, u64 *acc1) {
    int n = c->n, K = c->K, W = ell + K, F = c->L + K, beta = orc_beta(c, ell);
    PAR_LIMBS for (int t = 0; t < W; t++) {
        int qp = qp_of(c, ell, t);
        u64 q = c->mods[qp].q;
        for (int k = 0; k < n; k++) {
            u128 s0 = 0, s1 = 0; /* beta <= 8 products below 2^122 */
            for (int j = 0; j < beta; j++) {
                u64 d = digits[((size_t)j * W + t) * n + k];
                const u64 *kb = evk + (((size_t)j * 2 + 0) * F + qp) * n;
                const u64 *ka = evk + (((size_t)j * 2 + 1) * F + qp) * n;
                s0 += (u128)d * kb[k];
                s1 += (u128)d * ka[k];
            }
            acc0[(size_t)t * n + k] = reduce128(s0, q);
            acc1[(size_t)t * n + k] = reduce128(s1, q);
        }
    }
}
/* ModDown (approximate division by P): the P part is base-converted to Q and subtracted, then * P^-1. */
extern "C" void orc_moddown(const orc_ctx *c, int ell, const u64 *acc, u64 *out) {
    const LevelTables &T = level_tables(c, ell);
    int n = c->n, K = c->K, L = c->L;
    std::vector<u64> x((size_t)K * n);
    PAR_LIMBS for (int k = 0; k < K; k++) {
        u64 *p = &x[(size_t)k * n];
        memcpy(p, acc + (size_t)(ell + k) * n, sizeof(u64) * n);
        orc_ntt_inv(c, L + k, p);
        u64 pk = c->mods[L + k].q;
        for (int j = 0; j < n; j++) p[j] = mulmod(p[j], T.p_hat_inv[k], pk);
    }
    PAR_LIMBS for (int i = 0; i < ell; i++) {
        u64 q = c->mods[i].q;
        std::vector<u64> t(n);
        for (int j = 0; j < n; j++) {
            u128 s = 0;
            for (int k = 0; k < K; k++) s += (u128)x[(size_t)k * n + j] * T.p_hat[k][i];
            t[j] = reduce128(s, q);
        }
        orc_ntt_fwd(c, i, t.data());
        for (int j = 0; j < n; j++) out[(size_t)i * n + j] = mulmod(submod(acc[(size_t)i * n + j], t[j], q), T.p_inv[i], q);
    }
}
static void keyswitch_from_digits(const orc_ctx *c, int ell, const u64 *digits, const u64 *evk, u64 *ks0, u64 *ks1) {
    int n = c->n, W = ell + c->K;
    std::vector<u64> a0((size_t)W * n), a1((size_t)W * n);
    orc_inner_product(c, ell, digits, evk, a0.data(), a1.data());
    orc_moddown(c, ell, a0.data(), ks0);
    orc_moddown(c, ell, a1.data(), ks1);
}
extern "C" void orc_keyswitch(const orc_ctx *c, int ell, const u64 *cin, const u64 *evk, u64 *ks0, u64 *ks1) {
    int n = c->n, W = ell + c->K;
    std::vector<u64> digits((size_t)orc_beta(c, ell) * W * n);
    orc_modup(c, ell, cin, digits.data());
    keyswitch_from_digits(c, ell, digits.data(), evk, ks0, ks1);
}
extern "C" void orc_rotate_hoisted(const orc_ctx *c, int ell, uint32_t k, const u64 *c0, const u64 *digits,
                                   const u64 *evk, u64 *o0, u64 *o1) {
    size_t sz = (size_t)ell * c->n;
    std::vector<u64> k0(sz), k1(sz);
    keyswitch_from_digits(c, ell, digits, evk, k0.data(), k1.data());
    orc_add(c, ell, k0.data(), c0, k0.data());
    orc_automorphism_eval(c, k, ell, k0.data(), o0);
    orc_automorphism_eval(c, k, ell, k1.data(), o1);
}
extern "C" void orc_rotate(const orc_ctx *c, int ell, uint32_t k, const u64 *c0, const u64 *c1, const u64 *evk,
                           u64 *o0, u64 *o1) {
    int W = ell + c->K;
    std::vector<u64> digits((size_t)orc_beta(c, ell) * W * c->n);
    orc_modup(c, ell, c1, digits.data());
    orc_rotate_hoisted(c, ell, k, c0, digits.data(), evk, o0, o1);
}
extern "C" void orc_mult_relin(const orc_ctx *c, int ell, const u64 *a0, const u64 *a1, const u64 *b0, const u64 *b1,
                               const u64 *evk, u64 *o0, u64 *o1) {
    size_t sz = (size_t)ell * c->n;
    std::vector<u64> d0(sz), d1(sz), d2(sz), k0(sz), k1(sz);
    orc_tensor(c, ell, a0, a1, b0, b1, d0.data(), d1.data(), d2.data());
    orc_keyswitch(c, ell, d2.data(), evk, k0.data(), k1.data());
    orc_add(c, ell, d0.data(), k0.data(), o0);
    orc_add(c, ell, d1.data(), k1.data(), o1);
}

/* ------------------------------------------------------------------ deterministic randomness (spec: DESIGN.md) */
static inline u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
extern "C" uint64_t orc_rng_u64(u64 seed, u64 stream, u64 idx, u64 word) {
    u64 a = mix64(seed ^ (0x9E3779B97F4A7C15ULL * (stream + 1)));
    u64 b = mix64(a + idx);
    return mix64(b + 0xD1B54A32D192ED03ULL * (word + 1));
}
enum { ST_SK = 1, ST_PK_A = 2, ST_PK_E = 3, ST_ENC_V = 4, ST_ENC_E0 = 5, ST_ENC_E1 = 6, ST_EVK_A = 7, ST_EVK_E = 8 };
static inline u64 stream_id(u64 tag, u64 digit, u64 id) { return tag | (digit << 8) | (id << 16); }

static const u64 CDT[29] = {
    0x1001f9a1b2ca9468ULL, 0x2e7cf3ef07836cb5ULL, 0x48ca7e85d834a579ULL, 0x5d5d51778f760880ULL,
    0x6bf35598550421b0ULL, 0x7552dc90807bbed5ULL, 0x7ac8820561a4c974ULL, 0x7daa6596524b7ffdULL,
    0x7f0b8114bba4ecb9ULL, 0x7fa4a9fc5ea574e8ULL, 0x7fe0e1077aef3919ULL, 0x7ff6563f810f15d0ULL,
    0x7ffd4490999bc594ULL, 0x7fff4c0804cd386dULL, 0x7fffd5e18e7fda18ULL, 0x7ffff709c679c2edULL,
    0x7ffffe445c790f6bULL, 0x7fffffb20f7aa54aULL, 0x7ffffff3903d235dULL, 0x7ffffffe32b06d1dULL,
    0x7fffffffc35125ccULL, 0x7ffffffff8c11764ULL, 0x7fffffffff36fe33ULL, 0x7fffffffffec3bf7ULL,
    0x7ffffffffffe3c8bULL, 0x7fffffffffffdb75ULL, 0x7ffffffffffffd51ULL, 0x7fffffffffffffd2ULL,
    0x7ffffffffffffffdULL}; /* tools/gen_cdt.py, sigma = 3.19 */

/* uniform mod q from 128 random bits; element index = qp_index*n + j so every limb is independent */
extern "C" void orc_sample_uniform(const orc_ctx *c, u64 seed, u64 stream, int qp, u64 *out) {
    int n = c->n;
    u64 q = c->mods[qp].q;
    for (int j = 0; j < n; j++) {
        u64 idx = (u64)qp * n + j;
        u128 r = ((u128)orc_rng_u64(seed, stream, idx, 0) << 64) | orc_rng_u64(seed, stream, idx, 1);
        out[j] = (u64)(r % q);
    }
}
extern "C" void orc_sample_gauss(const orc_ctx *c, u64 seed, u64 stream, i64 *out) {
    for (int j = 0; j < c->n; j++) {
        u64 u = orc_rng_u64(seed, stream, (u64)j, 0), r = u >> 1;
        int mag = 0;
        for (int k = 0; k < 29; k++) mag += (r >= CDT[k]);
        out[j] = (u & 1) ? -(i64)mag : (i64)mag;
    }
}
extern "C" void orc_sample_ternary(const orc_ctx *c, u64 seed, u64 stream, i64 *out) {
    for (int j = 0; j < c->n; j++) {
        u64 v = orc_rng_u64(seed, stream, (u64)j, 0) % 3;
        out[j] = v == 2 ? -1 : (i64)v;
    }
}
static void small_to_limb(const orc_ctx *c, const i64 *v, int qp, u64 *out) { /* signed ints -> eval domain */
    u64 q = c->mods[qp].q;
    for (int j = 0; j < c->n; j++) out[j] = v[j] >= 0 ? (u64)v[j] % q : q - ((u64)(-v[j]) % q);
    orc_ntt_fwd(c, qp, out);
}

/* ------------------------------------------------------------------ keys (a9) */
extern "C" void orc_keygen_secret(const orc_ctx *c, u64 seed, u64 *s_qp) {
    std::vector<i64> s(c->n);
    orc_sample_ternary(c, seed, stream_id(ST_SK, 0, 0), s.data());
    PAR_LIMBS for (int i = 0; i < c->L + c->K; i++) small_to_limb(c, s.data(), i, s_qp + (size_t)i * c->n);
}
/* b = e - a s over the Q limbs */
extern "C" void orc_keygen_public(const orc_ctx *c, u64 seed, const u64 *s_qp, u64 *pk_b, u64 *pk_a) {
    int n = c->n;
    std::vector<i64> e(n);
    orc_sample_gauss(c, seed, stream_id(ST_PK_E, 0, 0), e.data());
    PAR_LIMBS for (int i = 0; i < c->L; i++) {
        u64 q = c->mods[i].q;
        u64 *a = pk_a + (size_t)i * n, *b = pk_b + (size_t)i * n;
        orc_sample_uniform(c, seed, stream_id(ST_PK_A, 0, 0), i, a);
        small_to_limb(c, e.data(), i, b);
        for (int j = 0; j < n; j++) b[j] = submod(b[j], mulmod(a[j], s_qp[(size_t)i * n + j], q), q);
    }
}
/* Hybrid (Han-Ki) switching key old -> new, one (b_j, a_j) pair per digit over all L+K limbs:
 *   b_j = -a_j s_new + e_j + [limb in digit j] * (P mod q_i) * s_old        (zero gadget on P limbs) */
extern "C" void orc_keygen_switch(const orc_ctx *c, u64 seed, u64 key_id, const u64 *s_old, const u64 *s_new, u64 *evk) {
    int n = c->n, L = c->L, K = c->K, F = L + K, beta = orc_beta(c, L);
    for (int j = 0; j < beta; j++) {
        std::vector<i64> e(n);
        orc_sample_gauss(c, seed, stream_id(ST_EVK_E, j, key_id), e.data());
        int s = j * c->alpha, d = std::min(c->alpha, L - s);
        PAR_LIMBS for (int i = 0; i < F; i++) {
            u64 q = c->mods[i].q;
            u64 *b = evk + (((size_t)j * 2 + 0) * F + i) * n, *a = evk + (((size_t)j * 2 + 1) * F + i) * n;
            orc_sample_uniform(c, seed, stream_id(ST_EVK_A, j, key_id), i, a);
            small_to_limb(c, e.data(), i, b);
            u64 pmod = 0;
            if (i >= s && i < s + d) {
                pmod = 1;
                for (int x = 0; x < K; x++) pmod = mulmod(pmod, c->mods[L + x].q % q, q);
            }
            for (int k = 0; k < n; k++) {
                u64 v = submod(b[k], mulmod(a[k], s_new[(size_t)i * n + k], q), q);
                if (pmod) v = addmod(v, mulmod(pmod, s_old[(size_t)i * n + k], q), q);
                b[k] = v;
            }
        }
    }
}
extern "C" void orc_keygen_mult(const orc_ctx *c, u64 seed, const u64 *s_qp, u64 *evk) { /* s^2 -> s, key id 0 */
    int n = c->n, F = c->L + c->K;
    std::vector<u64> s2((size_t)F * n);
    for (int i = 0; i < F; i++)
        for (int j = 0; j < n; j++) s2[(size_t)i * n + j] = mulmod(s_qp[(size_t)i * n + j], s_qp[(size_t)i * n + j], c->mods[i].q);
    orc_keygen_switch(c, seed, 0, s2.data(), s_qp, evk);
}
/* rotation key for Galois element k: switches s -> sigma_{k^-1}(s); the rotation then applies sigma_k
 * to the switched ciphertext (key switch first, automorphism second, as OpenFHE's EvalAutomorphism). */
extern "C" void orc_keygen_rot(const orc_ctx *c, u64 seed, uint32_t k, const u64 *s_qp, u64 *evk) {
    int n = c->n, F = c->L + c->K;
    u64 m = 2 * (u64)n, kinv = 1, base = k, e = m / 2 - 1; /* group order is n = m/2: k^(n-1) = k^-1 */
    while (e) {
        if (e & 1) kinv = kinv * base % m;
        base = base * base % m;
        e >>= 1;
    }
    std::vector<u64> sp((size_t)F * n);
    orc_automorphism_eval(c, (uint32_t)kinv, F, s_qp, sp.data());
    orc_keygen_switch(c, seed, k, s_qp, sp.data(), evk);
}

/* ------------------------------------------------------------------ CKKS encode / decode (a8) */
typedef std::complex<double> cplx;
static void bitrev_array(std::vector<cplx> &v) {
    int sz = (int)v.size(), lg = 0;
    while ((1 << lg) < sz) lg++;
    for (int i = 0; i < sz; i++) {
        int r = (int)bitrev((uint32_t)i, lg);
        if (i < r) std::swap(v[i], v[r]);
    }
}
/* "special" FFT over the orbit of 5 (decode direction): vals_j = m(zeta^(5^j)), zeta = exp(2 pi i / 4 size) */
static void fft_special(std::vector<cplx> &v) {
    int size = (int)v.size(), M = 4 * size;
    std::vector<int> rot(size);
    for (int j = 0, p = 1; j < size; j++, p = (int)((long long)p * 5 % M)) rot[j] = p;
    bitrev_array(v);
    for (int len = 2; len <= size; len <<= 1) {
        int lenh = len >> 1, lenq = len << 2;
        for (int i = 0; i < size; i += len)
            for (int j = 0; j < lenh; j++) {
                double ang = 2.0 * M_PI * (double)(rot[j] % lenq) / (double)lenq;
                cplx w(std::cos(ang), std::sin(ang));
                cplx u = v[i + j], t = v[i + j + lenh] * w;
                v[i + j] = u + t;
                v[i + j + lenh] = u - t;
            }
    }
}
static void fft_special_inv(std::vector<cplx> &v) {
    int size = (int)v.size(), M = 4 * size;
    std::vector<int> rot(size);
    for (int j = 0, p = 1; j < size; j++, p = (int)((long long)p * 5 % M)) rot[j] = p;
    for (int len = size; len >= 2; len >>= 1) {
        int lenh = len >> 1, lenq = len << 2;
        for (int i = 0; i < size; i += len)
            for (int j = 0; j < lenh; j++) {
                double ang = -2.0 * M_PI * (double)(rot[j] % lenq) / (double)lenq;
                cplx w(std::cos(ang), std::sin(ang));
                cplx u = v[i + j] + v[i + j + lenh], t = (v[i + j] - v[i + j + lenh]) * w;
                v[i + j] = u;
                v[i + j + lenh] = t;
            }
    }
    bitrev_array(v);
    for (auto &x : v) x /= (double)size;
}
static i128 round_ld(long double x) {
    bool neg = x < 0;
    if (neg) x = -x;
    i128 r;
    if (x < 4.0e18L)
        r = (i128)llroundl(x);
    else {
        int e;
        long double mant = frexpl(x, &e); /* x = mant * 2^e, mant in [0.5,1) */
        u64 mi = (u64)ldexpl(mant, 64);   /* exact: 64-bit mantissa */
        r = (e >= 64) ? ((i128)mi << (e - 64)) : ((i128)mi >> (64 - e));
    }
    return neg ? -r : r;
}
extern "C" void orc_encode(const orc_ctx *c, const double *vals, int nvals, int slots, double scale, int ell, u64 *pt) {
    int n = c->n, gap = (n / 2) / slots;
    std::vector<cplx> v(slots, cplx(0, 0));
    for (int i = 0; i < nvals && i < slots; i++) v[i] = cplx(vals[i], 0);
    fft_special_inv(v);
    std::vector<i128> coef(n, 0);
    for (int i = 0; i < slots; i++) {
        coef[(size_t)i * gap] = round_ld((long double)v[i].real() * (long double)scale);
        coef[(size_t)n / 2 + (size_t)i * gap] = round_ld((long double)v[i].imag() * (long double)scale);
    }
    PAR_LIMBS for (int l = 0; l < ell; l++) {
        u64 q = c->mods[l].q;
        u64 *dst = pt + (size_t)l * n;
        for (int j = 0; j < n; j++) {
            i128 x = coef[j] % (i128)q;
            if (x < 0) x += q;
            dst[j] = (u64)x;
        }
        orc_ntt_fwd(c, l, dst);
    }
}

/* c0 = b v + e0 + pt, c1 = a v + e1, v ternary, e gaussian (OpenFHE: PKE Encrypt with the public key) */
extern "C" void orc_encrypt(const orc_ctx *c, u64 seed, u64 nonce, int ell, const u64 *pk_b, const u64 *pk_a,
                            const u64 *pt, u64 *c0, u64 *c1) {
    int n = c->n;
    std::vector<i64> v(n), e0(n), e1(n);
    orc_sample_ternary(c, seed, stream_id(ST_ENC_V, 0, nonce), v.data());
    orc_sample_gauss(c, seed, stream_id(ST_ENC_E0, 0, nonce), e0.data());
    orc_sample_gauss(c, seed, stream_id(ST_ENC_E1, 0, nonce), e1.data());
    PAR_LIMBS for (int i = 0; i < ell; i++) {
        u64 q = c->mods[i].q;
        std::vector<u64> V(n), E0(n), E1(n);
        small_to_limb(c, v.data(), i, V.data());
        small_to_limb(c, e0.data(), i, E0.data());
        small_to_limb(c, e1.data(), i, E1.data());
        for (int j = 0; j < n; j++) {
            size_t x = (size_t)i * n + j;
            c0[x] = addmod(addmod(mulmod(pk_b[x], V[j], q), E0[j], q), pt[x], q);
            c1[x] = addmod(mulmod(pk_a[x], V[j], q), E1[j], q);
        }
    }
}
/* m = c0 + c1 s on the first k limbs (k minimal with log2(q_0..q_{k-1}) >= log2(scale) + 30: dropping
 * limbs of a CKKS ciphertext does not change the message while |m| < Q_k/2), Garner mixed-radix CRT with
 * an exact sign decision, divide by scale, special FFT. */
extern "C" void orc_decrypt(const orc_ctx *c, int ell, const u64 *c0, const u64 *c1, const u64 *s_qp, int slots,
                            double scale, double *out_re, double *out_im) {
    int n = c->n, k = 0;
    double bits = 0, need = std::log2(scale) + 30.0;
    while (k < ell && (bits < need || k < 1)) {
        bits += std::log2((double)c->mods[k].q);
        k++;
    }
    std::vector<u64> m((size_t)k * n);
    for (int i = 0; i < k; i++) {
        u64 q = c->mods[i].q;
        u64 *d = &m[(size_t)i * n];
        for (int j = 0; j < n; j++)
            d[j] = addmod(c0[(size_t)i * n + j], mulmod(c1[(size_t)i * n + j], s_qp[(size_t)i * n + j], q), q);
        orc_ntt_inv(c, i, d);
    }
    /* Garner constants: inv[i][j] = q_j^-1 mod q_i for j < i */
    std::vector<std::vector<u64>> inv(k, std::vector<u64>(k, 0));
    for (int i = 0; i < k; i++)
        for (int j = 0; j < i; j++) inv[i][j] = invmod(c->mods[j].q % c->mods[i].q, c->mods[i].q);
    int gap = (n / 2) / slots;
    std::vector<cplx> v(slots);
    auto centred = [&](int j) -> long double {
        std::vector<u64> d(k);
        for (int i = 0; i < k; i++) {
            u64 qi = c->mods[i].q, t = m[(size_t)i * n + j];
            for (int x = 0; x < i; x++) t = mulmod(submod(t, d[x] % qi, qi), inv[i][x], qi);
            d[i] = t;
        }
        bool neg = d[k - 1] > (c->mods[k - 1].q >> 1);
        if (neg)
            for (int i = 0; i < k; i++) d[i] = c->mods[i].q - 1 - d[i]; /* digits of Q-1-x */
        long double r = 0;
        for (int i = k - 1; i >= 0; i--) r = r * (long double)c->mods[i].q + (long double)d[i];
        return neg ? -(r + 1) : r;
    };
    for (int i = 0; i < slots; i++) {
        long double re = centred(i * gap) / (long double)scale, im = centred(n / 2 + i * gap) / (long double)scale;
        v[i] = cplx((double)re, (double)im);
    }
    fft_special(v);
    for (int i = 0; i < slots; i++) {
        out_re[i] = v[i].real();
        if (out_im) out_im[i] = v[i].imag();
    }
}

/* Chebyshev interpolation coefficients on [a,b] (OpenFHE EvalChebyshevCoefficients, used by the reference's
 * utils/generate_cheb_coeffs.cpp:17): deg+1 Chebyshev nodes, c_k = 2/(deg+1) sum f(x_i) cos(pi k (i+1/2)/(deg+1)) */
extern "C" void orc_cheb_coefficients(double (*f)(double, void *), void *user, double a, double b, int degree, double *out) {
    int N = degree + 1;
    double bma = 0.5 * (b - a), bpa = 0.5 * (b + a);
    std::vector<double> fv(N);
    for (int i = 0; i < N; i++) fv[i] = f(std::cos(M_PI * (i + 0.5) / N) * bma + bpa, user);
    for (int k = 0; k < N; k++) {
        double s = 0;
        for (int i = 0; i < N; i++) s += fv[i] * std::cos(M_PI * k * (i + 0.5) / N);
        out[k] = s * 2.0 / N;
    }
}
