// sampling.cu — device samplers for key generation / encryption, plaintext expansion, decrypt core.
// Counter-based generator and the 63-bit CDT of the folded discrete Gaussian (sigma = 3.19) are specified in
// DESIGN.md "Randomness"; the CPU oracle implements the same spec so keys are reproducible bit for bit.
#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

constexpr int SP_THREADS = 256;

__constant__ u64 c_cdt[29] = {
    0x1001f9a1b2ca9468ULL, 0x2e7cf3ef07836cb5ULL, 0x48ca7e85d834a579ULL, 0x5d5d51778f760880ULL,
    0x6bf35598550421b0ULL, 0x7552dc90807bbed5ULL, 0x7ac8820561a4c974ULL, 0x7daa6596524b7ffdULL,
    0x7f0b8114bba4ecb9ULL, 0x7fa4a9fc5ea574e8ULL, 0x7fe0e1077aef3919ULL, 0x7ff6563f810f15d0ULL,
    0x7ffd4490999bc594ULL, 0x7fff4c0804cd386dULL, 0x7fffd5e18e7fda18ULL, 0x7ffff709c679c2edULL,
    0x7ffffe445c790f6bULL, 0x7fffffb20f7aa54aULL, 0x7ffffff3903d235dULL, 0x7ffffffe32b06d1dULL,
    0x7fffffffc35125ccULL, 0x7ffffffff8c11764ULL, 0x7fffffffff36fe33ULL, 0x7fffffffffec3bf7ULL,
    0x7ffffffffffe3c8bULL, 0x7fffffffffffdb75ULL, 0x7ffffffffffffd51ULL, 0x7fffffffffffffd2ULL,
    0x7ffffffffffffffdULL}; // tools/gen_cdt.py

// uniform mod q from 128 random bits; element index = qp*n + j
__global__ void __launch_bounds__(SP_THREADS) sp_uniform(DeviceTables T, RngState R, u64 stream, int qp0, u64 *out) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x, qp = qp0 + blockIdx.y;
    if (j >= T.n) return;
    const u64 idx = (u64)qp * T.n + j;
    u64 hi, lo;
    rng_draw128(R, stream, idx, hi, lo);
    out[(size_t)blockIdx.y * T.n + j] = reduce128(hi, lo, T.mc[qp]);
}
__global__ void __launch_bounds__(SP_THREADS) sp_small(DeviceTables T, int kind, RngState R, u64 stream, int *out) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x;
    if (j >= T.n) return;
    u64 u, unused;
    rng_draw128(R, stream, (u64)j, u, unused);
    int v;
    if (kind == SMALL_TERNARY) {
        u64 r = u % 3;
        v = r == 2 ? -1 : (int)r;
    } else {
        u64 r = u >> 1;
        int mag = 0;
#pragma unroll
        for (int k = 0; k < 29; k++) mag += (r >= c_cdt[k]);
        v = (u & 1) ? -mag : mag;
    }
    out[j] = v;
}
__global__ void __launch_bounds__(SP_THREADS) sp_expand_small(DeviceTables T, const int *small, u64 *out, const uint32_t *desc) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x;
    if (j >= T.n) return;
    const uint32_t d = desc[blockIdx.y];
    const u64 q = T.mc[d & 0xff].q;
    int v = small[j];
    out[(size_t)(d >> 8) * T.n + j] = v >= 0 ? (u64)v : q - (u64)(-v);
}
// signed 128-bit coefficient (two's complement lo,hi) -> residue
__global__ void __launch_bounds__(SP_THREADS) sp_expand_i128(DeviceTables T, const u64 *coef, int ncoef, int gap, u64 *out) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x, limb = blockIdx.y;
    if (j >= T.n) return;
    u64 r = 0;
    if (j % gap == 0 && j / gap < ncoef) {
        const ModConst m = T.mc[limb];
        u64 lo = coef[2 * (size_t)(j / gap)], hi = coef[2 * (size_t)(j / gap) + 1];
        bool neg = (long long)hi < 0;
        if (neg) { // magnitude of the negative value
            lo = ~lo, hi = ~hi;
            lo += 1;
            hi += (lo == 0);
        }
        r = reduce128(hi, lo, m);
        if (neg) r = neg_mod(r, m.q);
    }
    out[(size_t)limb * T.n + j] = r;
}
__global__ void __launch_bounds__(SP_THREADS) sp_evk_finish(DeviceTables T, const u64 *a, const u64 *e, const u64 *s_new,
                                                            const u64 *s_old, LimbScalars pmod, u64 *b) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x, i = blockIdx.y;
    if (j >= T.n) return;
    const ModConst m = T.mc[i];
    const size_t x = (size_t)i * T.n + j;
    u64 v = sub_mod(e[x], mul_mod(a[x], s_new[x], m), m.q);
    if (pmod.v[i]) v = add_mod(v, shoup_mul(s_old[x], pmod.v[i], pmod.s[i], m.q), m.q);
    b[x] = v;
}
__global__ void __launch_bounds__(SP_THREADS) sp_encrypt_finish(DeviceTables T, const u64 *pkb, const u64 *pka, const u64 *v,
                                                                const u64 *e0, const u64 *e1, const u64 *pt, u64 *c0, u64 *c1) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x, i = blockIdx.y;
    if (j >= T.n) return;
    const ModConst m = T.mc[i];
    const size_t x = (size_t)i * T.n + j;
    u64 V = v[x];
    c0[x] = add_mod(add_mod(mul_mod(pkb[x], V, m), e0[x], m.q), pt[x], m.q);
    c1[x] = add_mod(mul_mod(pka[x], V, m), e1[x], m.q);
}
__global__ void __launch_bounds__(SP_THREADS) sp_decrypt_core(DeviceTables T, const u64 *c0, const u64 *c1, const u64 *s, u64 *mo) {
    const int j = blockIdx.x * SP_THREADS + threadIdx.x, i = blockIdx.y;
    if (j >= T.n) return;
    const ModConst m = T.mc[i];
    const size_t x = (size_t)i * T.n + j;
    mo[x] = add_mod(c0[x], mul_mod(c1[x], s[x], m), m.q);
}

static inline dim3 g1(const DeviceTables &T, int y) { return dim3((T.n + SP_THREADS - 1) / SP_THREADS, y); }

void k_sample_uniform(const DeviceTables &T, const RngState &R, u64 stream, int qp0, int limbs, u64 *out, cudaStream_t st) {
    g_launches++, sp_uniform<<<g1(T, limbs), SP_THREADS, 0, st>>>(T, R, stream, qp0, out);
}
void k_sample_small(const DeviceTables &T, SmallKind kind, const RngState &R, u64 stream, int *out, cudaStream_t st) {
    g_launches++, sp_small<<<g1(T, 1), SP_THREADS, 0, st>>>(T, (int)kind, R, stream, out);
}
void k_expand_small(const DeviceTables &T, const int *small, u64 *out, const uint32_t *d_desc, int count, cudaStream_t st) {
    g_launches++, sp_expand_small<<<g1(T, count), SP_THREADS, 0, st>>>(T, small, out, d_desc);
}
void k_expand_i128(const DeviceTables &T, const u64 *coef, int ncoef, int gap, int ell, u64 *out, cudaStream_t st) {
    g_launches++, sp_expand_i128<<<g1(T, ell), SP_THREADS, 0, st>>>(T, coef, ncoef, gap, out);
}
void k_evk_finish(const DeviceTables &T, int F, const u64 *a, const u64 *e, const u64 *s_new, const u64 *s_old,
                  const LimbScalars &pmod, u64 *b, cudaStream_t st) {
    g_launches++, sp_evk_finish<<<g1(T, F), SP_THREADS, 0, st>>>(T, a, e, s_new, s_old, pmod, b);
}
void k_encrypt_finish(const DeviceTables &T, int ell, const u64 *pkb, const u64 *pka, const u64 *v, const u64 *e0,
                      const u64 *e1, const u64 *pt, u64 *c0, u64 *c1, cudaStream_t st) {
    g_launches++, sp_encrypt_finish<<<g1(T, ell), SP_THREADS, 0, st>>>(T, pkb, pka, v, e0, e1, pt, c0, c1);
}
void k_decrypt_core(const DeviceTables &T, int k, const u64 *c0, const u64 *c1, const u64 *s, u64 *m, cudaStream_t st) {
    g_launches++, sp_decrypt_core<<<g1(T, k), SP_THREADS, 0, st>>>(T, c0, c1, s, m);
}

} // namespace fheb200
