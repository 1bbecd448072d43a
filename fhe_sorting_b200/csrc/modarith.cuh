// modarith.cuh — 64-bit modular arithmetic for sm_100a integer pipes (IMAD.WIDE based 64x64 products).
// All moduli satisfy q < 2^61, so lazy values up to 4q fit in a word (Harvey butterflies).
#pragma once
#include <cstdint>
#include "params.h"

namespace fheb200 {

__device__ __forceinline__ u64 mulhi64(u64 a, u64 b) { return __umul64hi(a, b); }
// c + a*b with 32-bit factors and a 64-bit accumulator: exactly one IMAD.WIDE.U32 (the C++ form `c += (u64)a * b`
// makes nvcc 12.9 append a dead carry add per term)
__device__ __forceinline__ u64 mad_wide(uint32_t a, uint32_t b, u64 c) {
    u64 r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
    return r;
}

// a*w mod q with precomputed ws = floor(w 2^64 / q). Any 64-bit a; result in [0, 2q).
__device__ __forceinline__ u64 shoup_lazy(u64 a, u64 w, u64 ws, u64 q) {
    u64 h = mulhi64(a, ws);
    return a * w - h * q;
}
// same with nq = 2^64 - q: a*w + h*nq keeps the whole expression on the IMAD accumulate path (no 64-bit subtract)
__device__ __forceinline__ u64 shoup_lazy_n(u64 a, u64 w, u64 ws, u64 nq) { return a * w + mulhi64(a, ws) * nq; }
// fully reduced variant
__device__ __forceinline__ u64 shoup_mul(u64 a, u64 w, u64 ws, u64 q) {
    u64 r = shoup_lazy(a, w, ws, q);
    return r >= q ? r - q : r;
}
__device__ __forceinline__ u64 add_mod(u64 a, u64 b, u64 q) {
    u64 s = a + b;
    return s >= q ? s - q : s;
}
__device__ __forceinline__ u64 sub_mod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
__device__ __forceinline__ u64 neg_mod(u64 a, u64 q) { return a ? q - a : 0; }

// single word x (any 64-bit value) mod q
__device__ __forceinline__ u64 reduce64(u64 x, const ModConst &m) {
    u64 r = x - mulhi64(x, m.r64) * m.q; // quotient estimate is at most one short: r in [0, 2q)
    return r >= m.q ? r - m.q : r;
}
// (hi:lo) mod q for any 128-bit value: hi is folded through 2^64 mod q (Shoup), lo by single-word Barrett, both lazy
__device__ __forceinline__ u64 reduce128(u64 hi, u64 lo, const ModConst &m) {
    const u64 nq = 0 - m.q;
    u64 r = shoup_lazy_n(hi, m.c64, m.c64s, nq) + (lo + mulhi64(lo, m.r64) * nq); // [0,2q) + [0,2q)
    const u64 q2 = m.q << 1;
    r = r >= q2 ? r - q2 : r;
    return r >= m.q ? r - m.q : r;
}
// (lo, hi) += ll + (lh + hl) * 2^30 + hh * 2^60   (recombination of a 30-bit split dot product)
__device__ __forceinline__ void acc_split30(u64 &lo, u64 &hi, u64 ll, u64 lh, u64 hl, u64 hh) {
    u64 c;
    lo += ll, hi += (lo < ll);
    c = lh << 30, lo += c, hi += (lh >> 34) + (lo < c);
    c = hl << 30, lo += c, hi += (hl >> 34) + (lo < c);
    c = hh << 60, lo += c, hi += (hh >> 4) + (lo < c);
}
// Barrett product of two reduced operands (a, b < q)
__device__ __forceinline__ u64 mul_mod(u64 a, u64 b, const ModConst &m) {
    u64 lo = a * b, hi = mulhi64(a, b);
    u64 z1 = (hi << (65 - m.bits)) | (lo >> (m.bits - 1)); // z >> (bits-1), fits: z < 2^(2 bits)
    u64 r = lo - mulhi64(z1, m.mu) * m.q;
    r = r >= m.q ? r - m.q : r;
    return r >= m.q ? r - m.q : r;
}
// 128-bit accumulator for lazy inner products (sum of < 2^4 products of 61-bit values)
struct Acc128 {
    u64 lo, hi;
    __device__ __forceinline__ Acc128() : lo(0), hi(0) {}
    __device__ __forceinline__ void mac(u64 a, u64 b) {
        u64 pl = a * b, ph = mulhi64(a, b);
        lo += pl;
        hi += ph + (lo < pl);
    }
};

// centred lift of x mod q_from (half = q_from >> 1) into modulus q_to
__device__ __forceinline__ u64 switch_modulus(u64 x, u64 q_from, u64 half, const ModConst &to) {
    if (x > half) {
        u64 d = reduce64(q_from - x, to);
        return d ? to.q - d : 0;
    }
    return reduce64(x, to);
}

// ---- exact modular arithmetic in double precision for moduli of at most 42 bits (see ntt.cu, "FP64 path"): values are
// exact signed integers held in doubles, a product is 6 DP instructions
constexpr double FP_MAGIC = 6755399441055744.0; // 1.5 * 2^52: (x + M) - M rounds |x| < 2^51 to the nearest integer
constexpr double FP_2P52 = 4503599627370496.0;
__device__ __forceinline__ double fp_rint(double x) { return __dadd_rn(__dadd_rn(x, FP_MAGIC), -FP_MAGIC); }
__device__ __forceinline__ double fp_mulmod(double y, double w, double qinv, double q) {
    const double h = __dmul_rn(y, w);
    const double l = __fma_rn(y, w, -h);
    const double c = __dadd_rn(__fma_rn(h, qinv, FP_MAGIC), -FP_MAGIC); // rint(h / q), off by at most one from rint(y w / q)
    const double d = __fma_rn(-c, q, h);
    return __dadd_rn(d, l);
}
__device__ __forceinline__ double fp_center(double v, double q, double qinv) { // v - q*rint(v/q), |result| <= q/2 + 1
    return __fma_rn(-fp_rint(__dmul_rn(v, qinv)), q, v);
}
__device__ __forceinline__ double fp_from_u64(u64 x) { // exact for x < 2^52
    return __dadd_rn(__longlong_as_double((long long)(x | 0x4330000000000000ull)), -FP_2P52);
}
__device__ __forceinline__ u64 fp_to_u64(double v, double q, double qinv) { // any |v| < 2^51 -> [0, q)
    double r = fp_center(v, q, qinv);
    r = r < 0.0 ? __dadd_rn(r, q) : r;
    r = r >= q ? __dadd_rn(r, -q) : r;
    return (u64)__double_as_longlong(__dadd_rn(r, FP_2P52)) & 0x000fffffffffffffull;
}
__device__ __forceinline__ bool fp_modulus_ok(const ModConst &m) { return m.bits <= 42; }

// ---- randomness ---------------------------------------------------------------------------------------------------
// Two generators behind one interface (RngState, params.h):
//   secure = 1 (default): ChaCha20 (RFC 8439 block function, 20 rounds) keyed with 256 bits of OS entropy; block counter =
//     element index, nonce = stream id (tag | digit | key id / encryption nonce).  Secret material (ternary secret,
//     Gaussian errors, encryption randomness) and public material (the uniform `a` polynomials of pk / evk) use
//     independent keys, so nothing public shares generator state with the secret key.
//   secure = 0: the counter-based mix64 generator below, seeded with 64 bits.  NOT cryptographic (the finaliser is
//     invertible): TEST-ONLY, used so that the CPU oracle can draw the same keys and noise (DESIGN.md "Randomness").
__host__ __device__ __forceinline__ u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ u64 rng_u64(u64 seed, u64 stream, u64 idx, u64 word) {
    u64 a = mix64(seed ^ (0x9E3779B97F4A7C15ULL * (stream + 1)));
    u64 b = mix64(a + idx);
    return mix64(b + 0xD1B54A32D192ED03ULL * (word + 1));
}
__host__ __device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
#define FHE_CHACHA_QR(a, b, c, d) \
    a += b, d ^= a, d = rotl32(d, 16), c += d, b ^= c, b = rotl32(b, 12), a += b, d ^= a, d = rotl32(d, 8), c += d, b ^= c, b = rotl32(b, 7)
// first 128 bits of the ChaCha20 block (key, counter = idx, nonce = stream)
__host__ __device__ __forceinline__ void chacha20_128(const uint32_t (&key)[8], u64 stream, u64 idx, u64 &w0, u64 &w1) {
    uint32_t x[16], in[16];
    in[0] = 0x61707865u, in[1] = 0x3320646eu, in[2] = 0x79622d32u, in[3] = 0x6b206574u;
#pragma unroll
    for (int i = 0; i < 8; i++) in[4 + i] = key[i];
    in[12] = (uint32_t)idx, in[13] = (uint32_t)(idx >> 32), in[14] = (uint32_t)stream, in[15] = (uint32_t)(stream >> 32);
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = in[i];
#pragma unroll
    for (int r = 0; r < 10; r++) {
        FHE_CHACHA_QR(x[0], x[4], x[8], x[12]);
        FHE_CHACHA_QR(x[1], x[5], x[9], x[13]);
        FHE_CHACHA_QR(x[2], x[6], x[10], x[14]);
        FHE_CHACHA_QR(x[3], x[7], x[11], x[15]);
        FHE_CHACHA_QR(x[0], x[5], x[10], x[15]);
        FHE_CHACHA_QR(x[1], x[6], x[11], x[12]);
        FHE_CHACHA_QR(x[2], x[7], x[8], x[13]);
        FHE_CHACHA_QR(x[3], x[4], x[9], x[14]);
    }
    w0 = (u64)(x[0] + in[0]) | ((u64)(x[1] + in[1]) << 32);
    w1 = (u64)(x[2] + in[2]) | ((u64)(x[3] + in[3]) << 32);
}
// 128 random bits for element idx of a stream
__host__ __device__ __forceinline__ void rng_draw128(const RngState &R, u64 stream, u64 idx, u64 &w0, u64 &w1) {
    if (R.secure) chacha20_128(R.key, stream, idx, w0, w1);
    else w0 = rng_u64(R.seed, stream, idx, 0), w1 = rng_u64(R.seed, stream, idx, 1);
}

} // namespace fheb200
