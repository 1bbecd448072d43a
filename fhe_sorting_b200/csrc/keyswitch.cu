// keyswitch.cu — hybrid (Han-Ki) key switching kernels: ModUp base conversion, evk inner product, ModDown.
// (OpenFHE entry points replaced: KeySwitchHYBRID::EvalKeySwitchPrecomputeCore / EvalFastKeySwitchCore and
//  DCRTPoly::ApproxSwitchCRTBasis / ApproxModDown, reached from the reference through EvalRotate,
//  EvalFastRotation* (src/rotation.h:224-350) and EvalMult/EvalSquare (src/sign.cpp:23-56).)
//
// Algorithmic bytes at ell active limbs, width = ell + K, beta digits (8n bytes per limb):
//   modup conversion : read ell, write beta*width
//   inner product    : read beta*width (digits) + 2*beta*width (evk), write 2*width      <- evk streaming, HBM bound
//   moddown conv     : read 2K, write 2 ell ; finish: read 2 ell + 2 ell (+ adds), write 2 ell
#include <stdexcept>

#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

constexpr int KS_THREADS = 256;
constexpr int CONV_THREADS = 128;
constexpr int MAXD = 24; // max limbs per digit / max K

// Fast base conversion core (ModUp and ModDown share it).  The DS source residues of one coefficient are multiplied by
// their (Q_j/q_i)^-1 constants, split into 30-bit halves and kept in registers; every target limb is then a DS-term dot
// product with the pre-split (Q_j/q_i mod p_t) table.  With both factors < 2^30 each partial product is < 2^60, so up
// to 16 terms accumulate in plain 64-bit IMAD.WIDE chains (4 per term, no carry handling); the four partial sums are
// recombined into 128 bits and reduced once.  Table layout: [target][DSP] uint2 (lo30, hi30), DSP = DS rounded up to even.
template <int DS>
struct ConvRegs {
    uint32_t lo[DS], hi[DS];
};
template <int DS>
__device__ __forceinline__ u64 conv_dot(const ConvRegs<DS> &x, const uint2 *__restrict__ row, const ModConst &m) {
    constexpr int DSP = DS + (DS & 1);
    u64 rlo = 0, rhi = 0;
#pragma unroll
    for (int i0 = 0; i0 < DS; i0 += 16) {
        u64 ll = 0, lh = 0, hl = 0, hh = 0;
#pragma unroll
        for (int i = i0; i < (i0 + 16 < DSP ? i0 + 16 : DSP); i += 2) {
            const uint4 h = __ldg(reinterpret_cast<const uint4 *>(row + i));
            ll = mad_wide(x.lo[i], h.x, ll), lh = mad_wide(x.lo[i], h.y, lh);
            hl = mad_wide(x.hi[i], h.x, hl), hh = mad_wide(x.hi[i], h.y, hh);
            if (i + 1 < DS) {
                ll = mad_wide(x.lo[i + 1], h.z, ll), lh = mad_wide(x.lo[i + 1], h.w, lh);
                hl = mad_wide(x.hi[i + 1], h.z, hl), hh = mad_wide(x.hi[i + 1], h.w, hh);
            }
        }
        acc_split30(rlo, rhi, ll, lh, hl, hh);
    }
    return reduce128(rhi, rlo, m);
}

// ModUp: one thread = one coefficient of one digit; own-digit limbs are copied from the evaluation-domain input
template <int DS>
__global__ void __launch_bounds__(CONV_THREADS) ks_modup(DeviceTables T, DevicePlan P, const u64 *__restrict__ coef,
                                                         const u64 *__restrict__ c_eval, u64 *__restrict__ digits) {
    const int j = blockIdx.y, k = blockIdx.x * CONV_THREADS + threadIdx.x, n = T.n;
    if (k >= n) return;
    constexpr int DSP = DS + (DS & 1);
    const int s = P.dstart[j], W = P.width;
    const u64 *hat_inv = P.hat_inv[j], *hat_inv_s = P.hat_inv_s[j];
    ConvRegs<DS> x;
#pragma unroll
    for (int i = 0; i < DS; i++) {
        u64 v = shoup_mul(coef[(size_t)(s + i) * n + k], __ldg(hat_inv + i), __ldg(hat_inv_s + i), T.mc[s + i].q);
        x.lo[i] = (uint32_t)v & 0x3fffffffu, x.hi[i] = (uint32_t)(v >> 30);
    }
    u64 *D = digits + (size_t)j * W * n + k;
    const uint2 *tab = P.hat2[j];
    for (int t = 0; t < W; t++) {
        if (t >= s && t < s + DS) {
            D[(size_t)t * n] = c_eval[(size_t)t * n + k];
            continue;
        }
        D[(size_t)t * n] = conv_dot<DS>(x, tab + (size_t)t * DSP, T.mc[P.qp[t]]);
    }
}

// ModDown: P -> Q base conversion of the (coefficient-domain) P limbs of accumulator blockIdx.y
template <int DS>
__global__ void __launch_bounds__(CONV_THREADS) ks_moddown_conv(DeviceTables T, DevicePlan P, const u64 *__restrict__ acc,
                                                                u64 *__restrict__ conv) {
    const int b = blockIdx.y, k = blockIdx.x * CONV_THREADS + threadIdx.x, n = T.n;
    if (k >= n) return;
    constexpr int DSP = DS + (DS & 1);
    const int ell = P.ell, W = P.width;
    const u64 *src = acc + ((size_t)b * W + ell) * n + k;
    ConvRegs<DS> x;
#pragma unroll
    for (int i = 0; i < DS; i++) {
        u64 v = shoup_mul(src[(size_t)i * n], __ldg(P.p_hat_inv + i), __ldg(P.p_hat_inv_s + i), T.mc[P.qp[ell + i]].q);
        x.lo[i] = (uint32_t)v & 0x3fffffffu, x.hi[i] = (uint32_t)(v >> 30);
    }
    u64 *dst = conv + (size_t)b * ell * n + k;
    for (int t = 0; t < ell; t++) dst[(size_t)t * n] = conv_dot<DS>(x, P.p_hat2 + (size_t)t * DSP, T.mc[t]);
}

// evk inner product: one thread = one coefficient of one level limb; both key halves accumulated over the digits as
// 30-bit split dot products (beta <= 8 terms).  HBM bound: streams beta digits + 2 beta key limbs per level limb.
__global__ void __launch_bounds__(KS_THREADS) ks_inner(DeviceTables T, DevicePlan P, const u64 *__restrict__ digits,
                                                       const u64 *__restrict__ evk, u64 *__restrict__ acc) {
    const int t = blockIdx.y, k = blockIdx.x * KS_THREADS + threadIdx.x, n = T.n;
    if (k >= n) return;
    const int W = P.width, qp = P.qp[t], F = T.F;
    const ModConst m = T.mc[qp];
    const u64 *dp = digits + (size_t)t * n + k, *kp = evk + (size_t)qp * n + k;
    if (fp_modulus_ok(m)) { // 40..42-bit scaling primes: 6 DP instructions per exact product instead of ~40 integer ones
        const double2 qd = T.qd[qp];
        double fb = 0.0, fa = 0.0;
#pragma unroll 3
        for (int j = 0; j < P.beta; j++) {
            const double dv = fp_from_u64(dp[(size_t)j * W * n]);
            const double kb = fp_from_u64(__ldcs(kp + (size_t)(2 * j) * F * n)), ka = fp_from_u64(__ldcs(kp + (size_t)(2 * j + 1) * F * n));
            fb = __dadd_rn(fb, fp_mulmod(dv, kb, qd.y, qd.x));
            fa = __dadd_rn(fa, fp_mulmod(dv, ka, qd.y, qd.x));
        }
        acc[((size_t)0 * W + t) * n + k] = fp_to_u64(fb, qd.x, qd.y);
        acc[((size_t)1 * W + t) * n + k] = fp_to_u64(fa, qd.x, qd.y);
        return;
    }
    u64 sb[4] = {0, 0, 0, 0}, sa[4] = {0, 0, 0, 0}; // ll, lh, hl, hh
#pragma unroll 3
    for (int j = 0; j < P.beta; j++) {
        const u64 dv = dp[(size_t)j * W * n];
        const u64 kb = __ldcs(kp + (size_t)(2 * j) * F * n), ka = __ldcs(kp + (size_t)(2 * j + 1) * F * n);
        const uint32_t dl = (uint32_t)dv & 0x3fffffffu, dh = (uint32_t)(dv >> 30);
        const uint32_t bl = (uint32_t)kb & 0x3fffffffu, bh = (uint32_t)(kb >> 30);
        const uint32_t al = (uint32_t)ka & 0x3fffffffu, ah = (uint32_t)(ka >> 30);
        sb[0] = mad_wide(dl, bl, sb[0]), sb[1] = mad_wide(dl, bh, sb[1]), sb[2] = mad_wide(dh, bl, sb[2]), sb[3] = mad_wide(dh, bh, sb[3]);
        sa[0] = mad_wide(dl, al, sa[0]), sa[1] = mad_wide(dl, ah, sa[1]), sa[2] = mad_wide(dh, al, sa[2]), sa[3] = mad_wide(dh, ah, sa[3]);
    }
    u64 lo = 0, hi = 0;
    acc_split30(lo, hi, sb[0], sb[1], sb[2], sb[3]);
    acc[((size_t)0 * W + t) * n + k] = reduce128(hi, lo, m);
    lo = 0, hi = 0;
    acc_split30(lo, hi, sa[0], sa[1], sa[2], sa[3]);
    acc[((size_t)1 * W + t) * n + k] = reduce128(hi, lo, m);
}

// out_b[i][p] = ((acc_b[i] - conv_b[i]) * P^-1 + add_b[i]) [src(p)], src = Galois gather (identity when g == 1)
__global__ void __launch_bounds__(KS_THREADS) ks_moddown_finish(DeviceTables T, DevicePlan P, const u64 *acc, const u64 *conv,
                                                                const u64 *add0, const u64 *add1, u64 *out0, u64 *out1,
                                                                uint32_t g) {
    const int i = blockIdx.y, b = blockIdx.z, n = T.n;
    const uint32_t p = blockIdx.x * KS_THREADS + threadIdx.x;
    if (p >= (uint32_t)n) return;
    uint32_t src = p;
    if (g != 1u) {
        const int sh = 32 - T.logn;
        uint32_t j = __brev(p) >> sh;
        uint32_t idx = (((2u * j + 1u) * g) & (2u * n - 1u)) >> 1;
        src = __brev(idx) >> sh;
    }
    const u64 q = T.mc[i].q;
    u64 a = acc[((size_t)b * P.width + i) * n + src], c = conv[((size_t)b * P.ell + i) * n + src];
    u64 r = shoup_mul(sub_mod(a, c, q), __ldg(P.p_inv + i), __ldg(P.p_inv_s + i), q);
    const u64 *add = b ? add1 : add0;
    if (add) r = add_mod(r, add[(size_t)i * n + src], q);
    (b ? out1 : out0)[(size_t)i * n + p] = r;
}

// S_b[t][p] += pt[t][p] * v_b[t][src(p)] (see kernels.h: k_dh_accumulate); one thread = one coefficient of one level limb
__global__ void __launch_bounds__(KS_THREADS) ks_dh_accumulate(DeviceTables T, DevicePlan P, const u64 *__restrict__ acc,
                                                               const u64 *__restrict__ cP, const u64 *__restrict__ pt,
                                                               u64 *__restrict__ S, uint32_t g) {
    const int t = blockIdx.y, n = T.n, W = P.width, ell = P.ell;
    const uint32_t p = blockIdx.x * KS_THREADS + threadIdx.x;
    if (p >= (uint32_t)n) return;
    uint32_t src = p;
    if (g != 1u) {
        const int sh = 32 - T.logn;
        uint32_t j = __brev(p) >> sh;
        uint32_t idx = (((2u * j + 1u) * g) & (2u * n - 1u)) >> 1;
        src = __brev(idx) >> sh;
    }
    if (!acc && t >= ell) return; // the unrotated term lives on the Q limbs only
    const ModConst m = T.mc[P.qp[t]];
    const u64 pv = pt[(size_t)t * n + p];
#pragma unroll
    for (int b = 0; b < 2; b++) {
        u64 v = acc ? acc[((size_t)b * W + t) * n + src] : 0;
        if (t < ell && (b == 0 || !acc)) v = add_mod(v, cP[((size_t)b * ell + t) * n + src], m.q);
        const size_t o = ((size_t)b * W + t) * n + p;
        S[o] = add_mod(S[o], mul_mod(v, pv, m), m.q);
    }
}

#define DS_SWITCH(ds, CALL)                                                                                              \
    switch (ds) {                                                                                                        \
        case 1: CALL(1); break; case 2: CALL(2); break; case 3: CALL(3); break; case 4: CALL(4); break;                 \
        case 5: CALL(5); break; case 6: CALL(6); break; case 7: CALL(7); break; case 8: CALL(8); break;                 \
        case 9: CALL(9); break; case 10: CALL(10); break; case 11: CALL(11); break; case 12: CALL(12); break;           \
        case 13: CALL(13); break; case 14: CALL(14); break; case 15: CALL(15); break; case 16: CALL(16); break;         \
        case 17: CALL(17); break; case 18: CALL(18); break; case 19: CALL(19); break; case 20: CALL(20); break;         \
        case 21: CALL(21); break; case 22: CALL(22); break; case 23: CALL(23); break; case 24: CALL(24); break;         \
        default: throw std::invalid_argument("fheb200: digit size / K beyond 24 limbs unsupported");                     \
    }

void k_modup(const DeviceTables &T, const DevicePlan &P, const u64 *coef, const u64 *c_eval, u64 *digits, cudaStream_t st) {
    // digits can differ in size (the last one may be shorter): one launch per distinct size
    const int gx = (T.n + CONV_THREADS - 1) / CONV_THREADS;
    for (int j = 0; j < P.beta; j++) {
        if (j > 0 && P.dsize[j] == P.dsize[j - 1]) continue;
        int cnt = 1;
        while (j + cnt < P.beta && P.dsize[j + cnt] == P.dsize[j]) cnt++;
        DevicePlan Q = P; // shift so that blockIdx.y = 0 is digit j
        for (int u = 0; u < cnt; u++) {
            Q.dstart[u] = P.dstart[j + u], Q.dsize[u] = P.dsize[j + u];
            Q.hat_inv[u] = P.hat_inv[j + u], Q.hat_inv_s[u] = P.hat_inv_s[j + u], Q.hat2[u] = P.hat2[j + u];
        }
        u64 *dj = digits + (size_t)j * P.width * T.n;
#define CALL(D) g_launches++, ks_modup<D><<<dim3(gx, cnt), CONV_THREADS, 0, st>>>(T, Q, coef, c_eval, dj)
        DS_SWITCH(P.dsize[j], CALL)
#undef CALL
    }
}
void k_inner_product(const DeviceTables &T, const DevicePlan &P, const u64 *digits, const u64 *evk, u64 *acc, cudaStream_t st) {
    g_launches++, ks_inner<<<dim3((T.n + KS_THREADS - 1) / KS_THREADS, P.width), KS_THREADS, 0, st>>>(T, P, digits, evk, acc);
}
void k_moddown_conv(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *conv, cudaStream_t st) {
    const int gx = (T.n + CONV_THREADS - 1) / CONV_THREADS;
#define CALL(D) g_launches++, ks_moddown_conv<D><<<dim3(gx, 2), CONV_THREADS, 0, st>>>(T, P, acc, conv)
    DS_SWITCH(P.K, CALL)
#undef CALL
}
void k_moddown_finish(const DeviceTables &T, const DevicePlan &P, const u64 *acc, const u64 *conv, const u64 *add0,
                      const u64 *add1, u64 *out0, u64 *out1, uint32_t galois, cudaStream_t st) {
    g_launches++, ks_moddown_finish<<<dim3((T.n + KS_THREADS - 1) / KS_THREADS, P.ell, 2), KS_THREADS, 0, st>>>(T, P, acc, conv, add0, add1,
                                                                                                  out0, out1, galois);
}
void k_dh_accumulate(const DeviceTables &T, const DevicePlan &P, const u64 *acc, const u64 *cP, const u64 *pt, u64 *S,
                     uint32_t galois, cudaStream_t st) {
    g_launches++, ks_dh_accumulate<<<dim3((T.n + KS_THREADS - 1) / KS_THREADS, P.width), KS_THREADS, 0, st>>>(T, P, acc, cP, pt, S, galois);
}

} // namespace fheb200
