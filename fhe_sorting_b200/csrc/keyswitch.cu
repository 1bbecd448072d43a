// keyswitch.cu — hybrid (Han-Ki) key switching kernels: ModUp base conversion, evk inner product, ModDown.
// (OpenFHE entry points replaced: KeySwitchHYBRID::EvalKeySwitchPrecomputeCore / EvalFastKeySwitchCore and
//  DCRTPoly::ApproxSwitchCRTBasis / ApproxModDown, reached from the reference through EvalRotate,
//  EvalFastRotation* (src/rotation.h:224-350) and EvalMult/EvalSquare (src/sign.cpp:23-56).)
//
// Algorithmic bytes at ell active limbs, width = ell + K, beta digits (8n bytes per limb):
//   modup conversion : read ell, write beta*width
//   inner product    : read beta*width (digits) + 2*beta*width (evk), write 2*width      <- evk streaming, HBM bound
//   moddown conv     : read 2K, write 2 ell ; finish: read 2 ell + 2 ell (+ adds), write 2 ell
#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

constexpr int KS_THREADS = 256;
constexpr int MAXD = 24; // max limbs per digit / max K held in registers

// one thread = one coefficient of one digit
__global__ void __launch_bounds__(KS_THREADS) ks_modup(DeviceTables T, DevicePlan P, const u64 *coef, const u64 *c_eval,
                                                       u64 *digits) {
    const int j = blockIdx.y, k = blockIdx.x * KS_THREADS + threadIdx.x, n = T.n;
    if (k >= n) return;
    const int s = P.dstart[j], d = P.dsize[j], W = P.width;
    const u64 *hat_inv = P.hat_inv[j], *hat_inv_s = P.hat_inv_s[j], *hat = P.hat[j];
    u64 x[MAXD];
#pragma unroll
    for (int i = 0; i < MAXD; i++) {
        x[i] = 0;
        if (i < d) {
            u64 v = coef[(size_t)(s + i) * n + k];
            x[i] = shoup_mul(v, __ldg(hat_inv + i), __ldg(hat_inv_s + i), T.mc[s + i].q);
        }
    }
    u64 *D = digits + (size_t)j * W * n;
    for (int t = 0; t < W; t++) {
        if (t >= s && t < s + d) {
            D[(size_t)t * n + k] = c_eval[(size_t)t * n + k];
            continue;
        }
        Acc128 acc;
#pragma unroll
        for (int i = 0; i < MAXD; i++)
            if (i < d) acc.mac(x[i], __ldg(hat + (size_t)i * W + t));
        D[(size_t)t * n + k] = reduce128(acc.hi, acc.lo, T.mc[P.qp[t]]);
    }
}

// one thread = two coefficients of one level limb; both key halves accumulated lazily in 128 bits
__global__ void __launch_bounds__(KS_THREADS) ks_inner(DeviceTables T, DevicePlan P, const u64 *digits, const u64 *evk,
                                                       u64 *acc) {
    const int t = blockIdx.y, k = (blockIdx.x * KS_THREADS + threadIdx.x) * 2, n = T.n;
    if (k >= n) return;
    const int W = P.width, qp = P.qp[t], F = T.F;
    const ModConst m = T.mc[qp];
    Acc128 b0, b1, a0, a1;
    for (int j = 0; j < P.beta; j++) {
        ulonglong2 dv = *reinterpret_cast<const ulonglong2 *>(digits + ((size_t)j * W + t) * n + k);
        ulonglong2 kb = *reinterpret_cast<const ulonglong2 *>(evk + (((size_t)j * 2 + 0) * F + qp) * n + k);
        ulonglong2 ka = *reinterpret_cast<const ulonglong2 *>(evk + (((size_t)j * 2 + 1) * F + qp) * n + k);
        b0.mac(dv.x, kb.x), b1.mac(dv.y, kb.y);
        a0.mac(dv.x, ka.x), a1.mac(dv.y, ka.y);
    }
    ulonglong2 rb, ra;
    rb.x = reduce128(b0.hi, b0.lo, m), rb.y = reduce128(b1.hi, b1.lo, m);
    ra.x = reduce128(a0.hi, a0.lo, m), ra.y = reduce128(a1.hi, a1.lo, m);
    *reinterpret_cast<ulonglong2 *>(acc + ((size_t)0 * W + t) * n + k) = rb;
    *reinterpret_cast<ulonglong2 *>(acc + ((size_t)1 * W + t) * n + k) = ra;
}

// P -> Q base conversion of the (coefficient-domain) P limbs of accumulator blockIdx.y
__global__ void __launch_bounds__(KS_THREADS) ks_moddown_conv(DeviceTables T, DevicePlan P, const u64 *acc, u64 *conv) {
    const int b = blockIdx.y, k = blockIdx.x * KS_THREADS + threadIdx.x, n = T.n;
    if (k >= n) return;
    const int ell = P.ell, K = P.K, W = P.width;
    const u64 *src = acc + ((size_t)b * W + ell) * n;
    u64 x[MAXD];
#pragma unroll
    for (int i = 0; i < MAXD; i++) {
        x[i] = 0;
        if (i < K) x[i] = shoup_mul(src[(size_t)i * n + k], __ldg(P.p_hat_inv + i), __ldg(P.p_hat_inv_s + i), T.mc[P.qp[ell + i]].q);
    }
    u64 *dst = conv + (size_t)b * ell * n;
    for (int t = 0; t < ell; t++) {
        Acc128 a;
#pragma unroll
        for (int i = 0; i < MAXD; i++)
            if (i < K) a.mac(x[i], __ldg(P.p_hat + (size_t)i * ell + t));
        dst[(size_t)t * n + k] = reduce128(a.hi, a.lo, T.mc[t]);
    }
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

void k_modup(const DeviceTables &T, const DevicePlan &P, const u64 *coef, const u64 *c_eval, u64 *digits, cudaStream_t st) {
    g_launches++, ks_modup<<<dim3((T.n + KS_THREADS - 1) / KS_THREADS, P.beta), KS_THREADS, 0, st>>>(T, P, coef, c_eval, digits);
}
void k_inner_product(const DeviceTables &T, const DevicePlan &P, const u64 *digits, const u64 *evk, u64 *acc, cudaStream_t st) {
    g_launches++, ks_inner<<<dim3((T.n / 2 + KS_THREADS - 1) / KS_THREADS, P.width), KS_THREADS, 0, st>>>(T, P, digits, evk, acc);
}
void k_moddown_conv(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *conv, cudaStream_t st) {
    g_launches++, ks_moddown_conv<<<dim3((T.n + KS_THREADS - 1) / KS_THREADS, 2), KS_THREADS, 0, st>>>(T, P, acc, conv);
}
void k_moddown_finish(const DeviceTables &T, const DevicePlan &P, const u64 *acc, const u64 *conv, const u64 *add0,
                      const u64 *add1, u64 *out0, u64 *out1, uint32_t galois, cudaStream_t st) {
    g_launches++, ks_moddown_finish<<<dim3((T.n + KS_THREADS - 1) / KS_THREADS, P.ell, 2), KS_THREADS, 0, st>>>(T, P, acc, conv, add0, add1,
                                                                                                  out0, out1, galois);
}

} // namespace fheb200
