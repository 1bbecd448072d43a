// elementwise.cu — HBM-bound per-limb kernels: add/sub/mul, scalar ops, tensor product, plaintext product,
// Galois automorphism gather, rescale lift/finish.  One thread handles two adjacent coefficients (128-bit
// loads/stores); grid = (n / 2 / 256, limbs).  Algorithmic bytes per limb: 8n x (#inputs + #outputs).
#include <cstdlib>
#include <stdexcept>

#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

constexpr int EW_THREADS = 256;
typedef ulonglong2 v2;

__device__ __forceinline__ v2 ld2(const u64 *p, size_t i) { return *reinterpret_cast<const v2 *>(p + i); }
__device__ __forceinline__ void st2(u64 *p, size_t i, v2 v) { *reinterpret_cast<v2 *>(p + i) = v; }
#define EW_INDEX                                                         \
    const int limb = blockIdx.y;                                         \
    const int j = (blockIdx.x * EW_THREADS + threadIdx.x) * 2;           \
    if (j >= T.n) return;                                                \
    const size_t x = (size_t)limb * T.n + j;

static inline dim3 ew_grid(const DeviceTables &T, int limbs) { return dim3((T.n / 2 + EW_THREADS - 1) / EW_THREADS, limbs); }

__global__ void __launch_bounds__(EW_THREADS) ew_binary(DeviceTables T, int op, const u64 *a, const u64 *b, u64 *o) {
    EW_INDEX
    const ModConst m = T.mc[limb];
    v2 A = ld2(a, x), B = ld2(b, x), R;
    if (op == OP_ADD) {
        R.x = add_mod(A.x, B.x, m.q), R.y = add_mod(A.y, B.y, m.q);
    } else if (op == OP_SUB) {
        R.x = sub_mod(A.x, B.x, m.q), R.y = sub_mod(A.y, B.y, m.q);
    } else {
        R.x = mul_mod(A.x, B.x, m), R.y = mul_mod(A.y, B.y, m);
    }
    st2(o, x, R);
}
__global__ void __launch_bounds__(EW_THREADS) ew_negate(DeviceTables T, const u64 *a, u64 *o) {
    EW_INDEX
    const u64 q = T.mc[limb].q;
    v2 A = ld2(a, x);
    A.x = neg_mod(A.x, q), A.y = neg_mod(A.y, q);
    st2(o, x, A);
}
__global__ void __launch_bounds__(EW_THREADS) ew_mul_scalar(DeviceTables T, const u64 *a, LimbScalars s, u64 *o) {
    EW_INDEX
    const u64 q = T.mc[limb].q, w = s.v[limb], ws = s.s[limb];
    v2 A = ld2(a, x);
    A.x = shoup_mul(A.x, w, ws, q), A.y = shoup_mul(A.y, w, ws, q);
    st2(o, x, A);
}
// io += a * s (per-limb scalar): lifts the unswitched polynomials of a relinearisation into the QP accumulator (times P)
__global__ void __launch_bounds__(EW_THREADS) ew_scalar_mad(DeviceTables T, const u64 *a, LimbScalars s, u64 *io) {
    EW_INDEX
    const u64 q = T.mc[limb].q, w = s.v[limb], ws = s.s[limb];
    v2 A = ld2(a, x), O = ld2(io, x);
    O.x = add_mod(O.x, shoup_mul(A.x, w, ws, q), q), O.y = add_mod(O.y, shoup_mul(A.y, w, ws, q), q);
    st2(io, x, O);
}
__global__ void __launch_bounds__(EW_THREADS) ew_add_scalar(DeviceTables T, const u64 *a, LimbScalars s, u64 *o) {
    EW_INDEX
    const u64 q = T.mc[limb].q, w = s.v[limb];
    v2 A = ld2(a, x);
    A.x = add_mod(A.x, w, q), A.y = add_mod(A.y, w, q);
    st2(o, x, A);
}
__global__ void __launch_bounds__(EW_THREADS) ew_mul_pt(DeviceTables T, const u64 *a0, const u64 *a1, const u64 *pt, u64 *o0,
                                                        u64 *o1, int acc) {
    EW_INDEX
    const ModConst m = T.mc[limb];
    v2 P = ld2(pt, x), A0 = ld2(a0, x), A1 = ld2(a1, x), R0, R1;
    R0.x = mul_mod(A0.x, P.x, m), R0.y = mul_mod(A0.y, P.y, m);
    R1.x = mul_mod(A1.x, P.x, m), R1.y = mul_mod(A1.y, P.y, m);
    if (acc) {
        v2 O0 = ld2(o0, x), O1 = ld2(o1, x);
        R0.x = add_mod(R0.x, O0.x, m.q), R0.y = add_mod(R0.y, O0.y, m.q);
        R1.x = add_mod(R1.x, O1.x, m.q), R1.y = add_mod(R1.y, O1.y, m.q);
    }
    st2(o0, x, R0);
    st2(o1, x, R1);
}
// d0 = a0 b0 ; d1 = a0 b1 + a1 b0 ; d2 = a1 b1   (d1 through one 128-bit accumulation)
__device__ __forceinline__ void tensor1(u64 a0, u64 a1, u64 b0, u64 b1, const ModConst &m, unsigned long long &d0,
                                        unsigned long long &d1, unsigned long long &d2) {
    d0 = mul_mod(a0, b0, m);
    d2 = mul_mod(a1, b1, m);
    Acc128 acc;
    acc.mac(a0, b1);
    acc.mac(a1, b0);
    d1 = reduce128(acc.hi, acc.lo, m);
}
__global__ void __launch_bounds__(EW_THREADS) ew_tensor(DeviceTables T, const u64 *a0, const u64 *a1, const u64 *b0,
                                                        const u64 *b1, u64 *d0, u64 *d1, u64 *d2) {
    EW_INDEX
    const ModConst m = T.mc[limb];
    v2 A0 = ld2(a0, x), A1 = ld2(a1, x), B0 = ld2(b0, x), B1 = ld2(b1, x), D0, D1, D2;
    tensor1(A0.x, A1.x, B0.x, B1.x, m, D0.x, D1.x, D2.x);
    tensor1(A0.y, A1.y, B0.y, B1.y, m, D0.y, D1.y, D2.y);
    st2(d0, x, D0);
    st2(d1, x, D1);
    st2(d2, x, D2);
}
// Galois automorphism in the evaluation domain (bit-reversed storage):
//   out[rev(j)] = in[rev(((2j+1) g mod 2n) >> 1)].  A warp's 32 consecutive outputs come from one aligned
//   32-word block of the input (only the high natural-order bits vary), so the gather stays coalesced.
__global__ void __launch_bounds__(EW_THREADS) ew_automorphism(DeviceTables T, uint32_t g, const u64 *in, u64 *out) {
    const int limb = blockIdx.y;
    const uint32_t p = blockIdx.x * EW_THREADS + threadIdx.x;
    if (p >= (uint32_t)T.n) return;
    const int sh = 32 - T.logn;
    uint32_t j = __brev(p) >> sh;
    uint32_t idx = (((2u * j + 1u) * g) & (2u * T.n - 1u)) >> 1;
    uint32_t src = __brev(idx) >> sh;
    out[(size_t)limb * T.n + p] = in[(size_t)limb * T.n + src];
}
// blockIdx.y = poly * (ell - 1) + limb; last: the dropped limb of poly p in coefficient form at last + p * last_stride
__global__ void __launch_bounds__(EW_THREADS) ew_rescale_lift(DeviceTables T, int ell, const u64 *last, size_t last_stride, u64 *tmp) {
    const int poly = blockIdx.y / (ell - 1), limb = blockIdx.y - poly * (ell - 1);
    const int j = (blockIdx.x * EW_THREADS + threadIdx.x) * 2;
    if (j >= T.n) return;
    const ModConst m = T.mc[limb];
    const u64 ql = T.mc[ell - 1].q, half = ql >> 1;
    v2 V = ld2(last + poly * last_stride, j), R;
    R.x = switch_modulus(V.x, ql, half, m), R.y = switch_modulus(V.y, ql, half, m);
    st2(tmp, (size_t)blockIdx.y * T.n + j, R);
}
// ModRaise (bootstrapping): coefficient-domain limb 0 of poly p at src + p * src_stride, taken as a centred integer in
// (-q_0/2, q_0/2], re-expressed on limbs 0..ell_out-1.  blockIdx.y = poly * ell_out + limb
__global__ void __launch_bounds__(EW_THREADS) ew_modraise_lift(DeviceTables T, int ell_out, const u64 *src, size_t src_stride, u64 *out) {
    const int poly = blockIdx.y / ell_out, limb = blockIdx.y - poly * ell_out;
    const int j = (blockIdx.x * EW_THREADS + threadIdx.x) * 2;
    if (j >= T.n) return;
    const ModConst m = T.mc[limb];
    const u64 q0 = T.mc[0].q, half = q0 >> 1;
    v2 V = ld2(src + poly * src_stride, j), R;
    if (limb == 0) R = V;
    else R.x = switch_modulus(V.x, q0, half, m), R.y = switch_modulus(V.y, q0, half, m);
    st2(out, (size_t)blockIdx.y * T.n + j, R);
}
// blockIdx.y = poly * ell_out + limb; in has ell_out + 1 limbs per poly, tmp and o have ell_out
__global__ void __launch_bounds__(EW_THREADS) ew_rescale_finish(DeviceTables T, int ell_out, const u64 *in, const u64 *tmp,
                                                                LimbScalars s, u64 *o) {
    const int poly = blockIdx.y / ell_out, limb = blockIdx.y - poly * ell_out;
    const int j = (blockIdx.x * EW_THREADS + threadIdx.x) * 2;
    if (j >= T.n) return;
    const size_t x = (size_t)blockIdx.y * T.n + j;
    const u64 q = T.mc[limb].q, w = s.v[limb], ws = s.s[limb];
    v2 A = ld2(in, x + (size_t)poly * T.n), B = ld2(tmp, x);
    A.x = shoup_mul(sub_mod(A.x, B.x, q), w, ws, q), A.y = shoup_mul(sub_mod(A.y, B.y, q), w, ws, q);
    st2(o, x, A);
}

// fused linear combination (Paterson-Stockmeyer leaves): each input coefficient is loaded once and feeds G accumulators.
// The input pointer table and this limb's weights are staged in shared memory first, so the data loads of the main loop
// do not hang off a pointer load each and can be issued eight at a time.
// CH chunks of EW_THREADS coefficients per CTA: the staging of the pointer table and of this limb's G * count weights
// (a strided gather and a barrier) is paid once per CH * 256 coefficients instead of once per 256
template <int G, int CH>
__global__ void __launch_bounds__(EW_THREADS) ew_lincomb(DeviceTables T, int ell, int count, const u64 *const *__restrict__ in,
                                                         const uint2 *__restrict__ w, u64 *const *__restrict__ out) {
    extern __shared__ u64 lsm[];
    const u64 **ptr = reinterpret_cast<const u64 **>(lsm);
    uint2 *wt = reinterpret_cast<uint2 *>(lsm + count); // [G][count]
    const int n = T.n, li = blockIdx.y % ell;
    for (int i = threadIdx.x; i < count; i += EW_THREADS) ptr[i] = in[i];
    for (int i = threadIdx.x; i < G * count; i += EW_THREADS) wt[i] = __ldg(w + (size_t)i * ell + li);
    __syncthreads();
    const ModConst m = T.mc[li];
#pragma unroll 1
    for (int chunk = 0; chunk < CH; chunk++) {
    const int k = (blockIdx.x * CH + chunk) * EW_THREADS + threadIdx.x;
    if (k >= n) return;
    const size_t idx = (size_t)blockIdx.y * n + k;
    u64 lo[G], hi[G];
#pragma unroll
    for (int g = 0; g < G; g++) lo[g] = 0, hi[g] = 0;
    for (int t0 = 0; t0 < count; t0 += 16) { // 16 products of 30-bit halves fit a 64-bit partial sum
        u64 a0[G], a1[G], a2[G], a3[G];
#pragma unroll
        for (int g = 0; g < G; g++) a0[g] = a1[g] = a2[g] = a3[g] = 0;
#pragma unroll
        for (int h = 0; h < 16; h += 8) {
            u64 a[8];
#pragma unroll
            for (int u = 0; u < 8; u++) a[u] = t0 + h + u < count ? ptr[t0 + h + u][idx] : 0;
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int t = t0 + h + u;
                if (t < count) {
                    const uint32_t al = (uint32_t)a[u] & 0x3fffffffu, ah = (uint32_t)(a[u] >> 30);
#pragma unroll
                    for (int g = 0; g < G; g++) {
                        const uint2 hw = wt[g * count + t];
                        a0[g] = mad_wide(al, hw.x, a0[g]), a1[g] = mad_wide(al, hw.y, a1[g]);
                        a2[g] = mad_wide(ah, hw.x, a2[g]), a3[g] = mad_wide(ah, hw.y, a3[g]);
                    }
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G; g++) acc_split30(lo[g], hi[g], a0[g], a1[g], a2[g], a3[g]);
    }
#pragma unroll
    for (int g = 0; g < G; g++) out[g][idx] = reduce128(hi[g], lo[g], m);
    }
}

// Same sums with the loads of the next eight inputs issued before the products of the current eight (register double
// buffer): twice the bytes in flight per thread.  Bit-identical to ew_lincomb (same partial sums in the same order).
template <int G>
__global__ void __launch_bounds__(EW_THREADS) ew_lincomb_pf(DeviceTables T, int ell, int count, const u64 *const *__restrict__ in,
                                                            const uint2 *__restrict__ w, u64 *const *__restrict__ out) {
    extern __shared__ u64 lsm[];
    const u64 **ptr = reinterpret_cast<const u64 **>(lsm);
    uint2 *wt = reinterpret_cast<uint2 *>(lsm + count); // [G][count]
    const int n = T.n, k = blockIdx.x * EW_THREADS + threadIdx.x, li = blockIdx.y % ell;
    for (int i = threadIdx.x; i < count; i += EW_THREADS) ptr[i] = in[i];
    for (int i = threadIdx.x; i < G * count; i += EW_THREADS) wt[i] = __ldg(w + (size_t)i * ell + li);
    __syncthreads();
    if (k >= n) return;
    const size_t idx = (size_t)blockIdx.y * n + k;
    u64 lo[G], hi[G], a0[G], a1[G], a2[G], a3[G];
#pragma unroll
    for (int g = 0; g < G; g++) lo[g] = hi[g] = a0[g] = a1[g] = a2[g] = a3[g] = 0;
    u64 cur[8], nxt[8];
#pragma unroll
    for (int u = 0; u < 8; u++) cur[u] = u < count ? __ldcs(ptr[u] + idx) : 0;
    for (int t0 = 0; t0 < count; t0 += 8) {
#pragma unroll
        for (int u = 0; u < 8; u++) nxt[u] = t0 + 8 + u < count ? __ldcs(ptr[t0 + 8 + u] + idx) : 0;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int t = t0 + u;
            if (t < count) {
                const uint32_t al = (uint32_t)cur[u] & 0x3fffffffu, ah = (uint32_t)(cur[u] >> 30);
#pragma unroll
                for (int g = 0; g < G; g++) {
                    const uint2 hw = wt[g * count + t];
                    a0[g] = mad_wide(al, hw.x, a0[g]), a1[g] = mad_wide(al, hw.y, a1[g]);
                    a2[g] = mad_wide(ah, hw.x, a2[g]), a3[g] = mad_wide(ah, hw.y, a3[g]);
                }
            }
        }
        if ((t0 & 8) || t0 + 8 >= count) { // 16 products of 30-bit halves fit a 64-bit partial sum
#pragma unroll
            for (int g = 0; g < G; g++) {
                acc_split30(lo[g], hi[g], a0[g], a1[g], a2[g], a3[g]);
                a0[g] = a1[g] = a2[g] = a3[g] = 0;
            }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) cur[u] = nxt[u];
    }
    const ModConst m = T.mc[li];
#pragma unroll
    for (int g = 0; g < G; g++) out[g][idx] = reduce128(hi[g], lo[g], m);
}

// Same sums with the inputs streamed through a shared-memory ring by cp.async (LDGSTS): every thread keeps LC_STAGES - 1
// eight-byte copies of its own coefficient in flight without holding registers for them, and only ever reads back what it
// copied itself, so the ring needs no barrier.  The loads of ew_lincomb wait in registers, eight at a time, and the
// kernel sits at a third of HBM peak on load latency; this one keeps 3 CTAs x LC_STAGES x 2 KB in flight per SM.
constexpr int LC_STAGES = 12;
template <int G>
__global__ void __launch_bounds__(EW_THREADS) ew_lincomb_async(DeviceTables T, int ell, int count, const u64 *const *__restrict__ in,
                                                               const uint2 *__restrict__ w, u64 *const *__restrict__ out) {
    extern __shared__ u64 lsm[];
    const u64 **ptr = reinterpret_cast<const u64 **>(lsm);
    uint2 *wt = reinterpret_cast<uint2 *>(lsm + count);            // [G][count]
    u64 *ring = lsm + (size_t)count * (1 + G) + threadIdx.x;       // [LC_STAGES][EW_THREADS], this thread's column
    const int n = T.n, k = blockIdx.x * EW_THREADS + threadIdx.x, li = blockIdx.y % ell;
    for (int i = threadIdx.x; i < count; i += EW_THREADS) ptr[i] = in[i];
    for (int i = threadIdx.x; i < G * count; i += EW_THREADS) wt[i] = __ldg(w + (size_t)i * ell + li);
    __syncthreads();
    if (k >= n) return;
    const size_t idx = (size_t)blockIdx.y * n + k;
    auto fetch = [&](int t) {
        if (t < count) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(ring + (size_t)(t % LC_STAGES) * EW_THREADS);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(ptr[t] + idx) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory"); // one group per input, empty past the end: the wait count stays fixed
    };
#pragma unroll
    for (int t = 0; t < LC_STAGES - 1; t++) fetch(t);
    u64 lo[G], hi[G], a0[G], a1[G], a2[G], a3[G];
#pragma unroll
    for (int g = 0; g < G; g++) lo[g] = hi[g] = a0[g] = a1[g] = a2[g] = a3[g] = 0;
    for (int t = 0; t < count; t++) {
        fetch(t + LC_STAGES - 1);
        asm volatile("cp.async.wait_group %0;" ::"n"(LC_STAGES - 1) : "memory"); // input t has landed
        const u64 a = ring[(size_t)(t % LC_STAGES) * EW_THREADS];
        const uint32_t al = (uint32_t)a & 0x3fffffffu, ah = (uint32_t)(a >> 30);
#pragma unroll
        for (int g = 0; g < G; g++) {
            const uint2 hw = wt[g * count + t];
            a0[g] = mad_wide(al, hw.x, a0[g]), a1[g] = mad_wide(al, hw.y, a1[g]);
            a2[g] = mad_wide(ah, hw.x, a2[g]), a3[g] = mad_wide(ah, hw.y, a3[g]);
        }
        if ((t & 15) == 15 || t + 1 == count) { // 16 products of 30-bit halves fit a 64-bit partial sum
#pragma unroll
            for (int g = 0; g < G; g++) {
                acc_split30(lo[g], hi[g], a0[g], a1[g], a2[g], a3[g]);
                a0[g] = a1[g] = a2[g] = a3[g] = 0;
            }
        }
    }
    const ModConst m = T.mc[li];
#pragma unroll
    for (int g = 0; g < G; g++) out[g][idx] = reduce128(hi[g], lo[g], m);
}

// The leaf sums without a single data-dependent branch: pointer table and weights are padded to a multiple of 16 inputs
// (pad weights are zero, pad pointers repeat input 0), so the body is straight-line code over groups of 16 inputs - the
// per-input bounds checks of ew_lincomb cost it 3x the instructions (ncu: 5.9e8 warp instructions, ALU pipe 50 %, DRAM
// 26 %).  Inputs arrive through a 16-deep cp.async ring (each thread reads back only its own copies: no barrier); ring
// slots, weight and pointer offsets inside a group are compile-time constants.  Same partial sums in the same order as
// ew_lincomb: bit-identical results.
template <int G>
__global__ void __launch_bounds__(EW_THREADS) ew_lincomb_ring(DeviceTables T, int ell, int count, const u64 *const *__restrict__ in,
                                                              const uint2 *__restrict__ w, u64 *const *__restrict__ out) {
    extern __shared__ u64 lsm[];
    const int cpad = (count + 15) & ~15;
    const u64 **ptr = reinterpret_cast<const u64 **>(lsm);          // [cpad]
    uint2 *wt = reinterpret_cast<uint2 *>(lsm + cpad);              // [G][cpad]
    u64 *ring = lsm + (size_t)cpad * (1 + G) + threadIdx.x;         // [16][EW_THREADS], this thread's column
    const int n = T.n, k = blockIdx.x * EW_THREADS + threadIdx.x, li = blockIdx.y % ell;
    for (int i = threadIdx.x; i < cpad; i += EW_THREADS) ptr[i] = in[i < count ? i : 0];
    for (int i = threadIdx.x; i < G * cpad; i += EW_THREADS) {
        const int g = i / cpad, t = i - g * cpad;
        wt[i] = t < count ? __ldg(w + (size_t)(g * count + t) * ell + li) : make_uint2(0u, 0u);
    }
    __syncthreads();
    if (k >= n) return;
    const size_t idx = (size_t)blockIdx.y * n + k;
    const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);
#pragma unroll
    for (int u = 0; u < 15; u++) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ring_s + u * EW_THREADS * 8), "l"(ptr[u] + idx) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    u64 lo[G], hi[G];
#pragma unroll
    for (int g = 0; g < G; g++) lo[g] = hi[g] = 0;
    for (int t0 = 0; t0 < cpad; t0 += 16) {
        u64 a0[G], a1[G], a2[G], a3[G];
#pragma unroll
        for (int g = 0; g < G; g++) a0[g] = a1[g] = a2[g] = a3[g] = 0;
        const u64 **pg = ptr + t0 + 15;
        const uint2 *wg = wt + t0;
#pragma unroll
        for (int u = 0; u < 16; u++) {
            if (t0 + u + 15 < cpad)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ring_s + ((u + 15) & 15) * EW_THREADS * 8), "l"(pg[u] + idx) : "memory");
            asm volatile("cp.async.commit_group;" ::: "memory"); // one group per input (empty past the end)
            asm volatile("cp.async.wait_group 15;" ::: "memory");  // input t0 + u has landed
            const u64 a = ring[u * EW_THREADS];
            const uint32_t al = (uint32_t)a & 0x3fffffffu, ah = (uint32_t)(a >> 30);
#pragma unroll
            for (int g = 0; g < G; g++) {
                const uint2 hw = wg[g * cpad + u];
                a0[g] = mad_wide(al, hw.x, a0[g]), a1[g] = mad_wide(al, hw.y, a1[g]);
                a2[g] = mad_wide(ah, hw.x, a2[g]), a3[g] = mad_wide(ah, hw.y, a3[g]);
            }
        }
#pragma unroll
        for (int g = 0; g < G; g++) acc_split30(lo[g], hi[g], a0[g], a1[g], a2[g], a3[g]); // 16 products fit 64 bits
    }
    const ModConst m = T.mc[li];
#pragma unroll
    for (int g = 0; g < G; g++) out[g][idx] = reduce128(hi[g], lo[g], m);
}
void k_lincomb(const DeviceTables &T, int ell, int limbs2, int count, int G, const u64 *const *in, const uint2 *w,
               u64 *const *out, cudaStream_t st) {
    dim3 grid((T.n + EW_THREADS - 1) / EW_THREADS, limbs2);
    const size_t smem = sizeof(u64) * (size_t)count * (1 + G);
    if (smem > 48 * 1024) throw std::invalid_argument("lincomb: too many inputs for one pass");
    g_launches++;
    // FHEB200_LINCOMB: 1 = one chunk of 256 coefficients per CTA, 2 = register-prefetching variant, 3 = cp.async ring, 5 = branch-free padded groups of 16 through a cp.async ring,
    // 4 / 8 = chunks per CTA
    static const int variant = [] { const char *e = getenv("FHEB200_LINCOMB"); return e ? atoi(e) : 1; }();
#define LC_LAUNCH(KERNEL, GRID)                                                                 \
    switch (G) {                                                                                \
    case 1: KERNEL(1)<<<GRID, EW_THREADS, smem, st>>>(T, ell, count, in, w, out); break;        \
    case 2: KERNEL(2)<<<GRID, EW_THREADS, smem, st>>>(T, ell, count, in, w, out); break;        \
    case 3: KERNEL(3)<<<GRID, EW_THREADS, smem, st>>>(T, ell, count, in, w, out); break;        \
    default: KERNEL(4)<<<GRID, EW_THREADS, smem, st>>>(T, ell, count, in, w, out); break;       \
    }
#define LC_RG(g) ew_lincomb_ring<g>
#define LC_AS(g) ew_lincomb_async<g>
#define LC_PF(g) ew_lincomb_pf<g>
#define LC_C1(g) ew_lincomb<g, 1>
#define LC_C4(g) ew_lincomb<g, 4>
#define LC_C8(g) ew_lincomb<g, 8>
    const int blocks = (T.n + EW_THREADS - 1) / EW_THREADS;
    if (variant == 5 && count >= 12) { // (short sums would mostly multiply padding)
        const int cpad = (count + 15) & ~15;
        const size_t smem = sizeof(u64) * ((size_t)cpad * (1 + G) + 16 * EW_THREADS);
        LC_LAUNCH(LC_RG, grid)
    } else if (variant == 3) {
        const size_t smem0 = smem;
        {
            const size_t smem = smem0 + sizeof(u64) * LC_STAGES * EW_THREADS; // shadows the table-only size in LC_LAUNCH
            LC_LAUNCH(LC_AS, grid)
        }
    } else if (variant == 2) { LC_LAUNCH(LC_PF, grid) }
    else if (variant == 8) { dim3 g8((blocks + 7) / 8, limbs2); LC_LAUNCH(LC_C8, g8) }
    else if (variant == 4) { dim3 g4((blocks + 3) / 4, limbs2); LC_LAUNCH(LC_C4, g4) }
    else { LC_LAUNCH(LC_C1, grid) }
}

__global__ void __launch_bounds__(EW_THREADS) ew_reduce_words(DeviceTables T, u64 *__restrict__ a) {
    const int k = blockIdx.x * EW_THREADS + threadIdx.x;
    if (k >= T.n) return;
    const size_t idx = (size_t)blockIdx.y * T.n + k;
    a[idx] = reduce64(a[idx], T.mc[blockIdx.y]);
}
void k_reduce_words(const DeviceTables &T, int ell, u64 *a, cudaStream_t st) {
    g_launches++, ew_reduce_words<<<dim3((T.n + EW_THREADS - 1) / EW_THREADS, ell), EW_THREADS, 0, st>>>(T, a);
}

void k_binary(const DeviceTables &T, BinOp op, int ell, const u64 *a, const u64 *b, u64 *out, cudaStream_t st) {
    g_launches++, ew_binary<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, (int)op, a, b, out);
}
void k_negate(const DeviceTables &T, int ell, const u64 *a, u64 *out, cudaStream_t st) {
    g_launches++, ew_negate<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a, out);
}
void k_mul_scalar(const DeviceTables &T, int ell, const u64 *a, const LimbScalars &s, u64 *out, cudaStream_t st) {
    g_launches++, ew_mul_scalar<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a, s, out);
}
void k_scalar_mad(const DeviceTables &T, int ell, const u64 *a, const LimbScalars &s, u64 *io, cudaStream_t st) {
    g_launches++, ew_scalar_mad<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a, s, io);
}
void k_add_scalar(const DeviceTables &T, int ell, const u64 *a, const LimbScalars &s, u64 *out, cudaStream_t st) {
    g_launches++, ew_add_scalar<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a, s, out);
}
void k_mul_pt(const DeviceTables &T, int ell, const u64 *a0, const u64 *a1, const u64 *pt, u64 *o0, u64 *o1, int acc,
              cudaStream_t st) {
    g_launches++, ew_mul_pt<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a0, a1, pt, o0, o1, acc);
}
void k_tensor(const DeviceTables &T, int ell, const u64 *a0, const u64 *a1, const u64 *b0, const u64 *b1, u64 *d0, u64 *d1,
              u64 *d2, cudaStream_t st) {
    g_launches++, ew_tensor<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a0, a1, b0, b1, d0, d1, d2);
}
void k_tensor_square(const DeviceTables &T, int ell, const u64 *a0, const u64 *a1, u64 *d0, u64 *d1, u64 *d2,
                     cudaStream_t st) {
    g_launches++, ew_tensor<<<ew_grid(T, ell), EW_THREADS, 0, st>>>(T, a0, a1, a0, a1, d0, d1, d2);
}
void k_automorphism(const DeviceTables &T, uint32_t g, int limbs, const u64 *in, u64 *out, cudaStream_t st) {
    g_launches++, ew_automorphism<<<dim3((T.n + EW_THREADS - 1) / EW_THREADS, limbs), EW_THREADS, 0, st>>>(T, g, in, out);
}
void k_rescale_lift(const DeviceTables &T, int ell, int npoly, const u64 *last, size_t last_stride, u64 *tmp, cudaStream_t st) {
    g_launches++, ew_rescale_lift<<<ew_grid(T, npoly * (ell - 1)), EW_THREADS, 0, st>>>(T, ell, last, last_stride, tmp);
}
void k_modraise_lift(const DeviceTables &T, int ell_out, int npoly, const u64 *src, size_t src_stride, u64 *out, cudaStream_t st) {
    g_launches++, ew_modraise_lift<<<ew_grid(T, npoly * ell_out), EW_THREADS, 0, st>>>(T, ell_out, src, src_stride, out);
}
void k_rescale_finish(const DeviceTables &T, int ell_out, int npoly, const u64 *in, const u64 *tmp, const LimbScalars &s, u64 *out,
                      cudaStream_t st) {
    g_launches++, ew_rescale_finish<<<ew_grid(T, npoly * ell_out), EW_THREADS, 0, st>>>(T, ell_out, in, tmp, s, out);
}

} // namespace fheb200
