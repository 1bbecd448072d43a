// baseconv.cu — fast base conversion (ModUp / ModDown of hybrid key switching) on the integer tensor cores.
//
//   out_t[k] = sum_j y_j[k] * C[j][t]  mod p_t          y_j = x_j * (Q/q_j)^-1 mod q_j  (already applied by the inverse NTT)
//
// is a small dense contraction per coefficient (14 sources x 37 targets at the N = 1024 parameters).  The scalar kernel
// (keyswitch.cu: ks_modup / ks_moddown_conv) spends ~5000 integer instructions per coefficient and digit on it and sits on
// the IMAD issue limit.  Here every 64-bit operand is cut into bytes: with y = sum_a y_a 2^(8a) and the constants
// C'[j][a][t] = C[j][t] 2^(8a) mod p_t precomputed and cut into byte planes b,
//
//   out_t = sum_b 2^(8b) * ( sum_(j,a) y_(j,a) * byte_b(C'[j][a][t]) )  mod p_t
//
// and the inner sums are u8 x u8 -> s32 matrix products (mma.sync.m16n8k32, < 2^23 each, exact).  M = 16 coefficients,
// K = (source, byte) in groups of 32 = 4 sources x 8 bytes, N = 8 targets; one accumulator tile per byte plane.  The A
// fragments are simply the 32-bit halves of the u64 coefficients as they lie in HBM (no byte shuffling); the B fragments
// are precomputed per lane (engine_ctx.cu: make_conv_table) and stay L1 resident.  The planes of one output sit in one
// lane, so the recombination is lane-local, and because the quotient of the recombined value by p_t is below 2^16 the
// reduction is one 32 x 32 multiplication for the quotient and one 17 x 64 for the remainder.
// Measured on B200 (tools/ubench_mma.cu): 572 TMAC/s for this MMA shape, i.e. ~25 us for the 14 GMAC of one ModUp at 41
// limbs; the kernel is bound by the epilogue and the 160 MB it moves instead.
#include <algorithm>
#include <cstdlib>
#include <stdexcept>

#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

#ifndef BC_MT_V
#define BC_MT_V 2
#endif
#ifndef BC_WARPS_V
#define BC_WARPS_V 4
#endif
constexpr int BC_WARPS = BC_WARPS_V; // warps per CTA
constexpr int BC_MT = BC_MT_V;       // 16-coefficient M tiles per warp (B fragments are loaded once per BC_MT MMAs)

__device__ __forceinline__ void imma_16832(int (&c)[4], const uint32_t (&a)[4], const uint2 b) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

// Recombination and reduction of one output.  The planes s[b] < 2^24 give V = sum_b s[b] 2^(8b) = sum over (source,
// byte) terms of (byte < 2^8) * (constant < p), so V / p < 2^16 whatever the size of p: the quotient comes from one 32 x 32
// multiplication of the top bits of V with mu = floor(2^(bits+31) / p) and is at most one short.
//   tc = (p lo, p hi, mu, bits)
__device__ __forceinline__ u64 planes_lo(const int (&s)[8]) { // s0 + s1 2^8 + s2 2^16 + s3 2^24 (< 2^49)
    return mad_wide((uint32_t)s[3], 1u << 24, mad_wide((uint32_t)s[2], 1u << 16, mad_wide((uint32_t)s[1], 1u << 8, (u64)(uint32_t)s[0])));
}
__device__ __forceinline__ u64 finish6(const int (&s)[8], const uint4 tc) { // planes 0..5, p < 2^48: V < 2^64
    const u64 hi = mad_wide((uint32_t)s[5], 1u << 8, (u64)(uint32_t)s[4]);
    const u64 v = planes_lo(s) + (hi << 32);
    const u64 p = ((u64)tc.y << 32) | tc.x;
    const uint32_t top = (uint32_t)(v >> (tc.w - 16));
    const uint32_t qe = (uint32_t)(((u64)top * tc.z) >> 47);
    const u64 r = v - (u64)qe * p;
    return r >= p ? r - p : r;
}
__device__ __forceinline__ u64 finish8(const int (&s)[8], const uint4 tc) { // planes 0..7, 2^48 < p < 2^63: V < 2^79
    const u64 lo = planes_lo(s);
    const u64 hi = mad_wide((uint32_t)s[7], 1u << 24, mad_wide((uint32_t)s[6], 1u << 16, mad_wide((uint32_t)s[5], 1u << 8, (u64)(uint32_t)s[4])));
    const u64 p = ((u64)tc.y << 32) | tc.x;
    const u64 v_hi32 = hi + (lo >> 32);                 // V >> 32
    const u64 v_lo = lo + (hi << 32);                   // V mod 2^64
    const uint32_t top = (uint32_t)(v_hi32 >> (tc.w - 48));
    const uint32_t qe = (uint32_t)(((u64)top * tc.z) >> 47);
    const u64 r = v_lo - (u64)qe * p;                   // V - qe p < 2 p < 2^64: the low words suffice
    return r >= p ? r - p : r;
}

template <int KS>
__global__ void __launch_bounds__(BC_WARPS * 32) bc_convert(DeviceTables T, ConvArgs A) {
    const ConvInstance I = A.inst[blockIdx.y];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, tq = lane & 3, n = T.n;
    const int coef0 = (blockIdx.x * BC_WARPS + warp) * (16 * BC_MT);
    if (coef0 >= n) return;
    const ConvTable &C = I.tab;

    // A fragments: lane (g, tq) holds the (tq & 1) half of sources 4 ks + (tq >> 1) and + 2, rows g and g + 8 of each M tile
    uint32_t a[BC_MT][KS][4];
    {
        const uint32_t *s32 = reinterpret_cast<const uint32_t *>(I.src), *e32 = reinterpret_cast<const uint32_t *>(I.extra);
        const int half = tq & 1, nreal = C.nsrc - (e32 ? 1 : 0);
        auto word = [&](int s, size_t r) -> uint32_t { // 32-bit half of coefficient r of source s
            if (s < nreal) return s32[2 * ((size_t)s * n + r) + half];
            return (e32 && s == nreal) ? e32[2 * r + half] : 0u;
        };
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
            const int s0 = 4 * ks + (tq >> 1), s1 = s0 + 2;
#pragma unroll
            for (int mt = 0; mt < BC_MT; mt++) {
                const size_t r0 = (size_t)(coef0 + 16 * mt + g), r1 = r0 + 8;
                a[mt][ks][0] = word(s0, r0), a[mt][ks][1] = word(s0, r1), a[mt][ks][2] = word(s1, r0), a[mt][ks][3] = word(s1, r1);
            }
        }
    }

    for (int nt = 0; nt < C.NT; nt++) {
        int acc[8][BC_MT][4];
#pragma unroll
        for (int b = 0; b < 8; b++)
#pragma unroll
            for (int mt = 0; mt < BC_MT; mt++) acc[b][mt][0] = acc[b][mt][1] = acc[b][mt][2] = acc[b][mt][3] = 0;
        const int np = C.planes[nt];
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
            const uint2 *bp = C.B + ((size_t)(ks * C.NT + nt) * 8) * 32 + lane;
#pragma unroll
            for (int b = 0; b < 8; b++) {
                if (b < 6 || np > 6) { // planes 6, 7 only exist for targets above 48 bits (uniform branch)
                    const uint2 bf = __ldg(bp + b * 32);
#pragma unroll
                    for (int mt = 0; mt < BC_MT; mt++) imma_16832(acc[b][mt], a[mt][ks], bf);
                }
            }
        }
        // epilogue: this lane owns targets 8 nt + 2 tq + {0, 1}, rows g and g + 8 of both M tiles
#pragma unroll
        for (int cpar = 0; cpar < 2; cpar++) {
            const int tp = 8 * nt + 2 * tq + cpar;
            const int slot = __ldg(C.tslot + tp);
            if (slot < 0) continue;
            const uint4 tc = __ldg(C.tconst + tp);
            u64 *o = I.dst + (size_t)slot * n + coef0 + g;
            // a real branch (not a select): a tile is all-wide or all-narrow except for at most one mixed tile per table
            if (tc.w > 48) {
#pragma unroll
                for (int mt = 0; mt < BC_MT; mt++)
#pragma unroll
                    for (int rh = 0; rh < 2; rh++) {
                        int sv[8];
#pragma unroll
                        for (int b = 0; b < 8; b++) sv[b] = acc[b][mt][2 * rh + cpar];
                        o[16 * mt + 8 * rh] = finish8(sv, tc);
                    }
            } else {
#pragma unroll
                for (int mt = 0; mt < BC_MT; mt++)
#pragma unroll
                    for (int rh = 0; rh < 2; rh++) {
                        int sv[8];
#pragma unroll
                        for (int b = 0; b < 8; b++) sv[b] = acc[b][mt][2 * rh + cpar];
                        o[16 * mt + 8 * rh] = finish6(sv, tc);
                    }
            }
        }
    }

    // ModUp: the digit's own limbs are the evaluation-domain input itself
    if (I.copy_limbs > 0) {
        if (lane < 16 * BC_MT)
            for (int t = 0; t < I.copy_limbs; t++) I.copy_dst[(size_t)t * n + coef0 + lane] = I.copy_src[(size_t)t * n + coef0 + lane];
    }
}

// Persistent form of bc_convert.  ncu on the kernel above: 56 % of the warp stalls are long-scoreboard waits - the B
// fragments (48 KB per digit, three digits per SM) fall out of L1 and every MMA pair waits for an L2 round trip, and the A
// loads of a 32-coefficient tile are only covered by the other warps.  Here a CTA of 8 warps stays on one conversion
// instance: it copies that instance's B fragments and target constants into shared memory once, and every warp walks its
// share of the 32-coefficient tiles, building the A fragments from a warp-private shared-memory stage that cp.async fills
// with the next tile while the current one is in the MMAs (source stride padded to 80 words as in lc_tc_ring).
// Measured on B200 (profiles/r01/baseconv_tiling_sweep.txt): 92.1 / 50.0 us (ModUp / ModDown at 41 limbs) against 84.6 / 44.3 us
// for bc_convert, bit-identical output - the waits were not the B fragments.  Kept behind FHEB200_BC_PERSIST=1 for the record.
constexpr int BCP_WARPS = 8, BCP_SRC_WORDS = 80;
template <int KS>
__global__ void __launch_bounds__(BCP_WARPS * 32, 2) bc_convert_p(DeviceTables T, ConvArgs A) {
    extern __shared__ uint4 bcsm[];
    const ConvInstance &I = A.inst[blockIdx.y];
    const ConvTable &C = I.tab;
    const int NT = C.NT, n = T.n;
    uint2 *Bs = reinterpret_cast<uint2 *>(bcsm);                      // [KS][NT][8][32]
    uint4 *tcs = reinterpret_cast<uint4 *>(Bs + (size_t)KS * NT * 256); // [NT * 8]
    int *slots = reinterpret_cast<int *>(tcs + NT * 8);                // [NT * 8]
    uint32_t *stage_all = reinterpret_cast<uint32_t *>(slots + NT * 8); // [BCP_WARPS][4 KS][BCP_SRC_WORDS]
    {
        const uint4 *Bg = reinterpret_cast<const uint4 *>(C.B);
        uint4 *Bd = reinterpret_cast<uint4 *>(Bs);
        for (int i = threadIdx.x; i < KS * NT * 128; i += BCP_WARPS * 32) Bd[i] = __ldg(Bg + i);
        for (int i = threadIdx.x; i < NT * 8; i += BCP_WARPS * 32) tcs[i] = __ldg(C.tconst + i), slots[i] = __ldg(C.tslot + i);
        for (int i = threadIdx.x; i < BCP_WARPS * 4 * KS * BCP_SRC_WORDS; i += BCP_WARPS * 32) stage_all[i] = 0u;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, tq = lane & 3, half = tq & 1;
    uint32_t *stage = stage_all + (size_t)warp * 4 * KS * BCP_SRC_WORDS;
    const unsigned stage_s = (unsigned)__cvta_generic_to_shared(stage);
    const int ntiles = n / (16 * BC_MT), step = gridDim.x * BCP_WARPS;
    // lane copies chunk c = lane & 15 (rows 2c, 2c + 1) of sources (lane >> 4) + 2 i
    const int csrc = lane >> 4, chunk = lane & 15;
    auto stage_in = [&](int tile) {
        if (tile < ntiles) {
            const u64 *gp = I.src + (size_t)tile * (16 * BC_MT) + 2 * chunk;
#pragma unroll
            for (int i = 0; i < 2 * KS; i++) {
                const int sidx = csrc + 2 * i;
                if (sidx < C.nsrc) {
                    const unsigned dst = stage_s + (unsigned)((sidx * BCP_SRC_WORDS + chunk * 4) * 4);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(gp + (size_t)sidx * n) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int tile = blockIdx.x * BCP_WARPS + warp;
    stage_in(tile);
    for (; tile < ntiles; tile += step) {
        const int coef0 = tile * (16 * BC_MT);
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        uint32_t a[BC_MT][KS][4];
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
            const uint32_t *s0 = stage + (4 * ks + (tq >> 1)) * BCP_SRC_WORDS + half, *s1 = s0 + 2 * BCP_SRC_WORDS;
#pragma unroll
            for (int mt = 0; mt < BC_MT; mt++) {
                const int r0 = 16 * mt + g, r1 = r0 + 8;
                a[mt][ks][0] = s0[2 * r0], a[mt][ks][1] = s0[2 * r1];
                a[mt][ks][2] = s1[2 * r0], a[mt][ks][3] = s1[2 * r1];
            }
        }
        __syncwarp();
        stage_in(tile + step);

        for (int nt = 0; nt < NT; nt++) {
            int acc[8][BC_MT][4];
#pragma unroll
            for (int b = 0; b < 8; b++)
#pragma unroll
                for (int mt = 0; mt < BC_MT; mt++) acc[b][mt][0] = acc[b][mt][1] = acc[b][mt][2] = acc[b][mt][3] = 0;
            const int np = C.planes[nt];
#pragma unroll
            for (int ks = 0; ks < KS; ks++) {
                const uint2 *bp = Bs + ((size_t)(ks * NT + nt) * 8) * 32 + lane;
#pragma unroll
                for (int b = 0; b < 8; b++) {
                    if (b < 6 || np > 6) {
                        const uint2 bf = bp[b * 32];
#pragma unroll
                        for (int mt = 0; mt < BC_MT; mt++) imma_16832(acc[b][mt], a[mt][ks], bf);
                    }
                }
            }
#pragma unroll
            for (int cpar = 0; cpar < 2; cpar++) {
                const int tp = 8 * nt + 2 * tq + cpar;
                const int slot = slots[tp];
                if (slot < 0) continue;
                const uint4 tc = tcs[tp];
                u64 *o = I.dst + (size_t)slot * n + coef0 + g;
                if (tc.w > 48) {
#pragma unroll
                    for (int mt = 0; mt < BC_MT; mt++)
#pragma unroll
                        for (int rh = 0; rh < 2; rh++) {
                            int sv[8];
#pragma unroll
                            for (int b = 0; b < 8; b++) sv[b] = acc[b][mt][2 * rh + cpar];
                            o[16 * mt + 8 * rh] = finish8(sv, tc);
                        }
                } else {
#pragma unroll
                    for (int mt = 0; mt < BC_MT; mt++)
#pragma unroll
                        for (int rh = 0; rh < 2; rh++) {
                            int sv[8];
#pragma unroll
                            for (int b = 0; b < 8; b++) sv[b] = acc[b][mt][2 * rh + cpar];
                            o[16 * mt + 8 * rh] = finish6(sv, tc);
                        }
                }
            }
        }
        if (I.copy_limbs > 0 && lane < 16 * BC_MT)
            for (int t = 0; t < I.copy_limbs; t++) I.copy_dst[(size_t)t * n + coef0 + lane] = I.copy_src[(size_t)t * n + coef0 + lane];
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}
static size_t bcp_smem(int KS, int NT) {
    return (size_t)KS * NT * 2048 + (size_t)NT * 8 * (sizeof(uint4) + sizeof(int)) + (size_t)BCP_WARPS * 4 * KS * BCP_SRC_WORDS * 4;
}

// ---------------------------------------------------------------------------------------------------------------------
// Weighted sums of many polynomials (the leaf sums of Paterson-Stockmeyer: up to 8 outputs = sum_t w[g][t] * in_t, 63
// inputs at degree 6510) are the same contraction as a base conversion with all targets on one modulus: K = (input,
// byte), N = 8 outputs, M = 16 coefficients.  The CUDA-core kernel (elementwise.cu: ew_lincomb) needs four 30 x 30-bit
// products per (input, output) pair and sits on the integer issue limit at a third of HBM bandwidth; here the products
// are u8 x u8 MMAs (5 byte planes for 40-bit moduli), 8 outputs per pass over the inputs instead of 4, and the kernel
// streams its inputs once.  A fragments are loaded per K step (nothing to reuse: one N tile), the B fragments of a limb
// (built on the device by lc_build_tables from the weight residues) are read by every warp of that limb and stay in L1/L2.
// The recombined value can reach 2^17 q, so the epilogue reduces the full 128-bit value (reduce128) instead of the
// 16-bit-quotient shortcut of the base conversion.  Bit-identical to ew_lincomb (exact integer sums either way).
constexpr int LC_WARPS = 4, LC_MT = 2;

// B[li][ks][b][lane] (uint2): lane (g, tq), reg r, byte i  <->  output g, input j = 4 ks + 2 r + (tq >> 1), input byte
// a = 4 (tq & 1) + i; the entry is byte b of w[g][j] 2^(8a) mod q_li.  One thread per (lane, reg).
__global__ void __launch_bounds__(64) lc_build_tables(DeviceTables T, int ell, int count, int G, int KS, const uint2 *__restrict__ w,
                                                      uint32_t *__restrict__ B) {
    const int ks = blockIdx.x, li = blockIdx.y, lane = threadIdx.x & 31, reg = threadIdx.x >> 5;
    const int g = lane >> 2, tq = lane & 3, j = 4 * ks + 2 * reg + (tq >> 1);
    const ModConst m = T.mc[li];
    uint32_t word[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (g < G && j < count) {
        const uint2 sp = __ldg(w + (size_t)(g * count + j) * ell + li); // (lo30, hi30) split residue
        const u64 c0 = (u64)sp.x | ((u64)sp.y << 30);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int a = 4 * (tq & 1) + i;
            const u64 cp = mul_mod(c0, reduce64(1ull << (8 * a), m), m);
#pragma unroll
            for (int b = 0; b < 8; b++) word[b] |= (uint32_t)((cp >> (8 * b)) & 0xff) << (8 * i);
        }
    }
#pragma unroll
    for (int b = 0; b < 8; b++) B[((((size_t)li * KS + ks) * 8 + b) * 32 + lane) * 2 + reg] = word[b];
}

__global__ void __launch_bounds__(LC_WARPS * 32) lc_tc(DeviceTables T, int ell, int count, int G, int KS, const u64 *const *__restrict__ in,
                                                       const uint2 *__restrict__ B, u64 *const *__restrict__ out) {
    extern __shared__ u64 lcsm[];
    const u64 **ptr = reinterpret_cast<const u64 **>(lcsm); // [4 KS], padded with input 0 (its weights are zero)
    for (int i = threadIdx.x; i < 4 * KS; i += LC_WARPS * 32) ptr[i] = in[i < count ? i : 0];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, tq = lane & 3, n = T.n;
    const int coef0 = (blockIdx.x * LC_WARPS + warp) * (16 * LC_MT);
    if (coef0 >= n) return;
    const int li = blockIdx.y % ell, half = tq & 1;
    const ModConst m = T.mc[li];
    const int np = (int)(m.bits + 7) >> 3; // byte planes of a constant below q
    const size_t base = (size_t)blockIdx.y * n;
    const uint2 *Bl = B + (size_t)li * KS * 8 * 32 + lane;

    int acc[8][LC_MT][4];
#pragma unroll
    for (int b = 0; b < 8; b++)
#pragma unroll
        for (int mt = 0; mt < LC_MT; mt++) acc[b][mt][0] = acc[b][mt][1] = acc[b][mt][2] = acc[b][mt][3] = 0;

    // A fragments of K step ks + 1 are requested before the MMAs of step ks (two steps of loads in flight per warp)
    uint32_t a[2][LC_MT][4];
    auto load_a = [&](int ks, uint32_t (&dst)[LC_MT][4]) {
        const uint32_t *s0 = reinterpret_cast<const uint32_t *>(ptr[4 * ks + (tq >> 1)] + base);
        const uint32_t *s1 = reinterpret_cast<const uint32_t *>(ptr[4 * ks + (tq >> 1) + 2] + base);
#pragma unroll
        for (int mt = 0; mt < LC_MT; mt++) {
            const size_t r0 = 2 * (size_t)(coef0 + 16 * mt + g) + half, r1 = r0 + 16;
            dst[mt][0] = __ldcs(s0 + r0), dst[mt][1] = __ldcs(s0 + r1), dst[mt][2] = __ldcs(s1 + r0), dst[mt][3] = __ldcs(s1 + r1);
        }
    };
    auto mma_step = [&](int ks, const uint32_t (&src)[LC_MT][4]) {
#pragma unroll
        for (int b = 0; b < 8; b++) {
            if (b < np) { // uniform: one modulus per CTA
                const uint2 bf = __ldg(Bl + (size_t)(ks * 8 + b) * 32);
#pragma unroll
                for (int mt = 0; mt < LC_MT; mt++) imma_16832(acc[b][mt], src[mt], bf);
            }
        }
    };
    load_a(0, a[0]);
    for (int ks = 0; ks < KS; ks += 2) {
        if (ks + 1 < KS) load_a(ks + 1, a[1]);
        mma_step(ks, a[0]);
        if (ks + 1 < KS) {
            if (ks + 2 < KS) load_a(ks + 2, a[0]);
            mma_step(ks + 1, a[1]);
        }
    }
    // this lane owns outputs 2 tq + {0, 1}, rows g and g + 8 of both M tiles
#pragma unroll
    for (int cpar = 0; cpar < 2; cpar++) {
        const int target = 2 * tq + cpar;
        if (target >= G) continue;
        u64 *o = out[target] + base + coef0 + g;
#pragma unroll
        for (int mt = 0; mt < LC_MT; mt++)
#pragma unroll
            for (int rh = 0; rh < 2; rh++) {
                int sv[8];
#pragma unroll
                for (int b = 0; b < 8; b++) sv[b] = acc[b][mt][2 * rh + cpar];
                const u64 lo4 = planes_lo(sv); // planes 0..3, < 2^57
                const u64 hi4 = mad_wide((uint32_t)sv[7], 1u << 24, mad_wide((uint32_t)sv[6], 1u << 16, mad_wide((uint32_t)sv[5], 1u << 8, (u64)(uint32_t)sv[4])));
                const u64 vlo = lo4 + (hi4 << 32), vhi = (hi4 >> 32) + (vlo < lo4); // V = lo4 + hi4 2^32
                o[16 * mt + 8 * rh] = reduce128(vhi, vlo, m);
            }
    }
}


// lc_tc with the inputs streamed through warp-private cp.async rings: every warp stages its own 32 rows of the four sources
// of a K step (1 KB, two 16-byte copies per lane) LC_RING - 1 steps ahead, so 7 KB per warp are in flight without holding
// registers, and builds its A fragments from shared memory (source stride padded to 80 words: the two sources a warp reads
// at once fall on disjoint banks).  Only the warp itself touches its ring: __syncwarp orders copies and reads.
constexpr int LC_RING = 8, LC_SRC_WORDS = 80 /* 32 rows x 2 words + 16 pad */, LC_STAGE_WORDS = 4 * LC_SRC_WORDS;
__global__ void __launch_bounds__(LC_WARPS * 32) lc_tc_ring(DeviceTables T, int ell, int count, int G, int KS, const u64 *const *__restrict__ in,
                                                            const uint2 *__restrict__ B, u64 *const *__restrict__ out) {
    extern __shared__ u64 lcsm[];
    const u64 **ptr = reinterpret_cast<const u64 **>(lcsm); // [4 KS], padded with input 0 (its weights are zero)
    for (int i = threadIdx.x; i < 4 * KS; i += LC_WARPS * 32) ptr[i] = in[i < count ? i : 0];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, tq = lane & 3, n = T.n;
    const int coef0 = (blockIdx.x * LC_WARPS + warp) * (16 * LC_MT);
    if (coef0 >= n) return;
    const int li = blockIdx.y % ell, half = tq & 1;
    const ModConst m = T.mc[li];
    const int np = (int)(m.bits + 7) >> 3;
    const size_t base = (size_t)blockIdx.y * n;
    const uint2 *Bl = B + (size_t)li * KS * 8 * 32 + lane;
    uint32_t *ring = reinterpret_cast<uint32_t *>(lcsm + 4 * KS) + (size_t)warp * LC_RING * LC_STAGE_WORDS;
    const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);

    int acc[8][LC_MT][4];
#pragma unroll
    for (int b = 0; b < 8; b++)
#pragma unroll
        for (int mt = 0; mt < LC_MT; mt++) acc[b][mt][0] = acc[b][mt][1] = acc[b][mt][2] = acc[b][mt][3] = 0;

    // lane copies chunk (16 bytes = rows 2c, 2c + 1) c = lane & 15 of sources (lane >> 4) and (lane >> 4) + 2
    const int csrc = lane >> 4, chunk = lane & 15;
    const size_t goff = base + coef0 + 2 * chunk;
    auto stage_in = [&](int ks) {
        if (ks < KS) {
            const unsigned dst = ring_s + (unsigned)(((ks % LC_RING) * LC_STAGE_WORDS + csrc * LC_SRC_WORDS + chunk * 4) * 4);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(ptr[4 * ks + csrc] + goff) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2 * LC_SRC_WORDS * 4), "l"(ptr[4 * ks + csrc + 2] + goff) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory"); // one group per K step (empty past the end)
    };
#pragma unroll
    for (int ks = 0; ks < LC_RING - 1; ks++) stage_in(ks);
    for (int ks = 0; ks < KS; ks++) {
        stage_in(ks + LC_RING - 1);
        asm volatile("cp.async.wait_group %0;" ::"n"(LC_RING - 1) : "memory");
        __syncwarp();
        const uint32_t *st = ring + (ks % LC_RING) * LC_STAGE_WORDS + (tq >> 1) * LC_SRC_WORDS + half;
        uint32_t a[LC_MT][4];
#pragma unroll
        for (int mt = 0; mt < LC_MT; mt++) {
            const int r0 = 2 * (16 * mt + g);
            a[mt][0] = st[r0], a[mt][1] = st[r0 + 16], a[mt][2] = st[2 * LC_SRC_WORDS + r0], a[mt][3] = st[2 * LC_SRC_WORDS + r0 + 16];
        }
        __syncwarp(); // every lane has its fragments before the slot is refilled by the next iteration's copies
#pragma unroll
        for (int b = 0; b < 8; b++) {
            if (b < np) {
                const uint2 bf = __ldg(Bl + (size_t)(ks * 8 + b) * 32);
#pragma unroll
                for (int mt = 0; mt < LC_MT; mt++) imma_16832(acc[b][mt], a[mt], bf);
            }
        }
    }
#pragma unroll
    for (int cpar = 0; cpar < 2; cpar++) {
        const int target = 2 * tq + cpar;
        if (target >= G) continue;
        u64 *o = out[target] + base + coef0 + g;
#pragma unroll
        for (int mt = 0; mt < LC_MT; mt++)
#pragma unroll
            for (int rh = 0; rh < 2; rh++) {
                int sv[8];
#pragma unroll
                for (int b = 0; b < 8; b++) sv[b] = acc[b][mt][2 * rh + cpar];
                const u64 lo4 = planes_lo(sv);
                const u64 hi4 = mad_wide((uint32_t)sv[7], 1u << 24, mad_wide((uint32_t)sv[6], 1u << 16, mad_wide((uint32_t)sv[5], 1u << 8, (u64)(uint32_t)sv[4])));
                const u64 vlo = lo4 + (hi4 << 32), vhi = (hi4 >> 32) + (vlo < lo4);
                o[16 * mt + 8 * rh] = reduce128(vhi, vlo, m);
            }
    }
}

// in: `count` input pointers, out: G <= 8 output pointers (device arrays), w: split weights [(g * count + t) * ell + limb],
// scratch: ell * ceil(count / 4) * 256 words for the B tables
void k_lincomb_tc(const DeviceTables &T, int ell, int limbs2, int count, int G, const u64 *const *in, const uint2 *w,
                  u64 *const *out, u64 *scratch, cudaStream_t st, int force_ring) {
    if (G < 1 || G > 8) throw std::invalid_argument("lincomb (tensor cores): 1..8 outputs per pass");
    const int KS = (count + 3) / 4;
    // tcgen05 / TMEM form (baseconv_tcgen05.cu): FHEB200_LC_TCGEN05=1; force_ring >= 0 (tests, kernel comparisons) keeps the mma.sync kernels
    if (force_ring < 0 && lc5_mode() == 1 && k_lincomb_tcgen05(T, ell, limbs2, count, G, in, w, out, scratch, st)) return;
    g_launches++, lc_build_tables<<<dim3(KS, ell), 64, 0, st>>>(T, ell, count, G, KS, w, reinterpret_cast<uint32_t *>(scratch));
    const dim3 grid((T.n + LC_WARPS * 16 * LC_MT - 1) / (LC_WARPS * 16 * LC_MT), limbs2);
    // FHEB200_LINCOMB_RING=0: A fragments straight from global memory (lc_tc) instead of through the cp.async rings
    static const bool ring_env = [] { const char *e = getenv("FHEB200_LINCOMB_RING"); return !(e && *e == '0'); }();
    const bool ring = force_ring < 0 ? ring_env : force_ring != 0;
    if (ring) {
        const size_t smem = sizeof(u64) * 4 * KS + sizeof(uint32_t) * LC_WARPS * LC_RING * LC_STAGE_WORDS;
        g_launches++, lc_tc_ring<<<grid, LC_WARPS * 32, smem, st>>>(T, ell, count, G, KS, in, reinterpret_cast<const uint2 *>(scratch), out);
        return;
    }
    g_launches++, lc_tc<<<grid, LC_WARPS * 32, sizeof(u64) * 4 * KS, st>>>(T, ell, count, G, KS, in, reinterpret_cast<const uint2 *>(scratch), out);
}

__global__ void __launch_bounds__(256) bc_lift_overflow(DeviceTables T, DevicePlan P, const u64 *__restrict__ acc, u64 *__restrict__ u) {
    const int b = blockIdx.y, n = T.n;
    const int x = blockIdx.x * 256 + threadIdx.x;
    if (x >= n) return;
    const u64 *src = acc + ((size_t)b * P.width + P.ell) * n + x;
    double v = 0.0;
    for (int j = 0; j < P.K; j++) v = fma((double)src[(size_t)j * n], __ldg(P.r_recip + j), v);
    u[(size_t)b * n + x] = (u64)__double2ll_rn(v);
}
void k_lift_overflow(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *u, cudaStream_t st) {
    g_launches++, bc_lift_overflow<<<dim3((T.n + 255) / 256, 2), 256, 0, st>>>(T, P, acc, u);
}

void k_base_convert(const DeviceTables &T, const ConvArgs &A, int count, int KS, cudaStream_t st) {
    // FHEB200_BC_PERSIST=1: persistent CTAs with the B fragments in shared memory (bc_convert_p); default: one
    // 32-coefficient tile per warp, operands straight from global memory (bc_convert)
    static const bool persist = [] { const char *e = getenv("FHEB200_BC_PERSIST"); return e && *e == '1'; }();
    if (bc5_enabled() && k_base_convert_tcgen05(T, A, count, KS, st)) return;
    int NT = 0;
    for (int i = 0; i < count; i++) NT = std::max(NT, A.inst[i].tab.NT);
    const size_t smem = bcp_smem(KS, NT);
    bool has_extra = false;
    for (int i = 0; i < count; i++) has_extra = has_extra || A.inst[i].extra != nullptr;
    if (persist && !has_extra && BC_MT == 2 && T.n % (16 * BC_MT * BCP_WARPS) == 0 && smem <= 110 * 1024) {
        static int sms = [] { int d = 0, v = 0; cudaGetDevice(&d); cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, d); return v; }();
        const int tiles = T.n / (16 * BC_MT * BCP_WARPS);
        const dim3 grid(std::max(1, std::min(tiles, 2 * sms / count)), count);
        switch (KS) {
#define CASE(k) case k: { \
            static bool attr = false; \
            if (!attr) cudaFuncSetAttribute(bc_convert_p<k>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024), attr = true; \
            g_launches++, bc_convert_p<k><<<grid, BCP_WARPS * 32, smem, st>>>(T, A); } break;
            CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6)
#undef CASE
            default: throw std::invalid_argument("fheb200: more than 24 limbs per digit unsupported");
        }
        return;
    }
    const dim3 grid((T.n + BC_WARPS * 16 * BC_MT - 1) / (BC_WARPS * 16 * BC_MT), count);
    switch (KS) {
#define CASE(k) case k: g_launches++, bc_convert<k><<<grid, BC_WARPS * 32, 0, st>>>(T, A); break;
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6)
#undef CASE
        default: throw std::invalid_argument("fheb200: more than 24 limbs per digit unsupported");
    }
}


// coef: ModUp's inverse-NTT output scaled by DevicePlan::up_scale; c_eval: the evaluation-domain input (own-digit copy)
void k_modup_tc(const DeviceTables &T, const DevicePlan &P, const u64 *coef, const u64 *c_eval, u64 *digits, cudaStream_t st) {
    const size_t n = T.n;
    for (int j = 0; j < P.beta;) { // one launch per run of digits with the same number of K steps
        ConvArgs A;
        int cnt = 0;
        const int KS = P.up[j].KS;
        while (j + cnt < P.beta && P.up[j + cnt].KS == KS) {
            const int d = j + cnt;
            ConvInstance &I = A.inst[cnt++];
            I.tab = P.up[d];
            I.src = coef + (size_t)P.dstart[d] * n;
            I.dst = digits + (size_t)d * P.width * n;
            I.copy_src = c_eval + (size_t)P.dstart[d] * n;
            I.copy_dst = digits + ((size_t)d * P.width + P.dstart[d]) * n;
            I.copy_limbs = P.dsize[d];
        }
        k_base_convert(T, A, cnt, KS, st);
        j += cnt;
    }
}
// acc: [2][width][n] with the P limbs already through the inverse NTT scaled by DevicePlan::down_scale
// extra (merged plans): [2][n] overflow counts, the table's last source
void k_moddown_conv_tc(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *conv, cudaStream_t st, const u64 *extra) {
    const size_t n = T.n;
    ConvArgs A;
    for (int b = 0; b < 2; b++) {
        ConvInstance &I = A.inst[b];
        I.tab = P.down;
        I.src = acc + ((size_t)b * P.width + P.ell) * n;
        I.dst = conv + (size_t)b * P.ell * n;
        I.extra = extra ? extra + (size_t)b * n : nullptr;
    }
    k_base_convert(T, A, 2, P.down.KS, st);
}

} // namespace fheb200
