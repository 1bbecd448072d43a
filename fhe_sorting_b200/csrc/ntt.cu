// ntt.cu — negacyclic NTT / iNTT over 64-bit prime moduli (q < 2^60), batched over limbs.
//
// Forward  : Cooley-Tukey, natural order in, bit-reversed order out, twiddles psi^bitrev(i); Harvey lazy butterflies
//            (values in [0,4q)).   Inverse: Gentleman-Sande mirror image, values in [0,2q), n^-1 applied on the way out.
// Work split: the log2(n) radix-2 stages run in at most two launches ("passes"); each CTA owns a tile of up to 4096
// coefficients and each thread 8 of them, so three stages at a time (a radix-8 step, 12 butterflies) run in registers and
// the tile only goes through shared memory between radix-8 steps (two exchanges for 9 stages, padded against bank
// conflicts).  The first load and the last store of a pass go straight between HBM and registers.
//   pass A : the SA = log2(n) - 9 large-stride stages; tile = 2^SA rows x C contiguous columns (C >= 16 words = 128 B)
//   pass B : the last 9 stages; tile = 4096 contiguous coefficients.   n <= 4096 is a single pass-B launch.
// Values stay lazy (unreduced) between the two passes.  Twiddles are stored interleaved with their Shoup companions
// (one 16-byte load per butterfly group).  Algorithmic traffic: 16 n bytes per limb per pass (+ twiddles, L2 resident
// when limbs of equal modulus are adjacent in the batch).
// Warp shuffles were considered for the 8x8 exchange between the last two radix-8 steps; 8 STS.64 + 8 LDS.64 per
// thread is fewer issue slots than the 48 SHFL a 64-bit 8x8 lane transpose needs, so shared memory is used.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <cooperative_groups.h>

#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

std::atomic<unsigned long long> g_launches{0};
std::atomic<unsigned long long> g_algo_bytes{0};
std::atomic<unsigned long long> g_op_count[4];

#ifndef NTT_TILE_LOG_V
#define NTT_TILE_LOG_V 11
#endif
#ifndef NTT_CLUSTER_MIN_DEFAULT
#define NTT_CLUSTER_MIN_DEFAULT 0 // 0 = off
#endif
#ifndef NTT_PERSIST_DEFAULT
#define NTT_PERSIST_DEFAULT 0
#endif
// stages of the last pass at n = 2^17 (the first pass runs the other 17 - NTT_SB on rows of 2^(NTT_TILE_LOG - 17 + NTT_SB) words)
#ifndef NTT_SB
#define NTT_SB 9
#endif
#define NTT_SA (17 - NTT_SB)
constexpr int NTT_TILE_LOG = NTT_TILE_LOG_V;
#ifndef NTT_CTAS_X
#define NTT_CTAS_X (8192 / (1 << NTT_TILE_LOG_V)) // 4 resident CTAs per SM: 1024 threads of 64 registers
#endif
constexpr int NTT_CTAS = NTT_CTAS_X; // (5: 48 registers, 6: 40 registers - the integer paths spill a few words)
constexpr int NTT_TILE = 1 << NTT_TILE_LOG;            // coefficients per CTA tile
constexpr int NTT_PAD_WORDS = NTT_TILE + NTT_TILE / 8; // e -> e + (e >> 3) keeps every exchange pattern conflict free

// LAZY (moduli with bits + log2(n) + 1 <= 63, i.e. every 40..45-bit scaling prime at n = 2^17): no conditional
// subtraction at all; forward values grow by 2q per stage (< 35q), inverse sums double per stage (< 2^logn q).
template <bool LAZY>
__device__ __forceinline__ void ct_bf(u64 &x, u64 &y, const ulonglong2 w, u64 nq, u64 q2) {
    u64 a = LAZY ? x : (x >= q2 ? x - q2 : x);
    u64 t = shoup_lazy_n(y, w.x, w.y, nq);
    x = a + t;
    y = a + q2 - t;
}
// c = multiple of q that bounds y: 2q in the strict variant, q << (stages already done) in the lazy one
template <bool LAZY>
__device__ __forceinline__ void gs_bf(u64 &x, u64 &y, const ulonglong2 w, u64 nq, u64 q2, u64 c) {
    u64 s = x + y;
    u64 d = x + (LAZY ? c : q2) - y;
    x = LAZY ? s : (s >= q2 ? s - q2 : s);
    y = shoup_lazy_n(d, w.x, w.y, nq);
}
__device__ __forceinline__ int pad(int e) { return e + (e >> 3); }

// Optional epilogue of a forward pass B whose 8 outputs per thread are contiguous:
//   rescale  : out = (in - ntt) * w mod q, stored by the pass itself               (`in` addressed like the output + in_off)
//   ModDown  : out[sigma^-1(g)] = (in - ntt) * w + add mod q, stored HERE through the inverse Galois map (scatter != null)
struct FwdEpilogue {
    const u64 *in = nullptr;
    size_t in_off = 0;
    u64 w = 0, ws = 0;
    const u64 *add = nullptr; // this limb of the polynomial that is added (may be null)
    u64 *scatter = nullptr;   // this limb of the output polynomial; null = the pass stores to its own destination
    uint32_t ginv = 1;        // inverse Galois element (1 = identity)
    int logn = 0;
};
// returns true when the values have been stored
__device__ __forceinline__ bool apply_epilogue(const FwdEpilogue &E, u64 (&o)[8], size_t at, uint32_t g0, u64 q) {
    const ulonglong2 *a = reinterpret_cast<const ulonglong2 *>(E.in + at + E.in_off);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const ulonglong2 t = a[k];
        o[2 * k] = shoup_mul(sub_mod(t.x, o[2 * k], q), E.w, E.ws, q);
        o[2 * k + 1] = shoup_mul(sub_mod(t.y, o[2 * k + 1], q), E.w, E.ws, q);
    }
    if (!E.scatter) return false;
    if (E.add) {
        const ulonglong2 *b = reinterpret_cast<const ulonglong2 *>(E.add + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const ulonglong2 t = b[k];
            o[2 * k] = add_mod(o[2 * k], t.x, q), o[2 * k + 1] = add_mod(o[2 * k + 1], t.y, q);
        }
    }
    if (E.ginv == 1u) {
        ulonglong2 *d = reinterpret_cast<ulonglong2 *>(E.scatter + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) d[k] = make_ulonglong2(o[2 * k], o[2 * k + 1]);
    } else { // value at bit-reversed position s belongs at p with s = src(p): p = brev(((2 brev(s) + 1) g^-1 mod 2n) >> 1)
        const int sh = 32 - E.logn;
        const uint32_t mask = (2u << E.logn) - 1u;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t idx = __brev(g0 + k) >> sh;
            const uint32_t j = (((2u * idx + 1u) * E.ginv) & mask) >> 1;
            E.scatter[__brev(j) >> sh] = o[k];
        }
    }
    return true;
}

// three (or the top NS of three) forward stages on the thread's 8 values; s = global stage of the triple's top bit
template <int NS, bool LAZY>
__device__ __forceinline__ void fwd_chunk(u64 (&v)[8], const ulonglong2 *__restrict__ tw, int s, int logn, uint32_t g0, u64 nq,
                                          u64 q2) {
    {
        const ulonglong2 w = __ldg(tw + (1u << s) + (g0 >> (logn - s)));
#pragma unroll
        for (int k = 0; k < 4; k++) ct_bf<LAZY>(v[k], v[k + 4], w, nq, q2);
    }
    if (NS >= 2) {
        const ulonglong2 *p = tw + (2u << s) + (g0 >> (logn - s - 1));
        const ulonglong2 w0 = __ldg(p), w1 = __ldg(p + 1);
        ct_bf<LAZY>(v[0], v[2], w0, nq, q2), ct_bf<LAZY>(v[1], v[3], w0, nq, q2);
        ct_bf<LAZY>(v[4], v[6], w1, nq, q2), ct_bf<LAZY>(v[5], v[7], w1, nq, q2);
    }
    if (NS >= 3) {
        const ulonglong2 *p = tw + (4u << s) + (g0 >> (logn - s - 2));
#pragma unroll
        for (int k = 0; k < 4; k++) ct_bf<LAZY>(v[2 * k], v[2 * k + 1], __ldg(p + k), nq, q2);
    }
}
template <int NS, bool LAZY>
__device__ __forceinline__ void inv_chunk(u64 (&v)[8], const ulonglong2 *__restrict__ tw, int s, int logn, uint32_t g0, u64 q,
                                          u64 nq, u64 q2) {
    // stage g has (logn - 1 - g) stages before it
    if (NS >= 3) {
        const ulonglong2 *p = tw + (4u << s) + (g0 >> (logn - s - 2));
        const u64 c = q << (logn - 3 - s);
#pragma unroll
        for (int k = 0; k < 4; k++) gs_bf<LAZY>(v[2 * k], v[2 * k + 1], __ldg(p + k), nq, q2, c);
    }
    if (NS >= 2) {
        const ulonglong2 *p = tw + (2u << s) + (g0 >> (logn - s - 1));
        const ulonglong2 w0 = __ldg(p), w1 = __ldg(p + 1);
        const u64 c = q << (logn - 2 - s);
        gs_bf<LAZY>(v[0], v[2], w0, nq, q2, c), gs_bf<LAZY>(v[1], v[3], w0, nq, q2, c);
        gs_bf<LAZY>(v[4], v[6], w1, nq, q2, c), gs_bf<LAZY>(v[5], v[7], w1, nq, q2, c);
    }
    {
        const ulonglong2 w = __ldg(tw + (1u << s) + (g0 >> (logn - s)));
        const u64 c = q << (logn - 1 - s);
#pragma unroll
        for (int k = 0; k < 4; k++) gs_bf<LAZY>(v[k], v[k + 4], w, nq, q2, c);
    }
}

// Geometry of one pass.  Tile index e (0..tile-1) -> coefficient index g inside the limb:
//   pass A: g = (e >> lc) * rowstride + blockIdx.x * C + (e & (C-1)),  lc = log2 C ; stage u acts on bit (S-1-u) + lc of e
//   pass B: g = blockIdx.x * tile + e                                           ; stage u acts on bit (S-1-u) of e
struct PassGeom {
    int logn, S, S0, lc, log_tile; // S stages starting at global stage S0; lc = 0 for pass B
    int pass_a;
};
__device__ __forceinline__ uint32_t tile_to_g(const PassGeom &G, int e, uint32_t tile) {
    if (G.pass_a) return ((uint32_t)(e >> G.lc) << (G.logn - G.S)) + (tile << G.lc) + (e & ((1 << G.lc) - 1));
    return (tile << G.log_tile) + e;
}
// CG: the word was written earlier in the same launch by another SM (fused two-phase kernels): read it from L2, never
// from this SM's L1, which may still hold the line as phase 0 of an in-place transform saw it
template <bool CG>
__device__ __forceinline__ u64 ld_word(const u64 *p) { return CG ? __ldcg(p) : *p; }
template <bool CG>
__device__ __forceinline__ ulonglong2 ld_pair(const ulonglong2 *p) { return CG ? __ldcg(p) : *p; }
__device__ __forceinline__ int triple_base(int tid, int p) { return ((tid >> p) << (p + 3)) | (tid & ((1 << p) - 1)); }

__device__ __forceinline__ uint32_t g_step(const PassGeom &G, int p) {
    return G.pass_a ? (1u << (p - G.lc + G.logn - G.S)) : (1u << p);
}

// ---------------------------------------------------------------- forward pass
// final_reduce: bring values to [0,q) on the way out (last pass), otherwise they stay in [0,4q)
// pre (optional): the 8 words of this thread's first access pattern, already loaded by the caller (persistent kernels
// fetch them one tile ahead with tile_prefetch); src is then not read
template <bool LAZY, bool CG>
__device__ __forceinline__ void fwd_pass_body(const DeviceTables &T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                              const uint32_t d, const PassGeom &G, int final_reduce, u64 *sm, uint32_t tile,
                                              const u64 *pre = nullptr, const FwdEpilogue *epi = nullptr) {
    const int qp = d & 0xff, n = T.n, tid = threadIdx.x;
    const size_t off = (size_t)(d >> 8) * n;
    const ModConst mc = T.mc[qp];
    const u64 q = mc.q, q2 = q << 1, nq = 0 - q;
    const ulonglong2 *tw = T.tw + (size_t)qp * n;
    const int lowbit = G.pass_a ? G.lc : 0;
    const int c_first = (G.S - 1) % 3 + 1;
    int p = G.S - 3 + lowbit; // low bit of the current triple
    int base = triple_base(tid, p);
    uint32_t g0 = tile_to_g(G, base, tile);
    u64 v[8];
    {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = pre ? pre[k] : ld_word<CG>(src + off + g0 + k * st);
    }
    int s = G.S0;
    if (c_first == 1) fwd_chunk<1, LAZY>(v, tw, s, G.logn, g0, nq, q2);
    else if (c_first == 2) fwd_chunk<2, LAZY>(v, tw, s, G.logn, g0, nq, q2);
    else fwd_chunk<3, LAZY>(v, tw, s, G.logn, g0, nq, q2);
    s += c_first;
    bool first = true;
#pragma unroll
    for (int pn = p - c_first; pn >= lowbit; pn -= 3) {
        if (!first) __syncthreads();
        first = false;
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        __syncthreads();
        p = pn;
        base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        g0 = tile_to_g(G, base, tile);
        fwd_chunk<3, LAZY>(v, tw, s, G.logn, g0, nq, q2);
        s += 3;
    }
    if (final_reduce) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            u64 x = v[k];
            if (LAZY) {
                v[k] = reduce64(x, mc);
            } else {
                x = x >= q2 ? x - q2 : x;
                v[k] = x >= q ? x - q : x;
            }
        }
    }
    if (!G.pass_a && p == 0) {
        if (epi && apply_epilogue(*epi, v, off + g0, g0, q)) return;
        ulonglong2 *o = reinterpret_cast<ulonglong2 *>(dst + off + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) o[k] = make_ulonglong2(v[2 * k], v[2 * k + 1]);
    } else {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = v[k];
    }
}

// ---------------------------------------------------------------- FP64 path (moduli of at most 42 bits)
// B200 issues DFMA at the same rate as IMAD (tools/ubench_int.cu: 3.4 vs 3.6 warp-inst/clk/SM), and a modular product
// of two < 2^53 integers is 7 double-precision instructions (error-free product + rounded quotient) against ~25
// integer ones for the 64-bit Shoup form.  Values are kept as exact signed integers in doubles:
//   h = RN(y*w) ; l = fma(y,w,-h) (exact) ; c = rint(h * (1/q)) ; d = fma(-c,q,h) (exact: |h - c q| < 2q + |l|) ; g = d + l
// (6 DP instructions, no per-twiddle companion value to load or compute)
// g == y*w (mod q) and |g| < 0.7 q for |y| < 2^50 (emulated exhaustively in tests/test_fp_modmul.py).  No conditional
// corrections inside the butterflies: forward values grow by < q per stage, inverse sums are re-centred every <= 8 stages.
// NTT_EXP (timing experiments only, results are wrong): 1 = constant twiddles, 2 = no shared-memory exchange, 3 = no butterfly
// arithmetic, in the FP64 forward path
#ifndef NTT_EXP
#define NTT_EXP 0
#endif
__device__ __forceinline__ void fp_ct(double &x, double &y, double w, double qinv, double q) {
    if (NTT_EXP == 3 || NTT_EXP == 4) {
        x = __dadd_rn(x, w);
        return;
    }
    const double g = fp_mulmod(y, w, qinv, q);
    y = __dadd_rn(x, -g);
    x = __dadd_rn(x, g);
}
__device__ __forceinline__ void fp_gs(double &x, double &y, double w, double qinv, double q) {
    const double d = __dadd_rn(x, -y);
    x = __dadd_rn(x, y);
    y = fp_mulmod(d, w, qinv, q);
}
template <int NS>
__device__ __forceinline__ void fp_fwd_chunk(double (&v)[8], const double *__restrict__ tw, int s, int logn, uint32_t g0,
                                             double q, double qinv) {
    {
        const double w = (NTT_EXP == 1 || NTT_EXP >= 4) ? 3.0 : __ldg(tw + (1u << s) + (g0 >> (logn - s)));
#pragma unroll
        for (int k = 0; k < 4; k++) fp_ct(v[k], v[k + 4], w, qinv, q);
    }
    if (NS >= 2) {
        const double2 w = (NTT_EXP == 1 || NTT_EXP >= 4) ? make_double2(5.0, 7.0) : __ldg(reinterpret_cast<const double2 *>(tw + (2u << s) + (g0 >> (logn - s - 1))));
        fp_ct(v[0], v[2], w.x, qinv, q), fp_ct(v[1], v[3], w.x, qinv, q);
        fp_ct(v[4], v[6], w.y, qinv, q), fp_ct(v[5], v[7], w.y, qinv, q);
    }
    if (NS >= 3) {
        const double2 *p = reinterpret_cast<const double2 *>(tw + (4u << s) + (g0 >> (logn - s - 2)));
        const double2 w0 = (NTT_EXP == 1 || NTT_EXP >= 4) ? make_double2(9.0, 11.0) : __ldg(p), w1 = (NTT_EXP == 1 || NTT_EXP >= 4) ? make_double2(13.0, 15.0) : __ldg(p + 1);
        fp_ct(v[0], v[1], w0.x, qinv, q), fp_ct(v[2], v[3], w0.y, qinv, q);
        fp_ct(v[4], v[5], w1.x, qinv, q), fp_ct(v[6], v[7], w1.y, qinv, q);
    }
}
template <int NS>
__device__ __forceinline__ void fp_inv_chunk(double (&v)[8], const double *__restrict__ tw, int s, int logn, uint32_t g0,
                                             double q, double qinv) {
    if (NS >= 3) {
        const double2 *p = reinterpret_cast<const double2 *>(tw + (4u << s) + (g0 >> (logn - s - 2)));
        const double2 w0 = __ldg(p), w1 = __ldg(p + 1);
        fp_gs(v[0], v[1], w0.x, qinv, q), fp_gs(v[2], v[3], w0.y, qinv, q);
        fp_gs(v[4], v[5], w1.x, qinv, q), fp_gs(v[6], v[7], w1.y, qinv, q);
    }
    if (NS >= 2) {
        const double2 w = __ldg(reinterpret_cast<const double2 *>(tw + (2u << s) + (g0 >> (logn - s - 1))));
        fp_gs(v[0], v[2], w.x, qinv, q), fp_gs(v[1], v[3], w.x, qinv, q);
        fp_gs(v[4], v[6], w.y, qinv, q), fp_gs(v[5], v[7], w.y, qinv, q);
    }
    {
        const double w = __ldg(tw + (1u << s) + (g0 >> (logn - s)));
#pragma unroll
        for (int k = 0; k < 4; k++) fp_gs(v[k], v[k + 4], w, qinv, q);
    }
}

// stages 14, 15, 16 of the last forward pass at n = 2^17 with factored twiddles (DeviceTables::twa): the table entries of tile T
// are those of tile 0 times one constant per stage, so the loads hit the 14 KB slice every tile of a modulus shares
__device__ __forceinline__ void fp_fwd_chunk_tail(double (&v)[8], const double *__restrict__ tw, const double *__restrict__ A,
                                                  uint32_t tile, uint32_t base, double q, double qinv) {
    const double a14 = __ldg(A + tile), a15 = __ldg(A + 64 + tile), a16 = __ldg(A + 128 + tile);
    {
        const double w = fp_mulmod(__ldg(tw + (1u << 14) + (base >> 3)), a14, qinv, q);
#pragma unroll
        for (int k = 0; k < 4; k++) fp_ct(v[k], v[k + 4], w, qinv, q);
    }
    {
        const double2 t = __ldg(reinterpret_cast<const double2 *>(tw + (2u << 14) + (base >> 2)));
        const double wx = fp_mulmod(t.x, a15, qinv, q), wy = fp_mulmod(t.y, a15, qinv, q);
        fp_ct(v[0], v[2], wx, qinv, q), fp_ct(v[1], v[3], wx, qinv, q);
        fp_ct(v[4], v[6], wy, qinv, q), fp_ct(v[5], v[7], wy, qinv, q);
    }
    {
        const double2 *p = reinterpret_cast<const double2 *>(tw + (4u << 14) + (base >> 1));
        const double2 t0 = __ldg(p), t1 = __ldg(p + 1);
        fp_ct(v[0], v[1], fp_mulmod(t0.x, a16, qinv, q), qinv, q), fp_ct(v[2], v[3], fp_mulmod(t0.y, a16, qinv, q), qinv, q);
        fp_ct(v[4], v[5], fp_mulmod(t1.x, a16, qinv, q), qinv, q), fp_ct(v[6], v[7], fp_mulmod(t1.y, a16, qinv, q), qinv, q);
    }
}
// the mirror image: stages 16, 15, 14 at the start of the first inverse pass (DeviceTables::twia)
__device__ __forceinline__ void fp_inv_chunk_head(double (&v)[8], const double *__restrict__ tw, const double *__restrict__ A,
                                                  uint32_t tile, uint32_t base, double q, double qinv) {
    const double a14 = __ldg(A + tile), a15 = __ldg(A + 64 + tile), a16 = __ldg(A + 128 + tile);
    {
        const double2 *p = reinterpret_cast<const double2 *>(tw + (4u << 14) + (base >> 1));
        const double2 t0 = __ldg(p), t1 = __ldg(p + 1);
        fp_gs(v[0], v[1], fp_mulmod(t0.x, a16, qinv, q), qinv, q), fp_gs(v[2], v[3], fp_mulmod(t0.y, a16, qinv, q), qinv, q);
        fp_gs(v[4], v[5], fp_mulmod(t1.x, a16, qinv, q), qinv, q), fp_gs(v[6], v[7], fp_mulmod(t1.y, a16, qinv, q), qinv, q);
    }
    {
        const double2 t = __ldg(reinterpret_cast<const double2 *>(tw + (2u << 14) + (base >> 2)));
        const double wx = fp_mulmod(t.x, a15, qinv, q), wy = fp_mulmod(t.y, a15, qinv, q);
        fp_gs(v[0], v[2], wx, qinv, q), fp_gs(v[1], v[3], wx, qinv, q);
        fp_gs(v[4], v[6], wy, qinv, q), fp_gs(v[5], v[7], wy, qinv, q);
    }
    {
        const double w = fp_mulmod(__ldg(tw + (1u << 14) + (base >> 3)), a14, qinv, q);
#pragma unroll
        for (int k = 0; k < 4; k++) fp_gs(v[k], v[k + 4], w, qinv, q);
    }
}

// first: the pass reads reduced u64 input (otherwise doubles left by the previous pass); last: writes reduced u64
template <bool CG>
__device__ __forceinline__ void fp_fwd_pass_body(const DeviceTables &T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                 const uint32_t d, const PassGeom &G, int first, int last, u64 *smu, uint32_t tile,
                                                 const u64 *pre = nullptr, const FwdEpilogue *epi = nullptr, const double2 *qdv = nullptr) {
    double *sm = reinterpret_cast<double *>(smu);
    const int qp = d & 0xff, n = T.n, tid = threadIdx.x;
    const size_t off = (size_t)(d >> 8) * n;
    const double2 qd = qdv ? *qdv : T.qd[qp];
    const double q = qd.x, qinv = qd.y;
    const double *tw = T.twf + (size_t)qp * n;
    const int lowbit = G.pass_a ? G.lc : 0;
    const int c_first = (G.S - 1) % 3 + 1;
    int p = G.S - 3 + lowbit;
    int base = triple_base(tid, p);
    uint32_t g0 = tile_to_g(G, base, tile);
    double v[8];
    {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const u64 x = pre ? pre[k] : ld_word<CG>(src + off + g0 + k * st);
            v[k] = first ? fp_from_u64(x) : __longlong_as_double((long long)x);
        }
    }
    int s = G.S0;
    if (c_first == 1) fp_fwd_chunk<1>(v, tw, s, G.logn, g0, q, qinv);
    else if (c_first == 2) fp_fwd_chunk<2>(v, tw, s, G.logn, g0, q, qinv);
    else fp_fwd_chunk<3>(v, tw, s, G.logn, g0, q, qinv);
    s += c_first;
    bool firstx = true;
#pragma unroll
    for (int pn = p - c_first; pn >= lowbit; pn -= 3) {
        if (NTT_EXP != 2 && NTT_EXP < 4) {
            if (!firstx) __syncthreads();
            firstx = false;
#pragma unroll
            for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
            __syncthreads();
        }
        p = pn;
        base = triple_base(tid, p);
        if (NTT_EXP != 2 && NTT_EXP < 4) {
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        }
        g0 = tile_to_g(G, base, tile);
        if (!G.pass_a && G.logn == 17 && s == 14 && p == 0 && T.twa) fp_fwd_chunk_tail(v, tw, T.twa + (size_t)qp * 192, tile, (uint32_t)base, q, qinv);
        else fp_fwd_chunk<3>(v, tw, s, G.logn, g0, q, qinv);
        s += 3;
    }
    u64 o[8];
#pragma unroll
    for (int k = 0; k < 8; k++) o[k] = last ? fp_to_u64(v[k], q, qinv) : (u64)__double_as_longlong(v[k]);
    if (!G.pass_a && p == 0) {
        if (epi && apply_epilogue(*epi, o, off + g0, g0, T.mc[qp].q)) return;
        ulonglong2 *op = reinterpret_cast<ulonglong2 *>(dst + off + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) op[k] = make_ulonglong2(o[2 * k], o[2 * k + 1]);
    } else {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = o[k];
    }
}

template <bool CG>
__device__ __forceinline__ void fp_inv_pass_body(const DeviceTables &T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                 const uint32_t d, const PassGeom &G, int first, int last, u64 *smu,
                                                 const u64 *__restrict__ custom, const u64 *__restrict__ custom_s, uint32_t tile,
                                                 uint32_t bidx, const u64 *pre = nullptr, const double2 *qdv = nullptr) {
    double *sm = reinterpret_cast<double *>(smu);
    const int qp = d & 0xff, n = T.n, tid = threadIdx.x;
    const size_t off = (size_t)(d >> 8) * n;
    const double2 qd = qdv ? *qdv : T.qd[qp];
    const double q = qd.x, qinv = qd.y;
    const double *tw = T.twif + (size_t)qp * n;
    const int lowbit = G.pass_a ? G.lc : 0;
    const int c_last = (G.S - 1) % 3 + 1, ptop = G.S - 3 + lowbit, m = (G.S - c_last) / 3;
    int p = m > 0 ? lowbit : ptop;
    int base = triple_base(tid, p);
    uint32_t g0 = tile_to_g(G, base, tile);
    double v[8];
    if (pre) {
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = first ? fp_from_u64(pre[k]) : __longlong_as_double((long long)pre[k]);
    } else if (!G.pass_a && p == 0) {
        const ulonglong2 *in = reinterpret_cast<const ulonglong2 *>(src + off + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const ulonglong2 t = ld_pair<CG>(in + k);
            v[2 * k] = first ? fp_from_u64(t.x) : __longlong_as_double((long long)t.x);
            v[2 * k + 1] = first ? fp_from_u64(t.y) : __longlong_as_double((long long)t.y);
        }
    } else {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const u64 x = ld_word<CG>(src + off + g0 + k * st);
            v[k] = first ? fp_from_u64(x) : __longlong_as_double((long long)x);
        }
    }
    int growth = 0; // stages since the values were last known to be within (-q, q)
    if (!first) {
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = fp_center(v[k], q, qinv);
    }
#pragma unroll
    for (int i = 0; i < m; i++) {
        if (growth + 3 > 8) {
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = fp_center(v[k], q, qinv);
            growth = 0;
        }
        const int s = G.S0 + G.S - 1 - (p + 2 - lowbit);
        if (!G.pass_a && G.logn == 17 && s == 14 && p == 0 && T.twia) fp_inv_chunk_head(v, tw, T.twia + (size_t)qp * 192, tile, (uint32_t)base, q, qinv);
        else fp_inv_chunk<3>(v, tw, s, G.logn, g0, q, qinv);
        growth += 3;
        if (i > 0) __syncthreads();
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        __syncthreads();
        p = i + 1 < m ? p + 3 : ptop;
        base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        g0 = tile_to_g(G, base, tile);
    }
    if (growth + c_last > 8) {
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = fp_center(v[k], q, qinv);
    }
    if (c_last == 1) fp_inv_chunk<1>(v, tw, G.S0, G.logn, g0, q, qinv);
    else if (c_last == 2) fp_inv_chunk<2>(v, tw, G.S0, G.logn, g0, q, qinv);
    else fp_inv_chunk<3>(v, tw, G.S0, G.logn, g0, q, qinv);
    const uint32_t st = g_step(G, p);
    if (last) {
        const double ni = (double)(custom ? custom[bidx] : T.ninv[qp]);
#pragma unroll
        for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = fp_to_u64(fp_mulmod(v[k], ni, qinv, q), q, qinv);
    } else {
#pragma unroll
        for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = (u64)__double_as_longlong(v[k]);
    }
}
__device__ __forceinline__ bool fp_ok(const DeviceTables &T, int qp) { return T.mc[qp].bits <= 42 && T.logn <= 17; }

__device__ __forceinline__ bool lazy_ok(const DeviceTables &T, int qp) { return T.mc[qp].bits + T.logn + 1 <= 63; }
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_fwd_pass(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                             const uint32_t *__restrict__ desc, PassGeom G, int first, int final_reduce) {
    extern __shared__ u64 sm[];
    const uint32_t d = desc[blockIdx.y];
    if (fp_ok(T, d & 0xff)) fp_fwd_pass_body<false>(T, src, dst, d, G, first, final_reduce, sm, blockIdx.x);
    else if (lazy_ok(T, d & 0xff)) fwd_pass_body<true, false>(T, src, dst, d, G, final_reduce, sm, blockIdx.x);
    else fwd_pass_body<false, false>(T, src, dst, d, G, final_reduce, sm, blockIdx.x);
}

// ---------------------------------------------------------------- inverse pass
// scale: multiply by n^-1 and reduce to [0,q) on the way out (last pass); otherwise values leave in [0,2q)
template <bool LAZY, bool CG>
__device__ __forceinline__ void inv_pass_body(const DeviceTables &T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                              const uint32_t d, const PassGeom &G, int scale, u64 *sm,
                                              const u64 *__restrict__ custom, const u64 *__restrict__ custom_s, uint32_t tile,
                                              uint32_t bidx, const u64 *pre = nullptr) {
    const int qp = d & 0xff, n = T.n, tid = threadIdx.x;
    const size_t off = (size_t)(d >> 8) * n;
    const u64 q = T.mc[qp].q, q2 = q << 1, nq = 0 - q;
    const ulonglong2 *tw = T.twi + (size_t)qp * n;
    const int lowbit = G.pass_a ? G.lc : 0;
    const int c_last = (G.S - 1) % 3 + 1, ptop = G.S - 3 + lowbit, m = (G.S - c_last) / 3; // m full radix-8 steps first
    int p = m > 0 ? lowbit : ptop;
    int base = triple_base(tid, p);
    uint32_t g0 = tile_to_g(G, base, tile);
    u64 v[8];
    if (pre) {
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = pre[k];
    } else if (!G.pass_a && p == 0) {
        const ulonglong2 *in = reinterpret_cast<const ulonglong2 *>(src + off + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            ulonglong2 t = ld_pair<CG>(in + k);
            v[2 * k] = t.x, v[2 * k + 1] = t.y;
        }
    } else {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = ld_word<CG>(src + off + g0 + k * st);
    }
#pragma unroll
    for (int i = 0; i < m; i++) {
        const int s = G.S0 + G.S - 1 - (p + 2 - lowbit); // global stage of the triple's top bit
        inv_chunk<3, LAZY>(v, tw, s, G.logn, g0, q, nq, q2);
        if (i > 0) __syncthreads();
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        __syncthreads();
        p = i + 1 < m ? p + 3 : ptop;
        base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        g0 = tile_to_g(G, base, tile);
    }
    // last (possibly partial) step on the top triple: stages S0 + c_last - 1 ... S0
    if (c_last == 1) inv_chunk<1, LAZY>(v, tw, G.S0, G.logn, g0, q, nq, q2);
    else if (c_last == 2) inv_chunk<2, LAZY>(v, tw, G.S0, G.logn, g0, q, nq, q2);
    else inv_chunk<3, LAZY>(v, tw, G.S0, G.logn, g0, q, nq, q2);
    if (scale) {
        const u64 ni = custom ? custom[bidx] : T.ninv[qp], nis = custom ? custom_s[bidx] : T.ninv[T.F + qp];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = shoup_mul(v[k], ni, nis, q);
    }
    const uint32_t st = g_step(G, p);
#pragma unroll
    for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = v[k];
}

__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_inv_pass(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                             const uint32_t *__restrict__ desc, PassGeom G, int first, int scale,
                                                             const u64 *__restrict__ custom, const u64 *__restrict__ custom_s) {
    extern __shared__ u64 sm[];
    const uint32_t d = desc[blockIdx.y];
    if (fp_ok(T, d & 0xff)) fp_inv_pass_body<false>(T, src, dst, d, G, first, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
    else if (lazy_ok(T, d & 0xff)) inv_pass_body<true, false>(T, src, dst, d, G, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
    else inv_pass_body<false, false>(T, src, dst, d, G, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
}

// Compile-time geometry for the production ring (n = 2^17): every shift, stride and loop trip count of the pass bodies
// becomes an immediate, which removes most of the index arithmetic around the butterflies.
template <int LOGN, int S, int S0, int LC, int PA>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_fwd_pass_c(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                  const uint32_t *__restrict__ desc, int first, int final_reduce) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    const uint32_t d = desc[blockIdx.y];
    if (fp_ok(T, d & 0xff)) fp_fwd_pass_body<false>(T, src, dst, d, G, first, final_reduce, sm, blockIdx.x);
    else if (lazy_ok(T, d & 0xff)) fwd_pass_body<true, false>(T, src, dst, d, G, final_reduce, sm, blockIdx.x);
    else fwd_pass_body<false, false>(T, src, dst, d, G, final_reduce, sm, blockIdx.x);
}
template <int LOGN, int S, int S0, int LC, int PA>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_inv_pass_c(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                  const uint32_t *__restrict__ desc, int first, int scale,
                                                                  const u64 *__restrict__ custom, const u64 *__restrict__ custom_s) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    const uint32_t d = desc[blockIdx.y];
    if (fp_ok(T, d & 0xff)) fp_inv_pass_body<false>(T, src, dst, d, G, first, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
    else if (lazy_ok(T, d & 0xff)) inv_pass_body<true, false>(T, src, dst, d, G, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
    else inv_pass_body<false, false>(T, src, dst, d, G, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
}

// ---------------------------------------------------------------- fused two-phase transform (n = 2^17)
// One launch runs both passes of a whole batch.  Work items (one CTA each) are ordered by blockIdx.x so that phase 1 of
// limb i is queued NTT_AHEAD limbs' worth of phase-0 items after phase 0 of limb i: by the time the hardware dispatches
// it (CTAs start in blockIdx order; the same assumption CUB's decoupled look-back scan makes) its 64 producers have long
// been running or are done, so the per-limb counter it polls is normally already complete, and the 1 MiB intermediate of
// a limb is consumed out of L2 a few microseconds after it was written instead of making a round trip through HBM.
// DRAM traffic per limb: 8n read + 8n written (the algorithmic 16n) + twiddles.  Phase-1 loads bypass L1 (ld.global.cg).
// done[limb]: +1 per finished phase-0 tile (64 = ready), +1 per finished phase-1 tile; the CTA that brings it to 128
// resets it to 0, so the scratch is clean for the next launch on the same stream without a memset.
constexpr int NTT_AHEAD = 10;        // limbs of phase-0 work between a limb's two phases (>= resident CTAs / 64 = 9.25)
constexpr int NTT_FUSED_MAX = 256;   // descriptors travel in the kernel parameters (constant bank)
struct NttBatch {
    uint32_t desc[NTT_FUSED_MAX];
};
__device__ __forceinline__ void fused_item(uint32_t b, int count, int &phase, int &limb) {
    const int D = NTT_AHEAD;
    if (count <= D) {
        phase = (int)b >= count;
        limb = phase ? (int)b - count : (int)b;
    } else if ((int)b < D) {
        phase = 0, limb = (int)b;
    } else {
        const int r = (int)b - D, pairs = count - D;
        if (r < 2 * pairs) phase = !(r & 1), limb = (r & 1) ? D + (r >> 1) : (r >> 1);
        else phase = 1, limb = pairs + (r - 2 * pairs);
    }
}
__device__ __forceinline__ void fused_publish(unsigned *done, int limb, int tiles) { // after the CTA's last global store
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned v = atomicAdd(done + limb, 1u) + 1u;
        if (v == 2u * (unsigned)tiles) atomicExch(done + limb, 0u);
    }
}
__device__ __forceinline__ void fused_wait(unsigned *done, int limb, int tiles) {
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        unsigned v;
        for (;;) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(done + limb) : "memory");
            if (v >= (unsigned)tiles) break;
            __nanosleep(64);
            if (clock64() - t0 > 8000000000ll) __trap(); // seconds: the dispatch-order assumption failed - fail loudly, never hang
        }
        __threadfence();
    }
    __syncthreads();
}
template <int LOGN>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_fwd_fused(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                 NttBatch B, int count, unsigned *done) {
    extern __shared__ u64 sm[];
    constexpr int TILES = (1 << LOGN) / NTT_TILE, LT = LOGN - NTT_TILE_LOG, SA = LOGN - NTT_SB;
    const uint32_t tile = blockIdx.x & (TILES - 1);
    int phase, limb;
    fused_item(blockIdx.x >> LT, count, phase, limb);
    const uint32_t d = B.desc[limb];
    const int qp = d & 0xff;
    if (phase == 0) {
        const PassGeom G{LOGN, SA, 0, NTT_TILE_LOG - SA, NTT_TILE_LOG, 1};
        if (fp_ok(T, qp)) fp_fwd_pass_body<false>(T, src, dst, d, G, 1, 0, sm, tile);
        else if (lazy_ok(T, qp)) fwd_pass_body<true, false>(T, src, dst, d, G, 0, sm, tile);
        else fwd_pass_body<false, false>(T, src, dst, d, G, 0, sm, tile);
        fused_publish(done, limb, TILES);
    } else {
        const PassGeom G{LOGN, NTT_SB, SA, 0, NTT_TILE_LOG, 0};
        fused_wait(done, limb, TILES);
        if (fp_ok(T, qp)) fp_fwd_pass_body<true>(T, dst, dst, d, G, 0, 1, sm, tile);
        else if (lazy_ok(T, qp)) fwd_pass_body<true, true>(T, dst, dst, d, G, 1, sm, tile);
        else fwd_pass_body<false, true>(T, dst, dst, d, G, 1, sm, tile);
        fused_publish(done, limb, TILES);
    }
}
template <int LOGN>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_inv_fused(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                 NttBatch B, int count, unsigned *done, const u64 *__restrict__ custom,
                                                                 const u64 *__restrict__ custom_s) {
    extern __shared__ u64 sm[];
    constexpr int TILES = (1 << LOGN) / NTT_TILE, LT = LOGN - NTT_TILE_LOG, SA = LOGN - NTT_SB;
    const uint32_t tile = blockIdx.x & (TILES - 1);
    int phase, limb;
    fused_item(blockIdx.x >> LT, count, phase, limb);
    const uint32_t d = B.desc[limb];
    const int qp = d & 0xff;
    if (phase == 0) { // the 9 small-stride stages first
        const PassGeom G{LOGN, NTT_SB, SA, 0, NTT_TILE_LOG, 0};
        if (fp_ok(T, qp)) fp_inv_pass_body<false>(T, src, dst, d, G, 1, 0, sm, nullptr, nullptr, tile, limb);
        else if (lazy_ok(T, qp)) inv_pass_body<true, false>(T, src, dst, d, G, 0, sm, nullptr, nullptr, tile, limb);
        else inv_pass_body<false, false>(T, src, dst, d, G, 0, sm, nullptr, nullptr, tile, limb);
        fused_publish(done, limb, TILES);
    } else {
        const PassGeom G{LOGN, SA, 0, NTT_TILE_LOG - SA, NTT_TILE_LOG, 1};
        fused_wait(done, limb, TILES);
        if (fp_ok(T, qp)) fp_inv_pass_body<true>(T, dst, dst, d, G, 0, 1, sm, custom, custom_s, tile, limb);
        else if (lazy_ok(T, qp)) inv_pass_body<true, true>(T, dst, dst, d, G, 1, sm, custom, custom_s, tile, limb);
        else inv_pass_body<false, true>(T, dst, dst, d, G, 1, sm, custom, custom_s, tile, limb);
        fused_publish(done, limb, TILES);
    }
}
// opt-in (FHEB200_NTT_FUSED=1).  Measured on B200 (profiles/r02/ntt_variants.txt): DRAM traffic per limb drops to the algorithmic
// 16n bytes, but 41 limbs take 61.5 us against 55.8 us for the two launches - the passes are bound by latency / issue, not by HBM.
static bool ntt_fused() {
    static const bool on = [] { const char *e = getenv("FHEB200_NTT_FUSED"); return e && *e == '1'; }();
    return on;
}
// per (device, stream) counters: launches on one stream run in order, so one self-cleaning scratch per stream suffices
static unsigned *fused_scratch(cudaStream_t st) {
    static std::mutex mu;
    static std::map<std::pair<int, cudaStream_t>, unsigned *> tab;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(mu);
    auto it = tab.find({dev, st});
    if (it != tab.end()) return it->second;
    unsigned *p = nullptr;
    if (cudaMalloc(&p, NTT_FUSED_MAX * sizeof(unsigned)) != cudaSuccess) throw std::runtime_error("fheb200: cudaMalloc failed (NTT counters)");
    cudaMemset(p, 0, NTT_FUSED_MAX * sizeof(unsigned)); // synchronous, once per stream
    tab[{dev, st}] = p;
    return p;
}

// ---------------------------------------------------------------- persistent passes with next-tile prefetch (n = 2^17)
// ncu on the one-tile-per-CTA kernels: the largest stall is the long scoreboard (a CTA's first global load is on its
// critical path; only the other three CTAs of the SM cover it).  Here a CTA walks a strided list of (limb, tile) items
// and issues the 8 loads of its NEXT tile before it starts the butterflies of the current one, so the HBM / L2 latency
// of a tile is covered by a whole tile of arithmetic of the same warp.  CTAS = resident CTAs per SM the kernel is
// compiled for (4: 64 registers, 3: 85 registers - the 16 prefetch registers fit without spills).
template <bool INV>
__device__ __forceinline__ void tile_prefetch(const u64 *__restrict__ src, uint32_t d, const PassGeom &G, uint32_t tile, int n,
                                              u64 (&r)[8]) {
    const int tid = threadIdx.x, lowbit = G.pass_a ? G.lc : 0, ptop = G.S - 3 + lowbit;
    int p = ptop;
    if (INV) {
        const int c_last = (G.S - 1) % 3 + 1, m = (G.S - c_last) / 3;
        p = m > 0 ? lowbit : ptop;
    }
    const int base = triple_base(tid, p);
    const uint32_t g0 = tile_to_g(G, base, tile);
    const size_t off = (size_t)(d >> 8) * n;
    if (!G.pass_a && p == 0) {
        const ulonglong2 *in = reinterpret_cast<const ulonglong2 *>(src + off + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const ulonglong2 t = in[k];
            r[2 * k] = t.x, r[2 * k + 1] = t.y;
        }
    } else {
        const uint32_t st = g_step(G, p);
#pragma unroll
        for (int k = 0; k < 8; k++) r[k] = src[off + g0 + k * st];
    }
}
template <int LOGN, int S, int S0, int LC, int PA, int CTAS>
__global__ void __launch_bounds__(NTT_TILE / 8, CTAS) ntt_fwd_persist(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                      NttBatch B, int count, int first, int final_reduce) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    constexpr int TILES = (1 << LOGN) / NTT_TILE, LT = LOGN - NTT_TILE_LOG;
    const int total = count * TILES;
    int item = blockIdx.x;
    if (item >= total) return;
    u64 nxt[8];
    tile_prefetch<false>(src, B.desc[item >> LT], G, item & (TILES - 1), T.n, nxt);
    for (; item < total; item += gridDim.x) {
        u64 cur[8];
#pragma unroll
        for (int k = 0; k < 8; k++) cur[k] = nxt[k];
        const int ni = item + gridDim.x;
        if (ni < total) tile_prefetch<false>(src, B.desc[ni >> LT], G, ni & (TILES - 1), T.n, nxt);
        const uint32_t d = B.desc[item >> LT], tile = item & (TILES - 1);
        const int qp = d & 0xff;
        if (fp_ok(T, qp)) fp_fwd_pass_body<false>(T, src, dst, d, G, first, final_reduce, sm, tile, cur);
        else if (lazy_ok(T, qp)) fwd_pass_body<true, false>(T, src, dst, d, G, final_reduce, sm, tile, cur);
        else fwd_pass_body<false, false>(T, src, dst, d, G, final_reduce, sm, tile, cur);
        __syncthreads(); // the tile's last shared-memory reads are done before the next tile's first writes
    }
}
template <int LOGN, int S, int S0, int LC, int PA, int CTAS>
__global__ void __launch_bounds__(NTT_TILE / 8, CTAS) ntt_inv_persist(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                      NttBatch B, int count, int first, int scale,
                                                                      const u64 *__restrict__ custom, const u64 *__restrict__ custom_s) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    constexpr int TILES = (1 << LOGN) / NTT_TILE, LT = LOGN - NTT_TILE_LOG;
    const int total = count * TILES;
    int item = blockIdx.x;
    if (item >= total) return;
    u64 nxt[8];
    tile_prefetch<true>(src, B.desc[item >> LT], G, item & (TILES - 1), T.n, nxt);
    for (; item < total; item += gridDim.x) {
        u64 cur[8];
#pragma unroll
        for (int k = 0; k < 8; k++) cur[k] = nxt[k];
        const int ni = item + gridDim.x;
        if (ni < total) tile_prefetch<true>(src, B.desc[ni >> LT], G, ni & (TILES - 1), T.n, nxt);
        const uint32_t d = B.desc[item >> LT], tile = item & (TILES - 1), bidx = item >> LT;
        const int qp = d & 0xff;
        if (fp_ok(T, qp)) fp_inv_pass_body<false>(T, src, dst, d, G, first, scale, sm, custom, custom_s, tile, bidx, cur);
        else if (lazy_ok(T, qp)) inv_pass_body<true, false>(T, src, dst, d, G, scale, sm, custom, custom_s, tile, bidx, cur);
        else inv_pass_body<false, false>(T, src, dst, d, G, scale, sm, custom, custom_s, tile, bidx, cur);
        __syncthreads();
    }
}
// ---------------------------------------------------------------- rescale fused into the forward NTT (n = 2^17)
// Rescale = inverse NTT of the dropped limb, centred lift onto the remaining limbs, forward NTT of those, then
// out = (in - lifted) * q_last^-1.  The lift is the PROLOGUE of pass A (each CTA reads its 8 x 256 coefficients of the
// dropped limb - 1 MiB per polynomial, L2 resident - and switches modulus in registers) and the last step the EPILOGUE of
// pass B: no lift kernel, no finish kernel, and the lifted polynomial never exists in HBM in coefficient form.
// desc[y] = (slot << 8) | limb with slot = poly * (ell - 1) + limb; `last` holds the dropped limb of poly p (coefficient
// domain) at slot p * ell; `in` has ell limbs per polynomial, tmp / out have ell - 1.
template <int LOGN, int S, int LC>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_rescale_pass_a(DeviceTables T, const u64 *__restrict__ last, u64 *__restrict__ tmp,
                                                                      const uint32_t *__restrict__ desc, int ell) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, 0, LC, NTT_TILE_LOG, 1};
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, poly = (int)(d >> 8) / (ell - 1);
    u64 pre[8];
    tile_prefetch<false>(last, ((uint32_t)(poly * ell) << 8) | (uint32_t)qp, G, blockIdx.x, T.n, pre);
    const ModConst m = T.mc[qp];
    const u64 ql = T.mc[ell - 1].q, half = ql >> 1;
#pragma unroll
    for (int k = 0; k < 8; k++) pre[k] = switch_modulus(pre[k], ql, half, m);
    if (fp_ok(T, qp)) fp_fwd_pass_body<false>(T, last, tmp, d, G, 1, 0, sm, blockIdx.x, pre);
    else if (lazy_ok(T, qp)) fwd_pass_body<true, false>(T, last, tmp, d, G, 0, sm, blockIdx.x, pre);
    else fwd_pass_body<false, false>(T, last, tmp, d, G, 0, sm, blockIdx.x, pre);
}
template <int LOGN, int S0>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_rescale_pass_b(DeviceTables T, const u64 *__restrict__ tmp, u64 *__restrict__ out,
                                                                      const uint32_t *__restrict__ desc, int ell,
                                                                      const u64 *__restrict__ in, LimbScalars qlinv) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, LOGN - S0, S0, 0, NTT_TILE_LOG, 0};
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, poly = (int)(d >> 8) / (ell - 1);
    FwdEpilogue E;
    E.in = in, E.in_off = (size_t)poly * T.n, E.w = qlinv.v[qp], E.ws = qlinv.s[qp]; // in: one more limb per polynomial
    if (fp_ok(T, qp)) fp_fwd_pass_body<false>(T, tmp, out, d, G, 0, 1, sm, blockIdx.x, nullptr, &E);
    else if (lazy_ok(T, qp)) fwd_pass_body<true, false>(T, tmp, out, d, G, 1, sm, blockIdx.x, nullptr, &E);
    else fwd_pass_body<false, false>(T, tmp, out, d, G, 1, sm, blockIdx.x, nullptr, &E);
}
// FHEB200_RESCALE_FUSED=0 keeps the separate lift / finish kernels
bool k_rescale_ntt_fused(const DeviceTables &T, int ell, int npoly, const u64 *last, const u64 *in, u64 *tmp, u64 *out,
                         const uint32_t *d_desc, const LimbScalars &qlinv, cudaStream_t st) {
    static const bool on = [] { const char *e = getenv("FHEB200_RESCALE_FUSED"); return !(e && *e == '0'); }();
    if (!on || T.logn != 17 || ell < 2) return false;
    const int tile = NTT_TILE, count = npoly * (ell - 1);
    const dim3 grid(T.n / tile, count);
    const size_t smem = (size_t)(tile + tile / 8) * sizeof(u64);
    g_launches++, ntt_rescale_pass_a<17, NTT_SA, NTT_TILE_LOG - NTT_SA><<<grid, tile / 8, smem, st>>>(T, last, tmp, d_desc, ell);
    g_launches++, ntt_rescale_pass_b<17, NTT_SA><<<grid, tile / 8, smem, st>>>(T, tmp, out, d_desc, ell, in, qlinv);
    return true;
}

// ---------------------------------------------------------------- ModDown's last step fused into the forward NTT (n = 2^17)
// conv[2][ell][n] (coefficient domain, the base-converted P part) -> forward NTT -> out_b = sigma_g((acc_b - conv_b) P^-1 + add_b):
// pass A is the plain kernel, pass B applies the rest as its epilogue and scatters through the inverse Galois map (coalesced
// reads of acc / add, one scattered store) instead of a separate kernel with three gathered reads.
template <int LOGN, int S0>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_moddown_pass_b(DeviceTables T, u64 *__restrict__ conv, const uint32_t *__restrict__ desc,
                                                                      int ell, int width, const u64 *__restrict__ p_inv,
                                                                      const u64 *__restrict__ p_inv_s, const u64 *__restrict__ acc,
                                                                      const u64 *add0, const u64 *add1, u64 *out0, u64 *out1, uint32_t ginv) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, LOGN - S0, S0, 0, NTT_TILE_LOG, 0};
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, slot = (int)(d >> 8), b = slot / ell; // slot = b * ell + limb, limb == qp
    const size_t n = T.n;
    FwdEpilogue E;
    E.in = acc, E.in_off = (size_t)b * (width - ell) * n; // acc is [2][width][n]: skip the K special limbs of the polynomials before
    E.w = __ldg(p_inv + qp), E.ws = __ldg(p_inv_s + qp);
    const u64 *add = b ? add1 : add0;
    E.add = add ? add + (size_t)qp * n : nullptr;
    E.scatter = (b ? out1 : out0) + (size_t)qp * n;
    E.ginv = ginv, E.logn = LOGN;
    if (fp_ok(T, qp)) fp_fwd_pass_body<false>(T, conv, conv, d, G, 0, 1, sm, blockIdx.x, nullptr, &E);
    else if (lazy_ok(T, qp)) fwd_pass_body<true, false>(T, conv, conv, d, G, 1, sm, blockIdx.x, nullptr, &E);
    else fwd_pass_body<false, false>(T, conv, conv, d, G, 1, sm, blockIdx.x, nullptr, &E);
}
// FHEB200_MODDOWN_FUSED=0 keeps the separate finish kernel
bool k_moddown_ntt_fused(const DeviceTables &T, const DevicePlan &P, u64 *conv, const u64 *acc, const u64 *add0, const u64 *add1,
                         u64 *out0, u64 *out1, uint32_t galois_inverse, cudaStream_t st) {
    static const bool on = [] { const char *e = getenv("FHEB200_MODDOWN_FUSED"); return !(e && *e == '0'); }();
    if (!on || T.logn != 17) return false;
    const int tile = NTT_TILE, count = 2 * P.ell;
    const dim3 grid(T.n / tile, count);
    const size_t smem = (size_t)(tile + tile / 8) * sizeof(u64);
    g_launches++, ntt_fwd_pass_c<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1><<<grid, tile / 8, smem, st>>>(T, conv, conv, P.desc_q2, 1, 0);
    g_launches++, ntt_moddown_pass_b<17, NTT_SA><<<grid, tile / 8, smem, st>>>(T, conv, P.desc_q2, P.ell, P.width, P.p_inv, P.p_inv_s, acc, add0,
                                                                          add1, out0, out1, galois_inverse);
    return true;
}

// FHEB200_NTT_PERSIST = 0 (one tile per CTA), 3 or 4 (persistent CTAs per SM)
static int ntt_persist() {
    static const int v = [] {
        const char *e = getenv("FHEB200_NTT_PERSIST");
        const int x = e ? atoi(e) : NTT_PERSIST_DEFAULT;
        return (x == 3 || x == 4) ? x : 0;
    }();
    return v;
}
static int sm_count() {
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return 148;
    if (!sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    return sms[dev] > 0 ? sms[dev] : 148;
}

// ---------------------------------------------------------------- constants in the parameter bank (n = 2^17)
// Ablation on B200 (profiles/r02/ntt_ablation.txt): the butterflies are free - a pass with the arithmetic, the twiddles and the
// exchanges removed still takes 74 % of the time, and a CTA lives only ~4 us.  Its prologue was a chain of three dependent
// global loads (descriptor -> modulus record -> (q, 1/q)) before the first data load could issue.  Here the batch descriptors
// and the per-modulus constants the FP64 path needs travel as kernel parameters (constant bank): the data loads go out first.
struct NttLimbConsts {
    double q[MAX_LIMBS], qinv[MAX_LIMBS];
    unsigned char kind[MAX_LIMBS]; // 0 = FP64 path, 1 = lazy integer, 2 = strict integer
};
template <int LOGN, int S, int S0, int LC, int PA>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_fwd_pass_k(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                  const NttBatch B, const NttLimbConsts C, int first, int final_reduce) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    const uint32_t d = B.desc[blockIdx.y];
    const int qp = d & 0xff, kind = C.kind[qp];
    if (kind == 0) {
        const double2 qd = make_double2(C.q[qp], C.qinv[qp]);
        fp_fwd_pass_body<false>(T, src, dst, d, G, first, final_reduce, sm, blockIdx.x, nullptr, nullptr, &qd);
    } else if (kind == 1) fwd_pass_body<true, false>(T, src, dst, d, G, final_reduce, sm, blockIdx.x);
    else fwd_pass_body<false, false>(T, src, dst, d, G, final_reduce, sm, blockIdx.x);
}
template <int LOGN, int S, int S0, int LC, int PA>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_inv_pass_k(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                  const NttBatch B, const NttLimbConsts C, int first, int scale,
                                                                  const u64 *__restrict__ custom, const u64 *__restrict__ custom_s) {
    extern __shared__ u64 sm[];
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    const uint32_t d = B.desc[blockIdx.y];
    const int qp = d & 0xff, kind = C.kind[qp];
    if (kind == 0) {
        const double2 qd = make_double2(C.q[qp], C.qinv[qp]);
        fp_inv_pass_body<false>(T, src, dst, d, G, first, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y, nullptr, &qd);
    } else if (kind == 1) inv_pass_body<true, false>(T, src, dst, d, G, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
    else inv_pass_body<false, false>(T, src, dst, d, G, scale, sm, custom, custom_s, blockIdx.x, blockIdx.y);
}
// Two tiles per CTA (FHEB200_NTT_PAIR=1, forward transform): the second tile's input is requested with cp.async (no registers held)
// before the first tile is transformed, so one of the two load latencies of the CTA is hidden behind arithmetic and there are twice
// as many bytes in flight per CTA at its start.  Each thread copies exactly the words of its own first access pattern.
template <int LOGN, int S, int S0, int LC, int PA>
__global__ void __launch_bounds__(NTT_TILE / 8, NTT_CTAS) ntt_fwd_pass_k2(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst,
                                                                   const NttBatch B, const NttLimbConsts C, int first, int final_reduce) {
    extern __shared__ u64 sm[];
    u64 *stage = sm + NTT_PAD_WORDS;
    const PassGeom G{LOGN, S, S0, LC, NTT_TILE_LOG, PA};
    const uint32_t d = B.desc[blockIdx.y];
    const int qp = d & 0xff, kind = C.kind[qp], tid = threadIdx.x;
    const uint32_t t0 = 2 * blockIdx.x, t1 = t0 + 1;
    {
        const int p = S - 3 + (PA ? LC : 0), base = triple_base(tid, p);
        const uint32_t g0 = tile_to_g(G, base, t1), st = g_step(G, p);
        const u64 *g = src + ((size_t)(d >> 8) << LOGN) + g0;
#pragma unroll
        for (int k = 0; k < 8; k++)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(stage + k * (NTT_TILE / 8) + tid)), "l"(g + (size_t)k * st) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    const double2 qd = make_double2(C.q[qp], C.qinv[qp]);
    if (kind == 0) fp_fwd_pass_body<false>(T, src, dst, d, G, first, final_reduce, sm, t0, nullptr, nullptr, &qd);
    else if (kind == 1) fwd_pass_body<true, false>(T, src, dst, d, G, final_reduce, sm, t0);
    else fwd_pass_body<false, false>(T, src, dst, d, G, final_reduce, sm, t0);
    asm volatile("cp.async.wait_all;" ::: "memory");
    u64 pre[8];
#pragma unroll
    for (int k = 0; k < 8; k++) pre[k] = stage[k * (NTT_TILE / 8) + tid];
    __syncthreads(); // the exchange buffer is reused by the second tile
    if (kind == 0) fp_fwd_pass_body<false>(T, src, dst, d, G, first, final_reduce, sm, t1, pre, nullptr, &qd);
    else if (kind == 1) fwd_pass_body<true, false>(T, src, dst, d, G, final_reduce, sm, t1, pre);
    else fwd_pass_body<false, false>(T, src, dst, d, G, final_reduce, sm, t1, pre);
}
static bool ntt_pair() {
    static const bool on = [] { const char *e = getenv("FHEB200_NTT_PAIR"); return e && *e == '1'; }();
    return on;
}
#include "ntt_cluster.cuh"
// FHEB200_NTT_CLUSTER=0 keeps the two-pass kernels everywhere; =k (k >= 1) uses the cluster transform for batches of at least k
// limbs (a cluster owns 8 SMs: small batches fill the GPU better with 64 independent tiles per limb)
static int ntt_cluster_min() {
    static const int v = [] {
        const char *e = getenv("FHEB200_NTT_CLUSTER");
        return e ? atoi(e) : NTT_CLUSTER_MIN_DEFAULT;
    }();
    return v;
}
static bool cluster_ready() {
    static const bool ok = [] {
        return cudaFuncSetAttribute(ntt_fwd_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CLN_FWD_SMEM) == cudaSuccess &&
               cudaFuncSetAttribute(ntt_inv_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CLN_INV_SMEM) == cudaSuccess;
    }();
    return ok;
}
static std::mutex g_consts_mu;
static std::map<const ModConst *, NttLimbConsts> g_consts;
void ntt_register_moduli(const DeviceTables &T, const ModConst *host_mc) {
    NttLimbConsts C;
    memset(&C, 0, sizeof(C));
    for (int i = 0; i < T.F && i < MAX_LIMBS; i++) {
        C.q[i] = (double)host_mc[i].q, C.qinv[i] = (double)(1.0L / (long double)host_mc[i].q); // as DeviceTables::qd
        C.kind[i] = (host_mc[i].bits <= 42 && T.logn <= 17) ? 0 : ((int)host_mc[i].bits + T.logn + 1 <= 63 ? 1 : 2);
    }
    std::lock_guard<std::mutex> lock(g_consts_mu);
    g_consts[T.mc] = C;
}
static bool limb_consts(const DeviceTables &T, NttLimbConsts &out) {
    std::lock_guard<std::mutex> lock(g_consts_mu);
    auto it = g_consts.find(T.mc);
    if (it == g_consts.end()) return false;
    out = it->second;
    return true;
}
// FHEB200_NTT_PARAMS=0: descriptors and constants from global memory (round-1 form)
static bool ntt_params() {
    static const bool on = [] { const char *e = getenv("FHEB200_NTT_PARAMS"); return !(e && *e == '0'); }();
    return on;
}

// host copies of the descriptor arrays (Context::desc registers them): the fused kernels take descriptors by value
static std::mutex g_desc_mu;
static std::map<const uint32_t *, std::vector<uint32_t>> g_desc_host;
void ntt_register_desc(const uint32_t *dev, const uint32_t *host, size_t count) {
    std::lock_guard<std::mutex> lock(g_desc_mu);
    g_desc_host[dev] = std::vector<uint32_t>(host, host + count);
}
static bool host_desc(const uint32_t *dev, int count, std::vector<uint32_t> &out) {
    std::lock_guard<std::mutex> lock(g_desc_mu);
    auto it = g_desc_host.find(dev);
    if (it == g_desc_host.end() || (int)it->second.size() < count) return false;
    out.assign(it->second.begin(), it->second.begin() + count);
    return true;
}

static int ntt_group() { // limbs per launch pair at n = 2^17: everything unless FHEB200_NTT_GROUP says otherwise
    static const int g = [] {
        const char *e = getenv("FHEB200_NTT_GROUP");
        const int v = e ? atoi(e) : 0;
        return v > 0 ? v : 1 << 20;
    }();
    return g;
}
static void passes(const DeviceTables &T, PassGeom &A, PassGeom &B, int &tile) {
    const int logn = T.logn;
    if (logn <= NTT_TILE_LOG) {
        tile = T.n;
        A = PassGeom{logn, 0, 0, 0, 0, 0};
        B = PassGeom{logn, logn, 0, 0, logn, 0};
    } else {
        tile = NTT_TILE;
        const int SB = logn == 17 ? NTT_SB : 9, SA = logn - SB;
        A = PassGeom{logn, SA, 0, NTT_TILE_LOG - SA, NTT_TILE_LOG, 1};
        B = PassGeom{logn, SB, SA, 0, NTT_TILE_LOG, 0};
    }
}

void k_ntt_fwd(const DeviceTables &T, const u64 *src, u64 *dst, const uint32_t *d_desc, int count, cudaStream_t st) {
    if (count <= 0) return;
    PassGeom A, B;
    int tile;
    passes(T, A, B, tile);
    dim3 grid(T.n / tile, count);
    const size_t smem = (size_t)(tile + tile / 8) * sizeof(u64);
    std::vector<uint32_t> hd;
    if (T.logn == 17 && ntt_fused() && host_desc(d_desc, count, hd)) {
        unsigned *done = fused_scratch(st);
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            g_launches++, ntt_fwd_fused<17><<<2 * gc * (T.n / tile), tile / 8, smem, st>>>(T, src, dst, B, gc, done);
        }
        return;
    }
    if (T.logn == 17 && ntt_persist() && host_desc(d_desc, count, hd)) {
        const int pc = ntt_persist();
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            const int gridp = std::min(gc * (T.n / tile), pc * sm_count());
            if (pc == 3) {
                g_launches++, ntt_fwd_persist<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1, 3><<<gridp, tile / 8, smem, st>>>(T, src, dst, B, gc, 1, 0);
                g_launches++, ntt_fwd_persist<17, NTT_SB, NTT_SA, 0, 0, 3><<<gridp, tile / 8, smem, st>>>(T, dst, dst, B, gc, 0, 1);
            } else {
                g_launches++, ntt_fwd_persist<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1, 4><<<gridp, tile / 8, smem, st>>>(T, src, dst, B, gc, 1, 0);
                g_launches++, ntt_fwd_persist<17, NTT_SB, NTT_SA, 0, 0, 4><<<gridp, tile / 8, smem, st>>>(T, dst, dst, B, gc, 0, 1);
            }
        }
        return;
    }
    NttLimbConsts LC_;
    if (T.logn == 17 && ntt_cluster_min() > 0 && count >= ntt_cluster_min() && host_desc(d_desc, count, hd) && limb_consts(T, LC_) &&
        cluster_ready()) {
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            g_launches++, ntt_fwd_cluster<<<dim3(CLN_CTAS, gc), CLN_THREADS, CLN_FWD_SMEM, st>>>(T, src, dst, B, LC_);
        }
        return;
    }
    if (T.logn == 17 && ntt_params() && host_desc(d_desc, count, hd) && limb_consts(T, LC_)) {
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            const dim3 gg(T.n / tile, gc);
            if (ntt_pair()) {
                const dim3 g2(T.n / tile / 2, gc);
                const size_t smem2 = smem + (size_t)tile * sizeof(u64);
                g_launches++, ntt_fwd_pass_k2<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1><<<g2, tile / 8, smem2, st>>>(T, src, dst, B, LC_, 1, 0);
                g_launches++, ntt_fwd_pass_k2<17, NTT_SB, NTT_SA, 0, 0><<<g2, tile / 8, smem2, st>>>(T, dst, dst, B, LC_, 0, 1);
                continue;
            }
            g_launches++, ntt_fwd_pass_k<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1><<<gg, tile / 8, smem, st>>>(T, src, dst, B, LC_, 1, 0);
            g_launches++, ntt_fwd_pass_k<17, NTT_SB, NTT_SA, 0, 0><<<gg, tile / 8, smem, st>>>(T, dst, dst, B, LC_, 0, 1);
        }
        return;
    }
    if (T.logn == 17) {
        // FHEB200_NTT_GROUP = g runs both passes over g limbs at a time so that the intermediate stays in L2.  Measured
        // (profiles/r01/ntt_group_sweep.txt): slower for every g (41 limbs: 55.7 us ungrouped, 67.7 us at g = 24, 104 us at
        // g = 8) - the passes are bound by the DP / issue pipes, not by HBM, and smaller launches only add tails.
        const int group = ntt_group();
        for (int c0 = 0; c0 < count; c0 += group) {
            const dim3 gg(T.n / tile, std::min(group, count - c0));
            g_launches++, ntt_fwd_pass_c<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1><<<gg, tile / 8, smem, st>>>(T, src, dst, d_desc + c0, 1, 0);
            g_launches++, ntt_fwd_pass_c<17, NTT_SB, NTT_SA, 0, 0><<<gg, tile / 8, smem, st>>>(T, dst, dst, d_desc + c0, 0, 1);
        }
        return;
    }
    if (A.S > 0) {
        g_launches++, ntt_fwd_pass<<<grid, tile / 8, smem, st>>>(T, src, dst, d_desc, A, 1, 0);
        src = dst;
    }
    g_launches++, ntt_fwd_pass<<<grid, tile / 8, smem, st>>>(T, src, dst, d_desc, B, A.S == 0, 1);
}
// custom (optional): [count] scalars + [count] Shoup companions replacing n^-1 in the final scaling, one per batch entry
// (ModUp / ModDown fold their (Q_j/q_i)^-1 factors into it)
void k_ntt_inv(const DeviceTables &T, const u64 *src, u64 *dst, const uint32_t *d_desc, int count, cudaStream_t st,
               const u64 *custom) {
    if (count <= 0) return;
    PassGeom A, B;
    int tile;
    passes(T, A, B, tile);
    dim3 grid(T.n / tile, count);
    const size_t smem = (size_t)(tile + tile / 8) * sizeof(u64);
    std::vector<uint32_t> hd;
    if (T.logn == 17 && ntt_fused() && host_desc(d_desc, count, hd)) {
        unsigned *done = fused_scratch(st);
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            g_launches++, ntt_inv_fused<17><<<2 * gc * (T.n / tile), tile / 8, smem, st>>>(
                T, src, dst, B, gc, done, custom ? custom + c0 : nullptr, custom ? custom + count + c0 : nullptr);
        }
        return;
    }
    if (T.logn == 17 && ntt_persist() && host_desc(d_desc, count, hd)) {
        const int pc = ntt_persist();
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            const int gridp = std::min(gc * (T.n / tile), pc * sm_count());
            const u64 *cu = custom ? custom + c0 : nullptr, *cs = custom ? custom + count + c0 : nullptr;
            if (pc == 3) {
                g_launches++, ntt_inv_persist<17, NTT_SB, NTT_SA, 0, 0, 3><<<gridp, tile / 8, smem, st>>>(T, src, dst, B, gc, 1, 0, nullptr, nullptr);
                g_launches++, ntt_inv_persist<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1, 3><<<gridp, tile / 8, smem, st>>>(T, dst, dst, B, gc, 0, 1, cu, cs);
            } else {
                g_launches++, ntt_inv_persist<17, NTT_SB, NTT_SA, 0, 0, 4><<<gridp, tile / 8, smem, st>>>(T, src, dst, B, gc, 1, 0, nullptr, nullptr);
                g_launches++, ntt_inv_persist<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1, 4><<<gridp, tile / 8, smem, st>>>(T, dst, dst, B, gc, 0, 1, cu, cs);
            }
        }
        return;
    }
    NttLimbConsts LC_;
    if (T.logn == 17 && ntt_cluster_min() > 0 && count >= ntt_cluster_min() && host_desc(d_desc, count, hd) && limb_consts(T, LC_) &&
        cluster_ready()) {
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            g_launches++, ntt_inv_cluster<<<dim3(CLN_CTAS, gc), CLN_THREADS, CLN_INV_SMEM, st>>>(
                T, src, dst, B, LC_, custom ? custom + c0 : nullptr, custom ? custom + count + c0 : nullptr);
        }
        return;
    }
    if (T.logn == 17 && ntt_params() && host_desc(d_desc, count, hd) && limb_consts(T, LC_)) {
        for (int c0 = 0; c0 < count; c0 += NTT_FUSED_MAX) {
            const int gc = std::min(NTT_FUSED_MAX, count - c0);
            NttBatch B;
            std::copy(hd.begin() + c0, hd.begin() + c0 + gc, B.desc);
            const dim3 gg(T.n / tile, gc);
            g_launches++, ntt_inv_pass_k<17, NTT_SB, NTT_SA, 0, 0><<<gg, tile / 8, smem, st>>>(T, src, dst, B, LC_, 1, 0, nullptr, nullptr);
            g_launches++, ntt_inv_pass_k<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1><<<gg, tile / 8, smem, st>>>(
                T, dst, dst, B, LC_, 0, 1, custom ? custom + c0 : nullptr, custom ? custom + count + c0 : nullptr);
        }
        return;
    }
    if (T.logn == 17) {
        const int group = ntt_group();
        for (int c0 = 0; c0 < count; c0 += group) {
            const int gc = std::min(group, count - c0);
            const dim3 gg(T.n / tile, gc);
            // custom scalars are indexed by blockIdx.y with the Shoup companions gridDim.y further on: hand each group its
            // own window [c0, c0 + gc) + [count + c0, ...) through the two-pointer form
            g_launches++, ntt_inv_pass_c<17, NTT_SB, NTT_SA, 0, 0><<<gg, tile / 8, smem, st>>>(T, src, dst, d_desc + c0, 1, 0, nullptr, nullptr);
            g_launches++, ntt_inv_pass_c<17, NTT_SA, 0, NTT_TILE_LOG - NTT_SA, 1><<<gg, tile / 8, smem, st>>>(
                T, dst, dst, d_desc + c0, 0, 1, custom ? custom + c0 : nullptr, custom ? custom + count + c0 : nullptr);
        }
        return;
    }
    const u64 *custom_s = custom ? custom + count : nullptr;
    g_launches++, ntt_inv_pass<<<grid, tile / 8, smem, st>>>(T, src, dst, d_desc, B, 1, A.S == 0, custom, custom_s);
    if (A.S > 0) g_launches++, ntt_inv_pass<<<grid, tile / 8, smem, st>>>(T, dst, dst, d_desc, A, 0, 1, custom, custom_s);
}

// ---------------------------------------------------------------- allocation-free forward NTT
// Same fused two-phase kernel as k_ntt_fwd at n = 2^17, but the per-limb dependency counters live in a file-scope device
// array instead of a cudaMalloc'd block, so the call path contains no cudaMalloc / cudaFree / pool allocation at all
// (usable under CUDA graph capture or in harnesses that forbid allocation).  Limits and rules:
//   * at most NTT_STATIC_MAX_LIMBS limbs per call (the array is 1 KiB; every GPU that loads the library holds it until exit);
//   * every call shares the array: it is cleared on the stream before each launch, and two calls that overlap on the GPU
//     (different streams, no ordering between them) race on it - callers must order such calls themselves.
constexpr int NTT_STATIC_MAX_LIMBS = NTT_FUSED_MAX;
static __device__ __align__(256) unsigned g_ntt_static_done[NTT_STATIC_MAX_LIMBS];
int ntt_static_max_limbs() { return NTT_STATIC_MAX_LIMBS; }
// the array's device address on the current device: in host code the symbol's name is only a host-side stand-in, never a
// device pointer, so it is resolved once per device with that device current and cached
static unsigned *ntt_static_scratch() {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static unsigned *addr[MAX_DEV] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) throw std::runtime_error("fheb200: no current CUDA device");
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        unsigned *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_ntt_static_done);
        if (e != cudaSuccess || !p) throw std::runtime_error(std::string("fheb200: cudaGetSymbolAddress(NTT counters): ") + cudaGetErrorString(e));
        addr[dev] = p;
    }
    return addr[dev];
}
void k_ntt_fwd_static(const DeviceTables &T, const u64 *src, u64 *dst, const uint32_t *host_desc, int count, cudaStream_t st) {
    if (T.logn != 17) throw std::invalid_argument("ntt_forward_noalloc: only ring dimension 2^17 (the fused two-phase kernel) is supported");
    if (count < 1 || count > NTT_STATIC_MAX_LIMBS)
        throw std::invalid_argument("ntt_forward_noalloc: 1.." + std::to_string(NTT_STATIC_MAX_LIMBS) + " limbs per call");
    unsigned *done = ntt_static_scratch();
    NttBatch B;
    std::copy(host_desc, host_desc + count, B.desc);
    const int tile = NTT_TILE;
    const size_t smem = (size_t)(tile + tile / 8) * sizeof(u64);
    // the counters must start at zero whatever an earlier (possibly aborted) call left behind
    if (cudaMemsetAsync(done, 0, sizeof(unsigned) * count, st) != cudaSuccess) throw std::runtime_error("fheb200: cudaMemsetAsync failed (NTT counters)");
    g_launches++, ntt_fwd_fused<17><<<2 * count * (T.n / tile), tile / 8, smem, st>>>(T, src, dst, B, count, done);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw std::runtime_error(std::string("fheb200: ntt_fwd_fused launch: ") + cudaGetErrorString(e));
}

} // namespace fheb200
