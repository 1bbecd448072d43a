// ntt_cluster.cuh — whole-limb NTT / iNTT at n = 2^17 on a thread-block cluster: one limb (1 MiB) per cluster of 8 CTAs.
// (included by ntt.cu after the pass bodies; shares their radix-8 chunk functions and index maps)
//
// The two-pass kernels move every limb through HBM twice (read + write per pass: 32n bytes against the algorithmic 16n), and
// the ablation of round 2 (profiles/r02/ntt_ablation.txt) shows that this data movement, not the butterflies, is 74 % of
// their run time.  Here the intermediate never leaves the chip: the limb is the 256 x 512 matrix g = row * 512 + col,
//   forward : CTA c runs pass A (stages 0..7, the column transforms) on columns [64c, 64c + 64) straight from HBM and leaves
//             the result in its own shared memory (130 KiB, "resident");  barrier.cluster;  CTA c runs pass B (stages 8..16,
//             the row transforms) on rows [32c, 32c + 32): each thread's first access pattern reads one word from the
//             resident block of EACH of the 8 CTAs (distributed shared memory, ld.shared::cluster through the generic
//             window), the result goes to HBM as contiguous 64-byte runs (or through the fused epilogue);
//   inverse : the mirror image - pass B on rows [32c, 32c + 32) from contiguous HBM into the resident block (144 KiB),
//             barrier.cluster, pass A on columns [64c, 64c + 64) gathering from the 8 resident blocks (each warp reads one
//             remote CTA), n^-1 (or the caller's per-limb scalar) applied on the way out.
// A last barrier.cluster keeps every CTA alive until its block is no longer read.  1024 threads per CTA = 4 groups of 256,
// each group runs the same 2048-coefficient tile code as the two-pass kernels on its own 18 KiB exchange buffer; the
// groups run independently (one named barrier each); the input of a CTA is fetched up front with asynchronous copies.  HBM traffic per limb: 8n read + 8n written (+ twiddles, L2-resident when
// limbs of equal modulus are adjacent) = the algorithmic bytes.
// Resident layouts (words of 8 bytes), chosen so that both the local stores and the remote loads are bank-conflict free:
//   forward [256 rows][64 cols], row stride 65 (odd: the 4 rows a warp stores to per instruction are 8 rows apart)
//   inverse [32 rows][pad(512)] with pad(e) = e + (e >> 3), row stride 577 (odd, same reason for the remote gathers)
#pragma once
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

constexpr int CLN_CTAS = 8, CLN_GROUPS = 4, CLN_THREADS = 256 * CLN_GROUPS;
constexpr int CLN_FRS = 65, CLN_IRS = 577;
constexpr int CLN_FWD_RES = 256 * CLN_FRS, CLN_INV_RES = 32 * CLN_IRS;
constexpr size_t CLN_FWD_SMEM = (size_t)(CLN_FWD_RES + CLN_GROUPS * NTT_PAD_WORDS) * sizeof(u64);
constexpr size_t CLN_INV_SMEM = (size_t)(CLN_INV_RES + CLN_GROUPS * NTT_PAD_WORDS) * sizeof(u64);
static_assert(NTT_TILE_LOG == 11 && NTT_SB == 9, "the cluster transform is written for 2048-coefficient tiles, 8 + 9 stages");

// the 4 groups of a CTA run independently: each synchronises on its own named barrier (ids 1..4, 256 threads)
__device__ __forceinline__ void grp_sync(int bar) { asm volatile("bar.sync %0, 256;" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void cp_async8(u64 *smem_dst, const u64 *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ PassGeom cl_geom_a() { return PassGeom{17, 8, 0, 3, 11, 1}; }
__device__ __forceinline__ PassGeom cl_geom_b() { return PassGeom{17, 9, 8, 0, 11, 0}; }

// ---------------------------------------------------------------- tile cores: registers in, registers out
// v enters in the first access pattern of the pass (p = S - 3 + lowbit), leaves in the last one (p = lowbit)
template <int PA>
__device__ __forceinline__ void cl_fp_fwd_core(double (&v)[8], const double *__restrict__ tw, double q, double qinv, double *sm,
                                               int tid, uint32_t tile, int bar) {
    const PassGeom G = PA ? cl_geom_a() : cl_geom_b();
    constexpr int S = PA ? 8 : 9, LOW = PA ? 3 : 0, CF = (S - 1) % 3 + 1;
    int p = S - 3 + LOW, base = triple_base(tid, p), s = G.S0;
    fp_fwd_chunk<CF>(v, tw, s, 17, tile_to_g(G, base, tile), q, qinv);
    s += CF;
    bool again = false;
#pragma unroll
    for (int pn = p - CF; pn >= LOW; pn -= 3) {
        if (again) grp_sync(bar);
        again = true;
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        grp_sync(bar);
        p = pn, base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        fp_fwd_chunk<3>(v, tw, s, 17, tile_to_g(G, base, tile), q, qinv);
        s += 3;
    }
}
template <int PA, bool LAZY>
__device__ __forceinline__ void cl_int_fwd_core(u64 (&v)[8], const ulonglong2 *__restrict__ tw, u64 nq, u64 q2, u64 *sm, int tid,
                                                uint32_t tile, int bar) {
    const PassGeom G = PA ? cl_geom_a() : cl_geom_b();
    constexpr int S = PA ? 8 : 9, LOW = PA ? 3 : 0, CF = (S - 1) % 3 + 1;
    int p = S - 3 + LOW, base = triple_base(tid, p), s = G.S0;
    fwd_chunk<CF, LAZY>(v, tw, s, 17, tile_to_g(G, base, tile), nq, q2);
    s += CF;
    bool again = false;
#pragma unroll
    for (int pn = p - CF; pn >= LOW; pn -= 3) {
        if (again) grp_sync(bar);
        again = true;
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        grp_sync(bar);
        p = pn, base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        fwd_chunk<3, LAZY>(v, tw, s, 17, tile_to_g(G, base, tile), nq, q2);
        s += 3;
    }
}
// inverse: v enters at p = lowbit, leaves at p = S - 3 + lowbit.  FP64: `centred` = the values are known to lie in (-q, q)
template <int PA>
__device__ __forceinline__ void cl_fp_inv_core(double (&v)[8], const double *__restrict__ tw, double q, double qinv, double *sm,
                                               int tid, uint32_t tile, int bar) {
    const PassGeom G = PA ? cl_geom_a() : cl_geom_b();
    constexpr int S = PA ? 8 : 9, LOW = PA ? 3 : 0, CL_ = (S - 1) % 3 + 1, PTOP = S - 3 + LOW, M = (S - CL_) / 3;
    int p = LOW, base = triple_base(tid, p), growth = 0;
    uint32_t g0 = tile_to_g(G, base, tile);
#pragma unroll
    for (int i = 0; i < M; i++) {
        if (growth + 3 > 8) {
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = fp_center(v[k], q, qinv);
            growth = 0;
        }
        const int s = G.S0 + S - 1 - (p + 2 - LOW);
        fp_inv_chunk<3>(v, tw, s, 17, g0, q, qinv);
        growth += 3;
        if (i > 0) grp_sync(bar);
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        grp_sync(bar);
        p = i + 1 < M ? p + 3 : PTOP;
        base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        g0 = tile_to_g(G, base, tile);
    }
    if (growth + CL_ > 8) {
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = fp_center(v[k], q, qinv);
    }
    fp_inv_chunk<CL_>(v, tw, G.S0, 17, g0, q, qinv);
}
template <int PA, bool LAZY>
__device__ __forceinline__ void cl_int_inv_core(u64 (&v)[8], const ulonglong2 *__restrict__ tw, u64 q, u64 nq, u64 q2, u64 *sm,
                                                int tid, uint32_t tile, int bar) {
    const PassGeom G = PA ? cl_geom_a() : cl_geom_b();
    constexpr int S = PA ? 8 : 9, LOW = PA ? 3 : 0, CL_ = (S - 1) % 3 + 1, PTOP = S - 3 + LOW, M = (S - CL_) / 3;
    int p = LOW, base = triple_base(tid, p);
    uint32_t g0 = tile_to_g(G, base, tile);
#pragma unroll
    for (int i = 0; i < M; i++) {
        const int s = G.S0 + S - 1 - (p + 2 - LOW);
        inv_chunk<3, LAZY>(v, tw, s, 17, g0, q, nq, q2);
        if (i > 0) grp_sync(bar);
#pragma unroll
        for (int k = 0; k < 8; k++) sm[pad(base | (k << p))] = v[k];
        grp_sync(bar);
        p = i + 1 < M ? p + 3 : PTOP;
        base = triple_base(tid, p);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sm[pad(base | (k << p))];
        g0 = tile_to_g(G, base, tile);
    }
    inv_chunk<CL_, LAZY>(v, tw, G.S0, 17, g0, q, nq, q2);
}

// ---------------------------------------------------------------- forward
// KIND: 0 = FP64 path (moduli <= 42 bits), 1 = lazy integer, 2 = strict integer (60-bit moduli)
template <int KIND>
__device__ __forceinline__ void cl_fwd_limb(const DeviceTables &T, const u64 *__restrict__ src, u64 *__restrict__ dst, uint32_t d,
                                            double qf, double qinv, u64 *res, u64 *scr, const FwdEpilogue *epi) {
    cg::cluster_group cluster = cg::this_cluster();
    const int c = (int)cluster.block_rank(), grp = threadIdx.x >> 8, tid = threadIdx.x & 255, bar = grp + 1;
    const int qp = d & 0xff;
    const size_t off = (size_t)(d >> 8) << 17;
    const PassGeom GA = cl_geom_a(), GB = cl_geom_b();
    const ModConst mc = T.mc[qp];
    const u64 q = mc.q, q2 = q << 1, nq = 0 - q;
    const double *twf = T.twf + ((size_t)qp << 17);
    const ulonglong2 *tw = T.tw + ((size_t)qp << 17);
    // ---- pass A: columns [64c, 64c + 64), two tiles of 8 columns per group, HBM -> resident (in place)
    // both tiles of the group are fetched up front with asynchronous copies (LDGSTS, no registers held): each thread copies
    // exactly the 16 words it will read back in its own first access pattern, so only its own wait_group orders them
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int j = r * CLN_GROUPS + grp;
        const int b0 = triple_base(tid, 8);
        const uint32_t g0 = tile_to_g(GA, b0, 8u * c + j), st = g_step(GA, 8);
#pragma unroll
        for (int k = 0; k < 8; k++) cp_async8(res + ((b0 >> 3) | (k << 5)) * CLN_FRS + j * 8 + (b0 & 7), src + off + g0 + k * st);
        cp_async_commit();
    }
#pragma unroll 1
    for (int r = 0; r < 2; r++) {
        const int j = r * CLN_GROUPS + grp;
        const uint32_t tile = 8u * c + j;
        const int b0 = triple_base(tid, 8);
        if (r) cp_async_wait<0>();
        else cp_async_wait<1>();
        u64 w[8];
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = res[((b0 >> 3) | (k << 5)) * CLN_FRS + j * 8 + (b0 & 7)];
        if (r) grp_sync(bar); // the exchange buffer of the previous tile has been read by every thread of the group
        if (KIND == 0) {
            double v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = fp_from_u64(w[k]);
            cl_fp_fwd_core<1>(v, twf, qf, qinv, reinterpret_cast<double *>(scr), tid, tile, bar);
#pragma unroll
            for (int k = 0; k < 8; k++) w[k] = (u64)__double_as_longlong(v[k]);
        } else {
            cl_int_fwd_core<1, KIND == 1>(w, tw, nq, q2, scr, tid, tile, bar);
        }
        const int b1 = triple_base(tid, 3); // e = b1 | k << 3 : row = e >> 3, column inside the tile = e & 7
#pragma unroll
        for (int k = 0; k < 8; k++) res[((b1 >> 3) | k) * CLN_FRS + j * 8 + (b1 & 7)] = w[k];
    }
    cluster.sync();
    // ---- pass B: rows [32c, 32c + 32), two tiles of 4 rows per group, resident blocks of the 8 CTAs -> HBM
    const u64 *peer[8];
#pragma unroll
    for (int k = 0; k < 8; k++) peer[k] = cluster.map_shared_rank(res, k);
#pragma unroll 1
    for (int r = 0; r < 2; r++) {
        const int j = r * CLN_GROUPS + grp;
        const uint32_t tile = 8u * c + j;
        const int row = 4 * (int)tile + (tid >> 6), col = tid & 63; // first pattern p = 6: e = (tid >> 6) << 9 | k << 6 | col
        u64 w[8];
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = peer[k][row * CLN_FRS + col];
        grp_sync(bar);
        if (KIND == 0) {
            double v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = __longlong_as_double((long long)w[k]);
            cl_fp_fwd_core<0>(v, twf, qf, qinv, reinterpret_cast<double *>(scr), tid, tile, bar);
#pragma unroll
            for (int k = 0; k < 8; k++) w[k] = fp_to_u64(v[k], qf, qinv);
        } else {
            cl_int_fwd_core<0, KIND == 1>(w, tw, nq, q2, scr, tid, tile, bar);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                u64 x = w[k];
                if (KIND == 1) {
                    w[k] = reduce64(x, mc);
                } else {
                    x = x >= q2 ? x - q2 : x;
                    w[k] = x >= q ? x - q : x;
                }
            }
        }
        const uint32_t g0 = (tile << 11) + ((uint32_t)tid << 3);
        if (epi && apply_epilogue(*epi, w, off + g0, g0, q)) continue;
        ulonglong2 *o = reinterpret_cast<ulonglong2 *>(dst + off + g0);
#pragma unroll
        for (int k = 0; k < 4; k++) o[k] = make_ulonglong2(w[2 * k], w[2 * k + 1]);
    }
    cluster.sync(); // no CTA may exit while a peer still reads its resident block
}

__global__ void __cluster_dims__(CLN_CTAS, 1, 1) __launch_bounds__(CLN_THREADS, 1)
    ntt_fwd_cluster(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst, const NttBatch B, const NttLimbConsts C) {
    extern __shared__ __align__(16) u64 cl_sm[];
    u64 *res = cl_sm, *scr = cl_sm + CLN_FWD_RES + (threadIdx.x >> 8) * NTT_PAD_WORDS;
    const uint32_t d = B.desc[blockIdx.y];
    const int qp = d & 0xff, kind = C.kind[qp];
    if (kind == 0) cl_fwd_limb<0>(T, src, dst, d, C.q[qp], C.qinv[qp], res, scr, nullptr);
    else if (kind == 1) cl_fwd_limb<1>(T, src, dst, d, 0.0, 0.0, res, scr, nullptr);
    else cl_fwd_limb<2>(T, src, dst, d, 0.0, 0.0, res, scr, nullptr);
}

// ---------------------------------------------------------------- inverse
template <int KIND>
__device__ __forceinline__ void cl_inv_limb(const DeviceTables &T, const u64 *__restrict__ src, u64 *__restrict__ dst, uint32_t d,
                                            double qf, double qinv, u64 *res, u64 *scr, const u64 *__restrict__ custom,
                                            const u64 *__restrict__ custom_s, uint32_t bidx) {
    cg::cluster_group cluster = cg::this_cluster();
    const int c = (int)cluster.block_rank(), grp = threadIdx.x >> 8, tid = threadIdx.x & 255, bar = grp + 1;
    const int qp = d & 0xff;
    const size_t off = (size_t)(d >> 8) << 17;
    const PassGeom GA = cl_geom_a();
    const u64 q = T.mc[qp].q, q2 = q << 1, nq = 0 - q;
    const double *twf = T.twif + ((size_t)qp << 17);
    const ulonglong2 *tw = T.twi + ((size_t)qp << 17);
    // ---- pass B: rows [32c, 32c + 32) from contiguous HBM -> resident (in place); asynchronous fetch as in the forward kernel
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int j = r * CLN_GROUPS + grp;
        const u64 *in = src + off + ((8u * c + j) << 11) + ((uint32_t)tid << 3);
        u64 *to = res + (4 * j + (tid >> 6)) * CLN_IRS + pad((tid & 63) << 3);
#pragma unroll
        for (int k = 0; k < 8; k++) cp_async8(to + k, in + k);
        cp_async_commit();
    }
#pragma unroll 1
    for (int r = 0; r < 2; r++) {
        const int j = r * CLN_GROUPS + grp;
        const uint32_t tile = 8u * c + j;
        if (r) cp_async_wait<0>();
        else cp_async_wait<1>();
        const u64 *from = res + (4 * j + (tid >> 6)) * CLN_IRS + pad((tid & 63) << 3);
        u64 w[8];
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = from[k];
        if (r) grp_sync(bar);
        if (KIND == 0) {
            double v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = fp_from_u64(w[k]);
            cl_fp_inv_core<0>(v, twf, qf, qinv, reinterpret_cast<double *>(scr), tid, tile, bar);
#pragma unroll
            for (int k = 0; k < 8; k++) w[k] = (u64)__double_as_longlong(v[k]);
        } else {
            cl_int_inv_core<0, KIND == 1>(w, tw, q, nq, q2, scr, tid, tile, bar);
        }
        const int row = 4 * j + (tid >> 6), col = tid & 63; // leaves in p = 6: e = (tid >> 6) << 9 | k << 6 | col
#pragma unroll
        for (int k = 0; k < 8; k++) res[row * CLN_IRS + pad(col | (k << 6))] = w[k];
    }
    cluster.sync();
    // ---- pass A: columns [64c, 64c + 64): rows ((tid >> 3) << 3 | k) live in CTA tid >> 5 (one remote CTA per warp)
    const u64 *peer = cluster.map_shared_rank(res, tid >> 5);
#pragma unroll 1
    for (int r = 0; r < 2; r++) {
        const int j = r * CLN_GROUPS + grp;
        const uint32_t tile = 8u * c + j;
        const int b0 = triple_base(tid, 3);
        const int lrow = (b0 >> 3) & 31, col = 8 * (int)tile + (b0 & 7);
        u64 w[8];
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = peer[(lrow | k) * CLN_IRS + pad(col)];
        grp_sync(bar);
        const int b1 = triple_base(tid, 8);
        const uint32_t g0 = tile_to_g(GA, b1, tile), st = g_step(GA, 8);
        if (KIND == 0) {
            double v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = fp_center(__longlong_as_double((long long)w[k]), qf, qinv);
            cl_fp_inv_core<1>(v, twf, qf, qinv, reinterpret_cast<double *>(scr), tid, tile, bar);
            const double ni = (double)(custom ? custom[bidx] : T.ninv[qp]);
#pragma unroll
            for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = fp_to_u64(fp_mulmod(v[k], ni, qinv, qf), qf, qinv);
        } else {
            cl_int_inv_core<1, KIND == 1>(w, tw, q, nq, q2, scr, tid, tile, bar);
            const u64 ni = custom ? custom[bidx] : T.ninv[qp], nis = custom ? custom_s[bidx] : T.ninv[T.F + qp];
#pragma unroll
            for (int k = 0; k < 8; k++) dst[off + g0 + k * st] = shoup_mul(w[k], ni, nis, q);
        }
    }
    cluster.sync();
}

__global__ void __cluster_dims__(CLN_CTAS, 1, 1) __launch_bounds__(CLN_THREADS, 1)
    ntt_inv_cluster(DeviceTables T, const u64 *__restrict__ src, u64 *__restrict__ dst, const NttBatch B, const NttLimbConsts C,
                    const u64 *__restrict__ custom, const u64 *__restrict__ custom_s) {
    extern __shared__ __align__(16) u64 cl_sm[];
    u64 *res = cl_sm, *scr = cl_sm + CLN_INV_RES + (threadIdx.x >> 8) * NTT_PAD_WORDS;
    const uint32_t d = B.desc[blockIdx.y];
    const int qp = d & 0xff, kind = C.kind[qp];
    if (kind == 0) cl_inv_limb<0>(T, src, dst, d, C.q[qp], C.qinv[qp], res, scr, custom, custom_s, blockIdx.y);
    else if (kind == 1) cl_inv_limb<1>(T, src, dst, d, 0.0, 0.0, res, scr, custom, custom_s, blockIdx.y);
    else cl_inv_limb<2>(T, src, dst, d, 0.0, 0.0, res, scr, custom, custom_s, blockIdx.y);
}
