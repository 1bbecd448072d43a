// kernels.h — launch wrappers of the sm_100a kernels (raw device pointers + stream; no scheme logic).
// Polynomials are [limb][n] u64.  "desc" arrays describe a batch of limbs: desc[b] = (slot << 8) | qp,
// slot = limb position inside the buffer, qp = modulus index (QP order).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include "params.h"

namespace fheb200 {

extern std::atomic<unsigned long long> g_launches; // kernels launched by this library (bench.py gpu_launches)
// ALGORITHMIC bytes of the scheme operations issued so far (operands read once, results written once, every
// intermediate assumed on chip): key switch 8n[4l + 2b(l+K)], HMult+relin 8n[6l + 2b(l+K)], rescale 8n(2l - 1) per
// polynomial, element-wise 8n * 3l per polynomial, leaf sums 8n * l (k + G) per polynomial.  bench.py divides the sum over
// one sort by its run time for the sort-level roofline fraction.
extern std::atomic<unsigned long long> g_algo_bytes;
inline void count_algo(double bytes) { g_algo_bytes += (unsigned long long)bytes; }
// scheme-operation counters (0: rescales, 1: relinearisations, 2: merged relinearise-and-rescale, 3: rotations / hoisted)
extern std::atomic<unsigned long long> g_op_count[4];

struct DeviceTables {            // resident per context
    int logn = 0, n = 0, F = 0;  // F = L + K moduli
    ModConst *mc = nullptr;      // [F]
    ulonglong2 *tw = nullptr;    // forward twiddles  [F][n] as (w, Shoup companion) pairs
    ulonglong2 *twi = nullptr;   // inverse twiddles  [F][n] as (w, Shoup companion) pairs
    double *twf = nullptr;       // forward twiddles as doubles [F][n] (FP64 NTT path, moduli <= 42 bits)
    double *twif = nullptr;      // inverse twiddles as doubles [F][n]
    double2 *qd = nullptr;       // [F] (q, 1/q) as doubles
    // n = 2^17, FP64 path: tw[2^s + T 2^(s-6) + r] = tw[2^s + r] * twa[s - 14][T] for the last three stages s = 14..16 (T = 2048-
    // coefficient tile, r = butterfly group inside the tile): the last pass then reads the 14 KB tile-0 slice of a modulus' table
    // instead of streaming 896 KB of it (ntt.cu: fp_fwd_chunk_tail).  [F][3][64] each; null = not available
    double *twa = nullptr, *twia = nullptr;
    u64 *ninv = nullptr;         // [F] n^-1, [F] Shoup
};

void ntt_register_moduli(const DeviceTables &T, const ModConst *host_mc); // host copy of the modulus records (ntt.cu)

constexpr int MAX_LIMBS = 96;    // upper bound on L+K (by-value kernel parameter blocks)
struct LimbScalars {             // per-limb scalar + Shoup companion, passed by value
    u64 v[MAX_LIMBS];
    u64 s[MAX_LIMBS];
};

// ---- NTT (batched over limbs) ----
// host copy of a descriptor array that lives at `dev` (the fused n = 2^17 kernels take descriptors as kernel parameters)
void ntt_register_desc(const uint32_t *dev, const uint32_t *host, size_t count);
// host copy of the modulus records of a context (the n = 2^17 kernels take (q, 1/q, path) per modulus as kernel parameters)

// src and dst share the slot layout of the descriptors; src == dst transforms in place
void k_ntt_fwd(const DeviceTables &T, const u64 *src, u64 *dst, const uint32_t *d_desc, int count, cudaStream_t st);
void k_ntt_inv(const DeviceTables &T, const u64 *src, u64 *dst, const uint32_t *d_desc, int count, cudaStream_t st,
               const u64 *custom_scale = nullptr);
inline void k_ntt_fwd(const DeviceTables &T, u64 *buf, const uint32_t *d_desc, int count, cudaStream_t st) {
    k_ntt_fwd(T, buf, buf, d_desc, count, st);
}
inline void k_ntt_inv(const DeviceTables &T, u64 *buf, const uint32_t *d_desc, int count, cudaStream_t st) {
    k_ntt_inv(T, buf, buf, d_desc, count, st);
}

// forward NTT at n = 2^17 with no allocation anywhere on the call path: the fused kernel's counters sit in a static device
// array (ntt.cu); host_desc is a HOST array of `count` <= ntt_static_max_limbs() descriptors; src/dst are device pointers
void k_ntt_fwd_static(const DeviceTables &T, const u64 *src, u64 *dst, const uint32_t *host_desc, int count, cudaStream_t st);
int ntt_static_max_limbs();

// ---- elementwise over ell limbs; qp of limb i is qp0 + i ----
enum BinOp { OP_ADD = 0, OP_SUB = 1, OP_MUL = 2 };
void k_binary(const DeviceTables &T, BinOp op, int ell, const u64 *a, const u64 *b, u64 *out, cudaStream_t st);
void k_negate(const DeviceTables &T, int ell, const u64 *a, u64 *out, cudaStream_t st);
void k_mul_scalar(const DeviceTables &T, int ell, const u64 *a, const LimbScalars &s, u64 *out, cudaStream_t st);
void k_add_scalar(const DeviceTables &T, int ell, const u64 *a, const LimbScalars &s, u64 *out, cudaStream_t st);
void k_scalar_mad(const DeviceTables &T, int ell, const u64 *a, const LimbScalars &s, u64 *io, cudaStream_t st); // io += a * s
// out0/out1 += a0/a1 * pt (pt eval domain) ; accumulate = 0 writes, 1 adds
void k_mul_pt(const DeviceTables &T, int ell, const u64 *a0, const u64 *a1, const u64 *pt, u64 *o0, u64 *o1,
              int accumulate, cudaStream_t st);
void k_tensor(const DeviceTables &T, int ell, const u64 *a0, const u64 *a1, const u64 *b0, const u64 *b1, u64 *d0,
              u64 *d1, u64 *d2, cudaStream_t st);
void k_tensor_square(const DeviceTables &T, int ell, const u64 *a0, const u64 *a1, u64 *d0, u64 *d1, u64 *d2,
                     cudaStream_t st);
// G <= 4 weighted sums of `count` polynomials in one pass: out[g][limb] = sum_t in[t][limb] * w[(g * count + t) * ell + limb % ell]
// over limbs2 = npoly * ell limbs; weights are residues pre-split as (lo30, hi30); all pointer tables live in HBM
void k_lincomb(const DeviceTables &T, int ell, int limbs2, int count, int G, const u64 *const *in, const uint2 *w,
               u64 *const *out, cudaStream_t st);
// the same sums on the integer tensor cores (baseconv.cu): G <= 8 outputs per pass, bit-identical results; scratch holds the
// per-limb B fragment tables (lincomb_tc_scratch_words of them), built on the device from w
// (force_ring: -1 = FHEB200_LINCOMB_RING decides [default: ring], 0 = lc_tc, 1 = lc_tc_ring)
void k_lincomb_tc(const DeviceTables &T, int ell, int limbs2, int count, int G, const u64 *const *in, const uint2 *w,
                  u64 *const *out, u64 *scratch, cudaStream_t st, int force_ring = -1);
// (the tcgen05 form keeps [ell][ceil(count / 16)] blocks of 64 rows x 128 bytes = 1024 words instead)
inline size_t lincomb_tc_scratch_words(int ell, int count) {
    const size_t frag = (size_t)ell * ((count + 3) / 4) * 256, umma = (size_t)ell * ((count + 15) / 16) * 1024;
    return frag > umma ? frag : umma;
}
int lc5_mode();
bool k_lincomb_tcgen05(const DeviceTables &T, int ell, int limbs2, int count, int G, const u64 *const *in, const uint2 *w, u64 *const *out,
                       u64 *scratch, cudaStream_t st);
// a[limb][k] <- a[limb][k] mod q_limb for arbitrary 64-bit words (sums of residues from several ranks)
void k_reduce_words(const DeviceTables &T, int ell, u64 *a, cudaStream_t st);
// Galois automorphism (eval domain, bit-reversed order) on `limbs` limbs
void k_automorphism(const DeviceTables &T, uint32_t galois, int limbs, const u64 *in, u64 *out, cudaStream_t st);

// ---- rescale ----
// last: coefficient-domain last limb [n]; tmp[ell-1][n] <- centred lift of last into q_i (then NTT it)
void k_rescale_lift(const DeviceTables &T, int ell, int npoly, const u64 *last_coeff, size_t last_stride, u64 *tmp, cudaStream_t st);
// out[p][i] = (in[p][i] - tmp[p][i]) * q_last^-1 mod q_i for all polys in one launch (in: ell_out + 1 limbs per poly)
void k_rescale_finish(const DeviceTables &T, int ell_out, int npoly, const u64 *in, const u64 *tmp, const LimbScalars &qlinv,
                      u64 *out, cudaStream_t st);

// lift + forward NTT + finish in two launches (ntt.cu, n = 2^17 only; false = not applicable, use the three calls above):
// last = dropped limbs in coefficient form (poly p at slot p * ell), d_desc = the (slot, limb) batch of the ell - 1 kept limbs
bool k_rescale_ntt_fused(const DeviceTables &T, int ell, int npoly, const u64 *last, const u64 *in, u64 *tmp, u64 *out,
                         const uint32_t *d_desc, const LimbScalars &qlinv, cudaStream_t st);

// ---- bootstrapping: ModRaise ----
// src: coefficient-domain limb 0 of each poly (poly p at src + p * src_stride); out[p][ell_out][n] <- the centred integer
// on limbs 0..ell_out-1 (coefficient domain; NTT it afterwards)
void k_modraise_lift(const DeviceTables &T, int ell_out, int npoly, const u64 *src, size_t src_stride, u64 *out, cudaStream_t st);

// ---- hybrid key switching ----
// ---- base conversion on the integer tensor cores (baseconv.cu) ----
struct ConvTable {
    const uint2 *B = nullptr;    // [KS][NT][8 planes][32 lanes] per-lane B fragments (bytes of C[j][t] 2^(8a) mod p_t)
    const int *tslot = nullptr;  // [NT * 8] output limb slot of each target column, -1 = padding
    const uint4 *tconst = nullptr; // [NT * 8] (p lo, p hi, floor(2^(bits+31) / p), bits) of each target column
    const unsigned char *planes = nullptr; // [NT] 6 or 8 byte planes
    const unsigned char *B5 = nullptr;     // the same constants as 16-target chunks in the UMMA K-major layout (baseconv_tcgen05.cu)
    int nsrc = 0, KS = 0, NT = 0;
};
struct ConvInstance {
    ConvTable tab;
    const u64 *src = nullptr;    // nsrc source limbs [nsrc][n], coefficient domain, already times (Q/q_j)^-1
    u64 *dst = nullptr;          // target limb of column c is dst + tslot[c] * n
    const u64 *extra = nullptr;  // optional last source (index nsrc - 1) that lives elsewhere: one limb [n] (merged ModDown: the
                                 // overflow count of the lift, Context::moddown_plan)
    const u64 *copy_src = nullptr; // copy_limbs limbs copied verbatim (ModUp: the digit's own limbs)
    u64 *copy_dst = nullptr;
    int copy_limbs = 0;
};
struct ConvArgs {
    ConvInstance inst[8];
};
void k_base_convert(const DeviceTables &T, const ConvArgs &A, int count, int KS, cudaStream_t st);

// tcgen05 / TMEM form of the same conversion (baseconv_tcgen05.cu, FHEB200_BC_TCGEN05=1); false = shape not covered
bool bc5_enabled();
unsigned char *bc5_make_table(const uint2 *frag, int KS, int NT);
bool k_base_convert_tcgen05(const DeviceTables &T, const ConvArgs &A, int count, int KS, cudaStream_t st);

struct DevicePlan {              // device copy of a LevelPlan
    int ell = 0, beta = 0, width = 0, K = 0;
    int *qp = nullptr;           // [width]
    int dstart[8] = {0}, dsize[8] = {0};
    u64 *hat_inv[8] = {nullptr}, *hat_inv_s[8] = {nullptr}; // per digit
    uint2 *hat2[8] = {nullptr};  // per digit: [width][dsize rounded up to even] (lo30, hi30) of Q_j/q_i mod p_t
    u64 *p_hat_inv = nullptr, *p_hat_inv_s = nullptr, *p_inv = nullptr, *p_inv_s = nullptr;
    uint2 *p_hat2 = nullptr;     // [ell][K rounded up to even] (lo30, hi30) of P/p_i mod q_t
    uint32_t *desc_q = nullptr;        // NTT batch: Q limbs 0..ell-1 in place
    uint32_t *desc_modup = nullptr;    // NTT batch over digits buffer, skipping own-digit limbs
    int n_modup = 0;
    uint32_t *desc_p2 = nullptr;       // iNTT batch: P limbs of two stacked accumulators [2][width]
    uint32_t *desc_q2 = nullptr;       // NTT batch: Q limbs of two stacked [2][ell] buffers
    ConvTable up[8], down;             // tensor-core base conversion tables per digit / for ModDown
    int exact_lift = 0;                // merged plan: down has one more source, u = round(sum_j y_j / r_j), with constant -R mod q_i
    double *r_recip = nullptr;         // [K] 1 / r_j of the special limbs (exact_lift)
    u64 *up_scale = nullptr;           // [ell] + [ell] Shoup: n^-1 (Q_j/q_i)^-1 mod q_i, final scaling of ModUp's inverse NTT
    u64 *down_scale = nullptr;         // [2K] + [2K] Shoup: n^-1 (P/p_k)^-1 mod p_k for the desc_p2 batch
};
// digits[j][t][k] for t outside digit j <- base conversion of coef (coefficient domain, ell limbs);
// own-digit limbs <- copy of c_eval.
void k_modup(const DeviceTables &T, const DevicePlan &P, const u64 *coef, const u64 *c_eval, u64 *digits,
             cudaStream_t st);
// tensor-core variants (baseconv.cu): inputs pre-scaled by the inverse NTT (DevicePlan::up_scale / down_scale)
void k_modup_tc(const DeviceTables &T, const DevicePlan &P, const u64 *coef, const u64 *c_eval, u64 *digits, cudaStream_t st);
void k_moddown_conv_tc(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *conv, cudaStream_t st, const u64 *extra = nullptr);
// u_b[x] = round(sum_j y_j[x] / r_j) over the K special limbs of accumulator b (coefficient domain, already times (R/r_j)^-1):
// how many multiples of R the fast base conversion of those residues overshoots the centred representative by (Halevi-Polyakov-Shoup)
void k_lift_overflow(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *u, cudaStream_t st);
// acc[2][width][n] = sum_j digits[j] * evk[j][{b,a}]   (evk stored at full chain: [digit][2][F][n])
void k_inner_product(const DeviceTables &T, const DevicePlan &P, const u64 *digits, const u64 *evk, u64 *acc,
                     cudaStream_t st);
// conv[2][ell][n] (coefficient domain) <- base conversion of the (already inverse-NTT'd) P limbs of acc
void k_moddown_conv(const DeviceTables &T, const DevicePlan &P, const u64 *acc, u64 *conv, cudaStream_t st);
// out_b = (acc_b[Q part] - conv_b) * P^-1 (+ add0 on b = 0 if non-null); optional automorphism on the way out
void k_moddown_finish(const DeviceTables &T, const DevicePlan &P, const u64 *acc, const u64 *conv, const u64 *add0,
                      const u64 *add1, u64 *out0, u64 *out1, uint32_t galois, cudaStream_t st);

// forward NTT of conv + k_moddown_finish in two launches (ntt.cu, n = 2^17 only; false = not applicable)
bool k_moddown_ntt_fused(const DeviceTables &T, const DevicePlan &P, u64 *conv, const u64 *acc, const u64 *add0, const u64 *add1,
                         u64 *out0, u64 *out1, uint32_t galois_inverse, cudaStream_t st);

// double hoisting (bootstrapping's linear transforms): S_b[t][p] += pt[t][p] * v_b[t][sigma_g(p)] over the level limbs t of
// the extended basis (Q 0..ell-1, then P), with v_0 = acc_0 + cP_0, v_1 = acc_1 (acc = an evk inner product) or, when acc
// is null, v_b = cP_b; cP = P * c_b on the Q limbs ([2][ell][n]).  The sum is brought back to Q by one ModDown per stage.
void k_dh_accumulate(const DeviceTables &T, const DevicePlan &P, const u64 *acc, const u64 *cP, const u64 *pt, u64 *S,
                     uint32_t galois, cudaStream_t st);

// ---- samplers / encode helpers ----
void k_sample_uniform(const DeviceTables &T, const RngState &R, u64 stream, int qp0, int limbs, u64 *out, cudaStream_t st);
enum SmallKind { SMALL_TERNARY = 0, SMALL_GAUSS = 1 };
void k_sample_small(const DeviceTables &T, SmallKind kind, const RngState &R, u64 stream, int *out, cudaStream_t st);
// signed small ints [n] -> residues for the limbs listed in desc (slot,qp); coefficient domain
void k_expand_small(const DeviceTables &T, const int *small, u64 *out, const uint32_t *d_desc, int count,
                    cudaStream_t st);
// 128-bit signed coefficients (lo,hi pairs) at stride `gap` -> residues on Q limbs 0..ell-1, zero elsewhere
void k_expand_i128(const DeviceTables &T, const u64 *coef_lohi, int ncoef, int gap, int ell, u64 *out,
                   cudaStream_t st);
// device CKKS encode: in-place special inverse FFT of v[size], then words (lo,hi pairs) <- round(v * scale / size) in the
// layout k_expand_i128 consumes (real parts first, imaginary parts after); ksi = exp(2 pi i k / 4 size), rot = 5^j mod 4 size
void k_encode(const DeviceTables &T, double2 *v, const double *real_in, int nv, int size, double scale, const double2 *ksi,
              const int *rot, int ell, u64 *out, cudaStream_t st, int extra_qp0 = 0, int extra = 0);
// (extra > 0: `extra` further limbs after the ell Q limbs, reduced modulo moduli extra_qp0, extra_qp0 + 1, ...)
// b = e - a*s_new + [limb in digit] * pmod * s_old   over F limbs (one digit of a switching key)
void k_evk_finish(const DeviceTables &T, int F, const u64 *a, const u64 *e, const u64 *s_new, const u64 *s_old,
                  const LimbScalars &pmod, u64 *b, cudaStream_t st);
// c0 = pkb*v + e0 + pt ; c1 = pka*v + e1
void k_encrypt_finish(const DeviceTables &T, int ell, const u64 *pkb, const u64 *pka, const u64 *v, const u64 *e0,
                      const u64 *e1, const u64 *pt, u64 *c0, u64 *c1, cudaStream_t st);
// m = c0 + c1*s on k limbs
void k_decrypt_core(const DeviceTables &T, int k, const u64 *c0, const u64 *c1, const u64 *s, u64 *m, cudaStream_t st);

} // namespace fheb200
