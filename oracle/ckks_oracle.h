/*
 * oracle/ckks_oracle.h — C API of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under fhe_sorting_b200/ may include, link or call this.
 * Allowed users: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.
 *
 * What it restates: the RNS-CKKS arithmetic that the reference (oksuman/FHE-Sorting) obtains from
 * OpenFHE through lbcrypto::CryptoContext<DCRTPoly>.  OpenFHE itself is
 * an unpinned, un-vendored find_package dependency (reference CMakeLists.txt:35) that is absent from
 * this image, and the reference's tests hold no RNS-level golden vectors, so:
 *
 *      PARITY UNPINNED at the integer-RNS level against OpenFHE.
 *
 * The oracle is anchored instead on (i) the published algorithms (CKKS-RNS: Cheon-Han-Kim-Kim-Song
 * 2018; hybrid key switching: Han-Ki 2020; FLEXIBLEAUTO(EXT) rescaling: Kim-Papadimitriou-Polyakov
 * 2020), (ii) size-independent algebraic properties (NTT vs schoolbook negacyclic convolution,
 * automorphism vs coefficient-domain X -> X^k, decrypt(keyswitch) == decrypt, ...) and (iii) the
 * reference's decrypted-value known-answer vectors (tests/golden/).
 *
 * All polynomial data are arrays of uint64 limbs, layout [limb][n], limb i reduced mod modulus i.
 * "QP index" means: 0..L-1 are the Q primes (q_0 first/60-bit, ..., q_{L-1} extra), L..L+K-1 the P primes.
 */
#ifndef CKKS_ORACLE_H
#define CKKS_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_ctx orc_ctx;

/* context: ring n = 2^logn, multiplicative depth, scaling-mod bits, first-mod bits, #digits (dnum). */
orc_ctx *orc_ctx_create(int logn, int depth, int scale_bits, int first_bits, int dnum);
void orc_ctx_free(orc_ctx *);
int orc_n(const orc_ctx *);
int orc_L(const orc_ctx *);     /* number of Q primes = depth + 2 (FLEXIBLEAUTOEXT extra level) */
int orc_K(const orc_ctx *);     /* number of P primes */
int orc_alpha(const orc_ctx *); /* limbs per digit */
uint64_t orc_modulus(const orc_ctx *, int qp_index);
uint64_t orc_psi(const orc_ctx *, int qp_index); /* primitive 2n-th root used by the NTT */
double orc_scale_real(const orc_ctx *, int level);    /* sf[level]     */
double orc_scale_big(const orc_ctx *, int level);     /* sfBig[level]  */
void orc_set_threads(int nthreads);                   /* OpenMP threads for the limb loops */

/* ---- a1/a3/a4 building blocks, all in place on one limb or a [limbs][n] block ---- */
void orc_ntt_fwd(const orc_ctx *, int qp_index, uint64_t *a);
void orc_ntt_inv(const orc_ctx *, int qp_index, uint64_t *a);
/* eval-domain Galois automorphism X -> X^k (k odd) on `limbs` limbs (any modulus: pure permutation) */
void orc_automorphism_eval(const orc_ctx *, uint32_t k, int limbs, const uint64_t *in, uint64_t *out);
/* coefficient-domain X -> X^k with sign flips, one limb (used to cross-check the eval-domain map) */
void orc_automorphism_coeff(const orc_ctx *, uint32_t k, int qp_index, const uint64_t *in, uint64_t *out);
uint32_t orc_rot_to_galois(const orc_ctx *, int rot_index); /* 5^i mod 2n, i<0 -> inverse */

/* elementwise on `ell` Q-limbs */
void orc_add(const orc_ctx *, int ell, const uint64_t *a, const uint64_t *b, uint64_t *out);
void orc_sub(const orc_ctx *, int ell, const uint64_t *a, const uint64_t *b, uint64_t *out);
void orc_mul(const orc_ctx *, int ell, const uint64_t *a, const uint64_t *b, uint64_t *out);
/* per-limb scalar ops: s[i] is already reduced mod q_i */
void orc_mul_scalar(const orc_ctx *, int ell, const uint64_t *a, const uint64_t *s, uint64_t *out);
void orc_add_scalar(const orc_ctx *, int ell, const uint64_t *a, const uint64_t *s, uint64_t *out);
/* tensor product of (a0,a1) x (b0,b1) -> (d0,d1,d2), each ell limbs, eval domain */
void orc_tensor(const orc_ctx *, int ell, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0,
                const uint64_t *b1, uint64_t *d0, uint64_t *d1, uint64_t *d2);
/* rescale (drop last limb and scale, rounding to nearest): in ell limbs -> out ell-1 limbs, eval domain */
void orc_rescale(const orc_ctx *, int ell, const uint64_t *in, uint64_t *out);

/* ---- hybrid key switching ---- */
/* evk layout: [digit][2: 0=b,1=a][L+K limbs (QP index)][n], generated at the full chain. */
uint64_t orc_evk_words(const orc_ctx *);
/* ModUp: c (ell limbs, eval domain) -> digits [beta][ell+K][n] (eval domain; limb order: Q 0..ell-1, then P) */
int orc_beta(const orc_ctx *, int ell);
void orc_modup(const orc_ctx *, int ell, const uint64_t *c, uint64_t *digits);
/* inner product of digits with evk -> acc0, acc1 over ell+K limbs */
void orc_inner_product(const orc_ctx *, int ell, const uint64_t *digits, const uint64_t *evk, uint64_t *acc0,
                       uint64_t *acc1);
/* ModDown: acc (ell+K limbs, eval) -> out (ell limbs, eval) */
void orc_moddown(const orc_ctx *, int ell, const uint64_t *acc, uint64_t *out);
/* full key switch of poly c under evk: (ks0, ks1), each ell limbs */
void orc_keyswitch(const orc_ctx *, int ell, const uint64_t *c, const uint64_t *evk, uint64_t *ks0, uint64_t *ks1);
/* EvalRotate core on (c0,c1): key switch c1, add c0, then automorphism k on both (OpenFHE order) */
void orc_rotate(const orc_ctx *, int ell, uint32_t k, const uint64_t *c0, const uint64_t *c1, const uint64_t *evk,
                uint64_t *o0, uint64_t *o1);
/* hoisted: digits from orc_modup(c1); equals orc_rotate bit for bit */
void orc_rotate_hoisted(const orc_ctx *, int ell, uint32_t k, const uint64_t *c0, const uint64_t *digits,
                        const uint64_t *evk, uint64_t *o0, uint64_t *o1);
/* EvalMult core + relinearise: (a0,a1)x(b0,b1) -> (o0,o1), no rescale */
void orc_mult_relin(const orc_ctx *, int ell, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0,
                    const uint64_t *b1, const uint64_t *evk_mult, uint64_t *o0, uint64_t *o1);

/* ---- deterministic samplers / keys (counter-based PRNG; spec in DESIGN.md "Randomness") ---- */
uint64_t orc_rng_u64(uint64_t seed, uint64_t stream, uint64_t idx, uint64_t word);
void orc_sample_uniform(const orc_ctx *, uint64_t seed, uint64_t stream, int qp_index, uint64_t *out);
void orc_sample_gauss(const orc_ctx *, uint64_t seed, uint64_t stream, int64_t *out);   /* n signed ints */
void orc_sample_ternary(const orc_ctx *, uint64_t seed, uint64_t stream, int64_t *out); /* n in {-1,0,1} */
/* secret key in eval domain over all L+K limbs */
void orc_keygen_secret(const orc_ctx *, uint64_t seed, uint64_t *s_qp);
/* public key (b,a) over the L Q-limbs, eval domain: b = -a s + e */
void orc_keygen_public(const orc_ctx *, uint64_t seed, const uint64_t *s_qp, uint64_t *pk_b, uint64_t *pk_a);
/* evaluation key switching from s_old (L+K limbs eval) to s. key_id domain-separates the randomness. */
void orc_keygen_switch(const orc_ctx *, uint64_t seed, uint64_t key_id, const uint64_t *s_old_qp,
                       const uint64_t *s_qp, uint64_t *evk);
void orc_keygen_mult(const orc_ctx *, uint64_t seed, const uint64_t *s_qp, uint64_t *evk);
void orc_keygen_rot(const orc_ctx *, uint64_t seed, uint32_t k, const uint64_t *s_qp, uint64_t *evk);

/* ---- encode / encrypt / decrypt / decode ---- */
/* encode `slots` real values at scale `scale` onto ell limbs (eval domain). */
void orc_encode(const orc_ctx *, const double *vals, int nvals, int slots, double scale, int ell, uint64_t *pt);
/* public-key encryption of pt (ell limbs): nonce domain-separates the randomness */
void orc_encrypt(const orc_ctx *, uint64_t seed, uint64_t nonce, int ell, const uint64_t *pk_b, const uint64_t *pk_a,
                 const uint64_t *pt, uint64_t *c0, uint64_t *c1);
/* decrypt + decode to `slots` real values (uses the first min(ell,3) limbs; see .cpp) */
void orc_decrypt(const orc_ctx *, int ell, const uint64_t *c0, const uint64_t *c1, const uint64_t *s_qp, int slots,
                 double scale, double *out_real, double *out_imag);

/* plain-double helpers that the reference's build-time generators rely on (utils/generate_cheb_*.cpp) */
void orc_cheb_coefficients(double (*f)(double, void *), void *user, double a, double b, int degree, double *out);

#ifdef __cplusplus
}
#endif
#endif
