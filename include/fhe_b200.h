/*
 * fhe_b200.h — C ABI of the B200 RNS-CKKS engine (libfheb200.so).
 *
 * This is the drop-in boundary for the hot path of oksuman/FHE-Sorting: the subset of OpenFHE's
 * lbcrypto::CryptoContext<DCRTPoly> API that the reference's host code calls.  Each entry
 * point names the OpenFHE method it replaces and one reference call site (paths relative to the reference
 * repository).  The header-only C++ shim under fhe_sorting_b200/shim/ re-exposes these as namespace
 * lbcrypto so that src/sort_algo.h, rotation.h, sign.cpp, comparison.cpp, encryption.cpp compile against it
 * (binding shown in INTEGRATION.md).
 *
 * Conventions: every function returns 0 on success, non-zero on failure; fheb200_last_error() returns a
 * thread-local message.  Handles are reference counted; *_release drops one reference.  Results are fresh
 * handles (the reference's *InPlace calls are expressed in the shim as "replace the handle").
 * Plain pointers and sizes only; no CUDA or torch types (streams travel as void*).
 * There is no CPU fallback: context creation fails when no CUDA device is present.
 */
#ifndef FHE_B200_H
#define FHE_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct fheb200_ctx fheb200_ctx;   /* lbcrypto::CryptoContext<DCRTPoly>          */
typedef struct fheb200_keys fheb200_keys; /* lbcrypto::KeyPair<DCRTPoly> (pk [+ sk])     */
typedef struct fheb200_ct fheb200_ct;     /* lbcrypto::Ciphertext<DCRTPoly>              */
typedef struct fheb200_pt fheb200_pt;     /* lbcrypto::Plaintext                         */
typedef struct fheb200_pre fheb200_pre;   /* std::shared_ptr<std::vector<DCRTPoly>> from EvalFastRotationPrecompute */

const char *fheb200_last_error(void);
const char *fheb200_version(void);
/* number of CUDA kernels this library has launched in the process so far */
unsigned long long fheb200_launch_count(void);
/* algorithmic bytes (SURVEY.md 8d formulas) of the scheme operations issued by this process so far: roofline numerator */
unsigned long long fheb200_algo_bytes(void);
/* process-wide scheme-operation counters for profiling: 0 rescales, 1 relinearisations (ModDown by P), 2 merged relinearise-and-rescale
 * divisions, 3 key-switched rotations (plain and hoisted) */
unsigned long long fheb200_op_count(int which);

/* ---- context: CCParams<CryptoContextCKKSRNS> setters + GenCryptoContext (tests/DirectSortTest.cpp:24-41) ---- */
typedef struct {
    int log_ring_dim;    /* SetRingDim(1 << log_ring_dim)            */
    int mult_depth;      /* SetMultiplicativeDepth                   */
    int scaling_mod_size;/* SetScalingModSize                        */
    int first_mod_size;  /* SetFirstModSize (OpenFHE default 60)     */
    int batch_size;      /* SetBatchSize (0 = ring_dim / 2)          */
    int num_large_digits;/* OpenFHE default 3                        */
    int device;          /* CUDA device ordinal                      */
} fheb200_params;
int fheb200_ctx_create(const fheb200_params *p, fheb200_ctx **out);
int fheb200_ctx_release(fheb200_ctx *ctx);
/* run every later operation of this context on an existing CUDA stream (cudaStream_t passed as void*) */
int fheb200_ctx_set_stream(fheb200_ctx *ctx, void *cuda_stream);
int fheb200_ctx_sync(fheb200_ctx *ctx);
/* Concurrent streams for independent sub-computations (the B batches of a DirectSort phase; not in OpenFHE, whose callers
 * use OpenMP threads instead - src/sort_algo.h:475, 713).  fork(k): k side streams continue from the main stream's
 * current point; use_stream(i): every following call runs on side stream i (i < 0: the main stream); join(): the main
 * stream waits for all of them.  Objects made inside the region stay on the stream that made them until join(); objects
 * made before fork() are read-only inside. */
int fheb200_ctx_fork(fheb200_ctx *ctx, int k);
int fheb200_ctx_use_stream(fheb200_ctx *ctx, int i);
int fheb200_ctx_join(fheb200_ctx *ctx);
/* Host threads inside a fork()ed region (what the reference's `#pragma omp parallel for` over batches becomes on one GPU): after
 * bind_thread(i) every call the CALLING THREAD makes on this context runs on side stream i (i < 0: the main stream), whatever other
 * threads do; unbind_thread() returns the thread to the fork() / use_stream() selection.  One launching thread cannot run more than
 * a launch queue (~1000 kernels) ahead of the GPU, so batch chains longer than that only overlap when several threads feed them. */
int fheb200_ctx_bind_thread(fheb200_ctx *ctx, int i);
int fheb200_ctx_unbind_thread(fheb200_ctx *ctx);
/* GetRingDimension / GetCyclotomicOrder (src/rotation.h:262, src/sort_algo.h:80) and chain geometry */
int fheb200_ctx_info(fheb200_ctx *ctx, int *ring_dim, int *num_q, int *num_p, int *digit_size, int *batch);
int fheb200_ctx_modulus(fheb200_ctx *ctx, int qp_index, uint64_t *q, uint64_t *psi);
int fheb200_ctx_scaling_factor(fheb200_ctx *ctx, int level, double *sf, double *sf_big);
int fheb200_ctx_key_bytes(fheb200_ctx *ctx, uint64_t *bytes);
/* host-only: the modulus chain for a parameter set, without touching a GPU (used by CPU-side tests) */
int fheb200_params_chain(const fheb200_params *p, uint64_t *moduli, uint64_t *psi, int capacity, int *num_q, int *num_p,
                         int *digit_size);

/* ---- keys: KeyGen / EvalMultKeyGen / EvalRotateKeyGen (tests/DirectSortTest.cpp:43-47) ---- */
/* Randomness (reference: OpenFHE's cryptographic PRNG behind KeyGen / Encrypt, tests/DirectSortTest.cpp:43).  By default
 * (seed = 0) all key material and encryption noise comes from ChaCha20 keyed with 2 x 256 bits of OS entropy drawn per key
 * pair; secret material and the public uniform polynomials use independent keys.  seed != 0 selects a 64-bit seeded,
 * NON-cryptographic generator that the CPU oracle shares - TEST ONLY, refused unless allow_deterministic_keys(ctx, 1) was
 * called first.  get/set_rng_keys copy the 64 bytes of key material so that the ranks of one job can replicate keys
 * (set before KeyGen on every rank; treat the bytes like the secret key). */
int fheb200_ctx_allow_deterministic_keys(fheb200_ctx *ctx, int allow);
int fheb200_ctx_get_rng_keys(fheb200_ctx *ctx, unsigned char *keys64);
int fheb200_ctx_set_rng_keys(fheb200_ctx *ctx, const unsigned char *keys64);
int fheb200_keygen(fheb200_ctx *ctx, uint64_t seed, fheb200_keys **out);
/* first 128 bits of the ChaCha20 block (key, counter = index, nonce = stream) as the samplers draw them; host-only, for tests */
int fheb200_rng_block(const uint32_t *key8, uint64_t stream, uint64_t index, uint64_t *out2);
int fheb200_keys_release(fheb200_keys *k);
int fheb200_eval_mult_keygen(fheb200_ctx *ctx, fheb200_keys *k);
int fheb200_eval_rotate_keygen(fheb200_ctx *ctx, fheb200_keys *k, const int *indices, size_t count);

/* ---- encode / encrypt / decrypt ---- */
/* MakeCKKSPackedPlaintext(values, noiseScaleDeg, level, nullptr, slots) (src/sort_algo.h:317,341,609,706) */
int fheb200_make_packed_plaintext(fheb200_ctx *ctx, const double *values, size_t count, int noise_scale_deg, int level,
                                  int slots, fheb200_pt **out);
/* Engine extension: the plaintext whose slot i is slot i + index of `pt` (same level, scale and slot count).  Its
 * encoding is the Galois automorphism X -> X^(5^index) of pt's encoding - one permutation kernel instead of an FFT and
 * one NTT per limb.  The sort drivers build their N-wide block masks (src/sort_algo.h:339-342, 571-574) this way. */
int fheb200_pt_rotate(fheb200_ctx *ctx, fheb200_pt *pt, int index, fheb200_pt **out);
int fheb200_pt_release(fheb200_pt *pt);
/* GetRealPackedValue / SetLength / GetLogPrecision (src/encryption.cpp:20, src/encryption.h:41) */
int fheb200_pt_values(fheb200_pt *pt, double *out, size_t capacity, size_t *count);
int fheb200_pt_set_length(fheb200_pt *pt, size_t length);
int fheb200_pt_log_precision(fheb200_pt *pt, double *bits);
/* Encrypt(publicKey, plaintext) (src/encryption.cpp:10) ; Decrypt(secretKey, ct, &pt) (src/encryption.cpp:19) */
int fheb200_encrypt(fheb200_ctx *ctx, fheb200_keys *k, fheb200_pt *pt, fheb200_ct **out);
int fheb200_decrypt(fheb200_ctx *ctx, fheb200_keys *k, fheb200_ct *ct, fheb200_pt **out);

/* ---- ciphertext object: Clone / GetLevel / GetSlots / SetSlots (src/sort_algo.h:429,501,614) ---- */
int fheb200_ct_release(fheb200_ct *ct);
int fheb200_ct_clone(fheb200_ctx *ctx, fheb200_ct *ct, fheb200_ct **out);
int fheb200_ct_info(fheb200_ct *ct, int *level, int *noise_scale_deg, int *slots, int *num_limbs, double *scale);
int fheb200_ct_set_slots(fheb200_ct *ct, int slots);

/* ---- arithmetic ---- */
/* EvalAdd / EvalSub / EvalAddInPlace(ct, ct) (src/sort_algo.h:454,495 ; src/comparison.cpp:13) */
int fheb200_eval_add(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct *b, fheb200_ct **out);
int fheb200_eval_sub(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct *b, fheb200_ct **out);
/* EvalAdd(ct, pt) / EvalSub(ct, pt) / EvalSub(pt, ct) (src/sort_algo.h:611,622,708,720) */
int fheb200_eval_add_plain(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out);
int fheb200_eval_sub_plain(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out);
/* EvalAdd(ct, double) / EvalSub(ct, double) (src/comparison.cpp:19,29-30 ; src/sort_algo.h:504) */
int fheb200_eval_add_const(fheb200_ctx *ctx, fheb200_ct *a, double c, fheb200_ct **out);
/* EvalNegate (src/mehp24/mehp24_utils.h) */
int fheb200_eval_negate(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out);
/* EvalAddMany (src/sort_algo.h:364) */
int fheb200_eval_add_many(fheb200_ctx *ctx, fheb200_ct *const *cts, size_t count, fheb200_ct **out);
/* EvalMult(ct, ct) / EvalMultAndRelinearize / EvalSquare when a == b (src/sign.cpp:23-33 ; src/sort_algo.h:730) */
int fheb200_eval_mult(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct *b, fheb200_ct **out);
/* EvalMult followed by the rescale FLEXIBLEAUTOEXT would apply before the product is multiplied again (ModReduceInternal), as ONE
 * division by P * q_last instead of ModDown then rescale: same level and scale as fheb200_eval_rescale(fheb200_eval_mult(a, b)),
 * integers within a few units of it.  Used by the Paterson-Stockmeyer evaluator behind fheb200_eval_chebyshev_series; no direct
 * counterpart in the reference's calls. */
int fheb200_eval_mult_rescaled(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct *b, fheb200_ct **out);
/* EvalMult(ct, pt) (src/sort_algo.h:319,344,359,576) */
int fheb200_eval_mult_plain(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out);
/* acc + a * p in one pass: EvalMult(ct, pt) followed by EvalAddInPlace (src/sort_algo.h:269-276, 616-624), bit-identical to
 * the two calls; acc may be modified in place and returned under the new handle - release acc afterwards, do not reuse it */
int fheb200_eval_mult_plain_add(fheb200_ctx *ctx, fheb200_ct *acc, fheb200_ct *a, fheb200_pt *p, fheb200_ct **out);
/* EvalMult(ct, double) / EvalMultInPlace(ct, double) (src/sign.cpp:26-30 ; src/sort_algo.h:723) */
int fheb200_eval_mult_const(fheb200_ctx *ctx, fheb200_ct *a, double c, fheb200_ct **out);
/* explicit rescale (ModReduce) — a no-op at the OpenFHE API level in FLEXIBLEAUTOEXT, exposed for tests */
int fheb200_eval_rescale(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out);

/* ---- rotation ---- */
/* EvalRotate(ct, index) (src/rotation.h:224,231,350) */
int fheb200_eval_rotate(fheb200_ctx *ctx, fheb200_ct *a, int index, fheb200_ct **out);
/* EvalFastRotationPrecompute(ct) / EvalFastRotation(ct, index, M, precomp) (src/rotation.h:286,344-346) */
int fheb200_fast_rotation_precompute(fheb200_ctx *ctx, fheb200_ct *a, fheb200_pre **out);
int fheb200_pre_release(fheb200_pre *pre);
int fheb200_eval_fast_rotation(fheb200_ctx *ctx, fheb200_ct *a, int index, fheb200_pre *pre, fheb200_ct **out);
int fheb200_has_rotation_key(fheb200_ctx *ctx, int index, int *present);

/* ---- polynomial evaluation ---- */
/* EvalChebyshevSeries / EvalChebyshevSeriesPS(ct, coeffs, a, b) (src/sort_algo.h:629,728 ; src/sign.cpp:76,448) */
int fheb200_eval_chebyshev_series(fheb200_ctx *ctx, fheb200_ct *x, const double *coeffs, size_t count, double a, double b,
                                  fheb200_ct **out);
/* EvalPolyLinear(ct, coeffs) (src/mehp24/mehp24_utils.cpp:256-259) */
int fheb200_eval_poly(fheb200_ctx *ctx, fheb200_ct *x, const double *coeffs, size_t count, fheb200_ct **out);
/* lbcrypto::EvalChebyshevCoefficients(f, a, b, degree) (utils/generate_cheb_coeffs.cpp:17); host only */
int fheb200_chebyshev_coefficients(double (*f)(double, void *), void *user, double a, double b, int degree, double *out);

/* ---- bootstrapping (sparsely packed ciphertexts, 2 * slots <= ring_dim / 2; uniform ternary secret) ---- */
/* EvalBootstrapSetup(levelBudget, bsgsDim, slots) (tests/k-way/KWaySort2Test.cpp:42-43): budgets are the levels spent on
 * CoeffsToSlots / SlotsToCoeffs (clamped to log2 slots); bsgsDim is not needed (every group has <= 2^(stages+1) diagonals,
 * evaluated with one hoisted ModUp).  slots = 0 selects the context's batch size. */
int fheb200_eval_bootstrap_setup(fheb200_ctx *ctx, int budget_enc, int budget_dec, int slots);
/* EvalBootstrapKeyGen(secretKey, slots) (tests/k-way/KWaySort2Test.cpp:44): rotation keys of the plan + the conjugation key */
int fheb200_eval_bootstrap_keygen(fheb200_ctx *ctx, fheb200_keys *k, int slots);
/* EvalBootstrap(ct) (src/k-way/EvalUtils.cpp:76 ; src/sign.cpp:168 ; src/sort_algo.h:1111).  The input needs one level
 * left (GetLevel() < multiplicative depth, or a pending rescale); the output is at level fheb200_bootstrap_depth - 1 with
 * noise scale degree 2. */
int fheb200_eval_bootstrap(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out);
/* EvalBootstrap(ct, numIterations, precision) (src/sort_algo.h:1110-1111): Meta-BTS, one extra bootstrapping of the scaled
 * error per extra iteration; precision (bits removed per extra pass) is capped at 12 */
int fheb200_eval_bootstrap_iterated(fheb200_ctx *ctx, fheb200_ct *a, int iterations, int precision, fheb200_ct **out);
int fheb200_bootstrap_depth(fheb200_ctx *ctx, int slots, int *levels);
/* test hook: the intermediate ciphertext after step 1 (scale-down to q_0), 2 (ModRaise + SubSum), 3 (CoeffsToSlots +
 * conjugate), 4 (EvalMod), 5 (ModRaise alone, before SubSum); 0 = the whole bootstrapping */
int fheb200_eval_bootstrap_partial(fheb200_ctx *ctx, fheb200_ct *a, int stop_after, fheb200_ct **out);
/* slot-wise complex conjugation (Galois element 2 * ring_dim - 1; key made by fheb200_eval_bootstrap_keygen) */
int fheb200_eval_conjugate(fheb200_ctx *ctx, fheb200_ct *a, fheb200_ct **out);
/* host only (no GPU): the bootstrapping plan for a parameter set - depth, the range K of the overflow polynomial, the
 * number of double-angle steps, the rotation indices, the Chebyshev series (in y = 2 u^2 - 1, u = slot - shift) of the
 * even scaled cosine and that shift; and the plan's CoeffsToSlots (which = 0)
 * or SlotsToCoeffs (which = 1) applied to 2 * slots complex values in the clear (re, im interleaved) */
int fheb200_bootstrap_plan_info(int ring_dim, int slots, int budget_enc, int budget_dec, double q0, double raise_scale,
                                int correction, int *depth, int *k_range, int *double_angles, int *rotations, int capacity,
                                int *num_rotations, double *cheb, int cheb_capacity, int *num_cheb, double *shift);
int fheb200_bootstrap_plan_apply(int ring_dim, int slots, int budget_enc, int budget_dec, double q0, double raise_scale,
                                 int correction, int which, const double *in_re_im, double *out_re_im);

/* ---- multi-GPU combine of partial ciphertexts: the device buffer [2][limbs][n] of a ciphertext, to be
 * summed element-wise across ranks as 64-bit integers (NCCL all-reduce; <= 16 summands of < 2^60 cannot wrap), then
 * brought back to [0, q_limb) by fheb200_ct_reduce_partial_sums.  The caller orders the streams (fheb200_ctx_sync). */
/* The engine's own NCCL communicator (multi-GPU extension, no OpenFHE counterpart; SURVEY.md 8e: partial rank / output
 * ciphertexts of a sharded sort, reference batches src/sort_algo.h:475-491, 713-742).  Rank 0 draws a 128-byte id and gives
 * it to the other ranks by any means; every rank then calls comm_init (world <= 16).  all_reduce_sum replaces the ciphertext
 * by the sum over all ranks, limb by limb modulo its modulus, on the engine's stream (ncclAllReduce + fix-up kernel; no host
 * synchronisation).  NCCL is loaded with dlopen on first use. */
int fheb200_comm_unique_id(unsigned char *id128);
int fheb200_ctx_comm_init(fheb200_ctx *ctx, const unsigned char *id128, int rank, int world);
int fheb200_ctx_comm_destroy(fheb200_ctx *ctx);
int fheb200_ctx_comm_info(fheb200_ctx *ctx, int *rank, int *world, int *nccl_version);
int fheb200_ct_all_reduce_sum(fheb200_ctx *ctx, fheb200_ct *ct);
int fheb200_ct_device_view(fheb200_ct *ct, void **device_ptr, size_t *words);
int fheb200_ct_reduce_partial_sums(fheb200_ctx *ctx, fheb200_ct *ct);

/* ---- raw RNS access: parity tests against the oracle and the primitive micro-benchmarks ----
 * Polynomials are uint64 [limb][ring_dim] host arrays, limb i reduced mod modulus i (evaluation domain).
 * Import preconditions (NOT validated: the caller is a test or a deserialiser that wrote the data itself): every residue is
 * below its modulus, noise_scale_deg is 1 or 2 and `scale` is the scaling factor the limbs were produced at. */
int fheb200_ct_export(fheb200_ctx *ctx, fheb200_ct *ct, uint64_t *c0, uint64_t *c1, size_t words_each);
int fheb200_ct_import(fheb200_ctx *ctx, const uint64_t *c0, const uint64_t *c1, int num_limbs, int noise_scale_deg,
                      double scale, int slots, fheb200_ct **out);
/* Asynchronous variants for pinned host buffers (cudaHostAlloc / torch pin_memory): the copies run on a copy stream of
 * the context and overlap the kernels of neighbouring work items (issue the import of item s+1 before the work of item
 * s); the import buffers must stay untouched, and the export buffers are valid, only after the next fheb200_ctx_sync. */
int fheb200_ct_import_async(fheb200_ctx *ctx, const uint64_t *c0, const uint64_t *c1, int limbs, int noise_scale_deg,
                            double scale, int slots, fheb200_ct **out);
/* makes the engine's stream wait for the copies of an fheb200_ct_import_async; call it before the first use of `ct` */
int fheb200_ct_ready(fheb200_ctx *ctx, fheb200_ct *ct);
int fheb200_ct_export_async(fheb200_ctx *ctx, fheb200_ct *ct, uint64_t *c0, uint64_t *c1, size_t words_each);
int fheb200_pt_export(fheb200_ctx *ctx, fheb200_pt *pt, uint64_t *out, size_t words);
/* keys: which = 0 secret [L+K][n] ; 1 public b [L][n] ; 2 public a [L][n] */
int fheb200_keys_export(fheb200_ctx *ctx, fheb200_keys *k, int which, uint64_t *out, size_t words);
/* evaluation key [digit][2][L+K][n]; galois = 0 selects the relinearisation key */
int fheb200_evk_export(fheb200_ctx *ctx, uint32_t galois, uint64_t *out, size_t words);
/* raw import of key material and context parameters: what Serial::DeserializeFromFile / DeserializeEvalMultKey /
 * DeserializeEvalAutomorphismKey need (reference src/sort.h:31-73) */
int fheb200_keys_import_public(fheb200_ctx *ctx, const uint64_t *b, const uint64_t *a, size_t words_each, fheb200_keys **out);
int fheb200_evk_import(fheb200_ctx *ctx, uint32_t galois, const uint64_t *data, size_t words);
int fheb200_evk_list(fheb200_ctx *ctx, uint32_t *galois_out, size_t capacity, size_t *count);
int fheb200_has_mult_key(fheb200_ctx *ctx, int *present);
int fheb200_ctx_evk_words(fheb200_ctx *ctx, size_t *words);
int fheb200_ctx_params(fheb200_ctx *ctx, fheb200_params *p);
int fheb200_galois_for_rotation(fheb200_ctx *ctx, int index, uint32_t *galois);
/* single primitives on host buffers (H2D, kernel, D2H) */
int fheb200_raw_ntt(fheb200_ctx *ctx, int inverse, int qp_index, uint64_t *poly);
int fheb200_raw_automorphism(fheb200_ctx *ctx, uint32_t galois, int limbs, const uint64_t *in, uint64_t *out);
int fheb200_raw_modup(fheb200_ctx *ctx, int num_limbs, const uint64_t *c, uint64_t *digits, size_t digit_words);
int fheb200_raw_rescale(fheb200_ctx *ctx, int num_limbs, const uint64_t *in, uint64_t *out);
/* single kernels on host-supplied limbs (limb-level parity tests against the oracle; not on any hot path).
 * raw_lincomb: out[g] = sum_t w[g][t][limb] * in[t] over limbs2 = npoly * num_limbs limbs (Paterson-Stockmeyer leaf sums inside
 * EvalChebyshevSeriesPS, reference src/sort_algo.h:557); variant 0 = tensor cores + cp.async rings, 1 = tensor cores, 2 = CUDA cores */
int fheb200_raw_lincomb(fheb200_ctx *ctx, int num_limbs, int limbs2, int count, int outputs, const uint64_t *in, const uint64_t *weights,
                        uint64_t *out, int variant);
int fheb200_raw_mul_pt(fheb200_ctx *ctx, int num_limbs, const uint64_t *a0, const uint64_t *a1, const uint64_t *pt, uint64_t *o0,
                       uint64_t *o1, int accumulate);
int fheb200_raw_tensor(fheb200_ctx *ctx, int num_limbs, const uint64_t *a0, const uint64_t *a1, const uint64_t *b0, const uint64_t *b1,
                       uint64_t *d0, uint64_t *d1, uint64_t *d2);
int fheb200_raw_dh_accumulate(fheb200_ctx *ctx, int num_limbs, const uint64_t *acc, const uint64_t *cP, const uint64_t *pt, uint64_t *S,
                              uint32_t galois);
int fheb200_raw_inner_product(fheb200_ctx *ctx, int num_limbs, uint32_t galois, const uint64_t *digits, uint64_t *acc);
int fheb200_raw_moddown(fheb200_ctx *ctx, int num_limbs, const uint64_t *acc, uint64_t *out0, uint64_t *out1);
/* Forward NTT with NO allocation on the call path.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * d_in / d_out are DEVICE pointers to [count][ring_dim] uint64 limbs (limb i modulo modulus qp_indices[i]; d_out may equal
 * d_in); ring dimension 2^17 only; count <= fheb200_ntt_forward_noalloc_max_limbs() (= 256; beyond it the call fails with an
 * error).  The kernel's scratch (per-limb dependency counters, 1 KiB) is a static device array that every GPU loading the
 * library holds until exit and that ALL calls share: it is cleared before each launch, and calls that overlap on the GPU
 * race on it - order them.  Runs on the context's stream; same limbs out as fheb200_raw_ntt. */
int fheb200_ntt_forward_noalloc(fheb200_ctx *ctx, const int *qp_indices, int count, const uint64_t *d_in, uint64_t *d_out);
int fheb200_ntt_forward_noalloc_max_limbs(void);
/* timing loops for bench.py: `reps` back-to-back launches on the context stream with inputs resident in HBM;
 * returns the elapsed device time in milliseconds measured with CUDA events on that stream. */
int fheb200_bench_rotate(fheb200_ctx *ctx, fheb200_ct *ct, int index, int reps, float *ms);
int fheb200_bench_rotations(fheb200_ctx *ctx, fheb200_ct *ct, const int *indices, size_t count, int reps, int hoisted,
                            float *ms);
int fheb200_bench_kernel(fheb200_ctx *ctx, const char *name, int num_limbs, int reps, float *ms, double *algo_bytes);

#ifdef __cplusplus
}
#endif
#endif
