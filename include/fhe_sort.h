/*
 * fhe_sort.h — C ABI of the host-side sort drivers (libfhesort.so): the reference's DirectSort test fixture as a
 * session object.  The drivers themselves are C++ (fhe_sorting_b200/host/: sort_algo.h, rotation.h, sign.h,
 * comparison.h, encryption.h — same class and method names as the reference's src/ headers) running over the
 * lbcrypto-shaped shim host/openfhe.h, which forwards every call to libfheb200.so (include/fhe_b200.h).
 */
#ifndef FHE_SORT_H
#define FHE_SORT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct fhesort_session fhesort_session;
const char *fhesort_last_error(void);

/* CCParams via DirectSort<N>::getSizeParameters, GenCryptoContext, KeyGen, EvalRotateKeyGen, EvalMultKeyGen, DirectSort<N>
 * (tests/DirectSortTest.cpp:24-51, 96-100).  mult_depth = 0 keeps the per-N depth; seed = 0 draws one from the OS. */
int fhesort_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out);
/* The same fixture with the parameters of DirectSort<N>::sort_hybrid (tests/DirectSortHTest.cpp:22-110: deeper chain, the
 * rotation steps of the MEHP24-style placement); run it with fhesort_run stage 3. */
int fhesort_hybrid_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out);
/* MEHP24 ranking sort session (tests/mehp24/Mehp24SortTest.cpp:22-93: batch = min(N^2, n/2), depth per N, scale 40,
 * rotation keys of mehp24::utils::getRotationIndices(N)); fhesort_run stage 0 runs mehp24::sortFG (N <= 256) or
 * sortLargeArrayFG (src/mehp24/mehp24_sort.cpp:246, 609) with dg_i = (log2 N + 1) / 2, df_i = 2. */
int fhesort_mehp24_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out);
/* HKC+21 k-way sorting network session (tests/k-way/KWaySort2Test.cpp:14-58): N = k^M values, batch = next power of two,
 * ring 2^log_ring_dim, depth 40 (mult_depth = 0), first modulus 60 / scaling 59 bits, rotation keys +-2^i < N,
 * EvalBootstrapSetup({4,4} or {5,5}) + EvalBootstrapKeyGen; fhesort_run stage 0 runs KWayAdapter<N>::sort
 * (src/kway_adapter.h:66) with bootstrapping enabled (SignConfig(compos, depth)); sign_n = 0 selects the configuration
 * the reference tests really run (dg = 2, df = the test table's d_g).  k = 2, 3 or 5. */
int fhesort_kway_session_create(int k, int M, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out);
/* BitonicSort<N> session (tests/BitonicSortTest.cpp:12-55): depth 58 (mult_depth = 0), 59-bit scaling, batch N, rotation keys
 * +-2^i, EvalBootstrapSetup({3,3}); inputs in [0, 255]; sign_n = 0 selects CompositeSign(4, 3, 3) as the reference test does */
int fhesort_bitonic_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out);
/* EvalBootstrap calls made by the network's level guards so far (the sign's own lazy bootstraps are not counted) */
int fhesort_session_bootstraps(fhesort_session *s, int *count);
int fhesort_session_free(fhesort_session *s);
int fhesort_session_info(fhesort_session *s, int *mult_depth, int *num_rotation_keys, double *setup_ms);
/* encrypt(input[N]) -> stage -> decrypt(output[N]).  stage 0: DirectSort<N>::sort (src/sort_algo.h:752) ;
 * 1: constructRank (:368) ; 2: rotationIndexCheckN (:658) on encrypted aux[N] ranks ; 3: sort_hybrid (:1050).  sign_n = 0 selects the
 * reference tests' CompositeSign configuration for N.  elapsed_ms covers encrypt + stage + decrypt (host clock). */
int fhesort_run(fhesort_session *s, int stage, const double *input, const double *aux, int sign_n, int dg, int df,
                double *output, double *elapsed_ms, int *out_level);
/* Multi-GPU: shard the independent batches of both phases of the sort over `world` processes (one per GPU, same seed so
 * that keys are replicated).  `fn` must sum `words` 64-bit integers at device_ptr in place over all processes (NCCL
 * all-reduce, SUM, int64) and return 0 after the result is visible to other streams. */
typedef int (*fhesort_allreduce_fn)(void *user, void *device_ptr, size_t words);
int fhesort_session_set_shard(fhesort_session *s, int rank, int world, fhesort_allreduce_fn fn, void *user);
/* Sharding over the engine's own NCCL communicator: id128 = fheb200_comm_unique_id() of rank 0, handed to every rank */
int fhesort_session_comm_init(fhesort_session *s, const unsigned char *id128, int rank, int world);
/* Concurrent CUDA streams over the batches this process evaluates (default 4, or FHESORT_STREAMS; 1 = one stream). */
int fhesort_session_set_streams(fhesort_session *s, int k);
/* RotationComposer<N>::rotate (src/rotation.h:215-233) */
int fhesort_composed_rotate(fhesort_session *s, const double *input, int count, int rotation, double *output);
/* Fixture of the reference's primitive-level tests (tests/RotationTest.cpp:30-54, tests/SignTest.cpp:22-39,
 * utils/sign_characterizer.cpp:26-38): context + keys with the given depth / scaling size / batch size / rotation keys. */
int fhesort_toolkit_session_create(int log_ring_dim, int mult_depth, int scaling_mod_size, int batch_size, const int *rotation_keys,
                                   int num_keys, int bootstrap_budget, uint64_t seed, int device, fhesort_session **out);
/* one kwaySort::EvalUtils / SortUtils / Sorter method on fresh encryptions (tests/k-way/EvalUtilsTest.cpp:62-158,
 * SortUtilsTest.cpp:68-113, SorterTest.cpp:67-260); op codes: see host/sort_cabi.cpp */
int fhesort_kway_unit(fhesort_session *s, int op, long k, long M, long param, int aux, const double *in0, const double *in1,
                      const double *in2, const int *idx0, const int *idx1, int count, double *out, int *level_before, int *level_after);
/* sign(x, cc, func, cfg) (src/sign.cpp:635-651): func 0 = CompositeSign(n, dg, df), 1 = SignumPolycircuit, 2 = Tanh, 3 = NaiveDiscrete */
int fhesort_sign_eval(fhesort_session *s, int func, int n, int dg, int df, const double *input, int count, double *output, int *level,
                      double *elapsed_ms);
/* RotationTree<N>::buildTree(start, end) + treeRotate for each requested rotation (src/rotation.h:240-357); N = 8 or 128;
 * algo 0 = NAF, 1 = BNAF, 2 = BINARY; outputs [nrot][count] */
int fhesort_tree_rotate(fhesort_session *s, int N, int algo, int start, int end, const double *input, int count, const int *rotations,
                        int nrot, double *outputs);
/* host-only (no GPU needed) */
int fhesort_rotation_indices(int N, int log_ring_dim, int *out, int capacity, int *count);
int fhesort_shard_batches(int N, int log_ring_dim, int rank, int world, int *out, int capacity, int *count, int *sharded);
/* k-way network geometry of one stage: geometry4 = (m, logDist, slope, rotate distance); indices0/1[num_slots] = sorter size
 * and 1-based position of every slot (src/k-way/Masking.cpp:26-167) */
int fhesort_kway_stage(int k, int M, int stage, int num_slots, int *geometry4, int *indices0, int *indices1);
int fhesort_decompose(const int *key_indices, int num_keys, int rotation, int wrap, int algo, int *steps, int capacity,
                      int *count);
int fhesort_doubled_sinc_coefficients(int N, double *out, int capacity, int *count);
/* selectCoefficients<N>() (src/generated_coeffs.h, written by utils/generate_cheb_coeffs.cpp) */
int fhesort_sinc_coefficients(int N, double *out, int capacity, int *count);

#ifdef __cplusplus
}
#endif
#endif
