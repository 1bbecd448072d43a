// sort_cabi.cpp — C entry points of the host-side sort drivers (libfhesort.so), declared in include/fhe_sort.h.
// A session owns what the reference's test fixture owns (tests/DirectSortTest.cpp:22-51): CryptoContext, key pair,
// rotation / relinearisation keys resident in HBM, a DebugEncryption and a DirectSort<N>.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>

// Built twice: (1) libfhesort.so from this repository's drivers (host/sort_algo.h ...); (2) with -DREFERENCE_DRIVERS and
// the reference's src/ first on the include path, oracle/_ref/libref_sort.so — the reference's own unmodified
// sort_algo.h / rotation.h / sign.cpp / comparison.cpp / encryption.cpp running on the B200 engine through the shim
// (built only where /root/reference exists; see oracle/Makefile).
#include <omp.h>

#include "../../include/fhe_sort.h"
#include "hybrid_params.h"
// the reference's rotation.h has no include guard: reach it through sort_algo.h only.  Angle brackets in the reference
// build: a quoted include would find this directory's own drivers first.
#ifdef REFERENCE_DRIVERS
#include <sort_algo.h>
#include <mehp24/mehp24_sort.h>
#include <kway_adapter.h>
#else
#include "sort_algo.h"
#include "mehp24.h"
#include "kway_adapter.h"
#endif

namespace {
thread_local std::string g_err;

struct SessionBase {
    CryptoContext<DCRTPoly> cc;
    KeyPair<DCRTPoly> keys;
    std::shared_ptr<DebugEncryption> enc;
    std::vector<int> rotations;
    int len = 0, depth = 0;
    bool seeded = false; // created with a non-zero (test-only, deterministic) key seed
    double setup_ms = 0;
    virtual ~SessionBase() = default;
    virtual Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) = 0;
    virtual Ciphertext<DCRTPoly> sort_hybrid(const Ciphertext<DCRTPoly> &, SignConfig &) { throw std::runtime_error("sort_hybrid: DirectSort sessions only"); }
    virtual Ciphertext<DCRTPoly> rank(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) = 0;
    virtual Ciphertext<DCRTPoly> place(const Ciphertext<DCRTPoly> &rank, const Ciphertext<DCRTPoly> &in) = 0;
    virtual Ciphertext<DCRTPoly> composed_rotate(const Ciphertext<DCRTPoly> &in, int r) = 0;
    virtual void shard(int rank, int world) = 0;
    virtual void streams(int) {}
    virtual bool own_sign_config(int, int, int, SignConfig &) { return false; }
    virtual int bootstraps() const { return 0; }
};
template <int N> struct Session : SessionBase {
    std::unique_ptr<DirectSort<N>> sorter;
    std::unique_ptr<RotationComposer<N>> composer;
    Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) override {
        return sorter->sort(in, SignFunc::CompositeSign, cfg);
    }
    Ciphertext<DCRTPoly> sort_hybrid(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) override {
        return sorter->sort_hybrid(in, SignFunc::CompositeSign, cfg, keys.secretKey);
    }
    Ciphertext<DCRTPoly> rank(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) override {
        return sorter->constructRank(in, SignFunc::CompositeSign, cfg);
    }
    Ciphertext<DCRTPoly> place(const Ciphertext<DCRTPoly> &rk, const Ciphertext<DCRTPoly> &in) override {
        return sorter->rotationIndexCheckN(rk, in);
    }
    Ciphertext<DCRTPoly> composed_rotate(const Ciphertext<DCRTPoly> &in, int r) override { return composer->rotate(in, r); }
#ifdef REFERENCE_DRIVERS
    void shard(int, int world) override {
        if (world > 1) throw std::runtime_error("the reference's DirectSort has no multi-GPU sharding");
    }
#else
    void shard(int rank, int world) override { sorter->setShard(rank, world); }
    void streams(int k) override { sorter->setStreams(k); }
#endif
};

// MEHP24 ranking sort (reference fixture: tests/mehp24/Mehp24SortTest.cpp:22-93): N values in the first row of an
// N x N slot matrix (N <= 256), longer arrays split into 256-value ciphertexts
struct Mehp24Session : SessionBase {
    std::unique_ptr<Comparison> comp;
    int side = 256;               // values per ciphertext (side x side slot matrix)
    int m_rank = 0, m_world = 1;  // multi-GPU: units of the several-ciphertext sort (host/mehp24.h)
    Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) override {
        const uint32_t dg_i = (uint32_t)((std::log2((double)len) + 1) / 2), df_i = 2;
        if (len <= side) return mehp24::sortFG(in, (size_t)len, SignFunc::CompositeSign, cfg, comp, dg_i, df_i, cc);
#ifdef REFERENCE_DRIVERS
        return mehp24::sortLargeArrayFG(in, (size_t)len, (size_t)side, SignFunc::CompositeSign, cfg, comp, dg_i, df_i, cc);
#else
        return mehp24::sortLargeArrayFG(in, (size_t)len, (size_t)side, SignFunc::CompositeSign, cfg, comp, dg_i, df_i, cc, m_rank, m_world);
#endif
    }
    Ciphertext<DCRTPoly> rank(const Ciphertext<DCRTPoly> &, SignConfig &) override { throw std::runtime_error("mehp24: sort only"); }
    Ciphertext<DCRTPoly> place(const Ciphertext<DCRTPoly> &, const Ciphertext<DCRTPoly> &) override {
        throw std::runtime_error("mehp24: sort only");
    }
    Ciphertext<DCRTPoly> composed_rotate(const Ciphertext<DCRTPoly> &, int) override { throw std::runtime_error("mehp24: sort only"); }
    void shard(int rank, int world) override {
#ifdef REFERENCE_DRIVERS
        if (world > 1) throw std::runtime_error("the reference's drivers have no multi-GPU sharding");
#endif
        if (world > 1 && len <= side) throw std::runtime_error("mehp24: the single-ciphertext sort is sequential (replicas only)");
        m_rank = rank, m_world = world;
    }
};
int mehp24_depth(int N) { // tests/mehp24/Mehp24SortTest.cpp:31-62
    switch (N) {
    case 4: return 31;
    case 8: case 16: return 35;
    case 32: case 64: return 42;
    case 128: return 46;
    case 256: return 49;
    case 512: return 57;
    case 1024: return 60;
    default: throw std::invalid_argument("mehp24: array length must be a power of two in 4..1024");
    }
}
SessionBase *make_mehp24_session(int N, int log_ring_dim, int depth_override, uint64_t seed, int device) {
    auto t0 = std::chrono::steady_clock::now();
    auto s = std::make_unique<Mehp24Session>();
    // values per ciphertext: N up to 256 (the reference's split length), less on small test rings whose slots hold a smaller matrix
    int side = N < 256 ? N : 256;
    while (side > 2 && side * side > (1 << log_ring_dim) / 2) side >>= 1;
    const int slots = side * side;
    s->side = side;
    CCParams<CryptoContextCKKSRNS> p;
    p.SetRingDim(1u << log_ring_dim);
    p.SetBatchSize(slots);
    p.SetMultiplicativeDepth(depth_override > 0 ? depth_override : mehp24_depth(N));
    p.SetScalingModSize(40);
    p.SetSecurityLevel(log_ring_dim >= 17 ? HEStd_128_classic : HEStd_NotSet);
    CryptoContextImpl::SetKeySeed(seed); // 0: OS entropy (ChaCha20); non-zero: TEST-ONLY deterministic keys
    s->seeded = seed != 0;
#ifdef REFERENCE_DRIVERS
    setenv("FHEB200_DEVICE", std::to_string(device).c_str(), 1);
    s->cc = GenCryptoContext(p);
#else
    s->cc = GenCryptoContext(p, device);
#endif
    s->cc->Enable(PKE), s->cc->Enable(KEYSWITCH), s->cc->Enable(LEVELEDSHE), s->cc->Enable(ADVANCEDSHE);
    s->keys = s->cc->KeyGen();
    #ifdef REFERENCE_DRIVERS
    s->rotations = mehp24::utils::getRotationIndices((size_t)N);
#else
    s->rotations = mehp24::utils::getRotationIndices((size_t)N, (size_t)side);
#endif
    s->cc->EvalRotateKeyGen(s->keys.secretKey, s->rotations);
    s->cc->EvalMultKeyGen(s->keys.secretKey);
    s->enc = std::make_shared<DebugEncryption>(s->cc, s->keys);
    s->comp = std::make_unique<Comparison>(s->enc);
    s->len = N, s->depth = (int)p.GetMultiplicativeDepth();
    s->cc->Sync();
    s->setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return s.release();
}

// HKC+21 k-way sorting network (reference fixture: tests/k-way/KWaySort2Test.cpp:14-58, KWaySort3Test.cpp): ring 2^17,
// depth 40, first modulus 60, scaling 59, rotation keys +-2^i, EvalBootstrapSetup(levelBudget, {0,0}, slots) + KeyGen
template <int N> struct KWaySession : SessionBase {
    std::unique_ptr<KWayAdapter<N>> sorter;
    int k = 2, M = 1;
    Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) override {
        return sorter->sort(in, SignFunc::CompositeSign, cfg);
    }
    Ciphertext<DCRTPoly> rank(const Ciphertext<DCRTPoly> &, SignConfig &) override { throw std::runtime_error("k-way: sort only"); }
    Ciphertext<DCRTPoly> place(const Ciphertext<DCRTPoly> &, const Ciphertext<DCRTPoly> &) override {
        throw std::runtime_error("k-way: sort only");
    }
    Ciphertext<DCRTPoly> composed_rotate(const Ciphertext<DCRTPoly> &, int) override { throw std::runtime_error("k-way: sort only"); }
    void shard(int, int world) override {
        if (world > 1) throw std::runtime_error("k-way: the stages of the network are sequential (replicas only)");
    }
    // The reference tests build CompositeSignConfig(3, d_f, d_g) for a constructor that reads (n, dg, df)
    // (tests/k-way/KWaySort2Test.cpp:149 vs src/sign.h:17): the network really runs dg = 2 and df = the table's d_g.
    bool own_sign_config(int n, int dg, int df, SignConfig &out) override {
        if (n == 0) {
            const int table = k == 2 ? (N <= 16 ? 2 : N <= 128 ? 3 : N <= 512 ? 4 : 5)
                            : k == 3 ? (N <= 9 ? 2 : N <= 81 ? 3 : N <= 243 ? 4 : 5) : (N <= 125 ? 3 : 4);
            n = 3, dg = 2, df = table;
        }
        out = SignConfig(CompositeSignConfig(n, dg, df), depth);
        return true;
    }
#ifndef REFERENCE_DRIVERS
    int bootstraps() const override { return sorter->bootstraps(); }
#endif
};
template <int N> SessionBase *make_kway_session(int k, int M, int log_ring_dim, int depth_override, uint64_t seed, int device) {
    auto t0 = std::chrono::steady_clock::now();
    auto s = std::make_unique<KWaySession<N>>();
    s->k = k, s->M = M;
    CCParams<CryptoContextCKKSRNS> p;
    std::vector<uint32_t> levelBudget;
    KWayAdapter<N>::getSizeParameters(p, s->rotations, levelBudget);
    p.SetRingDim(1u << log_ring_dim);
    if (depth_override > 0) p.SetMultiplicativeDepth(depth_override);
    p.SetSecurityLevel(log_ring_dim >= 17 ? HEStd_128_classic : HEStd_NotSet);
    CryptoContextImpl::SetKeySeed(seed); // 0: OS entropy (ChaCha20); non-zero: TEST-ONLY deterministic keys
    s->seeded = seed != 0;
#ifdef REFERENCE_DRIVERS
    setenv("FHEB200_DEVICE", std::to_string(device).c_str(), 1);
    s->cc = GenCryptoContext(p);
#else
    s->cc = GenCryptoContext(p, device);
#endif
    s->cc->Enable(PKE), s->cc->Enable(KEYSWITCH), s->cc->Enable(LEVELEDSHE), s->cc->Enable(ADVANCEDSHE), s->cc->Enable(FHE);
    const uint32_t slots = s->cc->GetEncodingParams()->GetBatchSize();
    s->keys = s->cc->KeyGen();
    s->cc->EvalRotateKeyGen(s->keys.secretKey, s->rotations);
    s->cc->EvalMultKeyGen(s->keys.secretKey);
    s->cc->EvalBootstrapSetup(levelBudget, {0, 0}, slots);
    s->cc->EvalBootstrapKeyGen(s->keys.secretKey, slots);
    s->enc = std::make_shared<DebugEncryption>(s->cc, s->keys);
    s->sorter = std::make_unique<KWayAdapter<N>>(s->cc, s->keys.publicKey, s->keys.secretKey, s->enc, k, M);
    s->len = N, s->depth = (int)p.GetMultiplicativeDepth();
    s->cc->Sync();
    s->setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return s.release();
}
SessionBase *make_kway_session_for(int k, int M, int log_ring_dim, int depth_override, uint64_t seed, int device) {
    long N = 1;
    for (int i = 0; i < M; i++) N *= k;
#define KWAY_CASE(n) case n: return make_kway_session<n>(k, M, log_ring_dim, depth_override, seed, device);
    switch (N) {
        KWAY_CASE(4) KWAY_CASE(8) KWAY_CASE(16) KWAY_CASE(32) KWAY_CASE(64) KWAY_CASE(128) KWAY_CASE(256) KWAY_CASE(512) KWAY_CASE(1024)
        KWAY_CASE(9) KWAY_CASE(27) KWAY_CASE(81) KWAY_CASE(243) KWAY_CASE(729)
        KWAY_CASE(25) KWAY_CASE(125) KWAY_CASE(625)
    default: throw std::invalid_argument("k-way: array length k^M must be one of 4..1024 (k = 2), 9..729 (k = 3), 25..625 (k = 5)");
    }
#undef KWAY_CASE
}

// Bitonic sorting network (reference fixture: tests/BitonicSortTest.cpp:12-55): depth 58, 59-bit scaling, batch N, rotation
// keys +-2^i, bootstrapping with level budget {3,3}; inputs in [0, 255], CompositeSign(4, 3, 3)
template <int N> struct BitonicSession : SessionBase {
    std::unique_ptr<BitonicSort<N>> sorter;
    Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &in, SignConfig &cfg) override {
        return sorter->sort(in, SignFunc::CompositeSign, cfg);
    }
    Ciphertext<DCRTPoly> rank(const Ciphertext<DCRTPoly> &, SignConfig &) override { throw std::runtime_error("bitonic: sort only"); }
    Ciphertext<DCRTPoly> place(const Ciphertext<DCRTPoly> &, const Ciphertext<DCRTPoly> &) override {
        throw std::runtime_error("bitonic: sort only");
    }
    Ciphertext<DCRTPoly> composed_rotate(const Ciphertext<DCRTPoly> &, int) override { throw std::runtime_error("bitonic: sort only"); }
    void shard(int, int world) override {
        if (world > 1) throw std::runtime_error("bitonic: the stages of the network are sequential (replicas only)");
    }
    bool own_sign_config(int n, int dg, int df, SignConfig &out) override {
        out = n == 0 ? SignConfig(CompositeSignConfig(4, 3, 3)) : SignConfig(CompositeSignConfig(n, dg, df));
        return true;
    }
#ifndef REFERENCE_DRIVERS
    int bootstraps() const override { return sorter->bootstraps; }
#endif
};
template <int N> SessionBase *make_bitonic_session(int log_ring_dim, int depth_override, uint64_t seed, int device) {
    auto t0 = std::chrono::steady_clock::now();
    auto s = std::make_unique<BitonicSession<N>>();
    CCParams<CryptoContextCKKSRNS> p;
    p.SetMultiplicativeDepth(depth_override > 0 ? depth_override : 58);
    p.SetScalingModSize(59);
    p.SetBatchSize(N);
    p.SetRingDim(1u << log_ring_dim);
    p.SetSecurityLevel(log_ring_dim >= 17 ? HEStd_128_classic : HEStd_NotSet);
    CryptoContextImpl::SetKeySeed(seed); // 0: OS entropy (ChaCha20); non-zero: TEST-ONLY deterministic keys
    s->seeded = seed != 0;
#ifdef REFERENCE_DRIVERS
    setenv("FHEB200_DEVICE", std::to_string(device).c_str(), 1);
    s->cc = GenCryptoContext(p);
#else
    s->cc = GenCryptoContext(p, device);
#endif
    s->cc->Enable(PKE), s->cc->Enable(KEYSWITCH), s->cc->Enable(LEVELEDSHE), s->cc->Enable(ADVANCEDSHE), s->cc->Enable(FHE);
    s->keys = s->cc->KeyGen();
    for (int i = 1; i < N; i *= 2) s->rotations.push_back(i), s->rotations.push_back(-i);
    s->cc->EvalRotateKeyGen(s->keys.secretKey, s->rotations);
    s->cc->EvalMultKeyGen(s->keys.secretKey);
    s->cc->EvalBootstrapSetup({3, 3}, {0, 0}, N);
    s->cc->EvalBootstrapKeyGen(s->keys.secretKey, N);
    s->enc = std::make_shared<DebugEncryption>(s->cc, s->keys);
    s->sorter = std::make_unique<BitonicSort<N>>(s->cc, s->keys.publicKey, s->rotations, s->enc);
    s->len = N, s->depth = (int)p.GetMultiplicativeDepth();
    s->cc->Sync();
    s->setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return s.release();
}

template <int N> SessionBase *make_session(int log_ring_dim, int depth_override, uint64_t seed, int device, bool hybrid = false) {
    auto t0 = std::chrono::steady_clock::now();
    auto s = std::make_unique<Session<N>>();
    CCParams<CryptoContextCKKSRNS> p;
    p.SetRingDim(1u << log_ring_dim);
    DirectSort<N>::getSizeParameters(p, s->rotations);
    if (hybrid) { // sort_hybrid: the deeper chain and the placement's rotation steps (tests/DirectSortHTest.cpp:22-110)
        p.SetMultiplicativeDepth(DirectSortHybridDepth(N));
        for (int r : DirectSortHybridExtraRotations(N))
            if (std::find(s->rotations.begin(), s->rotations.end(), r) == s->rotations.end()) s->rotations.push_back(r);
    }
    if (depth_override > 0) p.SetMultiplicativeDepth(depth_override);
    p.SetSecurityLevel(log_ring_dim >= 17 ? HEStd_128_classic : HEStd_NotSet);
    CryptoContextImpl::SetKeySeed(seed); // 0: OS entropy (ChaCha20); non-zero: TEST-ONLY deterministic keys
    s->seeded = seed != 0;
#ifdef REFERENCE_DRIVERS
    setenv("FHEB200_DEVICE", std::to_string(device).c_str(), 1);
    s->cc = GenCryptoContext(p);
#else
    s->cc = GenCryptoContext(p, device);
#endif
    s->cc->Enable(PKE), s->cc->Enable(KEYSWITCH), s->cc->Enable(LEVELEDSHE), s->cc->Enable(ADVANCEDSHE);
    s->keys = s->cc->KeyGen();
    s->cc->EvalRotateKeyGen(s->keys.secretKey, s->rotations);
    s->cc->EvalMultKeyGen(s->keys.secretKey);
    s->enc = std::make_shared<DebugEncryption>(s->cc, s->keys);
    s->sorter = std::make_unique<DirectSort<N>>(s->cc, s->keys.publicKey, s->rotations, s->enc);
    s->composer = std::make_unique<RotationComposer<N>>(s->cc, s->enc, s->rotations);
    s->len = N, s->depth = (int)p.GetMultiplicativeDepth();
    s->cc->Sync();
    s->setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return s.release();
}
// Fixture of the reference's primitive-level tests (tests/RotationTest.cpp:30-54, SignTest.cpp:22-39, SincTest.cpp,
// utils/sign_characterizer.cpp:26-38): a context with caller-chosen depth / scaling size / batch and rotation keys, no sorter.
struct ToolkitSession : SessionBase {
    std::unique_ptr<RotationComposer<128>> composer; // RotationTest.cpp:51: array_length = 128, NAF
    Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &, SignConfig &) override { throw std::runtime_error("toolkit session: no sorter"); }
    Ciphertext<DCRTPoly> rank(const Ciphertext<DCRTPoly> &, SignConfig &) override { throw std::runtime_error("toolkit session: no sorter"); }
    Ciphertext<DCRTPoly> place(const Ciphertext<DCRTPoly> &, const Ciphertext<DCRTPoly> &) override {
        throw std::runtime_error("toolkit session: no sorter");
    }
    Ciphertext<DCRTPoly> composed_rotate(const Ciphertext<DCRTPoly> &in, int r) override { return composer->rotate(in, r); }
    void shard(int, int world) override {
        if (world > 1) throw std::runtime_error("toolkit session: nothing to shard");
    }
};
SessionBase *make_toolkit_session(int log_ring_dim, int depth, int scaling_bits, int batch, const int *keys, int num_keys, uint64_t seed,
                                  int device, int boot_budget) {
    auto t0 = std::chrono::steady_clock::now();
    auto s = std::make_unique<ToolkitSession>();
    CCParams<CryptoContextCKKSRNS> p;
    p.SetRingDim(1u << log_ring_dim);
    p.SetMultiplicativeDepth(depth);
    p.SetScalingModSize(scaling_bits);
    p.SetBatchSize(batch);
    p.SetSecurityLevel(HEStd_NotSet);
    CryptoContextImpl::SetKeySeed(seed); // 0: OS entropy (ChaCha20); non-zero: TEST-ONLY deterministic keys
    s->seeded = seed != 0;
#ifdef REFERENCE_DRIVERS
    setenv("FHEB200_DEVICE", std::to_string(device).c_str(), 1);
    s->cc = GenCryptoContext(p);
#else
    s->cc = GenCryptoContext(p, device);
#endif
    s->cc->Enable(PKE), s->cc->Enable(KEYSWITCH), s->cc->Enable(LEVELEDSHE), s->cc->Enable(ADVANCEDSHE);
    if (boot_budget > 0) s->cc->Enable(FHE);
    s->keys = s->cc->KeyGen();
    if (boot_budget > 0) { // tests/k-way/SorterTest.cpp:40-43
        s->cc->EvalBootstrapSetup({(uint32_t)boot_budget, (uint32_t)boot_budget}, {0, 0}, (uint32_t)batch);
        s->cc->EvalBootstrapKeyGen(s->keys.secretKey, (uint32_t)batch);
    }
    s->rotations.assign(keys, keys + num_keys);
    if (num_keys > 0) s->cc->EvalRotateKeyGen(s->keys.secretKey, s->rotations);
    s->cc->EvalMultKeyGen(s->keys.secretKey);
    s->enc = std::make_shared<DebugEncryption>(s->cc, s->keys);
    if (num_keys > 0) s->composer = std::make_unique<RotationComposer<128>>(s->cc, s->enc, s->rotations, DecomposeAlgo::NAF);
    s->len = batch, s->depth = depth;
    s->cc->Sync();
    s->setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return s.release();
}
SignConfig sign_config(int n, int dg, int df, int N) {
    if (n == 0) { // the reference tests' choice per array length (tests/DirectSortTest.cpp:103-111)
        n = 3, df = 2;
        dg = N <= 16 ? 2 : N <= 128 ? 3 : N <= 512 ? 4 : 5;
    }
    return SignConfig(CompositeSignConfig(n, dg, df));
}
} // namespace

struct fhesort_session {
    std::unique_ptr<SessionBase> s;
};

#define GUARD(...)                  \
    try {                           \
        __VA_ARGS__;                \
        return 0;                   \
    } catch (const std::exception &e) { \
        g_err = e.what();           \
        return 1;                   \
    } catch (...) {                 \
        g_err = "unknown error";    \
        return 1;                   \
    }

// RotationTree<N>(cc, keys, algo).buildTree(start, end) then treeRotate(ct, r) for every r in `rotations`
// (tests/RotationTest.cpp:77-109); N in {8, 128}; outputs [nrot][count]
template <int N>
void tree_rotate_all(SessionBase &S, int algo, int start, int end, const Ciphertext<DCRTPoly> &ct, const int *rotations, int nrot,
                            int count, double *outputs) {
    RotationTree<N> tree(S.cc, S.rotations, algo == 0 ? DecomposeAlgo::NAF : algo == 1 ? DecomposeAlgo::BNAF : DecomposeAlgo::BINARY);
    tree.buildTree(start, end);
    for (int i = 0; i < nrot; i++) {
        auto v = S.enc->getPlaintext(tree.treeRotate(ct, rotations[i]), 0.0);
        for (int k = 0; k < count; k++) outputs[(size_t)i * count + k] = v[k];
    }
}

extern "C" {

const char *fhesort_last_error(void) { return g_err.c_str(); }

int fhesort_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out) {
    GUARD({
        if (!out) throw std::invalid_argument("null argument: out");
        SessionBase *b = nullptr;
        switch (N) {
        case 4: b = make_session<4>(log_ring_dim, mult_depth, seed, device); break;
        case 8: b = make_session<8>(log_ring_dim, mult_depth, seed, device); break;
        case 16: b = make_session<16>(log_ring_dim, mult_depth, seed, device); break;
        case 32: b = make_session<32>(log_ring_dim, mult_depth, seed, device); break;
        case 64: b = make_session<64>(log_ring_dim, mult_depth, seed, device); break;
        case 128: b = make_session<128>(log_ring_dim, mult_depth, seed, device); break;
        case 256: b = make_session<256>(log_ring_dim, mult_depth, seed, device); break;
        case 512: b = make_session<512>(log_ring_dim, mult_depth, seed, device); break;
        case 1024: b = make_session<1024>(log_ring_dim, mult_depth, seed, device); break;
        default: throw std::invalid_argument("DirectSort: array length must be a power of two in 4..1024");
        }
        *out = new fhesort_session{std::unique_ptr<SessionBase>(b)};
    })
}
int fhesort_hybrid_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out) {
    GUARD({
        if (!out) throw std::invalid_argument("null argument: out");
        SessionBase *b = nullptr;
        switch (N) {
        case 4: b = make_session<4>(log_ring_dim, mult_depth, seed, device, true); break;
        case 8: b = make_session<8>(log_ring_dim, mult_depth, seed, device, true); break;
        case 16: b = make_session<16>(log_ring_dim, mult_depth, seed, device, true); break;
        case 32: b = make_session<32>(log_ring_dim, mult_depth, seed, device, true); break;
        case 64: b = make_session<64>(log_ring_dim, mult_depth, seed, device, true); break;
        case 128: b = make_session<128>(log_ring_dim, mult_depth, seed, device, true); break;
        case 256: b = make_session<256>(log_ring_dim, mult_depth, seed, device, true); break;
        case 512: b = make_session<512>(log_ring_dim, mult_depth, seed, device, true); break;
        case 1024: b = make_session<1024>(log_ring_dim, mult_depth, seed, device, true); break;
        default: throw std::invalid_argument("DirectSort: array length must be a power of two in 4..1024");
        }
        *out = new fhesort_session{std::unique_ptr<SessionBase>(b)};
    })
}
int fhesort_mehp24_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out) {
    GUARD({
        if (!out) throw std::invalid_argument("null argument: out");
        *out = new fhesort_session{std::unique_ptr<SessionBase>(make_mehp24_session(N, log_ring_dim, mult_depth, seed, device))};
    })
}
int fhesort_kway_session_create(int k, int M, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out) {
    GUARD({
        if (!out) throw std::invalid_argument("null argument: out");
        if (k != 2 && k != 3 && k != 5) throw std::invalid_argument("k-way: k must be 2, 3 or 5");
        if (M < 1 || M > 10) throw std::invalid_argument("k-way: M out of range");
        *out = new fhesort_session{std::unique_ptr<SessionBase>(make_kway_session_for(k, M, log_ring_dim, mult_depth, seed, device))};
    })
}
int fhesort_bitonic_session_create(int N, int log_ring_dim, int mult_depth, uint64_t seed, int device, fhesort_session **out) {
    GUARD({
        if (!out) throw std::invalid_argument("null argument: out");
        SessionBase *b = nullptr;
        switch (N) {
        case 4: b = make_bitonic_session<4>(log_ring_dim, mult_depth, seed, device); break;
        case 8: b = make_bitonic_session<8>(log_ring_dim, mult_depth, seed, device); break;
        case 16: b = make_bitonic_session<16>(log_ring_dim, mult_depth, seed, device); break;
        case 32: b = make_bitonic_session<32>(log_ring_dim, mult_depth, seed, device); break;
        case 64: b = make_bitonic_session<64>(log_ring_dim, mult_depth, seed, device); break;
        case 128: b = make_bitonic_session<128>(log_ring_dim, mult_depth, seed, device); break;
        default: throw std::invalid_argument("BitonicSort: array length must be a power of two in 4..128");
        }
        *out = new fhesort_session{std::unique_ptr<SessionBase>(b)};
    })
}
int fhesort_session_bootstraps(fhesort_session *s, int *count) {
    GUARD({
        if (!s || !count) throw std::invalid_argument("null argument");
        *count = s->s->bootstraps();
    })
}
int fhesort_session_free(fhesort_session *s) {
    GUARD({ delete s; })
}
int fhesort_session_info(fhesort_session *s, int *mult_depth, int *num_rotation_keys, double *setup_ms) {
    GUARD({
        if (!s) throw std::invalid_argument("null argument: session");
        if (mult_depth) *mult_depth = s->s->depth;
        if (num_rotation_keys) *num_rotation_keys = (int)s->s->rotations.size();
        if (setup_ms) *setup_ms = s->s->setup_ms;
    })
}

int fhesort_session_set_shard(fhesort_session *s, int rank, int world, fhesort_allreduce_fn fn, void *user) {
    GUARD({
        if (!s) throw std::invalid_argument("null argument: session");
        if (world < 1 || rank < 0 || rank >= world) throw std::invalid_argument("set_shard: bad rank / world");
        if (world > 1 && !fn) throw std::invalid_argument("set_shard: an all-reduce callback is required when world > 1");
        // partial residues (< 2^60 for the first, extra and special limbs) are summed as int64: 16 summands cannot wrap
        if (world > 16) throw std::invalid_argument("set_shard: at most 16 ranks (the all-reduce sums raw residues below 2^60 in 64 bits)");
        // every rank must hold the same keys: with OS-entropy keys (seed 0) each rank would draw its own
        if (world > 1 && !s->s->seeded)
            throw std::invalid_argument("set_shard: sharding needs replicated keys - create the session on every rank with the same "
                                        "non-zero seed (the sessions of this library have no key-exchange step)");
        s->s->shard(rank, world);
        if (world > 1)
            s->s->cc->allReduceWords = [fn, user](void *p, size_t w) {
                if (fn(user, p, w) != 0) throw std::runtime_error("all-reduce callback failed");
            };
        else s->s->cc->allReduceWords = nullptr;
    })
}

// the same with the engine's own NCCL communicator (fheb200_comm_unique_id / fheb200_ctx_comm_init): no callback, no host
// synchronisation around the all-reduce
int fhesort_session_comm_init(fhesort_session *s, const unsigned char *id128, int rank, int world) {
    GUARD({
        if (!s) throw std::invalid_argument("null argument: session");
        if (world < 1 || rank < 0 || rank >= world) throw std::invalid_argument("comm_init: bad rank / world");
        if (world > 1 && !id128) throw std::invalid_argument("comm_init: the 128-byte NCCL id is required when world > 1");
        if (world > 1 && !s->s->seeded)
            throw std::invalid_argument("comm_init: sharding needs replicated keys - create the session on every rank with the same "
                                        "non-zero seed (the sessions of this library have no key-exchange step)");
#ifdef REFERENCE_DRIVERS
        if (world > 1) throw std::runtime_error("the reference's drivers have no multi-GPU sharding");
#else
        s->s->cc->CommInit(id128, rank, world);
        s->s->cc->allReduceWords = nullptr;
        s->s->shard(rank, world);
#endif
    })
}

int fhesort_session_set_streams(fhesort_session *s, int k) {
    GUARD({
        if (!s) throw std::invalid_argument("null argument: session");
        s->s->streams(k);
    })
}

// which: 0 = full sort, 1 = constructRank only (output = ranks), 2 = rotationIndexCheckN on plaintext ranks given in aux
int fhesort_run(fhesort_session *s, int which, const double *input, const double *aux, int sign_n, int dg, int df,
                double *output, double *elapsed_ms, int *out_level) {
    GUARD({
        if (!s || !input || !output) throw std::invalid_argument("null argument");
        SessionBase &S = *s->s;
        std::vector<double> in(input, input + S.len);
        SignConfig cfg = sign_config(sign_n, dg, df, S.len);
        S.own_sign_config(sign_n, dg, df, cfg);
        S.cc->Sync();
        auto t0 = std::chrono::steady_clock::now();
        auto ct = S.enc->encryptInput(in);
        Ciphertext<DCRTPoly> r;
        if (which == 0) r = S.sort(ct, cfg);
        else if (which == 1) r = S.rank(ct, cfg);
        else if (which == 3) r = S.sort_hybrid(ct, cfg);
        else if (which == 2) {
            if (!aux) throw std::invalid_argument("null argument: aux (ranks)");
            r = S.place(S.enc->encryptInput(std::vector<double>(aux, aux + S.len)), ct);
        } else throw std::invalid_argument("fhesort_run: unknown stage");
        if (getenv("FHESORT_TIMING")) { // how far the launching thread runs ahead of the GPU
            double host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            S.cc->Sync();
            double gpu_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            fprintf(stderr, "fhesort_run: all work enqueued after %.1f ms, GPU done after %.1f ms\n", host_ms, gpu_ms);
        }
        std::vector<double> v = S.enc->getPlaintext(r, 0.0); // decrypt synchronises the stream
        double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        for (int i = 0; i < S.len; i++) output[i] = i < (int)v.size() ? v[i] : 0.0;
        if (elapsed_ms) *elapsed_ms = ms;
        if (out_level) *out_level = (int)r->GetLevel();
    })
}

int fhesort_toolkit_session_create(int log_ring_dim, int mult_depth, int scaling_mod_size, int batch_size, const int *rotation_keys,
                                   int num_keys, int bootstrap_budget, uint64_t seed, int device, fhesort_session **out) {
    GUARD({
        if (!out || (num_keys > 0 && !rotation_keys)) throw std::invalid_argument("null argument");
        *out = new fhesort_session{std::unique_ptr<SessionBase>(
            make_toolkit_session(log_ring_dim, mult_depth, scaling_mod_size, batch_size, rotation_keys, num_keys, seed, device, bootstrap_budget))};
    })
}
// sign(x, cc, func, cfg) of src/sign.cpp:635-651 on a fresh encryption of `input`; func: 0 CompositeSign(n, dg, df),
// 1 SignumPolycircuit, 2 Tanh, 3 NaiveDiscrete
int fhesort_sign_eval(fhesort_session *s, int func, int n, int dg, int df, const double *input, int count, double *output, int *level,
                      double *elapsed_ms) {
    GUARD({
        if (!s || !input || !output) throw std::invalid_argument("null argument");
        if (func < 0 || func > 3) throw std::invalid_argument("sign: unknown function");
        SessionBase &S = *s->s;
        const SignFunc f = func == 0 ? SignFunc::CompositeSign : func == 1 ? SignFunc::SignumPolycircuit : func == 2 ? SignFunc::Tanh : SignFunc::NaiveDiscrete;
        SignConfig cfg = func == 0 ? SignConfig(CompositeSignConfig(n, dg, df)) : SignConfig();
        auto ct = S.enc->encryptInput(std::vector<double>(input, input + count));
        S.cc->Sync();
        auto t0 = std::chrono::steady_clock::now();
        auto r = sign(ct, S.cc, f, cfg);
        S.cc->Sync();
        if (elapsed_ms) *elapsed_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (level) *level = (int)r->GetLevel();
        auto v = S.enc->getPlaintext(r, 0.0);
        for (int i = 0; i < count; i++) output[i] = v[i];
    })
}
int fhesort_tree_rotate(fhesort_session *s, int N, int algo, int start, int end, const double *input, int count, const int *rotations,
                        int nrot, double *outputs) {
    GUARD({
        if (!s || !input || !rotations || !outputs) throw std::invalid_argument("null argument");
        SessionBase &S = *s->s;
        auto ct = S.enc->encryptInput(std::vector<double>(input, input + count));
        if (N == 8) tree_rotate_all<8>(S, algo, start, end, ct, rotations, nrot, count, outputs);
        else if (N == 128) tree_rotate_all<128>(S, algo, start, end, ct, rotations, nrot, count, outputs);
        else throw std::invalid_argument("tree_rotate: N must be 8 or 128");
    })
}

// One call of a kwaySort::EvalUtils / SortUtils / Sorter method on fresh encryptions of the given slot vectors - what the
// reference's tests/k-way/{EvalUtilsTest,SortUtilsTest,SorterTest}.cpp exercise.  op:
//   1 multByInt(in0, param)         2 multAndKillImage(in0, in1)     3 squareAndKillImage(in0)
//   4 leftRotate(in0, param)        5 rightRotate(in0, param)        6 checkLevelAndBoot(in0 encoded at level `aux`, param, depth)
//   10 fcnL(in0, in1, comp = in2)   11 twoSorter({in0, in1}, comp = in2) -> out[0..count) min, out[count..2count) max
//   22..25 run{Two,Three,Four,Five}Sorter(in0, indices, shift = param, comp1 = in1, comp2 = in2)
// out: count values (2 * count for op 11); levels: level of in0 before / of the result after.
int fhesort_kway_unit(fhesort_session *s, int op, long k, long M, long param, int aux, const double *in0, const double *in1,
                      const double *in2, const int *idx0, const int *idx1, int count, double *out, int *level_before, int *level_after) {
    GUARD({
        if (!s || !in0 || !out) throw std::invalid_argument("null argument");
        SessionBase &S = *s->s;
        const long slots = (long)S.cc->GetEncodingParams()->GetBatchSize();
        if (count < 1 || count > slots) throw std::invalid_argument("kway_unit: count must be 1..batch size");
        auto enc = [&](const double *v, int level) {
            if (!v) throw std::invalid_argument("kway_unit: missing input vector");
            auto pt = level > 0 ? S.cc->MakeCKKSPackedPlaintext(std::vector<double>(v, v + count), 1, (uint32_t)level, nullptr, (uint32_t)slots)
                                : S.cc->MakeCKKSPackedPlaintext(std::vector<double>(v, v + count));
            return S.cc->Encrypt(S.keys.publicKey, pt);
        };
        auto put = [&](const Ciphertext<DCRTPoly> &c, double *dst) {
            auto v = S.enc->getPlaintext(c, 0.0);
            for (int i = 0; i < count; i++) dst[i] = v[i];
        };
        kwaySort::Sorter sorter(S.cc, S.enc, slots, k, M);
        Ciphertext<DCRTPoly> a = enc(in0, op == 6 ? aux : 0), r;
        if (level_before) *level_before = (int)a->GetLevel();
        if (op == 1) sorter.multByInt(a, param, r);
        else if (op == 2) { auto b = enc(in1, 0); sorter.multAndKillImage(a, b, r); }
        else if (op == 3) sorter.squareAndKillImage(a, r);
        else if (op == 4) sorter.leftRotate(a, param, r);
        else if (op == 5) sorter.rightRotate(a, param, r);
        else if (op == 6) { sorter.checkLevelAndBoot(a, (int)param, S.depth); r = a; }
        else if (op == 10) { auto b = enc(in1, 0), c = enc(in2, 0); sorter.fcnL(a, b, c, r); }
        else if (op == 11) {
            Ciphertext<DCRTPoly> pair[2] = {a, enc(in1, 0)}, res[2];
            auto c = enc(in2, 0);
            sorter.twoSorter(pair, c, res);
            put(res[1], out + count);
            r = res[0];
        } else if (op >= 22 && op <= 25) {
            if (!idx0 || !idx1) throw std::invalid_argument("kway_unit: missing indices");
            std::vector<std::vector<int>> indices(2, std::vector<int>((size_t)slots, 0));
            for (int i = 0; i < count; i++) indices[0][i] = idx0[i], indices[1][i] = idx1[i];
            auto c1 = enc(in1, 0);
            if (op == 22) sorter.runTwoSorter(a, indices, param, c1, r);
            else if (op == 23) sorter.runThreeSorter(a, indices, param, c1, r);
            else { auto c2 = enc(in2, 0); op == 24 ? sorter.runFourSorter(a, indices, param, c1, c2, r) : sorter.runFiveSorter(a, indices, param, c1, c2, r); }
        } else throw std::invalid_argument("kway_unit: unknown op");
        if (level_after) *level_after = (int)r->GetLevel();
        put(r, out);
    })
}

// RotationComposer::rotate on a fresh encryption of `input` (count values, `slots` slots)
int fhesort_composed_rotate(fhesort_session *s, const double *input, int count, int rotation, double *output) {
    GUARD({
        if (!s || !input || !output) throw std::invalid_argument("null argument");
        SessionBase &S = *s->s;
        auto ct = S.enc->encryptInput(std::vector<double>(input, input + count));
        auto v = S.enc->getPlaintext(S.composed_rotate(ct, rotation), 0.0);
        for (int i = 0; i < count; i++) output[i] = v[i];
    })
}

#ifndef REFERENCE_DRIVERS
// host-only helpers (no GPU): the rotation-key list and Chebyshev table sizes the drivers would use
int fhesort_rotation_indices(int N, int log_ring_dim, int *out, int capacity, int *count) {
    GUARD({
        std::vector<int> r;
        switch (N) {
        case 4: r = DirectSort<4>::rotationIndices(1 << log_ring_dim); break;
        case 8: r = DirectSort<8>::rotationIndices(1 << log_ring_dim); break;
        case 16: r = DirectSort<16>::rotationIndices(1 << log_ring_dim); break;
        case 32: r = DirectSort<32>::rotationIndices(1 << log_ring_dim); break;
        case 64: r = DirectSort<64>::rotationIndices(1 << log_ring_dim); break;
        case 128: r = DirectSort<128>::rotationIndices(1 << log_ring_dim); break;
        case 256: r = DirectSort<256>::rotationIndices(1 << log_ring_dim); break;
        case 512: r = DirectSort<512>::rotationIndices(1 << log_ring_dim); break;
        case 1024: r = DirectSort<1024>::rotationIndices(1 << log_ring_dim); break;
        default: throw std::invalid_argument("unsupported N");
        }
        if (count) *count = (int)r.size();
        for (int i = 0; i < (int)r.size() && i < capacity; i++) out[i] = r[i];
    })
}
int fhesort_shard_batches(int N, int log_ring_dim, int rank, int world, int *out, int capacity, int *count, int *sharded) {
    GUARD({
        const DirectSortShape sh(N, (1 << log_ring_dim) / 2);
        std::vector<int> b = DirectSortBatchesOf(sh.B, rank, world);
        if (count) *count = (int)b.size();
        if (sharded) *sharded = world > 1 && sh.B % world == 0;
        for (int i = 0; i < (int)b.size() && i < capacity; i++) out[i] = b[i];
    })
}
#endif
// host-only: the network geometry of one k-way stage (kwaySort::sortType / genIndices / getRotateDistance); built from
// this repository's kway.h in libfhesort.so and from the reference's src/k-way/Masking.cpp in oracle/_ref/libref_sort.so
int fhesort_kway_stage(int k, int M, int stage, int num_slots, int *geometry4, int *indices0, int *indices1) {
    GUARD({
        if (!geometry4) throw std::invalid_argument("null argument: geometry");
        auto t = kwaySort::sortType(k, M, stage);
        const int m = std::get<0>(t), logDist = std::get<1>(t), slope = std::get<2>(t);
        geometry4[0] = m, geometry4[1] = logDist, geometry4[2] = slope;
        geometry4[3] = (int)kwaySort::getRotateDistance(k, logDist, slope);
        if (indices0 && indices1) {
            auto idx = kwaySort::genIndices(num_slots, k, M, m, logDist, slope);
            for (int i = 0; i < num_slots; i++) indices0[i] = idx[0][i], indices1[i] = idx[1][i];
        }
    })
}
int fhesort_decompose(const int *key_indices, int num_keys, int rotation, int wrap, int algo, int *steps, int capacity, int *count) {
    GUARD({
        Decomposer<1024> d(std::vector<int>(key_indices, key_indices + num_keys));
        auto st = d.decompose(rotation, wrap, algo == 0 ? DecomposeAlgo::NAF : algo == 1 ? DecomposeAlgo::BNAF : DecomposeAlgo::BINARY);
        if (count) *count = (int)st.size();
        for (int i = 0; i < (int)st.size() && i < capacity; i++) steps[i] = st[i].stepSize; // signed amount (reference convention)
    })
}
#ifndef REFERENCE_DRIVERS
int fhesort_doubled_sinc_coefficients(int N, double *out, int capacity, int *count) {
    GUARD({
        const std::vector<double> *c = nullptr;
        switch (N) {
        case 4: c = &selectDoubledSincCoefficients<4>(); break;
        case 8: c = &selectDoubledSincCoefficients<8>(); break;
        case 16: c = &selectDoubledSincCoefficients<16>(); break;
        case 32: c = &selectDoubledSincCoefficients<32>(); break;
        case 64: c = &selectDoubledSincCoefficients<64>(); break;
        case 128: c = &selectDoubledSincCoefficients<128>(); break;
        case 256: c = &selectDoubledSincCoefficients<256>(); break;
        case 512: c = &selectDoubledSincCoefficients<512>(); break;
        case 1024: c = &selectDoubledSincCoefficients<1024>(); break;
        default: throw std::invalid_argument("unsupported N");
        }
        if (count) *count = (int)c->size();
        for (int i = 0; i < (int)c->size() && i < capacity; i++) out[i] = (*c)[i];
    })
}

int fhesort_sinc_coefficients(int N, double *out, int capacity, int *count) {
    GUARD({
        const std::vector<double> *c = nullptr;
        switch (N) {
        case 4: c = &selectCoefficients<4>(); break;
        case 8: c = &selectCoefficients<8>(); break;
        case 16: c = &selectCoefficients<16>(); break;
        case 32: c = &selectCoefficients<32>(); break;
        case 64: c = &selectCoefficients<64>(); break;
        case 128: c = &selectCoefficients<128>(); break;
        case 256: c = &selectCoefficients<256>(); break;
        case 512: c = &selectCoefficients<512>(); break;
        case 1024: c = &selectCoefficients<1024>(); break;
        default: throw std::invalid_argument("unsupported N");
        }
        if (count) *count = (int)c->size();
        for (int i = 0; i < (int)c->size() && i < capacity; i++) out[i] = (*c)[i];
    })
}

#endif

} // extern "C"
