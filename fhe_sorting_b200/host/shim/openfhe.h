// openfhe.h — lbcrypto-shaped C++ shim over the C ABI of libfheb200.so (include/fhe_b200.h).
//
// This header gives the reference's host code (src/sort_algo.h, rotation.h, sign.cpp, comparison.cpp, encryption.cpp of
// oksuman/FHE-Sorting) the subset of OpenFHE's CryptoContext<DCRTPoly> API that it calls: same names,
// argument meaning and error behaviour (exceptions), every call forwarding to the CUDA engine.  There is no arithmetic
// in this file.  Semantics mirrored from OpenFHE: FLEXIBLEAUTOEXT level bookkeeping (done inside the engine), in-place
// operations mutate the shared ciphertext object, SetSlots is metadata on a shared handle (src/sort_algo.h:429,614).
#pragma once
#include <algorithm>
#include <cassert>
#include <chrono>
#include <cmath>
#include <fstream>
#include <iomanip>
#include <map>
#include <numeric>
#include <set>
#include <cstdint>
#include <cstring>
#include <functional>
#include <iostream>
#include <memory>
#include <random>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../include/fhe_b200.h"

namespace lbcrypto {

typedef unsigned int usint;
struct PrecompHandle;                // owner of a fheb200_pre (hoisted ModUp digits in HBM)
struct DCRTPoly {                    // tag: polynomials live in HBM behind the C ABI.  The one place the reference holds
    std::shared_ptr<PrecompHandle> pre; // DCRTPoly values is the result of EvalFastRotationPrecompute (src/rotation.h:247)
};
struct CryptoContextCKKSRNS {};
enum SecurityLevel { HEStd_128_classic, HEStd_192_classic, HEStd_256_classic, HEStd_NotSet };
// OpenFHE's serialisation selector (src/sort.h:34-71 uses SerType::BINARY only; JSON is not provided)
namespace SerType {
struct SERBINARY {};
struct SERJSON {};
static const SERBINARY BINARY{};
static const SERJSON JSON{};
} // namespace SerType
enum PKESchemeFeature { PKE = 1, KEYSWITCH = 2, PRE = 4, LEVELEDSHE = 8, ADVANCEDSHE = 16, MULTIPARTY = 32, FHE = 64 };

inline void fheb200_check(int rc) {
    if (rc != 0) throw std::runtime_error(std::string("fheb200: ") + fheb200_last_error());
}

template <typename Scheme> class CCParams {
    usint depth_ = 1, scale_ = 50, first_ = 60, batch_ = 0, ring_ = 0, dnum_ = 3;
    SecurityLevel sec_ = HEStd_128_classic;

  public:
    void SetMultiplicativeDepth(usint d) { depth_ = d; }
    void SetScalingModSize(usint s) { scale_ = s; }
    void SetFirstModSize(usint s) { first_ = s; }
    void SetBatchSize(usint b) { batch_ = b; }
    void SetRingDim(usint r) { ring_ = r; }
    void SetSecurityLevel(SecurityLevel s) { sec_ = s; }
    void SetNumLargeDigits(usint d) { dnum_ = d; }
    usint GetMultiplicativeDepth() const { return depth_; }
    usint GetScalingModSize() const { return scale_; }
    usint GetFirstModSize() const { return first_; }
    usint GetBatchSize() const { return batch_; }
    usint GetRingDim() const { return ring_; }
    usint GetNumLargeDigits() const { return dnum_; }
    SecurityLevel GetSecurityLevel() const { return sec_; }
};

class CryptoContextImpl;
typedef std::shared_ptr<CryptoContextImpl> ContextPtr;

// ---- plaintext
class PlaintextImpl {
  public:
    fheb200_pt *h = nullptr;
    ~PlaintextImpl() {
        if (h) fheb200_pt_release(h);
    }
    std::vector<double> GetRealPackedValue() const {
        size_t n = 0;
        fheb200_check(fheb200_pt_values(h, nullptr, 0, &n));
        std::vector<double> v(n);
        fheb200_check(fheb200_pt_values(h, v.data(), n, &n));
        return v;
    }
    void SetLength(size_t n) { fheb200_check(fheb200_pt_set_length(h, n)); }
    double GetLogPrecision() const {
        double b = 0;
        fheb200_check(fheb200_pt_log_precision(h, &b));
        return b;
    }
};
typedef std::shared_ptr<PlaintextImpl> Plaintext;
typedef std::shared_ptr<const PlaintextImpl> ConstPlaintext;
inline std::ostream &operator<<(std::ostream &os, const Plaintext &p) {
    os << "(";
    for (double v : p->GetRealPackedValue()) os << v << ", ";
    return os << " ... ); Estimated precision: " << p->GetLogPrecision() << " bits";
}

// ---- ciphertext: a nullable, copyable handle to a shared mutable object (as lbcrypto::Ciphertext<DCRTPoly>)
template <typename P> class Ciphertext;
class CiphertextImpl {
  public:
    fheb200_ct *h = nullptr;
    ContextPtr cc;
    CiphertextImpl(fheb200_ct *h_, ContextPtr cc_) : h(h_), cc(std::move(cc_)) {}
    CiphertextImpl(const CiphertextImpl &) = delete;
    ~CiphertextImpl() {
        if (h) fheb200_ct_release(h);
    }
    void replace(fheb200_ct *n) { // in-place operations swap the engine object under the shared handle
        if (h) fheb200_ct_release(h);
        h = n;
    }
    size_t GetLevel() const {
        int lv = 0;
        fheb200_check(fheb200_ct_info(h, &lv, nullptr, nullptr, nullptr, nullptr));
        return (size_t)lv;
    }
    usint GetSlots() const {
        int s = 0;
        fheb200_check(fheb200_ct_info(h, nullptr, nullptr, &s, nullptr, nullptr));
        return (usint)s;
    }
    usint GetNoiseScaleDeg() const {
        int d = 0;
        fheb200_check(fheb200_ct_info(h, nullptr, &d, nullptr, nullptr, nullptr));
        return (usint)d;
    }
    double GetScalingFactor() const {
        double s = 0;
        fheb200_check(fheb200_ct_info(h, nullptr, nullptr, nullptr, nullptr, &s));
        return s;
    }
    void SetSlots(usint s) const { fheb200_check(fheb200_ct_set_slots(h, (int)s)); }
    ContextPtr GetCryptoContext() const { return cc; }
    Ciphertext<DCRTPoly> Clone() const;
};
template <typename P> class Ciphertext {
    std::shared_ptr<CiphertextImpl> p_;

  public:
    Ciphertext() = default;
    Ciphertext(std::nullptr_t) {}
    explicit Ciphertext(std::shared_ptr<CiphertextImpl> p) : p_(std::move(p)) {}
    CiphertextImpl *operator->() const { return p_.get(); }
    CiphertextImpl &operator*() const { return *p_; }
    explicit operator bool() const { return (bool)p_; }
    bool operator==(std::nullptr_t) const { return !p_; }
    bool operator!=(std::nullptr_t) const { return (bool)p_; }
    CiphertextImpl *get() const { return p_.get(); }
};
template <typename P> using ConstCiphertext = Ciphertext<P>;

// ---- keys
struct KeySetImpl {
    fheb200_keys *h = nullptr;
    ~KeySetImpl() {
        if (h) fheb200_keys_release(h);
    }
};
template <typename P> class PublicKey {
  public:
    std::shared_ptr<KeySetImpl> k;
    explicit operator bool() const { return (bool)k; }
};
template <typename P> class PrivateKey {
  public:
    std::shared_ptr<KeySetImpl> k;
    explicit operator bool() const { return (bool)k; }
};
template <typename P> struct KeyPair {
    PublicKey<P> publicKey;
    PrivateKey<P> secretKey;
    bool good() const { return (bool)publicKey.k; }
};
struct EncodingParamsImpl {
    usint batch = 0;
    usint GetBatchSize() const { return batch; }
};
struct PrecompHandle {
    fheb200_pre *h = nullptr;
    ~PrecompHandle() {
        if (h) fheb200_pre_release(h);
    }
};
typedef std::vector<DCRTPoly> FastRotationPrecomp; // OpenFHE's type: std::shared_ptr<std::vector<DCRTPoly>>; element 0 owns the digits

// ---- the crypto context
class CryptoContextImpl : public std::enable_shared_from_this<CryptoContextImpl> {
    typedef Ciphertext<DCRTPoly> CT;
    CT wrap(fheb200_ct *h) { return CT(std::make_shared<CiphertextImpl>(h, shared_from_this())); }
    CT unary(int (*fn)(fheb200_ctx *, fheb200_ct *, fheb200_ct **), const CT &a) {
        need(a);
        fheb200_ct *o = nullptr;
        fheb200_check(fn(ctx, a->h, &o));
        return wrap(o);
    }
    static void need(const CT &a) {
        if (!a) throw std::runtime_error("null ciphertext passed to the crypto context");
    }

  public:
    fheb200_ctx *ctx = nullptr;
    usint ringDim = 0, batch = 0, depth = 0, features = 0;
    std::shared_ptr<KeySetImpl> evalKeys; // keys whose evaluation keys live in HBM (EvalMultKeyGen / EvalRotateKeyGen)
    ~CryptoContextImpl() {
        evalKeys.reset();
        if (ctx) fheb200_ctx_release(ctx);
    }
    void Enable(usint f) { features |= f; }
    void Enable(PKESchemeFeature f) { features |= (usint)f; }
    usint GetRingDimension() const { return ringDim; }
    usint GetCyclotomicOrder() const { return 2 * ringDim; }
    std::shared_ptr<EncodingParamsImpl> GetEncodingParams() const {
        auto e = std::make_shared<EncodingParamsImpl>();
        e->batch = batch;
        return e;
    }

    // ---- keys (tests/DirectSortTest.cpp:43-47)
    KeyPair<DCRTPoly> KeyGen() {
        auto k = std::make_shared<KeySetImpl>();
        const uint64_t seed = seed_next(); // 0 unless SetKeySeed was called: ChaCha20 from OS entropy
        if (seed) fheb200_check(fheb200_ctx_allow_deterministic_keys(ctx, 1));
        fheb200_check(fheb200_keygen(ctx, seed, &k->h));
        KeyPair<DCRTPoly> kp;
        kp.publicKey.k = k;
        kp.secretKey.k = k;
        return kp;
    }
    void EvalMultKeyGen(const PrivateKey<DCRTPoly> &sk) {
        fheb200_check(fheb200_eval_mult_keygen(ctx, sk.k->h));
        evalKeys = sk.k;
    }
    void EvalRotateKeyGen(const PrivateKey<DCRTPoly> &sk, const std::vector<int> &idx) {
        fheb200_check(fheb200_eval_rotate_keygen(ctx, sk.k->h, idx.data(), idx.size()));
        evalKeys = sk.k;
    }
    // ---- bootstrapping (tests/k-way/KWaySort2Test.cpp:42-44 ; src/k-way/EvalUtils.cpp:76 ; src/sign.cpp:168)
    // levelBudget = levels for CoeffsToSlots / SlotsToCoeffs; bsgsDim is unused (each group's few diagonals share one
    // hoisted ModUp); slots = 0 means the batch size.  Sparse packings only (2 * slots <= ring_dim / 2).
    void EvalBootstrapSetup(const std::vector<uint32_t> &levelBudget = {5, 4}, const std::vector<uint32_t> & = {0, 0},
                            uint32_t slots = 0, uint32_t = 0, bool = true) {
        if (levelBudget.size() != 2) throw std::invalid_argument("EvalBootstrapSetup: levelBudget needs two entries");
        if (!(features & FHE)) throw std::runtime_error("EvalBootstrapSetup: Enable(FHE) first");
        fheb200_check(fheb200_eval_bootstrap_setup(ctx, (int)levelBudget[0], (int)levelBudget[1], (int)slots));
    }
    void EvalBootstrapKeyGen(const PrivateKey<DCRTPoly> &sk, uint32_t slots) {
        fheb200_check(fheb200_eval_bootstrap_keygen(ctx, sk.k->h, (int)slots));
        evalKeys = sk.k;
    }
    // numIterations > 1: Meta-BTS, one more bootstrapping of the scaled error per extra iteration (engine: cabi.cu)
    CT EvalBootstrap(const CT &c, uint32_t numIterations = 1, uint32_t precision = 0) {
        need(c);
        fheb200_ct *o = nullptr;
        if (numIterations <= 1) fheb200_check(fheb200_eval_bootstrap(ctx, c->h, &o));
        else fheb200_check(fheb200_eval_bootstrap_iterated(ctx, c->h, (int)numIterations, (int)precision, &o));
        return wrap(o);
    }

    // ---- encode / encrypt / decrypt (src/encryption.cpp:5-27)
    Plaintext MakeCKKSPackedPlaintext(const std::vector<double> &v, size_t noiseScaleDeg = 1, uint32_t level = 0,
                                      const void *params = nullptr, usint slots = 0) {
        (void)params;
        auto p = std::make_shared<PlaintextImpl>();
        fheb200_check(fheb200_make_packed_plaintext(ctx, v.data(), v.size(), (int)noiseScaleDeg, (int)level, (int)slots, &p->h));
        return p;
    }
    // Engine extension (not an OpenFHE call): pt with its slots rotated left by `index`, encoded by a Galois permutation
    Plaintext RotatePlaintext(const Plaintext &pt, int index) {
        auto p = std::make_shared<PlaintextImpl>();
        fheb200_check(fheb200_pt_rotate(ctx, pt->h, index, &p->h));
        return p;
    }
    CT Encrypt(const PublicKey<DCRTPoly> &pk, const Plaintext &pt) {
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_encrypt(ctx, pk.k->h, pt->h, &o));
        return wrap(o);
    }
    void Decrypt(const PrivateKey<DCRTPoly> &sk, const CT &ct, Plaintext *out) {
        need(ct);
        auto p = std::make_shared<PlaintextImpl>();
        fheb200_check(fheb200_decrypt(ctx, sk.k->h, ct->h, &p->h));
        *out = p;
    }

    // ---- additions / subtractions
    CT EvalAdd(const CT &a, const CT &b) { return binary(fheb200_eval_add, a, b); }
    CT EvalSub(const CT &a, const CT &b) { return binary(fheb200_eval_sub, a, b); }
    CT EvalAdd(const CT &a, const Plaintext &p) { return plain(fheb200_eval_add_plain, a, p); }
    CT EvalAdd(const Plaintext &p, const CT &a) { return plain(fheb200_eval_add_plain, a, p); }
    CT EvalSub(const CT &a, const Plaintext &p) { return plain(fheb200_eval_sub_plain, a, p); }
    CT EvalSub(const Plaintext &p, const CT &a) { return EvalNegate(plain(fheb200_eval_sub_plain, a, p)); }
    CT EvalAdd(const CT &a, double c) { return constant(fheb200_eval_add_const, a, c); }
    CT EvalAdd(double c, const CT &a) { return constant(fheb200_eval_add_const, a, c); }
    CT EvalSub(const CT &a, double c) { return constant(fheb200_eval_add_const, a, -c); }
    CT EvalSub(double c, const CT &a) { return EvalNegate(constant(fheb200_eval_add_const, a, -c)); }
    CT EvalNegate(const CT &a) { return unary(fheb200_eval_negate, a); }
    void EvalNegateInPlace(CT &a) { inplace(a, EvalNegate(a)); }
    void EvalAddInPlace(CT &a, const CT &b) { inplace(a, EvalAdd(a, b)); }
    void EvalAddInPlace(CT &a, const Plaintext &p) { inplace(a, EvalAdd(a, p)); }
    void EvalAddInPlace(CT &a, double c) { inplace(a, EvalAdd(a, c)); }
    void EvalSubInPlace(CT &a, const CT &b) { inplace(a, EvalSub(a, b)); }
    void EvalSubInPlace(CT &a, double c) { inplace(a, EvalSub(a, c)); }
    CT EvalAddMany(const std::vector<CT> &v) {
        if (v.empty()) throw std::runtime_error("EvalAddMany: empty input");
        std::vector<fheb200_ct *> hs;
        for (auto &c : v) {
            need(c);
            hs.push_back(c->h);
        }
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_eval_add_many(ctx, hs.data(), hs.size(), &o));
        return wrap(o);
    }

    // ---- multiplications (src/sign.cpp:23-56, src/sort_algo.h:344,730)
    CT EvalMult(const CT &a, const CT &b) { return binary(fheb200_eval_mult, a, b); }
    CT EvalMultAndRelinearize(const CT &a, const CT &b) { return binary(fheb200_eval_mult, a, b); }
    CT EvalSquare(const CT &a) { return binary(fheb200_eval_mult, a, a); }
    // Extension (not in OpenFHE): the square already brought to degree 1, i.e. what the next multiplication's lazy ModReduce would
    // make of it, computed as one division by P * q_last (include/fhe_b200.h: fheb200_eval_mult_rescaled).  For powers that are only
    // ever multiplied again (x^2, x^4 of the composite-sign polynomials); the reference's drivers keep calling EvalSquare.
    CT EvalSquareReduced(const CT &a) { return binary(fheb200_eval_mult_rescaled, a, a); }
    CT EvalMult(const CT &a, const Plaintext &p) { return plain(fheb200_eval_mult_plain, a, p); }
    CT EvalMult(const Plaintext &p, const CT &a) { return plain(fheb200_eval_mult_plain, a, p); }
    // extension (not lbcrypto): acc += a * p in one pass; acc may be empty
    void EvalMultAddInPlace(CT &acc, const CT &a, const Plaintext &p) {
        if (!acc) {
            acc = EvalMult(a, p);
            return;
        }
        need(a);
        if (!p) throw std::runtime_error("null plaintext passed to the crypto context");
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_eval_mult_plain_add(ctx, acc->h, a->h, p->h, &o));
        acc->replace(o);
    }
    CT EvalMult(const CT &a, double c) { return constant(fheb200_eval_mult_const, a, c); }
    CT EvalMult(double c, const CT &a) { return constant(fheb200_eval_mult_const, a, c); }
    void EvalMultInPlace(CT &a, double c) { inplace(a, EvalMult(a, c)); }
    CT Rescale(const CT &a) { return unary(fheb200_eval_rescale, a); }

    // ---- rotations (src/rotation.h:224-350)
    CT EvalRotate(const CT &a, int index) {
        need(a);
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_eval_rotate(ctx, a->h, index, &o));
        return wrap(o);
    }
    std::shared_ptr<FastRotationPrecomp> EvalFastRotationPrecompute(const CT &a) {
        need(a);
        auto h = std::make_shared<PrecompHandle>();
        fheb200_check(fheb200_fast_rotation_precompute(ctx, a->h, &h->h));
        auto p = std::make_shared<FastRotationPrecomp>(1);
        (*p)[0].pre = h;
        return p;
    }
    CT EvalFastRotation(const CT &a, usint index, usint /*cyclotomic order*/, const std::shared_ptr<FastRotationPrecomp> &pre) {
        need(a);
        fheb200_ct *o = nullptr;
        if (!pre || pre->empty() || !(*pre)[0].pre) throw std::invalid_argument("EvalFastRotation: empty precomputation");
        fheb200_check(fheb200_eval_fast_rotation(ctx, a->h, (int)index, (*pre)[0].pre->h, &o));
        return wrap(o);
    }
    bool HasRotationKey(int index) const {
        int present = 0;
        fheb200_check(fheb200_has_rotation_key(ctx, index, &present));
        return present != 0;
    }

    // ---- polynomial evaluation (src/sort_algo.h:629,728 ; src/sign.cpp:76,448,616 ; src/mehp24/mehp24_utils.cpp:256)
    CT EvalChebyshevSeries(const CT &x, const std::vector<double> &coeffs, double a, double b) {
        need(x);
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_eval_chebyshev_series(ctx, x->h, coeffs.data(), coeffs.size(), a, b, &o));
        return wrap(o);
    }
    CT EvalChebyshevSeriesPS(const CT &x, const std::vector<double> &coeffs, double a, double b) {
        return EvalChebyshevSeries(x, coeffs, a, b);
    }
    CT EvalChebyshevFunction(std::function<double(double)> f, const CT &x, double a, double b, uint32_t degree);
    CT EvalPolyLinear(const CT &x, const std::vector<double> &coeffs) {
        need(x);
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_eval_poly(ctx, x->h, coeffs.data(), coeffs.size(), &o));
        return wrap(o);
    }
    CT EvalPoly(const CT &x, const std::vector<double> &coeffs) { return EvalPolyLinear(x, coeffs); }
    CT CloneCt(const CiphertextImpl &a) {
        fheb200_ct *o = nullptr;
        fheb200_check(fheb200_ct_clone(ctx, a.h, &o));
        return wrap(o);
    }
    void Sync() { fheb200_check(fheb200_ctx_sync(ctx)); }
    // Engine extension (not an OpenFHE call): concurrent streams for independent sub-computations, see fheb200_ctx_fork
    void ForkStreams(int k) { fheb200_check(fheb200_ctx_fork(ctx, k)); }
    void UseStream(int i) { fheb200_check(fheb200_ctx_use_stream(ctx, i)); }
    void BindThreadStream(int i) { fheb200_check(fheb200_ctx_bind_thread(ctx, i)); } // this host thread -> side stream i
    void UnbindThreadStream() { fheb200_check(fheb200_ctx_unbind_thread(ctx)); }
    void JoinStreams() { fheb200_check(fheb200_ctx_join(ctx)); }
    // Multi-GPU extension (not an OpenFHE call): sum a partial ciphertext over the ranks that share this job.  `reduce`
    // sums the device buffer in place across ranks as 64-bit integers (NCCL all-reduce supplied by the launcher).
    std::function<void(void *device_ptr, size_t words)> allReduceWords;
    // preferred: the engine's own NCCL communicator (fheb200_ctx_comm_init): the all-reduce and the fix-up are queued on the
    // engine's stream, nothing waits on the host.  id128 from fheb200_comm_unique_id on rank 0.
    bool nativeComm = false;
    void CommInit(const unsigned char *id128, int rank, int world) {
        fheb200_check(fheb200_ctx_comm_init(ctx, id128, rank, world));
        nativeComm = world > 1;
    }
    void AllReduceSum(CT &a) {
        need(a);
        if (nativeComm) {
            fheb200_check(fheb200_ct_all_reduce_sum(ctx, a->h));
            return;
        }
        if (!allReduceWords) return;
        void *ptr = nullptr;
        size_t words = 0;
        fheb200_check(fheb200_ct_device_view(a->h, &ptr, &words));
        Sync();
        allReduceWords(ptr, words);
        fheb200_check(fheb200_ct_reduce_partial_sums(ctx, a->h));
    }
    // ---- serialisation of evaluation keys (src/sort.h:47-66: DeserializeEvalMultKey / DeserializeEvalAutomorphismKey are
    // called through the context pointer).  Wire format: this engine's own raw-limb container (see Serial below), NOT OpenFHE's
    // cereal archives - those cannot be produced or checked on a machine without OpenFHE.
    bool SerializeEvalMultKey(std::ostream &os, const SerType::SERBINARY &) {
        int have = 0;
        fheb200_check(fheb200_has_mult_key(ctx, &have));
        if (!have) return false;
        return ser_evk_block(os, 0);
    }
    bool DeserializeEvalMultKey(std::istream &is, const SerType::SERBINARY &) { return deser_evk_stream(is, true); }
    bool SerializeEvalAutomorphismKey(std::ostream &os, const SerType::SERBINARY &) {
        size_t cnt = 0;
        fheb200_check(fheb200_evk_list(ctx, nullptr, 0, &cnt));
        std::vector<uint32_t> g(cnt);
        fheb200_check(fheb200_evk_list(ctx, g.data(), cnt, &cnt));
        ser_u64(os, SER_MAGIC_EVKS), ser_u64(os, cnt);
        for (uint32_t x : g)
            if (!ser_evk_block(os, x)) return false;
        return (bool)os;
    }
    bool DeserializeEvalAutomorphismKey(std::istream &is, const SerType::SERBINARY &) {
        if (deser_u64(is) != SER_MAGIC_EVKS) return false;
        const uint64_t cnt = deser_u64(is);
        for (uint64_t i = 0; i < cnt; i++)
            if (!deser_evk_stream(is, false)) return false;
        return true;
    }
    static constexpr uint64_t SER_MAGIC_CC = 0x3143433030324246ull, SER_MAGIC_PK = 0x314b503030324246ull,
                              SER_MAGIC_EVK = 0x4b56453030324246ull, SER_MAGIC_EVKS = 0x534b563030324246ull,
                              SER_MAGIC_CT = 0x3154433030324246ull;
    static void ser_u64(std::ostream &os, uint64_t v) { os.write(reinterpret_cast<const char *>(&v), 8); }
    static uint64_t deser_u64(std::istream &is) {
        uint64_t v = 0;
        is.read(reinterpret_cast<char *>(&v), 8);
        if (!is) throw std::runtime_error("Deserialize: truncated stream");
        return v;
    }
    bool ser_evk_block(std::ostream &os, uint32_t galois) {
        size_t words = 0;
        fheb200_check(fheb200_ctx_evk_words(ctx, &words));
        std::vector<uint64_t> buf(words);
        fheb200_check(fheb200_evk_export(ctx, galois, buf.data(), words));
        ser_u64(os, SER_MAGIC_EVK), ser_u64(os, galois), ser_u64(os, words);
        os.write(reinterpret_cast<const char *>(buf.data()), (std::streamsize)(words * 8));
        return (bool)os;
    }
    bool deser_evk_stream(std::istream &is, bool want_mult) {
        if (deser_u64(is) != SER_MAGIC_EVK) return false;
        const uint64_t galois = deser_u64(is), words = deser_u64(is);
        size_t expect = 0;
        fheb200_check(fheb200_ctx_evk_words(ctx, &expect));
        if (words != expect || (want_mult && galois != 0)) return false;
        std::vector<uint64_t> buf(words);
        is.read(reinterpret_cast<char *>(buf.data()), (std::streamsize)(words * 8));
        if (!is) return false;
        fheb200_check(fheb200_evk_import(ctx, (uint32_t)galois, buf.data(), words));
        return true;
    }

    // TEST ONLY (not an OpenFHE call): a non-zero seed makes the following KeyGen calls deterministic (seed, seed + 1, ...)
    // with the non-cryptographic generator the CPU oracle shares; 0 restores the default (OS entropy)
    static void SetKeySeed(uint64_t s) { seed_state() = s; }

  private:
    static uint64_t &seed_state() {
        static uint64_t s = 0;
        return s;
    }
    static uint64_t seed_next() {
        uint64_t &s = seed_state();
        return s == 0 ? 0 : s++;
    }
    CT binary(int (*fn)(fheb200_ctx *, fheb200_ct *, fheb200_ct *, fheb200_ct **), const CT &a, const CT &b) {
        need(a), need(b);
        fheb200_ct *o = nullptr;
        fheb200_check(fn(ctx, a->h, b->h, &o));
        return wrap(o);
    }
    CT plain(int (*fn)(fheb200_ctx *, fheb200_ct *, fheb200_pt *, fheb200_ct **), const CT &a, const Plaintext &p) {
        need(a);
        if (!p) throw std::runtime_error("null plaintext passed to the crypto context");
        fheb200_ct *o = nullptr;
        fheb200_check(fn(ctx, a->h, p->h, &o));
        return wrap(o);
    }
    CT constant(int (*fn)(fheb200_ctx *, fheb200_ct *, double, fheb200_ct **), const CT &a, double c) {
        need(a);
        fheb200_ct *o = nullptr;
        fheb200_check(fn(ctx, a->h, c, &o));
        return wrap(o);
    }
    void inplace(CT &a, const CT &r) { // move the result under a's shared handle
        fheb200_ct *n = r->h;
        r->h = nullptr;
        a->replace(n);
    }
};

inline Ciphertext<DCRTPoly> CiphertextImpl::Clone() const { return cc->CloneCt(*this); }

template <typename P> class CryptoContext {
    ContextPtr p_;

  public:
    CryptoContext() = default;
    CryptoContext(std::nullptr_t) {}
    CryptoContext(ContextPtr p) : p_(std::move(p)) {} // implicit: `CryptoContext<Element> cc = ct->GetCryptoContext();`
    CryptoContext(const ContextPtr &p, int) : p_(p) {}
    CryptoContextImpl *operator->() const { return p_.get(); }
    explicit operator bool() const { return (bool)p_; }
    const ContextPtr &ptr() const { return p_; }
};
// ct->GetCryptoContext() returns a ContextPtr; let it convert where the reference stores a CryptoContext<DCRTPoly>
template <typename P> inline CryptoContext<P> ToContext(const ContextPtr &p) { return CryptoContext<P>(p); }

// ciphertext operators of OpenFHE's ciphertext.h (used by src/mehp24/*.cpp: `c += c >> k`, `c1 * (1 - c2)`, `c1 - c2`)
template <typename P> inline Ciphertext<P> operator+(const Ciphertext<P> &a, const Ciphertext<P> &b) {
    return a->GetCryptoContext()->EvalAdd(a, b);
}
template <typename P> inline Ciphertext<P> operator-(const Ciphertext<P> &a, const Ciphertext<P> &b) {
    return a->GetCryptoContext()->EvalSub(a, b);
}
template <typename P> inline Ciphertext<P> operator*(const Ciphertext<P> &a, const Ciphertext<P> &b) {
    return a->GetCryptoContext()->EvalMult(a, b);
}
template <typename P> inline Ciphertext<P> operator-(const Ciphertext<P> &a) { return a->GetCryptoContext()->EvalNegate(a); }
template <typename P> inline const Ciphertext<P> &operator+=(Ciphertext<P> &a, const Ciphertext<P> &b) { return a = a + b; }
template <typename P> inline const Ciphertext<P> &operator-=(Ciphertext<P> &a, const Ciphertext<P> &b) { return a = a - b; }
template <typename P> inline const Ciphertext<P> &operator*=(Ciphertext<P> &a, const Ciphertext<P> &b) { return a = a * b; }

// lbcrypto::EvalChebyshevCoefficients(f, a, b, degree) (utils/generate_cheb_coeffs.cpp:17): host-only double math
inline std::vector<double> EvalChebyshevCoefficients(std::function<double(double)> f, double a, double b, uint32_t degree) {
    std::vector<double> out(degree + 1);
    auto tramp = [](double x, void *u) { return (*static_cast<std::function<double(double)> *>(u))(x); };
    fheb200_check(fheb200_chebyshev_coefficients(tramp, &f, a, b, (int)degree, out.data()));
    return out;
}
inline Ciphertext<DCRTPoly> CryptoContextImpl::EvalChebyshevFunction(std::function<double(double)> f, const CT &x, double a,
                                                                       double b, uint32_t degree) {
    return EvalChebyshevSeries(x, EvalChebyshevCoefficients(std::move(f), a, b, degree), a, b);
}

// ring dimension OpenFHE would pick for HEStd_128_classic (uniform ternary) when SetRingDim is not called:
// smallest power of two whose log(QP) bound (HomomorphicEncryption.org standard table, extended) covers the chain
inline usint StdRingDim128(double logQP) {
    static const struct { usint n; double bits; } tab[] = {{1024, 27},   {2048, 54},    {4096, 109},    {8192, 218},
                                                           {16384, 438}, {32768, 881}, {65536, 1747}, {131072, 3523}};
    for (auto &t : tab)
        if (logQP <= t.bits) return t.n;
    throw std::runtime_error("GenCryptoContext: modulus too large for 128-bit security at ring dimension <= 2^17");
}

// GenCryptoContext(CCParams<CryptoContextCKKSRNS>) (tests/DirectSortTest.cpp:24-41).  OpenFHE defaults for everything the
// reference does not set: FLEXIBLEAUTOEXT, HYBRID key switching with 3 digits, uniform ternary secret, first modulus 60.
inline CryptoContext<DCRTPoly> GenCryptoContext(const CCParams<CryptoContextCKKSRNS> &p, int device = -1) {
    usint ring = p.GetRingDim();
    const usint L = p.GetMultiplicativeDepth() + 2; // FLEXIBLEAUTOEXT keeps one extra level
    const usint dnum = p.GetNumLargeDigits();
    const double logQ = p.GetFirstModSize() + (double)(L - 1) * p.GetScalingModSize();
    const double logP = 60.0 * std::ceil(((double)((L + dnum - 1) / dnum) * p.GetScalingModSize() + 20) / 60.0);
    if (ring == 0) {
        if (p.GetSecurityLevel() == HEStd_NotSet) throw std::runtime_error("GenCryptoContext: SetRingDim is required with HEStd_NotSet");
        ring = StdRingDim128(logQ + logP);
        if (ring < 2 * p.GetBatchSize()) ring = 2 * p.GetBatchSize();
    } else if (p.GetSecurityLevel() == HEStd_128_classic && ring < StdRingDim128(logQ + logP)) {
        throw std::runtime_error("GenCryptoContext: ring dimension too small for HEStd_128_classic at this depth");
    }
    int logn = 0;
    while ((1u << logn) < ring) logn++;
    if ((1u << logn) != ring) throw std::runtime_error("GenCryptoContext: ring dimension must be a power of two");
    if (device < 0) {
        const char *e = getenv("FHEB200_DEVICE");
        device = e ? atoi(e) : 0;
    }
    fheb200_params q;
    q.log_ring_dim = logn, q.mult_depth = (int)p.GetMultiplicativeDepth(), q.scaling_mod_size = (int)p.GetScalingModSize();
    q.first_mod_size = (int)p.GetFirstModSize(), q.batch_size = (int)p.GetBatchSize(), q.num_large_digits = (int)dnum;
    q.device = device;
    auto impl = std::make_shared<CryptoContextImpl>();
    fheb200_check(fheb200_ctx_create(&q, &impl->ctx));
    impl->ringDim = ring, impl->depth = p.GetMultiplicativeDepth();
    impl->batch = p.GetBatchSize() ? p.GetBatchSize() : ring / 2;
    return CryptoContext<DCRTPoly>(impl);
}

// ---- Serial::SerializeToFile / DeserializeFromFile for the three object kinds the reference moves through files
// (src/sort.h:34-45, 69-73, 97-101): crypto context (parameters only), public key, ciphertext.  Container: 8-byte magic,
// fixed header of 64-bit fields, then raw little-endian limbs [poly][limb][n] as the engine exports them.  OpenFHE's cereal
// BINARY archives are a different format; reading them needs OpenFHE's class layout and is unverifiable on this machine.
namespace Serial {
inline void put(std::ostream &os, uint64_t v) { CryptoContextImpl::ser_u64(os, v); }
inline uint64_t get(std::istream &is) { return CryptoContextImpl::deser_u64(is); }
inline void put_double(std::ostream &os, double d) {
    uint64_t v;
    memcpy(&v, &d, 8);
    put(os, v);
}
inline double get_double(std::istream &is) {
    uint64_t v = get(is);
    double d;
    memcpy(&d, &v, 8);
    return d;
}
inline bool Serialize(const CryptoContext<DCRTPoly> &cc, std::ostream &os) {
    fheb200_params p;
    fheb200_check(fheb200_ctx_params(cc->ctx, &p));
    put(os, CryptoContextImpl::SER_MAGIC_CC);
    for (int v : {p.log_ring_dim, p.mult_depth, p.scaling_mod_size, p.first_mod_size, p.batch_size, p.num_large_digits}) put(os, (uint64_t)v);
    put(os, cc->features);
    return (bool)os;
}
inline bool Deserialize(CryptoContext<DCRTPoly> &cc, std::istream &is) {
    if (get(is) != CryptoContextImpl::SER_MAGIC_CC) return false;
    fheb200_params q;
    q.log_ring_dim = (int)get(is), q.mult_depth = (int)get(is), q.scaling_mod_size = (int)get(is), q.first_mod_size = (int)get(is);
    q.batch_size = (int)get(is), q.num_large_digits = (int)get(is);
    const uint64_t features = get(is);
    const char *e = getenv("FHEB200_DEVICE");
    q.device = e ? atoi(e) : 0;
    auto impl = std::make_shared<CryptoContextImpl>();
    fheb200_check(fheb200_ctx_create(&q, &impl->ctx));
    impl->ringDim = 1u << q.log_ring_dim, impl->depth = (usint)q.mult_depth;
    impl->batch = q.batch_size ? (usint)q.batch_size : impl->ringDim / 2;
    impl->features = (usint)features;
    cc = CryptoContext<DCRTPoly>(impl);
    return true;
}
// the public key file carries its context's parameters so that it can be checked against the context it is loaded into
inline bool Serialize(const PublicKey<DCRTPoly> &pk, std::ostream &os, const CryptoContext<DCRTPoly> &cc) {
    int n = 0, L = 0;
    fheb200_check(fheb200_ctx_info(cc->ctx, &n, &L, nullptr, nullptr, nullptr));
    const size_t w = (size_t)L * n;
    std::vector<uint64_t> b(w), a(w);
    fheb200_check(fheb200_keys_export(cc->ctx, pk.k->h, 1, b.data(), w));
    fheb200_check(fheb200_keys_export(cc->ctx, pk.k->h, 2, a.data(), w));
    put(os, CryptoContextImpl::SER_MAGIC_PK), put(os, w);
    os.write(reinterpret_cast<const char *>(b.data()), (std::streamsize)(w * 8));
    os.write(reinterpret_cast<const char *>(a.data()), (std::streamsize)(w * 8));
    return (bool)os;
}
inline bool Deserialize(PublicKey<DCRTPoly> &pk, std::istream &is, const CryptoContext<DCRTPoly> &cc) {
    if (get(is) != CryptoContextImpl::SER_MAGIC_PK) return false;
    const uint64_t w = get(is);
    std::vector<uint64_t> b(w), a(w);
    is.read(reinterpret_cast<char *>(b.data()), (std::streamsize)(w * 8));
    is.read(reinterpret_cast<char *>(a.data()), (std::streamsize)(w * 8));
    if (!is) return false;
    pk.k = std::make_shared<KeySetImpl>();
    fheb200_check(fheb200_keys_import_public(cc->ctx, b.data(), a.data(), w, &pk.k->h));
    return true;
}
inline bool Serialize(const Ciphertext<DCRTPoly> &ct, std::ostream &os) {
    int level = 0, nsd = 0, slots = 0, limbs = 0;
    double scale = 0;
    fheb200_check(fheb200_ct_info(ct->h, &level, &nsd, &slots, &limbs, &scale));
    int n = 0;
    fheb200_check(fheb200_ctx_info(ct->cc->ctx, &n, nullptr, nullptr, nullptr, nullptr));
    const size_t w = (size_t)limbs * n;
    std::vector<uint64_t> c0(w), c1(w);
    fheb200_check(fheb200_ct_export(ct->cc->ctx, ct->h, c0.data(), c1.data(), w));
    put(os, CryptoContextImpl::SER_MAGIC_CT), put(os, (uint64_t)limbs), put(os, (uint64_t)nsd), put(os, (uint64_t)slots), put(os, w);
    put_double(os, scale);
    os.write(reinterpret_cast<const char *>(c0.data()), (std::streamsize)(w * 8));
    os.write(reinterpret_cast<const char *>(c1.data()), (std::streamsize)(w * 8));
    return (bool)os;
}
inline bool Deserialize(Ciphertext<DCRTPoly> &ct, std::istream &is, const CryptoContext<DCRTPoly> &cc) {
    if (get(is) != CryptoContextImpl::SER_MAGIC_CT) return false;
    const int limbs = (int)get(is), nsd = (int)get(is), slots = (int)get(is);
    const uint64_t w = get(is);
    const double scale = get_double(is);
    std::vector<uint64_t> c0(w), c1(w);
    is.read(reinterpret_cast<char *>(c0.data()), (std::streamsize)(w * 8));
    is.read(reinterpret_cast<char *>(c1.data()), (std::streamsize)(w * 8));
    if (!is) return false;
    fheb200_ct *h = nullptr;
    fheb200_check(fheb200_ct_import(cc->ctx, c0.data(), c1.data(), limbs, nsd, scale, slots, &h));
    ct = Ciphertext<DCRTPoly>(std::make_shared<CiphertextImpl>(h, cc.ptr()));
    return true;
}
// OpenFHE deserialises keys and ciphertexts into the context registered for their parameters; here that is the context
// most recently deserialised or generated in this process (src/sort.h loads the context first)
inline CryptoContext<DCRTPoly> &CurrentContext() {
    static CryptoContext<DCRTPoly> cc;
    return cc;
}
inline bool SerializeToFile(const std::string &path, const CryptoContext<DCRTPoly> &cc, const SerType::SERBINARY &) {
    std::ofstream f(path, std::ios::binary);
    CurrentContext() = cc;
    return f.is_open() && Serialize(cc, f);
}
inline bool SerializeToFile(const std::string &path, const PublicKey<DCRTPoly> &pk, const SerType::SERBINARY &) {
    std::ofstream f(path, std::ios::binary);
    return f.is_open() && CurrentContext() && Serialize(pk, f, CurrentContext());
}
inline bool SerializeToFile(const std::string &path, const Ciphertext<DCRTPoly> &ct, const SerType::SERBINARY &) {
    std::ofstream f(path, std::ios::binary);
    return f.is_open() && Serialize(ct, f);
}
inline bool DeserializeFromFile(const std::string &path, CryptoContext<DCRTPoly> &cc, const SerType::SERBINARY &) {
    std::ifstream f(path, std::ios::binary);
    if (!f.is_open() || !Deserialize(cc, f)) return false;
    CurrentContext() = cc;
    return true;
}
inline bool DeserializeFromFile(const std::string &path, PublicKey<DCRTPoly> &pk, const SerType::SERBINARY &) {
    std::ifstream f(path, std::ios::binary);
    return f.is_open() && CurrentContext() && Deserialize(pk, f, CurrentContext());
}
inline bool DeserializeFromFile(const std::string &path, Ciphertext<DCRTPoly> &ct, const SerType::SERBINARY &) {
    std::ifstream f(path, std::ios::binary);
    return f.is_open() && CurrentContext() && Deserialize(ct, f, CurrentContext());
}
} // namespace Serial

} // namespace lbcrypto
