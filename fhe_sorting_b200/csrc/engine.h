// engine.h — B200 RNS-CKKS evaluation engine: context, device-resident keys, ciphertext/plaintext objects
// and the scheme operations the reference reaches through lbcrypto::CryptoContext<DCRTPoly>.
// Level / scale bookkeeping follows OpenFHE's FLEXIBLEAUTOEXT mode: lazy rescale before the next
// multiplication, exact real scaling factors per level, GetLevel() = number of dropped limbs.
#pragma once
#include <atomic>
#include <complex>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include <deque>

#include "bootstrap_plan.h"
#include "kernels.h"
#include "params.h"

namespace fheb200 {

class Context;

struct DevMem { // stream-ordered device allocation owned by a context
    u64 *p = nullptr;
    size_t words = 0;
    Context *ctx = nullptr;
    std::shared_ptr<Context> keep; // a buffer keeps its context (stream, memory pool) alive
    cudaStream_t stream = nullptr; // the stream it was allocated on (and is freed on)
    unsigned epoch = 0;            // Context::epoch at allocation (see Context::fork)
    ~DevMem();
};
typedef std::shared_ptr<DevMem> Mem;

struct Ciphertext {
    Mem mem;          // [npoly][ell][n], evaluation domain
    int npoly = 2;
    int ell = 0;      // active Q limbs = L - level
    int level = 0;    // dropped limbs
    int nsd = 1;      // noise scale degree (1 or 2; 3 transiently never kept)
    double scale = 0; // exact real scaling factor
    int slots = 0;
    // Rescaled copy, filled the first time a multiplication needs this (immutable) ciphertext at degree 1: the sort
    // drivers multiply one baby-step ciphertext by dozens of masks (src/sort_algo.h:339-345), each of which would
    // otherwise repeat the same rescale.  Never copied along with the object.
    mutable std::shared_ptr<Ciphertext> reduced_cache;
    u64 *poly(int i, int n) const { return mem->p + (size_t)i * ell * n; }
    Ciphertext() = default;
    Ciphertext(const Ciphertext &o)
        : mem(o.mem), npoly(o.npoly), ell(o.ell), level(o.level), nsd(o.nsd), scale(o.scale), slots(o.slots) {}
    Ciphertext &operator=(const Ciphertext &o) {
        mem = o.mem, npoly = o.npoly, ell = o.ell, level = o.level, nsd = o.nsd, scale = o.scale, slots = o.slots;
        reduced_cache.reset();
        return *this;
    }
};
typedef std::shared_ptr<Ciphertext> Ct;

struct Plaintext {
    std::vector<std::complex<double>> vals; // slot values (kept so the plaintext can be re-encoded at any level)
    std::vector<double> re;                 // ... or just their real parts (MakeCKKSPackedPlaintext of a real vector)
    bool real_only = false;
    std::shared_ptr<Plaintext> base;        // ... or "base rotated left by rot slots": encoded as a Galois permutation
    int rot = 0;                            //     of base's encoding instead of a transform of its own
    int slots = 0;
    int level = 0, nsd = 1;
    double scale = 0;
    int ell = 0;
    Mem mem;           // [ell][n] evaluation domain (may be null for decode-only plaintexts)
    size_t length = 0; // SetLength
    double log_precision = 0;
    std::vector<std::tuple<int, double, Mem>> alt; // re-encodings at other (limbs, scale), filled lazily
};
typedef std::shared_ptr<Plaintext> Pt;

struct KeySet {
    Mem s;            // secret key over L+K limbs (null on evaluation-only contexts)
    Mem pk;           // [2][L][n]: b then a
};
typedef std::shared_ptr<KeySet> Keys;

struct BootState { // EvalBootstrapSetup result for one slot count: the plan and its diagonals as (lazily encoded) plaintexts
    BootPlan plan;
    std::vector<std::vector<Pt>> cts_pt, stc_pt;
    // encodings of the diagonals over the extended basis (Q limbs of the level they are used at, then all P), filled on
    // first use: [stage][diagonal] -> (limbs, scale, memory)
    std::vector<std::vector<std::tuple<int, double, Mem>>> cts_qp, stc_qp;
};

struct Precomp { // EvalFastRotationPrecompute result: ModUp'ed digits of c1
    Mem digits;
    int ell = 0;
};

class Context : public std::enable_shared_from_this<Context> {
  public:
    static std::shared_ptr<Context> create(int logn, int depth, int scale_bits, int first_bits, int dnum, int batch,
                                           int device);
    ~Context();

    std::shared_ptr<Params> P;
    DeviceTables T;
    int device = 0;
    int batch = 0; // default slot count (CCParams::SetBatchSize)
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    std::recursive_mutex mu;
    // generator state (modarith.cuh "randomness"): secret material and the public uniform polynomials draw from
    // independent ChaCha20 keys (OS entropy); keygen(seed != 0) switches both to the TEST-ONLY deterministic generator and
    // is refused unless allow_deterministic was set (fheb200_ctx_allow_deterministic_keys)
    RngState rng_secret, rng_public;
    bool allow_deterministic = false;
    bool rng_pinned = false; // key material was installed by set_rng_keys (replicated keys): KeyGen must not redraw it
    std::atomic<u64> nonce{1};
    void reseed_from_os();                       // fresh 2 x 256 bits of OS entropy (throws when none is available)
    void set_rng_keys(const unsigned char *k64); // replicate key material on another rank (64 bytes from get_rng_keys)
    void get_rng_keys(unsigned char *k64) const;

    // ---- keys
    Keys keygen(u64 seed);
    void gen_mult_key(const Keys &k);
    void gen_rot_keys(const Keys &k, const std::vector<int> &indices);
    bool has_rot_key(uint32_t galois) const { return evk_rot.count(galois) != 0; }
    size_t key_bytes() const;

    // ---- encode / encrypt / decrypt
    Pt make_plaintext(const std::vector<std::complex<double>> &vals, int nsd, int level, int slots);
    Pt make_plaintext_real(const double *vals, size_t count, int nsd, int level, int slots);
    Pt rotate_plaintext(const Pt &base, int index); // slot i of the result = slot i + index of base
    Mem encode_at(const Plaintext &pt, int ell, double scale); // (re-)encode at a given limb count and scale
    Mem encoding_for(const Pt &pt, int ell, double scale);     // cached variant
    Ct encrypt(const Keys &k, const Pt &pt);
    Pt decrypt(const Keys &k, const Ct &ct);

    // ---- arithmetic (all return fresh objects)
    Ct clone(const Ct &a);
    Ct add(const Ct &a, const Ct &b);
    Ct sub(const Ct &a, const Ct &b);
    Ct add_pt(const Ct &a, const Pt &p, bool negate_pt = false);
    Ct add_const(const Ct &a, double c);
    Ct negate(const Ct &a);
    Ct mult(const Ct &a, const Ct &b);       // EvalMult / EvalMultAndRelinearize / EvalSquare (a.get()==b.get())
    Ct mult_lazy(const Ct &a, const Ct &b);  // tensor product only: 3 polynomials, relinearise() later
    Ct relinearize(const Ct &a);             // 3 -> 2 polynomials (no-op on 2)
    Ct relinearize_reduced(const Ct &a);     // == reduced(relinearize(a)) up to rounding: one ModDown by P * q_last
    Ct add_lazy(const Ct &a, const Ct &b);   // EvalAdd where either side may still carry its s^2 polynomial
    Ct mult_pt(const Ct &a, const Pt &p);
    Ct mult_pt_add(const Ct &acc, const Ct &a, const Pt &p); // acc + a * p, one pass
    Ct mult_const(const Ct &a, double c);
    Ct rotate(const Ct &a, int index);
    std::shared_ptr<Precomp> fast_rotation_precompute(const Ct &a);
    Ct fast_rotation(const Ct &a, int index, const std::shared_ptr<Precomp> &pre);
    Ct add_many(const std::vector<Ct> &v);
    Ct rescale(const Ct &a);                 // explicit ModReduce (one level)
    Ct level_reduce(const Ct &a, int levels);

    // ---- bootstrapping (bootstrap.cu): EvalBootstrapSetup / EvalBootstrapKeyGen / EvalBootstrap for sparsely packed
    // ciphertexts (src/k-way/EvalUtils.cpp:76, tests/k-way/KWaySort2Test.cpp:42-44)
    void bootstrap_setup(int budget_enc, int budget_dec, int slots);
    void bootstrap_keygen(const Keys &k, int slots);
    Ct bootstrap(const Ct &a, int stop_after = 0);      // stop_after 1..4: the intermediate ciphertext of that step (tests)
    int bootstrap_depth(int slots) const;               // levels one EvalBootstrap consumes (after the raise to level 1)
    Ct conjugate(const Ct &a);                          // slot-wise complex conjugation (Galois element 2n - 1)
    void gen_galois_keys(const Keys &k, const std::vector<uint32_t> &galois);
    Ct rotate_galois(const Ct &a, uint32_t g);
    Ct mod_raise(const Ct &a, int ell_out, double scale_label); // a: one limb; result on limbs 0..ell_out-1
    Ct linear_stage(const Ct &x, const BootStage &st, const std::vector<Pt> &diag,
                    std::vector<std::tuple<int, double, Mem>> *qp_cache = nullptr);
    Mem encode_qp(const Plaintext &pt, int ell, double scale);   // [ell + K][n]: Q limbs 0..ell-1, then the P limbs
    void moddown(int ell, u64 *acc, const u64 *add0, const u64 *add1, u64 *out0, u64 *out1, uint32_t galois);
    void moddown_plan(const DevicePlan &D, u64 *acc, const u64 *add0, const u64 *add1, u64 *out0, u64 *out1, uint32_t galois);
    std::map<int, std::shared_ptr<BootState>> boot;

    // ---- polynomial evaluation (poly_eval.cpp)
    Ct eval_chebyshev(const Ct &x, const std::vector<double> &coeffs, double a, double b);
    Ct eval_chebyshev_linear(const Ct &x, const std::vector<double> &coeffs, double a, double b);
    Ct eval_poly_linear(const Ct &x, const std::vector<double> &coeffs);

    // ---- raw access for parity tests and benchmarks
    const DevicePlan &plan(int ell);
    const DevicePlan &plan_merged(int ell); // ModDown by P * q_{ell-1}
    ConvTable make_conv_table(const std::vector<u64> &C, int cstride, int nsrc, std::vector<std::pair<int, int>> tg, const std::vector<int> &tpos);
    const uint32_t *desc(const std::vector<uint32_t> &d);
    Mem alloc(size_t words);
    void sync();
    // ---- concurrent streams for independent sub-computations (the B batches of a DirectSort phase): fork(k) lets k
    // side streams start from the main stream's current point, use_stream(i) directs every following call to side
    // stream i, join() makes the main stream wait for all of them and goes back to it.  Rules the caller keeps:
    // objects made inside the region are used on the stream that made them until join(); objects made before fork()
    // are read-only inside.  Device-side caches hanging off shared objects are only trusted if usable().
    void fork(int k);
    void use_stream(int i);
    void join();
    // ---- multi-GPU (comm.cu): the engine's own NCCL communicator
    void comm_init(const unsigned char *id128, int rank, int world);
    void comm_destroy();
    void all_reduce_sum(const Ct &a); // in place, on `stream`: 64-bit integer sum over the ranks, then every word mod q_limb
    void *comm_ = nullptr;            // ncclComm_t
    int comm_rank_ = 0, comm_world_ = 1;

    bool forked() const { return forked_; }
    bool park_if_shared(u64 *p, unsigned buf_epoch); // DevMem release inside a forked region (see DevMem::~DevMem)
    bool usable(const Mem &m) const { return !forked_ || m->stream == stream || m->epoch < fork_epoch_; }
    unsigned epoch = 1;
    // ---- asynchronous host <-> device transfers on a copy stream of their own (pinned host buffers), so that the copies
    // of one work item overlap the kernels of its neighbours; buffers a pending copy reads are kept alive until it is done
    void copy_in(void *dst, const void *src, size_t bytes, const Mem &owner);   // H2D; wait_copy_in(owner) before first use
    void wait_copy_in(const Mem &owner);                                         // current stream waits for owner's H2D copies
    void copy_out(void *dst, const void *src, size_t bytes, const Mem &owner);  // D2H after everything queued so far
    void drain_copies(bool wait);
    u64 *evk_mult = nullptr;
    std::map<uint32_t, u64 *> evk_rot;
    size_t evk_words() const { return (size_t)P->beta(P->L) * 2 * (P->L + P->K) * P->n; }

    // key switching core on raw device polys (c: ell limbs eval). Results written to out0/out1 (ell limbs each):
    //   out0 = sigma_g(ks0 + add0), out1 = sigma_g(ks1 + add1)   (add may be null, g = 1 means no automorphism)
    void modup(int ell, const u64 *c, u64 *digits);
    void keyswitch_digits(int ell, const u64 *digits, const u64 *evk, const u64 *add0, const u64 *add1, u64 *out0,
                          u64 *out1, uint32_t galois);
    void rescale_poly(int ell, int npoly, const u64 *in, u64 *out);

    // scheme helpers
    Ct mod_reduce(const Ct &a);
    Ct reduced(const Ct &a);                       // memoised mod_reduce of a degree-2 ciphertext
    Ct mult_core_const(const Ct &a, double c);     // multiply by round(c * sf[level]); nsd += 1
    // out[r] = sum_t round(w[r][t] * sf[level]) * in[t]  (all inputs at the same level and degree 1); null where all w vanish
    std::vector<Ct> lincomb(const std::vector<Ct> &in, const std::vector<std::vector<double>> &w);
    void adjust_pair(Ct &a, Ct &b);                // AdjustLevelsAndDepthInPlace
    void adjust_pair_to_one(Ct &a, Ct &b);         // AdjustLevelsAndDepthToOneInPlace
    LimbScalars scalars_for(long double value, int ell, bool shoup) const;

    // Host threads inside a forked region: a thread that called bind_thread(i) issues everything on side stream i (i < 0: the main
    // stream) whatever the other threads do; `stream` is re-pointed under the context mutex at the top of every operation
    // (CtxLock), so the many uses of `stream` below stay as they are.  Unbound threads keep the fork() / use_stream() selection.
    void bind_thread(int i);
    void unbind_thread();
    void bind_thread_stream(); // called with `mu` held

  private:
    friend struct DevMem;
    friend struct CtxLock;
    cudaStream_t main_stream_ = nullptr, fork_current_ = nullptr; // fork_current_: what fork() / use_stream() selected (unbound threads)
    std::vector<cudaStream_t> side_;
    std::vector<cudaEvent_t> side_done_;
    cudaEvent_t fork_ev_ = nullptr, free_ev_ = nullptr;
    std::mutex park_mu_;
    std::vector<u64 *> parked_; // pre-fork buffers dropped inside the forked region, freed by join()
    bool forked_ = false;
    int fork_k_ = 0;
    cudaStream_t copy_stream_ = nullptr, out_stream_ = nullptr; // H2D and D2H copies (one DMA engine per direction)
    struct PendingCopy { cudaEvent_t done; Mem keep; };
    std::deque<PendingCopy> pending_;
    std::map<const DevMem *, cudaEvent_t> inbound_; // last H2D copy into a buffer that nobody has waited for yet
    std::vector<cudaEvent_t> free_events_;
    cudaEvent_t take_event();
    unsigned fork_epoch_ = 0;
    // pinned staging ring for host -> device uploads of slot values: the host never waits for the stream unless it is a
    // whole ring ahead of it (a pageable cudaMemcpyAsync would stall the launching thread at every plaintext)
    struct Stage { void *host = nullptr; cudaEvent_t done = nullptr; };
    std::vector<Stage> stage_;
    size_t stage_bytes_ = 0, stage_next_ = 0;
    void upload(void *dst, const void *src, size_t bytes);
    Pt new_plaintext(size_t count, int nsd, int level, int slots);
    std::map<int, std::pair<double2 *, int *>> fft_tabs_; // per slot count: (ksi, rot) tables of the special FFT in HBM
    std::pair<double2 *, int *> fft_device_tables(int slots);
    std::map<int, DevicePlan> plans_, plans_m_;
    std::map<std::vector<uint32_t>, uint32_t *> descs_;
    std::vector<void *> owned_;
    void *owned_alloc(size_t bytes, const void *init);
    u64 *gen_switch_key(u64 key_id, const u64 *s_old, const u64 *s_new);
};

void fft_special(std::vector<std::complex<double>> &v);
void fft_special_inv(std::vector<std::complex<double>> &v);
std::vector<double> chebyshev_coefficients(const std::function<double(double)> &f, double a, double b, int degree);

#define FHE_CUDA(call)                                                                                       \
    do {                                                                                                     \
        cudaError_t e__ = (call);                                                                            \
        if (e__ != cudaSuccess)                                                                              \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e__) + " at " #call);  \
    } while (0)

// the context mutex + this thread's stream binding, for the duration of one operation
struct CtxLock {
    std::lock_guard<std::recursive_mutex> g;
    explicit CtxLock(Context &c) : g(c.mu) { c.bind_thread_stream(); }
};

} // namespace fheb200
