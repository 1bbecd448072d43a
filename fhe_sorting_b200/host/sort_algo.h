// sort_algo.h — encrypted sorting drivers: SortBase<N> and DirectSort<N>, the optimized rank sort of eprint 2025/1170.
// Interface and algorithm of the reference's src/sort_algo.h:36-774 (SortBase, DirectSort::getSizeParameters,
// constructRank, rotationIndexCheckN, sort), restated for a single-stream GPU engine:
//   * the reference spreads independent ciphertext work over OpenMP threads; here every call is an asynchronous launch
//     sequence on one CUDA stream that already fills the 148 SMs at ring 2^17, so the drivers are plain loops;
//   * baby-step rotations of one ciphertext share one hoisted ModUp (EvalFastRotationPrecompute), which is bit-identical
//     to separate EvalRotate calls;
//   * the rotation-key list is derived from the schedule instead of being tabulated (same sets as the reference's
//     tables up to indices that are multiples of the slot count, i.e. identity rotations).
#pragma once
#include <cstdlib>
#include <algorithm>
#include <cmath>
#include <iostream>
#include <memory>
#include <exception>
#include <set>
#include <thread>
#include <vector>

#include "comparison.h"
#include "encryption.h"
#include "generated_coeffs.h"
#include "openfhe.h"
#include "rotation.h"
#include "sign.h"

using namespace lbcrypto;

template <int N> class SortBase {
  protected:
    std::shared_ptr<Encryption> m_enc;
    const Ciphertext<DCRTPoly> m_zeroCache;
    virtual Ciphertext<DCRTPoly> createZeroCache() { return m_enc->encryptInput(std::vector<double>(N, 0.0)); }

  public:
    SortBase(std::shared_ptr<Encryption> enc) : m_enc(enc), m_zeroCache(createZeroCache()) {}
    virtual ~SortBase() = default;
    virtual Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &input_array, SignFunc SignFunc, SignConfig &Cfg) = 0;
    virtual const Ciphertext<DCRTPoly> &getZero() const { return m_zeroCache; }
    constexpr size_t getArraySize() const { return N; }
};

// Shape of the SIMD packing for array length N on a ring with max_batch slots (reference: sort_algo.h:378-416,665-703)
struct DirectSortShape {
    int P;       // copies of the array per ciphertext  = min(N, max_batch / N)
    int B;       // batches                             = N / P
    int S;       // slots in use                        = N * P
    int np_rank; // baby steps of the shifted-copies construction
    int np_idx;  // baby steps of the blind rotation
    static int baby(int N, bool index_phase) {
        if (index_phase && N >= 512) return N <= 1024 ? 8 : 4;
        if (N <= 8) return 2;
        if (N <= 32) return 4;
        if (N <= 128) return 8;
        if (N <= 512) return 16;
        return 32;
    }
    DirectSortShape(int N, int max_batch) {
        P = std::min(N, max_batch / N), B = N / P, S = N * P;
        np_rank = std::min(baby(N, false), P);
        np_idx = std::min(baby(N, true), P);
    }
};

// which of the B batches process `rank` of `world` evaluates (all of them when B is not a multiple of world: replicated)
inline std::vector<int> DirectSortBatchesOf(int B, int rank, int world) {
    std::vector<int> mine;
    for (int b = 0; b < B; b++)
        if (world <= 1 || B % world != 0 || b % world == rank) mine.push_back(b);
    return mine;
}

// The B batches of a phase are independent chains of small kernels (at N = 1024 the Paterson-Stockmeyer evaluations run on
// 4..17 limbs): one CUDA stream cannot fill 148 SMs with them.  The first batch runs on the main stream (it fills the
// caches that hang off the shared inputs: memoised rescales, plans), the others round-robin over k side streams; the
// destructor joins.  The reference reaches the same concurrency with OpenMP threads (src/sort_algo.h:475, 713).
class BatchStreams {
    CryptoContext<DCRTPoly> cc;
    int k, seen = 0;
    bool forked = false;

  public:
    BatchStreams(CryptoContext<DCRTPoly> c, int streams) : cc(c), k(streams) {}
    BatchStreams(const BatchStreams &) = delete;
    void next() { // call at the top of every batch
        if (k > 1 && seen >= 1) {
            if (!forked) cc->ForkStreams(k), forked = true;
            cc->UseStream((seen - 1) % k);
        }
        seen++;
    }
    ~BatchStreams() {
        if (forked) cc->JoinStreams();
    }
};
// The same with one host thread per stream (opt-in, FHESORT_THREADS=1): a launching thread cannot run more than a launch queue
// (~1000 kernels) ahead of the GPU, and one batch is several thousand kernels, so from one thread the streams overlap only at the
// seams between batches.  Measured on B200 (DESIGN.md, round-2 results): 1456 -> 1448 ms at N = 1024 - the batches do not have
// idle GPU time to give each other, so the single launching thread stays the default.  body(i) evaluates the i-th of `count` batches into its own result
// slot; the first batch runs alone on the calling thread (it fills the caches that hang off the shared inputs), the rest are
// dealt round-robin to k threads, each bound to its side stream (fheb200_ctx_bind_thread).
template <class Body> inline void forEachBatch(CryptoContext<DCRTPoly> cc, int streams, int count, Body body) {
    static const bool threads = [] { const char *e = getenv("FHESORT_THREADS"); return e && *e == '1'; }();
    if (count <= 0) return;
    if (!threads || streams <= 1 || count <= 2) {
        BatchStreams bs(cc, streams);
        for (int i = 0; i < count; i++) bs.next(), body(i);
        return;
    }
    body(0);
    const int k = std::min(streams, count - 1);
    cc->ForkStreams(k);
    std::vector<std::exception_ptr> err(k);
    std::vector<std::thread> pool;
    for (int t = 0; t < k; t++)
        pool.emplace_back([&, t] {
            try {
                cc->BindThreadStream(t);
                for (int i = 1 + t; i < count; i += k) body(i);
            } catch (...) {
                err[t] = std::current_exception();
            }
            try {
                cc->UnbindThreadStream();
            } catch (...) {
            }
        });
    for (auto &th : pool) th.join();
    cc->JoinStreams();
    for (auto &e : err)
        if (e) std::rethrow_exception(e);
}
inline int DirectSortDefaultStreams() {
    const char *e = getenv("FHESORT_STREAMS");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? (v > 16 ? 16 : v) : 4;
}

inline int ilog2(int v) {
    int l = 0;
    while ((1 << (l + 1)) <= v) l++;
    return l;
}

template <int N> class DirectSort : public SortBase<N> {
    CryptoContext<DCRTPoly> m_cc;
    PublicKey<DCRTPoly> m_PublicKey;
    Comparison comp;
    RotationComposer<N> rot;
    int max_batch; // ring dimension / 2
    std::set<int> m_keys;

    typedef Ciphertext<DCRTPoly> CT;
    int m_rank = 0, m_world = 1; // multi-GPU: this process evaluates the batches b with b % m_world == m_rank
    int m_streams = DirectSortDefaultStreams(); // concurrent CUDA streams over this process's batches

    bool mine(int b, int B) const {
        for (int x : DirectSortBatchesOf(B, m_rank, m_world))
            if (x == b) return true;
        return false;
    }
    bool sharded(int B) const { return m_world > 1 && B % m_world == 0; }

    // rotations 0..count-1 of `in`; one hoisted ModUp when every key exists
    std::vector<CT> babySteps(const CT &in, int count, int num_slots) {
        std::vector<CT> out(count);
        bool all = true;
        for (int i = 1; i < count; i++) all = all && m_keys.count(i);
        std::shared_ptr<FastRotationPrecomp> pre;
        if (all && count > 2) pre = m_cc->EvalFastRotationPrecompute(in);
        for (int i = 0; i < count; i++) {
            CT t = i == 0 ? in->Clone() : pre ? m_cc->EvalFastRotation(in, i, m_cc->GetCyclotomicOrder(), pre) : rot.rotate(in, i);
            t->SetSlots(num_slots);
            out[i] = t;
        }
        return out;
    }
    void foldCopies(CT &acc, int num_slots, int copies) { // sum the P copies: log2 P rotate-adds (serial by nature)
        for (int span = num_slots / 2; span >= num_slots / copies && span >= 1; span /= 2)
            m_cc->EvalAddInPlace(acc, rot.rotate(acc, span));
    }

  public:
    std::shared_ptr<Encryption> m_enc;

    DirectSort(CryptoContext<DCRTPoly> cc, PublicKey<DCRTPoly> publicKey, std::vector<int> rotIndices, std::shared_ptr<Encryption> enc)
        : SortBase<N>(enc), m_cc(cc), m_PublicKey(publicKey), comp(enc), rot(cc, enc, rotIndices),
          max_batch(cc->GetRingDimension() / 2), m_keys(rotIndices.begin(), rotIndices.end()), m_enc(enc) {}

    const std::set<int> &getRotationCalls() const { return rot.getRotationCalls(); }

    // Shard the B independent batches of both phases over `world` processes (one per GPU, keys replicated); partial rank
    // and output ciphertexts are summed with CryptoContextImpl::AllReduceSum.  Lengths with fewer batches than processes
    // (N <= 256 on ring 2^17: a single comparison ciphertext) run replicated.
    void setShard(int rank, int world) { m_rank = rank, m_world = world; }
    void setStreams(int k) { m_streams = k < 1 ? 1 : k; }

    // multiplicative depth per array length (reference table sort_algo.h:94-198), scaling modulus 40 bits, batch N, and
    // the rotation indices the schedule below uses on a ring of dimension `ringDim`
    static int multDepthFor() {
        switch (N) {
        case 4: return 23;
        case 8: return 24;
        case 16: return 25;
        case 32: return 28;
        case 64: return 29;
        case 128: return 30;
        case 256: return 34;
        case 512: return 35;
        case 1024: return 39;
        case 2048: return 52;
        default: return 39;
        }
    }
    static std::vector<int> rotationIndices(int ringDim = 1 << 17) {
        const DirectSortShape sh(N, ringDim / 2);
        std::set<int> r;
        for (int i = 1; i < sh.np_rank; i++) r.insert(i);
        for (int i = 1; i < sh.np_idx; i++) r.insert(i);
        for (int b = 0; b < sh.B; b++) {
            for (int j = 0; j < sh.P / sh.np_rank; j++) r.insert(b * sh.P + j * sh.np_rank);
            for (int i = 0; i < sh.P / sh.np_idx; i++) r.insert(b * sh.P + i * sh.np_idx);
        }
        for (int span = sh.S / 2; span >= N; span /= 2) r.insert(span);
        r.erase(0);
        return std::vector<int>(r.begin(), r.end());
    }
    static void getSizeParameters(CCParams<CryptoContextCKKSRNS> &parameters, std::vector<int> &rotations) {
        parameters.SetBatchSize(N);
        parameters.SetScalingModSize(40);
        parameters.SetMultiplicativeDepth(multDepthFor());
        rotations = rotationIndices(parameters.GetRingDim() ? parameters.GetRingDim() : 1 << 17);
    }

    // ---- plain vectors of the schedule
    std::vector<double> generateMaskVector(int num_slots, int k) { // ones on copy k
        std::vector<double> m(num_slots, 0.0);
        std::fill(m.begin() + (size_t)k * N, m.begin() + (size_t)(k + 1) * N, 1.0);
        return m;
    }
    std::vector<double> generateMaskVectorN(int num_slots, int k) { return generateMaskVector(num_slots, k); }
    std::vector<double> generateIndexVector() {
        std::vector<double> v(N);
        for (int i = 0; i < N; i++) v[i] = i;
        return v;
    }
    std::vector<double> generateCheckingVectorN(int num_slots, int k) { // copy c holds the constant (k + c) mod N
        std::vector<double> v(num_slots);
        for (int i = 0; i < num_slots; i++) v[i] = (k + i / N) % N;
        return v;
    }
    std::vector<double> vectorRotate(const std::vector<double> &vec, int rotateIndex) { // > 0 left, < 0 right
        std::vector<double> r(vec);
        if (r.empty()) return r;
        const int n = (int)r.size();
        int s = ((rotateIndex % n) + n) % n;
        std::rotate(r.begin(), r.begin() + s, r.end());
        return r;
    }

    // Shifted copies for batch `is`: copy c of the result holds the array rotated left by is*P + c.
    // Baby-step / giant-step: group j gathers, from the np pre-rotated inputs, the copies np*j .. np*j+np-1 (masks are
    // pre-rotated right by the giant step so that one rotation by is*P + j*np puts everything in place).
    // shard_groups (multi-GPU, lengths with a single batch): this rank evaluates the giant-step groups j = rank mod world
    // only and the partial sums are combined by AllReduceSum - the groups are independent (src/sort_algo.h:349-362)
    CT vecRotsOpt(const std::vector<CT> &preRotatedArrays, int num_partition, int num_slots, int np, int is, bool shard_groups = false) {
        std::vector<CT> groups;
        // every mask is the N-wide block of ones moved somewhere else: encode the block once, then each mask is a Galois
        // permutation of that encoding (RotatePlaintext) instead of a 65536-point FFT and one NTT per limb
        auto block0 = m_cc->MakeCKKSPackedPlaintext(generateMaskVector(num_slots, 0), 1, 0, nullptr, num_slots);
        for (int j = 0; j < num_partition / np; j++) {
            if (shard_groups && j % m_world != m_rank) continue;
            const int giant = is * num_partition + j * np;
            CT T;
            for (int i = 0; i < np; i++) {
                // = vectorRotate(generateMaskVector(num_slots, np * j + i), -giant)
                auto pmsk = m_cc->RotatePlaintext(block0, -((np * j + i) * N + giant));
                m_cc->EvalMultAddInPlace(T, preRotatedArrays[i], pmsk);
            }
            groups.push_back(rot.rotate(T, giant));
        }
        CT sum = groups.size() == 1 ? groups[0] : m_cc->EvalAddMany(groups);
        if (shard_groups) m_cc->AllReduceSum(sum);
        return sum;
    }
    // single-batch lengths are sharded over the giant-step groups when every rank gets the same number of them
    bool groupSharded(int B, int groups) const { return m_world > 1 && !sharded(B) && groups >= m_world && groups % m_world == 0; }

    // rank_i = #{ j : a_j < a_i } + 1/2 - 1/2  (all-pairs comparison, SIMD over P shifted copies, B batches)
    CT constructRank(const CT &input_array, SignFunc SignFunc, SignConfig &Cfg) {
        const DirectSortShape sh(N, max_batch);
        std::vector<CT> baby = babySteps(input_array, sh.np_rank, sh.S);
        const std::vector<int> batches = DirectSortBatchesOf(sh.B, m_rank, m_world);
        std::vector<CT> parts(batches.size());
        forEachBatch(m_cc, m_streams, (int)batches.size(), [&](int i) {
            const int b = batches[i];
            CT shifted = vecRotsOpt(baby, sh.P, sh.S, sh.np_rank, b, groupSharded(sh.B, sh.P / sh.np_rank));
            CT dup = input_array->Clone();
            dup->SetSlots(sh.S);
            SignConfig cfg = Cfg; // (a copy per batch: the evaluators take it by non-const reference)
            parts[i] = comp.compare(m_cc, dup, shifted, SignFunc, cfg);
        });
        CT rank = parts.size() == 1 ? parts[0] : m_cc->EvalAddMany(parts);
        parts.clear();
        if (sharded(sh.B)) m_cc->AllReduceSum(rank);
        foldCopies(rank, sh.S, sh.P);
        rank->SetSlots(N);
        return m_cc->EvalSub(rank, 0.5); // the self comparison contributed 1/2
    }

    // masked_inputs[j] = (one-hot mask * input) rotated left by j.  Copy c of the one-hot mask selects the elements whose
    // rank says "move left by ib*P + c"; group i collects copies np*i .. np*i+np-1 and rotates them by ib*P + i*np.
    CT blindRotationOptN(const std::vector<CT> &masked_inputs, int num_slots, int np, int ib, int num_partition, bool shard_groups = false) {
        CT result;
        auto block0 = m_cc->MakeCKKSPackedPlaintext(generateMaskVectorN(num_slots, 0), 1, masked_inputs[0]->GetLevel(), nullptr, num_slots);
        for (int i = 0; i < (num_slots / N) / np; i++) {
            if (shard_groups && i % m_world != m_rank) continue; // independent groups (src/sort_algo.h:561-584)
            CT tmp;
            for (int j = 0; j < np; j++) {
                // = vectorRotate(generateMaskVectorN(num_slots, np * i + j), j), see vecRotsOpt
                auto pmsk = m_cc->RotatePlaintext(block0, j - (np * i + j) * N);
                m_cc->EvalMultAddInPlace(tmp, masked_inputs[j], pmsk);
            }
            tmp = rot.rotate(tmp, ib * num_partition + i * np);
            if (result) m_cc->EvalAddInPlace(result, tmp);
            else result = tmp;
        }
        if (shard_groups) m_cc->AllReduceSum(result);
        return result;
    }

    // Places element i at position rank_i: for every rotation amount r, the doubled-sinc one-hot of (i - rank_i - r)/2N is 1
    // exactly where the element has to move left by r (or right by N - r), then a blind (masked) rotation applies it.
    CT rotationIndexCheckN(const CT &ctx_Rank, const CT &input_array) {
        const DirectSortShape sh(N, max_batch);
        auto index_vector = m_cc->MakeCKKSPackedPlaintext(generateIndexVector(), 1, ctx_Rank->GetLevel(), nullptr, N);
        CT index_minus_rank = m_cc->EvalSub(index_vector, ctx_Rank);
        index_minus_rank->SetSlots(sh.S);
        input_array->SetSlots(sh.S);
        static const std::vector<double> &coeffs = selectDoubledSincCoefficients<N>();
        const std::vector<int> batches = DirectSortBatchesOf(sh.B, m_rank, m_world);
        std::vector<CT> parts(batches.size());
        forEachBatch(m_cc, m_streams, (int)batches.size(), [&](int i) {
            const int b = batches[i];
            auto chk = m_cc->MakeCKKSPackedPlaintext(generateCheckingVectorN(sh.S, b * sh.P), 1, index_minus_rank->GetLevel(),
                                                     nullptr, sh.S);
            CT r = m_cc->EvalSub(index_minus_rank, chk);
            m_cc->EvalMultInPlace(r, 1.0 / N / 2); // (-2N, N) -> (-1, 1/2)
            r = m_cc->EvalChebyshevSeriesPS(r, coeffs, -1, 1);
            CT masked = m_cc->EvalMult(r, input_array);
            std::vector<CT> masked_inputs = babySteps(masked, sh.np_idx, sh.S);
            parts[i] = blindRotationOptN(masked_inputs, sh.S, sh.np_idx, b, sh.P, groupSharded(sh.B, (sh.S / N) / sh.np_idx));
        });
        CT output = parts.size() == 1 ? parts[0] : m_cc->EvalAddMany(parts);
        parts.clear();
        if (sharded(sh.B)) m_cc->AllReduceSum(output);
        foldCopies(output, sh.S, sh.P);
        output->SetSlots(N);
        return output;
    }

    CT sort(const CT &input_array, SignFunc SignFunc, SignConfig &Cfg) override {
        PRINT_PT(m_enc, input_array);
        CT ctx_Rank = constructRank(input_array, SignFunc, Cfg);
        PRINT_PT(m_enc, ctx_Rank);
        CT output_array = rotationIndexCheckN(ctx_Rank, input_array);
        PRINT_PT(m_enc, output_array);
        return output_array;
    }

    // ---- hybrid: the rank of this sort, the placement of MEHP24 (reference src/sort_algo.h:776-1064)
    // row sums of an m x m slot matrix gathered on column `col`: log2 m rotate-adds, towards the right where col's bit is set
    CT sumColumnsToTarget(CT c, size_t m, size_t col, bool maskOutput) {
        c->SetSlots(m * m);
        const int logm = ilog2((int)m);
        size_t step = m >> 1;
        for (int i = 0; i < logm; i++, step >>= 1) {
            const bool right = (col >> (logm - 1 - i)) & 1;
            c = m_cc->EvalAdd(c, rot.rotate(c, right ? -(int)step : (int)step));
        }
        if (maskOutput) {
            std::vector<double> msk(m * m, 0.0);
            for (size_t i = 0; i < m; i++) msk[m * i + col] = 1.0;
            c = m_cc->EvalMult(c, m_cc->MakeCKKSPackedPlaintext(msk, 1, c->GetLevel(), nullptr, m * m));
        }
        return c;
    }
    // column `row` of the matrix laid onto row `row`: element (i, row) moves to (row, i) by steps m(m-1)/2, m(m-1)/4, ...
    CT transposeColumnTarget(CT c, size_t m, size_t row, bool maskOutput) {
        c->SetSlots(m * m);
        const int logm = ilog2((int)m);
        size_t step = m * (m - 1) / 2;
        for (int i = 0; i < logm; i++, step >>= 1) {
            const bool right = (row >> (logm - 1 - i)) & 1;
            c = m_cc->EvalAdd(c, rot.rotate(c, right ? -(int)step : (int)step));
        }
        if (maskOutput) {
            std::vector<double> msk(m * m, 0.0);
            for (size_t i = 0; i < m; i++) msk[m * row + i] = 1.0;
            c = m_cc->EvalMult(c, m_cc->MakeCKKSPackedPlaintext(msk, 1, c->GetLevel(), nullptr, m * m));
        }
        return c;
    }
    // Slot (i, j) of block b asks "is the rank of the element in column j equal to b m + i?": a one-hot indicator of
    // (position - rank) / N (Chebyshev sinc below 256 values, the sign-based indicator from there on) times the input,
    // summed along the rows and transposed, leaves sorted[b m + i] on slot b m + i.  Arrays above 256 values are cut into
    // 256-value blocks on the full ring (every block sees every window of 256 elements through the shifts k * 256).
    CT rotationIndexCheckHybrid(const CT &ctx_Rank, const CT &input_array, PrivateKey<DCRTPoly> = PrivateKey<DCRTPoly>()) {
        const size_t m = std::min<size_t>(N, 256), batches = N / m;
        const size_t num_slots = N > 256 ? (size_t)max_batch : (size_t)N * N;
        if (num_slots > (size_t)max_batch) throw std::invalid_argument("sort_hybrid: N * N slots do not fit this ring");
        ctx_Rank->SetSlots(num_slots);
        input_array->SetSlots(num_slots);
        CT r = m_cc->EvalMult(ctx_Rank, 1.0 / N);
        std::vector<CT> shiftedRank(batches), shiftedInput(batches);
        for (size_t k = 0; k < batches; k++) {
            shiftedRank[k] = k ? rot.rotate(r, (int)(k * 256)) : r;
            shiftedInput[k] = k ? rot.rotate(input_array, (int)(k * 256)) : input_array;
        }
        std::vector<CT> placed(batches);
        for (size_t b = 0; b < batches; b++) {
            std::vector<double> position(num_slots, 0.0);
            for (size_t i = 0; i < m; i++)
                for (size_t j = 0; j < m; j++) position[i * m + j] = (double)(b * m + i) / (double)N;
            Plaintext positionPlain = m_cc->MakeCKKSPackedPlaintext(position, 1, ctx_Rank->GetLevel(), nullptr, num_slots);
            CT acc;
            for (size_t k = 0; k < batches; k++) {
                CT hit = m_cc->EvalSub(positionPlain, shiftedRank[k]);
                if (N < 256) {
                    hit = m_cc->EvalChebyshevSeriesPS(hit, selectCoefficients<N>(), -1, 1);
                } else {
                    SignConfig Cfg = SignConfig(CompositeSignConfig(3, N < 512 ? 4 : 5, 2));
                    hit = comp.indicator(m_cc, hit, 0.5 / N, SignFunc::CompositeSign, Cfg);
                }
                CT term = m_cc->EvalMult(shiftedInput[k], hit);
                acc = k ? m_cc->EvalAdd(acc, term) : term;
            }
            acc = sumColumnsToTarget(acc, m, b, true);
            placed[b] = transposeColumnTarget(acc, m, b, true);
        }
        return batches == 1 ? placed[0] : m_cc->EvalAddMany(placed);
    }
    CT sort_hybrid(const CT &input_array, SignFunc SignFunc, SignConfig &Cfg, PrivateKey<DCRTPoly> sk = PrivateKey<DCRTPoly>()) {
        CT ctx_Rank = constructRank(input_array, SignFunc, Cfg);
        return rotationIndexCheckHybrid(ctx_Rank, input_array, sk);
    }
};

// Bitonic sorting network on one ciphertext (interface and behaviour of the reference's BitonicSort<N>,
// src/sort_algo.h:1066-1160): inputs in [0, 255] are normalised by 1/255; stage (k, j) pairs slot i with i ^ j, ascending
// where (i & k) == 0; the ciphertext is bootstrapped whenever its level exceeds 29.  Per stage: four masks split the slots
// into lower/upper members of ascending/descending pairs, four rotations by -+j bring every member's partner onto its
// slot, one comparison (partner > own) selects between "what to keep if the partner is greater" and "... if it is not".
template <int N> class BitonicSort : public SortBase<N> {
    typedef Ciphertext<DCRTPoly> CT;
    CryptoContext<DCRTPoly> m_cc;
    PublicKey<DCRTPoly> m_PublicKey;
    Comparison comp;
    RotationComposer<N> rot;

  public:
    std::shared_ptr<Encryption> m_enc;
    int bootstraps = 0;

    BitonicSort(CryptoContext<DCRTPoly> cc, PublicKey<DCRTPoly> publicKey, std::vector<int> rotIndices, std::shared_ptr<Encryption> enc)
        : SortBase<N>(enc), m_cc(cc), m_PublicKey(publicKey), comp(enc), rot(cc, enc, rotIndices), m_enc(enc) {}

    CT sort(const CT &input_array, SignFunc SignFunc, SignConfig &Cfg) override {
        CT result = m_cc->EvalMult(input_array, 1.0 / 255);
        for (size_t k = 2; k <= (size_t)N; k *= 2)
            for (size_t j = k / 2; j > 0; j /= 2) {
                if (result->GetLevel() > 29) {
                    result = m_cc->EvalBootstrap(result, 2, 20);
                    bootstraps++;
                }
                // member kinds: 0 lower of an ascending pair, 1 its upper, 2 lower of a descending pair, 3 its upper
                std::vector<std::vector<double>> mask(4, std::vector<double>(N, 0.0));
                for (size_t i = 0; i < (size_t)N; i++) {
                    const size_t l = i ^ j;
                    if (i >= l) continue;
                    const int kind = (i & k) == 0 ? 0 : 2;
                    mask[kind][i] = 1, mask[kind + 1][l] = 1;
                }
                CT part[4], moved[4];
                for (int t = 0; t < 4; t++) {
                    part[t] = m_cc->EvalMult(result, m_cc->MakeCKKSPackedPlaintext(mask[t]));
                    moved[t] = rot.rotate(part[t], (t & 1) ? (int)j : -(int)j); // lowers go up to their partner's slot, uppers down
                }
                CT partner = m_cc->EvalAdd(m_cc->EvalAdd(moved[0], moved[2]), m_cc->EvalAdd(moved[1], moved[3]));
                // if the partner is greater, an ascending pair's lower slot keeps its own value and its upper slot takes the
                // partner's (descending pairs the other way round); if it is not, the roles swap
                CT ifGreater = m_cc->EvalAdd(m_cc->EvalAdd(moved[0], part[0]), m_cc->EvalAdd(moved[3], part[3]));
                CT ifNot = m_cc->EvalAdd(m_cc->EvalAdd(moved[2], part[2]), m_cc->EvalAdd(moved[1], part[1]));
                CT greater = comp.compare(m_cc, partner, result, SignFunc, Cfg);
                result = m_cc->EvalAdd(m_cc->EvalMult(greater, ifGreater), m_cc->EvalMult(m_cc->EvalSub(1, greater), ifNot));
            }
        return m_cc->EvalMult(result, 255.0);
    }
};
