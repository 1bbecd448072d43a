// kway_adapter.h — the k-way sorting network behind the SortBase<N> interface, with the parameter set the reference
// fixes for it (interface and values of the reference's src/kway_adapter.h:23-73).
#pragma once
#include <memory>
#include <stdexcept>

#include "kway.h"
#include "sort_algo.h"

constexpr int next_power_of_two(int n) {
    int p = 1;
    while (p < n) p <<= 1;
    return p;
}

template <int N> class KWayAdapter : public SortBase<N> {
    CryptoContext<DCRTPoly> m_cc;
    PublicKey<DCRTPoly> m_PublicKey;
    std::shared_ptr<Encryption> m_enc;
    std::unique_ptr<kwaySort::Sorter> m_sorter;

  public:
    KWayAdapter(CryptoContext<DCRTPoly> cc, PublicKey<DCRTPoly> publicKey, PrivateKey<DCRTPoly> privateKey,
                std::shared_ptr<Encryption> enc, int k, int M)
        : SortBase<N>(enc), m_cc(cc), m_PublicKey(publicKey), m_enc(enc) {
        if (kwaySort::ipow(k, M) != N) throw std::invalid_argument("KWayAdapter: k^M must equal the array length N");
        m_sorter = std::make_unique<kwaySort::Sorter>(cc, enc, (long)N, k, M, privateKey, publicKey);
    }
    // batch = next power of two >= N, first modulus 60 bits, scaling 59 bits, depth 40; rotation keys +-2^i < N;
    // bootstrapping level budget {4,4} up to N = 128, {5,5} beyond
    static void getSizeParameters(CCParams<CryptoContextCKKSRNS> &parameters, std::vector<int> &rotations,
                                  std::vector<uint32_t> &levelBudget) {
        parameters.SetBatchSize(next_power_of_two(N));
        parameters.SetFirstModSize(60);
        parameters.SetScalingModSize(59);
        for (int step = 1; step < N; step *= 2) rotations.push_back(step), rotations.push_back(-step);
        levelBudget = N <= 128 ? std::vector<uint32_t>{4, 4} : std::vector<uint32_t>{5, 5};
        parameters.SetMultiplicativeDepth(40);
    }
    Ciphertext<DCRTPoly> sort(const Ciphertext<DCRTPoly> &input_array, SignFunc, SignConfig &Cfg) override {
        Ciphertext<DCRTPoly> in = input_array->Clone(), out;
        m_sorter->sorter(in, out, Cfg);
        return out;
    }
    int bootstraps() const { return m_sorter->bootstraps; }
};
