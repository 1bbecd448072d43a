// rotation.h — rotation-index decomposition onto the available rotation keys, composed rotations, and a rotation tree
// with hoisted key switching.  Interface of the reference's src/rotation.h:12-358 (Step, Decomposer<N>,
// RotationComposer<N>, RotationTree<N>), restated.
#pragma once
#include <algorithm>
#include <cstdint>
#include <map>
#include <memory>
#include <set>
#include <vector>

#include "encryption.h"
#include "openfhe.h"

using namespace lbcrypto;

struct Step {
    int8_t value; // -1, 0 or 1 (signed-digit forms); 0 or 1 for binary
    int stepSize; // rotation amount of the digit (a power of two or an available key index)
    Step(int8_t v, int s) : value(v), stepSize(s) {}
    Step(int s) : value(1), stepSize(s) {}
};
enum class DecomposeAlgo { NAF, BNAF, BINARY };

// Splits a rotation amount into steps for which keys exist (the key list is assumed to hold the powers of two below
// the range it wants decomposed, as the reference's lists do).
template <int N> class Decomposer {
    std::vector<int> rotIndices;
    int maxDecomposed; // every amount up to this is reachable with the power-of-two keys at the front of the list

  public:
    Decomposer(std::vector<int> rot) : rotIndices(std::move(rot)) {
        std::sort(rotIndices.begin(), rotIndices.end());
        maxDecomposed = 0;
        int prev = 1;
        for (int idx : rotIndices) { // sum of the doubling chain 2, 4, 8, ... present in the list
            if (prev == idx / 2) maxDecomposed += idx;
            prev = idx;
        }
    }
    std::vector<int> getRotIndices() { return rotIndices; }

    // Conventions of the reference (src/rotation.h:104-165): Step::stepSize is the SIGNED rotation amount of the digit,
    // Step::value its digit (-1, 0, 1), and the steps come largest magnitude first.
    // plain binary digits below N
    std::vector<Step> decomposeBinary(int rotation) const {
        std::vector<Step> s;
        for (int bit = 30; bit >= 0; bit--) {
            const int w = 1 << bit;
            if (w < N && (rotation & w)) s.emplace_back((int8_t)1, w);
        }
        return s;
    }
    // non-adjacent form: digits in {-1, 0, 1}, no two adjacent non-zero (works on negative amounts too: two's complement
    // bits and an arithmetic shift).  A digit worth -N/2 is emitted as +N/2: the two rotations coincide on N slots.
    std::vector<Step> decomposeNAF(int rotation) const {
        std::vector<Step> s;
        for (int bit = 0; rotation != 0; bit++, rotation >>= 1) {
            if (!(rotation & 1)) continue;
            const int d = (rotation & 2) ? -1 : 1;
            const int amount = d * (1 << bit);
            if (amount == -N / 2) s.emplace_back((int8_t)-d, -amount);
            else s.emplace_back((int8_t)d, amount);
            rotation -= d;
        }
        std::reverse(s.begin(), s.end());
        return s;
    }
    // balanced recoding in base 2: a digit above B/2 - or equal to B/2 with the next digit at least B/2 - borrows from
    // the next position (src/rotation.h:138-165)
    std::vector<Step> decomposeBNAF(int k) const {
        const int B = 2;
        std::vector<int> digit;
        for (int rest = k; rest != 0;) {
            int d = rest % B;
            rest = (rest - d) / B;
            if (d > B / 2 || (d == B / 2 && rest % B >= B / 2)) d -= B, rest += 1;
            digit.push_back(d);
        }
        std::vector<Step> s;
        for (size_t i = digit.size(); i-- > 0;)
            if (digit[i] != 0) s.emplace_back((int8_t)digit[i], digit[i] * (1 << i));
        return s;
    }

    std::vector<Step> decompose(int rotation, int wrapN, DecomposeAlgo algo) {
        std::vector<Step> steps;
        if (rotIndices.empty()) return steps;
        const int largest = rotIndices.back();
        while (rotation >= largest) { // peel the largest key off until the remainder is below it
            steps.emplace_back(largest);
            rotation -= largest;
        }
        if (rotation == 0) return steps; // (as the reference: whole-key amounts are returned as they are, wrap steps included)
        while (rotation > maxDecomposed) { // then the largest key strictly below the remainder
            auto it = std::lower_bound(rotIndices.begin(), rotIndices.end(), rotation);
            if (it == rotIndices.begin()) break;
            int legal = *(it - 1);
            steps.emplace_back(legal);
            rotation -= legal;
        }
        if (rotation == 0) return steps;
        {
            std::vector<Step> rest = algo == DecomposeAlgo::NAF    ? decomposeNAF(rotation)
                                     : algo == DecomposeAlgo::BNAF ? decomposeBNAF(rotation)
                                                                   : decomposeBinary(rotation);
            steps.insert(steps.end(), rest.begin(), rest.end());
        }
        steps.erase(std::remove_if(steps.begin(), steps.end(), [wrapN](const Step &s) { return s.stepSize % wrapN == 0; }),
                    steps.end());
        return steps;
    }
};

template <int N> class RotationComposer {
    CryptoContext<DCRTPoly> m_cc;
    std::shared_ptr<Encryption> m_enc;
    Decomposer<N> m_decomposer;
    DecomposeAlgo m_algo;
    std::set<int> rotation_calls;
    std::set<int> available_indices;

  public:
    RotationComposer(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, std::vector<int> rotIndices,
                     DecomposeAlgo algo = DecomposeAlgo::BINARY)
        : m_cc(cc), m_enc(enc), m_decomposer(rotIndices), m_algo(algo),
          available_indices(rotIndices.begin(), rotIndices.end()) {}

    // identity when the amount is a multiple of the slot count; one EvalRotate when the key exists; a chain otherwise
    Ciphertext<DCRTPoly> rotate(const Ciphertext<DCRTPoly> &input, int rotation) {
        if (rotation % (int)input->GetSlots() == 0) return input->Clone();
        if (available_indices.count(rotation)) return m_cc->EvalRotate(input, rotation);
        Ciphertext<DCRTPoly> r = input->Clone();
        for (const Step &s : m_decomposer.decompose(rotation, (int)input->GetSlots(), m_algo))
            r = m_cc->EvalRotate(r, s.stepSize);
        return r;
    }
    const std::set<int> &getRotationCalls() const { return rotation_calls; }
};

// Memoising rotation tree: rotations that share a prefix of decomposed steps share the intermediate ciphertexts, and
// a node with more than two children rotates them all from one hoisted ModUp (EvalFastRotationPrecompute).
template <int N> class RotationTree {
    struct Node {
        Ciphertext<DCRTPoly> ct;
        std::map<int, std::unique_ptr<Node>> child; // signed step -> node
    };
    CryptoContext<DCRTPoly> m_cc;
    Decomposer<N> m_decomposer;
    DecomposeAlgo m_algo;
    Node root;
    std::map<int, std::vector<int>> paths; // rotation -> signed steps
    uint32_t M;

    void fill(Node &node) {
        std::vector<int> todo;
        for (auto &kv : node.child)
            if (!kv.second->ct) todo.push_back(kv.first);
        if (todo.size() > 2) { // hoisting pays once the ModUp is shared by several rotations (reference: rotation.h:342)
            auto pre = m_cc->EvalFastRotationPrecompute(node.ct);
            for (int s : todo) node.child[s]->ct = m_cc->EvalFastRotation(node.ct, s, M, pre);
        } else {
            for (int s : todo) node.child[s]->ct = m_cc->EvalRotate(node.ct, s);
        }
        for (auto &kv : node.child) fill(*kv.second);
    }

  public:
    RotationTree(CryptoContext<DCRTPoly> cc, std::vector<int> rotIndices, DecomposeAlgo algo = DecomposeAlgo::NAF)
        : m_cc(cc), m_decomposer(rotIndices), m_algo(algo), M(cc->GetCyclotomicOrder()) {}

    // register rotations start..end-1 (negative amounts rotate right)
    void buildTree(int start, int end) {
        for (int r = start; r <= end; r++) { // end inclusive, as the reference
            if (r == 0) continue;
            std::vector<int> p;
            for (const Step &s : m_decomposer.decompose(std::abs(r), N * N, m_algo)) {
                p.push_back(s.stepSize * (r < 0 ? -1 : 1));
            }
            std::sort(p.begin(), p.end(), [](int a, int b) { return std::abs(a) > std::abs(b); });
            paths[r] = p;
            Node *n = &root;
            for (int v : p) {
                auto &c = n->child[v];
                if (!c) c = std::make_unique<Node>();
                n = c.get();
            }
        }
    }
    // evaluate (or reuse) every registered rotation of `input`, then return the requested one
    Ciphertext<DCRTPoly> treeRotate(const Ciphertext<DCRTPoly> &input, int rotation) {
        if (rotation == 0) return input->Clone();
        if (root.ct.get() != input.get()) {
            std::function<void(Node &)> clear = [&](Node &n) {
                n.ct = Ciphertext<DCRTPoly>();
                for (auto &kv : n.child) clear(*kv.second);
            };
            clear(root);
            root.ct = input;
            fill(root);
        }
        auto it = paths.find(rotation);
        if (it == paths.end()) throw std::out_of_range("RotationTree: rotation was not registered with buildTree");
        Node *n = &root;
        for (int v : it->second) n = n->child[v].get();
        return n->ct;
    }
};
