// kway.h — the HKC+21 k-way sorting network on one packed ciphertext, with bootstrapping between stages.
// Interface of the reference's src/k-way/{Masking,EvalUtils,SortUtils,Sorter}.h (namespace kwaySort, same class and
// method names and argument meaning); the bodies are restated from the behaviour of src/k-way/*.cpp:
//   Masking   sortType (Masking.cpp:26-47), genIndices (:49-144), genMask (:146-156), getRotateDistance (:158-167)
//   EvalUtils leftRotate / rightRotate by binary chains over the +-2^i keys (EvalUtils.cpp:106-147), checkLevelAndBoot
//             (:57-86), flipCtxt (:96-104)
//   SortUtils fcnL, compareMax/Min, two/threeSorter (SortUtils.cpp:5-78), slotMatching2/3 (:202-237), slotAssemble (:425)
//   Sorter    rightRotateForSort (Sorter.cpp:180-249), comparisonForSort (:251-262), runTwo/ThreeSorter (:8-69),
//             sorter (:284-404)
//   k = 5     four/fiveSorter (SortUtils.cpp:80-200), slotMatching4/5/2345 (:239-423), runFour/Five/2345Sorter
//             (Sorter.cpp:70-178), comparisonForSort2 (:264-282).  Here the pairwise bits of every sorter size come from one
//             rule (pairBit below) instead of one hand-written table per size.
// One ciphertext, strictly sequential stages: nothing to shard across GPUs (replicas only).
#pragma once
#include <algorithm>
#include <cmath>
#include <memory>
#include <stdexcept>
#include <tuple>
#include <vector>

#include "comparison.h"
#include "encryption.h"
#include "openfhe.h"
#include "sign.h"

namespace kwaySort {

// ------------------------------------------------------------------ network geometry (plain integers)
inline long ipow(long b, long e) {
    long r = 1;
    while (e-- > 0) r *= b;
    return r;
}

// stage -> (m, logDist, slope).  Merge round r (merging runs of k^r values) holds 1 + r * ceil(k/2) stages: one k-sorter
// stage (slope 0) and, for every earlier level m = 1..r, ceil(k/2) clean-up stages with slopes 1..ceil(k/2).
inline std::tuple<int, int, int> sortType(int k, int M, int stage) {
    (void)M;
    const int per = (k + 1) / 2;
    int round = 0, first = 0; // first = index of the round's first stage
    while (stage >= first + 1 + round * per) first += 1 + round * per, round++;
    const int off = stage - first;
    const int m = (off + per - 1) / per;
    return std::make_tuple(m, round - m, off == 0 ? 0 : (off - 1) % per + 1);
}

// indices[0][slot] = size of the sorter the slot takes part in at this stage (0: the slot rests), indices[1][slot] = its
// 1-based position inside that sorter.  Slots are laid out as blocks of dist * k^(m+1) values; inside a block, element
// (row, col) of a k^m x k grid sits at dist * (col + k * row) (+ d < dist); at slope 0 the grid is taken column-major
// with k^m rows instead (row + k^m * col).
inline std::vector<std::vector<int>> genIndices(long numSlots, long k, long M, long m, long logDist, long slope) {
    std::vector<std::vector<int>> idx(2, std::vector<int>((size_t)numSlots, 0));
    const long rows = ipow(k, m), dist = ipow(k, logDist), block = dist * ipow(k, m + 1), total = ipow(k, M);
    for (long base = 0; base < total; base += block) {
        if (slope == 0) { // every k^m-strided k-tuple goes through a k-sorter
            for (long row = 0; row < rows; row++)
                for (long col = 0; col < k; col++)
                    for (long d = 0; d < dist; d++) {
                        const long at = base + dist * (row + rows * col) + d;
                        idx[0][at] = (int)k, idx[1][at] = (int)col + 1;
                    }
        } else if (slope > k / 2) { // (k-1)-sorters across neighbouring rows, starting at column k - k/2
            const long col0 = k - k / 2;
            for (long t = 0; t + 1 < rows; t++)
                for (long pos = 1; pos < k; pos++)
                    for (long d = 0; d < dist; d++) {
                        const long at = base + dist * (col0 + k * t + pos - 1) + d;
                        idx[0][at] = (int)k - 1, idx[1][at] = (int)pos;
                    }
        } else { // anti-diagonals of the grid: from (row0, col0) step (+1, -slope) until the grid ends
            auto walk = [&](long row0, long col0) {
                long row = row0, col = col0;
                for (int len = 1; row < rows && col >= 0; len++, row++, col -= slope)
                    for (long d = 0; d < dist; d++) {
                        idx[0][base + dist * (col + k * row) + d] = len;
                        if (row == rows - 1 || col - slope < 0) // last element: number the whole diagonal backwards
                            for (int i = 0; i < len; i++) {
                                const long at = base + dist * ((col + i * slope) + k * (row - i)) + d;
                                idx[1][at] = len - i;
                                idx[0][at] += i;
                            }
                    }
            };
            for (long t = slope; t < k; t++) walk(0, t);
            for (long s = 1; s + 1 < rows; s++)
                for (long t = k - slope; t < k; t++) walk(s, t);
        }
    }
    return idx;
}

inline void genMask(const std::vector<std::vector<int>> &indices, long index0, long index1, std::vector<double> &mask) {
    const size_t slots = indices[0].size();
    mask.resize(slots, 0.0);
    for (size_t i = 0; i < slots; i++)
        if (indices[0][i] == index0 && indices[1][i] == index1) mask[i] = 1.0;
}

inline long getRotateDistance(long k, long logDist, long slope) {
    const long dist = ipow(k, logDist);
    return (slope == 0 || slope == k / 2 + 1) ? dist : dist * (k - slope);
}

// ------------------------------------------------------------------ rotations, bootstrapping guard, 1 - x
class EvalUtils {
  public:
    EvalUtils(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc) : m_cc(cc), m_enc(enc) {}
    EvalUtils(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, const PrivateKey<DCRTPoly> &, const PublicKey<DCRTPoly> &)
        : m_cc(cc), m_enc(enc) {}

    // rotation by r >= 0 as the chain of its binary digits (keys exist for +-2^i only, src/kway_adapter.h:48-51)
    void leftRotate(Ciphertext<DCRTPoly> &ctxt, long r, Ciphertext<DCRTPoly> &ctxt_out) { chain(ctxt, r, +1, ctxt_out); }
    void rightRotate(Ciphertext<DCRTPoly> &ctxt, long r, Ciphertext<DCRTPoly> &ctxt_out) { chain(ctxt, r, -1, ctxt_out); }

    // bootstrap when fewer than level + 1 levels remain (a ciphertext at the last level cannot be bootstrapped)
    void checkLevelAndBoot(Ciphertext<DCRTPoly> &ctxt, int level, int multDepth, bool verbose = false) {
        (void)verbose;
        if (multDepth - (int)ctxt->GetLevel() < level + 1) {
            ctxt = m_cc->EvalBootstrap(ctxt);
            bootstraps++;
        }
    }
    void checkLevelAndBoot2(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &b, long level, long multDepth, bool verbose = false) {
        checkLevelAndBoot(a, (int)level, (int)multDepth, verbose);
        checkLevelAndBoot(b, (int)level, (int)multDepth, verbose);
    }
    // integer multiple without consuming a level: double-and-add over the bits of |coeff| (src/k-way/EvalUtils.cpp:18-44)
    void multByInt(Ciphertext<DCRTPoly> &ctxt, long coeff, Ciphertext<DCRTPoly> &ctxt_out) {
        if (coeff == 0) throw std::invalid_argument("kwaySort::multByInt: coefficient must be non-zero");
        const Ciphertext<DCRTPoly> unit = coeff < 0 ? m_cc->EvalNegate(ctxt) : ctxt;
        unsigned long mag = (unsigned long)(coeff < 0 ? -coeff : coeff);
        int top = 0;
        while ((mag >> (top + 1)) != 0) top++;
        Ciphertext<DCRTPoly> acc = unit;
        for (int bit = top - 1; bit >= 0; bit--) {
            acc = m_cc->EvalAdd(acc, acc);
            if ((mag >> bit) & 1ul) acc = m_cc->EvalAdd(acc, unit);
        }
        ctxt_out = acc;
    }
    // HEAAN's "kill imaginary part" products: the packing here is real, so they are the plain product and square
    void multAndKillImage(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &b, Ciphertext<DCRTPoly> &out) { out = m_cc->EvalMult(a, b); }
    void squareAndKillImage(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &out) { out = m_cc->EvalSquare(a); }
    void flipCtxt(Ciphertext<DCRTPoly> &ctxt) { ctxt = m_cc->EvalSub(1.0, ctxt); }
    void flipCtxt(Ciphertext<DCRTPoly> &ctxt, const Plaintext &mask) { ctxt = m_cc->EvalSub(mask, ctxt); }

    int bootstraps = 0; // EvalBootstrap calls made by the level guards of this object

  protected:
    CryptoContext<DCRTPoly> m_cc;
    std::shared_ptr<Encryption> m_enc;

  private:
    void chain(const Ciphertext<DCRTPoly> &in, long r, int dir, Ciphertext<DCRTPoly> &out) {
        if (r < 0) throw std::invalid_argument("kwaySort: rotation amounts are non-negative");
        Ciphertext<DCRTPoly> acc = in;
        for (long bit = 1; bit <= r; bit <<= 1)
            if (r & bit) acc = m_cc->EvalRotate(acc, (int)(dir * bit));
        out = acc;
    }
};

// ------------------------------------------------------------------ small sorters on slot-aligned ciphertexts
class SortUtils : public EvalUtils {
  public:
    SortUtils(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, long numSlots, long k, long M)
        : EvalUtils(cc, enc), m_numSlots(numSlots), m_k(k), m_M(M) {}
    SortUtils(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, long numSlots, long k, long M,
              const PrivateKey<DCRTPoly> &sk, const PublicKey<DCRTPoly> &pk)
        : EvalUtils(cc, enc, sk, pk), m_numSlots(numSlots), m_k(k), m_M(M) {}

    // comp = (a > b):  comp * (a - b) + b = max(a, b)
    void fcnL(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &b, Ciphertext<DCRTPoly> &comp, Ciphertext<DCRTPoly> &out) {
        out = m_cc->EvalAdd(m_cc->EvalMult(m_cc->EvalSub(a, b), comp), b);
    }
    void compareMax(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &b, Ciphertext<DCRTPoly> &comp, Ciphertext<DCRTPoly> &out) {
        fcnL(a, b, comp, out);
    }
    void compareMin(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &b, Ciphertext<DCRTPoly> &comp, Ciphertext<DCRTPoly> &out) {
        fcnL(b, a, comp, out);
    }
    // [a, b], comp = (a > b)  ->  [min, max]; the minimum is a + b - max (no second multiplication)
    void twoSorter(Ciphertext<DCRTPoly> &a, Ciphertext<DCRTPoly> &b, Ciphertext<DCRTPoly> &comp, Ciphertext<DCRTPoly> *out) {
        fcnL(a, b, comp, out[1]);
        out[0] = m_cc->EvalSub(m_cc->EvalAdd(a, b), out[1]);
    }
    void twoSorter(Ciphertext<DCRTPoly> *ctxt, Ciphertext<DCRTPoly> &comp, Ciphertext<DCRTPoly> *out) {
        twoSorter(ctxt[0], ctxt[1], comp, out);
    }
    // [a, b, c], comp = [(a > b), (a > c), (b > c)]  ->  [min, middle, max].  The comparison bits of max(a,b) and
    // min(a,b) against c are themselves selected from (a > c), (b > c) by the two-sorter on bits.
    void threeSorter(Ciphertext<DCRTPoly> *ctxt, Ciphertext<DCRTPoly> *comp, Ciphertext<DCRTPoly> *out) {
        Ciphertext<DCRTPoly> ab[2], abVsC[2];
        twoSorter(ctxt[0], ctxt[1], comp[0], ab);
        twoSorter(comp[1], comp[2], comp[0], abVsC);
        compareMax(ab[1], ctxt[2], abVsC[1], out[2]);
        compareMin(ab[0], ctxt[2], abVsC[0], out[0]);
        auto sum = m_cc->EvalAdd(m_cc->EvalAdd(ctxt[0], ctxt[1]), ctxt[2]);
        out[1] = m_cc->EvalSub(m_cc->EvalSub(sum, out[0]), out[2]);
    }
    // [a, b, c, d], comp = [(a>b), (a>c), (a>d), (b>c), (b>d), (c>d)] -> ascending.  Sort the halves {a,b}, {c,d}; the
    // bits between the halves' extremes are selected from the input bits by the same two-sorters (a bit vector sorted by
    // (a>b) is [bit of min(a,b), bit of max(a,b)]); then merge.
    void fourSorter(Ciphertext<DCRTPoly> *ctxt, Ciphertext<DCRTPoly> *comp, Ciphertext<DCRTPoly> *out) {
        Ciphertext<DCRTPoly> lo[2], hi[2], loVsC[2], loVsD[2], bigVsHi[2], smallVsHi[2];
        twoSorter(ctxt[0], ctxt[1], comp[0], lo);            // lo = [min(a,b), max(a,b)]
        twoSorter(ctxt[2], ctxt[3], comp[5], hi);            // hi = [min(c,d), max(c,d)]
        twoSorter(comp[1], comp[3], comp[0], loVsC);         // [(lo0 > c), (lo1 > c)]
        twoSorter(comp[2], comp[4], comp[0], loVsD);         // [(lo0 > d), (lo1 > d)]
        twoSorter(loVsC[1], loVsD[1], comp[5], bigVsHi);     // [(lo1 > hi0), (lo1 > hi1)]
        twoSorter(loVsC[0], loVsD[0], comp[5], smallVsHi);   // [(lo0 > hi0), (lo0 > hi1)]
        compareMax(lo[1], hi[1], bigVsHi[1], out[3]);
        Ciphertext<DCRTPoly> ifLoWins, ifHiWins;             // runner-up: max(lo0, hi1) if lo1 is the maximum, else max(lo1, hi0)
        compareMax(lo[0], hi[1], smallVsHi[1], ifLoWins);
        compareMax(lo[1], hi[0], bigVsHi[0], ifHiWins);
        compareMax(ifLoWins, ifHiWins, bigVsHi[1], out[2]);
        compareMin(lo[0], hi[0], smallVsHi[0], out[0]);
        restByTotal(ctxt, 4, out, 1);
    }
    // [a..e], comp = [a>b a>c a>d a>e b>c b>d b>e c>d c>e d>e] -> ascending: 3-sorter on {a,b,c}, 2-sorter on {d,e}, the
    // bits of the sorted triple against d and e by the same 3-sorter applied to the bit vectors, then merge
    void fiveSorter(Ciphertext<DCRTPoly> *ctxt, Ciphertext<DCRTPoly> *comp, Ciphertext<DCRTPoly> *out) {
        Ciphertext<DCRTPoly> tripleBits[3] = {comp[0], comp[1], comp[4]}, triple[3], pair[2];
        threeSorter(ctxt, tripleBits, triple);
        twoSorter(ctxt[3], ctxt[4], comp[9], pair);
        Ciphertext<DCRTPoly> vsD[3] = {comp[2], comp[5], comp[7]}, vsE[3] = {comp[3], comp[6], comp[8]}, tVsD[3], tVsE[3];
        threeSorter(vsD, tripleBits, tVsD);                  // (triple[i] > d)
        threeSorter(vsE, tripleBits, tVsE);                  // (triple[i] > e)
        Ciphertext<DCRTPoly> vsPair[3][2];                   // [(triple[i] > pair[0]), (triple[i] > pair[1])]
        for (int i = 0; i < 3; i++) twoSorter(tVsD[i], tVsE[i], comp[9], vsPair[i]);
        compareMax(triple[2], pair[1], vsPair[2][1], out[4]);
        compareMin(triple[0], pair[0], vsPair[0][0], out[0]);
        Ciphertext<DCRTPoly> x, y;
        compareMax(triple[1], pair[1], vsPair[1][1], x);     // fourth: max(triple1, pair1) if triple2 is the maximum ...
        compareMax(triple[2], pair[0], vsPair[2][0], y);     // ... else max(triple2, pair0)
        compareMax(x, y, vsPair[2][1], out[3]);
        compareMin(triple[1], pair[0], vsPair[1][0], x);     // second: min(triple1, pair0) if pair0 is not the minimum ...
        compareMin(triple[0], pair[1], vsPair[0][1], y);     // ... else min(triple0, pair1)
        compareMin(x, y, vsPair[0][0], out[1]);
        restByTotal(ctxt, 5, out, 2);
    }

    // bring partner p of every sorter onto the slot of its first element (rotation left by p * shift)
    void slotMatching2(Ciphertext<DCRTPoly> &ctxt, Ciphertext<DCRTPoly> &comp, std::vector<std::vector<int>> &indices, long shift,
                       Ciphertext<DCRTPoly> *out, Ciphertext<DCRTPoly> &comp_out) {
        (void)indices;
        out[0] = ctxt;
        leftRotate(ctxt, shift, out[1]);
        comp_out = comp;
    }
    // comp holds (x_i > x_{i-1}) style bits of neighbours at distance shift (and the wrap-around pair); at the first slot
    // of a sorter that gives (a > c); one and two shifts further (b > a) and (c > b), which are flipped into (a > b), (b > c)
    void slotMatching3(Ciphertext<DCRTPoly> &ctxt, Ciphertext<DCRTPoly> &comp, std::vector<std::vector<int>> &indices, long shift,
                       Ciphertext<DCRTPoly> *out, Ciphertext<DCRTPoly> *comp_out) {
        std::vector<double> first;
        genMask(indices, 3, 1, first);
        auto firstPlain = m_cc->MakeCKKSPackedPlaintext(first);
        for (int p = 0; p < 3; p++) leftRotate(ctxt, p * shift, out[p]);
        comp_out[1] = comp;
        leftRotate(comp, shift, comp_out[0]);
        leftRotate(comp, 2 * shift, comp_out[2]);
        flipCtxt(comp_out[0], firstPlain);
        flipCtxt(comp_out[2], firstPlain);
    }
    // The comparison ciphertexts hold, at member p of a cyclic s-sorter, comp1 = (x_p > x_{p-1}) and comp2 = (x_p > x_{p-2})
    // (indices mod s; Sorter::comparisonForSort2).  The bit (x_i > x_j), i < j, on the sorter's first slot is therefore
    //   comp_{s-D} rotated left by i * shift              if s - D is 1 or 2 (member i looks back onto j), D = j - i
    //   mask - (comp_D rotated left by j * shift)         otherwise (member j looks back onto i: flipped)
    // masked = true multiplies by `mask` (first slots of the sorters the bit is meant for) instead of leaving stray values
    Ciphertext<DCRTPoly> pairBit(int s, int i, int j, Ciphertext<DCRTPoly> &comp1, Ciphertext<DCRTPoly> &comp2, long shift,
                                 const Plaintext &mask, bool masked) {
        const int D = j - i, back = s - D;
        const bool direct = back == 1 || back == 2;
        Ciphertext<DCRTPoly> bit;
        leftRotate((direct ? back : D) == 1 ? comp1 : comp2, (direct ? i : j) * shift, bit);
        if (masked) bit = m_cc->EvalMult(bit, mask);
        if (!direct) flipCtxt(bit, mask);
        return bit;
    }
    void slotMatching4(Ciphertext<DCRTPoly> &ctxt, Ciphertext<DCRTPoly> &comp1, Ciphertext<DCRTPoly> &comp2,
                       std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> *arr, Ciphertext<DCRTPoly> *comp_arr) {
        std::vector<double> first;
        genMask(indices, 4, 1, first);
        auto firstPlain = m_cc->MakeCKKSPackedPlaintext(first);
        int at = 0;
        for (int i = 0; i < 4; i++)
            for (int j = i + 1; j < 4; j++) comp_arr[at++] = pairBit(4, i, j, comp1, comp2, shift, firstPlain, false);
        for (int p = 0; p < 4; p++) { // the four-sorter's outputs are not masked afterwards: mask its inputs
            leftRotate(ctxt, p * shift, arr[p]);
            arr[p] = m_cc->EvalMult(arr[p], firstPlain);
        }
    }
    void slotMatching5(Ciphertext<DCRTPoly> &ctxt, Ciphertext<DCRTPoly> &comp1, Ciphertext<DCRTPoly> &comp2,
                       std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> *arr, Ciphertext<DCRTPoly> *comp_arr) {
        std::vector<double> first;
        genMask(indices, 5, 1, first);
        auto firstPlain = m_cc->MakeCKKSPackedPlaintext(first);
        for (int p = 0; p < 5; p++) leftRotate(ctxt, p * shift, arr[p]);
        int at = 0;
        for (int i = 0; i < 5; i++)
            for (int j = i + 1; j < 5; j++) comp_arr[at++] = pairBit(5, i, j, comp1, comp2, shift, firstPlain, false);
    }
    // sorters of 2, 3, 4 and 5 members in one stage, all run through the five-sorter: a bit involving a member the sorter
    // does not have is 0 ("the missing member is larger"), so the real members come out in the low positions
    void slotMatching2345(Ciphertext<DCRTPoly> &ctxt, Ciphertext<DCRTPoly> &comp1, Ciphertext<DCRTPoly> &comp2,
                          std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> *arr, Ciphertext<DCRTPoly> *comp_arr) {
        for (int p = 0; p < 5; p++) leftRotate(ctxt, p * shift, arr[p]);
        std::vector<std::vector<double>> first(6);
        for (int s = 2; s <= 5; s++) genMask(indices, s, 1, first[s]);
        int at = 0;
        for (int i = 0; i < 5; i++)
            for (int j = i + 1; j < 5; j++, at++) {
                // sizes that have member j, grouped by where their bit comes from (the rule of pairBit)
                Ciphertext<DCRTPoly> sum;
                bool done[6] = {false, false, false, false, false, false};
                for (int s = j + 1; s <= 5; s++) {
                    if (done[s]) continue;
                    const int D = j - i;
                    auto source = [&](int size) { const int back = size - D; return (back == 1 || back == 2) ? back : -D; };
                    std::vector<double> group((size_t)m_numSlots, 0.0);
                    for (int t = s; t <= 5; t++)
                        if (!done[t] && source(t) == source(s)) {
                            done[t] = true;
                            for (long x = 0; x < m_numSlots; x++) group[x] += first[t][x];
                        }
                    Ciphertext<DCRTPoly> term = pairBit(s, i, j, comp1, comp2, shift, m_cc->MakeCKKSPackedPlaintext(group), true);
                    sum = sum ? m_cc->EvalAdd(sum, term) : term;
                }
                comp_arr[at] = sum;
            }
    }
    // sorted element i goes back i * shift slots to the right
    void slotAssemble(Ciphertext<DCRTPoly> *sorted, long num, long shift, Ciphertext<DCRTPoly> &out) {
        out = sorted[0];
        for (long i = 1; i < num; i++) {
            Ciphertext<DCRTPoly> moved;
            rightRotate(sorted[i], i * shift, moved);
            out = m_cc->EvalAdd(out, moved);
        }
    }

  protected:
    // out[which] = sum of the inputs - sum of the other outputs (the one element the merge did not have to select)
    void restByTotal(Ciphertext<DCRTPoly> *ctxt, int count, Ciphertext<DCRTPoly> *out, int which) {
        Ciphertext<DCRTPoly> acc = ctxt[0];
        for (int i = 1; i < count; i++) acc = m_cc->EvalAdd(acc, ctxt[i]);
        for (int i = 0; i < count; i++)
            if (i != which) acc = m_cc->EvalSub(acc, out[i]);
        out[which] = acc;
    }
    long m_numSlots, m_k, m_M;
};

// ------------------------------------------------------------------ the network
class Sorter : public SortUtils {
  public:
    Sorter(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, long numSlots, long k, long M)
        : SortUtils(cc, enc, numSlots, k, M), comp(enc) {}
    Sorter(CryptoContext<DCRTPoly> cc, std::shared_ptr<Encryption> enc, long numSlots, long k, long M,
           const PrivateKey<DCRTPoly> &sk, const PublicKey<DCRTPoly> &pk)
        : SortUtils(cc, enc, numSlots, k, M, sk, pk), comp(enc) {}

    // ctxt_rot: at the first slot of every sorter the value of its last element, at every other member the value of its
    // predecessor (so that one compare yields all neighbour bits); ctxt_fix: the slots that rest in this stage
    void rightRotateForSort(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long logDist, long slope,
                            Ciphertext<DCRTPoly> &ctxt_rot, Ciphertext<DCRTPoly> &ctxt_fix) {
        const long k = m_k;
        std::vector<double> inner((size_t)m_numSlots, 0.0); // members that have a successor: they move right by rot
        std::vector<std::vector<double>> last((size_t)k, std::vector<double>((size_t)m_numSlots, 0.0)); // last of an s-sorter, by s
        for (long i = 0; i < m_numSlots; i++) {
            const int size = indices[0][i], pos = indices[1][i];
            if (pos < size) inner[i] = 1.0;
            if (size > 0 && pos == size) last[size - 1][i] = 1.0;
        }
        const long rot = getRotateDistance(k, logDist, slope);
        auto masked = [&](const std::vector<double> &m) { return m_cc->EvalMult(ctxt, m_cc->MakeCKKSPackedPlaintext(m)); };
        Ciphertext<DCRTPoly> left = masked(inner);
        rightRotate(left, rot, ctxt_rot);
        if (slope == 0 || slope == k / 2 + 1) { // one sorter size in the stage: k, or k - 1 for the middle slope
            const long size = slope == 0 ? k : k - 1;
            Ciphertext<DCRTPoly> right = masked(last[size - 1]), wrapped;
            if (slope != 0) ctxt_fix = m_cc->EvalSub(m_cc->EvalSub(ctxt, left), right);
            leftRotate(right, (size - 1) * rot, wrapped);
            ctxt_rot = m_cc->EvalAdd(ctxt_rot, wrapped);
        } else { // diagonals of every length 1..k
            std::vector<Ciphertext<DCRTPoly>> right((size_t)k);
            ctxt_fix = m_cc->EvalSub(ctxt, left);
            for (long s = 0; s < k; s++) {
                right[s] = masked(last[s]);
                ctxt_fix = m_cc->EvalSub(ctxt_fix, right[s]);
            }
            for (long s = 1; s < k; s++) {
                Ciphertext<DCRTPoly> wrapped;
                leftRotate(right[s], s * rot, wrapped);
                ctxt_rot = m_cc->EvalAdd(ctxt_rot, wrapped);
            }
        }
    }
    void comparisonForSort(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long logDist, long slope,
                           Ciphertext<DCRTPoly> &ctxt_comp, Ciphertext<DCRTPoly> &ctxt_fix, SignConfig &Cfg) {
        Ciphertext<DCRTPoly> ctxt_rot;
        rightRotateForSort(ctxt, indices, logDist, slope, ctxt_rot, ctxt_fix);
        ctxt_comp = comp.compare(m_cc, ctxt, ctxt_rot, SignFunc::CompositeSign, Cfg);
    }
    // comp1 = (x_p > x_{p-1}), comp2 = (x_p > x_{p-2}) inside every cyclic sorter: the shifted copy shifted once more
    void comparisonForSort2(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long logDist, long slope,
                            Ciphertext<DCRTPoly> &ctxt_comp1, Ciphertext<DCRTPoly> &ctxt_comp2, Ciphertext<DCRTPoly> &ctxt_fix,
                            SignConfig &Cfg) {
        Ciphertext<DCRTPoly> once, twice, unused;
        rightRotateForSort(ctxt, indices, logDist, slope, once, ctxt_fix);
        rightRotateForSort(once, indices, logDist, slope, twice, unused);
        ctxt_comp1 = comp.compare(m_cc, ctxt, once, SignFunc::CompositeSign, Cfg);
        ctxt_comp2 = comp.compare(m_cc, ctxt, twice, SignFunc::CompositeSign, Cfg);
    }

    void runTwoSorter(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> &ctxt_comp,
                      Ciphertext<DCRTPoly> &ctxt_out) {
        std::vector<double> first;
        genMask(indices, 2, 1, first);
        auto firstPlain = m_cc->MakeCKKSPackedPlaintext(first);
        Ciphertext<DCRTPoly> aligned[2], bits, sorted[2];
        slotMatching2(ctxt, ctxt_comp, indices, shift, aligned, bits);
        twoSorter(aligned, bits, sorted);
        for (auto &c : sorted) c = m_cc->EvalMult(c, firstPlain);
        slotAssemble(sorted, 2, shift, ctxt_out);
    }
    void runThreeSorter(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> &ctxt_comp,
                        Ciphertext<DCRTPoly> &ctxt_out) {
        std::vector<double> first;
        genMask(indices, 3, 1, first);
        auto firstPlain = m_cc->MakeCKKSPackedPlaintext(first);
        Ciphertext<DCRTPoly> aligned[3], bits[3], sorted[3];
        slotMatching3(ctxt, ctxt_comp, indices, shift, aligned, bits);
        threeSorter(aligned, bits, sorted);
        for (auto &c : sorted) c = m_cc->EvalMult(c, firstPlain);
        slotAssemble(sorted, 3, shift, ctxt_out);
    }
    void runFourSorter(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> &ctxt_comp1,
                       Ciphertext<DCRTPoly> &ctxt_comp2, Ciphertext<DCRTPoly> &ctxt_out) {
        Ciphertext<DCRTPoly> aligned[4], bits[6], sorted[4];
        slotMatching4(ctxt, ctxt_comp1, ctxt_comp2, indices, shift, aligned, bits);
        fourSorter(aligned, bits, sorted);
        slotAssemble(sorted, 4, shift, ctxt_out);
    }
    void runFiveSorter(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> &ctxt_comp1,
                       Ciphertext<DCRTPoly> &ctxt_comp2, Ciphertext<DCRTPoly> &ctxt_out) {
        std::vector<double> first;
        genMask(indices, 5, 1, first);
        auto firstPlain = m_cc->MakeCKKSPackedPlaintext(first);
        Ciphertext<DCRTPoly> aligned[5], bits[10], sorted[5];
        slotMatching5(ctxt, ctxt_comp1, ctxt_comp2, indices, shift, aligned, bits);
        fiveSorter(aligned, bits, sorted);
        for (auto &c : sorted) c = m_cc->EvalMult(c, firstPlain);
        slotAssemble(sorted, 5, shift, ctxt_out);
    }
    // mixed sizes: output position i is kept on the first slots of the sorters that have more than i members
    void run2345Sorter(Ciphertext<DCRTPoly> &ctxt, std::vector<std::vector<int>> &indices, long shift, Ciphertext<DCRTPoly> &ctxt_comp1,
                       Ciphertext<DCRTPoly> &ctxt_comp2, Ciphertext<DCRTPoly> &ctxt_out) {
        Ciphertext<DCRTPoly> aligned[5], bits[10], sorted[5];
        slotMatching2345(ctxt, ctxt_comp1, ctxt_comp2, indices, shift, aligned, bits);
        fiveSorter(aligned, bits, sorted);
        for (int i = 0; i < 5; i++) {
            std::vector<double> keep((size_t)m_numSlots, 0.0);
            for (long x = 0; x < m_numSlots; x++)
                if (indices[1][x] == 1 && indices[0][x] > std::max(i, 1)) keep[x] = 1.0;
            sorted[i] = m_cc->EvalMult(sorted[i], m_cc->MakeCKKSPackedPlaintext(keep));
        }
        slotAssemble(sorted, 5, shift, ctxt_out);
    }

    // M + M(M-1)/2 * ceil(k/2) stages; before each compare and each sorter the operands are bootstrapped if fewer levels
    // remain than the step needs (m_level, src/k-way/Sorter.h:85-93)
    void sorter(Ciphertext<DCRTPoly> &ctxt, Ciphertext<DCRTPoly> &ctxt_out, SignConfig &Cfg) {
        if (m_k != 2 && m_k != 3 && m_k != 5) throw std::invalid_argument("kwaySort::Sorter: k must be 2, 3 or 5");
        static const int need[6] = {0, 1, 3, 5, 6, 7};
        const int depth = Cfg.multDepth, k = (int)m_k, M = (int)m_M;
        const int stages = M + M * (M - 1) / 2 * ((k + 1) / 2);
        for (int stage = 0; stage < stages; stage++) {
            int m, logDist, slope;
            std::tie(m, logDist, slope) = sortType(k, M, stage);
            const long shift = getRotateDistance(k, logDist, slope);
            auto indices = genIndices(m_numSlots, k, M, m, logDist, slope);
            Ciphertext<DCRTPoly> bits, bits2, rest;
            if (k == 5) { // two comparison ciphertexts per stage (neighbours at distance 1 and 2)
                if (slope == 0) {
                    checkLevelAndBoot(ctxt, need[5], depth);
                    comparisonForSort2(ctxt, indices, logDist, slope, bits, bits2, rest, Cfg);
                    checkLevelAndBoot2(bits, bits2, need[5], depth);
                    runFiveSorter(ctxt, indices, shift, bits, bits2, ctxt);
                } else if (slope == 3) { // middle slope: 4-sorters
                    checkLevelAndBoot(ctxt, need[4], depth);
                    comparisonForSort2(ctxt, indices, logDist, slope, bits, bits2, rest, Cfg);
                    checkLevelAndBoot(ctxt, need[4], depth);
                    checkLevelAndBoot2(bits, bits2, need[4], depth);
                    runFourSorter(ctxt, indices, shift, bits, bits2, ctxt);
                    ctxt = m_cc->EvalAdd(ctxt, rest);
                } else if (slope == 1) { // diagonals of 2..5 members
                    checkLevelAndBoot(ctxt, need[5], depth);
                    comparisonForSort2(ctxt, indices, logDist, slope, bits, bits2, rest, Cfg);
                    checkLevelAndBoot2(bits, bits2, need[5], depth);
                    run2345Sorter(ctxt, indices, shift, bits, bits2, ctxt);
                    ctxt = m_cc->EvalAdd(ctxt, rest);
                } else { // slope 2: diagonals of 2 and 3 members
                    Ciphertext<DCRTPoly> twos, threes;
                    checkLevelAndBoot(ctxt, need[3], depth);
                    comparisonForSort(ctxt, indices, logDist, slope, bits, rest, Cfg);
                    checkLevelAndBoot(bits, need[2], depth);
                    runTwoSorter(ctxt, indices, shift, bits, twos);
                    checkLevelAndBoot(bits, need[3], depth);
                    runThreeSorter(ctxt, indices, shift, bits, threes);
                    ctxt = m_cc->EvalAdd(m_cc->EvalAdd(twos, rest), threes);
                }
            } else if (slope == 0) { // k-sorters
                checkLevelAndBoot(ctxt, need[k], depth);
                comparisonForSort(ctxt, indices, logDist, slope, bits, rest, Cfg);
                checkLevelAndBoot(bits, need[k], depth);
                if (k == 2) runTwoSorter(ctxt, indices, shift, bits, ctxt);
                else runThreeSorter(ctxt, indices, shift, bits, ctxt);
            } else if (k == 3 && slope == 2) { // middle slope: 2-sorters, the rest passes through
                checkLevelAndBoot(ctxt, need[2], depth);
                comparisonForSort(ctxt, indices, logDist, slope, bits, rest, Cfg);
                checkLevelAndBoot(bits, need[2], depth);
                runTwoSorter(ctxt, indices, shift, bits, ctxt);
                ctxt = m_cc->EvalAdd(ctxt, rest);
            } else if (k == 3) { // slope 1: diagonals of length 2 and 3 in one stage
                Ciphertext<DCRTPoly> twos, threes;
                checkLevelAndBoot(ctxt, need[3], depth);
                comparisonForSort(ctxt, indices, logDist, slope, bits, rest, Cfg);
                checkLevelAndBoot(bits, need[2], depth);
                runTwoSorter(ctxt, indices, shift, bits, twos);
                checkLevelAndBoot(bits, need[3], depth);
                runThreeSorter(ctxt, indices, shift, bits, threes);
                ctxt = m_cc->EvalAdd(m_cc->EvalAdd(twos, rest), threes);
            } else { // k == 2, slope 1
                Ciphertext<DCRTPoly> twos;
                checkLevelAndBoot(ctxt, need[2], depth);
                comparisonForSort(ctxt, indices, logDist, slope, bits, rest, Cfg);
                checkLevelAndBoot(bits, need[2], depth);
                runTwoSorter(ctxt, indices, shift, bits, twos);
                ctxt = m_cc->EvalAdd(twos, rest);
            }
        }
        ctxt_out = ctxt;
    }

  protected:
    Comparison comp;
};

} // namespace kwaySort
