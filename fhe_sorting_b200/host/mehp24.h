// mehp24.h — ranking-based sort of Mazzone, Everts, Hahn, Peter ("Efficient Ranking, Order Statistics, and Sorting under
// CKKS", arXiv 2412.15126) as the reference drives it: src/mehp24/mehp24_sort.{h,cpp} (sortFG for one and for several
// ciphertexts, sortLargeArrayFG) and src/mehp24/mehp24_utils.{h,cpp} (matrix replicate / sum / transpose folds, signAdv,
// indicatorAdv, split / combine).  Own restatement against the lbcrypto shim: the vector of length m lives in the first
// row of an m x m matrix packed row-major in the slots; every fold is log2(m) rotate-and-add steps.
//
// Same names, arguments and level consumption as the reference (namespace mehp24 / mehp24::utils); the OpenMP pragmas of
// the reference are dropped (one CUDA stream executes the calls in order) and the two sortFG overloads share one body.
#pragma once

#include <cmath>
#include <memory>
#include <stdexcept>
#include <vector>

#include "comparison.h"
#include "openfhe.h"
#include "sign.h"

namespace mehp24 {

using namespace lbcrypto;
typedef Ciphertext<DCRTPoly> CT;

namespace utils {

inline size_t log2ceil(size_t x) {
    size_t l = 0;
    while ((size_t(1) << l) < x) l++;
    return l;
}

// c + (c rotated by each of `steps` in turn, accumulating): the doubling fold behind every replicate / sum below
inline CT fold(CT c, const std::vector<int32_t> &steps) {
    auto cc = c->GetCryptoContext();
    for (int32_t s : steps) c = cc->EvalAdd(c, cc->EvalRotate(c, s));
    return c;
}
inline CT timesMask(const CT &c, const std::vector<double> &mask) {
    auto cc = c->GetCryptoContext();
    return cc->EvalMult(c, cc->MakeCKKSPackedPlaintext(mask));
}

// reference: mehp24_utils.cpp:22-44
inline CT maskRow(const CT &c, size_t m, size_t row) {
    if (row >= m) throw std::invalid_argument("maskRow: invalid row index");
    std::vector<double> mask(m * m, 0.0);
    for (size_t i = 0; i < m; i++) mask[m * row + i] = 1.0;
    return timesMask(c, mask);
}
inline CT maskColumn(const CT &c, size_t m, size_t col) {
    if (col >= m) throw std::invalid_argument("maskColumn: invalid column index");
    std::vector<double> mask(m * m, 0.0);
    for (size_t i = 0; i < m; i++) mask[m * i + col] = 1.0;
    return timesMask(c, mask);
}
// reference: mehp24_utils.cpp:46-62 — rotations to the right by m, 2m, 4m, ... (rows) / 1, 2, 4, ... (columns)
inline CT replicateRow(const CT &c, size_t m) {
    std::vector<int32_t> st;
    for (size_t i = 0; i < log2ceil(m); i++) st.push_back(-(int32_t)(1u << (log2ceil(m) + i)));
    return fold(c, st);
}
inline CT replicateColumn(const CT &c, size_t m) {
    std::vector<int32_t> st;
    for (size_t i = 0; i < log2ceil(m); i++) st.push_back(-(int32_t)(1u << i));
    return fold(c, st);
}
// reference: mehp24_utils.cpp:64-85
inline CT sumRows(const CT &c, size_t m, bool maskOutput = false, size_t outputRow = 0) {
    CT r = replicateRow(c, m);
    return maskOutput ? maskRow(r, m, outputRow) : r;
}
inline CT sumColumns(const CT &c, size_t m, bool maskOutput = false) {
    std::vector<int32_t> st;
    for (size_t i = 0; i < log2ceil(m); i++) st.push_back((int32_t)(1u << i));
    CT r = fold(c, st);
    return maskOutput ? maskColumn(r, m, 0) : r;
}
// reference: mehp24_utils.cpp:87-109 — first row <-> first column by steps of m (m - 1) / 2^i
inline CT transposeRow(const CT &c, size_t m, bool maskOutput = false) {
    std::vector<int32_t> st;
    for (size_t i = 1; i <= log2ceil(m); i++) st.push_back(-(int32_t)(m * (m - 1) / (size_t(1) << i)));
    CT r = fold(c, st);
    return maskOutput ? maskColumn(r, m, 0) : r;
}
inline CT transposeColumn(const CT &c, size_t m, bool maskOutput = false) {
    std::vector<int32_t> st;
    for (size_t i = 1; i <= log2ceil(m); i++) st.push_back((int32_t)(m * (m - 1) / (size_t(1) << i)));
    CT r = fold(c, st);
    return maskOutput ? maskRow(r, m, 0) : r;
}

// reference: mehp24_utils.cpp:255-271 — dg times g3, df - 1 times f3, then (f3 + 1) / 2, each by EvalPolyLinear
inline CT signAdv(CT c, size_t dg, size_t df) {
    static const std::vector<double> f3 = {0, 35.0 / 16, 0, -35.0 / 16, 0, 21.0 / 16, 0, -5.0 / 16};
    static const std::vector<double> f3_last = {0.5, 35.0 / 32, 0, -35.0 / 32, 0, 21.0 / 32, 0, -5.0 / 32};
    static const std::vector<double> g3 = {0, 4589.0 / 1024, 0, -16577.0 / 1024, 0, 25614.0 / 1024, 0, -12860.0 / 1024};
    auto cc = c->GetCryptoContext();
    for (size_t d = 0; d < dg; d++) c = cc->EvalPolyLinear(c, g3);
    for (size_t d = 0; d + 1 < df; d++) c = cc->EvalPolyLinear(c, f3);
    return cc->EvalPolyLinear(c, f3_last);
}
// reference: mehp24_utils.cpp:138-143
inline CT compareAdv(const CT &a, const CT &b, size_t dg, size_t df) {
    return signAdv(a->GetCryptoContext()->EvalSub(a, b), dg, df);
}
// reference: mehp24_utils.cpp:165-173 — 1 on (-1/2, 1/2), 0 elsewhere on (-b, b): step(x + 1/2) * (1 - step(x - 1/2))
inline CT indicatorAdv(const CT &c, double b, size_t dg, size_t df) {
    auto cc = c->GetCryptoContext();
    CT t = cc->EvalMult(c, 1.0 / b);
    CT up = signAdv(cc->EvalAdd(t, 0.5 / b), dg, df);
    CT dn = signAdv(cc->EvalSub(t, 0.5 / b), dg, df);
    return cc->EvalMult(up, cc->EvalSub(1.0, dn));
}
// reference: mehp24_utils.cpp:175-183
inline CT indicatorAdvShifted(const CT &c, double b, size_t dg, size_t df) {
    auto cc = c->GetCryptoContext();
    const double k = 2.0 / (b + 1);
    CT up = signAdv(cc->EvalAdd(cc->EvalMult(c, k), k - 1.0), dg, df);
    CT dn = signAdv(cc->EvalAdd(cc->EvalMult(c, -k), k + 1.0), dg, df);
    return cc->EvalMult(up, dn);
}

// Chebyshev-function variants of the original MEHP24 code (reference: mehp24_utils.cpp:111-163)
inline CT equal(const CT &a, const CT &b, double lo, double hi, uint32_t degree, double error = 0.00001) {
    auto cc = a->GetCryptoContext();
    return cc->EvalChebyshevFunction([error](double x) -> double { return (x > error || x < -error) ? 0.0 : 1.0; },
                                     cc->EvalSub(a, b), lo, hi, degree);
}
inline CT compare(const CT &a, const CT &b, double lo, double hi, uint32_t degree, double error = 0.00001) {
    auto cc = a->GetCryptoContext();
    return cc->EvalChebyshevFunction([error](double x) -> double { return x > error ? 1.0 : x >= -error ? 0.5 : 0.0; },
                                     cc->EvalSub(a, b), lo, hi, degree);
}
inline CT compareGt(const CT &a, const CT &b, double lo, double hi, uint32_t degree, double error = 0.00001) {
    auto cc = a->GetCryptoContext();
    return cc->EvalChebyshevFunction([error](double x) -> double { return x > error ? 1.0 : 0.0; }, cc->EvalSub(a, b), lo,
                                     hi, degree);
}
inline CT indicator(const CT &c, double a1, double b1, double lo, double hi, uint32_t degree) {
    return c->GetCryptoContext()->EvalChebyshevFunction(
        [a1, b1](double x) -> double { return (x < a1 || x > b1) ? 0.0 : 1.0; }, c, lo, hi, degree);
}

// reference: mehp24_utils.cpp:185-214 — every index the folds above use for side length m (capped at 256 per ciphertext),
// plus multiples of 256 for split / combine of longer arrays
// (sideMax: values per ciphertext; 256 in the reference, less on small test rings)
inline std::vector<int32_t> getRotationIndices(size_t m, size_t sideMax = 256) {
    std::vector<int32_t> idx;
    size_t side = m;
    if (m > sideMax) {
        for (size_t i = 1; i < m / sideMax; i++) idx.push_back((int32_t)(i * sideMax)), idx.push_back(-(int32_t)(i * sideMax));
        side = sideMax;
    }
    const size_t lg = log2ceil(side);
    for (size_t i = 0; i < lg; i++) {
        idx.push_back((int32_t)(1u << i)), idx.push_back(-(int32_t)(1u << i));
        idx.push_back(-(int32_t)(1u << (lg + i)));
        const int32_t t = (int32_t)(side * (side - 1) / (size_t(1) << (i + 1)));
        idx.push_back(t), idx.push_back(-t);
    }
    return idx;
}
// reference: mehp24_utils.cpp:216-247
inline unsigned depth2degree(unsigned depth) {
    static const int table[] = {2, 5, 13, 27, 59, 119, 247, 495, 1007, 2031, 4031, 8127};
    return depth >= 3 && depth <= 14 ? (unsigned)table[depth - 3] : (unsigned)-1;
}

// reference: mehp24_utils.cpp:273-303
inline std::vector<CT> splitCiphertext(const CT &c, size_t totalLength, size_t subLength, CryptoContext<DCRTPoly> &cc) {
    std::vector<CT> parts(totalLength / subLength);
    for (size_t i = 0; i < parts.size(); i++) {
        std::vector<double> mask(totalLength, 0.0);
        for (size_t j = 0; j < subLength; j++) mask[i * subLength + j] = 1.0;
        CT part = cc->EvalMult(c, cc->MakeCKKSPackedPlaintext(mask));
        parts[i] = i ? cc->EvalRotate(part, (int32_t)(i * subLength)) : part;
    }
    return parts;
}
inline CT combineCiphertext(const std::vector<CT> &parts, size_t subLength, CryptoContext<DCRTPoly> &cc) {
    CT r = parts.at(0);
    for (size_t i = 1; i < parts.size(); i++) r = cc->EvalAdd(r, cc->EvalRotate(parts[i], -(int32_t)(i * subLength)));
    return r;
}

} // namespace utils

// How the m x m comparison matrix is produced: MEHP24's own f/g composition, or the reference's Comparison object with a
// SignConfig (the variant tests/mehp24/Mehp24SortTest.cpp:133-141 runs)
struct Comparator {
    uint32_t dg = 0, df = 0;
    Comparison *comp = nullptr;
    SignFunc func = SignFunc::CompositeSign;
    SignConfig *cfg = nullptr;
    CryptoContext<DCRTPoly> cc;
    int rank = 0, world = 1; // multi-GPU: this process evaluates the units u of the several-ciphertext sort with u % world == rank
    CT operator()(const CT &a, const CT &b) const {
        return comp ? comp->compare(cc, a, b, func, *cfg) : utils::compareAdv(a, b, dg, df);
    }
};

// -1 * (offset + row) - 1/2 in every slot of the row: R + this is zero-centred exactly where rank == row
inline std::vector<double> rankOffsets(size_t m, size_t offset) {
    std::vector<double> v(m * m);
    for (size_t i = 0; i < m; i++)
        for (size_t j = 0; j < m; j++) v[i * m + j] = -1.0 * (double)(offset + i) - 0.5;
    return v;
}

// reference: mehp24_sort.cpp:213-277 (both overloads) — one ciphertext, m = vectorLength <= sqrt(slots)
inline CT sortFGWith(const CT &c, size_t m, const Comparator &cmp, uint32_t dg_i, uint32_t df_i) {
    using namespace utils;
    auto cc = c->GetCryptoContext();
    CT VR = replicateRow(c, m);                                    // v_j in column j of every row
    CT VC = replicateColumn(transposeRow(c, m, true), m);          // v_i in row i of every column
    CT C = cmp(VR, VC);                                            // C[i][j] = (v_j > v_i)
    CT R = sumRows(C, m);                                          // rank of v_j in column j of every row
    CT M = indicatorAdv(cc->EvalAdd(R, cc->MakeCKKSPackedPlaintext(rankOffsets(m, 0))), (double)m, dg_i, df_i);
    CT S = sumColumns(cc->EvalMult(M, VR), m, true);               // row i, column 0: the value of rank i
    return transposeColumn(S, m, true);
}
inline CT sortFG(const CT &c, size_t vectorLength, uint32_t dg_c, uint32_t df_c, uint32_t dg_i, uint32_t df_i,
                 CryptoContext<DCRTPoly> cc) {
    Comparator cmp;
    cmp.dg = dg_c, cmp.df = df_c, cmp.cc = cc;
    return sortFGWith(c, vectorLength, cmp, dg_i, df_i);
}
inline CT sortFG(const CT &c, size_t vectorLength, SignFunc func, SignConfig &cfg, std::unique_ptr<Comparison> &comp,
                 uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> cc) {
    Comparator cmp;
    cmp.comp = comp.get(), cmp.func = func, cmp.cfg = &cfg, cmp.cc = cc;
    return sortFGWith(c, vectorLength, cmp, dg_i, df_i);
}

// reference: mehp24_sort.cpp:279-575 — the vector is split over several ciphertexts of m = subVectorLength values each.
// Block (j, k) of the comparison matrix compares part j (as a row) with part k (as a column); only j <= k is evaluated,
// the mirror block is 1 - C.  `owner(unit)` says whether this rank evaluates a unit (multi-GPU sharding, all units by
// default); partial sums over units are combined by the caller.
inline std::vector<CT> sortFGWith(const std::vector<CT> &c, size_t m, const Comparator &cmp, uint32_t dg_i, uint32_t df_i) {
    using namespace utils;
    const size_t parts = c.size(), total = m * parts;
    auto cc = c.at(0)->GetCryptoContext();
    auto accumulate = [&](CT &acc, const CT &x) { acc = acc ? cc->EvalAdd(acc, x) : x; };
    std::vector<CT> replR(parts), replC(parts);
    for (size_t j = 0; j < parts; j++) {
        replR[j] = replicateRow(c[j], m);
        replC[j] = replicateColumn(transposeRow(c[j], m, true), m);
    }
    // multi-GPU ("MEHP24 N = 1024"): the parts (parts + 1) / 2 comparison blocks and the parts^2 indicator
    // products are independent; rank r evaluates unit u when u % world == r, a rank without a contribution to some sum puts
    // in a zero of the same level and scale, and the sums are all-reduced (partial Cv / Ch after phase 1, partial `sorted`
    // after phase 2; the folds in between are cheap and run replicated)
    const bool sharded = cmp.world > 1;
    auto mine = [&](size_t unit) { return !sharded || (int)(unit % (size_t)cmp.world) == cmp.rank; };
    auto reduce = [&](std::vector<CT> &v, size_t first, const CT &sample) {
        if (!sharded) return;
        if (!sample) throw std::runtime_error("mehp24: more ranks than independent units");
        CT zero = cc->EvalSub(sample, sample);
        for (size_t j = first; j < v.size(); j++) {
            if (!v[j]) v[j] = zero->Clone();
            cc->AllReduceSum(v[j]);
        }
    };
    std::vector<CT> Cv(parts), Ch(parts);
    {
        size_t unit = 0;
        CT sample;
        for (size_t j = 0; j < parts; j++)
            for (size_t k = j; k < parts; k++) {
                if (!mine(unit++)) continue;
                CT Cjk = cmp(replR[j], replC[k]);
                sample = Cjk;
                accumulate(Cv[j], Cjk);
                if (j != k) accumulate(Ch[k], cc->EvalSub(1.0, Cjk));
            }
        reduce(Cv, 0, sample);
        reduce(Ch, 1, sample);
    }
    // s[j]: rank of every element of part j among all parts, replicated over the rows
    std::vector<CT> s(parts);
    for (size_t j = 0; j < parts; j++) {
        accumulate(s[j], sumRows(Cv[j], m));
        if (j > 0) accumulate(s[j], replicateRow(transposeColumn(sumColumns(Ch[j], m, true), m, true), m));
    }
    // s[0] has no column-sum term and sits one mask multiplication higher than the others: bring it to their level, so that every
    // indicator product - whichever rank evaluates it - has the same level and scale and the partial sums can be all-reduced
    if (sharded && parts > 1) s[0] = cc->EvalAdd(s[0], cc->EvalSub(s[1], s[1]));
    std::vector<CT> sorted(parts);
    {
        size_t unit = 0;
        CT sample;
        for (size_t j = 0; j < parts; j++) {
            Plaintext offs = cc->MakeCKKSPackedPlaintext(rankOffsets(m, j * m));
            for (size_t k = 0; k < parts; k++) {
                if (!mine(unit++)) continue;
                CT term = cc->EvalMult(indicatorAdv(cc->EvalAdd(s[k], offs), (double)total, dg_i, df_i), replR[k]);
                sample = term;
                accumulate(sorted[j], term);
            }
        }
        reduce(sorted, 0, sample);
    }
    for (size_t j = 0; j < parts; j++) sorted[j] = transposeColumn(sumColumns(sorted[j], m, true), m, true);
    return sorted;
}
inline std::vector<CT> sortFG(const std::vector<CT> &c, size_t subVectorLength, uint32_t dg_c, uint32_t df_c, uint32_t dg_i,
                              uint32_t df_i) {
    Comparator cmp;
    cmp.dg = dg_c, cmp.df = df_c, cmp.cc = c.at(0)->GetCryptoContext();
    return sortFGWith(c, subVectorLength, cmp, dg_i, df_i);
}
inline std::vector<CT> sortFG(const std::vector<CT> &c, size_t subVectorLength, SignFunc func, SignConfig &cfg,
                              std::unique_ptr<Comparison> &comp, uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> cc,
                              int rank = 0, int world = 1) {
    Comparator cmp;
    cmp.comp = comp.get(), cmp.func = func, cmp.cfg = &cfg, cmp.cc = cc, cmp.rank = rank, cmp.world = world;
    return sortFGWith(c, subVectorLength, cmp, dg_i, df_i);
}

// reference: mehp24_sort.cpp:577-645
inline CT sortLargeArrayFG(const CT &c, size_t totalLength, size_t subLength, uint32_t dg_c, uint32_t df_c, uint32_t dg_i,
                           uint32_t df_i, CryptoContext<DCRTPoly> cc) {
    return utils::combineCiphertext(sortFG(utils::splitCiphertext(c, totalLength, subLength, cc), subLength, dg_c, df_c, dg_i, df_i),
                                    subLength, cc);
}
inline CT sortLargeArrayFG(const CT &c, size_t totalLength, size_t subLength, SignFunc func, SignConfig &cfg,
                           std::unique_ptr<Comparison> &comp, uint32_t dg_i, uint32_t df_i, CryptoContext<DCRTPoly> cc, int rank = 0,
                           int world = 1) {
    return utils::combineCiphertext(
        sortFG(utils::splitCiphertext(c, totalLength, subLength, cc), subLength, func, cfg, comp, dg_i, df_i, cc, rank, world),
        subLength, cc);
}

// reference: mehp24_sort.cpp:23-47 — Chebyshev-function variant (degrees instead of f/g compositions)
inline CT sort(const CT &c, size_t m, double leftBoundC, double rightBoundC, uint32_t degreeC, uint32_t degreeI) {
    using namespace utils;
    auto cc = c->GetCryptoContext();
    CT VR = replicateRow(c, m);
    CT VC = replicateColumn(transposeRow(c, m, true), m);
    CT R = sumRows(compare(VR, VC, leftBoundC, rightBoundC, degreeC), m);
    CT M = indicator(cc->EvalAdd(R, cc->MakeCKKSPackedPlaintext(rankOffsets(m, 0))), -0.5, 0.5, -1.0 * (double)m,
                     1.0 * (double)m, degreeI);
    return sumColumns(cc->EvalMult(M, VR), m);
}

// multiplicative depth the reference's test fixture picks per length (tests/mehp24/Mehp24SortTest.cpp:31-62)
inline int multDepthFor(size_t N) {
    switch (N) {
    case 4: return 31;
    case 8: case 16: return 35;
    case 32: case 64: return 42;
    case 128: return 46;
    case 256: return 49;
    case 512: return 57;
    case 1024: return 60;
    case 2048: return 64;
    default: throw std::invalid_argument("mehp24: array length must be a power of two in 4..2048");
    }
}

} // namespace mehp24
