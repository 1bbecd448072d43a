// hybrid_params.h — parameters of DirectSort<N>::sort_hybrid (rank construction + MEHP24-style indicator placement,
// reference src/sort_algo.h:776-1064).  The reference keeps them in its test fixture (tests/DirectSortHTest.cpp:22-110:
// a multiplicative depth and a hand-written rotation-key list per N); here the depth table is that fixture's and the
// extra rotation indices are derived from the schedule of rotationIndexCheckHybrid.
#pragma once
#include <algorithm>
#include <set>
#include <vector>

inline int DirectSortHybridDepth(int N) { // tests/DirectSortHTest.cpp:29-101
    switch (N) {
    case 4: return 24;
    case 8: case 16: return 25;
    case 32: return 29;
    case 64: return 30;
    case 128: return 31;
    case 256: return 44;
    case 512: return 47;
    default: return 50;
    }
}
// signed rotation steps of the placement: the block shifts b * 256, the column fold towards column b (steps m/2 .. 1, right
// where b's bit is set) and the move of column b onto row b (steps m(m-1)/2, halved each time)
inline std::vector<int> DirectSortHybridExtraRotations(int N) {
    const int m = std::min(N, 256), batches = N / m;
    int logm = 0;
    while ((1 << logm) < m) logm++;
    std::set<int> r;
    for (int b = 0; b < batches; b++) {
        if (b) r.insert(b * 256);
        int fold = m / 2, move = m * (m - 1) / 2;
        for (int i = 0; i < logm; i++, fold >>= 1, move >>= 1) {
            const bool right = (b >> (logm - 1 - i)) & 1;
            if (fold) r.insert(right ? -fold : fold);
            if (move) r.insert(right ? -move : move);
        }
    }
    return std::vector<int>(r.begin(), r.end());
}
