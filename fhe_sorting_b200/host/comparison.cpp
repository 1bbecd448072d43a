// comparison.cpp — see comparison.h (reference: src/comparison.cpp:4-40).
#include "comparison.h"

Ciphertext<DCRTPoly> Comparison::compare(const CryptoContext<DCRTPoly> &cc, const Ciphertext<DCRTPoly> &a,
                                         const Ciphertext<DCRTPoly> &b, SignFunc SignFunc, SignConfig &Cfg) {
    auto s = sign(cc->EvalSub(a, b), cc, SignFunc, Cfg);
    return cc->EvalMult(cc->EvalAdd(s, 1), 0.5);
}

Ciphertext<DCRTPoly> Comparison::indicator(const CryptoContext<DCRTPoly> &cc, const Ciphertext<DCRTPoly> &x, const double c,
                                           SignFunc SignFunc, SignConfig &Cfg) {
    auto above = cc->EvalMult(cc->EvalAdd(sign(cc->EvalAdd(x, c), cc, SignFunc, Cfg), 1.0), 0.5);
    auto below = cc->EvalMult(cc->EvalAdd(sign(cc->EvalSub(x, c), cc, SignFunc, Cfg), 1.0), 0.5);
    return cc->EvalMultAndRelinearize(above, cc->EvalSub(1.0, below));
}
