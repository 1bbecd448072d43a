// sign.cpp — composite sign f_n^(df) o g_n^(dg) of Cheon, Kim, Kim (eprint 2019/1234) and the Chebyshev fall-backs.
// Restates the evaluation schedule of the reference (src/sign.cpp:15-185, 610-651) so that level consumption is the
// same: each g_3 / f_3 costs 3 levels (2 squarings, 3 relinearised products, 4 scalar products), g_4 / f_4 cost 4.
#include "sign.h"
#include "signum_polycircuit_table.h"

#include <cstdlib>

namespace {

typedef Ciphertext<DCRTPoly> CT;

// odd degree-7 polynomial c1 x + c3 x^3 + c5 x^5 + c7 x^7 in depth 3:
//   y = c1 x + (c3 x) x^2 + ((c5 x) + (c7 x) x^2) x^4       (reference: src/sign.cpp:23-35)
CT odd7(const CT &x, const CryptoContext<DCRTPoly> &cc, double c1, double c3, double c5, double c7) {
    CT x2 = cc->EvalSquareReduced(x); // the even powers are only ever multiplicands: squared and rescaled in one division
    CT x4 = cc->EvalSquareReduced(x2);
    CT y = cc->EvalMult(x, c1);
    cc->EvalAddInPlace(y, cc->EvalMultAndRelinearize(cc->EvalMult(x, c3), x2));
    CT hi = cc->EvalAdd(cc->EvalMult(x, c5), cc->EvalMultAndRelinearize(cc->EvalMult(x, c7), x2));
    cc->EvalAddInPlace(y, cc->EvalMultAndRelinearize(hi, x4));
    return y;
}

template <int n> struct Family;
template <> struct Family<3> {
    static constexpr int g_depth = 3, f_depth = 3;
    // g_3 = (4589 x - 16577 x^3 + 25614 x^5 - 12860 x^7) / 2^10 ; f_3 = (35 x - 35 x^3 + 21 x^5 - 5 x^7) / 2^4
    static CT g(const CT &x, const CryptoContext<DCRTPoly> &cc) {
        return odd7(x, cc, 4589.0 / 1024, -16577.0 / 1024, 25614.0 / 1024, -12860.0 / 1024);
    }
    static CT f(const CT &x, const CryptoContext<DCRTPoly> &cc) { return odd7(x, cc, 35.0 / 16, -35.0 / 16, 21.0 / 16, -5.0 / 16); }
};
template <> struct Family<4> {
    static constexpr int g_depth = 4, f_depth = 4;
    // g_4: degree-27 odd Chebyshev expansion, coefficient values from the reference (src/sign.cpp:66-74), evaluated
    // with Paterson-Stockmeyer like there
    static CT g(const CT &x, const CryptoContext<DCRTPoly> &cc) {
        static const double odd[14] = {1.077117252745569,   -0.36166113998402755, 0.2137420717859748,   -0.15635204788780485,
                                       0.11749645501187332, -0.10074154666447852, 0.08002086947825496,  -0.07533558758484624,
                                       0.059514472116534836, -0.06146663712787884, 0.04570084927999001, -0.05403683682999072,
                                       0.03364293851188723, -0.054459493266273494};
        std::vector<double> c(28, 0.0);
        for (int i = 0; i < 14; i++) c[2 * i + 1] = odd[i];
        return cc->EvalChebyshevSeriesPS(x, c, -1, 1);
    }
    // f_4 = sum_{i<=4} 4^-i C(2i,i) x (1-x^2)^i = (315 x - 420 x^3 + 378 x^5 - 180 x^7 + 35 x^9)/128 in the paper; the
    // reference evaluates a degree-15 variant (src/sign.cpp:80-87) whose coefficients are kept for identical depth:
    //   y = c1x + c3x x^2 + (c5x + c7x x^2) x^4 + ((c9x + c11x x^2) + (c13x + c15x x^2) x^4) x^8
    static CT f(const CT &x, const CryptoContext<DCRTPoly> &cc) {
        static const double c[8] = {3.14208984375,  -7.33154296875, 13.19677734375, -15.71044921875,
                                    12.21923828125, -5.99853515625, 1.69189453125,  -0.20947265625};
        CT x2 = cc->EvalSquareReduced(x);
        CT t[8];
        for (int i = 0; i < 8; i++) t[i] = cc->EvalMult(x, c[i]);
        CT x4 = cc->EvalSquareReduced(x2);
        for (int i = 1; i < 8; i += 2) t[i] = cc->EvalMultAndRelinearize(t[i], x2); // c3 x^3, c7 x^3, c11 x^3, c15 x^3
        CT x8 = cc->EvalSquareReduced(x4);
        CT y = cc->EvalAdd(t[0], t[1]);
        cc->EvalAddInPlace(y, cc->EvalMultAndRelinearize(cc->EvalAdd(t[2], t[3]), x4));
        CT hi = cc->EvalAdd(t[4], t[5]);
        cc->EvalAddInPlace(hi, cc->EvalMultAndRelinearize(cc->EvalAdd(t[6], t[7]), x4));
        cc->EvalAddInPlace(y, cc->EvalMultAndRelinearize(hi, x8));
        return y;
    }
};

} // namespace

// g_n is applied max(dg, 1) times, then f_n df times (reference: src/sign.cpp:160-185, the first g_n is unconditional);
// before each polynomial the ciphertext is bootstrapped if fewer than depth + 2 levels remain
template <int n> Ciphertext<DCRTPoly> compositeSign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, const SignConfig &Cfg) {
    auto refresh = [&](CT &c, int need) {
        if (Cfg.multDepth - (int)c->GetLevel() < need + 2) c = cc->EvalBootstrap(c);
    };
    refresh(x, Family<n>::g_depth);
    CT y = Family<n>::g(x, cc);
    for (int i = 1; i < Cfg.compos.dg; i++) {
        refresh(y, Family<n>::g_depth);
        y = Family<n>::g(y, cc);
    }
    for (int i = 0; i < Cfg.compos.df; i++) {
        refresh(y, Family<n>::f_depth);
        y = Family<n>::f(y, cc);
    }
    return y;
}
template Ciphertext<DCRTPoly> compositeSign<3>(Ciphertext<DCRTPoly>, CryptoContext<DCRTPoly>, const SignConfig &);
template Ciphertext<DCRTPoly> compositeSign<4>(Ciphertext<DCRTPoly>, CryptoContext<DCRTPoly>, const SignConfig &);

// degree-119 Chebyshev fit of the step function (reference: src/sign.cpp:610-624)
Ciphertext<DCRTPoly> naive_discrete_sign(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc) {
    return cc->EvalChebyshevFunction([](double x) -> double { return x >= 0 ? 1 : 0; }, a, -1, 1, 119);
}
// degree-1006 Chebyshev fit of tanh(RAND_MAX x) (reference: src/sign.cpp:626-633)
Ciphertext<DCRTPoly> tanh(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc) {
    return cc->EvalChebyshevFunction([](double x) -> double { return std::tanh(x * RAND_MAX); }, a, -1, 1, 1006);
}

// Degree-1023 odd Chebyshev series of the polycircuit library (reference: src/sign.cpp:188-608, which evaluates T_0..T_1008
// with EvalChebyshevSeries and adds T_1009..T_1023 through a hand-unrolled product tree).  Same polynomial here, in one
// Paterson-Stockmeyer evaluation over the whole coefficient vector: 10 levels.
Ciphertext<DCRTPoly> signum_polycircuit(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc) {
    static const std::vector<double> coeffs = [] {
        std::vector<double> c(1024, 0.0);
        for (int i = 0; i < 512; i++) c[2 * i + 1] = kSignumOddChebyshev[i];
        return c;
    }();
    return cc->EvalChebyshevSeries(x, coeffs, -1, 1);
}

Ciphertext<DCRTPoly> sign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, SignFunc func, const SignConfig &Cfg) {
    switch (func) {
    case SignFunc::Tanh: return tanh(x, cc);
    case SignFunc::NaiveDiscrete: return naive_discrete_sign(x, cc);
    case SignFunc::SignumPolycircuit: return signum_polycircuit(x, cc);
    case SignFunc::CompositeSign:
    default:
        if (Cfg.compos.n == 3) return compositeSign<3>(x, cc, Cfg);
        if (Cfg.compos.n == 4) return compositeSign<4>(x, cc, Cfg);
        throw std::invalid_argument("sign: CompositeSignConfig.n must be 3 or 4");
    }
}
