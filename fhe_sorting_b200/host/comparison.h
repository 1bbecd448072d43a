// comparison.h — encrypted comparison and equality indicator built on the sign approximations, plus the plain-double
// sinc targets the Chebyshev fits are made from.  Interface of the reference's src/comparison.h:11-100.
#pragma once
#include <cmath>
#include <memory>

#include "encryption.h"
#include "openfhe.h"
#include "sign.h"

template <int N> struct Sinc {
    static double simple_sinc(double x) { return std::abs(x) < 0.5 ? 1.0 : 0.0; }
    static double sinc(double x) { return std::abs(x) < 1e-10 ? 1.0 : std::sin(M_PI * x) / M_PI * x; }
    // sin(pi N x) / (pi N x): one-hot at x = 0 on the grid k / N
    static double scaled_sinc(double x) {
        const double t = M_PI * N * x;
        return std::abs(x) < 1e-10 ? 1.0 : std::sin(t) / t;
    }
    static double scaled_sinc_j(double x, int j) {
        const double d1 = N * M_PI * x - j * M_PI, d2 = d1 + N * M_PI;
        return (std::abs(d1) < 1e-10 ? 1.0 : std::sin(d1) / d1) + (std::abs(d2) < 1e-10 ? 1.0 : std::sin(d2) / d2);
    }
    // scaled_sinc(x) + scaled_sinc(x + 1/2): one-hot at x = 0 and x = -1/2 (rank offsets 0 and -N after scaling by 1/2N)
    static double doubled_sinc(double x) { return scaled_sinc(x) + scaled_sinc(x + 0.5); }
};

class Comparison {
    std::shared_ptr<Encryption> m_enc;

  public:
    Comparison(std::shared_ptr<Encryption> enc) : m_enc(enc) {}
    Comparison() : m_enc(nullptr) {}
    // (sign(a - b) + 1) / 2 : 1 where a > b, 0 where a < b, 1/2 where equal
    Ciphertext<DCRTPoly> compare(const CryptoContext<DCRTPoly> &cc, const Ciphertext<DCRTPoly> &a, const Ciphertext<DCRTPoly> &b,
                                 SignFunc SignFunc, SignConfig &Cfg);
    // 1 where |x| < c, 0 elsewhere (MEHP24 style): cmp(x + c) * (1 - cmp(x - c))
    Ciphertext<DCRTPoly> indicator(const CryptoContext<DCRTPoly> &cc, const Ciphertext<DCRTPoly> &x, const double c,
                                   SignFunc SignFunc, SignConfig &Cfg);
};
