// generated_coeffs.h — Chebyshev coefficient tables of the sinc one-hot functions used by DirectSort.
// The reference generates these at build time (utils/generate_cheb_coeffs.cpp:11-64 -> generated_coeffs.h,
// utils/generate_cheb_doubled_coeffs.cpp:11-49 -> generated_doubled_sinc_coeffs.h): a degree-13011 Chebyshev fit on
// [-1,1] of Sinc<2N>::scaled_sinc (even coefficients >= 1e-6 kept, odd zeroed) and of Sinc<2N>::doubled_sinc
// (coefficients >= 1e-8 kept), trailing zeros trimmed, and every value written with the default ostream precision,
// i.e. rounded to 6 significant digits.  Here the same tables are produced on first use (same fit, same thresholds,
// same 6-digit rounding) and cached per N.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <vector>

#include "comparison.h"
#include "openfhe.h"

namespace sinc_tables {
constexpr int kFitDegree = 13011;
inline double six_digits(double v) { // what `ostream << double` followed by a compiler's strtod gives back
    char buf[64];
    snprintf(buf, sizeof buf, "%g", v);
    return strtod(buf, nullptr);
}
inline std::vector<double> finish(std::vector<double> c, double trim) {
    while (!c.empty() && std::abs(c.back()) < trim) c.pop_back();
    for (double &v : c) v = six_digits(v);
    return c;
}
template <int N> std::vector<double> scaled() {
    auto c = lbcrypto::EvalChebyshevCoefficients([](double x) { return Sinc<2 * N>::scaled_sinc(x); }, -1.0, 1.0, kFitDegree);
    for (size_t i = 0; i < c.size(); i++)
        if (i % 2 == 1 || std::abs(c[i]) < 1e-6) c[i] = 0.0;
    return finish(std::move(c), 1e-15);
}
template <int N> std::vector<double> doubled() {
    auto c = lbcrypto::EvalChebyshevCoefficients([](double x) { return Sinc<2 * N>::doubled_sinc(x); }, -1.0, 1.0, kFitDegree);
    for (double &v : c)
        if (std::abs(v) < 1e-8) v = 0.0;
    return finish(std::move(c), 1e-8);
}
} // namespace sinc_tables

template <std::size_t N> const std::vector<double> &selectCoefficients() {
    static const std::vector<double> table = sinc_tables::scaled<(int)N>();
    return table;
}
template <std::size_t N> const std::vector<double> &selectDoubledSincCoefficients() {
    static const std::vector<double> table = sinc_tables::doubled<(int)N>();
    return table;
}
