// bootstrap_plan.h — host-side plan of one CKKS bootstrapping (no CUDA in this header): the grouped FFT-like factors
// of the homomorphic encoding / decoding transforms as generalised diagonals, the Chebyshev series of the scaled
// cosine and the constants that tie them to the modulus chain.
//
// Reference call sites: EvalBootstrapSetup / EvalBootstrapKeyGen (tests/k-way/KWaySort2Test.cpp:42-44),
// EvalBootstrap (src/k-way/EvalUtils.cpp:76, src/sign.cpp:168, src/sort_algo.h:1111).  OpenFHE is absent, so this
// restates the published algorithm: Cheon-Han-Kim-Kim-Song 2018 (ModRaise, SubSum, CoeffsToSlots, EvalMod,
// SlotsToCoeffs), Chen-Chillotti-Song 2019 (level-collapsed FFT factors) and Han-Ki 2020 (cosine + double angle),
// for sparsely packed ciphertexts (2 * slots <= ring_dim / 2) and uniform ternary secrets.
#pragma once
#include <complex>
#include <vector>

namespace fheb200 {

struct BootStage {                                    // out = sum_i diag[i] (.) rotate(in, rot[i]) on 2 * slots slots
    std::vector<int> rot;                             // rotation indices in [0, slots)
    std::vector<std::vector<std::complex<double>>> diag; // [rot.size()][2 * slots]
};

struct BootPlan {
    int n = 0, slots = 0;
    int K = 512;          // |I| < K for the overflow polynomial of a uniform ternary secret at ring_dim <= 2^17
    int r = 6;            // double-angle iterations
    int degree = 44;      // Chebyshev degree IN y = 2u^2 - 1 of the even function cos(w u), w = 2 pi K (1 + shift') / 2^r
    double shift = 0;     // u = slot value - shift: the -1/4 of cos(2 pi (K x - 1/4) / 2^r) moved into the argument so
                          // that the function is even in u and only every other Chebyshev term has to be evaluated
    int correction = 10;  // the message is brought to q0 * 2^-correction before ModRaise
    double q0 = 0;        // first modulus
    double msg_scale = 0; // q0 * 2^-correction
    double raise_scale = 0; // scale label the raised ciphertext carries (sf of the level it is raised to)
    std::vector<BootStage> cts, stc;
    std::vector<double> cheb; // series convention c_0 / 2 + sum c_i T_i
    std::vector<int> rotations() const; // every rotation index the evaluation needs (SubSum, stages, final fold)
    bool even = true;     // FHEB200_BOOT_EVEN=0: the plain degree-88 series in u (one bit more precise, 6 ms slower)
    int cheb_depth() const { int d = 0; while ((1 << d) < degree + 1) d++; return d + (even ? 1 : 0); } // + the square for y
    int depth() const { return (int)cts.size() + cheb_depth() + r + (int)stc.size(); }
};

// b_enc / b_dec = level budgets of CoeffsToSlots / SlotsToCoeffs (clamped to [1, log2 slots])
BootPlan make_boot_plan(int n, int slots, int b_enc, int b_dec, double q0, double raise_scale, int correction);

} // namespace fheb200
