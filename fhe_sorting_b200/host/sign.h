// sign.h — homomorphic sign approximations used by the comparisons of the sorts.
// Interface of the reference's src/sign.h:6-45; evaluation schedules restated in sign.cpp.
#pragma once
#include "encryption.h"
#include "openfhe.h"

enum class SignFunc { CompositeSign, SignumPolycircuit, Tanh, NaiveDiscrete };

struct CompositeSignConfig {
    int n, dg, df; // polynomial family (3 or 4), number of g_n and f_n compositions
    CompositeSignConfig(int n_, int dg_, int df_) : n(n_), dg(dg_), df(df_) {}
};

struct SignConfig {
    CompositeSignConfig compos;
    int multDepth;
    SignConfig() : compos(0, 0, 0), multDepth(100) {}
    SignConfig(CompositeSignConfig c) : compos(c), multDepth(100) {}            // no bootstrapping
    SignConfig(CompositeSignConfig c, int depth) : compos(c), multDepth(depth) {} // bootstrap when levels run out
};

template <int n>
Ciphertext<DCRTPoly> compositeSign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, const SignConfig &Cfg);
Ciphertext<DCRTPoly> sign(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc, SignFunc func, const SignConfig &Cfg);
Ciphertext<DCRTPoly> signum_polycircuit(Ciphertext<DCRTPoly> x, CryptoContext<DCRTPoly> cc);
Ciphertext<DCRTPoly> naive_discrete_sign(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc);
Ciphertext<DCRTPoly> tanh(Ciphertext<DCRTPoly> a, CryptoContext<DCRTPoly> cc);
