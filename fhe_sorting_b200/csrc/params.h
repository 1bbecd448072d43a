// params.h — host-side RNS-CKKS parameter set of the B200 engine (no CUDA in this header).
//
// Mirrors what the reference obtains from OpenFHE's GenCryptoContext(CCParams<CryptoContextCKKSRNS>)
// with only SetMultiplicativeDepth / SetScalingModSize / SetFirstModSize / SetBatchSize / SetRingDim /
// SetSecurityLevel touched (reference: tests/DirectSortTest.cpp:24-51, src/sort_algo.h:87-201,
// src/kway_adapter.h:41-64), i.e. OpenFHE defaults: FLEXIBLEAUTOEXT rescaling, HYBRID key switching with
// 3 digits, uniform ternary secrets, sigma = 3.19, 64-bit native moduli.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

namespace fheb200 {

typedef uint64_t u64;
typedef unsigned __int128 u128;
typedef __int128 i128;

// Per-modulus constants consumed by the device arithmetic (modarith.cuh).
struct ModConst {
    u64 q;        // modulus, q < 2^61
    u64 mu;       // floor(2^(bits+63) / q)  : Barrett constant for products of two reduced operands
    u64 r64;      // floor(2^64 / q)         : single-word Barrett constant
    u64 c64;      // 2^64 mod q
    u64 c64s;     // Shoup companion of c64: floor(c64 * 2^64 / q)
    uint32_t bits; // bit length of q
    uint32_t pad;
};

// generator state handed to the sampling kernels by value (modarith.cuh "randomness")
struct RngState {
    u64 seed = 0;            // secure == 0: 64-bit seed of the test-only deterministic generator
    uint32_t key[8] = {0};   // secure == 1: 256-bit ChaCha20 key
    int secure = 0;
    int pad = 0;
};

u64 mulmod_h(u64 a, u64 b, u64 q);
u64 powmod_h(u64 a, u64 e, u64 q);
u64 invmod_h(u64 a, u64 q);
inline u64 shoup_h(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
ModConst make_modconst(u64 q);
inline uint32_t bitrev_h(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}

// Base-conversion constants for key switching at `ell` active Q limbs (level tables).
struct LevelPlan {
    int ell = 0, beta = 0, width = 0;   // width = ell + K ("level limbs": Q 0..ell-1 then all P)
    std::vector<int> qp;                // level limb t -> QP index
    std::vector<int> dstart, dsize;     // digit j = Q limbs [dstart, dstart+dsize)
    // ModUp: hat_inv[j][i], hat_inv_s (Shoup), hat[j][i*width + t]
    std::vector<std::vector<u64>> hat_inv, hat_inv_s, hat;
    // ModDown: p_hat_inv[k] (+Shoup), p_hat[k*ell + i], p_inv[i] (+Shoup)
    std::vector<u64> p_hat_inv, p_hat_inv_s, p_hat, p_inv, p_inv_s;
    // rescale from ell -> ell-1: ql_inv[i] = q_{ell-1}^-1 mod q_i (+Shoup)
    std::vector<u64> ql_inv, ql_inv_s;
};

struct Params {
    int logn = 0, n = 0;
    int depth = 0, scale_bits = 0, first_bits = 0, dnum = 3;
    int L = 0, K = 0, alpha = 0;        // L = depth + 2 Q primes (last = FLEXIBLEAUTOEXT extra), K P primes
    std::vector<u64> moduli;            // QP order
    std::vector<u64> psi;               // primitive 2n-th roots (minimal), QP order
    std::vector<ModConst> mc;           // QP order
    std::vector<double> sf, sf_big;     // real scaling factors per level (FLEXIBLEAUTOEXT)
    std::vector<u64> n_inv, n_inv_s;    // n^-1 mod q (+Shoup)

    // twiddle tables: psi^bitrev(i) and psi^-bitrev(i) with Shoup companions, generated per modulus on demand
    void twiddles(int qp, std::vector<u64> &w, std::vector<u64> &ws, std::vector<u64> &wi, std::vector<u64> &wis) const;

    int beta(int ell) const { return (ell + alpha - 1) / alpha; }
    const LevelPlan &plan(int ell) const;
    uint32_t galois_for_rotation(int rot) const;   // 5^rot mod 2n (negative: inverse)
    uint32_t galois_inverse(uint32_t k) const;

    static std::shared_ptr<Params> create(int logn, int depth, int scale_bits, int first_bits, int dnum);

  private:
    mutable std::map<int, LevelPlan> plans_;
    mutable std::mutex mu_;
};

} // namespace fheb200
