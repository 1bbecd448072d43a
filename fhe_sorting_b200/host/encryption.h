// encryption.h — crypto-context facade of the sort drivers: holds the context and the public key, encrypts input
// vectors; DebugEncryption additionally holds the secret key for decryption in tests.
// Mirrors the interface of the reference's src/encryption.h:56-94 / src/encryption.cpp:5-27.
#pragma once
#include <cmath>
#include <memory>
#include <stdexcept>
#include <vector>

#include "openfhe.h"

using namespace lbcrypto;

class Encryption {
  public:
    Encryption(CryptoContext<DCRTPoly> cc, PublicKey<DCRTPoly> pk) : m_cc(cc), m_PublicKey(pk) {}
    virtual ~Encryption() = default;
    // MakeCKKSPackedPlaintext (default slots = batch size) + Encrypt
    Ciphertext<DCRTPoly> encryptInput(std::vector<double> input) {
        if (input.size() > m_cc->GetRingDimension() / 2)
            throw std::invalid_argument("Input size is larger than the maximum available batch size");
        return m_cc->Encrypt(m_PublicKey, m_cc->MakeCKKSPackedPlaintext(input));
    }
    virtual std::vector<double> getPlaintext(const Ciphertext<DCRTPoly> &, double = 1e-10) const {
        throw std::runtime_error("Decryption not available in base Encryption class");
    }
    virtual Plaintext getDecrypt(const Ciphertext<DCRTPoly> &) const {
        throw std::runtime_error("Decryption not available in base Encryption class");
    }
    CryptoContext<DCRTPoly> m_cc;
    PublicKey<DCRTPoly> m_PublicKey;
};

class DebugEncryption : public Encryption {
  public:
    DebugEncryption(CryptoContext<DCRTPoly> cc, KeyPair<DCRTPoly> kp) : Encryption(cc, kp.publicKey), m_PrivateKey(kp.secretKey) {}
    Plaintext getDecrypt(const Ciphertext<DCRTPoly> &ct) const override {
        Plaintext pt;
        m_cc->Decrypt(m_PrivateKey, ct, &pt);
        return pt;
    }
    // decrypted slot values with |v| < threshold flushed to zero
    std::vector<double> getPlaintext(const Ciphertext<DCRTPoly> &ct, double threshold = 1e-10) const override {
        std::vector<double> v = getDecrypt(ct)->GetRealPackedValue();
        for (double &x : v)
            if (std::abs(x) < threshold) x = 0.0;
        return v;
    }

  private:
    PrivateKey<DCRTPoly> m_PrivateKey;
};

#ifdef ENABLE_PRINT_PT
#define PRINT_PT(enc, ct)                                                                                         \
    do {                                                                                                          \
        if (dynamic_cast<const DebugEncryption *>((enc).get()) != nullptr) {                                      \
            auto pt__ = (enc)->getDecrypt((ct));                                                                  \
            std::cout << pt__ << ": " << #ct << " Level: " << (ct)->GetLevel() << "\n";                           \
        }                                                                                                         \
    } while (0)
#else
#define PRINT_PT(enc, ct)
#endif
