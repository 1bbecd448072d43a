// fherma_roundtrip.cpp — test driver for the reference's file harness (src/sort.h:15-103 + src/main.cpp:9-44, the "FHERMA
// challenge" entry point): plays the key owner around the reference's own, unmodified `main`.
//   1. GenCryptoContext with the values of src/config.json (depth 44, scaling 40, batch 128; ring from argv), KeyGen, the
//      rotation keys of config.json, EvalMultKeyGen; everything is serialised to files with Serial::SerializeToFile /
//      SerializeEvalMultKey / SerializeEvalAutomorphismKey;
//   2. runs <ref_main> --cc .. --key_pub .. --key_mult .. --key_rot .. --input .. --output ..  (a separate process that
//      holds no secret key: it deserialises, runs DirectSort<128>, serialises the result);
//   3. deserialises the output ciphertext, decrypts it with the secret key that never left this process and prints the
//      largest deviation from std::sort.
// Test infrastructure (built by oracle/Makefile `ref` next to the reference's main).  usage:
//   fherma_roundtrip <path to ref_main> <work dir> <log2 ring dim>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <numeric>
#include <random>
#include <string>
#include <vector>

#include "openfhe.h"

using namespace lbcrypto;

int main(int argc, char **argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s <ref_main> <work dir> <log2 ring dim>\n", argv[0]);
        return 2;
    }
    const std::string ref_main = argv[1], dir = argv[2];
    const int logn = atoi(argv[3]);
    const int N = 128;
    CCParams<CryptoContextCKKSRNS> p; // src/config.json
    p.SetMultiplicativeDepth(44);
    p.SetScalingModSize(40);
    p.SetBatchSize(N);
    p.SetRingDim(1u << logn);
    p.SetSecurityLevel(logn >= 17 ? HEStd_128_classic : HEStd_NotSet);
    auto cc = GenCryptoContext(p);
    cc->Enable(PKE), cc->Enable(KEYSWITCH), cc->Enable(LEVELEDSHE), cc->Enable(ADVANCEDSHE);
    auto keys = cc->KeyGen();
    const std::vector<int> rot = {-1, -2, -4, -8, -16, -32, 1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384};
    std::vector<int> usable;
    for (int r : rot)
        if (std::abs(r) < (1 << (logn - 1))) usable.push_back(r);
    cc->EvalRotateKeyGen(keys.secretKey, usable);
    cc->EvalMultKeyGen(keys.secretKey);

    std::vector<double> x(N);
    std::vector<int> perm(N);
    std::iota(perm.begin(), perm.end(), 0);
    std::shuffle(perm.begin(), perm.end(), std::mt19937(N));
    for (int i = 0; i < N; i++) x[i] = (double)perm[i] / N; // tests/utils.h:28-51 getVectorWithMinDiff(N, 0, 1, 1/N)
    auto ct = cc->Encrypt(keys.publicKey, cc->MakeCKKSPackedPlaintext(x));

    const std::string f_cc = dir + "/cc.bin", f_pk = dir + "/pk.bin", f_mult = dir + "/mult.bin", f_rot = dir + "/rot.bin",
                      f_in = dir + "/in.bin", f_out = dir + "/out.bin";
    bool ok = Serial::SerializeToFile(f_cc, cc, SerType::BINARY) && Serial::SerializeToFile(f_pk, keys.publicKey, SerType::BINARY) &&
              Serial::SerializeToFile(f_in, ct, SerType::BINARY);
    {
        std::ofstream m(f_mult, std::ios::binary), r(f_rot, std::ios::binary);
        ok = ok && cc->SerializeEvalMultKey(m, SerType::BINARY) && cc->SerializeEvalAutomorphismKey(r, SerType::BINARY);
    }
    if (!ok) {
        fprintf(stderr, "serialisation failed\n");
        return 1;
    }
    cc->Sync();
    const std::string cmd = ref_main + " --cc " + f_cc + " --key_pub " + f_pk + " --key_mult " + f_mult + " --key_rot " + f_rot +
                            " --input " + f_in + " --output " + f_out;
    const int rc = system(cmd.c_str());
    if (rc != 0) {
        fprintf(stderr, "reference main failed: %d\n", rc);
        return 1;
    }
    Ciphertext<DCRTPoly> out;
    if (!Serial::DeserializeFromFile(f_out, out, SerType::BINARY)) {
        fprintf(stderr, "cannot read the output ciphertext\n");
        return 1;
    }
    Plaintext res;
    cc->Decrypt(keys.secretKey, out, &res);
    res->SetLength(N);
    std::vector<double> got = res->GetRealPackedValue(), want(x);
    std::sort(want.begin(), want.end());
    double err = 0;
    for (int i = 0; i < N; i++) err = std::max(err, std::abs(got[i] - want[i]));
    printf("fherma roundtrip: N=%d ring 2^%d output level %zu max_abs_err %.3e\n", N, logn, out->GetLevel(), err);
    return err < 0.01 ? 0 : 3;
}
