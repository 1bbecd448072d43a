// encode.cu — CKKS encoding on the device: special inverse FFT over the slots (double complex), scaling and rounding
// to 128-bit signed coefficients.  Replaces the host FFT in MakeCKKSPackedPlaintext, which the sort drivers call once
// per mask inside their inner loops (reference: src/sort_algo.h:339-342, 571-574 — ~2 N plaintexts of up to 65536 slots
// per sort); on the host those FFTs cost more than all the key switching of a sort on a B200.
// One launch per radix-2 stage (the transform is at most 2^16 points = 1 MiB, latency bound; 16 launches ~ 50 us).
#include "kernels.h"

namespace fheb200 {

constexpr int ENC_THREADS = 256;

// Gentleman-Sande stage of the special inverse FFT (same butterfly order as the host fft_special_inv):
//   for blocks of `len`: u = a + b ; x = (a - b) * ksi[(lenq - rot[j] % lenq) * gap]
__global__ void __launch_bounds__(ENC_THREADS) enc_fft_inv_stage(double2 *__restrict__ v, int size, int len,
                                                                 const double2 *__restrict__ ksi, const int *__restrict__ rot) {
    const int id = blockIdx.x * ENC_THREADS + threadIdx.x;
    if (id >= size / 2) return;
    const int lenh = len >> 1, lenq = len << 2, gap = (4 * size) / lenq;
    const int blk = id / lenh, j = id - blk * lenh;
    const int i0 = blk * len + j, i1 = i0 + lenh;
    const double2 w = ksi[(lenq - (rot[j] & (lenq - 1))) * gap];
    const double2 a = v[i0], b = v[i1];
    const double dr = a.x - b.x, di = a.y - b.y;
    v[i0] = make_double2(a.x + b.x, a.y + b.y);
    v[i1] = make_double2(dr * w.x - di * w.y, dr * w.y + di * w.x);
}

// round(x) for |x| < 2^126 as a two's complement 128-bit integer (lo, hi); ties away from zero below 2^63, exact above
// 2^52 anyway (a double has no fraction bits there)
__device__ __forceinline__ void double_to_i128(double x, u64 &lo, u64 &hi) {
    const bool neg = x < 0;
    const double a = fabs(x);
    u64 mlo, mhi;
    if (a < 9.0e18) {
        mlo = (u64)llround(a), mhi = 0;
    } else {
        const u64 bits = (u64)__double_as_longlong(a);
        const int e = (int)(bits >> 52) - 1075; // a = mant * 2^e, e >= 10 here
        const u64 mant = (bits & 0x000fffffffffffffull) | 0x0010000000000000ull;
        if (e >= 64) mlo = 0, mhi = mant << (e - 64);
        else mlo = mant << e, mhi = mant >> (64 - e);
    }
    if (neg) {
        mlo = ~mlo + 1;
        mhi = ~mhi + (mlo == 0);
    }
    lo = mlo, hi = mhi;
}

// bit-reversal, 1/size, scaling and rounding: words[2 i] / words[2 (size + i)] <- real / imaginary part of slot i
__global__ void __launch_bounds__(ENC_THREADS) enc_finish(const double2 *__restrict__ v, int size, int logsize, double scale,
                                                          u64 *__restrict__ words) {
    const int i = blockIdx.x * ENC_THREADS + threadIdx.x;
    if (i >= size) return;
    const int r = logsize ? (int)(__brev((unsigned)i) >> (32 - logsize)) : 0;
    const double2 x = v[r];
    const double s = scale / (double)size;
    u64 lo, hi;
    double_to_i128(x.x * s, lo, hi);
    words[2 * (size_t)i] = lo, words[2 * (size_t)i + 1] = hi;
    double_to_i128(x.y * s, lo, hi);
    words[2 * (size_t)(size + i)] = lo, words[2 * (size_t)(size + i) + 1] = hi;
}

void k_encode_fft(double2 *v, int size, double scale, const double2 *ksi, const int *rot, u64 *words, cudaStream_t st) {
    int logsize = 0;
    while ((1 << logsize) < size) logsize++;
    const int g = (size / 2 + ENC_THREADS - 1) / ENC_THREADS;
    for (int len = size; len >= 2; len >>= 1) g_launches++, enc_fft_inv_stage<<<g, ENC_THREADS, 0, st>>>(v, size, len, ksi, rot);
    g_launches++, enc_finish<<<(size + ENC_THREADS - 1) / ENC_THREADS, ENC_THREADS, 0, st>>>(v, size, logsize, scale, words);
}

} // namespace fheb200
