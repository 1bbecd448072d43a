// encode.cu — CKKS encoding on the device: special inverse FFT over the slots (double complex), scaling, rounding to
// 128-bit signed coefficients and reduction onto the RNS limbs.  Replaces the host FFT in MakeCKKSPackedPlaintext, which
// the sort drivers call once per mask inside their inner loops (reference: src/sort_algo.h:339-342, 571-574 — ~2 N
// plaintexts of up to 65536 slots per sort); on the host those FFTs cost more than all the key switching of a sort.
//
// Three launches per plaintext (the forward NTT of the limbs follows in ntt.cu):
//   enc_fft_pass (A) : the large-stride radix-2 stages, tile = R rows x C contiguous columns in shared memory
//   enc_fft_pass (B) : the last <= 10 stages, tile = 1024 contiguous points
//   enc_round_expand : bit-reversal, 1/size, scale, round to i128, residue on every limb, zero fill between the gaps
// Butterfly arithmetic and order of operations are those of the host fft_special_inv (bit-identical results).
#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

constexpr int ENC_THREADS = 256;
constexpr int ENC_TILE_LOG = 10, ENC_TILE = 1 << ENC_TILE_LOG;

// Stages u0 .. u0+nu-1 of the Gentleman-Sande special inverse FFT; stage u works on blocks of len = size >> u:
//   u = a + b ; x = (a - b) * ksi[(4 len - rot[j] mod 4 len) * (size / len)],  j = position inside the half block.
// pass A: tile element e = r * C + c  <->  point g = r * (size >> nu) + blockIdx.x * C + c     (u0 = 0)
// pass B: tile element e               <->  point g = blockIdx.x * tile + e
// real_in != nullptr (first pass only): the slot values are real doubles, entries >= nv are zero.
__global__ void __launch_bounds__(ENC_THREADS) enc_fft_pass(double2 *__restrict__ v, const double *__restrict__ real_in, int nv,
                                                            int logsize, int u0, int nu, int logc, int pass_a, int log_tile,
                                                            const double2 *__restrict__ ksi, const int *__restrict__ rot) {
    __shared__ double2 sm[ENC_TILE];
    const int tile = 1 << log_tile, size = 1 << logsize, tid = threadIdx.x;
    const int rowshift = logsize - nu; // pass A: log2 of the row stride
    auto to_g = [&](int e) -> int {
        return pass_a ? ((e >> logc) << rowshift) + (blockIdx.x << logc) + (e & ((1 << logc) - 1)) : (blockIdx.x << log_tile) + e;
    };
    for (int e = tid; e < tile; e += ENC_THREADS) {
        const int g = to_g(e);
        sm[e] = real_in ? make_double2(g < nv ? real_in[g] : 0.0, 0.0) : v[g];
    }
    __syncthreads();
    for (int u = u0; u < u0 + nu; u++) {
        const int loglenh = logsize - u - 1;                         // log2 of the butterfly distance in point space
        const int p = pass_a ? loglenh - rowshift + logc : loglenh;  // ... and in tile space
        const int lenq_mask = (4 << (loglenh + 1)) - 1, gapsh = u;   // lenq = 4 len ; gap = size / len = 2^u
        for (int b = tid; b < tile / 2; b += ENC_THREADS) {
            const int i0 = ((b >> p) << (p + 1)) | (b & ((1 << p) - 1)), i1 = i0 + (1 << p);
            const int j = to_g(i0) & ((1 << loglenh) - 1);
            const double2 w = ksi[((lenq_mask + 1) - (rot[j] & lenq_mask)) << gapsh];
            const double2 a = sm[i0], c = sm[i1];
            const double dr = a.x - c.x, di = a.y - c.y;
            sm[i0] = make_double2(a.x + c.x, a.y + c.y);
            sm[i1] = make_double2(dr * w.x - di * w.y, dr * w.y + di * w.x);
        }
        __syncthreads();
    }
    for (int e = tid; e < tile; e += ENC_THREADS) v[to_g(e)] = sm[e];
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

// coefficient j of the plaintext polynomial, on every limb: j = i * gap carries round(scale/size * Re or Im of slot
// bitrev(i mod size)) (real parts first, then imaginary parts), every other coefficient is zero
__global__ void __launch_bounds__(ENC_THREADS) enc_round_expand(DeviceTables T, const double2 *__restrict__ v, int logsize,
                                                                double scale, int loggap, int ell, u64 *__restrict__ out,
                                                                int extra_qp0, int extra) {
    const int j = blockIdx.x * ENC_THREADS + threadIdx.x;
    if (j >= T.n) return;
    const int size = 1 << logsize;
    bool live = (j & ((1 << loggap) - 1)) == 0;
    u64 lo = 0, hi = 0;
    bool neg = false;
    if (live) {
        const int i = j >> loggap, s = i & (size - 1);
        const int r = logsize ? (int)(__brev((unsigned)s) >> (32 - logsize)) : 0;
        const double2 x = v[r];
        const double f = scale / (double)size;
        double_to_i128((i < size ? x.x : x.y) * f, lo, hi);
        neg = (long long)hi < 0;
        if (neg) { // magnitude of the negative value
            lo = ~lo, hi = ~hi;
            lo += 1;
            hi += (lo == 0);
        }
        live = (lo | hi) != 0;
    }
    for (int limb = 0; limb < ell + extra; limb++) { // limbs past ell: moduli extra_qp0.. (the P primes of a QP encoding)
        u64 r = 0;
        if (live) {
            const ModConst m = T.mc[limb < ell ? limb : extra_qp0 + limb - ell];
            r = reduce128(hi, lo, m);
            if (neg) r = neg_mod(r, m.q);
        }
        out[(size_t)limb * T.n + j] = r;
    }
}

// v: device buffer of `size` double2 (work space; holds the complex slot values unless real_in is given)
void k_encode(const DeviceTables &T, double2 *v, const double *real_in, int nv, int size, double scale, const double2 *ksi,
              const int *rot, int ell, u64 *out, cudaStream_t st, int extra_qp0, int extra) {
    int logsize = 0;
    while ((1 << logsize) < size) logsize++;
    int loggap = 0;
    while ((2 * size) << loggap < T.n) loggap++;
    if (logsize > 0) {
        const int log_tile = logsize < ENC_TILE_LOG ? logsize : ENC_TILE_LOG;
        const int nb = logsize < ENC_TILE_LOG ? logsize : ENC_TILE_LOG, na = logsize - nb;
        const int grid = size >> log_tile;
        if (na > 0) {
            g_launches++, enc_fft_pass<<<grid, ENC_THREADS, 0, st>>>(v, real_in, nv, logsize, 0, na, log_tile - na, 1, log_tile, ksi, rot);
            g_launches++, enc_fft_pass<<<grid, ENC_THREADS, 0, st>>>(v, nullptr, 0, logsize, na, nb, 0, 0, log_tile, ksi, rot);
        } else {
            g_launches++, enc_fft_pass<<<grid, ENC_THREADS, 0, st>>>(v, real_in, nv, logsize, 0, nb, 0, 0, log_tile, ksi, rot);
        }
    } else if (real_in) { // a single slot: no butterflies, just widen the value
        g_launches++, enc_fft_pass<<<1, ENC_THREADS, 0, st>>>(v, real_in, nv, 0, 0, 0, 0, 0, 0, ksi, rot);
    }
    g_launches++, enc_round_expand<<<(T.n + ENC_THREADS - 1) / ENC_THREADS, ENC_THREADS, 0, st>>>(T, v, logsize, scale, loggap, ell, out, extra_qp0, extra);
}

} // namespace fheb200
