// ntt.cu — negacyclic NTT / iNTT over 64-bit prime moduli, batched over limbs.
//
// Forward  : Cooley-Tukey, natural order in, bit-reversed order out, twiddles psi^bitrev(i) (Harvey lazy
//            butterflies, values kept in [0,4q), q < 2^61).
// Inverse  : Gentleman-Sande mirror image, values kept in [0,2q), n^-1 folded into the last store.
// The log2(n) radix-2 stages are split in two passes so that each pass works on a shared-memory tile:
//   pass A : the SA large-stride stages; a tile is 2^SA rows x C contiguous columns (coalesced 8C-byte rows)
//   pass B : the SB small-stride stages; a tile is TE contiguous coefficients (TE / 2^SB independent blocks)
// n <= 2^11 runs as a single pass-B launch.  Algorithmic traffic: 16 n bytes per limb per pass.
#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

std::atomic<unsigned long long> g_launches{0};

constexpr int NTT_THREADS = 256;
constexpr int NTT_TILE = 4096; // elements per CTA tile (32 KB of shared memory)

__device__ __forceinline__ void ct_butterfly(u64 &x, u64 &y, u64 w, u64 ws, u64 q, u64 q2) {
    u64 a = x >= q2 ? x - q2 : x;
    u64 t = shoup_lazy(y, w, ws, q);
    x = a + t;
    y = a + q2 - t;
}
__device__ __forceinline__ void gs_butterfly(u64 &x, u64 &y, u64 w, u64 ws, u64 q, u64 q2) {
    u64 s = x + y;
    u64 d = x + q2 - y;
    x = s >= q2 ? s - q2 : s;
    y = shoup_lazy(d, w, ws, q);
}

// ---------------------------------------------------------------- forward
__global__ void __launch_bounds__(NTT_THREADS) ntt_fwd_passA(DeviceTables T, u64 *buf, const uint32_t *desc, int SA) {
    extern __shared__ u64 sm[];
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, n = T.n;
    u64 *data = buf + (size_t)(d >> 8) * n;
    const u64 q = T.mc[qp].q, q2 = q << 1;
    const u64 *tw = T.tw + (size_t)qp * n, *tws = T.tws + (size_t)qp * n;
    const int R = 1 << SA, C = NTT_TILE >> SA, rowstride = n >> SA, col0 = blockIdx.x * C;
    for (int e = threadIdx.x; e < NTT_TILE; e += NTT_THREADS) {
        int r = e / C, c = e - r * C;
        sm[e] = data[(size_t)r * rowstride + col0 + c];
    }
    __syncthreads();
    for (int u = 0; u < SA; u++) {
        const int sh = SA - 1 - u;
        for (int id = threadIdx.x; id < NTT_TILE / 2; id += NTT_THREADS) {
            int h = id / C, c = id - h * C;
            int hi = h >> sh, lo = h & ((1 << sh) - 1);
            int r = (hi << (sh + 1)) | lo;
            int i0 = r * C + c, i1 = i0 + (C << sh);
            int ti = (1 << u) + hi;
            u64 x = sm[i0], y = sm[i1];
            ct_butterfly(x, y, __ldg(tw + ti), __ldg(tws + ti), q, q2);
            sm[i0] = x;
            sm[i1] = y;
        }
        __syncthreads();
    }
    for (int e = threadIdx.x; e < NTT_TILE; e += NTT_THREADS) {
        int r = e / C, c = e - r * C;
        data[(size_t)r * rowstride + col0 + c] = sm[e];
    }
    (void)R;
}

// tile = TE contiguous elements; SA = number of stages already done (twiddle offset), SB stages here
__global__ void __launch_bounds__(NTT_THREADS) ntt_fwd_passB(DeviceTables T, u64 *buf, const uint32_t *desc, int SA, int SB,
                                                             int TE) {
    extern __shared__ u64 sm[];
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, n = T.n;
    u64 *data = buf + (size_t)(d >> 8) * n + (size_t)blockIdx.x * TE;
    const u64 q = T.mc[qp].q, q2 = q << 1;
    const u64 *tw = T.tw + (size_t)qp * n, *tws = T.tws + (size_t)qp * n;
    for (int e = threadIdx.x; e < TE; e += NTT_THREADS) sm[e] = data[e];
    __syncthreads();
    const int blocks_per_tile = TE >> SB, halfB = 1 << (SB - 1);
    for (int u = 0; u < SB; u++) {
        const int sh = SB - 1 - u;
        for (int id = threadIdx.x; id < TE / 2; id += NTT_THREADS) {
            int p = id >> (SB - 1), h = id & (halfB - 1);
            int hi = h >> sh, lo = h & ((1 << sh) - 1);
            int e = (p << SB) | (hi << (sh + 1)) | lo;
            int b = blockIdx.x * blocks_per_tile + p;
            int ti = (((1 << SA) + b) << u) + hi;
            u64 x = sm[e], y = sm[e + (1 << sh)];
            ct_butterfly(x, y, __ldg(tw + ti), __ldg(tws + ti), q, q2);
            sm[e] = x;
            sm[e + (1 << sh)] = y;
        }
        __syncthreads();
    }
    for (int e = threadIdx.x; e < TE; e += NTT_THREADS) {
        u64 x = sm[e];
        x = x >= q2 ? x - q2 : x;
        data[e] = x >= q ? x - q : x;
    }
}

// ---------------------------------------------------------------- inverse
__global__ void __launch_bounds__(NTT_THREADS) ntt_inv_passB(DeviceTables T, u64 *buf, const uint32_t *desc, int SA, int SB,
                                                             int TE, int last) {
    extern __shared__ u64 sm[];
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, n = T.n;
    u64 *data = buf + (size_t)(d >> 8) * n + (size_t)blockIdx.x * TE;
    const u64 q = T.mc[qp].q, q2 = q << 1;
    const u64 *tw = T.twi + (size_t)qp * n, *tws = T.twis + (size_t)qp * n;
    for (int e = threadIdx.x; e < TE; e += NTT_THREADS) sm[e] = data[e];
    __syncthreads();
    const int blocks_per_tile = TE >> SB, halfB = 1 << (SB - 1);
    for (int u = SB - 1; u >= 0; u--) {
        const int sh = SB - 1 - u;
        for (int id = threadIdx.x; id < TE / 2; id += NTT_THREADS) {
            int p = id >> (SB - 1), h = id & (halfB - 1);
            int hi = h >> sh, lo = h & ((1 << sh) - 1);
            int e = (p << SB) | (hi << (sh + 1)) | lo;
            int b = blockIdx.x * blocks_per_tile + p;
            int ti = (((1 << SA) + b) << u) + hi;
            u64 x = sm[e], y = sm[e + (1 << sh)];
            gs_butterfly(x, y, __ldg(tw + ti), __ldg(tws + ti), q, q2);
            sm[e] = x;
            sm[e + (1 << sh)] = y;
        }
        __syncthreads();
    }
    if (last) {
        const u64 ni = T.ninv[qp], nis = T.ninv[T.F + qp];
        for (int e = threadIdx.x; e < TE; e += NTT_THREADS) data[e] = shoup_mul(sm[e], ni, nis, q);
    } else {
        for (int e = threadIdx.x; e < TE; e += NTT_THREADS) data[e] = sm[e];
    }
}

__global__ void __launch_bounds__(NTT_THREADS) ntt_inv_passA(DeviceTables T, u64 *buf, const uint32_t *desc, int SA) {
    extern __shared__ u64 sm[];
    const uint32_t d = desc[blockIdx.y];
    const int qp = d & 0xff, n = T.n;
    u64 *data = buf + (size_t)(d >> 8) * n;
    const u64 q = T.mc[qp].q, q2 = q << 1;
    const u64 *tw = T.twi + (size_t)qp * n, *tws = T.twis + (size_t)qp * n;
    const int C = NTT_TILE >> SA, rowstride = n >> SA, col0 = blockIdx.x * C;
    for (int e = threadIdx.x; e < NTT_TILE; e += NTT_THREADS) {
        int r = e / C, c = e - r * C;
        sm[e] = data[(size_t)r * rowstride + col0 + c];
    }
    __syncthreads();
    for (int u = SA - 1; u >= 0; u--) {
        const int sh = SA - 1 - u;
        for (int id = threadIdx.x; id < NTT_TILE / 2; id += NTT_THREADS) {
            int h = id / C, c = id - h * C;
            int hi = h >> sh, lo = h & ((1 << sh) - 1);
            int r = (hi << (sh + 1)) | lo;
            int i0 = r * C + c, i1 = i0 + (C << sh);
            int ti = (1 << u) + hi;
            u64 x = sm[i0], y = sm[i1];
            gs_butterfly(x, y, __ldg(tw + ti), __ldg(tws + ti), q, q2);
            sm[i0] = x;
            sm[i1] = y;
        }
        __syncthreads();
    }
    const u64 ni = T.ninv[qp], nis = T.ninv[T.F + qp];
    for (int e = threadIdx.x; e < NTT_TILE; e += NTT_THREADS) {
        int r = e / C, c = e - r * C;
        data[(size_t)r * rowstride + col0 + c] = shoup_mul(sm[e], ni, nis, q);
    }
}

static void split(int logn, int &SA, int &SB, int &TE) {
    if (logn <= 11) {
        SA = 0, SB = logn, TE = 1 << logn;
    } else {
        SB = 9, SA = logn - SB, TE = NTT_TILE; // SA = 3..8: a pass-A row is C = 4096 >> SA >= 16 contiguous words
    }
}

void k_ntt_fwd(const DeviceTables &T, u64 *buf, const uint32_t *d_desc, int count, cudaStream_t st) {
    if (count <= 0) return;
    int SA, SB, TE;
    split(T.logn, SA, SB, TE);
    dim3 grid(T.n / TE, count);
    if (SA > 0) g_launches++, ntt_fwd_passA<<<grid, NTT_THREADS, NTT_TILE * sizeof(u64), st>>>(T, buf, d_desc, SA);
    g_launches++, ntt_fwd_passB<<<grid, NTT_THREADS, TE * sizeof(u64), st>>>(T, buf, d_desc, SA, SB, TE);
}
void k_ntt_inv(const DeviceTables &T, u64 *buf, const uint32_t *d_desc, int count, cudaStream_t st) {
    if (count <= 0) return;
    int SA, SB, TE;
    split(T.logn, SA, SB, TE);
    dim3 grid(T.n / TE, count);
    g_launches++, ntt_inv_passB<<<grid, NTT_THREADS, TE * sizeof(u64), st>>>(T, buf, d_desc, SA, SB, TE, SA == 0);
    if (SA > 0) g_launches++, ntt_inv_passA<<<grid, NTT_THREADS, NTT_TILE * sizeof(u64), st>>>(T, buf, d_desc, SA);
}

} // namespace fheb200
