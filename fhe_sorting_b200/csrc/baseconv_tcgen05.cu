// baseconv_tcgen05.cu — the base conversion of baseconv.cu on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators
// in tensor memory).  Same exact integer algorithm and the same tables as bc_convert (mma.sync): opt-in with
// FHEB200_BC_TCGEN05=1, results bit-identical (tests/test_gpu_fullsize.py runs under the switch in tools/ab_env.sh).
//
//   D[128 coefficients][(plane b, target t)]  +=  A[128 coefficients][(source j, byte a)] x B[(source j, byte a)][(plane b, target t)]
//
// A tile (M = 128 coefficients, K = 8 bytes x up to 16 sources = 128 bytes per row) is the raw little-endian memory of the source
// limbs: every thread loads the u64 of its coefficient from each source and stores it into shared memory in the canonical K-major
// 128-byte-swizzled UMMA layout (8-row groups of 1024 bytes, 16-byte chunk index XOR row mod 8).  The constants are cut into
// chunks of 16 targets: B chunk = 128 rows (plane b x 16 targets) x 128 bytes, converted once per table from bc_convert's
// per-lane fragments (bc5_build) and copied verbatim to shared memory by every CTA.  One chunk = 4 (K = 32) x tcgen05.mma
// M128 N128 into a 128-column TMEM buffer; two buffers alternate, so the MMAs of chunk c + 1 run under the epilogue of chunk c.
// Epilogue: thread = coefficient (TMEM lane), tcgen05.ld 32x32b.x8 brings plane b of 8 targets; the 6 or 8 planes of a target are
// recombined and reduced exactly as in bc_convert (finish6 / finish8) and stored with consecutive lanes on consecutive
// coefficients (256-byte runs per warp and target).
#include <algorithm>
#include <cstdlib>
#include <stdexcept>

#include "kernels.h"
#include "modarith.cuh"

namespace fheb200 {

constexpr int BC5_M = 128;              // coefficients per tile = TMEM lanes
constexpr int BC5_CHUNK_T = 16;         // targets per B chunk
constexpr int BC5_N = 8 * BC5_CHUNK_T;  // MMA N = TMEM columns per buffer
constexpr int BC5_TILE_BYTES = 128 * 128;
constexpr int BC5_MAX_CHUNKS = 4;       // up to 64 targets

__device__ __forceinline__ uint32_t bc5_swz(int row, int kbyte) { // canonical K-major SWIZZLE_128B, rows of 128 bytes
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((kbyte >> 4) ^ (row & 7)) & 7) << 4) + (kbyte & 15));
}
// recombination / reduction: as baseconv.cu
__device__ __forceinline__ u64 bc5_lo(const uint32_t (&s)[8]) {
    return mad_wide(s[3], 1u << 24, mad_wide(s[2], 1u << 16, mad_wide(s[1], 1u << 8, (u64)s[0])));
}
__device__ __forceinline__ u64 bc5_finish6(const uint32_t (&s)[8], const uint4 tc) {
    const u64 hi = mad_wide(s[5], 1u << 8, (u64)s[4]);
    const u64 v = bc5_lo(s) + (hi << 32);
    const u64 p = ((u64)tc.y << 32) | tc.x;
    const uint32_t top = (uint32_t)(v >> (tc.w - 16));
    const uint32_t qe = (uint32_t)(((u64)top * tc.z) >> 47);
    const u64 r = v - (u64)qe * p;
    return r >= p ? r - p : r;
}
__device__ __forceinline__ u64 bc5_finish8(const uint32_t (&s)[8], const uint4 tc) {
    const u64 lo = bc5_lo(s);
    const u64 hi = mad_wide(s[7], 1u << 24, mad_wide(s[6], 1u << 16, mad_wide(s[5], 1u << 8, (u64)s[4])));
    const u64 p = ((u64)tc.y << 32) | tc.x;
    const u64 v_hi32 = hi + (lo >> 32);
    const u64 v_lo = lo + (hi << 32);
    const uint32_t top = (uint32_t)(v_hi32 >> (tc.w - 48));
    const uint32_t qe = (uint32_t)(((u64)top * tc.z) >> 47);
    const u64 r = v_lo - (u64)qe * p;
    return r >= p ? r - p : r;
}

// ---- B chunks from bc_convert's fragment table (one thread per byte)
__global__ void bc5_build(const uint2 *__restrict__ frag, int KS, int NT, int chunks, unsigned char *__restrict__ out) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= chunks * BC5_TILE_BYTES) return;
    const int chunk = id / BC5_TILE_BYTES, r = (id % BC5_TILE_BYTES) / 128, c = id % 128;
    const int b = r / BC5_CHUNK_T, t = chunk * BC5_CHUNK_T + r % BC5_CHUNK_T;
    const int nt = t >> 3, g = t & 7, j = c >> 3, a = c & 7, ks = j >> 2;
    unsigned char v = 0;
    if (nt < NT && ks < KS) {
        const int reg = (j >> 1) & 1, tq = ((j & 1) << 1) | (a >> 2), i = a & 3;
        const uint2 w = frag[((size_t)(ks * NT + nt) * 8 + b) * 32 + g * 4 + tq];
        v = (unsigned char)(((reg ? w.y : w.x) >> (8 * i)) & 0xff);
    }
    out[(size_t)chunk * BC5_TILE_BYTES + bc5_swz(r, c)] = v;
}

struct Bc5Tables {
    const unsigned char *bt[8];
};

// ---- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t bc5_desc(uint32_t saddr) { // K-major, SWIZZLE_128B, 8-row groups 1024 bytes apart, version 1
    return (uint64_t)((saddr & 0x3ffff) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void bc5_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
        :
        : "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void bc5_commit(uint32_t mbar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}\n"
            : "=r"(done)
            : "r"(mbar), "r"(parity)
            : "memory");
        if (done) break;
        if (++spins > (1u << 22)) __trap(); // a tensor-core operation that never completes must not hang the device
    }
}
__device__ __forceinline__ void bc5_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}

__device__ __forceinline__ void bc5_ld4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void bc5_cp8(uint32_t sdst, const u64 *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sdst), "l"(gsrc) : "memory");
}

// 512 threads = 16 warps: warp w reads TMEM lanes 32 (w & 3) .. + 31 (coefficient = lane) and, of every 16-target chunk, the four
// targets 4 (w >> 2) .. + 3, so the epilogue of a tile is spread over four times as many warps as there are lane quarters.
// Two CTAs per SM (2 x 256 TMEM columns, 64 registers per thread).  The A tile is double buffered and filled with cp.async.
constexpr int BC5_THREADS = 512;
template <int KS>
__global__ void __launch_bounds__(BC5_THREADS, 2) bc_convert_t5(DeviceTables T, ConvArgs A, Bc5Tables BT, int chunks) {
    extern __shared__ unsigned char bc5_raw[];
    __shared__ __align__(8) unsigned long long mbar_store[2];
    __shared__ uint32_t tmem_slot;
    __shared__ uint4 s_tconst[BC5_MAX_CHUNKS * BC5_CHUNK_T];
    __shared__ int s_tslot[BC5_MAX_CHUNKS * BC5_CHUNK_T];
    const ConvInstance I = A.inst[blockIdx.y];
    const ConvTable &C = I.tab;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, n = T.n;
    const int row = 32 * (warp & 3) + lane, q = warp >> 2; // epilogue role
    const int lrow = tid & 127, jg = tid >> 7;             // loader role: sources jg, jg + 4, ...
    unsigned char *base = bc5_raw + ((1024u - (smem_u32(bc5_raw) & 1023u)) & 1023u); // swizzle atoms are 1024-byte aligned
    unsigned char *sA = base, *sB = base + 2 * BC5_TILE_BYTES;
    const uint32_t mbar0 = smem_u32(&mbar_store[0]), mbar1 = smem_u32(&mbar_store[1]);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(2 * BC5_N) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) mbar_init(mbar0, 1), mbar_init(mbar1, 1);
    { // the B chunks of this conversion, already in the swizzled layout, and the per-target constants
        const uint4 *g = reinterpret_cast<const uint4 *>(BT.bt[blockIdx.y]);
        uint4 *s = reinterpret_cast<uint4 *>(sB);
        for (int i = tid; i < chunks * (BC5_TILE_BYTES / 16); i += BC5_THREADS) s[i] = g[i];
        if (tid < BC5_MAX_CHUNKS * BC5_CHUNK_T) {
            const bool in = tid < 8 * C.NT;
            s_tslot[tid] = in ? C.tslot[tid] : -1;
            s_tconst[tid] = in ? C.tconst[tid] : make_uint4(0, 0, 0, 0);
        }
    }
    auto fetch = [&](int tile, int buf) { // this thread's share of the A tile: coefficient lrow of sources jg + 4 i
        const u64 *g = I.src + (size_t)tile * BC5_M + lrow;
        const uint32_t sdst = smem_u32(sA + buf * BC5_TILE_BYTES);
#pragma unroll
        for (int i = 0; i < KS; i++) {
            const int j = jg + 4 * i;
            if (j < C.nsrc) bc5_cp8(sdst + bc5_swz(lrow, 8 * j), g + (size_t)j * n);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (blockIdx.x * BC5_M < n) fetch(blockIdx.x, 0);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    // instruction descriptor: D = s32, A = B = unsigned 8 bit, both K-major, N = 128, M = 128
    const uint32_t idesc = (2u << 4) | ((uint32_t)(BC5_N >> 3) << 17) | ((uint32_t)(BC5_M >> 4) << 24);
    const uint32_t baddr = smem_u32(sB);
    uint32_t phase0 = 0, phase1 = 0;

    auto issue = [&](uint32_t aaddr, int chunk, int buf) { // thread 0 only
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int ks = 0; ks < KS; ks++)
            bc5_mma(tmem + buf * BC5_N, bc5_desc(aaddr + ks * 32), bc5_desc(baddr + chunk * BC5_TILE_BYTES + ks * 32), idesc, ks > 0);
        bc5_commit(buf ? mbar1 : mbar0);
    };

    int ab = 0;
    for (int tile = blockIdx.x; tile * BC5_M < n; tile += gridDim.x, ab ^= 1) {
        const size_t x = (size_t)tile * BC5_M + row;
        const uint32_t aaddr = smem_u32(sA + ab * BC5_TILE_BYTES);
        asm volatile("cp.async.wait_all;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the copies of this thread -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads(); // A tile complete; every thread is done with the TMEM buffers of the previous tile
        if (tid == 0) {
            issue(aaddr, 0, 0);
            if (chunks > 1) issue(aaddr, 1, 1);
        }
        // the other A buffer was last read by the MMAs of the previous tile, whose completion every thread has waited for
        if ((tile + (int)gridDim.x) * BC5_M < n) fetch(tile + gridDim.x, ab ^ 1);
        for (int c = 0; c < chunks; c++) {
            const int buf = c & 1;
            if (buf) mbar_wait(mbar1, phase1), phase1 ^= 1;
            else mbar_wait(mbar0, phase0), phase0 ^= 1;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int t0 = BC5_CHUNK_T * c + 4 * q; // this warp's four targets of the chunk
            if (t0 < 8 * C.NT) {
                const int np = C.planes[t0 >> 3];
                const uint32_t tbase = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + buf * BC5_N + 4 * q;
                uint32_t s[8][4];
#pragma unroll
                for (int b = 0; b < 8; b++) {
                    if (b < 6 || np > 6) bc5_ld4(tbase + b * BC5_CHUNK_T, s[b]);
                    else s[b][0] = s[b][1] = s[b][2] = s[b][3] = 0;
                }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int slot = s_tslot[t0 + k];
                    if (slot < 0) continue;
                    const uint4 tc = s_tconst[t0 + k];
                    uint32_t sv[8];
#pragma unroll
                    for (int b = 0; b < 8; b++) sv[b] = s[b][k];
                    I.dst[(size_t)slot * n + x] = tc.w > 48 ? bc5_finish8(sv, tc) : bc5_finish6(sv, tc);
                }
            }
            if (c + 2 < chunks) { // this buffer is free again once every thread has read it
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads();
                if (tid == 0) issue(aaddr, c + 2, buf);
            }
        }
        for (int t = jg; t < I.copy_limbs; t += 4) // ModUp: the digit's own limbs are the evaluation-domain input itself
            I.copy_dst[(size_t)t * n + (size_t)tile * BC5_M + lrow] = I.copy_src[(size_t)t * n + (size_t)tile * BC5_M + lrow];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * BC5_N) : "memory");
}

// ---- host side
// B chunks of one conversion table in the UMMA layout (called by the context when it creates the table; the context owns and
// frees the buffer).  nullptr = shape outside what bc_convert_t5 covers.
unsigned char *bc5_make_table(const uint2 *frag, int KS, int NT) {
    if (KS < 1 || KS > 4 || (NT + 1) / 2 > BC5_MAX_CHUNKS) return nullptr;
    unsigned char *d = nullptr;
    const int total = BC5_MAX_CHUNKS * BC5_TILE_BYTES; // always the full size: zero beyond the table's own targets
    if (cudaMalloc(&d, (size_t)total) != cudaSuccess) throw std::runtime_error("fheb200: cudaMalloc failed (tcgen05 conversion table)");
    g_launches++, bc5_build<<<(total + 255) / 256, 256>>>(frag, KS, NT, BC5_MAX_CHUNKS, d);
    if (cudaStreamSynchronize(0) != cudaSuccess) throw std::runtime_error("fheb200: building the tcgen05 conversion table failed");
    return d;
}
// FHEB200_BC_TCGEN05: 0 = never, 1 = every shape the kernel covers, unset = where it measured faster than bc_convert on B200
// (profiles/r02/baseconv_tcgen05.txt: full digits of 14 sources with at least 17 targets, i.e. ModUp at ell >= 24)
int bc5_mode() {
    static const int m = [] { const char *e = getenv("FHEB200_BC_TCGEN05"); return e ? (*e == '1' ? 1 : 0) : 2; }();
    return m;
}
bool bc5_enabled() { return bc5_mode() != 0; }
// returns false when the shape is outside what the kernel covers (the caller falls back to bc_convert)
bool k_base_convert_tcgen05(const DeviceTables &T, const ConvArgs &A, int count, int KS, cudaStream_t st) {
    if (KS < 1 || KS > 4 || T.n % BC5_M != 0 || count > 8) return false;
    int NT = 0;
    for (int i = 0; i < count; i++) NT = std::max(NT, A.inst[i].tab.NT);
    const int chunks = (NT + 1) / 2;
    if (chunks > BC5_MAX_CHUNKS) return false;
    if (bc5_mode() == 2 && !(KS == 4 && NT >= 3)) return false;
    Bc5Tables BT;
    for (int i = 0; i < count; i++) {
        if (A.inst[i].tab.KS != KS) return false;
        if (!A.inst[i].tab.B5) return false;
        BT.bt[i] = A.inst[i].tab.B5;
    }
    static int sms = [] { int d = 0, v = 0; cudaGetDevice(&d); cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, d); return v; }();
    // two CTAs per SM hold all 512 TMEM columns; a third resident CTA would sit in tcgen05.alloc until one of them exits, so the
    // shared-memory request is padded to just over a third of the SM's 227 KB
    const size_t smem = std::max((size_t)(2 + chunks) * BC5_TILE_BYTES + 1024, (size_t)78 * 1024);
    const int tiles = T.n / BC5_M;
    const dim3 grid(std::max(1, std::min(tiles, 2 * sms / count)), count);
    switch (KS) {
#define CASE(k) case k: { \
        static bool attr = false; \
        if (!attr) cudaFuncSetAttribute(bc_convert_t5<k>, cudaFuncAttributeMaxDynamicSharedMemorySize, std::max((2 + BC5_MAX_CHUNKS) * BC5_TILE_BYTES + 1024, 78 * 1024)), attr = true; \
        g_launches++, bc_convert_t5<k><<<grid, BC5_THREADS, smem, st>>>(T, A, BT, chunks); } break;
        CASE(1) CASE(2) CASE(3) CASE(4)
#undef CASE
    }
    return true;
}

} // namespace fheb200
