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
        if (!A.inst[i].tab.B5 || A.inst[i].extra) return false;
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

// ---------------------------------------------------------------------------------------------------------------------
// Weighted sums of many polynomials (Paterson-Stockmeyer leaf sums: up to 8 outputs = sum_t w[g][t] in_t over up to 63 inputs) on
// tcgen05: the contraction of baseconv.cu's lc_tc_ring with the accumulators in tensor memory.
//   D[128 coefficients][(plane b, output g)] += A[128 coefficients][(input t, byte a)] x B[(input t, byte a)][(plane b, output g)]
// K = 8 bytes x up to 64 inputs is cut into K blocks of 16 inputs: one "granule" = the [128 rows x 128 bytes] swizzled A block of
// one tile and K block (16 KB).  A CTA (512 threads, one per SM) streams its contiguous range of (limb, tile) work through a ring of
// 12 granules filled by cp.async (11 in flight: ~176 KB per SM, no registers held), issues the 4 MMAs (M128 N64 K32) of a granule as
// soon as it has landed, and after the last K block of a tile reads the 64 TMEM columns back (tcgen05.ld 32x32b.x2: plane b of the two
// outputs of this warp), recombines the planes and reduces the 128-bit value exactly as lc_tc_ring does.  The B operand of the
// current limb (8 planes x 8 outputs rows, 32 KB) lives in shared memory and is reloaded when the range crosses into the next limb.
constexpr int LC5_THREADS = 512, LC5_RING = 12, LC5_N = 64;
constexpr int LC5_GRAN = BC5_TILE_BYTES;      // 16 KB
constexpr int LC5_BBLOCK = 64 * 128;          // B rows of one K block
// w: split weights [(g * count + t) * ell + limb] (lo30, hi30); out: [ell][KB][64 rows x 128 bytes] swizzled
__global__ void __launch_bounds__(512) lc5_build(DeviceTables T, int ell, int count, int G, int KB, const uint2 *__restrict__ w,
                                                 unsigned char *__restrict__ out) {
    const int li = blockIdx.x, j = threadIdx.x >> 3, g = threadIdx.x & 7;
    if (j >= 16 * KB) return;
    const ModConst m = T.mc[li];
    u64 word[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (g < G && j < count) {
        const uint2 sp = __ldg(w + (size_t)(g * count + j) * ell + li);
        const u64 c0 = (u64)sp.x | ((u64)sp.y << 30);
#pragma unroll
        for (int a = 0; a < 8; a++) {
            const u64 cp = mul_mod(c0, reduce64(1ull << (8 * a), m), m);
#pragma unroll
            for (int b = 0; b < 8; b++) word[b] |= ((cp >> (8 * b)) & 0xffull) << (8 * a);
        }
    }
    unsigned char *blk = out + ((size_t)li * KB + (j >> 4)) * LC5_BBLOCK;
#pragma unroll
    for (int b = 0; b < 8; b++) *reinterpret_cast<u64 *>(blk + bc5_swz(b * 8 + g, 8 * (j & 15))) = word[b];
}
__device__ __forceinline__ void bc5_ld2(uint32_t taddr, uint32_t (&r)[2]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t mbar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}\n" ::"r"(mbar) : "memory");
}
// Roles.  Warps 0..15 (thread = coefficient row tid & 127, group g = tid >> 7): per tile, group g copies K block g (its 16 inputs,
// cp.async 8 bytes each) two tiles ahead, signals "K block g of tile i landed" (full[(i % 3) * 4 + g], 128 arrivals), and retires
// tile i - 1: waits for done[(i - 1) & 3] (tcgen05.commit after the tile's last MMA: its ring slots are free again), reads its 64
// TMEM columns back (tcgen05.ld 32x32b.x2: plane b of outputs 2g, 2g + 1), signals tfree[buffer] (512 arrivals), recombines the
// planes and reduces the 128-bit value exactly as lc_tc_ring does.  Thread 512 only issues MMAs: per tile it waits for tfree of the
// buffer (tile i - 2 read back) and for each full[.], so no CTA-wide barrier sits between a block landing and its MMAs.
// CTA-wide barriers only where the B operand changes (a CTA's contiguous range crosses into the next limb at most once or twice).
__global__ void __launch_bounds__(LC5_THREADS + 32, 1) lc_tc_t5(DeviceTables T, int ell, int limbs2, int count, int G, int KS,
                                                                const u64 *const *__restrict__ in, const unsigned char *__restrict__ Bt,
                                                                u64 *const *__restrict__ out) {
    extern __shared__ unsigned char lc5_raw[];
    __shared__ __align__(8) unsigned long long mbar_store[LC5_RING + 4 + 2];
    __shared__ uint32_t tmem_slot;
    __shared__ const u64 *ptr[64];
    const int tid = threadIdx.x, warp = tid >> 5, n = T.n;
    const bool producer = tid < LC5_THREADS;
    const int KB = (KS + 3) >> 2, tl_log = T.logn - 7; // tiles per limb = n / 128 (a power of two)
    const int row = tid & 127, g = (tid >> 7) & 3;     // coefficient row (= TMEM lane), K block copied / output pair read back
    unsigned char *base = lc5_raw + ((1024u - (smem_u32(lc5_raw) & 1023u)) & 1023u);
    unsigned char *sB = base, *sA = base + 4 * LC5_BBLOCK;
    const uint32_t full0 = smem_u32(&mbar_store[0]), done0 = full0 + 8 * LC5_RING, tfree0 = done0 + 8 * 4;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(2 * LC5_N) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid < LC5_RING) mbar_init(full0 + 8 * tid, 128);
    else if (tid < LC5_RING + 4) mbar_init(full0 + 8 * tid, 1);
    else if (tid < LC5_RING + 6) mbar_init(full0 + 8 * tid, LC5_THREADS);
    if (tid < 64) ptr[tid] = in[tid < count ? tid : 0]; // padding reads input 0 (its weights are zero)
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    const uint32_t idesc = (2u << 4) | ((uint32_t)(LC5_N >> 3) << 17) | ((uint32_t)(BC5_M >> 4) << 24);
    // this CTA's contiguous range of tiles in (limb-polynomial, tile) order
    const long long total_tiles = (long long)limbs2 << tl_log;
    const int t_begin = (int)(total_tiles * blockIdx.x / gridDim.x), t_end = (int)(total_tiles * (blockIdx.x + 1) / gridDim.x);
    const int ntiles = t_end - t_begin;
    const int ncopy = min(16, 4 * KS - 16 * g); // inputs of this thread's K block (<= 0: the group only takes part in the epilogue)
    // ring slot of (tile i, K block g): ((i % 3) * 4 + g); row base inside the slot with the swizzle term of this row folded in
    const uint32_t a_row = smem_u32(sA) + g * LC5_GRAN + (row >> 3) * 1024 + (row & 7) * 128 + ((row & 7) << 4);

    auto fetch = [&](int i) { // K block g of tile i -> its ring slot
        if (i < ntiles && ncopy > 0) {
            const int tile = t_begin + i;
            const size_t goff = (size_t)(tile >> tl_log) * n + (size_t)(tile & ((1 << tl_log) - 1)) * BC5_M + row;
            const uint32_t sdst = a_row + (uint32_t)(i % 3) * (4 * LC5_GRAN);
            const u64 *const *pp = ptr + 16 * g;
#pragma unroll
            for (int jl = 0; jl < 16; jl++) // input jl of the block: 16-byte chunk jl >> 1 (XOR row & 7, folded into a_row), half jl & 1
                if (jl < ncopy) bc5_cp8((sdst ^ (uint32_t)((jl >> 1) << 4)) + ((jl & 1) << 3), pp[jl] + goff);
        }
        asm volatile("cp.async.commit_group;" ::: "memory"); // one group per tile (empty past the end)
    };
    if (producer) fetch(0), fetch(1);

    int cur_lp = -1;
    ModConst m = T.mc[0];
    int np = 8;
    auto retire = [&](int i) { // tile i: wait for its MMAs (its slots are free again), read the accumulators back, store
        mbar_wait(done0 + 8 * (i & 3), (i >> 2) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int tile = t_begin + i;
        const uint32_t tbase = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + (i & 1) * LC5_N + 2 * g;
        uint32_t s[8][2];
#pragma unroll
        for (int b = 0; b < 8; b++) {
            if (b < np) bc5_ld2(tbase + b * 8, s[b]);
            else s[b][0] = s[b][1] = 0;
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(tfree0 + 8 * (i & 1)); // this thread is done with the TMEM buffer
        const size_t at = (size_t)(tile >> tl_log) * n + (size_t)(tile & ((1 << tl_log) - 1)) * BC5_M + row;
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int o = 2 * g + k;
            if (o >= G) continue;
            uint32_t sv[8];
#pragma unroll
            for (int b = 0; b < 8; b++) sv[b] = s[b][k];
            const u64 lo4 = bc5_lo(sv);
            const u64 hi4 = mad_wide(sv[7], 1u << 24, mad_wide(sv[6], 1u << 16, mad_wide(sv[5], 1u << 8, (u64)sv[4])));
            const u64 vlo = lo4 + (hi4 << 32), vhi = (hi4 >> 32) + (vlo < lo4);
            out[o][at] = reduce128(vhi, vlo, m);
        }
    };
    int retired = 0, i3 = 0, phase3 = 0; // i3 = i % 3, phase3 = (i / 3) & 1
#pragma unroll 1
    for (int i = 0; i < ntiles + 1; i++, (++i3 == 3 ? (i3 = 0, phase3 ^= 1) : 0)) {
        const int lp = (t_begin + i) >> tl_log;
        if (i < ntiles && lp != cur_lp) { // new limb: everything that used the old B operand and modulus finishes first
            if (producer)
                while (retired < i) retire(retired++);
            __syncthreads();
            cur_lp = lp;
            const int li = lp % ell;
            m = T.mc[li];
            np = (int)(m.bits + 7) >> 3;
            if (producer) {
                const uint4 *gsrc = reinterpret_cast<const uint4 *>(Bt + (size_t)li * KB * LC5_BBLOCK);
                uint4 *sdst = reinterpret_cast<uint4 *>(sB);
                for (int k = tid; k < KB * (LC5_BBLOCK / 16); k += LC5_THREADS) sdst[k] = gsrc[k];
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            __syncthreads();
        }
        if (producer) {
            if (i < ntiles && ncopy > 0) { // this thread's part of tile i has landed (2 + i groups committed so far)
                asm volatile("cp.async.wait_group 1;" ::: "memory");
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(full0 + 8 * (i3 * 4 + g));
            }
            while (retired < i) retire(retired++); // tile i - 1
            fetch(i + 2);                           // into the slots tile i - 1 occupied
        } else if (tid == LC5_THREADS && i < ntiles) {
            if (i >= 2) mbar_wait(tfree0 + 8 * (i & 1), ((i >> 1) - 1) & 1); // tile i - 2 has been read out of this TMEM buffer
            for (int kb = 0; kb < KB; kb++) {
                mbar_wait(full0 + 8 * (i3 * 4 + kb), phase3);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t aaddr = smem_u32(sA) + (i3 * 4 + kb) * LC5_GRAN, baddr = smem_u32(sB + kb * LC5_BBLOCK);
                const int steps = min(4, KS - 4 * kb);
                for (int ks = 0; ks < steps; ks++)
                    bc5_mma(tmem + (i & 1) * LC5_N, bc5_desc(aaddr + ks * 32), bc5_desc(baddr + ks * 32), idesc, (kb | ks) != 0);
            }
            bc5_commit(done0 + 8 * (i & 3));
        }
        if (!producer) __syncwarp(); // the idle lanes of the issuing warp stay with lane 0
    }
    if (producer) asm volatile("cp.async.wait_all;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * LC5_N) : "memory");
}

// FHEB200_LC_TCGEN05: 0 = never, unset or 1 = whenever the shape is covered (the default since it measured faster than lc_tc_ring
// at every size tried: profiles/r02/lincomb_tcgen05.txt)
int lc5_mode() {
    static const int m = [] { const char *e = getenv("FHEB200_LC_TCGEN05"); return e ? (*e == '0' ? 0 : 1) : 1; }();
    return m;
}
// scratch: at least lincomb_tc_scratch_words(ell, count) words
bool k_lincomb_tcgen05(const DeviceTables &T, int ell, int limbs2, int count, int G, const u64 *const *in, const uint2 *w, u64 *const *out,
                       u64 *scratch, cudaStream_t st) {
    const int KS = (count + 3) / 4, KB = (KS + 3) / 4;
    if (count < 1 || count > 64 || G < 1 || G > 8 || T.n % BC5_M != 0) return false;
    unsigned char *Bt = reinterpret_cast<unsigned char *>(scratch);
    g_launches++, lc5_build<<<ell, 512, 0, st>>>(T, ell, count, G, KB, w, Bt);
    static int sms = [] { int d = 0, v = 0; cudaGetDevice(&d); cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, d); return v; }();
    const size_t smem = (size_t)4 * LC5_BBLOCK + (size_t)LC5_RING * LC5_GRAN + 1024;
    static bool attr = false;
    if (!attr) cudaFuncSetAttribute(lc_tc_t5, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), attr = true;
    const long long tiles = (long long)limbs2 * (T.n / BC5_M);
    const int grid = (int)std::min<long long>(tiles, sms);
    g_launches++, lc_tc_t5<<<grid, LC5_THREADS + 32, smem, st>>>(T, ell, limbs2, count, G, KS, in, Bt, out);
    return true;
}

} // namespace fheb200
