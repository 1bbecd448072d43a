// comm.cu — the engine's own NCCL communicator: all-reduce of partial ciphertexts over the ranks of one job, on the
// engine's stream, with the modular fix-up as the next kernel on the same stream (no host synchronisation, no callback).
//
// NCCL is resolved at run time (dlopen of libnccl.so.2: the copy a hosting process such as PyTorch has already loaded, else
// the system one), so libfheb200.so carries no link-time dependency on it and single-GPU users never touch it.
// Bootstrap: rank 0 calls fheb200_comm_unique_id and hands the 128 bytes to the other ranks by any means (MPI, a file,
// torch.distributed.broadcast); every rank then calls fheb200_ctx_comm_init(ctx, id, rank, world).
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

#include "../../include/fhe_b200.h"
#include "engine.h"

using namespace fheb200;

struct fheb200_ctx { std::shared_ptr<Context> c; };
struct fheb200_ct { Ct c; };
namespace fheb200 { void set_last_error(const char *); }

namespace {
struct NcclApi {
    decltype(&ncclGetUniqueId) get_unique_id = nullptr;
    decltype(&ncclCommInitRank) comm_init_rank = nullptr;
    decltype(&ncclAllReduce) all_reduce = nullptr;
    decltype(&ncclCommDestroy) comm_destroy = nullptr;
    decltype(&ncclGetErrorString) error_string = nullptr;
    decltype(&ncclGetVersion) get_version = nullptr;
};
const NcclApi &nccl() {
    static NcclApi api;
    static std::once_flag once;
    static std::string err;
    std::call_once(once, [] {
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); // already in the process (PyTorch's copy)?
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) {
            err = std::string("fheb200: NCCL not found (dlopen libnccl.so.2): ") + dlerror();
            return;
        }
        auto sym = [&](const char *name) {
            void *p = dlsym(h, name);
            if (!p && err.empty()) err = std::string("fheb200: NCCL symbol missing: ") + name;
            return p;
        };
        api.get_unique_id = (decltype(api.get_unique_id))sym("ncclGetUniqueId");
        api.comm_init_rank = (decltype(api.comm_init_rank))sym("ncclCommInitRank");
        api.all_reduce = (decltype(api.all_reduce))sym("ncclAllReduce");
        api.comm_destroy = (decltype(api.comm_destroy))sym("ncclCommDestroy");
        api.error_string = (decltype(api.error_string))sym("ncclGetErrorString");
        api.get_version = (decltype(api.get_version))sym("ncclGetVersion");
    });
    if (!err.empty()) throw std::runtime_error(err);
    return api;
}
void nccl_check(ncclResult_t r, const char *what) {
    if (r != ncclSuccess) throw std::runtime_error(std::string("NCCL error in ") + what + ": " + nccl().error_string(r));
}
} // namespace

#define GUARD(...)                                \
    try {                                         \
        __VA_ARGS__;                              \
        return 0;                                 \
    } catch (const std::exception &e) {           \
        fheb200::set_last_error(e.what());        \
        return 1;                                 \
    } catch (...) {                               \
        fheb200::set_last_error("unknown error"); \
        return 1;                                 \
    }

namespace fheb200 {
void Context::comm_init(const unsigned char *id128, int rank, int world) {
    CtxLock g(*this);
    if (world < 1 || rank < 0 || rank >= world) throw std::invalid_argument("comm_init: bad rank / world");
    // residues below 2^60 (first, extra and special limbs) are summed as 64-bit integers before the fix-up
    if (world > 16) throw std::invalid_argument("comm_init: at most 16 ranks (raw residues are summed in 64 bits)");
    comm_destroy();
    if (world == 1) return;
    FHE_CUDA(cudaSetDevice(device));
    ncclUniqueId id;
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(&id, id128, sizeof(id));
    ncclComm_t c = nullptr;
    nccl_check(nccl().comm_init_rank(&c, world, id, rank), "ncclCommInitRank");
    comm_ = c, comm_rank_ = rank, comm_world_ = world;
}
void Context::comm_destroy() {
    CtxLock g(*this);
    if (comm_) {
        cudaStreamSynchronize(stream);
        nccl().comm_destroy((ncclComm_t)comm_);
    }
    comm_ = nullptr, comm_rank_ = 0, comm_world_ = 1;
}
// a <- sum over ranks of a (every limb modulo its modulus), in place, on the engine's stream
void Context::all_reduce_sum(const Ct &a) {
    CtxLock g(*this);
    if (!a) throw std::invalid_argument("AllReduceSum: null ciphertext");
    if (!comm_) throw std::runtime_error("AllReduceSum: no communicator (fheb200_ctx_comm_init)");
    if (forked_) throw std::runtime_error("AllReduceSum: not inside a forked stream region");
    FHE_CUDA(cudaSetDevice(device));
    const int n = P->n;
    const size_t words = (size_t)a->npoly * a->ell * n;
    a->reduced_cache.reset();
    nccl_check(nccl().all_reduce(a->mem->p, a->mem->p, words, ncclUint64, ncclSum, (ncclComm_t)comm_, stream), "ncclAllReduce");
    for (int p = 0; p < a->npoly; p++) k_reduce_words(T, a->ell, a->poly(p, n), stream);
}
} // namespace fheb200

extern "C" {
int fheb200_comm_unique_id(unsigned char *id128) {
    GUARD({
        if (!id128) throw std::invalid_argument("null argument: id");
        ncclUniqueId id;
        nccl_check(nccl().get_unique_id(&id), "ncclGetUniqueId");
        memcpy(id128, &id, sizeof(id));
    })
}
int fheb200_ctx_comm_init(fheb200_ctx *ctx, const unsigned char *id128, int rank, int world) {
    GUARD({
        if (!ctx || (world > 1 && !id128)) throw std::invalid_argument("null argument");
        ctx->c->comm_init(id128, rank, world);
    })
}
int fheb200_ctx_comm_destroy(fheb200_ctx *ctx) {
    GUARD({
        if (!ctx) throw std::invalid_argument("null argument: ctx");
        ctx->c->comm_destroy();
    })
}
int fheb200_ctx_comm_info(fheb200_ctx *ctx, int *rank, int *world, int *nccl_version) {
    GUARD({
        if (!ctx) throw std::invalid_argument("null argument: ctx");
        if (rank) *rank = ctx->c->comm_rank_;
        if (world) *world = ctx->c->comm_world_;
        if (nccl_version) {
            *nccl_version = 0;
            if (ctx->c->comm_) nccl().get_version(nccl_version);
        }
    })
}
int fheb200_ct_all_reduce_sum(fheb200_ctx *ctx, fheb200_ct *ct) {
    GUARD({
        if (!ctx || !ct) throw std::invalid_argument("null argument");
        ctx->c->all_reduce_sum(ct->c);
    })
}
}
