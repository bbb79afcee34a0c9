// The row-sharded build driven from C++ over NCCL [not in reference; one process per GPU].
//
// gravo-mg_b200/sharded.py drives the phases of csrc/shard.cu from Python and exchanges through torch.distributed:
// ~230 phase / exchange steps per build, each a ctypes call or a tensor wrapper (~20-65 us of host time apiece).
// gmg_shard_build_nccl runs the SAME loop (the one documented in include/gravomg_b200.h) in one call, with the
// all-gathers issued as ncclAllGather on the context's stream, in place on the library's exchange buffers.
//
// NCCL is not linked: libnccl.so.2 is opened at run time (the copy torch has already loaded when there is one, the
// system's otherwise), so the library still loads on a box without NCCL and the single-GPU path never touches it.
// The communicator is this library's own: rank 0 calls gmg_nccl_unique_id, the caller broadcasts the 128 bytes any
// way it likes (the Python driver uses torch.distributed's store), every rank calls gmg_nccl_init.
#include "hierarchy_types.h"
#include "shard_types.h"

#include <dlfcn.h>

#include <mutex>

using namespace gmg;

namespace {

// the few NCCL entry points used, with the types of nccl.h (2.x ABI)
struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
constexpr int kNcclChar = 0;          // ncclInt8 / ncclChar
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string error;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) { api.error = std::string("libnccl.so.2 not found: ") + dlerror(); return; }
        auto sym = [&](const char* s) {
            void* p = dlsym(api.handle, s);
            if (!p && api.error.empty()) api.error = std::string("missing NCCL symbol ") + s;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    });
    return api;
}

void nccl_check(int rc, const char* what) {
    if (rc == 0) return;
    const char* msg = nccl().GetErrorString ? nccl().GetErrorString(rc) : "?";
    throw Error{GMG_ERR_NCCL, std::string(what) + ": " + msg};
}

}  // namespace

struct gmg_nccl {
    NcclComm comm = nullptr;
    int rank = 0, world = 1;
    DevBuf<unsigned char> staging;       // padded slices of a variable-size gather
    int64_t exchanges = 0, bytes = 0;
};

namespace {

// G(buffer): every rank's owned slice of a replicated buffer to every rank, in place
void gather_fixed(gmg_ctx* c, gmg_shard* s, gmg_nccl* n, int which) {
    void* ptr = nullptr;
    int32_t eb = 0;
    int64_t slice = 0;
    const int rc = gmg_shard_buffer(s, which, &ptr, &eb, &slice);
    if (rc != GMG_OK) throw Error{rc, "gmg_shard_buffer"};
    const size_t bytes = size_t(eb) * size_t(slice);
    ++n->exchanges;
    n->bytes += int64_t(bytes) * n->world;
    if (bytes == 0) return;
    unsigned char* full = static_cast<unsigned char*>(ptr);
    nccl_check(nccl().AllGather(full + size_t(n->rank) * bytes, full, bytes, kNcclChar, n->comm, c->stream), "ncclAllGather");
}

// slices of different sizes (packed rows), each already in place in the final array: they are nearly equal, so
// they travel as ONE equal-size all-gather through a padded staging buffer and are copied into place (NCCL's
// grouped broadcasts for uneven sizes measured 2.4x slower on 8 B200)
void gather_var(gmg_ctx* c, gmg_shard* s, gmg_nccl* n, int which) {
    void* ptr = nullptr;
    int32_t eb = 0;
    std::vector<int64_t> off(size_t(n->world) + 1);
    const int rc = gmg_shard_buffer_v(s, which, &ptr, &eb, off.data());
    if (rc != GMG_OK) throw Error{rc, "gmg_shard_buffer_v"};
    unsigned char* full = static_cast<unsigned char*>(ptr);
    size_t smax = 0;
    for (int r = 0; r < n->world; ++r) smax = std::max(smax, size_t(off[r + 1] - off[r]) * size_t(eb));
    ++n->exchanges;
    n->bytes += off[n->world] * int64_t(eb);
    if (smax == 0) return;
    smax = (smax + 15) / 16 * 16;
    if (n->staging.n < int64_t(smax) * n->world) n->staging.alloc(c, int64_t(smax) * n->world);
    unsigned char* st = n->staging.p;
    const size_t mine = size_t(off[n->rank + 1] - off[n->rank]) * size_t(eb);
    if (mine) GMG_CUDA(cudaMemcpyAsync(st + size_t(n->rank) * smax, full + size_t(off[n->rank]) * eb, mine, cudaMemcpyDeviceToDevice, c->stream));
    nccl_check(nccl().AllGather(st + size_t(n->rank) * smax, st, smax, kNcclChar, n->comm, c->stream), "ncclAllGather");
    for (int r = 0; r < n->world; ++r) {
        const size_t len = size_t(off[r + 1] - off[r]) * size_t(eb);
        if (r != n->rank && len)
            GMG_CUDA(cudaMemcpyAsync(full + size_t(off[r]) * eb, st + size_t(r) * smax, len, cudaMemcpyDeviceToDevice, c->stream));
    }
}

void gather(gmg_ctx* c, gmg_shard* s, gmg_nccl* n, int which) {
    if (which == GMG_XB_ROWS_ID || which == GMG_XB_ROWS_VAL || which == GMG_XB_P_COL || which == GMG_XB_P_VAL) gather_var(c, s, n, which);
    else gather_fixed(c, s, n, which);
}

int64_t phase(gmg_ctx* c, gmg_shard* s, int ph, int64_t* io) {
    const int rc = gmg_shard_phase(c, s, ph, io);
    if (rc != GMG_OK) throw Error{rc, c->last_error};
    return io[0];
}

}  // namespace

extern "C" {

int gmg_nccl_unique_id(void* out128) {
    if (!out128) return GMG_ERR_INVALID;
    NcclApi& api = nccl();
    if (!api.error.empty()) return GMG_ERR_NCCL;
    NcclUniqueId id;
    if (api.GetUniqueId(&id) != 0) return GMG_ERR_NCCL;
    memcpy(out128, id.internal, sizeof id.internal);
    return GMG_OK;
}

int gmg_nccl_init(gmg_ctx* c, const void* id128, int rank, int world, gmg_nccl** out) {
    if (!c || !out || !id128) return GMG_ERR_INVALID;
    *out = nullptr;
    try {
        NcclApi& api = nccl();
        if (!api.error.empty()) throw Error{GMG_ERR_NCCL, api.error};
        GMG_REQUIRE(world >= 1 && rank >= 0 && rank < world, GMG_ERR_INVALID, "gmg_nccl_init: bad rank / world");
        GMG_CUDA(cudaSetDevice(c->device));
        std::unique_ptr<gmg_nccl> n(new gmg_nccl());
        n->rank = rank; n->world = world;
        NcclUniqueId id;
        memcpy(id.internal, id128, sizeof id.internal);
        nccl_check(api.CommInitRank(&n->comm, world, id, rank), "ncclCommInitRank");
        *out = n.release();
        c->last_error.clear();
        return GMG_OK;
    } catch (const Error& e) {
        c->last_error = e.what;
        return e.status;
    }
}

void gmg_nccl_destroy(gmg_nccl* n) {
    if (!n) return;
    if (n->comm && nccl().CommDestroy) nccl().CommDestroy(n->comm);
    delete n;
}

/* The whole row-sharded build of `s` (created with gmg_shard_create, same options on every rank): the driver loop of
 * the section above, exchanges over `comm` on the context's stream. stats (optional): {exchanges, bytes gathered}. */
int gmg_shard_build_nccl(gmg_ctx* c, gmg_shard* s, gmg_nccl* n, int build_galerkin, int64_t* stats) {
    if (!c || !s || !n) return GMG_ERR_INVALID;
    try {
        GMG_CUDA(cudaSetDevice(c->device));
        int64_t io[8];
        while (phase(c, s, GMG_PH_LEVEL_BEGIN, io)) {
            gather(c, s, n, GMG_XB_LEN_PARTIAL);
            phase(c, s, GMG_PH_RADIUS, io);
            phase(c, s, GMG_PH_SAMPLER_BUILD, io);
            gather(c, s, n, GMG_XB_STATE);
            while (phase(c, s, GMG_PH_SAMPLER_CHECK, io) > 0) {
                phase(c, s, GMG_PH_SAMPLER_ROUND, io);
                gather(c, s, n, GMG_XB_STATE);
            }
            phase(c, s, GMG_PH_SAMPLES, io);
            do {
                phase(c, s, GMG_PH_PARENTS_PULL, io);
                gather(c, s, n, GMG_XB_DIST);
                gather(c, s, n, GMG_XB_CHANGED);
            } while (phase(c, s, GMG_PH_PARENTS_CHECK, io) != 0);
            phase(c, s, GMG_PH_PARENTS_PRED, io);
            gather(c, s, n, GMG_XB_PRED);
            phase(c, s, GMG_PH_PARENTS_RESOLVE, io);
            gather(c, s, n, GMG_XB_PARENTS);
            phase(c, s, GMG_PH_COARSE_ROWS, io);
            gather(c, s, n, GMG_XB_COARSE_COUNT);
            phase(c, s, GMG_PH_COARSE_PACK, io);
            gather(c, s, n, GMG_XB_ROWS_ID);
            gather(c, s, n, GMG_XB_ROWS_VAL);
            phase(c, s, GMG_PH_COARSE_ASSEMBLE, io);
            phase(c, s, GMG_PH_PROLONG, io);
            if (build_galerkin) {
                phase(c, s, GMG_PH_P_COUNT, io);
                gather(c, s, n, GMG_XB_P_COUNT);
                phase(c, s, GMG_PH_P_PACK, io);
                gather(c, s, n, GMG_XB_P_COL);
                gather(c, s, n, GMG_XB_P_VAL);
                phase(c, s, GMG_PH_GALERKIN_ROWS, io);
                gather(c, s, n, GMG_XB_COARSE_COUNT);
                phase(c, s, GMG_PH_GALERKIN_PACK, io);
                gather(c, s, n, GMG_XB_ROWS_ID);
                gather(c, s, n, GMG_XB_ROWS_VAL);
                phase(c, s, GMG_PH_GALERKIN_ASSEMBLE, io);
            }
            phase(c, s, GMG_PH_LEVEL_STATUS, io);
            gather(c, s, n, GMG_XB_STATUS);
            phase(c, s, GMG_PH_LEVEL_END, io);
        }
        if (stats) { stats[0] = n->exchanges; stats[1] = n->bytes; }
        c->last_error.clear();
        return GMG_OK;
    } catch (const Error& e) {
        c->last_error = e.what;
        abort_sync(c);
        return e.status;
    }
}

}  // extern "C"
