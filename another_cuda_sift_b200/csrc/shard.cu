// Query-row sharded brute-force matching over the GPUs of one box (BASELINE.json configs[4]).
//
// The semantic that is sharded is sift_cuda::matchBruteForce (sift_func/Match.cu:8-33, kernel :35-177): the result of a
// query row needs EVERY train column (the reference's 32 lane chunks and their merge, quirk Q15, run over the whole
// train set), so the train set is replicated and the QUERY ROWS are split: rank r owns rows [r*per, min(nq, (r+1)*per)),
// per = ceil(nq / world). One call enqueues, on ONE stream and without any host synchronisation,
//     ncclBroadcast   of the train set from its owner (skipped when every rank already holds it)
//     the matcher     (tcgen05 filter + exact re-rank, match_tc.cu) on this rank's rows
//     ncclAllGather   of the int32 indices (`per` per rank, -1 padded)
// so the collectives queue directly behind / in front of the kernels they depend on.
//
// NCCL is resolved at run time (dlopen of libnccl.so.2: in a torch process that is the copy torch already loaded), so
// libb200sift.so itself has no link-time dependency on it and single-GPU users never touch it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/b200sift.h"
#include "common.cuh"

namespace b200sift {
int set_error(int code, const char* fmt, ...);  // engine.cu
}
using b200sift::set_error;

namespace {

struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    std::string error;
};

NcclApi* nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) {
            api.error = std::string("cannot load NCCL (libnccl.so.2): ") + (dlerror() ? dlerror() : "not found");
            return;
        }
        auto sym = [&](const char* n) -> void* {
            void* p = dlsym(api.lib, n);
            if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + n;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
    });
    return &api;
}

#define NCCLCHECK(call)                                                                                          \
    do {                                                                                                         \
        ncclResult_t r_ = (call);                                                                                \
        if (r_ != ncclSuccess) return set_error(B200SIFT_ECUDA, "%s:%d NCCL: %s", __FILE__, __LINE__, api->GetErrorString(r_)); \
    } while (0)
#define CUCHECK(call)                                                                                            \
    do {                                                                                                         \
        cudaError_t e_ = (call);                                                                                 \
        if (e_ != cudaSuccess) return set_error(B200SIFT_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, cudaGetErrorName(e_), cudaGetErrorString(e_)); \
    } while (0)

__global__ void fill_i32_kernel(int* p, int n, int v) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

}  // namespace

struct b200sift_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
    void* scratch = nullptr;       // matcher scratch, private to this communicator (no sharing with b200sift_match_device)
    size_t scratch_bytes = 0;
    int* local = nullptr;          // this rank's block of indices (per ints)
    size_t local_cap = 0;
};

extern "C" {

int b200sift_nccl_unique_id(uint8_t* id128) {
    if (!id128) return set_error(B200SIFT_EINVAL, "null argument");
    NcclApi* api = nccl();
    if (!api->error.empty()) return set_error(B200SIFT_ECUDA, "%s", api->error.c_str());
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NCCLCHECK(api->GetUniqueId(&id));
    std::memcpy(id128, &id, sizeof(id));
    return 0;
}

int b200sift_comm_create(const uint8_t* id128, int rank, int world, b200sift_comm** out) {
    if (!id128 || !out || world < 1 || rank < 0 || rank >= world) return set_error(B200SIFT_EINVAL, "bad argument");
    *out = nullptr;
    NcclApi* api = nccl();
    if (!api->error.empty()) return set_error(B200SIFT_ECUDA, "%s", api->error.c_str());
    b200sift_comm* c = new b200sift_comm;
    c->rank = rank;
    c->world = world;
    if (cudaGetDevice(&c->device) != cudaSuccess) {
        delete c;
        return set_error(B200SIFT_ECUDA, "no CUDA device: this engine has no CPU fallback");
    }
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    ncclResult_t r = api->CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
        delete c;
        return set_error(B200SIFT_ECUDA, "ncclCommInitRank: %s", api->GetErrorString(r));
    }
    *out = c;
    return 0;
}

int b200sift_comm_destroy(b200sift_comm* c) {
    if (!c) return 0;
    cudaDeviceSynchronize();
    if (c->comm) nccl()->CommDestroy(c->comm);
    if (c->scratch) cudaFree(c->scratch);
    if (c->local) cudaFree(c->local);
    delete c;
    return 0;
}

int b200sift_comm_info(const b200sift_comm* c, int* rank, int* world, int* nccl_version) {
    if (!c) return set_error(B200SIFT_EINVAL, "null argument");
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    if (nccl_version) {
        *nccl_version = 0;
        NcclApi* api = nccl();
        if (api->GetVersion) api->GetVersion(nccl_version);
    }
    return 0;
}

int b200sift_shard_rows(int n_rows, int world, int rank, int* begin, int* end) {
    if (world < 1 || rank < 0 || rank >= world || n_rows < 0) return set_error(B200SIFT_EINVAL, "bad argument");
    const int per = n_rows > 0 ? (n_rows + world - 1) / world : 0;
    const int lo = std::min(n_rows, rank * per);
    if (begin) *begin = lo;
    if (end) *end = std::min(n_rows, lo + per);
    return per;
}

}  // extern "C"

// ev (profiling only): six events -- start, after the broadcast, after norms, after the filter, after the re-rank, after the all-gather
static int match_sharded_impl(b200sift_comm* c, const uint16_t* d_query_block, int nq_total, uint16_t* d_train, int nt, int train_root,
                              int* d_out_all, void* stream, cudaEvent_t* ev) {
    if (!c || nq_total < 0 || nt < 0 || (nq_total > 0 && !d_out_all) || (nt > 0 && !d_train) || train_root >= c->world)
        return set_error(B200SIFT_EINVAL, "bad argument");
    if (nq_total == 0) return 0;
    NcclApi* api = nccl();
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int lo = 0, hi = 0;
    const int per = b200sift_shard_rows(nq_total, c->world, c->rank, &lo, &hi);
    const int mine = hi - lo;
    if (mine > 0 && !d_query_block) return set_error(B200SIFT_EINVAL, "this rank owns %d query rows but d_query_block is null", mine);
    // buffers of this communicator, grown outside the enqueue path when the problem grows
    const size_t need = b200sift::match_scratch_bytes(std::max(mine, 1), nt);
    if (c->scratch_bytes < need) {
        CUCHECK(cudaStreamSynchronize(s));
        if (c->scratch) CUCHECK(cudaFree(c->scratch));
        c->scratch = nullptr;
        c->scratch_bytes = 0;
        CUCHECK(cudaMalloc(&c->scratch, need));
        c->scratch_bytes = need;
    }
    if (c->local_cap < (size_t)per) {
        CUCHECK(cudaStreamSynchronize(s));
        if (c->local) CUCHECK(cudaFree(c->local));
        c->local = nullptr;
        c->local_cap = 0;
        CUCHECK(cudaMalloc(&c->local, (size_t)per * sizeof(int)));
        c->local_cap = (size_t)per;
    }
    if (ev) CUCHECK(cudaEventRecord(ev[0], s));
    // 1. the train set travels from its owner to everybody (25.6 MB at 100k rows), in place in d_train
    if (train_root >= 0 && c->world > 1 && nt > 0)
        NCCLCHECK(api->Broadcast(d_train, d_train, (size_t)nt * 128, ncclFloat16, train_root, c->comm, s));
    if (ev) CUCHECK(cudaEventRecord(ev[1], s));
    // 2. this rank's rows; rows of the padded block beyond the query set report -1
    int* dst = c->world > 1 ? c->local : d_out_all;
    if (mine < per) fill_i32_kernel<<<(per - mine + 255) / 256, 256, 0, s>>>(dst + mine, per - mine, -1);
    if (mine > 0) {
        const int rc = b200sift::match_launch(reinterpret_cast<const __half*>(d_query_block), mine, reinterpret_cast<const __half*>(d_train), nt,
                                              dst, c->scratch, c->scratch_bytes, s, nullptr, ev ? ev + 2 : nullptr);
        if (rc) return set_error(B200SIFT_EINVAL, rc == -2 ? "train descriptors: tensor map rejected (the pointer must be 16-byte aligned)" : "matcher scratch too small");
    }
    CUCHECK(cudaGetLastError());
    // 3. everybody gets every block: world * per ints (the first nq_total are the result)
    if (ev && mine <= 0)
        for (int i = 2; i < 5; i++) CUCHECK(cudaEventRecord(ev[i], s));
    if (c->world > 1) NCCLCHECK(api->AllGather(c->local, d_out_all, (size_t)per, ncclInt32, c->comm, s));
    if (ev) CUCHECK(cudaEventRecord(ev[5], s));
    return 0;
}

extern "C" {

int b200sift_match_sharded(b200sift_comm* c, const uint16_t* d_query_block, int nq_total, uint16_t* d_train, int nt, int train_root,
                           int* d_out_all, void* stream) {
    return match_sharded_impl(c, d_query_block, nq_total, d_train, nt, train_root, d_out_all, stream, nullptr);
}

int b200sift_match_sharded_profile(b200sift_comm* c, const uint16_t* d_query_block, int nq_total, uint16_t* d_train, int nt, int train_root,
                                   int* d_out_all, void* stream, float* ms5) {
    if (!ms5) return set_error(B200SIFT_EINVAL, "null argument");
    cudaEvent_t ev[6];
    for (auto& e : ev) CUCHECK(cudaEventCreate(&e));
    int rc = match_sharded_impl(c, d_query_block, nq_total, d_train, nt, train_root, d_out_all, stream, ev);
    if (rc == 0) {
        CUCHECK(cudaEventSynchronize(ev[5]));
        for (int i = 0; i < 5; i++) CUCHECK(cudaEventElapsedTime(&ms5[i], ev[i], ev[i + 1]));
    }
    for (auto& e : ev) cudaEventDestroy(e);
    return rc;
}

}  // extern "C"
