// mft_comm_* — NCCL behind the C ABI (SURVEY §8b last row, §8e): the one exchange step of the data-parallel training
// step, so a C / C++ caller drives N ranks without torch.distributed. One process per GPU; rank 0 creates the unique id and
// hands it to the others (through a file here; any side channel works with mft_comm_init_rank). The all-reduce is enqueued
// on the trainer's own stream, between the forward/backward phase that packs the gradients into the bucket and the update
// phase that reads it, so no host synchronisation separates the three.
// libnccl is loaded at run time (dlopen "libnccl.so.2"): a process that already carries one (PyTorch's) shares it, a plain
// C++ process picks up the system library; the rest of libmft_host has no link-time dependency on NCCL.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <unistd.h>

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>

#include "mft_trainer.h"
#include "runtime.h"

namespace {

// the few NCCL declarations used here (nccl.h: ncclUniqueId is 128 opaque bytes; ncclFloat = 7, ncclSum = 0)
struct NcclUniqueId { char internal[128]; };
using ncclComm_t = void*;
using FnGetUniqueId = int (*)(NcclUniqueId*);
using FnCommInitRank = int (*)(ncclComm_t*, int, NcclUniqueId, int);
using FnAllReduce = int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
using FnCommDestroy = int (*)(ncclComm_t);
using FnGetErrorString = const char* (*)(int);
using FnGetVersion = int (*)(int*);

struct Nccl {
    void* handle = nullptr;
    FnGetUniqueId get_unique_id = nullptr;
    FnCommInitRank comm_init_rank = nullptr;
    FnAllReduce all_reduce = nullptr;
    FnCommDestroy comm_destroy = nullptr;
    FnGetErrorString error_string = nullptr;
    FnGetVersion get_version = nullptr;
};

thread_local char g_cerr[512] = "";
int cfail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_cerr, sizeof(g_cerr), fmt, ap);
    va_end(ap);
    return 1;
}

Nccl* nccl() {
    static Nccl n;
    static bool tried = false;
    if (tried) return n.handle ? &n : nullptr;
    tried = true;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
        n.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (n.handle) break;
    }
    if (!n.handle) return nullptr;
    n.get_unique_id = reinterpret_cast<FnGetUniqueId>(dlsym(n.handle, "ncclGetUniqueId"));
    n.comm_init_rank = reinterpret_cast<FnCommInitRank>(dlsym(n.handle, "ncclCommInitRank"));
    n.all_reduce = reinterpret_cast<FnAllReduce>(dlsym(n.handle, "ncclAllReduce"));
    n.comm_destroy = reinterpret_cast<FnCommDestroy>(dlsym(n.handle, "ncclCommDestroy"));
    n.error_string = reinterpret_cast<FnGetErrorString>(dlsym(n.handle, "ncclGetErrorString"));
    n.get_version = reinterpret_cast<FnGetVersion>(dlsym(n.handle, "ncclGetVersion"));
    if (!n.get_unique_id || !n.comm_init_rank || !n.all_reduce || !n.comm_destroy) { dlclose(n.handle); n.handle = nullptr; return nullptr; }
    return &n;
}

const char* nccl_err(Nccl* n, int rc) { return n->error_string ? n->error_string(rc) : "NCCL error"; }

}  // namespace

struct mft_comm {
    ncclComm_t comm = nullptr;
    int world = 1, rank = 0;
};

extern "C" const char* mft_comm_last_error(void) { return g_cerr; }

extern "C" int mft_comm_nccl_version(void) {
    Nccl* n = nccl();
    int v = 0;
    if (n && n->get_version) n->get_version(&v);
    return v;
}

extern "C" int mft_comm_unique_id(void* out128) {
    Nccl* n = nccl();
    if (!n) return cfail("mft_comm_unique_id: libnccl.so.2 could not be loaded: %s", dlerror());
    if (!out128) return cfail("mft_comm_unique_id: null argument");
    NcclUniqueId id;
    const int rc = n->get_unique_id(&id);
    if (rc != 0) return cfail("ncclGetUniqueId: %s", nccl_err(n, rc));
    std::memcpy(out128, &id, sizeof(id));
    return 0;
}

extern "C" int mft_comm_init_rank(mft_comm** out, int world_size, int rank, const void* unique_id_128) {
    Nccl* n = nccl();
    if (!n) return cfail("mft_comm_init_rank: libnccl.so.2 could not be loaded");
    if (!out || !unique_id_128 || world_size < 1 || rank < 0 || rank >= world_size) return cfail("mft_comm_init_rank: bad argument");
    NcclUniqueId id;
    std::memcpy(&id, unique_id_128, sizeof(id));
    auto c = new mft_comm();
    c->world = world_size; c->rank = rank;
    const int rc = n->comm_init_rank(&c->comm, world_size, id, rank);      // binds to the CURRENT CUDA device
    if (rc != 0) { delete c; return cfail("ncclCommInitRank: %s", nccl_err(n, rc)); }
    *out = c;
    return 0;
}

// Rendezvous through a file: rank 0 writes the 128-byte id to `path` (atomically, via rename); the others poll for it.
extern "C" int mft_comm_init_from_file(mft_comm** out, int world_size, int rank, const char* path, int timeout_s) {
    if (!path) return cfail("mft_comm_init_from_file: null path");
    char id[128];
    if (rank == 0) {
        if (int rc = mft_comm_unique_id(id)) return rc;
        const std::string tmp = std::string(path) + ".tmp";
        { std::ofstream f(tmp, std::ios::binary); f.write(id, sizeof(id)); if (!f) return cfail("mft_comm_init_from_file: cannot write %s", tmp.c_str()); }
        if (std::rename(tmp.c_str(), path) != 0) return cfail("mft_comm_init_from_file: cannot publish %s", path);
    } else {
        const auto t0 = std::chrono::steady_clock::now();
        for (;;) {
            std::ifstream f(path, std::ios::binary);
            if (f && f.read(id, sizeof(id)) && f.gcount() == std::streamsize(sizeof(id))) break;
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s)
                return cfail("mft_comm_init_from_file: rank %d timed out waiting for %s", rank, path);
            std::this_thread::sleep_for(std::chrono::milliseconds(20));
        }
    }
    return mft_comm_init_rank(out, world_size, rank, id);
}

extern "C" int mft_comm_world_size(mft_comm* c) { return c ? c->world : 0; }
extern "C" int mft_comm_rank(mft_comm* c) { return c ? c->rank : -1; }

extern "C" int mft_comm_allreduce_sum(mft_comm* c, float* dev_buf, int64_t n, void* stream) {
    Nccl* nc = nccl();
    if (!nc || !c || !c->comm) return cfail("mft_comm_allreduce_sum: no communicator");
    if (n == 0) return 0;
    if (!dev_buf) return cfail("mft_comm_allreduce_sum: null buffer");
    const int rc = nc->all_reduce(dev_buf, dev_buf, size_t(n), /*ncclFloat*/ 7, /*ncclSum*/ 0, c->comm, static_cast<cudaStream_t>(stream));
    if (rc != 0) return cfail("ncclAllReduce: %s", nccl_err(nc, rc));
    return 0;
}

extern "C" int mft_comm_destroy(mft_comm* c) {
    if (!c) return 0;
    Nccl* nc = nccl();
    if (nc && c->comm) nc->comm_destroy(c->comm);
    delete c;
    return 0;
}
