// Multi-GPU exchange at the C ABI: one process (or thread) per GPU, NCCL as the transport.  The reference's
// aggregations (Aggregator.merge, treeAggregate, reduceByKey) become sums / minima / maxima of
// small arrays -- partial sums, count tables, per-bucket sums.  NCCL is loaded at run time (libnccl.so.2, the copy a
// host process such as PyTorch has already mapped if there is one), so the library itself does not depend on it.
#include <dlfcn.h>
#include <string.h>

#include <new>

#include "common.cuh"

namespace {

struct NcclId {
    char bytes[128];
};
typedef void* NcclComm;
typedef int (*GetUniqueIdFn)(NcclId*);
typedef int (*CommInitRankFn)(NcclComm*, int, NcclId, int);
typedef int (*AllReduceFn)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*CommDestroyFn)(NcclComm);
typedef const char* (*GetErrorStringFn)(int);

struct NcclApi {
    void* handle = nullptr;
    GetUniqueIdFn get_unique_id = nullptr;
    CommInitRankFn comm_init_rank = nullptr;
    AllReduceFn all_reduce = nullptr;
    CommDestroyFn comm_destroy = nullptr;
    GetErrorStringFn error_string = nullptr;
};

NcclApi g_nccl;

int load_nccl() {
    if (g_nccl.handle) return TMOG_OK;
    const char* override_path = getenv("TMOG_NCCL_LIB");
    void* h = dlopen(override_path ? override_path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    TMOG_REQUIRE(h != nullptr, "libnccl.so.2 cannot be loaded (%s); set TMOG_NCCL_LIB", dlerror());
    NcclApi a;
    a.handle = h;
    a.get_unique_id = (GetUniqueIdFn)dlsym(h, "ncclGetUniqueId");
    a.comm_init_rank = (CommInitRankFn)dlsym(h, "ncclCommInitRank");
    a.all_reduce = (AllReduceFn)dlsym(h, "ncclAllReduce");
    a.comm_destroy = (CommDestroyFn)dlsym(h, "ncclCommDestroy");
    a.error_string = (GetErrorStringFn)dlsym(h, "ncclGetErrorString");
    TMOG_REQUIRE(a.get_unique_id && a.comm_init_rank && a.all_reduce && a.comm_destroy, "libnccl.so.2 lacks an entry point");
    g_nccl = a;
    return TMOG_OK;
}

#define TMOG_NCCL(call)                                                                          \
    do {                                                                                         \
        const int _r = (call);                                                                   \
        if (_r != 0) {                                                                           \
            ::tmog::set_error("NCCL error %d (%s)", _r, g_nccl.error_string ? g_nccl.error_string(_r) : "?"); \
            return TMOG_ECUDA;                                                                   \
        }                                                                                        \
    } while (0)

}  // namespace

struct tmog_comm {
    NcclComm comm = nullptr;
    int rank = 0, size = 1, device = 0;
    cudaStream_t stream = nullptr;
    void* staging = nullptr;
    size_t staging_cap = 0;
};

extern "C" int tmog_comm_unique_id(uint8_t* id128) {
    TMOG_REQUIRE(id128 != nullptr, "id is NULL");
    if (int rc = load_nccl()) return rc;
    NcclId id;
    TMOG_NCCL(g_nccl.get_unique_id(&id));
    memcpy(id128, id.bytes, sizeof(id.bytes));
    return TMOG_OK;
}

extern "C" int tmog_comm_init(const uint8_t* id128, int rank, int size, int device, tmog_comm** out) {
    TMOG_REQUIRE(id128 != nullptr && out != nullptr, "NULL argument");
    TMOG_REQUIRE(size >= 1 && rank >= 0 && rank < size, "bad rank %d of %d", rank, size);
    if (int rc = load_nccl()) return rc;
    TMOG_CUDA(cudaSetDevice(device));
    tmog_comm* c = new (std::nothrow) tmog_comm();
    TMOG_REQUIRE(c != nullptr, "out of host memory");
    c->rank = rank;
    c->size = size;
    c->device = device;
    NcclId id;
    memcpy(id.bytes, id128, sizeof(id.bytes));
    TMOG_NCCL(g_nccl.comm_init_rank(&c->comm, size, id, rank));
    TMOG_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    *out = c;
    return TMOG_OK;
}

extern "C" int tmog_comm_destroy(tmog_comm* c) {
    if (c == nullptr) return TMOG_OK;
    cudaSetDevice(c->device);
    if (c->comm) g_nccl.comm_destroy(c->comm);
    if (c->staging) cudaFree(c->staging);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return TMOG_OK;
}

// dtype: 0 = int64, 1 = float64;  op: 0 = sum, 1 = max, 2 = min
static int nccl_codes(int dtype, int op, int* dt, int* ro) {
    TMOG_REQUIRE(dtype == 0 || dtype == 1, "dtype must be 0 (int64) or 1 (float64)");
    TMOG_REQUIRE(op >= 0 && op <= 2, "op must be 0 (sum), 1 (max) or 2 (min)");
    *dt = dtype == 0 ? 4 /* ncclInt64 */ : 8 /* ncclFloat64 */;
    *ro = op == 0 ? 0 /* ncclSum */ : op == 1 ? 2 /* ncclMax */ : 3 /* ncclMin */;
    return TMOG_OK;
}

extern "C" int tmog_comm_allreduce_dev(tmog_comm* c, void* buf, int64_t count, int dtype, int op, void* stream) {
    TMOG_REQUIRE(c != nullptr && (buf != nullptr || count == 0) && count >= 0, "bad arguments");
    if (count == 0 || c->size == 1) return TMOG_OK;
    int dt, ro;
    if (int rc = nccl_codes(dtype, op, &dt, &ro)) return rc;
    TMOG_NCCL(g_nccl.all_reduce(buf, buf, (size_t)count, dt, ro, c->comm, (cudaStream_t)stream));
    return TMOG_OK;
}

extern "C" int tmog_comm_allreduce_host(tmog_comm* c, void* buf, int64_t count, int dtype, int op) {
    TMOG_REQUIRE(c != nullptr && (buf != nullptr || count == 0) && count >= 0, "bad arguments");
    if (count == 0 || c->size == 1) return TMOG_OK;
    TMOG_CUDA(cudaSetDevice(c->device));
    const size_t bytes = (size_t)count * 8;
    if (bytes > c->staging_cap) {
        if (c->staging) cudaFree(c->staging);
        c->staging = nullptr;
        c->staging_cap = 0;
        TMOG_CUDA(cudaMalloc(&c->staging, bytes));
        c->staging_cap = bytes;
    }
    TMOG_CUDA(cudaMemcpyAsync(c->staging, buf, bytes, cudaMemcpyHostToDevice, c->stream));
    if (int rc = tmog_comm_allreduce_dev(c, c->staging, count, dtype, op, c->stream)) return rc;
    TMOG_CUDA(cudaMemcpyAsync(buf, c->staging, bytes, cudaMemcpyDeviceToHost, c->stream));
    TMOG_CUDA(cudaStreamSynchronize(c->stream));
    return TMOG_OK;
}
