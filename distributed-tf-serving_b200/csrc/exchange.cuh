// exchange.cuh -- BASELINE config 4 as north star (d) words it: row-sharded tables with an NCCL all-to-all over NVLink.
//
// The reference fans a request out to model replicas (DCNClient.java:148-156) and has no sharded tables; this is the
// new functionality SURVEY.md section 8(e) describes, built next to the peer-load variant (gather kernels with
// kSharded: remote rows fetched by plain loads through NVLink peer mappings) so that the two can be raced:
//
//   bucket   every device turns its chunk's (candidate, field) ids into rows, counts them per owner (row mod G),
//            and writes the LOCAL row indices bucket by bucket; it remembers in which slot of the reply each
//            position's row will come back
//   ids      all-to-all #1 (ncclSend / ncclRecv inside one group): 4 bytes per lookup
//   serve    every owner copies the rows it was asked for out of its own shard, at local HBM speed
//   rows     all-to-all #2: 128 bytes (fp32 rows) or 64 bytes (f16 rows) per lookup
//   score    the ordinary gather kernel (compact layout: uint32 rows + weights) over the returned rows as its
//            "table", then the tower -- the same kernels, so the scores are the bits the replicated engine gives
//
// One host thread drives all G devices of the engine in lockstep (single-process NCCL, ncclCommInitAll).  NCCL is
// bound at run time (dlopen of libnccl.so.2: the copy torch already loaded when called from Python), so the
// library has no link-time dependency on it and every other entry point works without it.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return a;
        a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(dlsym(h, "ncclCommInitAll"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(dlsym(h, "ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(dlsym(h, "ncclGroupEnd"));
        a.Send = reinterpret_cast<decltype(a.Send)>(dlsym(h, "ncclSend"));
        a.Recv = reinterpret_cast<decltype(a.Recv)>(dlsym(h, "ncclRecv"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        a.ok = a.CommInitAll && a.CommDestroy && a.GroupStart && a.GroupEnd && a.Send && a.Recv && a.GetErrorString;
        return a;
    }();
    return api;
}

#define NCCL_TRY(expr)                                                                                     \
    do {                                                                                                   \
        ncclResult_t _r = (expr);                                                                          \
        if (_r != ncclSuccess) return fail(CTR_ERR_CUDA, std::string(#expr) + ": " + nccl_api().GetErrorString(_r)); \
    } while (0)

constexpr uint32_t kNoRow = 0xFFFFFFFFu;

// ---- bucket: count the lookups of a chunk per owner.  counts[G]
__global__ void exch_count_kernel(const int64_t* ids, int64_t n_elems, uint64_t num_rows, int hashed, int G, uint32_t* counts) {
    __shared__ uint32_t local[CTR_MAX_DEVICES];
    if (threadIdx.x < CTR_MAX_DEVICES) local[threadIdx.x] = 0;
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_elems; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t id = (uint64_t)ids[i];
        const uint64_t row = hashed ? (mix64(id) & (num_rows - 1)) : id;
        if (row < num_rows) atomicAdd(&local[row % (uint32_t)G], 1u);
    }
    __syncthreads();
    if (threadIdx.x < G && local[threadIdx.x]) atomicAdd(&counts[threadIdx.x], local[threadIdx.x]);
}

// ---- bucket: write each lookup's local row index into its owner's bucket (cursor[G] starts at the bucket offsets)
// and remember the reply slot of every position (kNoRow: the id is outside a direct table)
__global__ void exch_scatter_kernel(const int64_t* ids, int64_t n_elems, uint64_t num_rows, int hashed, int G, uint32_t* cursor,
                                    uint32_t* send_rows, uint32_t* slot_of_pos) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_elems; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t id = (uint64_t)ids[i];
        const uint64_t row = hashed ? (mix64(id) & (num_rows - 1)) : id;
        uint32_t slot = kNoRow;
        if (row < num_rows) {
            slot = atomicAdd(&cursor[row % (uint32_t)G], 1u);
            send_rows[slot] = (uint32_t)(row / (uint32_t)G);
        }
        slot_of_pos[i] = slot;
    }
}

// ---- serve: copy the requested rows out of this device's shard (16 bytes per thread, coalesced within a row)
__global__ void exch_serve_kernel(const uint8_t* table, int row_shift, const uint32_t* req_rows, uint32_t n_req, uint8_t* reply) {
    const int per_row = 1 << (row_shift - 4);                 // 16-byte pieces per row: 8 (fp32) or 4 (f16)
    const uint64_t total = (uint64_t)n_req * per_row;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(i / per_row), piece = (uint32_t)(i % per_row);
        const uint4 v = *reinterpret_cast<const uint4*>(table + ((uint64_t)req_rows[r] << row_shift) + piece * 16);
        *reinterpret_cast<uint4*>(reply + ((uint64_t)r << row_shift) + piece * 16) = v;
    }
}

// Buffers of one device for the exchange.  What the score stage reads (slot_of_pos, got) exists twice: chunk k+1 is
// exchanged while chunk k is scored.
struct ExchangeBufs {
    uint32_t* counts = nullptr;      // [G] lookups this device sends to each owner; then the scatter cursors
    uint32_t* send_rows = nullptr;   // [chunk * F] local row indices, bucket after bucket
    uint32_t* slot_of_pos[2] = {nullptr, nullptr};   // [chunk * F]
    uint32_t* req_rows = nullptr;    // [cap_req] what the peers ask this device for
    uint8_t* reply = nullptr;        // [cap_req] rows served
    uint8_t* got[2] = {nullptr, nullptr};            // [chunk * F] rows that came back: the score stage's table
    uint32_t* h_counts = nullptr;    // pinned [G]
    cudaEvent_t scored[2] = {nullptr, nullptr};      // the score stage is done with buffer set i
    cudaEvent_t exchanged = nullptr;                 // the exchange of the current chunk is complete on this device
};

}  // namespace

struct ExchangeState {
    bool ready = false;
    int chunk = 0;
    size_t cap_req = 0;
    std::vector<ncclComm_t> comms;
    std::vector<ExchangeBufs> bufs;
    std::mutex mu;
};
