// engine_state.cuh -- what the engine holds: the model on each device, slots (a stream with its private
// activation buffers), per-device worker queues, tickets, and struct ctr_engine itself.
#pragma once
namespace {
// ------------------------------------------------------------------ per-device model state
struct GemmLayer {
    int N = 0, Npad = 0, Kpad = 0;
    __half* W = nullptr;      // [Npad, Kpad]
    float* bias = nullptr;    // [Npad]
};

struct DeviceModel {
    int dev = -1;
    int num_sms = 148;
    uint8_t* table = nullptr;          // packed rows (local shard when row-sharded)
    uint64_t table_rows_local = 0;
    size_t table_bytes = 0;
    std::vector<GemmLayer> mlp, cross;
    float* w_out_h = nullptr;          // [Npad of last hidden] head weights for the hidden part
    float* w_out_x = nullptr;          // [d_pad] head weights for x_L (DCN-v2)
    float b0 = 0.f, b_out = 0.f;
    bool loaded = false;
};

// How the features of a request lie in memory (the same description serves the host side of a shard and the
// device side of a chunk).  Reference layout (DC:97-108): ids int64 [n, F], wts float [n, F].  Compact layouts
// (ctr_predict_shared_async / ctr_predict_packed_async): the first n_ctx fields once per request (host lists),
// per candidate the remaining Fk ids as int64 or uint32 and Wk weights (only the fields in weighted_mask).
struct InputView {
    const void* ids = nullptr;
    const float* wts = nullptr;
    int Fk = 0, Wk = 0, id_bytes = 8, n_ctx = 0;
    bool ids_are_rows = false;               // values index the table directly (hashed by the host packer already)
    uint64_t weighted_mask = ~0ull;
    uint64_t id_mask = ~0ull;                // bit f: per-candidate field f has an id column (clear: its id is given once, in ctx_ids)
    const int64_t* ctx_ids = nullptr;        // host memory: n_ctx context ids, then the ids of the fixed-id fields in field order
    const float* ctx_wts = nullptr;
    const uint8_t* table_override = nullptr; // the rows to gather from instead of the device's table (exchange path:
    uint64_t rows_override = 0;              //   the rows the owners sent back), and how many there are
    bool reference_layout(int F) const { return n_ctx == 0 && id_bytes == 8 && Fk == F && Wk == F && !ids_are_rows; }
    InputView at(int64_t c0) const {         // the same view, starting at candidate c0
        InputView v = *this;
        v.ids = static_cast<const uint8_t*>(ids) + (size_t)c0 * Fk * id_bytes;
        v.wts = wts ? wts + (size_t)c0 * Wk : nullptr;
        return v;
    }
    size_t ids_bytes(int64_t n) const { return (size_t)n * Fk * id_bytes; }
    size_t wts_bytes(int64_t n) const { return (size_t)n * Wk * sizeof(float); }
};

InputView reference_view(const void* ids, const float* wts, int F) {
    InputView v;
    v.ids = ids; v.wts = wts; v.Fk = F; v.Wk = F;
    return v;
}

// One launch of tower_chain_kernel: a sequence of layers, each feeding the next.
struct Chain {
    int BN = 256;
    int n_layers = 0;
    ChainMaps maps;
    ChainParams params;       // template: M / score pointer are filled per launch
};

// Device buffers of one sort (csrc/sort.cuh): capacity `cap` candidates.
struct SortBufs {
    int64_t cap = 0;
    float* sorted = nullptr;
    uint32_t* keys[2] = {nullptr, nullptr};
    int32_t* idx[2] = {nullptr, nullptr};
    uint32_t* counts = nullptr;
};

// One stream with its private activation buffers ("slot"): everything a shard needs on the device.
// The gather and the tower on disjoint sets of SMs (sm_split.cuh); off unless CTR_SM_SPLIT asks for it.
struct SmSplit {
    bool on = false;
    int gather_sms = 0, tower_sms = 0;
    CUgreenCtx gather_ctx = nullptr, tower_ctx = nullptr;
};

struct Slot {
    int dev = -1;
    cudaStream_t stream = nullptr;
    // SM split: the gather runs on g_stream (the gather's SMs) after ev_in, the chains on t_stream (the tower's SMs)
    // after ev_g, and `stream` continues after ev_t -- callers of score_chunk still only see `stream`
    cudaStream_t g_stream = nullptr, t_stream = nullptr;
    cudaEvent_t ev_in = nullptr, ev_g = nullptr, ev_t = nullptr;
    bool owns_stream = true;
    bool owns_buffers = true;  // false: the buffers are slices of the fixed scratch of ctr_score_device_static
    int64_t cap = 0;           // candidate capacity of the id/weight/score buffers
    int x_rows = 0;            // candidates per launch: rows of X / wide / extra (multiple of 256)
    int scratch_rows = 0;      // resident CTA pairs x 256: bounds the pairs of a launch
    int scratch_tiles = 0;     // 128-row tiles each scratch buffer holds
    int64_t* d_ids = nullptr; float* d_wts = nullptr; float* d_scores = nullptr;   // d_ids / d_wts also take the compact layouts
    float* d_wide = nullptr; float* d_extra = nullptr; int32_t* d_err = nullptr;
    __half* buf[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int ld[6] = {0, 0, 0, 0, 0, 0};
    std::vector<Chain> chains;
    unsigned long long* slice_bitmap = nullptr;    // the device's PairScratch bitmap (nullptr: slices by pair index)
    int n_slices = 0;
    // pinned staging for pageable callers
    int64_t* h_ids = nullptr; float* h_wts = nullptr; float* h_scores = nullptr; int32_t* h_err = nullptr;
    SortBufs srt;              // host-request slots: the sort of a one-shard request runs on this slot's stream
    int32_t* h_order = nullptr;
    // stage timing: a ring of event sets, one set per timed call (first chunk of the call only)
    static constexpr int kTimeSets = 32, kTimeMarks = 12;
    cudaEvent_t (*tev)[kTimeMarks] = nullptr;      // [kTimeSets][kTimeMarks], created when timing is switched on
    int mark_stage[kTimeMarks]; int n_marks = 0;   // stage id of mark i (same sequence every call)
    int cur_mark = 0; int sets_recorded = 0; bool time_this = false;
    cudaEvent_t join_ev = nullptr;
};

struct Request {
    std::atomic<int> pending{0};
    std::atomic<int> status{CTR_OK};
    std::string error;
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    std::atomic<bool> done_flag{false};   // same as `done`, readable without the mutex (ctr_wait polls it briefly first)
    ctr_callback cb = nullptr;
    void* user = nullptr;
    ctr_ticket ticket = 0;
};

struct Task {
    std::shared_ptr<Request> req;
    InputView in;                 // this shard's features in the caller's (host) memory
    int64_t n = 0; float* scores = nullptr;
    // A6 (DC:195) on the shard's own stream: scores come back sorted, order_out (may be null) gets the
    // candidate index of each.  Only set on requests that run as ONE shard.
    bool sort = false; int32_t* order_out = nullptr;
    float* d_merge = nullptr;     // a multi-shard sorted request: this shard's slice of the merge buffer on the engine's first GPU
    bool stop = false;
};

// The activation scratch of the tower (tower.cuh): ONE allocation per device, cut into num_sms / 2 slices of
// 2 x 256 rows.  Every slot (stream) of the device works out of it; the tower's CTA pairs take and return slices
// themselves, so the footprint that cycles through L2 does not grow with the number of streams.
struct PairScratch {
    uint8_t* base = nullptr;           // one cudaMalloc: [bitmap 256 B | buf 1 | buf 2 | buf 3 | buf 4 | buf 5]
    size_t bytes = 0;
    unsigned long long* bitmap = nullptr;
    int n_slices = 0;
    __half* buf[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int ld[6] = {0, 0, 0, 0, 0, 0};
};

struct DeviceCtx {
    DeviceModel model;
    PairScratch scratch;
    SmSplit split;
    size_t l2_persist_bytes = 0, l2_window_max = 0;    // persisting-L2 set-aside of the device / largest access window
    std::vector<std::unique_ptr<Slot>> worker_slots;
    std::vector<std::unique_ptr<Slot>> dev_slots;     // for ctr_score_device
    std::mutex dev_mu;                                // serialises ctr_score_device enqueues per device
    cudaStream_t main_stream = nullptr;
    cudaEvent_t fork_ev = nullptr;
    // device sort scratch (ctr_predict_sorted), allocated on first use
    std::mutex sort_mu;
    int64_t sort_cap = 0;
    float* srt_scores = nullptr;
    SortBufs srt;
    std::deque<Task> queue;
    std::atomic<int> active{0};                       // shard tasks being processed on this device right now (sizes small tower launches)
    std::atomic<int> queued{0};                       // queue.size(), readable without q_mu (workers poll it briefly before sleeping)
    std::mutex q_mu;
    std::condition_variable q_cv;
    std::vector<std::thread> threads;
};

}  // namespace

constexpr size_t kKeepCalledBack = 4096;

struct ctr_engine {
    ctr_config cfg;
    ModelDesc m;
    bool loaded = false;
    bool timing = false;
    int chunk_rows = 0;          // 0 = num_sms * 128
    std::vector<std::unique_ptr<DeviceCtx>> devs;
    std::mutex t_mu;
    std::unordered_map<ctr_ticket, std::shared_ptr<Request>> tickets;
    std::deque<ctr_ticket> called_back;     // completed requests that had a callback, oldest first (guarded by t_mu)
    std::atomic<int64_t> next_ticket{1};
    std::atomic<int64_t> launches{0};
    std::atomic<int64_t> h2d_bytes{0}, d2h_bytes{0};    // request path: bytes that crossed the host link
    std::mutex pin_mu;
    std::map<uintptr_t, size_t> pinned;     // base -> bytes (buffers from ctr_alloc_pinned)
    std::atomic<unsigned> next_first_device{0};   // rotation of the first GPU of requests that use fewer than all
    std::atomic<bool> shutting_down{false};
    uint8_t* shard_base[CTR_MAX_DEVICES] = {nullptr};   // row-sharded tables: every device's shard
    std::atomic<uint64_t> narrow_weighted{0};            // host narrowing: fields seen with a weight != 1.0 so far (sticky)
    int host_pack = 0;                                   // cfg.host_pack, CTR_HOST_PACK overrides
    ExchangeState exch;                                  // NCCL all-to-all variant of the row-sharded lookup (exchange.cuh)
};

namespace {

// DC:46-55 applied to candidates
void split_bounds(int64_t n, int parts, int64_t* b, int64_t* e) {
    const int64_t range = n / parts;
    for (int i = 0; i < parts - 1; ++i) { b[i] = i * range; e[i] = (i + 1) * range; }
    b[parts - 1] = (int64_t)(parts - 1) * range;
    e[parts - 1] = n;
}

bool is_pinned(ctr_engine* e, const void* p, size_t bytes) {
    std::lock_guard<std::mutex> lk(e->pin_mu);
    if (e->pinned.empty()) return false;
    auto it = e->pinned.upper_bound((uintptr_t)p);
    if (it == e->pinned.begin()) return false;
    --it;
    return (uintptr_t)p >= it->first && (uintptr_t)p + bytes <= it->first + it->second;
}

}  // namespace
