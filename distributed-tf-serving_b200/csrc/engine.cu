// engine.cu -- host runtime + C ABI of the B200-native CTR scoring engine.
//
// What this file replaces in the reference (DC = src/main/java/tensorflow/serving/DCNClient.java):
//   processARequest  DC:137-203  -> ctr_predict_async / ctr_wait: split, fan-out to devices, ordered merge
//   partitionList    DC:46-55    -> split_bounds(): range = n / parts, remainder to the last part
//   sendRequest      DC:91-115   -> Worker::run_task(): pinned staging + cudaMemcpyAsync instead of protobuf
//                                   marshalling, then the kernels instead of the TF-Serving RPC
//   threadPool       DC:42       -> worker threads, one CUDA stream each, per device
//   getChannels      DC:118-125  -> ctr_engine_create (one "channel" per GPU)
// The model arithmetic the reference delegates to an external SavedModel is the CUDA path in
// gather.cuh / tower.cuh.  There is no CPU fallback anywhere.
#include "../../include/ctr_engine.h"
#include "common.cuh"
#include "gather.cuh"
#include "tower.cuh"
#include "sort.cuh"
#include "wire.hpp"

#include <sched.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

using namespace ctr;

thread_local std::string g_last_error;

int32_t fail(int32_t code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define CUDA_TRY(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t _e = (expr);                                                                         \
        if (_e != cudaSuccess) {                                                                         \
            (void)cudaGetLastError();                                                                    \
            return fail(CTR_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));               \
        }                                                                                                \
    } while (0)

inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

// ------------------------------------------------------------------ driver entry point for tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// half matrix [rows, cols] with leading dimension ld (elements), box {64, box_rows}, 128-byte swizzle
int32_t make_map_f16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(CTR_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * sizeof(__half)};
    cuuint32_t box[2] = {(cuuint32_t)kBK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(CTR_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return CTR_OK;
}

// 3-D map over a tile-major activation buffer of `ld` columns (tower.cuh): per 128-row tile the K blocks
// are contiguous 16 KB runs, described as 32 lines of 512 B; no swizzle, so a box lands in shared memory
// byte for byte.  Coordinates: (0, 32 * k_block, tile).
int32_t make_map_tile_major(CUtensorMap* map, const void* base, uint64_t ld, uint64_t n_tiles) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(CTR_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[3] = {256, ld / 2, n_tiles};
    cuuint64_t strides[2] = {512, (cuuint64_t)kBM * ld * sizeof(__half)};
    cuuint32_t box[3] = {256, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(CTR_ERR_CUDA, "cuTensorMapEncodeTiled (tile-major) failed with CUresult " + std::to_string((int)r));
    return CTR_OK;
}

// ------------------------------------------------------------------ weight file (format: DESIGN.md)
constexpr char kMagic[8] = {'C', 'T', 'R', 'W', '0', '0', '0', '1'};
constexpr size_t kHeaderBytes = 128;
constexpr uint64_t kTagE = 1, kTagW1 = 2;

struct ModelDesc {
    int32_t kind = 0, F = 0, K = 0, id_mode = 0, table_dtype = 0, table_mode = 0, n_hidden = 0;
    int32_t hidden[CTR_MAX_HIDDEN] = {0, 0, 0, 0};
    int32_t n_cross = 0;
    int64_t V = 0;
    uint64_t seed = 0;
    float scale_e = 0, scale_w1 = 0;
    int d() const { return F * K; }
    int d_pad() const { return round_up(F * K, 128); }
    int head_in() const { return kind == CTR_MODEL_LR ? 0 : (n_hidden ? hidden[n_hidden - 1] : 0) + (kind == CTR_MODEL_DCN_V2 ? d() : 0); }
};

struct HostWeights {
    ModelDesc m;
    std::vector<float> blob;        // everything after the header
    size_t off_b0 = 0, off_w1 = 0, off_emb = 0, off_wout = 0, off_bout = 0;
    std::vector<size_t> off_mlp_w, off_mlp_b, off_cross_w, off_cross_b;
};

int32_t read_weight_file(const char* path, HostWeights* hw) {
    FILE* f = fopen(path, "rb");
    if (!f) return fail(CTR_ERR_IO, std::string("cannot open weight file ") + path);
    unsigned char hdr[kHeaderBytes];
    if (fread(hdr, 1, kHeaderBytes, f) != kHeaderBytes || memcmp(hdr, kMagic, 8) != 0) {
        fclose(f);
        return fail(CTR_ERR_IO, "weight file: bad header / magic");
    }
    ModelDesc& m = hw->m;
    int32_t iv[12];
    memcpy(iv, hdr + 8, sizeof(iv));
    m.kind = iv[0]; m.F = iv[1]; m.K = iv[2]; m.id_mode = iv[3]; m.table_dtype = iv[4]; m.table_mode = iv[5]; m.n_hidden = iv[6];
    for (int i = 0; i < CTR_MAX_HIDDEN; ++i) m.hidden[i] = iv[7 + i];
    m.n_cross = iv[11];
    memcpy(&m.V, hdr + 56, 8);
    memcpy(&m.seed, hdr + 64, 8);
    memcpy(&m.scale_e, hdr + 72, 4);
    memcpy(&m.scale_w1, hdr + 76, 4);

    auto bad = [&](const char* why) { fclose(f); return fail(CTR_ERR_IO, std::string("weight file: ") + why); };
    if (m.kind < 0 || m.kind > 3) return bad("unknown model kind");
    if (m.K != kEmbDim) return bad("embedding dim must be 16");
    if (m.F < 1 || m.F > CTR_MAX_FIELDS) return bad("field count out of range (1..48)");
    if (m.V < 1) return bad("row count must be positive");
    if (m.id_mode == CTR_ID_HASHED && (m.V & (m.V - 1))) return bad("hashed tables need a power-of-two row count");
    if (m.n_hidden < 0 || m.n_hidden > CTR_MAX_HIDDEN) return bad("too many hidden layers");
    if (m.kind != CTR_MODEL_LR && m.n_hidden < 1) return bad("tower needs at least one hidden layer");
    for (int i = 0; i < m.n_hidden; ++i)
        if (m.hidden[i] < 8 || m.hidden[i] % 8) return bad("hidden sizes must be positive multiples of 8");
    if (m.kind != CTR_MODEL_DCN_V2) m.n_cross = 0;
    if (m.n_cross < 0 || m.n_cross > 4) return bad("cross layer count out of range (0..4)");
    if (m.kind == CTR_MODEL_DCN_V2 && m.n_cross < 1) return bad("DCN-v2 needs at least one cross layer");

    size_t n = 1;
    hw->off_b0 = 0;
    if (m.table_mode == 0) { hw->off_w1 = n; n += (size_t)m.V; hw->off_emb = n; n += (size_t)m.V * m.K; }
    if (m.kind != CTR_MODEL_LR) {
        int fan_in = m.d();
        for (int l = 0; l < m.n_hidden; ++l) {
            hw->off_mlp_w.push_back(n); n += (size_t)m.hidden[l] * fan_in;
            hw->off_mlp_b.push_back(n); n += (size_t)m.hidden[l];
            fan_in = m.hidden[l];
        }
        for (int l = 0; l < m.n_cross; ++l) {
            hw->off_cross_w.push_back(n); n += (size_t)m.d() * m.d();
            hw->off_cross_b.push_back(n); n += (size_t)m.d();
        }
        hw->off_wout = n; n += (size_t)m.head_in();
        hw->off_bout = n; n += 1;
    }
    hw->blob.resize(n);
    size_t got = fread(hw->blob.data(), sizeof(float), n, f);
    unsigned char extra;
    bool trailing = fread(&extra, 1, 1, f) == 1;
    fclose(f);
    if (got != n) return fail(CTR_ERR_IO, "weight file: truncated tensor data");
    if (trailing) return fail(CTR_ERR_IO, "weight file: trailing bytes after the last tensor");
    return CTR_OK;
}

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

// One launch of tower_chain_kernel: a sequence of layers, each feeding the next.
struct Chain {
    int BN = 256;
    int n_layers = 0;
    ChainMaps maps;
    ChainParams params;       // template: M / score pointer are filled per launch
};

// One stream with its private activation buffers ("slot"): everything a shard needs on the device.
struct Slot {
    int dev = -1;
    cudaStream_t stream = nullptr;
    bool owns_stream = true;
    int64_t cap = 0;           // candidate capacity of the id/weight/score buffers
    int x_rows = 0;            // candidates per launch: rows of X / wide / extra (multiple of 256)
    int scratch_rows = 0;      // rows of the per-pair scratch (resident CTA pairs x 256)
    int64_t* d_ids = nullptr; float* d_wts = nullptr; float* d_scores = nullptr;
    int64_t* d_cids = nullptr; float* d_cwts = nullptr;     // compact lists of a shared-context request (host slots)
    float* d_wide = nullptr; float* d_extra = nullptr; int32_t* d_err = nullptr;
    __half* buf[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int ld[6] = {0, 0, 0, 0, 0, 0};
    std::vector<Chain> chains;
    // pinned staging for pageable callers
    int64_t* h_ids = nullptr; float* h_wts = nullptr; float* h_scores = nullptr; int32_t* h_err = nullptr;
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
    ctr_callback cb = nullptr;
    void* user = nullptr;
    ctr_ticket ticket = 0;
};

struct Task {
    std::shared_ptr<Request> req;
    const int64_t* ids; const float* wts; int64_t n; float* scores;
    // shared-context request: the first n_ctx fields of every candidate come from one shared list and
    // ids / wts hold only the remaining F - n_ctx fields per candidate
    const int64_t* ctx_ids = nullptr; const float* ctx_wts = nullptr; int32_t n_ctx = 0;
    bool stop = false;
};

struct DeviceCtx {
    DeviceModel model;
    std::vector<std::unique_ptr<Slot>> worker_slots;
    std::vector<std::unique_ptr<Slot>> dev_slots;     // for ctr_score_device
    std::mutex dev_mu;                                // serialises ctr_score_device enqueues per device
    cudaStream_t main_stream = nullptr;
    cudaEvent_t fork_ev = nullptr;
    // device sort scratch (ctr_predict_sorted), allocated on first use
    std::mutex sort_mu;
    int64_t sort_cap = 0;
    float* srt_scores = nullptr; float* srt_sorted = nullptr; uint32_t* srt_keys[2] = {nullptr, nullptr};
    int32_t* srt_idx[2] = {nullptr, nullptr}; uint32_t* srt_counts = nullptr;
    std::deque<Task> queue;
    std::mutex q_mu;
    std::condition_variable q_cv;
    std::vector<std::thread> threads;
};

}  // namespace

struct ctr_engine {
    ctr_config cfg;
    ModelDesc m;
    bool loaded = false;
    bool timing = false;
    int chunk_rows = 0;          // 0 = num_sms * 128
    std::vector<std::unique_ptr<DeviceCtx>> devs;
    std::mutex t_mu;
    std::unordered_map<ctr_ticket, std::shared_ptr<Request>> tickets;
    std::atomic<int64_t> next_ticket{1};
    std::atomic<int64_t> launches{0};
    std::mutex pin_mu;
    std::map<uintptr_t, size_t> pinned;     // base -> bytes (buffers from ctr_alloc_pinned)
    std::atomic<unsigned> next_first_device{0};   // rotation of the first GPU of requests that use fewer than all
    std::atomic<bool> shutting_down{false};
    uint8_t* shard_base[CTR_MAX_DEVICES] = {nullptr};   // row-sharded tables: every device's shard
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

// ------------------------------------------------------------------ device set-up
template <typename Kern>
int32_t set_smem_attr(Kern k, int bytes) {
    CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    return CTR_OK;
}

int32_t setup_kernel_attrs() {
    int32_t rc;
    if ((rc = set_smem_attr(tower_chain_kernel<256>, ChainCfg<256>::kSmemBytes)) != CTR_OK) return rc;
    return CTR_OK;
}

// Output widths are padded to 128 columns: the tower walks a layer in 256-column tiles (a third less L2
// traffic per flop than 128-column ones: 22.0k against 35.3k cycles of MMA loop for one 18944 x 640 x 640
// layer) and finishes a width that is an odd multiple of 128 with one 128-column tile.
int32_t upload_gemm_layer(const float* W, const float* b, int N, int K, GemmLayer* L, cudaStream_t st) {
    L->N = N;
    L->Npad = round_up(N, 128);
    L->Kpad = round_up(K, kBK);
    float* tmp = nullptr;
    CUDA_TRY(cudaMalloc(&tmp, (size_t)N * K * sizeof(float)));
    CUDA_TRY(cudaMemcpyAsync(tmp, W, (size_t)N * K * sizeof(float), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMalloc(&L->W, (size_t)L->Npad * L->Kpad * sizeof(__half)));
    convert_pad_f16_kernel<<<1024, 256, 0, st>>>(tmp, N, K, L->W, L->Npad, L->Kpad);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMalloc(&L->bias, (size_t)L->Npad * sizeof(float)));
    CUDA_TRY(cudaMemsetAsync(L->bias, 0, (size_t)L->Npad * sizeof(float), st));
    CUDA_TRY(cudaMemcpyAsync(L->bias, b, (size_t)N * sizeof(float), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    CUDA_TRY(cudaFree(tmp));
    return CTR_OK;
}

int32_t upload_padded_vec(const float* src, int n, int n_pad, float** dst, cudaStream_t st) {
    CUDA_TRY(cudaMalloc(dst, (size_t)n_pad * sizeof(float)));
    CUDA_TRY(cudaMemsetAsync(*dst, 0, (size_t)n_pad * sizeof(float), st));
    if (n) CUDA_TRY(cudaMemcpyAsync(*dst, src, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return CTR_OK;
}

void free_model(DeviceModel& dm) {
    cudaSetDevice(dm.dev);
    cudaFree(dm.table);
    for (auto* v : {&dm.mlp, &dm.cross})
        for (auto& L : *v) { cudaFree(L.W); cudaFree(L.bias); }
    cudaFree(dm.w_out_h);
    cudaFree(dm.w_out_x);
    dm = DeviceModel();
}

int32_t load_model_on_device(ctr_engine* e, int di, const HostWeights& hw) {
    DeviceCtx& dc = *e->devs[di];
    DeviceModel& dm = dc.model;
    const ModelDesc& m = hw.m;
    const int dev = e->cfg.device_ids[di];
    CUDA_TRY(cudaSetDevice(dev));
    dm.dev = dev;
    CUDA_TRY(cudaDeviceGetAttribute(&dm.num_sms, cudaDevAttrMultiProcessorCount, dev));
    cudaStream_t st = dc.main_stream;
    const float* blob = hw.blob.data();
    dm.b0 = blob[hw.off_b0];

    const bool sharded = e->cfg.shard_mode == CTR_TABLES_ROW_SHARDED && e->cfg.n_devices > 1;
    const uint64_t G = sharded ? (uint64_t)e->cfg.n_devices : 1;
    const uint64_t local_rows = ((uint64_t)m.V + G - 1) / G;
    const int shift = m.table_dtype == CTR_TABLE_FP32 ? 7 : 6;
    dm.table_rows_local = local_rows;
    dm.table_bytes = (size_t)local_rows << shift;
    CUDA_TRY(cudaMalloc(&dm.table, dm.table_bytes));
    if (m.table_mode == 1) {
        const uint64_t ke = proc_key(m.seed, kTagE), kw = proc_key(m.seed, kTagW1);
        const int grid = dm.num_sms * 8;
        if (m.table_dtype == CTR_TABLE_FP32)
            fill_table_procedural_kernel<0><<<grid, 256, 0, st>>>(dm.table, local_rows, G, (uint64_t)di % G, ke, kw, m.scale_e, m.scale_w1);
        else
            fill_table_procedural_kernel<1><<<grid, 256, 0, st>>>(dm.table, local_rows, G, (uint64_t)di % G, ke, kw, m.scale_e, m.scale_w1);
        CUDA_TRY(cudaGetLastError());
    } else {
        // raw table: stream it through a bounded device staging buffer and pack on the device
        const uint64_t chunk = 1u << 20;
        float *t_emb = nullptr, *t_w1 = nullptr;
        CUDA_TRY(cudaMalloc(&t_emb, chunk * kEmbDim * sizeof(float)));
        CUDA_TRY(cudaMalloc(&t_w1, chunk * sizeof(float)));
        std::vector<float> h_emb, h_w1;
        for (uint64_t l0 = 0; l0 < local_rows; l0 += chunk) {
            const uint64_t cnt = std::min<uint64_t>(chunk, local_rows - l0);
            const float *src_e, *src_w;
            uint64_t valid = cnt;
            if (G == 1) {
                src_e = blob + hw.off_emb + l0 * kEmbDim;
                src_w = blob + hw.off_w1 + l0;
            } else {                                   // local slot l holds global row l*G + di
                h_emb.assign(cnt * kEmbDim, 0.f);
                h_w1.assign(cnt, 0.f);
                for (uint64_t l = 0; l < cnt; ++l) {
                    const uint64_t row = (l0 + l) * G + (uint64_t)di;
                    if (row >= (uint64_t)m.V) continue;
                    memcpy(&h_emb[l * kEmbDim], blob + hw.off_emb + row * kEmbDim, kEmbDim * sizeof(float));
                    h_w1[l] = blob[hw.off_w1 + row];
                }
                src_e = h_emb.data();
                src_w = h_w1.data();
            }
            CUDA_TRY(cudaMemcpyAsync(t_emb, src_e, valid * kEmbDim * sizeof(float), cudaMemcpyHostToDevice, st));
            CUDA_TRY(cudaMemcpyAsync(t_w1, src_w, valid * sizeof(float), cudaMemcpyHostToDevice, st));
            if (m.table_dtype == CTR_TABLE_FP32) pack_table_kernel<0><<<1024, 256, 0, st>>>(dm.table, l0, t_emb, t_w1, cnt);
            else pack_table_kernel<1><<<1024, 256, 0, st>>>(dm.table, l0, t_emb, t_w1, cnt);
            CUDA_TRY(cudaGetLastError());
            CUDA_TRY(cudaStreamSynchronize(st));
        }
        CUDA_TRY(cudaFree(t_emb));
        CUDA_TRY(cudaFree(t_w1));
    }
    CUDA_TRY(cudaStreamSynchronize(st));

    if (m.kind != CTR_MODEL_LR) {
        int32_t rc;
        int fan_in = m.d();
        dm.mlp.resize(m.n_hidden);
        for (int l = 0; l < m.n_hidden; ++l) {
            // layer 0 reads X (K padded to d_pad); later layers read the previous padded width
            if ((rc = upload_gemm_layer(blob + hw.off_mlp_w[l], blob + hw.off_mlp_b[l], m.hidden[l], fan_in, &dm.mlp[l], st)) != CTR_OK) return rc;
            fan_in = m.hidden[l];
        }
        dm.cross.resize(m.n_cross);
        for (int l = 0; l < m.n_cross; ++l)
            if ((rc = upload_gemm_layer(blob + hw.off_cross_w[l], blob + hw.off_cross_b[l], m.d(), m.d(), &dm.cross[l], st)) != CTR_OK) return rc;
        const float* w_out = blob + hw.off_wout;
        const int h_last = m.hidden[m.n_hidden - 1];
        const int x_part = m.kind == CTR_MODEL_DCN_V2 ? m.d() : 0;
        if ((rc = upload_padded_vec(w_out + x_part, h_last, dm.mlp.back().Npad, &dm.w_out_h, st)) != CTR_OK) return rc;
        if (x_part && (rc = upload_padded_vec(w_out, x_part, dm.cross.back().Npad, &dm.w_out_x, st)) != CTR_OK) return rc;
        dm.b_out = blob[hw.off_bout];
    }
    dm.loaded = true;
    return CTR_OK;
}

// ------------------------------------------------------------------ host locality
// The CPUs on the socket a GPU hangs off (sysfs local_cpulist of its PCIe function), intersected with
// what this process may run on.  Worker threads run there and the pinned staging buffers are allocated
// from there (first touch), so H2D traffic does not cross the inter-socket link.  CTR_NO_CPU_BIND=1
// turns it off; any failure just leaves the default placement.
bool gpu_local_cpus(int dev, cpu_set_t* out) {
    static const bool off = getenv("CTR_NO_CPU_BIND") != nullptr;
    if (off) return false;
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof(bdf), dev) != cudaSuccess) { cudaGetLastError(); return false; }
    for (char* c = bdf; *c; ++c) *c = (char)tolower(*c);
    const std::string path = std::string("/sys/bus/pci/devices/") + bdf + "/local_cpulist";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return false;
    char buf[1024] = {0};
    const bool ok = fgets(buf, sizeof(buf), f) != nullptr;
    fclose(f);
    if (!ok) return false;
    cpu_set_t allowed, local;
    CPU_ZERO(&local);
    if (sched_getaffinity(0, sizeof(allowed), &allowed) != 0) return false;
    for (char* tok = strtok(buf, ",\n"); tok; tok = strtok(nullptr, ",\n")) {   // "0-31,64-95"
        int a = 0, b = 0;
        const int n = sscanf(tok, "%d-%d", &a, &b);
        if (n < 1) continue;
        if (n == 1) b = a;
        for (int c = a; c <= b && c < CPU_SETSIZE; ++c) if (c >= 0 && CPU_ISSET(c, &allowed)) CPU_SET(c, &local);
    }
    if (CPU_COUNT(&local) == 0) return false;
    *out = local;
    return true;
}

// Runs the calling thread on the GPU-local CPUs for the lifetime of this object.
struct ScopedGpuLocality {
    cpu_set_t saved;
    bool active = false;
    explicit ScopedGpuLocality(int dev) {
        cpu_set_t local;
        if (gpu_local_cpus(dev, &local) && sched_getaffinity(0, sizeof(saved), &saved) == 0)
            active = sched_setaffinity(0, sizeof(local), &local) == 0;
    }
    ~ScopedGpuLocality() { if (active) sched_setaffinity(0, sizeof(saved), &saved); }
};

// ------------------------------------------------------------------ slots
void free_slot(Slot& s) {
    cudaSetDevice(s.dev);
    cudaFree(s.d_ids); cudaFree(s.d_wts); cudaFree(s.d_scores); cudaFree(s.d_cids); cudaFree(s.d_cwts); cudaFree(s.d_wide); cudaFree(s.d_extra); cudaFree(s.d_err);
    for (auto*& b : s.buf) { cudaFree(b); b = nullptr; }
    cudaFreeHost(s.h_ids); cudaFreeHost(s.h_wts); cudaFreeHost(s.h_scores); cudaFreeHost(s.h_err);
    if (s.tev) {
        for (int i = 0; i < Slot::kTimeSets; ++i) for (int j = 0; j < Slot::kTimeMarks; ++j) cudaEventDestroy(s.tev[i][j]);
        delete[] s.tev;
    }
    if (s.join_ev) cudaEventDestroy(s.join_ev);
    if (s.stream && s.owns_stream) cudaStreamDestroy(s.stream);
    s = Slot();
}

// Builds the chains of a slot from buffers that already exist (no allocation in here: the
// static-scratch entry point calls this too).
//   buf[0] X [x_rows, d_pad] row-major; tile-major pair scratch (tower.cuh), 2 * scratch_rows rows each:
//   buf[1], buf[2] MLP activations (maxn columns)   buf[3], buf[4] x_l (cross width)   buf[5] copy of x0
int32_t build_chains(ctr_engine* e, DeviceModel& dm, Slot& s) {
    const ModelDesc& m = e->m;
    s.chains.clear();
    if (m.kind == CTR_MODEL_LR) return CTR_OK;
    int32_t rc;
    auto make_chain = [&](const std::vector<GemmLayer>& layers, bool is_cross) -> int32_t {
        Chain c;
        memset(&c.params, 0, sizeof(c.params));
        c.n_layers = (int)layers.size();
        if (c.n_layers > kMaxChain) return fail(CTR_ERR_INVALID, "a chain holds at most 4 layers");
        c.BN = 256;
        int a = 0;                                            // buffer id of the current layer's A operand
        for (int l = 0; l < c.n_layers; ++l) {
            const GemmLayer& L = layers[l];
            const bool last = l == c.n_layers - 1;
            const int out = is_cross ? 3 + (l & 1) : 1 + (l & 1);
            if (s.ld[a] < L.Kpad) return fail(CTR_ERR_INVALID, "internal: activation buffer narrower than the layer's K");
            if (a == 0) rc = make_map_f16(&c.maps.a[l], s.buf[0], (uint64_t)s.x_rows, (uint64_t)L.Kpad, (uint64_t)s.ld[0], kBM);
            else rc = make_map_tile_major(&c.maps.a[l], s.buf[a], (uint64_t)s.ld[a], (uint64_t)(2 * s.scratch_rows / kBM));   // two interleaved tiles per CTA pair
            if (rc != CTR_OK) return rc;
            if ((rc = make_map_f16(&c.maps.w[l], L.W, (uint64_t)L.Npad, (uint64_t)L.Kpad, (uint64_t)L.Kpad, c.BN / 2)) != CTR_OK) return rc;
            ChainLayer& cl = c.params.layer[l];
            cl.N = L.Npad; cl.K = L.Kpad; cl.a_global = a == 0; cl.bias = L.bias; cl.a_ptr = s.buf[a]; cl.lda = s.ld[a];
            cl.n_valid = is_cross ? m.d_pad() : L.Npad;
            cl.out_global = 0;
            if (last) {
                cl.epi = is_cross ? EPI_CROSS_DOT : EPI_HEAD;
                cl.w_out = is_cross ? dm.w_out_x : dm.w_out_h;
                cl.out = nullptr; cl.ldo = 0;
            } else {
                cl.epi = is_cross ? EPI_CROSS : EPI_RELU;
                cl.out = s.buf[out]; cl.ldo = s.ld[out];
            }
            a = out;
        }
        c.params.n_layers = c.n_layers;
        c.params.x0 = s.buf[0]; c.params.ldx0 = s.ld[0];
        if (is_cross && c.n_layers > 1) { c.params.x0e = s.buf[5]; c.params.ldx0e = s.ld[5]; }
        c.params.b_out = dm.b_out;
        if (is_cross) {
            c.params.extra_out = s.d_extra;                   // w_x . x_L, consumed by the MLP chain's head
        } else {
            c.params.wide = m.kind == CTR_MODEL_DCN_V2 ? nullptr : s.d_wide;
            c.params.extra_in = m.kind == CTR_MODEL_DCN_V2 ? s.d_extra : nullptr;
        }
        s.chains.push_back(c);
        return CTR_OK;
    };
    if (m.kind == CTR_MODEL_DCN_V2 && (rc = make_chain(dm.cross, true)) != CTR_OK) return rc;
    return make_chain(dm.mlp, false);
}

// Allocates the model-dependent buffers of a slot and builds its plan.
int32_t build_slot_plan(ctr_engine* e, DeviceCtx& dc, Slot& s, bool host_staging) {
    const ModelDesc& m = e->m;
    DeviceModel& dm = dc.model;
    CUDA_TRY(cudaSetDevice(dm.dev));
    s.cap = e->cfg.max_candidates;
    const int pairs = dm.num_sms / 2;
    // rows per launch: host-request slots hold one request; device-resident slots default to 16 tiles per
    // CTA pair (16 x 74 x 256 = 303 104 candidates, X = 388 MB; measured 239 / 264 / 278 / 284 / 293 M cand/s
    // at 1 / 2 / 4 / 8 / 16 tiles per pair), CTR_CHUNK_ROWS overrides
    const int want = host_staging ? (int)s.cap : (e->chunk_rows > 0 ? e->chunk_rows : 16 * pairs * 2 * kBM);
    s.x_rows = round_up(std::max(want, 1), 2 * kBM);
    s.scratch_rows = std::min(pairs, s.x_rows / (2 * kBM)) * 2 * kBM;
    CUDA_TRY(cudaMalloc(&s.d_ids, (size_t)s.cap * m.F * sizeof(int64_t)));
    CUDA_TRY(cudaMalloc(&s.d_wts, (size_t)s.cap * m.F * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_scores, (size_t)s.cap * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_wide, (size_t)s.x_rows * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_extra, (size_t)s.x_rows * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_err, sizeof(int32_t)));
    CUDA_TRY(cudaMemset(s.d_err, 0, sizeof(int32_t)));
    if (host_staging) {
        CUDA_TRY(cudaMalloc(&s.d_cids, ((size_t)s.cap * m.F + CTR_MAX_FIELDS) * sizeof(int64_t)));
        CUDA_TRY(cudaMalloc(&s.d_cwts, ((size_t)s.cap * m.F + CTR_MAX_FIELDS) * sizeof(float)));
        CUDA_TRY(cudaMallocHost(&s.h_ids, (size_t)s.cap * m.F * sizeof(int64_t)));
        CUDA_TRY(cudaMallocHost(&s.h_wts, (size_t)s.cap * m.F * sizeof(float)));
        CUDA_TRY(cudaMallocHost(&s.h_scores, (size_t)s.cap * sizeof(float)));
        CUDA_TRY(cudaMallocHost(&s.h_err, sizeof(int32_t)));
    }
    if (m.kind == CTR_MODEL_LR) return CTR_OK;

    const int dpad = m.d_pad();
    int maxn = 0;
    for (auto& L : dm.mlp) maxn = std::max(maxn, L.Npad);
    auto alloc_buf = [&](int id, int rows, int ldv) -> int32_t {
        s.ld[id] = ldv;
        CUDA_TRY(cudaMalloc(&s.buf[id], (size_t)rows * ldv * sizeof(__half)));
        CUDA_TRY(cudaMemset(s.buf[id], 0, (size_t)rows * ldv * sizeof(__half)));   // pad columns stay zero
        return CTR_OK;
    };
    int32_t rc;
    if ((rc = alloc_buf(0, s.x_rows, dpad)) != CTR_OK) return rc;
    if ((rc = alloc_buf(1, 2 * s.scratch_rows, maxn)) != CTR_OK) return rc;        // 2 tiles in flight per CTA pair
    if ((rc = alloc_buf(2, 2 * s.scratch_rows, maxn)) != CTR_OK) return rc;
    if (m.kind == CTR_MODEL_DCN_V2) {
        const int xl_ld = dm.cross.empty() ? dpad : dm.cross.back().Npad;      // cross outputs may be padded past d_pad
        if ((rc = alloc_buf(3, 2 * s.scratch_rows, xl_ld)) != CTR_OK) return rc;
        if ((rc = alloc_buf(4, 2 * s.scratch_rows, xl_ld)) != CTR_OK) return rc;
        if ((rc = alloc_buf(5, 2 * s.scratch_rows, xl_ld)) != CTR_OK) return rc;
    }
    return build_chains(e, dm, s);
}

int32_t init_slot_common(Slot& s, int dev, cudaStream_t stream /*nullptr -> create*/) {
    CUDA_TRY(cudaSetDevice(dev));
    s.dev = dev;
    if (stream) { s.stream = stream; s.owns_stream = false; }
    else CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&s.join_ev, cudaEventDisableTiming));
    return CTR_OK;
}

// ------------------------------------------------------------------ the device path for one chunk
template <int TD, bool H, bool E, bool L, bool FM, bool S>
void launch_gather_t(const GatherParams& gp, int grid, size_t smem, cudaStream_t st) {
    gather_pool_kernel<TD, H, E, L, FM, S><<<grid, kGatherThreads, smem, st>>>(gp);
}

template <int TD, bool H, bool S>
void launch_gather_kind(int kind, const GatherParams& gp, int grid, size_t smem, cudaStream_t st) {
    switch (kind) {
        case CTR_MODEL_LR: launch_gather_t<TD, H, false, true, false, S>(gp, grid, smem, st); break;
        case CTR_MODEL_WIDE_DEEP: launch_gather_t<TD, H, true, true, false, S>(gp, grid, smem, st); break;
        case CTR_MODEL_DEEPFM: launch_gather_t<TD, H, true, true, true, S>(gp, grid, smem, st); break;
        default: launch_gather_t<TD, H, true, false, false, S>(gp, grid, smem, st); break;
    }
}

void launch_gather(const ModelDesc& m, bool sharded, const GatherParams& gp, int grid, size_t smem, cudaStream_t st) {
    const bool hashed = m.id_mode == CTR_ID_HASHED;
    if (m.table_dtype == CTR_TABLE_FP32) {
        if (sharded) { if (hashed) launch_gather_kind<0, true, true>(m.kind, gp, grid, smem, st); else launch_gather_kind<0, false, true>(m.kind, gp, grid, smem, st); }
        else { if (hashed) launch_gather_kind<0, true, false>(m.kind, gp, grid, smem, st); else launch_gather_kind<0, false, false>(m.kind, gp, grid, smem, st); }
    } else {
        if (sharded) { if (hashed) launch_gather_kind<1, true, true>(m.kind, gp, grid, smem, st); else launch_gather_kind<1, false, true>(m.kind, gp, grid, smem, st); }
        else { if (hashed) launch_gather_kind<1, true, false>(m.kind, gp, grid, smem, st); else launch_gather_kind<1, false, false>(m.kind, gp, grid, smem, st); }
    }
}

// One chain on `st`: persistent CTA pairs, at most one pair per TPC and per scratch slice.
int32_t launch_chain(const Chain& c, const Slot& s, int n, float* d_scores, int num_sms, cudaStream_t st, long long* prof = nullptr) {
    ChainParams p = c.params;
    p.M = n;
    p.prof = prof;
    if (!p.extra_out) p.scores = d_scores;
    const int tiles = (n + 2 * kBM - 1) / (2 * kBM);
    const int pairs = std::max(1, std::min(std::min(tiles, num_sms / 2), s.scratch_rows / (2 * kBM)));
    tower_chain_kernel<256><<<2 * pairs, kGemmThreads, ChainCfg<256>::kSmemBytes, st>>>(c.maps, p);
    CUDA_TRY(cudaGetLastError());
    return CTR_OK;
}

// Records a CUDA event on the launching stream in front of a stage (stage -2 = end of chunk).
void mark(ctr_engine* e, Slot& s, int stage) {
    if (!e->timing || !s.time_this || !s.tev || s.sets_recorded >= Slot::kTimeSets || s.cur_mark >= Slot::kTimeMarks) return;
    s.mark_stage[s.cur_mark] = stage;
    cudaEventRecord(s.tev[s.sets_recorded][s.cur_mark++], s.stream);
    if (stage == -2) { s.n_marks = s.cur_mark; s.cur_mark = 0; ++s.sets_recorded; s.time_this = false; }
}

// Scores rows [0, n) of (d_ids, d_wts) into d_scores on s.stream; n <= s.x_rows for tower models.
int32_t score_chunk(ctr_engine* e, DeviceCtx& dc, Slot& s, const int64_t* d_ids, const float* d_wts, int n, float* d_scores) {
    const ModelDesc& m = e->m;
    DeviceModel& dm = dc.model;
    cudaStream_t st = s.stream;
    const bool sharded = e->cfg.shard_mode == CTR_TABLES_ROW_SHARDED && e->cfg.n_devices > 1;

    GatherParams gp;
    memset(&gp, 0, sizeof(gp));
    gp.ids = d_ids; gp.wts = d_wts; gp.n = n; gp.F = m.F;
    gp.bulk_ok = ((reinterpret_cast<uintptr_t>(d_ids) | reinterpret_cast<uintptr_t>(d_wts)) & 15) == 0;
    gp.table = dm.table; gp.num_rows = (uint64_t)m.V; gp.b0 = dm.b0;
    gp.error_flag = s.d_err;
    gp.n_shards = sharded ? e->cfg.n_devices : 1;
    for (int g = 0; g < CTR_MAX_DEVICES; ++g) gp.shard_base[g] = e->shard_base[g];
    if (m.kind == CTR_MODEL_LR) { gp.scores = d_scores; }
    else { gp.X = s.buf[0]; gp.ldx = s.ld[0]; gp.wide = m.kind == CTR_MODEL_DCN_V2 ? nullptr : s.d_wide; }
    const int n_tiles = (n + kGatherTile - 1) / kGatherTile;
    static const int gather_ctas_per_sm = [] { const char* c = getenv("CTR_GATHER_CTAS"); return c ? atoi(c) : CTR_GATHER_MIN_CTAS; }();
    const int ggrid = std::max(1, std::min(n_tiles, dm.num_sms * gather_ctas_per_sm));
    const size_t gsmem = (size_t)2 * kGatherTile * m.F * 12;
    mark(e, s, 0);
    launch_gather(m, sharded, gp, ggrid, gsmem, st);
    CUDA_TRY(cudaGetLastError());
    e->launches.fetch_add(1, std::memory_order_relaxed);
    int li = 0;
    static const bool chain_prof = getenv("CTR_CHAIN_PROF") != nullptr;      // bring-up: in-kernel cycle counters to stderr
    for (const Chain& c : s.chains) {
        mark(e, s, ++li);
        long long* d_prof = nullptr;
        if (chain_prof) {
            CUDA_TRY(cudaMalloc(&d_prof, (size_t)dm.num_sms * 8 * sizeof(long long)));
            CUDA_TRY(cudaMemsetAsync(d_prof, 0, (size_t)dm.num_sms * 8 * sizeof(long long), st));
        }
        int32_t rc = launch_chain(c, s, n, d_scores, dm.num_sms, st, d_prof);
        if (rc != CTR_OK) return rc;
        e->launches.fetch_add(1, std::memory_order_relaxed);
        if (chain_prof) {
            std::vector<long long> h((size_t)dm.num_sms * 8);
            CUDA_TRY(cudaMemcpyAsync(h.data(), d_prof, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            const char* names[8] = {"producer wait empty", "mma wait acc_empty", "mma wait full", "mma issue+commit", "mma loop total",
                                    "epi wait acc_full", "epi body", "producer wait layer"};
            fprintf(stderr, "[chain prof] chain %d: %d layers, BN=%d, n=%d (cycles, mean over leader CTAs)\n", li, c.n_layers, c.BN, n);
            for (int i = 0; i < 8; ++i) {
                double mean = 0; int cnt = 0;
                for (int cta = 0; cta < dm.num_sms; cta += 2) { mean += (double)h[cta * 8 + i]; ++cnt; }
                fprintf(stderr, "  %-22s %12.0f\n", names[i], mean / cnt);
            }
            double cyc = 0, ns = 0;
            for (int cta = 0; cta + 1 < dm.num_sms; cta += 2) { cyc += (double)h[cta * 8 + 4]; ns += (double)h[(cta + 1) * 8 + 4]; }
            if (ns > 0) fprintf(stderr, "  %-22s %12.0f MHz (cycles / globaltimer over the MMA loop)\n", "effective SM clock", cyc / ns * 1000.0);
            cudaFree(d_prof);
        }
    }
    mark(e, s, -2);
    return CTR_OK;
}

// Scores n candidates (device pointers) on slot s, chunked so activations stay in their buffers.
int32_t score_on_slot(ctr_engine* e, DeviceCtx& dc, Slot& s, const int64_t* d_ids, const float* d_wts, int64_t n, float* d_scores) {
    const int F = e->m.F;
    const int64_t step = e->m.kind == CTR_MODEL_LR ? (int64_t)1 << 30 : s.x_rows;
    for (int64_t c0 = 0; c0 < n; c0 += step) {
        const int cnt = (int)std::min<int64_t>(step, n - c0);
        s.time_this = false;                   // stage timing is for ctr_score_device calls
        int32_t rc = score_chunk(e, dc, s, d_ids + c0 * F, d_wts + c0 * F, cnt, d_scores + c0);
        if (rc != CTR_OK) return rc;
    }
    return CTR_OK;
}

// Mean device time of `stage` over the recorded sets (caller has synchronised the device).
double mean_stage_ms(Slot& s, int stage) {
    if (!s.tev || s.sets_recorded == 0 || s.n_marks < 2) return -1.0;
    double sum = 0.0; int cnt = 0;
    for (int set = 0; set < s.sets_recorded; ++set) {
        double acc = 0.0; bool any = false;
        for (int i = 0; i + 1 < s.n_marks; ++i) {
            if (stage != -1 && s.mark_stage[i] != stage) continue;
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, s.tev[set][i], s.tev[set][i + 1]) != cudaSuccess) { (void)cudaGetLastError(); any = false; break; }
            acc += ms; any = true;
        }
        if (any) { sum += acc; ++cnt; }
    }
    return cnt ? sum / cnt : -1.0;
}

// ------------------------------------------------------------------ request completion
void complete_shard(ctr_engine* e, const std::shared_ptr<Request>& r, int32_t status, const std::string& err) {
    if (status != CTR_OK) {
        int expected = CTR_OK;
        if (r->status.compare_exchange_strong(expected, status)) {
            std::lock_guard<std::mutex> lk(r->m);
            r->error = err;
        }
    }
    if (r->pending.fetch_sub(1) == 1) {
        {
            std::lock_guard<std::mutex> lk(r->m);
            r->done = true;
        }
        r->cv.notify_all();
        if (r->cb) r->cb(r->user, r->ticket, r->status.load());
    }
}

// sendRequest (DC:91-115) for one shard: stage -> H2D -> kernels -> D2H -> hand back.
void run_task(ctr_engine* e, DeviceCtx& dc, Slot& s, const Task& t) {
    const ModelDesc& m = e->m;
    auto body = [&]() -> int32_t {
        CUDA_TRY(cudaSetDevice(s.dev));
        const int Fk = m.F - t.n_ctx;                          // per-candidate fields actually shipped
        const size_t ib = (size_t)t.n * Fk * sizeof(int64_t), wb = (size_t)t.n * Fk * sizeof(float), sb = (size_t)t.n * sizeof(float);
        // The feature packer: caller buffers that came from ctr_alloc_pinned are DMA'd in place;
        // pageable ones (a Java heap array, a numpy array) are packed into this slot's pinned slab.
        const int64_t* src_ids = t.ids;
        const float* src_wts = t.wts;
        if (!is_pinned(e, t.ids, ib)) { memcpy(s.h_ids, t.ids, ib); src_ids = s.h_ids; }
        if (!is_pinned(e, t.wts, wb)) { memcpy(s.h_wts, t.wts, wb); src_wts = s.h_wts; }
        const bool out_pinned = is_pinned(e, t.scores, sb);
        float* dst_scores = out_pinned ? t.scores : s.h_scores;
        if (t.n_ctx == 0) {
            CUDA_TRY(cudaMemcpyAsync(s.d_ids, src_ids, ib, cudaMemcpyHostToDevice, s.stream));
            CUDA_TRY(cudaMemcpyAsync(s.d_wts, src_wts, wb, cudaMemcpyHostToDevice, s.stream));
        } else {
            // shared context: ship the per-candidate fields once and the context list once, rebuild the
            // reference layout [n, F] on the device (HBM bandwidth is cheap, the host link is not)
            CUDA_TRY(cudaMemcpyAsync(s.d_cids, src_ids, ib, cudaMemcpyHostToDevice, s.stream));
            CUDA_TRY(cudaMemcpyAsync(s.d_cwts, src_wts, wb, cudaMemcpyHostToDevice, s.stream));
            int64_t* d_ctx_ids = s.d_cids + (size_t)s.cap * m.F;
            float* d_ctx_wts = s.d_cwts + (size_t)s.cap * m.F;
            CUDA_TRY(cudaMemcpyAsync(d_ctx_ids, t.ctx_ids, (size_t)t.n_ctx * sizeof(int64_t), cudaMemcpyHostToDevice, s.stream));
            CUDA_TRY(cudaMemcpyAsync(d_ctx_wts, t.ctx_wts, (size_t)t.n_ctx * sizeof(float), cudaMemcpyHostToDevice, s.stream));
            const int64_t total = t.n * m.F;
            expand_context_kernel<<<(int)std::min<int64_t>((total + 255) / 256, dc.model.num_sms * 16), 256, 0, s.stream>>>(
                d_ctx_ids, d_ctx_wts, t.n_ctx, s.d_cids, s.d_cwts, t.n, m.F, s.d_ids, s.d_wts);
            CUDA_TRY(cudaGetLastError());
            e->launches.fetch_add(1, std::memory_order_relaxed);
        }
        int32_t rc = score_on_slot(e, dc, s, s.d_ids, s.d_wts, t.n, s.d_scores);
        if (rc != CTR_OK) return rc;
        CUDA_TRY(cudaMemcpyAsync(dst_scores, s.d_scores, sb, cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaMemcpyAsync(s.h_err, s.d_err, sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        if (*s.h_err) {
            CUDA_TRY(cudaMemsetAsync(s.d_err, 0, sizeof(int32_t), s.stream));
            return fail(CTR_ERR_INVALID, "feature id out of range for a direct-indexed table");
        }
        if (!out_pinned) memcpy(t.scores, s.h_scores, sb);
        return CTR_OK;
    };
    int32_t rc = body();
    complete_shard(e, t.req, rc, rc == CTR_OK ? std::string() : g_last_error);
}

void worker_main(ctr_engine* e, DeviceCtx* dc, Slot* s) {
    cudaSetDevice(s->dev);
    cpu_set_t local;
    if (gpu_local_cpus(s->dev, &local)) sched_setaffinity(0, sizeof(local), &local);   // stay next to the GPU
    for (;;) {
        Task t;
        {
            std::unique_lock<std::mutex> lk(dc->q_mu);
            dc->q_cv.wait(lk, [&] { return !dc->queue.empty(); });
            t = dc->queue.front();
            if (t.stop) return;                 // leave the stop marker for the other workers
            dc->queue.pop_front();
        }
        run_task(e, *dc, *s, t);
    }
}

void stop_workers(ctr_engine* e) {
    for (auto& dcp : e->devs) {
        DeviceCtx& dc = *dcp;
        if (dc.threads.empty()) continue;
        {
            std::lock_guard<std::mutex> lk(dc.q_mu);
            Task t; t.stop = true; t.ids = nullptr; t.wts = nullptr; t.n = 0; t.scores = nullptr;
            dc.queue.push_back(t);
        }
        dc.q_cv.notify_all();
        for (auto& th : dc.threads) th.join();
        dc.threads.clear();
        dc.queue.clear();
    }
}

void free_device_state(ctr_engine* e) {
    for (auto& dcp : e->devs) {
        DeviceCtx& dc = *dcp;
        for (auto& s : dc.worker_slots) free_slot(*s);
        for (auto& s : dc.dev_slots) free_slot(*s);
        dc.worker_slots.clear();
        dc.dev_slots.clear();
        if (dc.model.dev >= 0) free_model(dc.model);
        if (dc.sort_cap) {
            cudaFree(dc.srt_scores); cudaFree(dc.srt_sorted); cudaFree(dc.srt_counts);
            for (int i = 0; i < 2; ++i) { cudaFree(dc.srt_keys[i]); cudaFree(dc.srt_idx[i]); }
            dc.sort_cap = 0;
        }
    }
}

}  // namespace

// =========================================================================== C ABI
extern "C" {

const char* ctr_last_error(void) { return g_last_error.c_str(); }
int32_t ctr_abi_version(void) { return CTR_ABI_VERSION; }

int32_t ctr_partition_bounds(int64_t n, int32_t part_num, int64_t* begin, int64_t* end) {
    if (part_num <= 0 || n < 0 || !begin || !end) return fail(CTR_ERR_INVALID, "ctr_partition_bounds: part_num must be positive");
    split_bounds(n, part_num, begin, end);
    return CTR_OK;
}

int32_t ctr_engine_create(const ctr_config* cfg, ctr_engine** out) {
    if (!cfg || !out) return fail(CTR_ERR_INVALID, "ctr_engine_create: null argument");
    if (cfg->struct_size != (int32_t)sizeof(ctr_config)) return fail(CTR_ERR_INVALID, "ctr_engine_create: ctr_config size mismatch (ABI)");
    if (cfg->n_devices < 1 || cfg->n_devices > CTR_MAX_DEVICES) return fail(CTR_ERR_INVALID, "ctr_engine_create: n_devices must be 1..8");
    if (cfg->max_candidates < 1) return fail(CTR_ERR_INVALID, "ctr_engine_create: max_candidates must be positive");
    int count = 0;
    cudaError_t ce = cudaGetDeviceCount(&count);
    if (ce != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        return fail(CTR_ERR_CUDA, std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(ce));
    }
    for (int i = 0; i < cfg->n_devices; ++i) {
        if (cfg->device_ids[i] < 0 || cfg->device_ids[i] >= count) return fail(CTR_ERR_INVALID, "ctr_engine_create: device id out of range");
        cudaDeviceProp prop;
        CUDA_TRY(cudaGetDeviceProperties(&prop, cfg->device_ids[i]));
        if (prop.major != 10) return fail(CTR_ERR_CUDA, std::string("device ") + prop.name + " is not sm_100: this engine has no other code path");
    }
    std::unique_ptr<ctr_engine> e(new ctr_engine());
    e->cfg = *cfg;
    if (e->cfg.workers_per_device <= 0) e->cfg.workers_per_device = 2;
    if (const char* c = getenv("CTR_CHUNK_ROWS")) { int v = atoi(c); if (v >= 128) e->chunk_rows = round_up(v, 128); }
    for (int i = 0; i < cfg->n_devices; ++i) {
        std::unique_ptr<DeviceCtx> dc(new DeviceCtx());
        CUDA_TRY(cudaSetDevice(cfg->device_ids[i]));
        CUDA_TRY(cudaStreamCreateWithFlags(&dc->main_stream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&dc->fork_ev, cudaEventDisableTiming));
        int32_t rc = setup_kernel_attrs();
        if (rc != CTR_OK) return rc;
        dc->model.dev = -1;
        e->devs.push_back(std::move(dc));
    }
    if (cfg->shard_mode == CTR_TABLES_ROW_SHARDED && cfg->n_devices > 1) {
        for (int i = 0; i < cfg->n_devices; ++i)
            for (int j = 0; j < cfg->n_devices; ++j) {
                if (i == j) continue;
                int can = 0;
                CUDA_TRY(cudaDeviceCanAccessPeer(&can, cfg->device_ids[i], cfg->device_ids[j]));
                if (!can) return fail(CTR_ERR_CUDA, "row-sharded tables need peer access between all devices");
                CUDA_TRY(cudaSetDevice(cfg->device_ids[i]));
                cudaError_t pe = cudaDeviceEnablePeerAccess(cfg->device_ids[j], 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) CUDA_TRY(pe);
                (void)cudaGetLastError();
            }
    }
    *out = e.release();
    return CTR_OK;
}

int32_t ctr_engine_destroy(ctr_engine* e) {
    if (!e) return CTR_OK;
    e->shutting_down = true;
    stop_workers(e);
    free_device_state(e);
    for (size_t i = 0; i < e->devs.size(); ++i) {
        cudaSetDevice(e->cfg.device_ids[i]);
        if (e->devs[i]->main_stream) cudaStreamDestroy(e->devs[i]->main_stream);
        if (e->devs[i]->fork_ev) cudaEventDestroy(e->devs[i]->fork_ev);
    }
    {
        std::lock_guard<std::mutex> lk(e->pin_mu);
        for (auto& kv : e->pinned) cudaFreeHost(reinterpret_cast<void*>(kv.first));
        e->pinned.clear();
    }
    delete e;
    return CTR_OK;
}

int32_t ctr_engine_load_weights(ctr_engine* e, const char* path) {
    if (!e || !path) return fail(CTR_ERR_INVALID, "ctr_engine_load_weights: null argument");
    HostWeights hw;
    int32_t rc = read_weight_file(path, &hw);
    if (rc != CTR_OK) return rc;
    stop_workers(e);
    free_device_state(e);
    e->loaded = false;
    e->m = hw.m;
    for (int i = 0; i < e->cfg.n_devices; ++i) {
        if ((rc = load_model_on_device(e, i, hw)) != CTR_OK) return rc;
        e->shard_base[i] = e->devs[i]->model.table;
    }
    for (int i = 0; i < e->cfg.n_devices; ++i) {
        DeviceCtx& dc = *e->devs[i];
        const int dev = e->cfg.device_ids[i];
        ScopedGpuLocality near_gpu(dev);                       // pinned staging of this GPU is allocated on its socket
        for (int w = 0; w < e->cfg.workers_per_device; ++w) {
            std::unique_ptr<Slot> s(new Slot());
            if ((rc = init_slot_common(*s, dev, nullptr)) != CTR_OK) return rc;
            if ((rc = build_slot_plan(e, dc, *s, true)) != CTR_OK) return rc;
            dc.worker_slots.push_back(std::move(s));
        }
        // one stream by default: gather and tower both fill the machine, and a second stream only
        // interleaves them (+4% throughput, measured) at the price of blurred per-kernel timing
        int n_dev_slots = 1;
        if (const char* c = getenv("CTR_DEVICE_STREAMS")) { int v = atoi(c); if (v >= 1 && v <= 8) n_dev_slots = v; }
        for (int w = 0; w < n_dev_slots; ++w) {
            std::unique_ptr<Slot> s(new Slot());
            if ((rc = init_slot_common(*s, dev, nullptr)) != CTR_OK) return rc;
            if ((rc = build_slot_plan(e, dc, *s, false)) != CTR_OK) return rc;
            dc.dev_slots.push_back(std::move(s));
        }
        CUDA_TRY(cudaDeviceSynchronize());
    }
    e->loaded = true;
    for (int i = 0; i < e->cfg.n_devices; ++i) {
        DeviceCtx& dc = *e->devs[i];
        for (auto& s : dc.worker_slots) dc.threads.emplace_back(worker_main, e, &dc, s.get());
    }
    return CTR_OK;
}

int32_t ctr_engine_model_info(ctr_engine* e, ctr_model_info* out) {
    if (!e || !out) return fail(CTR_ERR_INVALID, "ctr_engine_model_info: null argument");
    if (!e->loaded) return fail(CTR_ERR_STATE, "weights not loaded");
    memset(out, 0, sizeof(*out));
    out->model_kind = e->m.kind; out->num_fields = e->m.F; out->emb_dim = e->m.K; out->id_mode = e->m.id_mode;
    out->table_dtype = e->m.table_dtype; out->n_hidden = e->m.n_hidden;
    for (int i = 0; i < CTR_MAX_HIDDEN; ++i) out->hidden[i] = e->m.hidden[i];
    out->n_cross = e->m.n_cross; out->d_padded = e->m.d_pad(); out->num_rows = e->m.V;
    out->table_bytes_per_device = (int64_t)e->devs[0]->model.table_bytes;
    out->chunk_rows = e->devs[0]->dev_slots.empty() ? 0 : e->devs[0]->dev_slots[0]->x_rows;
    return CTR_OK;
}

static int32_t submit_request(ctr_engine* e, const char* who, const int64_t* ctx_ids, const float* ctx_wts, int32_t n_ctx,
                              const int64_t* ids, const float* wts, int64_t n_cand, float* scores_out, ctr_callback cb, void* user,
                              ctr_ticket* ticket) {
    if (!e || !ticket) return fail(CTR_ERR_INVALID, std::string(who) + ": null argument");
    if (!e->loaded || e->shutting_down) return fail(CTR_ERR_STATE, std::string(who) + ": weights not loaded or engine shutting down");
    if (n_cand < 0 || (n_cand > 0 && (!ids || !wts || !scores_out))) return fail(CTR_ERR_INVALID, std::string(who) + ": null buffer or negative count");
    if (n_ctx < 0 || n_ctx >= e->m.F || (n_ctx > 0 && (!ctx_ids || !ctx_wts)))
        return fail(CTR_ERR_INVALID, std::string(who) + ": the shared context must cover 0 .. F-1 leading fields");
    // How many GPUs this request is cut over.  The reference always cuts hostsList.size() ways (DC:38,
    // DC:148); scores do not depend on the cut, and for a small request the fan-out only costs host
    // work and tail latency (8 GPUs, 64-4096 candidates: p99 0.42-0.5 ms against 0.16 on one GPU), so a
    // shard holds at least min_shard candidates and the first GPU rotates from request to request.
    const int G = e->cfg.n_devices;
    const int64_t min_shard = e->cfg.min_shard_candidates > 0 ? e->cfg.min_shard_candidates : 1024;
    int P = (int)std::max<int64_t>(1, std::min<int64_t>(G, n_cand / min_shard));
    const int64_t need = (n_cand + e->cfg.max_candidates - 1) / std::max(1, e->cfg.max_candidates);   // capacity of one shard
    if (need > G) return fail(CTR_ERR_CAPACITY, std::string(who) + ": a shard exceeds max_candidates of this engine");
    P = (int)std::max<int64_t>(P, need);
    while (P < G && n_cand - (int64_t)(P - 1) * (n_cand / P) > e->cfg.max_candidates) ++P;   // the last shard takes the remainder
    const int first = P < G ? (int)(e->next_first_device.fetch_add(1, std::memory_order_relaxed) % (unsigned)G) : 0;
    int64_t b[CTR_MAX_DEVICES], en[CTR_MAX_DEVICES];
    split_bounds(n_cand, P, b, en);
    for (int i = 0; i < P; ++i)
        if (en[i] - b[i] > e->cfg.max_candidates)
            return fail(CTR_ERR_CAPACITY, std::string(who) + ": a shard exceeds max_candidates of this engine");
    auto req = std::make_shared<Request>();
    req->cb = cb; req->user = user;
    req->ticket = e->next_ticket.fetch_add(1);
    int shards = 0;
    for (int i = 0; i < P; ++i) shards += en[i] > b[i];
    {
        std::lock_guard<std::mutex> lk(e->t_mu);
        e->tickets[req->ticket] = req;
    }
    *ticket = req->ticket;
    if (shards == 0) {                      // empty request: nothing to score, completes at once
        req->pending = 1;
        complete_shard(e, req, CTR_OK, "");
        return CTR_OK;
    }
    req->pending = shards;
    const int Fk = e->m.F - n_ctx;
    for (int i = 0; i < P; ++i) {
        if (en[i] == b[i]) continue;
        Task t;
        t.req = req; t.ids = ids + b[i] * Fk; t.wts = wts + b[i] * Fk; t.n = en[i] - b[i]; t.scores = scores_out + b[i];
        t.ctx_ids = ctx_ids; t.ctx_wts = ctx_wts; t.n_ctx = n_ctx;
        DeviceCtx& dc = *e->devs[(first + i) % G];
        {
            std::lock_guard<std::mutex> lk(dc.q_mu);
            dc.queue.push_back(t);
        }
        dc.q_cv.notify_one();
    }
    return CTR_OK;
}

int32_t ctr_predict_async(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand, float* scores_out,
                          ctr_callback cb, void* user, ctr_ticket* ticket) {
    return submit_request(e, "ctr_predict_async", nullptr, nullptr, 0, ids, wts, n_cand, scores_out, cb, user, ticket);
}

int32_t ctr_predict_shared_async(ctr_engine* e, const int64_t* ctx_ids, const float* ctx_wts, int32_t n_ctx_fields,
                                 const int64_t* cand_ids, const float* cand_wts, int64_t n_cand, float* scores_out,
                                 ctr_callback cb, void* user, ctr_ticket* ticket) {
    return submit_request(e, "ctr_predict_shared_async", ctx_ids, ctx_wts, n_ctx_fields, cand_ids, cand_wts, n_cand, scores_out, cb, user, ticket);
}

int32_t ctr_wait(ctr_engine* e, ctr_ticket ticket, int32_t timeout_ms) {
    if (!e) return fail(CTR_ERR_INVALID, "ctr_wait: null engine");
    std::shared_ptr<Request> r;
    {
        std::lock_guard<std::mutex> lk(e->t_mu);
        auto it = e->tickets.find(ticket);
        if (it == e->tickets.end()) return fail(CTR_ERR_STATE, "ctr_wait: unknown or already released ticket");
        r = it->second;
    }
    {
        std::unique_lock<std::mutex> lk(r->m);
        if (timeout_ms < 0) r->cv.wait(lk, [&] { return r->done; });
        else if (!r->cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), [&] { return r->done; }))
            return fail(CTR_ERR_TIMEOUT, "ctr_wait: timed out");
    }
    {
        std::lock_guard<std::mutex> lk(e->t_mu);
        e->tickets.erase(ticket);
    }
    const int32_t st = r->status.load();
    if (st != CTR_OK) return fail(st, r->error);
    return CTR_OK;
}

int32_t ctr_predict(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand, float* scores_out) {
    ctr_ticket t;
    int32_t rc = ctr_predict_async(e, ids, wts, n_cand, scores_out, nullptr, nullptr, &t);
    if (rc != CTR_OK) return rc;
    return ctr_wait(e, t, -1);
}

// DC:195 on the device.  The merged scores (host) go to device 0 of the engine, are sorted there with
// their candidate indices, and come back; 8 bytes per candidate each way next to 468 going in.
int32_t ctr_predict_sorted(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand, float* sorted_scores_out,
                           int32_t* order_out) {
    if (!e || (n_cand > 0 && !sorted_scores_out)) return fail(CTR_ERR_INVALID, "ctr_predict_sorted: null argument");
    if (n_cand > INT32_MAX / 2) return fail(CTR_ERR_CAPACITY, "ctr_predict_sorted: too many candidates");
    int32_t rc = ctr_predict(e, ids, wts, n_cand, sorted_scores_out);          // unsorted, ordered merge (DC:161-164)
    if (rc != CTR_OK || n_cand == 0) return rc;
    DeviceCtx& dc = *e->devs[0];
    std::lock_guard<std::mutex> lk(dc.sort_mu);
    CUDA_TRY(cudaSetDevice(dc.model.dev));
    const int n = (int)n_cand;
    const int n_blocks = (n + kSortTile - 1) / kSortTile;
    if (dc.sort_cap < n_cand) {
        if (dc.sort_cap) {
            cudaFree(dc.srt_scores); cudaFree(dc.srt_sorted); cudaFree(dc.srt_counts);
            for (int i = 0; i < 2; ++i) { cudaFree(dc.srt_keys[i]); cudaFree(dc.srt_idx[i]); }
            dc.sort_cap = 0;
        }
        const int64_t cap = std::max<int64_t>(n_cand, e->cfg.max_candidates * (int64_t)e->cfg.n_devices);
        CUDA_TRY(cudaMalloc(&dc.srt_scores, cap * sizeof(float)));
        CUDA_TRY(cudaMalloc(&dc.srt_sorted, cap * sizeof(float)));
        for (int i = 0; i < 2; ++i) {
            CUDA_TRY(cudaMalloc(&dc.srt_keys[i], cap * sizeof(uint32_t)));
            CUDA_TRY(cudaMalloc(&dc.srt_idx[i], cap * sizeof(int32_t)));
        }
        CUDA_TRY(cudaMalloc(&dc.srt_counts, (size_t)256 * ((cap + kSortTile - 1) / kSortTile) * sizeof(uint32_t)));
        dc.sort_cap = cap;
    }
    cudaStream_t st = dc.main_stream;
    CUDA_TRY(cudaMemcpyAsync(dc.srt_scores, sorted_scores_out, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, st));
    const int g1 = std::max(1, std::min((n + 255) / 256, dc.model.num_sms * 8));
    sort_make_keys_kernel<<<g1, 256, 0, st>>>(dc.srt_scores, n, dc.srt_keys[0], dc.srt_idx[0]);
    int cur = 0;
    for (int pass = 0; pass < 4; ++pass) {
        sort_hist_kernel<<<n_blocks, kSortThreads, 0, st>>>(dc.srt_keys[cur], n, pass * 8, dc.srt_counts);
        sort_scan_kernel<<<1, 1024, 0, st>>>(dc.srt_counts, 256 * n_blocks);
        sort_scatter_kernel<<<n_blocks, kSortThreads, 0, st>>>(dc.srt_keys[cur], dc.srt_idx[cur], n, pass * 8, dc.srt_counts,
                                                              dc.srt_keys[cur ^ 1], dc.srt_idx[cur ^ 1]);
        cur ^= 1;
    }
    sort_gather_scores_kernel<<<g1, 256, 0, st>>>(dc.srt_scores, dc.srt_idx[cur], n, dc.srt_sorted);
    CUDA_TRY(cudaGetLastError());
    e->launches.fetch_add(14, std::memory_order_relaxed);
    CUDA_TRY(cudaMemcpyAsync(sorted_scores_out, dc.srt_sorted, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (order_out) CUDA_TRY(cudaMemcpyAsync(order_out, dc.srt_idx[cur], (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return CTR_OK;
}

// The RPC itself (DC:112): serialised PredictRequest in, serialised PredictResponse out.
int32_t ctr_predict_wire(ctr_engine* e, const void* request, int64_t request_len, void* response, int64_t response_cap,
                         int64_t* response_len) {
    if (!e || !request || request_len < 0 || !response_len) return fail(CTR_ERR_INVALID, "ctr_predict_wire: null argument");
    if (!e->loaded) return fail(CTR_ERR_STATE, "ctr_predict_wire: weights not loaded");
    std::vector<int64_t> ids;
    std::vector<float> wts;
    std::string model, sig, why;
    if (!ctr_wire::parse_predict_request(static_cast<const uint8_t*>(request), (size_t)request_len, e->m.F, &ids, &wts, &model, &sig, &why))
        return fail(CTR_ERR_INVALID, "ctr_predict_wire: " + why);
    const int64_t n = (int64_t)ids.size() / e->m.F;
    std::vector<float> scores((size_t)n);
    int32_t rc = ctr_predict(e, ids.data(), wts.data(), n, scores.data());
    if (rc != CTR_OK) return rc;
    const std::string out = ctr_wire::build_predict_response(scores.data(), n, model, sig);
    *response_len = (int64_t)out.size();
    if (!response || response_cap < (int64_t)out.size()) return fail(CTR_ERR_CAPACITY, "ctr_predict_wire: response buffer too small (needed size is in *response_len)");
    memcpy(response, out.data(), out.size());
    return CTR_OK;
}

// Wire helpers without an engine (pure host code): what ctr_predict_wire does on either side of the scoring.
int32_t ctr_wire_parse_request(const void* request, int64_t request_len, int32_t field_num, int64_t* ids_out, float* wts_out,
                               int64_t cap_elems, int64_t* n_cand) {
    if (!request || request_len < 0 || field_num < 1 || !n_cand) return fail(CTR_ERR_INVALID, "ctr_wire_parse_request: bad argument");
    std::vector<int64_t> ids;
    std::vector<float> wts;
    std::string model, sig, why;
    if (!ctr_wire::parse_predict_request(static_cast<const uint8_t*>(request), (size_t)request_len, field_num, &ids, &wts, &model, &sig, &why))
        return fail(CTR_ERR_INVALID, "ctr_wire_parse_request: " + why);
    *n_cand = (int64_t)ids.size() / field_num;
    if ((int64_t)ids.size() > cap_elems || !ids_out || !wts_out) return fail(CTR_ERR_CAPACITY, "ctr_wire_parse_request: output buffers too small");
    memcpy(ids_out, ids.data(), ids.size() * sizeof(int64_t));
    memcpy(wts_out, wts.data(), wts.size() * sizeof(float));
    return CTR_OK;
}

int32_t ctr_wire_build_response(const float* scores, int64_t n_cand, void* response, int64_t response_cap, int64_t* response_len) {
    if ((n_cand > 0 && !scores) || n_cand < 0 || !response_len) return fail(CTR_ERR_INVALID, "ctr_wire_build_response: bad argument");
    const std::string out = ctr_wire::build_predict_response(scores, n_cand, "DCN", "serving_default");
    *response_len = (int64_t)out.size();
    if (!response || response_cap < (int64_t)out.size()) return fail(CTR_ERR_CAPACITY, "ctr_wire_build_response: response buffer too small");
    memcpy(response, out.data(), out.size());
    return CTR_OK;
}

int32_t ctr_alloc_pinned(ctr_engine* e, int64_t bytes, void** out) {
    if (!e || !out || bytes <= 0) return fail(CTR_ERR_INVALID, "ctr_alloc_pinned: bad argument");
    void* p = nullptr;
    CUDA_TRY(cudaSetDevice(e->cfg.device_ids[0]));
    // a single-GPU engine knows which GPU will read the buffer: allocate it on that GPU's socket
    std::unique_ptr<ScopedGpuLocality> near_gpu(e->cfg.n_devices == 1 ? new ScopedGpuLocality(e->cfg.device_ids[0]) : nullptr);
    CUDA_TRY(cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocPortable));
    std::lock_guard<std::mutex> lk(e->pin_mu);
    e->pinned[(uintptr_t)p] = (size_t)bytes;
    *out = p;
    return CTR_OK;
}

int32_t ctr_free_pinned(ctr_engine* e, void* p) {
    if (!e || !p) return fail(CTR_ERR_INVALID, "ctr_free_pinned: null argument");
    {
        std::lock_guard<std::mutex> lk(e->pin_mu);
        auto it = e->pinned.find((uintptr_t)p);
        if (it == e->pinned.end()) return fail(CTR_ERR_INVALID, "ctr_free_pinned: pointer was not allocated by ctr_alloc_pinned");
        e->pinned.erase(it);
    }
    CUDA_TRY(cudaFreeHost(p));
    return CTR_OK;
}

int32_t ctr_score_device(ctr_engine* e, int32_t device_index, const void* d_ids, const void* d_wts, int64_t n_cand,
                         void* d_scores, void* stream) {
    if (!e) return fail(CTR_ERR_INVALID, "ctr_score_device: null engine");
    if (!e->loaded) return fail(CTR_ERR_STATE, "ctr_score_device: weights not loaded");
    if (device_index < 0 || device_index >= e->cfg.n_devices) return fail(CTR_ERR_INVALID, "ctr_score_device: device index out of range");
    if (n_cand < 0 || (n_cand > 0 && (!d_ids || !d_wts || !d_scores))) return fail(CTR_ERR_INVALID, "ctr_score_device: null buffer or negative count");
    if (n_cand == 0) return CTR_OK;
    DeviceCtx& dc = *e->devs[device_index];
    std::lock_guard<std::mutex> lk(dc.dev_mu);
    CUDA_TRY(cudaSetDevice(dc.model.dev));
    cudaStream_t user = reinterpret_cast<cudaStream_t>(stream);
    cudaStream_t origin = user ? user : dc.main_stream;
    const int F = e->m.F;
    const int64_t* ids = static_cast<const int64_t*>(d_ids);
    const float* wts = static_cast<const float*>(d_wts);
    float* scores = static_cast<float*>(d_scores);
    const int n_slots = (int)dc.dev_slots.size();
    const int64_t step = e->m.kind == CTR_MODEL_LR ? n_cand : dc.dev_slots[0]->x_rows;
    const int64_t n_chunks = (n_cand + step - 1) / step;
    const int used = (int)std::min<int64_t>(n_slots, n_chunks);
    // fork: the engine's streams start after whatever the caller queued on `origin`
    CUDA_TRY(cudaEventRecord(dc.fork_ev, origin));
    for (int w = 0; w < used; ++w) CUDA_TRY(cudaStreamWaitEvent(dc.dev_slots[w]->stream, dc.fork_ev, 0));
    for (int64_t c = 0; c < n_chunks; ++c) {
        Slot& s = *dc.dev_slots[c % used];
        const int64_t c0 = c * step;
        const int cnt = (int)std::min<int64_t>(step, n_cand - c0);
        if (c == 0) { s.time_this = true; s.cur_mark = 0; }  // stage timing: first chunk (slot 0) only
        int32_t rc = score_chunk(e, dc, s, ids + c0 * F, wts + c0 * F, cnt, scores + c0);
        if (rc != CTR_OK) return rc;
    }
    // join: `origin` continues after every engine stream has finished
    for (int w = 0; w < used; ++w) {
        CUDA_TRY(cudaEventRecord(dc.dev_slots[w]->join_ev, dc.dev_slots[w]->stream));
        CUDA_TRY(cudaStreamWaitEvent(origin, dc.dev_slots[w]->join_ev, 0));
    }
    if (!user) {
        CUDA_TRY(cudaStreamSynchronize(origin));
        int32_t err = 0;
        for (int w = 0; w < used; ++w) {
            int32_t v = 0;
            CUDA_TRY(cudaMemcpy(&v, dc.dev_slots[w]->d_err, sizeof(v), cudaMemcpyDeviceToHost));
            if (v) { err = 1; CUDA_TRY(cudaMemset(dc.dev_slots[w]->d_err, 0, sizeof(v))); }
        }
        if (err) return fail(CTR_ERR_INVALID, "feature id out of range for a direct-indexed table");
    }
    return CTR_OK;
}

int64_t ctr_engine_launch_count(ctr_engine* e) { return e ? e->launches.load() : 0; }

int32_t ctr_engine_set_timing(ctr_engine* e, int32_t enabled) {
    if (!e) return fail(CTR_ERR_INVALID, "null engine");
    if (!e->loaded) return fail(CTR_ERR_STATE, "weights not loaded");
    for (auto& dcp : e->devs) {
        std::lock_guard<std::mutex> lk(dcp->dev_mu);
        Slot& s = *dcp->dev_slots[0];
        CUDA_TRY(cudaSetDevice(s.dev));
        if (enabled && !s.tev) {
            s.tev = new cudaEvent_t[Slot::kTimeSets][Slot::kTimeMarks];
            for (int i = 0; i < Slot::kTimeSets; ++i) for (int j = 0; j < Slot::kTimeMarks; ++j) CUDA_TRY(cudaEventCreate(&s.tev[i][j]));
        }
        s.sets_recorded = 0; s.cur_mark = 0; s.n_marks = 0; s.time_this = false;    // (re)start the averaging window
    }
    e->timing = enabled != 0;
    return CTR_OK;
}

double ctr_engine_stage_ms(ctr_engine* e, int32_t device_index, int32_t stage) {
    if (!e || device_index < 0 || device_index >= (int)e->devs.size() || e->devs[device_index]->dev_slots.empty()) return -1.0;
    DeviceCtx& dc = *e->devs[device_index];
    std::lock_guard<std::mutex> lk(dc.dev_mu);
    cudaSetDevice(dc.model.dev);
    return mean_stage_ms(*dc.dev_slots[0], stage);
}

int32_t ctr_debug_read_rows(ctr_engine* e, int32_t device_index, const int64_t* rows, int64_t n, float* emb_out, float* w1_out) {
    if (!e || !rows || !emb_out || !w1_out || n < 0) return fail(CTR_ERR_INVALID, "ctr_debug_read_rows: bad argument");
    if (!e->loaded) return fail(CTR_ERR_STATE, "weights not loaded");
    if (device_index < 0 || device_index >= e->cfg.n_devices) return fail(CTR_ERR_INVALID, "device index out of range");
    if (n == 0) return CTR_OK;
    DeviceModel& dm = e->devs[device_index]->model;
    const bool sharded = e->cfg.shard_mode == CTR_TABLES_ROW_SHARDED && e->cfg.n_devices > 1;
    const int64_t G = sharded ? e->cfg.n_devices : 1;
    std::vector<int64_t> slots(n);
    for (int64_t i = 0; i < n; ++i) {
        if (rows[i] < 0 || rows[i] >= e->m.V) return fail(CTR_ERR_INVALID, "row out of range");
        if (sharded && rows[i] % G != device_index) return fail(CTR_ERR_INVALID, "row is not held by this device's shard");
        slots[i] = rows[i] / G;
    }
    CUDA_TRY(cudaSetDevice(dm.dev));
    int64_t* d_slots; float *d_e, *d_w;
    CUDA_TRY(cudaMalloc(&d_slots, n * sizeof(int64_t)));
    CUDA_TRY(cudaMalloc(&d_e, n * kEmbDim * sizeof(float)));
    CUDA_TRY(cudaMalloc(&d_w, n * sizeof(float)));
    CUDA_TRY(cudaMemcpy(d_slots, slots.data(), n * sizeof(int64_t), cudaMemcpyHostToDevice));
    if (e->m.table_dtype == CTR_TABLE_FP32) read_rows_kernel<0><<<256, 256>>>(dm.table, d_slots, n, d_e, d_w);
    else read_rows_kernel<1><<<256, 256>>>(dm.table, d_slots, n, d_e, d_w);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy(emb_out, d_e, n * kEmbDim * sizeof(float), cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(w1_out, d_w, n * sizeof(float), cudaMemcpyDeviceToHost));
    cudaFree(d_slots); cudaFree(d_e); cudaFree(d_w);
    return CTR_OK;
}

int32_t ctr_debug_gemm(ctr_engine* e, int32_t device_index, const float* A, const float* W, const float* bias,
                       int32_t M, int32_t N, int32_t K, int32_t relu, float* out) {
    if (!e || !A || !W || !bias || !out || M < 1 || N < 1 || K < 1) return fail(CTR_ERR_INVALID, "ctr_debug_gemm: bad argument");
    if (device_index < 0 || device_index >= e->cfg.n_devices) return fail(CTR_ERR_INVALID, "device index out of range");
    (void)relu;
    const int dev = e->cfg.device_ids[device_index];
    CUDA_TRY(cudaSetDevice(dev));
    int num_sms = 148;
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    cudaStream_t st = e->devs[device_index]->main_stream;
    GemmLayer L;
    int32_t rc = upload_gemm_layer(W, bias, N, K, &L, st);
    if (rc != CTR_OK) return rc;
    const int Mpad = round_up(M, 2 * kBM);
    float* dA32 = nullptr; __half *dA = nullptr, *dO = nullptr;
    CUDA_TRY(cudaMalloc(&dA32, (size_t)M * K * sizeof(float)));
    CUDA_TRY(cudaMalloc(&dA, (size_t)Mpad * L.Kpad * sizeof(__half)));
    CUDA_TRY(cudaMalloc(&dO, (size_t)Mpad * L.Npad * sizeof(__half)));
    CUDA_TRY(cudaMemcpyAsync(dA32, A, (size_t)M * K * sizeof(float), cudaMemcpyHostToDevice, st));
    convert_pad_f16_kernel<<<1024, 256, 0, st>>>(dA32, M, K, dA, Mpad, L.Kpad);
    CUDA_TRY(cudaGetLastError());
    Chain c;
    memset(&c.params, 0, sizeof(c.params));
    c.BN = 256;
    c.n_layers = 1;
    if ((rc = make_map_f16(&c.maps.a[0], dA, (uint64_t)Mpad, (uint64_t)L.Kpad, (uint64_t)L.Kpad, kBM)) != CTR_OK) return rc;
    if ((rc = make_map_f16(&c.maps.w[0], L.W, (uint64_t)L.Npad, (uint64_t)L.Kpad, (uint64_t)L.Kpad, c.BN / 2)) != CTR_OK) return rc;
    ChainLayer& cl = c.params.layer[0];
    cl.N = L.Npad; cl.K = L.Kpad; cl.n_valid = L.Npad; cl.epi = EPI_RELU; cl.a_global = 1; cl.bias = L.bias; cl.a_ptr = dA; cl.lda = L.Kpad;
    cl.out = dO; cl.ldo = L.Npad; cl.out_global = 1;
    c.params.n_layers = 1;
    c.params.extra_out = nullptr;
    Slot fake;
    fake.scratch_rows = Mpad;
    long long* d_prof = nullptr;
    const bool prof = getenv("CTR_GEMM_PROF") != nullptr;
    if (prof) {
        CUDA_TRY(cudaMalloc(&d_prof, (size_t)num_sms * 8 * sizeof(long long)));
        CUDA_TRY(cudaMemsetAsync(d_prof, 0, (size_t)num_sms * 8 * sizeof(long long), st));
        if ((rc = launch_chain(c, fake, M, nullptr, num_sms, st, d_prof)) != CTR_OK) return rc;     // warm-up
    }
    if ((rc = launch_chain(c, fake, M, nullptr, num_sms, st, d_prof)) != CTR_OK) return rc;
    e->launches.fetch_add(1, std::memory_order_relaxed);
    if (prof) {
        std::vector<long long> h((size_t)num_sms * 8);
        CUDA_TRY(cudaMemcpyAsync(h.data(), d_prof, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        const char* names[8] = {"producer wait empty", "mma wait acc_empty", "mma wait full", "mma issue+commit", "mma loop total",
                                "epi wait acc_full", "epi body", "producer wait layer"};
        fprintf(stderr, "[chain prof] M=%d N=%d K=%d BN=%d (cycles; CTA0 leader / CTA1 peer / CTA2 / mean over leaders)\n", M, N, K, c.BN);
        for (int i = 0; i < 8; ++i) {
            double mean = 0; int cnt = 0;
            for (int cta = 0; cta < num_sms; cta += 2) if (h[cta * 8 + 4]) { mean += (double)h[cta * 8 + i]; ++cnt; }
            fprintf(stderr, "  %-22s %10lld %10lld %10lld   mean %.0f\n", names[i], h[i], h[8 + i], h[16 + i], cnt ? mean / cnt : 0.0);
        }
        cudaFree(d_prof);
    }
    std::vector<__half> h((size_t)M * L.Npad);
    CUDA_TRY(cudaMemcpyAsync(h.data(), dO, h.size() * sizeof(__half), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) out[(size_t)m * N + n] = __half2float(h[(size_t)m * L.Npad + n]);
    cudaFree(dA32); cudaFree(dA); cudaFree(dO); cudaFree(L.W); cudaFree(L.bias);
    return CTR_OK;
}

int32_t ctr_debug_gather(ctr_engine* e, int32_t device_index, const int64_t* ids, const float* wts, int64_t n,
                         float* x_out, float* wide_out) {
    if (!e || !ids || !wts || n < 1) return fail(CTR_ERR_INVALID, "ctr_debug_gather: bad argument");
    if (!e->loaded) return fail(CTR_ERR_STATE, "weights not loaded");
    if (device_index < 0 || device_index >= e->cfg.n_devices) return fail(CTR_ERR_INVALID, "device index out of range");
    DeviceCtx& dc = *e->devs[device_index];
    std::lock_guard<std::mutex> lk(dc.dev_mu);
    Slot& s = *dc.dev_slots[0];
    const ModelDesc& m = e->m;
    if ((e->m.kind != CTR_MODEL_LR && n > s.x_rows) || n > s.cap) return fail(CTR_ERR_CAPACITY, "ctr_debug_gather: n exceeds the chunk size");
    CUDA_TRY(cudaSetDevice(s.dev));
    CUDA_TRY(cudaMemcpyAsync(s.d_ids, ids, (size_t)n * m.F * sizeof(int64_t), cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.d_wts, wts, (size_t)n * m.F * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    // run the whole chunk (the gather is its first kernel), then read the gather's outputs back
    s.time_this = false;
    int32_t rc = score_chunk(e, dc, s, s.d_ids, s.d_wts, (int)n, s.d_scores);
    if (rc != CTR_OK) return rc;
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    if (x_out && m.kind != CTR_MODEL_LR) {
        std::vector<__half> h((size_t)n * s.ld[0]);
        CUDA_TRY(cudaMemcpy(h.data(), s.buf[0], h.size() * sizeof(__half), cudaMemcpyDeviceToHost));
        for (int64_t b = 0; b < n; ++b)
            for (int j = 0; j < m.d(); ++j) x_out[b * m.d() + j] = __half2float(h[(size_t)b * s.ld[0] + j]);
    }
    if (wide_out && m.kind != CTR_MODEL_LR && m.kind != CTR_MODEL_DCN_V2)
        CUDA_TRY(cudaMemcpy(wide_out, s.d_wide, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
    return CTR_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// ctr_score_device_static: device-resident scoring that allocates nothing.
//
// Who needs it.  ctr_score_device works out of activation buffers the engine cudaMalloc'ed when the
// weights were loaded.  Some hosts cannot afford that: a serving process whose HBM is already handed
// out in full to another allocator (a framework's caching pool, a memory-budgeted sidecar), or a
// request thread that must never enter the driver's allocation path because a cudaMalloc can stall
// every stream of the device.  For those callers the tower's intermediate activations live here, in
// one fixed array that is part of the loaded library image: every GPU that loads libctr_b200.so
// holds these bytes from load until the process exits, whether or not this entry point is used.
//
// Limits.  At most CTR_STATIC_MAX_CANDIDATES candidates per call, d_padded <= 768 and hidden widths
// <= 1024 (all the configurations of BASELINE.json fit).  The array is shared by every call, every
// engine and every stream on that GPU and keeps whatever the previous call left in it: the regions
// that must start at zero (the padding columns of the gathered activations, the out-of-range flag)
// are cleared on the caller's stream at the start of each call, and two calls whose GPU work overlaps
// in time race on it -- callers serialise them (same stream, or an event between streams).
constexpr int kStaticMaxCandidates = CTR_STATIC_MAX_CANDIDATES;
constexpr int kStaticMaxDpad = 768;
constexpr int kStaticMaxHidden = 1024;
constexpr size_t kStaticXBytes = (size_t)kStaticMaxCandidates * kStaticMaxDpad * sizeof(__half);     // X, XL0, XL1, X0 copy
constexpr size_t kStaticHBytes = (size_t)kStaticMaxCandidates * kStaticMaxHidden * sizeof(__half);   // H0, H1
constexpr size_t kStaticVecBytes = (size_t)kStaticMaxCandidates * sizeof(float);                     // wide, extra
constexpr size_t kStaticScratchBytes = 4 * kStaticXBytes + 2 * kStaticHBytes + 2 * kStaticVecBytes + 256;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {
// In host code `g_static_scratch` is only a handle for the runtime, not an address a kernel can use.
// The device address is looked up once per GPU, with that GPU current, and cached here.
struct StaticScratchAddr { std::mutex mu; unsigned char* ptr[64] = {nullptr}; };
StaticScratchAddr g_static_addr;

int32_t static_scratch_for_device(int dev, unsigned char** out) {
    if (dev < 0 || dev >= 64) return fail(CTR_ERR_INVALID, "ctr_score_device_static: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_static_addr.mu);
    if (!g_static_addr.ptr[dev]) {
        void* p = nullptr;
        CUDA_TRY(cudaSetDevice(dev));
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail(CTR_ERR_CUDA, "ctr_score_device_static: cudaGetSymbolAddress returned a null address");
        g_static_addr.ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_static_addr.ptr[dev];
    return CTR_OK;
}
}  // namespace

extern "C" int32_t ctr_score_device_static(ctr_engine* e, int32_t device_index, const void* d_ids, const void* d_wts,
                                           int64_t n_cand, void* d_scores, void* stream) {
    if (!e) return fail(CTR_ERR_INVALID, "ctr_score_device_static: null engine");
    if (!e->loaded) return fail(CTR_ERR_STATE, "ctr_score_device_static: weights not loaded");
    if (device_index < 0 || device_index >= e->cfg.n_devices) return fail(CTR_ERR_INVALID, "ctr_score_device_static: device index out of range");
    if (n_cand < 0 || (n_cand > 0 && (!d_ids || !d_wts || !d_scores))) return fail(CTR_ERR_INVALID, "ctr_score_device_static: null buffer or negative count");
    if (n_cand > kStaticMaxCandidates)
        return fail(CTR_ERR_CAPACITY, "ctr_score_device_static: " + std::to_string(n_cand) + " candidates exceed the fixed scratch (CTR_STATIC_MAX_CANDIDATES = " +
                                          std::to_string(kStaticMaxCandidates) + "); split the request or use ctr_score_device");
    const ModelDesc& m = e->m;
    DeviceCtx& dc = *e->devs[device_index];
    DeviceModel& dm = dc.model;
    if (m.d_pad() > kStaticMaxDpad) return fail(CTR_ERR_CAPACITY, "ctr_score_device_static: model input width exceeds the fixed scratch (d_padded <= 768)");
    for (auto& L : dm.mlp)
        if (L.Npad > kStaticMaxHidden) return fail(CTR_ERR_CAPACITY, "ctr_score_device_static: hidden layer wider than the fixed scratch (<= 1024)");
    if (n_cand == 0) return CTR_OK;

    unsigned char* base = nullptr;
    int32_t rc = static_scratch_for_device(dm.dev, &base);      // first call on this GPU resolves the address
    if (rc != CTR_OK) return rc;
    CUDA_TRY(cudaSetDevice(dm.dev));
    cudaStream_t user = reinterpret_cast<cudaStream_t>(stream);
    cudaStream_t st = user ? user : dc.main_stream;

    // A slot whose buffers are slices of the fixed array; nothing in it is allocated or freed.
    Slot s;
    s.dev = dm.dev; s.stream = st; s.owns_stream = false;
    s.cap = kStaticMaxCandidates; s.x_rows = kStaticMaxCandidates; s.scratch_rows = kStaticMaxCandidates / 2;   // 2 CTA pairs x 2 tiles each
    unsigned char* p = base;
    s.buf[0] = reinterpret_cast<__half*>(p); p += kStaticXBytes;
    s.buf[1] = reinterpret_cast<__half*>(p); p += kStaticHBytes;
    s.buf[2] = reinterpret_cast<__half*>(p); p += kStaticHBytes;
    s.buf[3] = reinterpret_cast<__half*>(p); p += kStaticXBytes;
    s.buf[4] = reinterpret_cast<__half*>(p); p += kStaticXBytes;
    s.buf[5] = reinterpret_cast<__half*>(p); p += kStaticXBytes;
    s.d_wide = reinterpret_cast<float*>(p); p += kStaticVecBytes;
    s.d_extra = reinterpret_cast<float*>(p); p += kStaticVecBytes;
    s.d_err = reinterpret_cast<int32_t*>(p);
    int maxn = 0;
    for (auto& L : dm.mlp) maxn = std::max(maxn, L.Npad);
    const int xl_ld = dm.cross.empty() ? m.d_pad() : dm.cross.back().Npad;
    s.ld[0] = m.d_pad(); s.ld[1] = maxn; s.ld[2] = maxn; s.ld[3] = xl_ld; s.ld[4] = xl_ld; s.ld[5] = xl_ld;
    if (m.kind != CTR_MODEL_LR) {
        if ((rc = build_chains(e, dm, s)) != CTR_OK) return rc;      // tensor maps over the slices (host-side encode only)
        // the previous call (any engine, any model) may have left anything here: clear what must be zero
        const size_t x_bytes = (size_t)round_up((int)n_cand, 2 * kBM) * s.ld[0] * sizeof(__half);
        CUDA_TRY(cudaMemsetAsync(s.buf[0], 0, x_bytes, st));
    }
    CUDA_TRY(cudaMemsetAsync(s.d_err, 0, sizeof(int32_t), st));

    rc = score_chunk(e, dc, s, static_cast<const int64_t*>(d_ids), static_cast<const float*>(d_wts), (int)n_cand,
                     static_cast<float*>(d_scores));
    s.stream = nullptr;                       // nothing of this slot is owned: keep free_slot-style cleanup away from it
    if (rc != CTR_OK) return rc;
    if (!user) {
        CUDA_TRY(cudaStreamSynchronize(st));
        int32_t flag = 0;
        CUDA_TRY(cudaMemcpy(&flag, base + (kStaticScratchBytes - 256), sizeof(flag), cudaMemcpyDeviceToHost));
        if (flag) return fail(CTR_ERR_INVALID, "feature id out of range for a direct-indexed table");
    }
    return CTR_OK;
}
