// engine.cu -- host runtime + C ABI of the B200-native CTR scoring engine (one translation unit).
//
//   engine_base.cuh     error reporting, tensor-map encoders          device_path.cuh     gather + tower chains of one chunk
//   weight_file.cuh     CTRW0001 reader                               workers.cuh         shard tasks, worker threads, tickets
//   engine_state.cuh    model / slot / queue / engine structs         debug_hooks.cuh     ctr_debug_* test hooks
//   device_setup.cuh    table + weight upload, slots, chain plans     static_scratch.cuh  ctr_score_device_static
//   exchange.cuh / exchange_host.cuh   row-sharded lookup as an NCCL all-to-all (ctr_score_devices_exchange)
//   this file           the C ABI (include/ctr_engine.h)
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

#include "engine_base.cuh"
#include "weight_file.cuh"
#include "exchange.cuh"
#include "engine_state.cuh"
#include "sm_split.cuh"
#include "device_setup.cuh"
#include "device_path.cuh"
#include "workers.cuh"

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
    e->host_pack = cfg->host_pack;
    if (const char* c = getenv("CTR_HOST_PACK")) e->host_pack = atoi(c);
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
    if (cfg->n_devices > 1) {       // peer access: row-sharded tables need it; replicated engines use it for the device-side merge
        for (int i = 0; i < cfg->n_devices; ++i)
            for (int j = 0; j < cfg->n_devices; ++j) {
                if (i == j) continue;
                int can = 0;
                CUDA_TRY(cudaDeviceCanAccessPeer(&can, cfg->device_ids[i], cfg->device_ids[j]));
                if (!can) {
                    if (cfg->shard_mode == CTR_TABLES_ROW_SHARDED) return fail(CTR_ERR_CUDA, "row-sharded tables need peer access between all devices");
                    continue;               // replicated tables: peer copies then go through the host, nothing else changes
                }
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
    int32_t rc;
    try {                                   // nothing may unwind through the C boundary
        rc = read_weight_file(path, &hw);
    } catch (const std::exception& ex) {
        return fail(CTR_ERR_IO, std::string("weight file: ") + ex.what());
    }
    if (rc != CTR_OK) return rc;
    stop_workers(e);
    free_device_state(e);
    e->loaded = false;
    e->m = hw.m;
    for (int i = 0; i < e->cfg.n_devices; ++i) {
        if ((rc = load_model_on_device(e, i, hw)) != CTR_OK) return rc;
        e->shard_base[i] = e->devs[i]->model.table;
        if ((rc = alloc_pair_scratch(e, *e->devs[i])) != CTR_OK) return rc;
        if (const char* c = getenv("CTR_SM_SPLIT")) {
            const int g = atoi(c);
            if (g >= 8 && g <= e->devs[i]->model.num_sms - 16 && e->m.kind != CTR_MODEL_LR &&
                (rc = setup_sm_split(e->devs[i]->split, e->cfg.device_ids[i], g)) != CTR_OK) return rc;
        }
    }
    for (int i = 0; i < e->cfg.n_devices; ++i) {
        DeviceCtx& dc = *e->devs[i];
        const int dev = e->cfg.device_ids[i];
        ScopedGpuLocality near_gpu(dev);                       // pinned staging of this GPU is allocated on its socket
        for (int w = 0; w < e->cfg.workers_per_device; ++w) {
            std::unique_ptr<Slot> s(new Slot());
            if ((rc = init_slot_common(*s, dev, nullptr)) != CTR_OK) return rc;
            if ((rc = init_slot_split(dc.split, *s)) != CTR_OK) return rc;
            if ((rc = build_slot_plan(e, dc, *s, true)) != CTR_OK) return rc;
            dc.worker_slots.push_back(std::move(s));
        }
        // One engine stream by default.  With two (CTR_DEVICE_STREAMS=2) chunks alternate between them and chunk k+1's
        // gather runs beside chunk k's tower -- the two kernels can share an SM since round 2 (tower.cuh) -- but they
        // trade time almost one for one (+2 % under a sustained power cap, +3-5 % in short runs),
        // and a kernel's event-to-event time then includes its neighbour, which blurs the per-kernel roofline figures.
        int n_dev_slots = 1;
        if (const char* c = getenv("CTR_DEVICE_STREAMS")) { int v = atoi(c); if (v >= 1 && v <= 8) n_dev_slots = v; }
        for (int w = 0; w < n_dev_slots; ++w) {
            std::unique_ptr<Slot> s(new Slot());
            if ((rc = init_slot_common(*s, dev, nullptr)) != CTR_OK) return rc;
            if ((rc = init_slot_split(dc.split, *s)) != CTR_OK) return rc;
            if ((rc = build_slot_plan(e, dc, *s, false)) != CTR_OK) return rc;
            dc.dev_slots.push_back(std::move(s));
        }
        CUDA_TRY(cudaDeviceSynchronize());
    }
    if (e->cfg.n_devices > 1) {             // merge + sort buffers of multi-GPU sorted requests, on the first GPU
        DeviceCtx& d0 = *e->devs[0];
        const int64_t cap = (int64_t)e->cfg.max_candidates * e->cfg.n_devices;
        CUDA_TRY(cudaSetDevice(d0.model.dev));
        CUDA_TRY(cudaMalloc(&d0.srt_scores, (size_t)cap * sizeof(float)));
        if ((rc = alloc_sort_bufs(d0.srt, cap)) != CTR_OK) return rc;
        d0.sort_cap = cap;
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

// Checks a compact layout against the loaded model and fills the view's derived fields.
static int32_t make_compact_view(ctr_engine* e, const char* who, int32_t n_ctx, uint64_t weighted_mask, uint64_t fixed_id_mask, bool narrow_rows,
                                 InputView* v) {
    if (!e) return fail(CTR_ERR_INVALID, std::string(who) + ": null engine");
    if (!e->loaded) return fail(CTR_ERR_STATE, std::string(who) + ": weights not loaded");
    const int F = e->m.F;
    if (n_ctx < 0 || n_ctx >= F) return fail(CTR_ERR_INVALID, std::string(who) + ": the shared context must cover 0 .. F-1 leading fields");
    const uint64_t field_bits = F >= 64 ? ~0ull : ((1ull << F) - 1ull);
    const uint64_t cand_bits = field_bits & ~((1ull << n_ctx) - 1ull);
    v->n_ctx = n_ctx;
    v->id_mask = field_bits & ~(fixed_id_mask & cand_bits);
    v->Fk = __builtin_popcountll(v->id_mask & cand_bits);
    if (v->Fk < 1) return fail(CTR_ERR_INVALID, std::string(who) + ": at least one field must keep a per-candidate id column");
    v->weighted_mask = weighted_mask & field_bits;
    v->Wk = __builtin_popcountll(weighted_mask & cand_bits);
    v->id_bytes = narrow_rows ? 4 : 8;
    v->ids_are_rows = narrow_rows;
    return CTR_OK;
}

// The request itself (processARequest, DC:137-203): cut the candidates of the host-side view `in` over the
// devices, queue one shard task per device, hand back a ticket.
static int32_t submit_request(ctr_engine* e, const char* who, const InputView& in, int64_t n_cand, float* scores_out, ctr_callback cb,
                              void* user, ctr_ticket* ticket, bool want_sort = false, int32_t* order_out = nullptr, bool* sorted = nullptr,
                              float* d_merge = nullptr) {
    if (!e || !ticket) return fail(CTR_ERR_INVALID, std::string(who) + ": null argument");
    if (!e->loaded || e->shutting_down) return fail(CTR_ERR_STATE, std::string(who) + ": weights not loaded or engine shutting down");
    if (n_cand < 0 || (n_cand > 0 && (!in.ids || (in.Wk > 0 && !in.wts) || !scores_out))) return fail(CTR_ERR_INVALID, std::string(who) + ": null buffer or negative count");
    if ((in.n_ctx > 0 && (!in.ctx_ids || !in.ctx_wts)) || (in.Fk + in.n_ctx < e->m.F && !in.ctx_ids)) return fail(CTR_ERR_INVALID, std::string(who) + ": null context list");
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
            return fail(CTR_ERR_CAPACITY, std::string(who) + ": the last shard (it takes the remainder, DC:46-55) exceeds max_candidates of this engine");
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
    const bool sort_in_stream = want_sort && shards == 1;    // a one-shard request is sorted on its own stream
    if (sorted) *sorted = sort_in_stream;
    for (int i = 0; i < P; ++i) {
        if (en[i] == b[i]) continue;
        Task t;
        t.req = req; t.in = in.at(b[i]); t.n = en[i] - b[i]; t.scores = scores_out + b[i];
        t.sort = sort_in_stream; t.order_out = order_out;
        if (want_sort && !sort_in_stream) t.d_merge = d_merge ? d_merge + b[i] : nullptr;   // ordered merge on GPU 0 (DC:161-164)
        DeviceCtx& dc = *e->devs[(first + i) % G];
        {
            std::lock_guard<std::mutex> lk(dc.q_mu);
            dc.queue.push_back(t);
            dc.queued.fetch_add(1, std::memory_order_release);
        }
        dc.q_cv.notify_one();
    }
    return CTR_OK;
}

int32_t ctr_predict_async(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand, float* scores_out,
                          ctr_callback cb, void* user, ctr_ticket* ticket) {
    if (!e) return fail(CTR_ERR_INVALID, "ctr_predict_async: null engine");
    return submit_request(e, "ctr_predict_async", reference_view(ids, wts, e->m.F), n_cand, scores_out, cb, user, ticket);
}

int32_t ctr_predict_shared_async(ctr_engine* e, const int64_t* ctx_ids, const float* ctx_wts, int32_t n_ctx_fields,
                                 const int64_t* cand_ids, const float* cand_wts, int64_t n_cand, float* scores_out,
                                 ctr_callback cb, void* user, ctr_ticket* ticket) {
    InputView v;
    int32_t rc = make_compact_view(e, "ctr_predict_shared_async", n_ctx_fields, ~0ull, 0, false, &v);
    if (rc != CTR_OK) return rc;
    v.ids = cand_ids; v.wts = cand_wts; v.ctx_ids = ctx_ids; v.ctx_wts = ctx_wts;
    return submit_request(e, "ctr_predict_shared_async", v, n_cand, scores_out, cb, user, ticket);
}

int32_t ctr_predict_packed_async(ctr_engine* e, const ctr_packed_layout* layout, const int64_t* ctx_ids, const float* ctx_wts,
                                 const uint32_t* cand_rows, const float* cand_wts, int64_t n_cand, float* scores_out,
                                 ctr_callback cb, void* user, ctr_ticket* ticket) {
    if (!layout || (layout->struct_size != (int32_t)sizeof(ctr_packed_layout) && layout->struct_size != 16)) return fail(CTR_ERR_INVALID, "ctr_predict_packed_async: ctr_packed_layout missing or of the wrong size (ABI)");
    InputView v;
    int32_t rc = make_compact_view(e, "ctr_predict_packed_async", layout->n_ctx_fields, layout->weighted_mask, layout->struct_size == 16 ? 0 : layout->fixed_id_mask, true, &v);
    if (rc != CTR_OK) return rc;
    if ((uint64_t)e->m.V > 0xFFFFFFFFull) return fail(CTR_ERR_CAPACITY, "ctr_predict_packed_async: the table has more than 2^32 - 1 rows, 32-bit row indices cannot address it");
    v.ids = cand_rows; v.wts = cand_wts; v.ctx_ids = ctx_ids; v.ctx_wts = ctx_wts;
    return submit_request(e, "ctr_predict_packed_async", v, n_cand, scores_out, cb, user, ticket);
}

// The marshaller for the packed layout (what replaces the addAllInt64Val / addAllFloatVal loops of DC:97-108 for a
// caller that keeps the reference's lists): pure host code.
int32_t ctr_pack_candidates(ctr_engine* e, const ctr_packed_layout* layout, const int64_t* ids, const float* wts, int64_t n_cand,
                            uint32_t* cand_rows_out, float* cand_wts_out, int64_t* ctx_ids_out, float* ctx_wts_out) {
    if (!layout || (layout->struct_size != (int32_t)sizeof(ctr_packed_layout) && layout->struct_size != 16)) return fail(CTR_ERR_INVALID, "ctr_pack_candidates: ctr_packed_layout missing or of the wrong size (ABI)");
    InputView v;
    int32_t rc = make_compact_view(e, "ctr_pack_candidates", layout->n_ctx_fields, layout->weighted_mask, layout->struct_size == 16 ? 0 : layout->fixed_id_mask, true, &v);
    if (rc != CTR_OK) return rc;
    if (n_cand < 0 || (n_cand > 0 && (!ids || !wts || !cand_rows_out || (v.Wk > 0 && !cand_wts_out)))) return fail(CTR_ERR_INVALID, "ctr_pack_candidates: null buffer or negative count");
    if (v.n_ctx > 0 && n_cand > 0 && (!ctx_ids_out || !ctx_wts_out)) return fail(CTR_ERR_INVALID, "ctr_pack_candidates: null context output");
    const ModelDesc& m = e->m;
    if ((uint64_t)m.V > 0xFFFFFFFFull) return fail(CTR_ERR_CAPACITY, "ctr_pack_candidates: the table has more than 2^32 - 1 rows");
    const int F = m.F, nc = v.n_ctx, Fk = v.Fk, Wk = v.Wk;
    const int n_fixed = (F - nc) - Fk;
    if ((nc > 0 || n_fixed > 0) && n_cand > 0 && !ctx_ids_out) return fail(CTR_ERR_INVALID, "ctr_pack_candidates: null context output");
    const bool hashed = m.id_mode == CTR_ID_HASHED;
    const uint64_t row_mask = (uint64_t)m.V - 1;
    int idf[CTR_MAX_FIELDS], wf[CTR_MAX_FIELDS], fxf[CTR_MAX_FIELDS];        // fields of the id columns / weight columns / fixed ids
    int ni = 0, nw = 0, nx = 0;
    for (int f = nc; f < F; ++f) {
        if ((v.id_mask >> f) & 1ull) idf[ni++] = f; else fxf[nx++] = f;
        if ((v.weighted_mask >> f) & 1ull) wf[nw++] = f;
    }
    if (n_cand > 0) {
        for (int f = 0; f < nc; ++f) { ctx_ids_out[f] = ids[f]; ctx_wts_out[f] = wts[f]; }
        for (int k = 0; k < nx; ++k) ctx_ids_out[nc + k] = ids[fxf[k]];
    }
    for (int64_t b = 0; b < n_cand; ++b) {
        const int64_t* idr = ids + b * F;
        const float* wtr = wts + b * F;
        for (int f = 0; f < nc; ++f)
            if (idr[f] != ids[f] || memcmp(&wtr[f], &wts[f], sizeof(float)) != 0)
                return fail(CTR_ERR_INVALID, "ctr_pack_candidates: context field " + std::to_string(f) + " of candidate " + std::to_string(b) + " differs from candidate 0");
        for (int k = 0; k < nx; ++k)
            if (idr[fxf[k]] != ids[fxf[k]])
                return fail(CTR_ERR_INVALID, "ctr_pack_candidates: field " + std::to_string(fxf[k]) + " is declared fixed-id but candidate " + std::to_string(b) + " has another id than candidate 0");
        uint32_t* ro = cand_rows_out + b * Fk;
        for (int k = 0; k < ni; ++k) {
            const uint64_t id = (uint64_t)idr[idf[k]];
            // a direct id the table cannot hold becomes row 0xFFFFFFFF: the kernel reports it exactly like the reference layout
            ro[k] = hashed ? (uint32_t)(mix64(id) & row_mask) : (id > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)id);
        }
        for (int f = nc; f < F; ++f)
            if (!((v.weighted_mask >> f) & 1ull) && wtr[f] != 1.0f)
                return fail(CTR_ERR_INVALID, "ctr_pack_candidates: field " + std::to_string(f) + " of candidate " + std::to_string(b) + " has weight " +
                                                 std::to_string(wtr[f]) + " but the layout declares it unweighted (1.0)");
        float* wo = cand_wts_out + b * Wk;
        for (int k = 0; k < Wk; ++k) wo[k] = wtr[wf[k]];
    }
    return CTR_OK;
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
    if (timeout_ms != 0) poll_briefly([&] { return r->done_flag.load(std::memory_order_acquire); });
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

// DC:195 on the device: sorted scores and the candidate index of each.
int32_t ctr_predict_sorted(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand, float* sorted_scores_out,
                           int32_t* order_out) {
    if (!e || (n_cand > 0 && !sorted_scores_out)) return fail(CTR_ERR_INVALID, "ctr_predict_sorted: null argument");
    if (n_cand > INT32_MAX / 2) return fail(CTR_ERR_CAPACITY, "ctr_predict_sorted: too many candidates");
    // One shard (one GPU, or a request below the fan-out threshold): scored and sorted on the same stream, one
    // round trip.  Several shards: every shard copies its scores over NVLink into its slice of one buffer on the
    // engine's first GPU (the ordered merge of DC:161-164, on the device), which sorts them; one copy back.
    DeviceCtx& dc = *e->devs[0];
    const bool multi = e->loaded && e->cfg.n_devices > 1 && dc.sort_cap >= n_cand;
    std::unique_lock<std::mutex> lk(dc.sort_mu, std::defer_lock);
    if (multi) lk.lock();                    // the merge buffer is one per engine: multi-GPU sorted requests take turns
    ctr_ticket tk;
    bool sorted = false;
    int32_t rc = submit_request(e, "ctr_predict_sorted", reference_view(ids, wts, e->m.F), n_cand, sorted_scores_out, nullptr, nullptr, &tk,
                                true, order_out, &sorted, multi ? dc.srt_scores : nullptr);
    if (rc != CTR_OK) return rc;
    rc = ctr_wait(e, tk, -1);
    if (rc != CTR_OK || n_cand == 0 || sorted) return rc;
    if (!multi) lk.lock();
    CUDA_TRY(cudaSetDevice(dc.model.dev));
    const int n = (int)n_cand;
    cudaStream_t st = dc.main_stream;
    if (!multi) {                            // a request beyond what was sized at load: merged on the host, sorted here
        if (dc.sort_cap < n_cand) {
            cudaFree(dc.srt_scores);
            dc.srt_scores = nullptr;
            dc.sort_cap = 0;
            CUDA_TRY(cudaMalloc(&dc.srt_scores, (size_t)n_cand * sizeof(float)));
            if ((rc = alloc_sort_bufs(dc.srt, n_cand)) != CTR_OK) return rc;
            dc.sort_cap = n_cand;
        }
        CUDA_TRY(cudaMemcpyAsync(dc.srt_scores, sorted_scores_out, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, st));
    }
    const int32_t* d_order = nullptr;
    if ((rc = launch_sort(e, dc.model, dc.srt, dc.srt_scores, n, st, &d_order)) != CTR_OK) return rc;
    CUDA_TRY(cudaMemcpyAsync(sorted_scores_out, dc.srt.sorted, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (order_out) CUDA_TRY(cudaMemcpyAsync(order_out, d_order, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
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
    for (int w = 0; w < used; ++w) {
        CUDA_TRY(cudaStreamWaitEvent(dc.dev_slots[w]->stream, dc.fork_ev, 0));
        // a call on a caller's stream never reads the flag: clear it here so that a bad id of that call cannot
        // fail a later, clean call on the engine's stream
        CUDA_TRY(cudaMemsetAsync(dc.dev_slots[w]->d_err, 0, sizeof(int32_t), dc.dev_slots[w]->stream));
    }
    for (int64_t c = 0; c < n_chunks; ++c) {
        Slot& s = *dc.dev_slots[c % used];
        const int64_t c0 = c * step;
        const int cnt = (int)std::min<int64_t>(step, n_cand - c0);
        if (c == 0) { s.time_this = true; s.cur_mark = 0; }  // stage timing: first chunk (slot 0) only
        int32_t rc = score_chunk(e, dc, s, reference_view(ids + c0 * F, wts + c0 * F, F), cnt, scores + c0);
        if (rc != CTR_OK) return rc;
    }
    // join: `origin` continues after every engine stream has finished
    for (int w = 0; w < used; ++w) {
        CUDA_TRY(cudaEventRecord(dc.dev_slots[w]->join_ev, dc.dev_slots[w]->stream));
        CUDA_TRY(cudaStreamWaitEvent(origin, dc.dev_slots[w]->join_ev, 0));
    }
    if (g_timeline_on) { CUDA_TRY(cudaStreamSynchronize(origin)); timeline_dump(); }
    if (!user) {
        CUDA_TRY(cudaStreamSynchronize(origin));
        int32_t err = 0;
        for (int w = 0; w < used; ++w) {
            int32_t v = 0;
            CUDA_TRY(cudaMemcpy(&v, dc.dev_slots[w]->d_err, sizeof(v), cudaMemcpyDeviceToHost));
            if (v) err = 1;
        }
        if (err) return fail(CTR_ERR_INVALID, "feature id out of range for a direct-indexed table");
    }
    return CTR_OK;
}

int64_t ctr_engine_launch_count(ctr_engine* e) { return e ? e->launches.load() : 0; }
int64_t ctr_engine_h2d_bytes(ctr_engine* e) { return e ? e->h2d_bytes.load() : 0; }
int64_t ctr_engine_d2h_bytes(ctr_engine* e) { return e ? e->d2h_bytes.load() : 0; }

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

}  // extern "C"

#include "debug_hooks.cuh"
#include "static_scratch.cuh"
#include "exchange_host.cuh"
