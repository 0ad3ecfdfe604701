// workers.cuh -- the host side of a request: shard tasks, worker threads (the reference's threadPool,
// DC:42), staging + copies + kernels per shard (sendRequest, DC:91-115), completion and tickets.
#pragma once
namespace {
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
            r->done_flag.store(true, std::memory_order_release);
        }
        r->cv.notify_all();
        if (r->cb) {
            r->cb(r->user, r->ticket, r->status.load());
            // A caller that completes its future in the callback never calls ctr_wait: release such tickets
            // here, keeping the most recent ones so that a wait after the callback still finds its ticket.
            std::lock_guard<std::mutex> lk(e->t_mu);
            e->called_back.push_back(r->ticket);
            while (e->called_back.size() > kKeepCalledBack) {
                e->tickets.erase(e->called_back.front());
                e->called_back.pop_front();
            }
        }
    }
}

// Host narrowing of a shard that arrived in the reference's layout (DC:97-108: an int64 id and a float weight per field
// slot): rows[n, F] uint32 + weights[n, Wk] of the fields in `mask`, written behind each other into `slab`.  The loops are
// plain and branch-free so that the compiler vectorises them.  Returns the fields that carry a weight != 1.0 but are not
// in `mask` (the caller adds them and packs again); *wide is set when some id does not fit 32 bits.
constexpr int kNarrowWhenActive = 5;
constexpr int kNarrowMinWorkers = 8;         // the automatic mode needs this many worker threads (and so cores) on the GPU
inline int narrow_when_active() {            // CTR_NARROW_AT=<shards in flight on the GPU from which a worker narrows its shard>
    static const int v = [] { const char* c = getenv("CTR_NARROW_AT"); const int n = c ? atoi(c) : 0; return n >= 1 ? n : kNarrowWhenActive; }();
    return v;
}

uint64_t narrow_reference(const int64_t* ids, const float* wts, int64_t n, int F, uint64_t mask, uint32_t* rows, float* pw, bool* wide) {
    int wf[CTR_MAX_FIELDS], Wk = 0;
    for (int f = 0; f < F; ++f) if ((mask >> f) & 1ull) wf[Wk++] = f;
    uint32_t notone[CTR_MAX_FIELDS] = {0};
    uint64_t hi = 0;
    for (int64_t b = 0; b < n; ++b) {
        const int64_t* idr = ids + b * F;
        const float* wtr = wts + b * F;
        uint32_t* ro = rows + b * F;
        float* wo = pw + b * Wk;
        // the pass is bound by the core's DRAM read rate: touching the lines 16 / 32 candidates ahead keeps more of them in
        // flight than the hardware prefetcher does on its own (0.62 -> 0.51 ms per 10 000 candidates measured)
        for (int o = 0; o < F * 8; o += 64) __builtin_prefetch(reinterpret_cast<const char*>(idr + 16 * F) + o, 0, 3);
        for (int o = 0; o < F * 4; o += 64) __builtin_prefetch(reinterpret_cast<const char*>(wtr + 32 * F) + o, 0, 3);
        for (int f = 0; f < F; ++f) { ro[f] = (uint32_t)idr[f]; hi |= (uint64_t)idr[f]; }
        for (int f = 0; f < F; ++f) notone[f] |= (uint32_t)(wtr[f] != 1.0f);
        for (int k = 0; k < Wk; ++k) wo[k] = wtr[wf[k]];
    }
    *wide = (hi >> 32) != 0;
    uint64_t more = 0;
    for (int f = 0; f < F; ++f) if (notone[f] && !((mask >> f) & 1ull)) more |= 1ull << f;
    return more;
}

// sendRequest (DC:91-115) for one shard: stage -> H2D -> kernels -> D2H -> hand back.
void run_task(ctr_engine* e, DeviceCtx& dc, Slot& s, const Task& t) {
    auto body = [&]() -> int32_t {
        CUDA_TRY(cudaSetDevice(s.dev));
        InputView in = t.in;
        // Host narrowing (ctr_config.host_pack): a reference-layout shard for a direct-id table becomes uint32 rows +
        // the weights of the fields that ever had one, packed into this slot's pinned slab -- 208 instead of 468 bytes
        // per candidate for the Criteo shape, the same kernels and the same bits as ctr_predict_packed_async.  A core
        // narrows ~9.5 GB/s of input (0.49 ms per 10 000 candidates, bound by its own DRAM read rate), the link moves
        // 41-48 GB/s shared by the shards in flight.  Measured end to end on one GPU (profiles/r02_narrow_sweep.txt):
        // 3 workers saturate the link at 103 M candidates/s and narrowing their shards gives 62 M, so few shards in flight
        // travel as they are; 12 workers: 103 M never narrowing, 164-174 M always, 179-181 M from 5 shards in flight on
        // (173-175 M from 8), which is the automatic mode.  A rule driven by a clock of the copies queued on the link was
        // tried as well and mixed worse at 12 workers (143-157 M), better at 6 (121-127 against 103-111 M).  The cores have
        // to be there: 4 GPUs with 7 workers each on a 32-core host lose with it (284 against 396 M: all the workers narrow
        // at once and the host's memory cannot feed them; narrowing only the shards that find the link busy gives 360 M there),
        // so the automatic mode also wants 8 or more workers on the GPU.
        const int hp = e->host_pack;
        if (in.reference_layout(e->m.F) && e->m.id_mode == CTR_ID_DIRECT && hp != 1 && t.n > 0 &&
            (hp == 2 || (e->cfg.workers_per_device >= kNarrowMinWorkers && dc.active.load(std::memory_order_relaxed) >= narrow_when_active()))) {
            const int F = e->m.F;
            uint32_t* rows = reinterpret_cast<uint32_t*>(s.h_ids);
            float* pw = reinterpret_cast<float*>(rows + (size_t)t.n * F);
            const uint64_t all = F >= 64 ? ~0ull : (1ull << F) - 1ull;
            for (int attempt = 0; attempt <= F; ++attempt) {
                const uint64_t mask = e->narrow_weighted.load(std::memory_order_relaxed) & all;
                bool wide = false;
                const uint64_t more = narrow_reference(static_cast<const int64_t*>(t.in.ids), t.in.wts, t.n, F, mask, rows, pw, &wide);
                if (wide) break;                               // an id beyond 32 bits: this shard travels as it is
                if (more) { e->narrow_weighted.fetch_or(more, std::memory_order_relaxed); continue; }
                in = InputView();
                in.ids = rows; in.wts = pw; in.Fk = F; in.Wk = __builtin_popcountll(mask); in.id_bytes = 4;
                in.ids_are_rows = true; in.weighted_mask = mask;
                break;
            }
        }
        const size_t ib = in.ids_bytes(t.n), wb = in.wts_bytes(t.n), sb = (size_t)t.n * sizeof(float);
        // The feature packer: caller buffers that came from ctr_alloc_pinned are DMA'd in place;
        // pageable ones (a Java heap array, a numpy array) are packed into this slot's pinned slab.
        const void* src_ids = in.ids;
        const float* src_wts = in.wts;
        const bool in_slab = in.ids == static_cast<const void*>(s.h_ids);          // narrowed above: already in pinned memory
        if (!in_slab && !is_pinned(e, in.ids, ib)) { memcpy(s.h_ids, in.ids, ib); src_ids = s.h_ids; }
        if (!in_slab && wb && !is_pinned(e, in.wts, wb)) { memcpy(s.h_wts, in.wts, wb); src_wts = s.h_wts; }
        const bool out_pinned = is_pinned(e, t.scores, sb);
        float* dst_scores = out_pinned ? t.scores : s.h_scores;
        // Whatever the layout (reference, shared context, packed rows), the lists cross the host link as they
        // are and the gather kernel reads them as they are: the shared context rides in the kernel parameters.
        InputView dv = in;
        dv.ids = s.d_ids; dv.wts = s.d_wts;
        const bool one_copy = wb && reinterpret_cast<const uint8_t*>(src_wts) == static_cast<const uint8_t*>(src_ids) + ib &&
                              ib + wb <= (size_t)s.cap * e->m.F * sizeof(int64_t);
        if (one_copy) {
            // the weights start right where the ids end (a caller that packs a request into one slab): one DMA instead of
            // two -- small copies are what keeps the compact layouts below the link rate (37 against 49 GB/s measured)
            CUDA_TRY(cudaMemcpyAsync(s.d_ids, src_ids, ib + wb, cudaMemcpyHostToDevice, s.stream));
            dv.wts = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(s.d_ids) + ib);
        } else {
            CUDA_TRY(cudaMemcpyAsync(s.d_ids, src_ids, ib, cudaMemcpyHostToDevice, s.stream));
            if (wb) CUDA_TRY(cudaMemcpyAsync(s.d_wts, src_wts, wb, cudaMemcpyHostToDevice, s.stream));
        }
        e->h2d_bytes.fetch_add((int64_t)(ib + wb), std::memory_order_relaxed);
        e->d2h_bytes.fetch_add((int64_t)sb + (t.sort && t.order_out ? (int64_t)sb : 0), std::memory_order_relaxed);
        int32_t rc = score_on_slot(e, dc, s, dv, t.n, s.d_scores);
        if (rc != CTR_OK) return rc;
        const float* d_result = s.d_scores;
        if (t.sort) {                                          // DC:195 before the scores leave the device
            const int32_t* d_order = nullptr;
            if ((rc = alloc_sort_bufs(s.srt, s.cap)) != CTR_OK) return rc;
            if (!s.h_order) CUDA_TRY(cudaMallocHost(&s.h_order, (size_t)s.cap * sizeof(int32_t)));
            if ((rc = launch_sort(e, dc.model, s.srt, s.d_scores, (int)t.n, s.stream, &d_order)) != CTR_OK) return rc;
            d_result = s.srt.sorted;
            if (t.order_out) CUDA_TRY(cudaMemcpyAsync(s.h_order, d_order, (size_t)t.n * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        }
        if (t.d_merge) {                                       // ordered merge on the device: straight into GPU 0's buffer
            CUDA_TRY(cudaMemcpyPeerAsync(t.d_merge, e->devs[0]->model.dev, d_result, s.dev, sb, s.stream));
        } else {
            CUDA_TRY(cudaMemcpyAsync(dst_scores, d_result, sb, cudaMemcpyDeviceToHost, s.stream));
        }
        CUDA_TRY(cudaMemcpyAsync(s.h_err, s.d_err, sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        if (t.sort && t.order_out) memcpy(t.order_out, s.h_order, (size_t)t.n * sizeof(int32_t));
        if (*s.h_err) {
            CUDA_TRY(cudaMemsetAsync(s.d_err, 0, sizeof(int32_t), s.stream));
            return fail(CTR_ERR_INVALID, "feature id out of range for a direct-indexed table");
        }
        if (!out_pinned && !t.d_merge) memcpy(t.scores, s.h_scores, sb);
        return CTR_OK;
    };
    dc.active.fetch_add(1, std::memory_order_relaxed);
    int32_t rc = body();
    dc.active.fetch_sub(1, std::memory_order_relaxed);
    complete_shard(e, t.req, rc, rc == CTR_OK ? std::string() : g_last_error);
}

void worker_main(ctr_engine* e, DeviceCtx* dc, Slot* s) {
    cudaSetDevice(s->dev);
    cpu_set_t local;
    if (gpu_local_cpus(s->dev, &local)) sched_setaffinity(0, sizeof(local), &local);   // stay next to the GPU
    for (;;) {
        Task t;
        poll_briefly([&] { return dc->queued.load(std::memory_order_acquire) > 0; });
        {
            std::unique_lock<std::mutex> lk(dc->q_mu);
            dc->q_cv.wait(lk, [&] { return !dc->queue.empty(); });
            t = dc->queue.front();
            if (t.stop) return;                 // leave the stop marker for the other workers
            dc->queue.pop_front();
            dc->queued.fetch_sub(1, std::memory_order_relaxed);
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
            Task t; t.stop = true;
            dc.queue.push_back(t);
            dc.queued.fetch_add(1, std::memory_order_release);
        }
        dc.q_cv.notify_all();
        for (auto& th : dc.threads) th.join();
        dc.threads.clear();
        // a request that raced the stop marker must still complete, or its ctr_wait(-1) would never return
        std::deque<Task> left;
        {
            std::lock_guard<std::mutex> lk(dc.q_mu);
            left.swap(dc.queue);
            dc.queued.store(0);
        }
        for (Task& t : left)
            if (!t.stop && t.req) complete_shard(e, t.req, CTR_ERR_STATE, "engine stopped (weights reloaded or engine destroyed) before the request ran");
    }
}

void free_exchange(ctr_engine* e);          // exchange_host.cuh

void free_device_state(ctr_engine* e) {
    free_exchange(e);
    for (auto& dcp : e->devs) {
        DeviceCtx& dc = *dcp;
        for (auto& s : dc.worker_slots) free_slot(*s);
        for (auto& s : dc.dev_slots) free_slot(*s);
        dc.worker_slots.clear();
        dc.dev_slots.clear();
        free_pair_scratch(dc);
        free_sm_split(dc.split);
        if (dc.model.dev >= 0) free_model(dc.model);
        if (dc.sort_cap) {
            cudaFree(dc.srt_scores);
            free_sort_bufs(dc.srt);
            dc.sort_cap = 0;
        }
    }
}

}  // namespace
