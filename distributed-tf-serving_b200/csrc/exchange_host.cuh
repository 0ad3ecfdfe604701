// exchange_host.cuh -- host side of the NCCL exchange path (kernels and buffers: exchange.cuh).
#pragma once
namespace {

void free_exchange(ctr_engine* e) {
    ExchangeState& x = e->exch;
    for (size_t d = 0; d < x.bufs.size(); ++d) {
        cudaSetDevice(e->cfg.device_ids[d]);
        ExchangeBufs& b = x.bufs[d];
        cudaFree(b.counts); cudaFree(b.send_rows); cudaFree(b.req_rows); cudaFree(b.reply);
        for (int i = 0; i < 2; ++i) { cudaFree(b.slot_of_pos[i]); cudaFree(b.got[i]); if (b.scored[i]) cudaEventDestroy(b.scored[i]); }
        if (b.exchanged) cudaEventDestroy(b.exchanged);
        cudaFreeHost(b.h_counts);
    }
    x.bufs.clear();
    if (nccl_api().ok) for (ncclComm_t c : x.comms) nccl_api().CommDestroy(c);
    x.comms.clear();
    x.ready = false;
}

// Communicators (once per engine) and the buffers of one chunk on every device (once per loaded model).
int32_t setup_exchange(ctr_engine* e) {
    ExchangeState& x = e->exch;
    if (x.ready) return CTR_OK;
    NcclApi& nc = nccl_api();
    if (!nc.ok) return fail(CTR_ERR_STATE, "ctr_score_devices_exchange: libnccl.so.2 could not be loaded");
    const int G = e->cfg.n_devices;
    const ModelDesc& m = e->m;
    if (x.comms.empty()) {
        x.comms.resize(G);
        NCCL_TRY(nc.CommInitAll(x.comms.data(), G, e->cfg.device_ids));
    }
    // a chunk of the exchange: as many candidates as the tower scores per launch, but at most 2^16 (the reply
    // buffers hold chunk * F rows of up to 128 bytes)
    x.chunk = std::min(e->devs[0]->dev_slots[0]->x_rows, 1 << 16);
    const size_t elems = (size_t)x.chunk * m.F;
    x.cap_req = elems * 2;                          // what one device may be asked for: twice its even share of G chunks
    const int shift = m.table_dtype == CTR_TABLE_FP32 ? 7 : 6;
    x.bufs.assign(G, ExchangeBufs());
    for (int d = 0; d < G; ++d) {
        CUDA_TRY(cudaSetDevice(e->cfg.device_ids[d]));
        ExchangeBufs& b = x.bufs[d];
        CUDA_TRY(cudaMalloc(&b.counts, CTR_MAX_DEVICES * sizeof(uint32_t)));
        CUDA_TRY(cudaMalloc(&b.send_rows, elems * sizeof(uint32_t)));
        CUDA_TRY(cudaMalloc(&b.req_rows, x.cap_req * sizeof(uint32_t)));
        CUDA_TRY(cudaMalloc(&b.reply, x.cap_req << shift));
        for (int i = 0; i < 2; ++i) {
            CUDA_TRY(cudaMalloc(&b.slot_of_pos[i], elems * sizeof(uint32_t)));
            CUDA_TRY(cudaMalloc(&b.got[i], elems << shift));
            CUDA_TRY(cudaEventCreateWithFlags(&b.scored[i], cudaEventDisableTiming));
        }
        CUDA_TRY(cudaEventCreateWithFlags(&b.exchanged, cudaEventDisableTiming));
        CUDA_TRY(cudaMallocHost(&b.h_counts, CTR_MAX_DEVICES * sizeof(uint32_t)));
    }
    x.ready = true;
    return CTR_OK;
}

}  // namespace

extern "C" int32_t ctr_score_devices_exchange(ctr_engine* e, const void* const* d_ids, const void* const* d_wts, const int64_t* n_cand,
                                              void* const* d_scores) {
    if (!e || !d_ids || !d_wts || !n_cand || !d_scores) return fail(CTR_ERR_INVALID, "ctr_score_devices_exchange: null argument");
    if (!e->loaded) return fail(CTR_ERR_STATE, "ctr_score_devices_exchange: weights not loaded");
    const int G = e->cfg.n_devices;
    if (e->cfg.shard_mode != CTR_TABLES_ROW_SHARDED || G < 2) return fail(CTR_ERR_STATE, "ctr_score_devices_exchange: the engine's tables are not row-sharded over several devices");
    for (int d = 0; d < G; ++d)
        if (n_cand[d] < 0 || (n_cand[d] > 0 && (!d_ids[d] || !d_wts[d] || !d_scores[d]))) return fail(CTR_ERR_INVALID, "ctr_score_devices_exchange: null buffer or negative count");
    std::lock_guard<std::mutex> lk(e->exch.mu);
    int32_t rc = setup_exchange(e);
    if (rc != CTR_OK) return rc;
    ExchangeState& x = e->exch;
    NcclApi& nc = nccl_api();
    const ModelDesc& m = e->m;
    const int F = m.F;
    const int shift = m.table_dtype == CTR_TABLE_FP32 ? 7 : 6;
    const int hashed = m.id_mode == CTR_ID_HASHED;
    int64_t most = 0;
    for (int d = 0; d < G; ++d) most = std::max(most, n_cand[d]);
    std::vector<std::unique_lock<std::mutex>> dev_locks;          // this call owns every device's engine streams
    for (int d = 0; d < G; ++d) dev_locks.emplace_back(e->devs[d]->dev_mu);

    for (int d = 0; d < G; ++d) {                                   // the out-of-range flag covers the whole call
        CUDA_TRY(cudaSetDevice(e->devs[d]->model.dev));
        CUDA_TRY(cudaMemsetAsync(e->devs[d]->dev_slots[0]->d_err, 0, sizeof(int32_t), e->devs[d]->main_stream));
    }
    uint32_t cnt[CTR_MAX_DEVICES][CTR_MAX_DEVICES];               // cnt[d][p]: lookups device d sends to owner p
    int64_t chunk_no = 0;
    for (int64_t c0 = 0; c0 < most; c0 += x.chunk, ++chunk_no) {
        const int pb = (int)(chunk_no & 1);             // buffer set of this chunk; chunk_no - 2 used it last
        int n_here[CTR_MAX_DEVICES];
        // ---- bucket (count)
        for (int d = 0; d < G; ++d) {
            n_here[d] = (int)std::max<int64_t>(0, std::min<int64_t>(x.chunk, n_cand[d] - c0));
            DeviceCtx& dc = *e->devs[d];
            ExchangeBufs& b = x.bufs[d];
            CUDA_TRY(cudaSetDevice(dc.model.dev));
            cudaStream_t st = dc.main_stream;
            if (chunk_no >= 2) CUDA_TRY(cudaStreamWaitEvent(st, b.scored[pb], 0));      // its score stage has let go of set pb
            CUDA_TRY(cudaMemsetAsync(b.counts, 0, CTR_MAX_DEVICES * sizeof(uint32_t), st));
            if (n_here[d]) {
                const int64_t elems = (int64_t)n_here[d] * F;
                exch_count_kernel<<<dc.model.num_sms * 4, 256, 0, st>>>(static_cast<const int64_t*>(d_ids[d]) + c0 * F, elems, (uint64_t)m.V, hashed, G, b.counts);
                CUDA_TRY(cudaGetLastError());
                e->launches.fetch_add(1, std::memory_order_relaxed);
            }
            CUDA_TRY(cudaMemcpyAsync(b.h_counts, b.counts, CTR_MAX_DEVICES * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        }
        for (int d = 0; d < G; ++d) {
            CUDA_TRY(cudaSetDevice(e->devs[d]->model.dev));
            CUDA_TRY(cudaStreamSynchronize(e->devs[d]->main_stream));
            for (int p = 0; p < G; ++p) cnt[d][p] = x.bufs[d].h_counts[p];
        }
        for (int p = 0; p < G; ++p) {
            size_t asked = 0;
            for (int d = 0; d < G; ++d) asked += cnt[d][p];
            if (asked > x.cap_req) return fail(CTR_ERR_CAPACITY, "ctr_score_devices_exchange: one shard is asked for more rows than its reply buffer holds (ids too skewed towards one owner)");
        }
        // ---- bucket (scatter): cursors start at the bucket offsets
        for (int d = 0; d < G; ++d) {
            if (!n_here[d]) continue;
            DeviceCtx& dc = *e->devs[d];
            ExchangeBufs& b = x.bufs[d];
            CUDA_TRY(cudaSetDevice(dc.model.dev));
            cudaStream_t st = dc.main_stream;
            uint32_t off[CTR_MAX_DEVICES], acc = 0;
            for (int p = 0; p < CTR_MAX_DEVICES; ++p) { off[p] = acc; if (p < G) acc += cnt[d][p]; }
            memcpy(b.h_counts, off, sizeof(off));
            CUDA_TRY(cudaMemcpyAsync(b.counts, b.h_counts, sizeof(off), cudaMemcpyHostToDevice, st));
            exch_scatter_kernel<<<dc.model.num_sms * 4, 256, 0, st>>>(static_cast<const int64_t*>(d_ids[d]) + c0 * F, (int64_t)n_here[d] * F, (uint64_t)m.V, hashed, G,
                                                                       b.counts, b.send_rows, b.slot_of_pos[pb]);
            CUDA_TRY(cudaGetLastError());
            e->launches.fetch_add(1, std::memory_order_relaxed);
        }
        // ---- all-to-all #1: row indices to their owners
        NCCL_TRY(nc.GroupStart());
        for (int d = 0; d < G; ++d) {
            cudaStream_t st = e->devs[d]->main_stream;
            size_t soff = 0, roff = 0;
            for (int p = 0; p < G; ++p) {
                if (cnt[d][p]) NCCL_TRY(nc.Send(x.bufs[d].send_rows + soff, cnt[d][p], ncclUint32, p, x.comms[d], st));
                if (cnt[p][d]) NCCL_TRY(nc.Recv(x.bufs[d].req_rows + roff, cnt[p][d], ncclUint32, p, x.comms[d], st));
                soff += cnt[d][p];
                roff += cnt[p][d];
            }
        }
        NCCL_TRY(nc.GroupEnd());
        // ---- serve: owners copy the requested rows out of their shard
        for (int d = 0; d < G; ++d) {
            uint32_t n_req = 0;
            for (int p = 0; p < G; ++p) n_req += cnt[p][d];
            if (!n_req) continue;
            DeviceCtx& dc = *e->devs[d];
            CUDA_TRY(cudaSetDevice(dc.model.dev));
            exch_serve_kernel<<<dc.model.num_sms * 8, 256, 0, dc.main_stream>>>(dc.model.table, shift, x.bufs[d].req_rows, n_req, x.bufs[d].reply);
            CUDA_TRY(cudaGetLastError());
            e->launches.fetch_add(1, std::memory_order_relaxed);
        }
        // ---- all-to-all #2: the rows back to whoever asked
        NCCL_TRY(nc.GroupStart());
        for (int d = 0; d < G; ++d) {
            cudaStream_t st = e->devs[d]->main_stream;
            size_t soff = 0, roff = 0;
            for (int p = 0; p < G; ++p) {
                if (cnt[p][d]) NCCL_TRY(nc.Send(x.bufs[d].reply + (roff << shift), (size_t)cnt[p][d] << shift, ncclUint8, p, x.comms[d], st));
                if (cnt[d][p]) NCCL_TRY(nc.Recv(x.bufs[d].got[pb] + (soff << shift), (size_t)cnt[d][p] << shift, ncclUint8, p, x.comms[d], st));
                soff += cnt[d][p];
                roff += cnt[p][d];
            }
        }
        NCCL_TRY(nc.GroupEnd());
        // ---- score: gather over the returned rows (compact layout: uint32 slots + the caller's weights), then the tower
        for (int d = 0; d < G; ++d) {
            if (!n_here[d]) continue;
            DeviceCtx& dc = *e->devs[d];
            Slot& s = *dc.dev_slots[0];
            CUDA_TRY(cudaSetDevice(dc.model.dev));
            CUDA_TRY(cudaEventRecord(x.bufs[d].exchanged, dc.main_stream));
            CUDA_TRY(cudaStreamWaitEvent(s.stream, x.bufs[d].exchanged, 0));
            InputView v;
            v.ids = x.bufs[d].slot_of_pos[pb]; v.wts = static_cast<const float*>(d_wts[d]) + c0 * F;
            v.Fk = F; v.Wk = F; v.id_bytes = 4; v.ids_are_rows = true; v.weighted_mask = ~0ull;
            v.table_override = x.bufs[d].got[pb]; v.rows_override = (uint64_t)n_here[d] * F;
            s.time_this = false;
            rc = score_chunk(e, dc, s, v, n_here[d], static_cast<float*>(d_scores[d]) + c0);
            if (rc != CTR_OK) return rc;
            CUDA_TRY(cudaEventRecord(x.bufs[d].scored[pb], s.stream));       // the exchange of chunk_no + 2 waits for this
        }
    }
    for (int d = 0; d < G; ++d) {                                   // everything scored before the call returns
        DeviceCtx& dc = *e->devs[d];
        CUDA_TRY(cudaSetDevice(dc.model.dev));
        CUDA_TRY(cudaStreamSynchronize(dc.dev_slots[0]->stream));
    }
    int32_t err = 0;
    for (int d = 0; d < G; ++d) {
        DeviceCtx& dc = *e->devs[d];
        CUDA_TRY(cudaSetDevice(dc.model.dev));
        CUDA_TRY(cudaStreamSynchronize(dc.main_stream));
        int32_t v = 0;
        CUDA_TRY(cudaMemcpy(&v, dc.dev_slots[0]->d_err, sizeof(v), cudaMemcpyDeviceToHost));
        err |= v;
    }
    if (err) return fail(CTR_ERR_INVALID, "feature id out of range for a direct-indexed table");
    return CTR_OK;
}
