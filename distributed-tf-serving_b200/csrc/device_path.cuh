// device_path.cuh -- the kernels of one chunk of candidates on a slot's stream: gather, then the chains.
#pragma once
namespace {
// ------------------------------------------------------------------ the device path for one chunk
// Shared-memory carve-out of the gather kernels.  Measured on B200 (tools/probes/coresidency_probe.cu, profiles/r02_*):
// a gather CTA only becomes resident beside the tower's CTA when BOTH kernels prefer the largest carve-out (an SM
// configured for the tower's 196 KB cannot take the gather's 30 KB, and the other way round, until it drains) --
// but with the largest carve-out the gather loses its L1 (it re-reads the line of a row for the first-order weight
// and keeps the dense fields' rows there) and runs 13 % slower (340 against 299 us per 303 104 candidates), and the
// two kernels side by side only trade time with each other (tower 298 -> 443 us per chunk while a 175 us gather
// hides under it).  Net: 330 M candidates/s with the driver's default carve-out against 319-325 M with co-residency,
// so the default stays; CTR_GATHER_MAX_CARVEOUT=1 switches the co-resident configuration on.
template <typename Kern>
void prefer_max_carveout(Kern k) {
    static const bool on = getenv("CTR_GATHER_MAX_CARVEOUT") != nullptr;
    if (!on) return;
    static std::atomic<uint64_t> done{0};                     // one bit per device ordinal, per kernel instantiation
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
    if (done.load(std::memory_order_relaxed) & (1ull << dev)) return;
    if (cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) != cudaSuccess) (void)cudaGetLastError();
    done.fetch_or(1ull << dev, std::memory_order_relaxed);
}

// block-tile kernel: the reference layout (C = false) or a compact layout expanded per tile (C = true)
template <typename Kern>
void allow_smem(Kern k, size_t smem) {           // more than the 48 KB a kernel gets without asking (48 fields of int64 ids + weights)
    if (smem > 48 * 1024 && cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) (void)cudaGetLastError();
}
template <int TD, bool H, bool E, bool L, bool FM, bool C>
void launch_tile_t(const GatherParams& gp, int grid, size_t smem, cudaStream_t st) {
    if (gp.n_shards > 1) {
        prefer_max_carveout(gather_tile_kernel<TD, H, E, L, FM, true, C>);
        allow_smem(gather_tile_kernel<TD, H, E, L, FM, true, C>, smem);
        gather_tile_kernel<TD, H, E, L, FM, true, C><<<grid, kTileThreads, smem, st>>>(gp);
    } else {
        prefer_max_carveout(gather_tile_kernel<TD, H, E, L, FM, false, C>);
        allow_smem(gather_tile_kernel<TD, H, E, L, FM, false, C>, smem);
        gather_tile_kernel<TD, H, E, L, FM, false, C><<<grid, kTileThreads, smem, st>>>(gp);
    }
}
template <int TD, bool H, bool C>
void launch_tile_kind(int kind, const GatherParams& gp, int grid, size_t smem, cudaStream_t st) {
    switch (kind) {
        case CTR_MODEL_LR: launch_tile_t<TD, H, false, true, false, C>(gp, grid, smem, st); break;
        case CTR_MODEL_WIDE_DEEP: launch_tile_t<TD, H, true, true, false, C>(gp, grid, smem, st); break;
        case CTR_MODEL_DEEPFM: launch_tile_t<TD, H, true, true, true, C>(gp, grid, smem, st); break;
        default: launch_tile_t<TD, H, true, false, false, C>(gp, grid, smem, st); break;
    }
}

// The gather of n candidates on `st`: 32-candidate tiles, min(tiles, 4 x SMs) persistent CTAs.
void launch_gather(const ModelDesc& m, bool compact, const GatherParams& gp, int n, int num_sms, cudaStream_t st) {
    static const int ctas_per_sm = [] { const char* c = getenv("CTR_GATHER_CTAS"); return c ? atoi(c) : 4; }();
    const bool f32 = m.table_dtype == CTR_TABLE_FP32;
    const int n_tiles = (n + kGatherTile - 1) / kGatherTile;
    const int grid = std::max(1, std::min(n_tiles, num_sms * ctas_per_sm));
    if (compact) {
        const size_t smem = (size_t)kGatherTile * (2 * ((size_t)gp.Fk * gp.id_bytes + (size_t)gp.Wk * 4) + (size_t)m.F * 8);
        if (f32) launch_tile_kind<0, false, true>(m.kind, gp, grid, smem, st);
        else launch_tile_kind<1, false, true>(m.kind, gp, grid, smem, st);
    } else {
        const size_t smem = (size_t)2 * kGatherTile * m.F * 12;
        const bool hashed = gp.hashed != 0;
        if (f32) { if (hashed) launch_tile_kind<0, true, false>(m.kind, gp, grid, smem, st); else launch_tile_kind<0, false, false>(m.kind, gp, grid, smem, st); }
        else { if (hashed) launch_tile_kind<1, true, false>(m.kind, gp, grid, smem, st); else launch_tile_kind<1, false, false>(m.kind, gp, grid, smem, st); }
    }
}

// One chain on `st`: persistent CTA pairs, at most one pair per TPC and per scratch slice.
int32_t launch_chain(const Chain& c, const Slot& s, int n, float* d_scores, int num_sms, cudaStream_t st, long long* prof = nullptr, bool busy = true) {
    ChainParams p = c.params;
    p.M = n;
    p.tiles_in_flight = tiles_in_flight();
    p.prof = prof;
    if (!p.extra_out) p.scores = d_scores;
    const int tiles = (n + 2 * kBM - 1) / (2 * kBM);
    // A pair that holds two tiles alternates between them layer by layer (no idle tensor pipe at a layer boundary), so
    // when other requests are in flight on this GPU a small launch takes half as many pairs with two tiles each rather
    // than one tile on twice the SMs: the other half of the machine is free for the next request's tower (a request of
    // 10 000 candidates is 40 tiles; measured: packed + shared-context requests 268 -> 297 M candidates/s).  A request
    // that has the GPU to itself spreads out, one tile per pair: 0.185 instead of 0.206 ms for a lone 10 000-candidate
    // request.  CTR_TOWER_SPREAD=1 / =0 force one or the other.
    static const int spread_env = [] { const char* c = getenv("CTR_TOWER_SPREAD"); return c ? (atoi(c) ? 1 : 0) : -1; }();
    const bool spread = spread_env >= 0 ? spread_env == 1 : !busy;
    const int want = (tiles_in_flight() == 2 && !spread) ? (tiles + 1) / 2 : tiles;
    const int pairs = std::max(1, std::min(std::min(std::min(want, num_sms / 2), s.scratch_rows / (2 * kBM)), tower_pairs_cap()));
    tower_chain_kernel<256><<<2 * pairs, kGemmThreads, ChainCfg<256>::kSmemBytes, st>>>(c.maps, p);
    CUDA_TRY(cudaGetLastError());
    return CTR_OK;
}

// Bring-up: CTR_TIMELINE=1 records an event in front of and behind every kernel of a ctr_score_device call (all
// streams) and prints when each ran, relative to the first one -- which kernels really overlap in time.
struct TimelineMark { cudaEvent_t ev; const char* what; int slot; };
static std::vector<TimelineMark> g_timeline;
static const bool g_timeline_on = getenv("CTR_TIMELINE") != nullptr;
void timeline_mark(const char* what, int slot, cudaStream_t st) {
    if (!g_timeline_on || g_timeline.size() >= 256) return;
    cudaEvent_t ev;
    if (cudaEventCreate(&ev) != cudaSuccess) return;
    cudaEventRecord(ev, st);
    g_timeline.push_back({ev, what, slot});
}
void timeline_dump() {
    if (!g_timeline_on || g_timeline.empty()) return;
    static int dumps = 0;
    const bool print = ++dumps == 8;            // one call, after warm-up
    for (size_t i = 0; i < g_timeline.size(); ++i) {
        float ms = 0.f;
        if (print && cudaEventElapsedTime(&ms, g_timeline[0].ev, g_timeline[i].ev) == cudaSuccess)
            fprintf(stderr, "[timeline] stream %d %-14s %9.1f us\n", g_timeline[i].slot, g_timeline[i].what, ms * 1e3);
    }
    for (auto& m : g_timeline) cudaEventDestroy(m.ev);
    g_timeline.clear();
}

// Records a CUDA event on the launching stream in front of a stage (stage -2 = end of chunk).
void mark(ctr_engine* e, Slot& s, int stage, cudaStream_t on = nullptr) {
    if (!e->timing || !s.time_this || !s.tev || s.sets_recorded >= Slot::kTimeSets || s.cur_mark >= Slot::kTimeMarks) return;
    s.mark_stage[s.cur_mark] = stage;
    cudaEventRecord(s.tev[s.sets_recorded][s.cur_mark++], on ? on : s.stream);
    if (stage == -2) { s.n_marks = s.cur_mark; s.cur_mark = 0; ++s.sets_recorded; s.time_this = false; }
}

// Scores rows [0, n) of the device-side view `in` into d_scores on s.stream; n <= s.x_rows for tower models.
int32_t score_chunk(ctr_engine* e, DeviceCtx& dc, Slot& s, const InputView& in, int n, float* d_scores) {
    const ModelDesc& m = e->m;
    DeviceModel& dm = dc.model;
    // with an SM split the gather goes to the gather's SMs and the chains to the tower's; `s.stream` joins at the end
    const bool split = dc.split.on && s.g_stream && m.kind != CTR_MODEL_LR;
    cudaStream_t st = split ? s.g_stream : s.stream;
    const int gather_sms = split ? dc.split.gather_sms : dm.num_sms, tower_sms = split ? dc.split.tower_sms : dm.num_sms;
    if (split) {
        CUDA_TRY(cudaEventRecord(s.ev_in, s.stream));
        CUDA_TRY(cudaStreamWaitEvent(s.g_stream, s.ev_in, 0));
    }
    const bool sharded = e->cfg.shard_mode == CTR_TABLES_ROW_SHARDED && e->cfg.n_devices > 1 && !in.table_override;
    const bool compact = !in.reference_layout(m.F);

    GatherParams gp;
    memset(&gp, 0, sizeof(gp));
    gp.ids = in.ids; gp.wts = in.wts; gp.n = n; gp.F = m.F;
    gp.Fk = in.Fk; gp.Wk = in.Wk; gp.id_bytes = in.id_bytes; gp.n_ctx = in.n_ctx; gp.weighted_mask = in.weighted_mask;
    gp.hashed = m.id_mode == CTR_ID_HASHED && !in.ids_are_rows;
    gp.id_mask = in.id_mask;
    {   // context ids and fixed ids always arrive as ids: bring them to what the per-candidate list holds, indexed by field
        const bool to_row = in.ids_are_rows && m.id_mode == CTR_ID_HASHED;
        auto conv = [&](int64_t id) { return to_row ? (int64_t)(mix64((uint64_t)id) & ((uint64_t)m.V - 1)) : id; };
        for (int f = 0; f < in.n_ctx; ++f) { gp.ctx_ids[f] = conv(in.ctx_ids[f]); gp.ctx_wts[f] = in.ctx_wts[f]; }
        int k = in.n_ctx;
        if (compact)
            for (int f = in.n_ctx; f < m.F; ++f)
                if (!((in.id_mask >> f) & 1ull)) gp.ctx_ids[f] = conv(in.ctx_ids[k++]);
    }
    gp.bulk_ok = ((reinterpret_cast<uintptr_t>(in.ids) | reinterpret_cast<uintptr_t>(in.wts)) & 15) == 0;
    gp.table = in.table_override ? in.table_override : dm.table;
    gp.num_rows = in.table_override ? in.rows_override : (uint64_t)m.V;
    gp.b0 = dm.b0;
    gp.error_flag = s.d_err;
    gp.n_shards = sharded ? e->cfg.n_devices : 1;
    for (int g = 0; g < CTR_MAX_DEVICES; ++g) gp.shard_base[g] = e->shard_base[g];
    if (m.kind == CTR_MODEL_LR) { gp.scores = d_scores; }
    else { gp.X = s.buf[0]; gp.ldx = s.ld[0]; gp.wide = m.kind == CTR_MODEL_DCN_V2 ? nullptr : s.d_wide; }
    mark(e, s, 0, st);
    const int tl_slot = (int)(reinterpret_cast<uintptr_t>(s.stream) >> 4 & 0xff);
    timeline_mark("gather begin", tl_slot, st);
    launch_gather(m, compact, gp, n, gather_sms, st);
    timeline_mark("gather end", tl_slot, st);
    CUDA_TRY(cudaGetLastError());
    e->launches.fetch_add(1, std::memory_order_relaxed);
    if (split) {
        mark(e, s, -3, st);                                    // end of the gather on its own stream (stage -3: not a stage)
        CUDA_TRY(cudaEventRecord(s.ev_g, s.g_stream));
        CUDA_TRY(cudaStreamWaitEvent(s.t_stream, s.ev_g, 0));
        st = s.t_stream;
        timeline_mark("tower begin", tl_slot, st);
    }
    int li = 0;
    static const bool chain_prof_env = getenv("CTR_CHAIN_PROF") != nullptr;  // bring-up: in-kernel cycle counters to stderr
    const bool chain_prof = chain_prof_env && s.owns_buffers;                // it allocates: never on the no-allocation path
    for (const Chain& c : s.chains) {
        mark(e, s, ++li, st);
        long long* d_prof = nullptr;
        if (chain_prof) {
            CUDA_TRY(cudaMalloc(&d_prof, (size_t)dm.num_sms * 8 * sizeof(long long)));
            CUDA_TRY(cudaMemsetAsync(d_prof, 0, (size_t)dm.num_sms * 8 * sizeof(long long), st));
        }
        int32_t rc = launch_chain(c, s, n, d_scores, tower_sms, st, d_prof, dc.active.load(std::memory_order_relaxed) > 1);
        if (rc != CTR_OK) { if (d_prof) cudaFree(d_prof); return rc; }
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
    timeline_mark("tower end", tl_slot, st);
    mark(e, s, -2, st);
    if (split) {
        CUDA_TRY(cudaEventRecord(s.ev_t, s.t_stream));
        CUDA_TRY(cudaStreamWaitEvent(s.stream, s.ev_t, 0));
    }
    return CTR_OK;
}

// Scores n candidates (device-side view) on slot s, chunked so activations stay in their buffers.
int32_t score_on_slot(ctr_engine* e, DeviceCtx& dc, Slot& s, const InputView& in, int64_t n, float* d_scores) {
    const int64_t step = e->m.kind == CTR_MODEL_LR ? (int64_t)1 << 30 : s.x_rows;
    for (int64_t c0 = 0; c0 < n; c0 += step) {
        const int cnt = (int)std::min<int64_t>(step, n - c0);
        s.time_this = false;                   // stage timing is for ctr_score_device calls
        int32_t rc = score_chunk(e, dc, s, in.at(c0), cnt, d_scores + c0);
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

// ------------------------------------------------------------------ A6: the sort of processARequest (DC:195)
// Sorts d_scores[0, n) on `st` (ascending, Float.compareTo order, ties by candidate index): the sorted scores
// land in b.sorted, the candidate index of each in *d_order (one of b.idx).
int32_t launch_sort(ctr_engine* e, const DeviceModel& dm, SortBufs& b, const float* d_scores, int n, cudaStream_t st, const int32_t** d_order) {
    if (n <= kBlockSortMax) {                                   // one block, one launch
        sort_block_kernel<<<1, kBlockSortThreads, 0, st>>>(d_scores, n, b.keys[0], b.idx[0], b.keys[1], b.idx[1], b.sorted);
        CUDA_TRY(cudaGetLastError());
        e->launches.fetch_add(1, std::memory_order_relaxed);
        *d_order = b.idx[0];
        return CTR_OK;
    }
    const int n_blocks = (n + kSortTile - 1) / kSortTile;
    const int g1 = std::max(1, std::min((n + 255) / 256, dm.num_sms * 8));
    sort_make_keys_kernel<<<g1, 256, 0, st>>>(d_scores, n, b.keys[0], b.idx[0]);
    int cur = 0;
    for (int pass = 0; pass < 4; ++pass) {
        sort_hist_kernel<<<n_blocks, kSortThreads, 0, st>>>(b.keys[cur], n, pass * 8, b.counts);
        sort_scan_kernel<<<1, 1024, 0, st>>>(b.counts, 256 * n_blocks);
        sort_scatter_kernel<<<n_blocks, kSortThreads, 0, st>>>(b.keys[cur], b.idx[cur], n, pass * 8, b.counts, b.keys[cur ^ 1], b.idx[cur ^ 1]);
        cur ^= 1;
    }
    sort_gather_scores_kernel<<<g1, 256, 0, st>>>(d_scores, b.idx[cur], n, b.sorted);
    CUDA_TRY(cudaGetLastError());
    e->launches.fetch_add(14, std::memory_order_relaxed);
    *d_order = b.idx[cur];
    return CTR_OK;
}

}  // namespace
