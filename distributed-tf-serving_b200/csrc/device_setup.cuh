// device_setup.cuh -- model upload (tables, tower weights), GPU-local host placement, slot buffers and the
// chains of tower layers built over them.
#pragma once
namespace {
// ------------------------------------------------------------------ device set-up
template <typename Kern>
int32_t set_smem_attr(Kern k, int bytes) {
    CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    // the whole 228 KB as shared memory, so that a gather CTA (23 KB) fits beside this kernel's 194 KB (device_path.cuh)
    CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
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
// buffers of the device sort (launch_sort, device_path.cuh)
void free_sort_bufs(SortBufs& b) {
    cudaFree(b.sorted); cudaFree(b.counts);
    for (int i = 0; i < 2; ++i) { cudaFree(b.keys[i]); cudaFree(b.idx[i]); }
    b = SortBufs();
}

int32_t alloc_sort_bufs(SortBufs& b, int64_t cap) {
    if (b.cap >= cap) return CTR_OK;
    free_sort_bufs(b);
    CUDA_TRY(cudaMalloc(&b.sorted, cap * sizeof(float)));
    for (int i = 0; i < 2; ++i) {
        CUDA_TRY(cudaMalloc(&b.keys[i], cap * sizeof(uint32_t)));
        CUDA_TRY(cudaMalloc(&b.idx[i], cap * sizeof(int32_t)));
    }
    CUDA_TRY(cudaMalloc(&b.counts, (size_t)256 * ((cap + kSortTile - 1) / kSortTile) * sizeof(uint32_t)));
    b.cap = cap;
    return CTR_OK;
}

// Sizes and allocates the device's pair scratch for the loaded model.  Buffer 1 takes the outputs of the even MLP
// layers, buffer 2 of the odd ones (the last layer has no output: its epilogue is the head), each as wide as the
// widest layer it holds -- 1024 + 512 columns for 624->1024->512->256, not 2 x 1024; buffers 3-5 are DCN-v2's
// x_l ping-pong and the tile-major copy of x0.
int32_t alloc_pair_scratch(ctr_engine* e, DeviceCtx& dc) {
    const ModelDesc& m = e->m;
    DeviceModel& dm = dc.model;
    PairScratch& ps = dc.scratch;
    if (m.kind == CTR_MODEL_LR) return CTR_OK;
    CUDA_TRY(cudaSetDevice(dm.dev));
    ps.n_slices = std::min(dm.num_sms / 2, tower_pairs_cap());
    const size_t rows = (size_t)ps.n_slices * tiles_in_flight() * 2 * kBM;     // one or two 256-row tiles in flight per slice
    for (int l = 0; l + 1 < (int)dm.mlp.size(); ++l) ps.ld[1 + (l & 1)] = std::max(ps.ld[1 + (l & 1)], dm.mlp[l].Npad);
    if (m.kind == CTR_MODEL_DCN_V2) {
        const int xl_ld = dm.cross.empty() ? m.d_pad() : dm.cross.back().Npad;      // cross outputs may be padded past d_pad
        ps.ld[3] = ps.ld[4] = ps.ld[5] = xl_ld;
    }
    size_t off[6] = {0, 0, 0, 0, 0, 0}, total = 256;
    for (int i = 1; i < 6; ++i) { off[i] = total; total += rows * ps.ld[i] * sizeof(__half); }
    CUDA_TRY(cudaMalloc(&ps.base, total));
    CUDA_TRY(cudaMemset(ps.base, 0, total));                          // bitmap clear, pad columns zero
    ps.bytes = total;
    ps.bitmap = reinterpret_cast<unsigned long long*>(ps.base);
    for (int i = 1; i < 6; ++i) ps.buf[i] = ps.ld[i] ? reinterpret_cast<__half*>(ps.base + off[i]) : nullptr;
    // Keep the scratch in L2: it is written and re-read within microseconds by the same SM pair, but anything
    // streaming through (X, table rows) would push its dirty lines out to HBM and back.  The device sets aside
    // what it can for persisting lines; each stream's access window (set in init_slot_common) covers the scratch.
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, dm.dev));
    // measured (profiles/r02_*): with the window on, W&D device-resident throughput FELL from 324 to 303 M candidates/s --
    // the part of the scratch that does not fit the set-aside is treated as streaming and evicted first.  Off unless asked for.
    static const bool persist = getenv("CTR_L2_PERSIST") != nullptr;
    if (persist && prop.persistingL2CacheMaxSize > 0) {
        const size_t want = std::min<size_t>((size_t)prop.persistingL2CacheMaxSize, total);
        if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) != cudaSuccess) (void)cudaGetLastError();
        else dc.l2_persist_bytes = want;
        dc.l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
    }
    return CTR_OK;
}

void free_pair_scratch(DeviceCtx& dc) {
    if (dc.scratch.base) cudaFree(dc.scratch.base);
    dc.scratch = PairScratch();
}

// Points a stream's L2 access-policy window at the device's pair scratch (persisting hits; misses stream).
void set_scratch_window(DeviceCtx& dc, cudaStream_t st) {
    const PairScratch& ps = dc.scratch;
    if (!ps.base || !dc.l2_persist_bytes || !dc.l2_window_max) return;
    cudaStreamAttrValue v;
    memset(&v, 0, sizeof(v));
    v.accessPolicyWindow.base_ptr = ps.base;
    v.accessPolicyWindow.num_bytes = std::min(ps.bytes, dc.l2_window_max);
    v.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)dc.l2_persist_bytes / (double)v.accessPolicyWindow.num_bytes);
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) (void)cudaGetLastError();
}

void free_slot(Slot& s) {
    cudaSetDevice(s.dev);
    cudaFree(s.d_ids); cudaFree(s.d_wts); cudaFree(s.d_scores); cudaFree(s.d_wide); cudaFree(s.d_extra); cudaFree(s.d_err);
    cudaFree(s.buf[0]);                        // X; buffers 1-5 are slices of the device's pair scratch
    for (auto*& b : s.buf) b = nullptr;
    cudaFreeHost(s.h_ids); cudaFreeHost(s.h_wts); cudaFreeHost(s.h_scores); cudaFreeHost(s.h_err); cudaFreeHost(s.h_order);
    free_sort_bufs(s.srt);
    if (s.tev) {
        for (int i = 0; i < Slot::kTimeSets; ++i) for (int j = 0; j < Slot::kTimeMarks; ++j) cudaEventDestroy(s.tev[i][j]);
        delete[] s.tev;
    }
    if (s.join_ev) cudaEventDestroy(s.join_ev);
    if (s.g_stream) cudaStreamDestroy(s.g_stream);
    if (s.t_stream) cudaStreamDestroy(s.t_stream);
    for (cudaEvent_t ev : {s.ev_in, s.ev_g, s.ev_t}) if (ev) cudaEventDestroy(ev);
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
            else rc = make_map_tile_major(&c.maps.a[l], s.buf[a], (uint64_t)s.ld[a], (uint64_t)s.scratch_tiles);
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
        c.params.slice_bitmap = s.slice_bitmap; c.params.n_slices = s.n_slices;
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
    // at 1 / 2 / 4 / 8 / 16 tiles per pair in round 1), CTR_CHUNK_ROWS overrides
    const int want = host_staging ? (int)s.cap : (e->chunk_rows > 0 ? e->chunk_rows : 16 * pairs * 2 * kBM);
    s.x_rows = round_up(std::max(want, 1), 2 * kBM);
    CUDA_TRY(cudaMalloc(&s.d_ids, (size_t)s.cap * m.F * sizeof(int64_t)));
    CUDA_TRY(cudaMalloc(&s.d_wts, (size_t)s.cap * m.F * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_scores, (size_t)s.cap * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_wide, (size_t)s.x_rows * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_extra, (size_t)s.x_rows * sizeof(float)));
    CUDA_TRY(cudaMalloc(&s.d_err, sizeof(int32_t)));
    CUDA_TRY(cudaMemset(s.d_err, 0, sizeof(int32_t)));
    if (host_staging) {
        CUDA_TRY(cudaMallocHost(&s.h_ids, (size_t)s.cap * m.F * sizeof(int64_t)));
        CUDA_TRY(cudaMallocHost(&s.h_wts, (size_t)s.cap * m.F * sizeof(float)));
        CUDA_TRY(cudaMallocHost(&s.h_scores, (size_t)s.cap * sizeof(float)));
        CUDA_TRY(cudaMallocHost(&s.h_err, sizeof(int32_t)));
        CUDA_TRY(cudaMallocHost(&s.h_order, (size_t)s.cap * sizeof(int32_t)));
        int32_t src = alloc_sort_bufs(s.srt, s.cap);           // 24 B per candidate: ctr_predict_sorted never allocates on the request path
        if (src != CTR_OK) return src;
    }
    if (m.kind == CTR_MODEL_LR) return CTR_OK;

    // X is the slot's own; the layer activations live in the device's pair scratch
    s.ld[0] = m.d_pad();
    CUDA_TRY(cudaMalloc(&s.buf[0], (size_t)s.x_rows * s.ld[0] * sizeof(__half)));
    CUDA_TRY(cudaMemset(s.buf[0], 0, (size_t)s.x_rows * s.ld[0] * sizeof(__half)));   // pad columns stay zero
    const PairScratch& ps = dc.scratch;
    for (int i = 1; i < 6; ++i) { s.buf[i] = ps.buf[i]; s.ld[i] = ps.ld[i]; }
    s.slice_bitmap = ps.bitmap; s.n_slices = ps.n_slices;
    s.scratch_rows = ps.n_slices * 2 * kBM;
    s.scratch_tiles = ps.n_slices * tiles_in_flight() * 2;      // 128-row tiles in each scratch buffer
    set_scratch_window(dc, s.stream);
    return build_chains(e, dm, s);
}

// The slot's two streams on the device's SM partitions (no-op without a split).
int32_t init_slot_split(const SmSplit& split, Slot& s) {
    if (!split.on) return CTR_OK;
    int32_t rc;
    if ((rc = split_stream(split, true, &s.g_stream)) != CTR_OK) return rc;
    if ((rc = split_stream(split, false, &s.t_stream)) != CTR_OK) return rc;
    CUDA_TRY(cudaEventCreateWithFlags(&s.ev_in, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&s.ev_g, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&s.ev_t, cudaEventDisableTiming));
    return CTR_OK;
}

int32_t init_slot_common(Slot& s, int dev, cudaStream_t stream /*nullptr -> create*/) {
    CUDA_TRY(cudaSetDevice(dev));
    s.dev = dev;
    if (stream) { s.stream = stream; s.owns_stream = false; }
    else CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&s.join_ev, cudaEventDisableTiming));
    return CTR_OK;
}

}  // namespace
