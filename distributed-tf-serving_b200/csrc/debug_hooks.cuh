// debug_hooks.cuh -- test hooks of the C ABI (ctr_debug_*): table rows, one GEMM layer alone, the gather alone.
#pragma once
extern "C" {
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
    int32_t rc = score_chunk(e, dc, s, reference_view(s.d_ids, s.d_wts, m.F), (int)n, s.d_scores);
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
