#pragma once
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
    // Host threads that call in at the same time must not interleave their enqueues (one call's clearing
    // memset would land between another call's gather and tower): the whole enqueue -- and, on the engine's
    // own stream, the wait and the flag read as well -- runs under the device's mutex, like ctr_score_device.
    std::lock_guard<std::mutex> lk(dc.dev_mu);
    CUDA_TRY(cudaSetDevice(dm.dev));
    cudaStream_t user = reinterpret_cast<cudaStream_t>(stream);
    cudaStream_t st = user ? user : dc.main_stream;

    // A slot whose buffers are slices of the fixed array; nothing in it is allocated or freed.
    Slot s;
    s.dev = dm.dev; s.stream = st; s.owns_stream = false;
    s.cap = kStaticMaxCandidates; s.x_rows = kStaticMaxCandidates; s.scratch_rows = kStaticMaxCandidates / 2;   // 2 CTA pairs x 2 tiles each
    s.scratch_tiles = kStaticMaxCandidates / kBM;
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

    s.owns_buffers = false;                   // bring-up profiling (which allocates) is refused for this slot
    rc = score_chunk(e, dc, s, reference_view(d_ids, static_cast<const float*>(d_wts), m.F), (int)n_cand, static_cast<float*>(d_scores));
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
