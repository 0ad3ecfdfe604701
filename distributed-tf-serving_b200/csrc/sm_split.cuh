// sm_split.cuh -- the gather and the tower on disjoint sets of SMs (green contexts), so that one chunk's gather
// runs while another chunk's tower does without the two sharing an SM.
//
// Why: the two kernels are bound by different things (the gather by HBM's random-row rate, the tower by the tensor
// pipes), but side by side on the SAME SM they only trade time (device_path.cuh), and with the driver's default
// shared-memory carve-out a gather CTA cannot become resident beside a tower CTA at all, so two engine streams overlap
// little more than the tails.  HBM's random-row rate needs few warps (tools/probes/random_rows_probe.cu: 16 warps on
// every SM reach it), and the tower loses less than proportionally when it gets fewer SMs (it runs power-capped: 66 of
// 74 pairs cost 5 %).  A green context is CUDA's way to give a stream a fixed set of SMs.
//
// Measured (profiles/r02_split_sweep.txt): best split 48 / 100 SMs, 305 M candidates/s against 321-326 M
// for one kernel after the other -- the gather needs ~60 SMs for its full rate and the tower pays for what it gives up -- so
// this stays off by default.
//
// CTR_SM_SPLIT=<SMs for the gather> switches it on (a multiple of 8; the tower gets the rest).  Driver entry points
// come from cudaGetDriverEntryPoint: the library does not link libcuda.
#pragma once
namespace {

struct GreenApi {
    CUresult (*getDevResource)(CUdevice, CUdevResource*, CUdevResourceType) = nullptr;
    CUresult (*splitByCount)(CUdevResource*, unsigned int*, const CUdevResource*, CUdevResource*, unsigned int, unsigned int) = nullptr;
    CUresult (*generateDesc)(CUdevResourceDesc*, CUdevResource*, unsigned int) = nullptr;
    CUresult (*ctxCreate)(CUgreenCtx*, CUdevResourceDesc, CUdevice, unsigned int) = nullptr;
    CUresult (*ctxDestroy)(CUgreenCtx) = nullptr;
    CUresult (*streamCreate)(CUstream*, CUgreenCtx, unsigned int, int) = nullptr;
    CUresult (*deviceGet)(CUdevice*, int) = nullptr;
    bool ok = false;
};

const GreenApi& green_api() {
    static const GreenApi api = [] {
        GreenApi a;
        auto get = [](const char* name) -> void* {
            void* p = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { (void)cudaGetLastError(); return nullptr; }
            return p;
        };
        a.getDevResource = reinterpret_cast<decltype(a.getDevResource)>(get("cuDeviceGetDevResource"));
        a.splitByCount = reinterpret_cast<decltype(a.splitByCount)>(get("cuDevSmResourceSplitByCount"));
        a.generateDesc = reinterpret_cast<decltype(a.generateDesc)>(get("cuDevResourceGenerateDesc"));
        a.ctxCreate = reinterpret_cast<decltype(a.ctxCreate)>(get("cuGreenCtxCreate"));
        a.ctxDestroy = reinterpret_cast<decltype(a.ctxDestroy)>(get("cuGreenCtxDestroy"));
        a.streamCreate = reinterpret_cast<decltype(a.streamCreate)>(get("cuGreenCtxStreamCreate"));
        a.deviceGet = reinterpret_cast<decltype(a.deviceGet)>(get("cuDeviceGet"));
        a.ok = a.getDevResource && a.splitByCount && a.generateDesc && a.ctxCreate && a.ctxDestroy && a.streamCreate && a.deviceGet;
        return a;
    }();
    return api;
}

// Splits device `dev` into `gather_sms` SMs and the rest.  Leaves split.on false (and says why) when the driver cannot.
int32_t setup_sm_split(SmSplit& split, int dev, int gather_sms) {
    const GreenApi& g = green_api();
    if (!g.ok) return fail(CTR_ERR_CUDA, "CTR_SM_SPLIT: this driver has no green-context entry points");
    CUDA_TRY(cudaSetDevice(dev));
    CUDA_TRY(cudaFree(nullptr));                               // the primary context exists from here on
    CUdevice cudev;
    if (g.deviceGet(&cudev, dev) != CUDA_SUCCESS) return fail(CTR_ERR_CUDA, "cuDeviceGet failed");
    CUdevResource all, part, rest;
    CUresult r = g.getDevResource(cudev, &all, CU_DEV_RESOURCE_TYPE_SM);
    if (r != CUDA_SUCCESS) return fail(CTR_ERR_CUDA, "cuDeviceGetDevResource failed with CUresult " + std::to_string((int)r));
    unsigned int groups = 1;
    r = g.splitByCount(&part, &groups, &all, &rest, 0, (unsigned int)gather_sms);
    if (r != CUDA_SUCCESS || groups != 1) return fail(CTR_ERR_CUDA, "cuDevSmResourceSplitByCount failed with CUresult " + std::to_string((int)r));
    CUdevResourceDesc d_part, d_rest;
    if ((r = g.generateDesc(&d_part, &part, 1)) != CUDA_SUCCESS || (r = g.generateDesc(&d_rest, &rest, 1)) != CUDA_SUCCESS)
        return fail(CTR_ERR_CUDA, "cuDevResourceGenerateDesc failed with CUresult " + std::to_string((int)r));
    if ((r = g.ctxCreate(&split.gather_ctx, d_part, cudev, CU_GREEN_CTX_DEFAULT_STREAM)) != CUDA_SUCCESS)
        return fail(CTR_ERR_CUDA, "cuGreenCtxCreate (gather) failed with CUresult " + std::to_string((int)r));
    if ((r = g.ctxCreate(&split.tower_ctx, d_rest, cudev, CU_GREEN_CTX_DEFAULT_STREAM)) != CUDA_SUCCESS) {
        g.ctxDestroy(split.gather_ctx);
        split.gather_ctx = nullptr;
        return fail(CTR_ERR_CUDA, "cuGreenCtxCreate (tower) failed with CUresult " + std::to_string((int)r));
    }
    split.gather_sms = (int)part.sm.smCount;
    split.tower_sms = (int)rest.sm.smCount;
    split.on = true;
    return CTR_OK;
}

int32_t split_stream(const SmSplit& split, bool gather, cudaStream_t* out) {
    CUstream st = nullptr;
    CUresult r = green_api().streamCreate(&st, gather ? split.gather_ctx : split.tower_ctx, CU_STREAM_NON_BLOCKING, 0);
    if (r != CUDA_SUCCESS) return fail(CTR_ERR_CUDA, "cuGreenCtxStreamCreate failed with CUresult " + std::to_string((int)r));
    *out = reinterpret_cast<cudaStream_t>(st);
    return CTR_OK;
}

void free_sm_split(SmSplit& split) {
    if (split.gather_ctx) green_api().ctxDestroy(split.gather_ctx);
    if (split.tower_ctx) green_api().ctxDestroy(split.tower_ctx);
    split = SmSplit();
}

}  // namespace
