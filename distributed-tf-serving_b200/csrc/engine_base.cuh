// engine_base.cuh -- error reporting, small helpers and the tensor-map encoders of the host runtime
// (included by engine.cu only: one translation unit, everything here has internal linkage).
#pragma once
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

// 256-row tiles a tower CTA pair keeps in flight (tower.cuh): 2 hides the layer-boundary bubble, 1 halves the
// activation scratch that cycles through L2.  CTR_TOWER_TILES=1|2.
inline int tiles_in_flight() {
    static const int v = [] { const char* c = getenv("CTR_TOWER_TILES"); return c && atoi(c) == 1 ? 1 : 2; }();
    return v;
}

// CTA pairs a tower launch may use (default: one per TPC).  The tower runs at the power cap, so fewer pairs at a higher
// clock cost little, and the pair scratch (one slice per pair) shrinks with them.  CTR_TOWER_PAIRS=<n>.
inline int tower_pairs_cap() {
    static const int v = [] { const char* c = getenv("CTR_TOWER_PAIRS"); const int n = c ? atoi(c) : 0; return n >= 1 ? n : 1 << 20; }();
    return v;
}

// A thread that is about to block on a condition variable first polls for this long: waking a sleeping
// thread costs 10-20 us on each side of a request whose whole GPU time is ~60 us.  CTR_SPIN_US=0 turns
// the polling off (hosts with fewer cores than engine threads).
inline int spin_us() {
    static const int v = getenv("CTR_SPIN_US") ? atoi(getenv("CTR_SPIN_US")) : 30;
    return v;
}
template <typename Pred>
inline bool poll_briefly(Pred ready) {
    const int us = spin_us();
    if (us <= 0) return ready();
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        for (int i = 0; i < 64; ++i) {
            if (ready()) return true;
            __builtin_ia32_pause();
        }
        if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(us)) return ready();
    }
}

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

}  // namespace
