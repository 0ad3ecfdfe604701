// gather.cuh -- hashed sparse-feature embedding gather / pool kernel (north star (b)).
//
// Replaces the part of the served model that TF-Serving ran behind sendRequest
// (DCNClient.java:91-115): look up the row of every (candidate, field) id, scale it by the
// field weight, and pool.  Builder-defined math, SURVEY.md section 8(a-M):
//     x[b,f,:] = E[row] * wt          -> X[b, f*16 .. f*16+15]  (half, the GEMM's A operand)
//     lin[b]   = b0 + sum_f W1[row]*wt           (+ 0.5*sum_k[(sum_f x)^2 - sum_f x^2] for DeepFM)
//
// Data layout in HBM ("packed rows"): one table row holds the 16 embedding values AND the
// first-order weight, so a lookup is ONE place in DRAM and ONE load instruction:
//     fp32 row (128 B): [e0..e15 fp32 | w1 fp32 | pad]      f16 row (64 B): [e0..e15 half | w1 fp32 | pad]
// B200's DRAM moves 64-byte granules: an fp32 row (68 useful bytes) costs two of them, an f16 row one.
//
// Mapping (round 2).  A warp owns "warp tiles" of 4 consecutive candidates and never synchronises
// with the other warps of its CTA.  A row is fetched by ONE 128-bit load instruction spread over the
// lanes of a lane group (8 lanes per fp32 row: 4 embedding quarters + the lane that takes w1; 4 lanes
// per f16 row), so a load instruction touches one 128-byte line per row -- the first version issued a
// second instruction for w1 and paid the L1 line-lookup cost of every row twice, which is what bounded
// it (time did not change with the table size or dtype).  The loads of candidate c+1 are issued slot by
// slot while candidate c's rows are consumed from the same registers (a rolling window of kIters loads
// per lane, always full), so the warp never drains between candidates or between tiles.  The id / weight
// lists of a warp tile are staged into the warp's own shared-memory buffers with cp.async.bulk (SASS
// UBLKCP) on the warp's own mbarriers, double buffered.  Pooling (first-order sum, FM sums) is a fixed
// per-lane order plus a fixed warp-shuffle tree: the same candidate gets bit-identical results at any
// position, in any chunk, on any GPU, from any input format.
//
// Input formats.  (1) the reference's: ids int64 [n, F], wts float [n, F] (DC:97-108).  (2) compact
// (kCompact): the first n_ctx fields of every candidate come from one shared list (kernel parameters),
// the per-candidate list holds the remaining Fk = F - n_ctx ids as int64 or as uint32 ROW indices
// (already hashed by the host packer), and weights only for the fields flagged in weighted_mask (the
// others are 1.0) -- what ctr_predict_shared_async / ctr_predict_packed_async ship over the host link.
//
// Footprint: 256 threads x 64 registers and ~30 KB of shared memory per CTA, so that one CTA fits on an SM
// NEXT TO the resident tower CTA (tower.cuh) and a chunk's gather runs under the previous chunk's tower.
//
// Roofline: HBM.  Algorithmic bytes per candidate (F=39: 13 dense + 26 sparse, fp32 table):
// 468 (ids+wts) + 26*(64+4) (rows; the 13 dense-field rows stay in L1/L2) + 4 = 2240 B.
#pragma once
#include "common.cuh"

namespace ctr {

constexpr int kWarpTile = 4;           // candidates per warp tile
constexpr int kGatherThreads = 256;    // 8 independent warps
constexpr int kGatherWarps = kGatherThreads / 32;

struct GatherParams {
    const void* ids;           // [n, Fk] int64 (id_bytes 8) or uint32 row indices (id_bytes 4)
    const float* wts;          // [n, Wk]
    int64_t n;
    int32_t F;                 // fields of the model
    int32_t Fk;                // id columns per candidate (F - n_ctx)
    int32_t Wk;                // weight columns per candidate (popcount of weighted_mask over the per-candidate fields)
    int32_t id_bytes;
    int32_t n_ctx;             // leading fields taken from ctx_ids / ctx_wts (compact input only)
    int32_t hashed;            // row = mix64(id) & (num_rows - 1); 0 for direct tables and for ids that already are rows
    int32_t bulk_ok;           // ids/wts base pointers are 16-byte aligned -> bulk copies allowed
    uint64_t weighted_mask;    // bit f: per-candidate field f has a weight column (compact input only)
    const uint8_t* table;      // packed rows
    uint64_t num_rows;         // direct mode: bound; hashed mode: power of two
    float b0;
    __half* X;                 // [n, ldx] or nullptr
    int32_t ldx;
    float* wide;               // [n] lin (+FM) or nullptr
    float* scores;             // [n] LR only: sigmoid(lin) written directly
    int32_t* error_flag;       // set to 1 when a direct id is out of range
    // row-sharded tables (rows striped over devices, row mod n_shards): peer base pointers
    int32_t n_shards;
    const uint8_t* shard_base[8];
    int64_t ctx_ids[48];       // shared context (compact input): ids (or rows) and weights of fields 0 .. n_ctx-1
    float ctx_wts[48];
};

template <int kTableDtype>
struct RowLayout;
template <>
struct RowLayout<0> { static constexpr int shift = 7; static constexpr int w1_off = 64; static constexpr int lanes = 8; static constexpr int emb_lanes = 4; static constexpr int nv = 4; };
template <>
struct RowLayout<1> { static constexpr int shift = 6; static constexpr int w1_off = 32; static constexpr int lanes = 4; static constexpr int emb_lanes = 2; static constexpr int nv = 8; };

__device__ __forceinline__ uint4 ldg_nc_u4(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// kTableDtype: 0 fp32 rows, 1 f16 rows.  kIters: load slots per lane and candidate (kIters * rows-per-load >= F).
// kCompact: compact input (shared context / uint32 rows / sparse weights).  kEmb: write X.  kLin: first-order
// sum.  kFM: DeepFM second-order term.
template <int kTableDtype, int kIters, bool kCompact, bool kEmb, bool kLin, bool kFM>
__global__ void __launch_bounds__(kGatherThreads, (kTableDtype == 0 && kIters > 10) ? 3 : 4)
gather_pool_kernel(const __grid_constant__ GatherParams p) {
    using RL = RowLayout<kTableDtype>;
    constexpr int LPR = RL::lanes;              // lanes per row
    constexpr int RPL = 32 / LPR;               // rows per load instruction
    constexpr int NV = RL::nv;                  // embedding values per embedding lane
    extern __shared__ __align__(128) uint8_t gsm[];
    __shared__ __align__(8) uint64_t full_bar[kGatherWarps][2];
    __shared__ int64_t s_ctx_id[kCompact ? 48 : 1];
    __shared__ float s_ctx_wt[kCompact ? 48 : 1];

    const int F = p.F, Fk = p.Fk, Wk = p.Wk;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane / LPR, q = lane % LPR;
    const uint32_t ids_bytes = (uint32_t)(kWarpTile * Fk * p.id_bytes), wts_bytes = (uint32_t)(kWarpTile * Wk * 4);   // multiples of 16
    const uint32_t buf_bytes = ids_bytes + wts_bytes;
    uint8_t* const wbase = gsm + (size_t)warp * 2 * buf_bytes;
    const int64_t n_tiles = (p.n + kWarpTile - 1) / kWarpTile;

    if (kCompact) {
        for (int f = tid; f < p.n_ctx; f += kGatherThreads) { s_ctx_id[f] = p.ctx_ids[f]; s_ctx_wt[f] = p.ctx_wts[f]; }
    }
    if (lane == 0) {
        mbar_init(&full_bar[warp][0], 1);
        mbar_init(&full_bar[warp][1], 1);
        mbar_fence_init();
    }
    __syncthreads();                        // the only block-wide barrier: from here on every warp is on its own

    // Stage the id / weight lists of one warp tile into this warp's buffer `b`.  Full, aligned tiles: one lane
    // issues two bulk copies that complete on the warp's mbarrier.  The ragged last tile (or unaligned caller
    // pointers) is copied with ordinary loads by the warp.
    const uint8_t* const gids = static_cast<const uint8_t*>(p.ids);
    auto tile_is_bulk = [&](int64_t tile) { return p.bulk_ok && (tile + 1) * kWarpTile <= p.n; };
    auto stage = [&](int64_t tile, int b) {
        const int64_t c0 = tile * kWarpTile;
        uint8_t* dst = wbase + b * buf_bytes;
        __syncwarp();                                          // every lane is done reading what `b` held before
        if (tile_is_bulk(tile)) {
            if (lane == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&full_bar[warp][b], buf_bytes);
                bulk_g2s(dst, gids + (size_t)c0 * Fk * p.id_bytes, ids_bytes, &full_bar[warp][b]);
                if (wts_bytes) bulk_g2s(dst + ids_bytes, p.wts + (size_t)c0 * Wk, wts_bytes, &full_bar[warp][b]);
            }
        } else {
            const int cnt = (int)min((int64_t)kWarpTile, p.n - c0);
            const uint32_t* si = reinterpret_cast<const uint32_t*>(gids + (size_t)c0 * Fk * p.id_bytes);
            uint32_t* di = reinterpret_cast<uint32_t*>(dst);
            for (int i = lane; i < cnt * Fk * (p.id_bytes >> 2); i += 32) di[i] = si[i];
            float* dw = reinterpret_cast<float*>(dst + ids_bytes);
            for (int i = lane; i < cnt * Wk; i += 32) dw[i] = p.wts[(size_t)c0 * Wk + i];
            __syncwarp();
        }
    };

    // id and weight of (candidate c of the tile in buffer `buf`, field f)
    auto fetch_id = [&](const uint8_t* buf, int c, int f) -> uint64_t {
        if (kCompact) {
            if (f < p.n_ctx) return (uint64_t)s_ctx_id[f];
            const int col = c * Fk + (f - p.n_ctx);
            return p.id_bytes == 4 ? (uint64_t)reinterpret_cast<const uint32_t*>(buf)[col] : (uint64_t)reinterpret_cast<const int64_t*>(buf)[col];
        }
        return (uint64_t)reinterpret_cast<const int64_t*>(buf)[c * F + f];
    };
    auto fetch_wt = [&](const uint8_t* buf, int c, int f) -> float {
        const float* swt = reinterpret_cast<const float*>(buf + ids_bytes);
        if (kCompact) {
            if (f < p.n_ctx) return s_ctx_wt[f];
            const uint64_t m = p.weighted_mask >> p.n_ctx;
            const int j = f - p.n_ctx;
            if (!((m >> j) & 1ull)) return 1.0f;
            return swt[c * Wk + __popcll(m & ((1ull << j) - 1ull))];
        }
        return swt[c * F + f];
    };

    // the rolling window of row loads: slot j holds the 16 bytes this lane fetched of row (field j*RPL + g)
    uint4 ev[kIters];
    const bool emb_lane = q < RL::emb_lanes, w1_lane = q == RL::emb_lanes;
    const bool load_lane = (emb_lane && (kEmb || kFM)) || (w1_lane && kLin);

    // issue the load of slot j for candidate c of the tile in `buf`; returns 1 when the id was out of range
    auto issue = [&](int j, const uint8_t* buf, int c) -> uint32_t {
        const int f = j * RPL + g;
        ev[j] = make_uint4(0u, 0u, 0u, 0u);
        uint32_t bad = 0;
        if (j * RPL < F && f < F) {
            const uint64_t id = fetch_id(buf, c, f);
            uint64_t row;
            if (p.hashed) {
                row = mix64(id) & (p.num_rows - 1);
            } else {
                row = id;
                if (row >= p.num_rows) { row = 0; bad = 1; *p.error_flag = 1; }
            }
            const uint8_t* base = p.table;
            if (p.n_shards > 1) {
                const uint32_t owner = (uint32_t)(row % (uint32_t)p.n_shards);
                row = row / (uint32_t)p.n_shards;
                base = p.shard_base[owner];
            }
            if (load_lane) ev[j] = ldg_nc_u4(base + (row << RL::shift) + q * 16);
        }
        return bad;
    };

    // per-candidate accumulators (of the candidate being consumed)
    float lin, s2, s1[NV];
    auto reset_acc = [&]() {
        lin = 0.f; s2 = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) s1[k] = 0.f;
    };
    // consume slot j of candidate (global index b, candidate c of the tile in `buf`); `bad`: its id was out of range
    auto consume = [&](int j, const uint8_t* buf, int c, int64_t b, uint32_t bad) {
        const int f = j * RPL + g;
        if (j * RPL < F) {
            float wt = (f < F && !bad) ? fetch_wt(buf, c, f) : 0.f;
            if (emb_lane && (kEmb || kFM)) {
                float x[NV];
                if (kTableDtype == 0) {
                    x[0] = __uint_as_float(ev[j].x) * wt; x[1] = __uint_as_float(ev[j].y) * wt;
                    x[2] = __uint_as_float(ev[j].z) * wt; x[3] = __uint_as_float(ev[j].w) * wt;
                } else {
                    const uint32_t w4[4] = {ev[j].x, ev[j].y, ev[j].z, ev[j].w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&w4[k]));
                        x[2 * k] = a.x * wt; x[2 * k + 1] = a.y * wt;
                    }
                }
                if (kFM) {
#pragma unroll
                    for (int k = 0; k < NV; ++k) { s1[k] += x[k]; s2 = fmaf(x[k], x[k], s2); }
                }
                if (kEmb && f < F) {
                    uint32_t o[NV / 2];
#pragma unroll
                    for (int k = 0; k < NV / 2; ++k) o[k] = pack_half2_sat(x[2 * k], x[2 * k + 1]);
                    __half* dst = p.X + b * p.ldx + f * kEmbDim + q * NV;
                    if (NV == 4) *reinterpret_cast<uint2*>(dst) = make_uint2(o[0], o[1]);
                    else *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[NV / 2 - 2], o[NV / 2 - 1]);
                }
            } else if (kLin && w1_lane) {
                lin = fmaf(__uint_as_float(ev[j].x), wt, lin);
            }
        }
    };
    // after the last slot of a candidate: pool over the warp and write lin (+FM) / the LR score
    auto finish = [&](int64_t b) {
        if (kLin || kFM) {
            float acc = lin;                                  // only the w1 lanes hold first-order terms
            if (kFM) {
                // field sums per embedding coordinate: reduce over the lane groups (same q, different g)
#pragma unroll
                for (int off = LPR; off <= 16; off <<= 1) {
#pragma unroll
                    for (int k = 0; k < NV; ++k) s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], off);
                }
                float sq = 0.f;
#pragma unroll
                for (int k = 0; k < NV; ++k) sq = fmaf(s1[k], s1[k], sq);
                acc += (g == 0 && emb_lane ? 0.5f * sq : 0.f) - 0.5f * s2;
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
            if (lane == 0) {
                const float z = acc + p.b0;
                if (p.wide) p.wide[b] = z;
                if (p.scores) p.scores[b] = sigmoidf_fast(z);
            }
        }
    };

    const int64_t tile_stride = (int64_t)gridDim.x * kGatherWarps;
    int64_t tile = (int64_t)blockIdx.x * kGatherWarps + warp;
    if (tile >= n_tiles) return;
    stage(tile, 0);

    uint32_t phase_bits = 0;                               // bit b = parity to wait for on full_bar[warp][b]
    bool have_prev = false;                                // a candidate's loads are in flight in ev[]
    const uint8_t* prev_buf = nullptr;
    int prev_c = 0;
    int64_t prev_b = 0;
    uint32_t prev_bad = 0;                                 // bit j: slot j of the candidate in flight had a bad id (per lane)
    for (int it = 0; tile < n_tiles; tile += tile_stride, ++it) {
        const int bsel = it & 1;
        const uint8_t* buf = wbase + bsel * buf_bytes;
        if (tile_is_bulk(tile)) {
            mbar_wait(&full_bar[warp][bsel], (phase_bits >> bsel) & 1u);
            phase_bits ^= 1u << bsel;
        }
        const int64_t c0 = tile * kWarpTile;
        const int n_here = (int)min((int64_t)kWarpTile, p.n - c0);
        for (int c = 0; c < n_here; ++c) {
            uint32_t bad = 0;
            if (have_prev) {
                reset_acc();
#pragma unroll
                for (int j = 0; j < kIters; ++j) {
                    consume(j, prev_buf, prev_c, prev_b, (prev_bad >> j) & 1u);
                    bad |= issue(j, buf, c) << j;
                }
                finish(prev_b);
            } else {
#pragma unroll
                for (int j = 0; j < kIters; ++j) bad |= issue(j, buf, c) << j;
            }
            have_prev = true; prev_buf = buf; prev_c = c; prev_b = c0 + c; prev_bad = bad;
            // the previous tile's buffer has been read for the last time: stage the next tile into it
            if (c == 0 && tile + tile_stride < n_tiles) stage(tile + tile_stride, bsel ^ 1);
        }
    }
    if (have_prev) {
        reset_acc();
#pragma unroll
        for (int j = 0; j < kIters; ++j) consume(j, prev_buf, prev_c, prev_b, (prev_bad >> j) & 1u);
        finish(prev_b);
    }
}

// ------------------------------------------------------------------ table construction
// Procedural table: row r, coordinate k = proc_unit(key_e, r*16+k) * scale_e (exactly the
// oracle's formula), first-order weight = proc_unit(key_w, r) * scale_w1.
template <int kTableDtype>
__global__ void fill_table_procedural_kernel(uint8_t* table, uint64_t local_rows, uint64_t row_stride, uint64_t row_offset,
                                             uint64_t key_e, uint64_t key_w, float scale_e, float scale_w1) {
    const uint64_t total = local_rows * 4;       // one thread per quarter row
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t lr = i >> 2;
        const int q = (int)(i & 3);
        const uint64_t row = lr * row_stride + row_offset;          // global row held at local slot lr
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = __fmul_rn(proc_unit(key_e, row * kEmbDim + q * 4 + j), scale_e);
        uint8_t* rp = table + (lr << RowLayout<kTableDtype>::shift);
        if (kTableDtype == 0) {
            *reinterpret_cast<float4*>(rp + q * 16) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            const __half2 a = __floats2half2_rn(v[0], v[1]), c = __floats2half2_rn(v[2], v[3]);
            uint2 o;
            o.x = *reinterpret_cast<const uint32_t*>(&a);
            o.y = *reinterpret_cast<const uint32_t*>(&c);
            *reinterpret_cast<uint2*>(rp + q * 8) = o;
        }
        if (q == 0) *reinterpret_cast<float*>(rp + RowLayout<kTableDtype>::w1_off) = __fmul_rn(proc_unit(key_w, row), scale_w1);
    }
}

// Raw table from a weight file: pack [rows,16] fp32 + [rows] fp32 into packed rows.
template <int kTableDtype>
__global__ void pack_table_kernel(uint8_t* table, uint64_t dst_row0, const float* emb, const float* w1, uint64_t rows) {
    const uint64_t total = rows * 4;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i >> 2;
        const int q = (int)(i & 3);
        const float4 v = *reinterpret_cast<const float4*>(emb + r * kEmbDim + q * 4);
        uint8_t* rp = table + ((dst_row0 + r) << RowLayout<kTableDtype>::shift);
        if (kTableDtype == 0) {
            *reinterpret_cast<float4*>(rp + q * 16) = v;
        } else {
            const __half2 a = __floats2half2_rn(v.x, v.y), c = __floats2half2_rn(v.z, v.w);
            uint2 o;
            o.x = *reinterpret_cast<const uint32_t*>(&a);
            o.y = *reinterpret_cast<const uint32_t*>(&c);
            *reinterpret_cast<uint2*>(rp + q * 8) = o;
        }
        if (q == 0) *reinterpret_cast<float*>(rp + RowLayout<kTableDtype>::w1_off) = w1[r];
    }
}

template <int kTableDtype>
__global__ void read_rows_kernel(const uint8_t* table, const int64_t* slots, int64_t n, float* emb_out, float* w1_out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n * kEmbDim; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / kEmbDim;
        const int k = (int)(i % kEmbDim);
        const uint8_t* rp = table + ((uint64_t)slots[r] << RowLayout<kTableDtype>::shift);
        emb_out[i] = kTableDtype == 0 ? reinterpret_cast<const float*>(rp)[k] : __half2float(reinterpret_cast<const __half*>(rp)[k]);
        if (k == 0) w1_out[r] = *reinterpret_cast<const float*>(rp + RowLayout<kTableDtype>::w1_off);
    }
}

}  // namespace ctr
