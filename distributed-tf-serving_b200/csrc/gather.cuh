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
// Mapping (round 2).  A warp owns "warp tiles" of 4 consecutive candidates and never synchronises with the other
// warps of its CTA.  Two phases per tile:
//   expansion  the tile's id / weight lists, staged into the warp's own shared memory with cp.async.bulk (SASS
//              UBLKCP) on the warp's own mbarrier, are turned into one 8-byte entry {table row, weight} per
//              (candidate, field): input format, hash, range check and shared context are handled ONCE per id by
//              one lane, one candidate of the NEXT tile between two candidates of the current one.
//   gather     4 lanes per row, 8 rows per load instruction: an fp32 row is fetched with ONE 256-bit load per lane
//              (SASS LDG.E.256: lanes 0-1 the embedding halves, lane 2 the first-order weight), an f16 row with one
//              128-bit load.  The loads of candidate c+1 are issued slot by slot while candidate c's rows are
//              consumed out of the same registers (a rolling window of 5-6 loads = 40-48 rows per warp, always
//              full), so a warp never drains between candidates or tiles.  Inside a candidate the code is straight
//              line: predicated loads and stores, no branch.
// Pooling (first-order sum, FM sums) is a fixed per-lane order plus a fixed warp-shuffle tree: the same candidate
// gets bit-identical results at any position, in any chunk, on any GPU, from any input format.
//
// Input formats.  (1) the reference's: ids int64 [n, F], wts float [n, F] (DC:97-108).  (2) compact
// (kCompact): the first n_ctx fields of every candidate come from one shared list (kernel parameters),
// the per-candidate list holds the remaining Fk = F - n_ctx ids as int64 or as uint32 ROW indices
// (already hashed by the host packer), and weights only for the fields flagged in weighted_mask (the
// others are 1.0) -- what ctr_predict_shared_async / ctr_predict_packed_async ship over the host link.
//
// Footprint: 192 threads x 80 registers and ~23 KB of shared memory per CTA, so that one CTA fits on an SM
// NEXT TO the resident tower CTA (tower.cuh) and a chunk's gather runs under the previous chunk's tower.
//
// Roofline: HBM.  Algorithmic bytes per candidate (F=39: 13 dense + 26 sparse, fp32 table):
// 468 (ids+wts) + 26*(64+4) (rows; the 13 dense-field rows stay in L1/L2) + 4 = 2240 B.
#pragma once
#include "common.cuh"

namespace ctr {

constexpr int kWarpTile = 4;           // candidates per warp tile
constexpr int kGatherThreads = 192;    // 6 independent warps x 80 registers: 15 K registers, what the tower leaves free on its SM
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
struct RowLayout<0> { static constexpr int shift = 7; static constexpr int w1_off = 64; static constexpr int lane_bytes = 32; };
template <>
struct RowLayout<1> { static constexpr int shift = 6; static constexpr int w1_off = 32; static constexpr int lane_bytes = 16; };

constexpr int kLanesPerRow = 4;                    // lane q of a group: 0, 1 = embedding halves, 2 = first-order weight, 3 idle
constexpr int kRowsPerLoad = 32 / kLanesPerRow;    // 8 rows per load instruction

// One lane's piece of a packed row.  fp32 rows: 32 bytes (one 256-bit load, SASS LDG.E.256) -- lanes 0-2 of a group
// cover [e0..e7 | e8..e15 | w1, pad]; f16 rows: 16 bytes -- [e0..e7 | e8..e15 | w1, pad].  The load is predicated
// (no branch: a lane that does not load keeps what it has) and does not allocate in L1: a row is used once.
template <int kTableDtype>
struct RowPiece;
template <>
struct RowPiece<0> {
    uint32_t r[8];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = 0u;
    }
    __device__ __forceinline__ void load(const void* p, bool on) {
        asm volatile("{\n\t.reg .pred P;\n\tsetp.ne.u32 P, %9, 0;\n\t"
                     "@P ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n\t}"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]) : "l"(p), "r"((uint32_t)on));
    }
    __device__ __forceinline__ float val(int k) const { return __uint_as_float(r[k]); }
    __device__ __forceinline__ float w1() const { return __uint_as_float(r[0]); }
};
template <>
struct RowPiece<1> {
    uint32_t r[4];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = 0u;
    }
    __device__ __forceinline__ void load(const void* p, bool on) {
        asm volatile("{\n\t.reg .pred P;\n\tsetp.ne.u32 P, %5, 0;\n\t"
                     "@P ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]) : "l"(p), "r"((uint32_t)on));
    }
    __device__ __forceinline__ float val(int k) const {
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&r[k >> 1]));
        return (k & 1) ? a.y : a.x;
    }
    __device__ __forceinline__ float w1() const { return __uint_as_float(r[0]); }
};

// Shared memory of one warp (bytes): the raw id / weight lists as they arrive, and the expanded {row, weight} entries.
// Reference layout: two raw buffers of 4*F*12 bytes, expanded IN PLACE (an entry takes the 8 bytes of its int64 id).
// Compact layouts: one raw buffer (the lists are consumed by the expansion) + two expanded buffers of 4*F*8 bytes.
__host__ __device__ inline uint32_t gather_raw_bytes(int Fk, int Wk, int id_bytes) { return (uint32_t)(kWarpTile * (Fk * id_bytes + Wk * 4)); }
__host__ __device__ inline uint32_t gather_warp_smem(bool compact, int F, int Fk, int Wk, int id_bytes) {
    return compact ? gather_raw_bytes(Fk, Wk, id_bytes) + 2u * kWarpTile * F * 8u : 2u * gather_raw_bytes(F, F, 8);
}

// kTableDtype: 0 fp32 rows, 1 f16 rows.  kIters: load slots per lane and candidate (kIters * 8 >= F).
// kCompact: compact input (shared context / uint32 rows / sparse weights).  kEmb: write X.  kLin: first-order
// sum.  kFM: DeepFM second-order term.
template <int kTableDtype, int kIters, bool kCompact, bool kEmb, bool kLin, bool kFM>
__global__ void __launch_bounds__(kGatherThreads, 4)
gather_pool_kernel(const __grid_constant__ GatherParams p) {
    using RL = RowLayout<kTableDtype>;
    using Piece = RowPiece<kTableDtype>;
    extern __shared__ __align__(128) uint8_t gsm[];
    __shared__ __align__(8) uint64_t full_bar[kGatherWarps][2];
    __shared__ int64_t s_ctx_id[kCompact ? 48 : 1];
    __shared__ float s_ctx_wt[kCompact ? 48 : 1];

    const int F = p.F, Fk = p.Fk, Wk = p.Wk;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t ids_bytes = (uint32_t)(kWarpTile * Fk * p.id_bytes), wts_bytes = (uint32_t)(kWarpTile * Wk * 4);   // multiples of 16
    const uint32_t raw_bytes = ids_bytes + wts_bytes;
    const uint32_t exp_bytes = (uint32_t)(kWarpTile * F * 8);
    uint8_t* const wbase = gsm + (size_t)warp * gather_warp_smem(kCompact, F, Fk, Wk, p.id_bytes);
    // buffer b of the raw lists / of the expanded entries
    auto raw_buf = [&](int b) -> uint8_t* { return kCompact ? wbase : wbase + b * raw_bytes; };
    auto exp_buf = [&](int b) -> uint8_t* { return kCompact ? wbase + raw_bytes + b * exp_bytes : wbase + b * raw_bytes; };

    if (kCompact) {
        for (int f = tid; f < p.n_ctx; f += kGatherThreads) { s_ctx_id[f] = p.ctx_ids[f]; s_ctx_wt[f] = p.ctx_wts[f]; }
    }
    if (lane == 0) {
        mbar_init(&full_bar[warp][0], 1);
        mbar_init(&full_bar[warp][1], 1);
        mbar_fence_init();
    }
    __syncthreads();                        // the only block-wide barrier: from here on every warp is on its own

    // Stage the raw lists of one warp tile into raw buffer `b`.  Full, aligned tiles: one lane issues two bulk
    // copies that complete on the warp's mbarrier.  The ragged last tile (or unaligned caller pointers) is copied
    // with ordinary loads by the warp.
    const uint8_t* const gids = static_cast<const uint8_t*>(p.ids);
    auto tile_is_bulk = [&](int tile) { return p.bulk_ok && (int64_t)(tile + 1) * kWarpTile <= p.n; };
    auto stage = [&](int tile, int b) {
        const int64_t c0 = (int64_t)tile * kWarpTile;
        uint8_t* dst = raw_buf(b);
        uint64_t* bar = &full_bar[warp][kCompact ? 0 : b];
        __syncwarp();                                          // every lane is done with what the buffer held before
        if (tile_is_bulk(tile)) {
            if (lane == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(bar, raw_bytes);
                bulk_g2s(dst, gids + (size_t)c0 * Fk * p.id_bytes, ids_bytes, bar);
                if (wts_bytes) bulk_g2s(dst + ids_bytes, p.wts + (size_t)c0 * Wk, wts_bytes, bar);
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

    // Expansion: every (candidate, field) of a tile becomes one 8-byte entry {table row, weight}.  This is where
    // the input format, the hash, the range check and the shared context are dealt with -- ONCE per id, by one lane
    // (the gather loop below is the same handful of instructions whatever the format).  An id outside a direct
    // table becomes row 0 with weight 0 and raises the error flag.
    auto expand_one = [&](const uint8_t* raw, uint8_t* ex, int c) {
        const float* swt = reinterpret_cast<const float*>(raw + ids_bytes);
        for (int f = lane; f < F; f += 32) {
            uint64_t id;
            float wt;
            if (kCompact) {
                if (f < p.n_ctx) {
                    id = (uint64_t)s_ctx_id[f];
                    wt = s_ctx_wt[f];
                } else {
                    const int j = f - p.n_ctx, col = c * Fk + j;
                    id = p.id_bytes == 4 ? (uint64_t)reinterpret_cast<const uint32_t*>(raw)[col] : (uint64_t)reinterpret_cast<const int64_t*>(raw)[col];
                    const uint64_t m = p.weighted_mask >> p.n_ctx;
                    wt = ((m >> j) & 1ull) ? swt[c * Wk + __popcll(m & ((1ull << j) - 1ull))] : 1.0f;
                }
            } else {
                id = (uint64_t)reinterpret_cast<const int64_t*>(raw)[c * F + f];
                wt = swt[c * F + f];
            }
            uint64_t row = p.hashed ? (mix64(id) & (p.num_rows - 1)) : id;
            if (row >= p.num_rows) { row = 0; wt = 0.f; *p.error_flag = 1; }
            reinterpret_cast<uint2*>(ex)[c * F + f] = make_uint2((uint32_t)row, __float_as_uint(wt));
        }
    };

    // ---- gather phase: straight-line code, no branch inside a candidate.
    Piece ev[kIters];                                          // the rolling window: slot j = row of field j*8 + g
    const bool emb_lane = q < 2, w1_lane = q == 2;
    const bool load_lane = (emb_lane && (kEmb || kFM)) || (w1_lane && kLin);
    const uint32_t lane_off = (uint32_t)q * RL::lane_bytes;
    uint32_t slot_mask = 0;                                    // bit j: field j*8 + g exists
#pragma unroll
    for (int j = 0; j < kIters; ++j) slot_mask |= (j * kRowsPerLoad + g < F ? 1u : 0u) << j;
#pragma unroll
    for (int j = 0; j < kIters; ++j) ev[j].zero();             // lanes / slots that never load keep zeros for good

    // the entries {row, weight} of the candidate at shared address e0, one per slot of this lane; a slot that does
    // not exist for this lane reads as {0, 0.0f}.  All slots are fetched together: one shared-memory latency per
    // candidate, not one in front of every load
    auto lds_entries = [&](uint32_t e0, uint2 (&ent)[kIters]) {
#pragma unroll
        for (int j = 0; j < kIters; ++j) {
            ent[j] = make_uint2(0u, 0u);
            asm volatile("{\n\t.reg .pred P;\n\tsetp.ne.u32 P, %3, 0;\n\t@P ld.shared.v2.u32 {%0,%1}, [%2];\n\t}"
                         : "+r"(ent[j].x), "+r"(ent[j].y) : "r"(e0 + j * (kRowsPerLoad * 8)), "r"((slot_mask >> j) & 1u));
        }
    };
    // issue the load of slot j: table row `row`
    auto issue = [&](int j, uint32_t row) {
        const uint8_t* base = p.table;
        if (p.n_shards > 1) {
            base = p.shard_base[row % (uint32_t)p.n_shards];
            row = row / (uint32_t)p.n_shards;
        }
        ev[j].load(base + (((uint64_t)row << RL::shift) | lane_off), load_lane && ((slot_mask >> j) & 1u));
    };

    // per-candidate accumulators (of the candidate being consumed)
    float lin, s2, s1[8];
    auto reset_acc = [&]() {
        lin = 0.f; s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) s1[k] = 0.f;
    };
    // consume slot j (weight wt; 0 for a slot this lane does not have) of the candidate whose X piece starts at xrow
    auto consume = [&](int j, float wt, __half* xrow) {
        if (kEmb || kFM) {
            float x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = ev[j].val(k) * wt;      // every lane: the idle lanes' products go nowhere
            if (kFM) {
                if (emb_lane) {                                        // (the w1 lane holds pad bytes, not embedding values)
#pragma unroll
                    for (int k = 0; k < 8; ++k) { s1[k] += x[k]; s2 = fmaf(x[k], x[k], s2); }
                }
            }
            if (kEmb) {
                uint4 o;
                o.x = pack_half2_sat(x[0], x[1]); o.y = pack_half2_sat(x[2], x[3]);
                o.z = pack_half2_sat(x[4], x[5]); o.w = pack_half2_sat(x[6], x[7]);
                if (emb_lane && ((slot_mask >> j) & 1u)) *reinterpret_cast<uint4*>(xrow + j * (kRowsPerLoad * kEmbDim)) = o;
            }
        }
        if (kLin) lin = fmaf(ev[j].w1(), w1_lane ? wt : 0.f, lin);     // embedding lanes: a finite value times zero
    };
    // after the last slot of a candidate: pool over the warp and write lin (+FM) / the LR score
    auto finish = [&](int b) {
        if (kLin || kFM) {
            float acc = lin;                                  // only the w1 lanes hold first-order terms
            if (kFM) {
                // field sums per embedding coordinate: reduce over the 8 lane groups (same q, different g)
#pragma unroll
                for (int off = kLanesPerRow; off <= 16; off <<= 1) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], off);
                }
                float sq = 0.f;
#pragma unroll
                for (int k = 0; k < 8; ++k) sq = fmaf(s1[k], s1[k], sq);
                acc += (g == 0 && emb_lane ? 0.5f * sq : 0.f) - 0.5f * s2;
            }
            // first-order terms live in the q == 2 lanes only: without FM three steps over the lane groups are enough
#pragma unroll
            for (int off = 16; off >= (kFM ? 1 : kLanesPerRow); off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
            if (lane == (kFM ? 0 : 2)) {
                const float z = acc + p.b0;
                if (p.wide) p.wide[b] = z;
                if (p.scores) p.scores[b] = sigmoidf_fast(z);
            }
        }
    };

    const int n_tiles = (int)((p.n + kWarpTile - 1) / kWarpTile);           // a launch scores fewer than 2^31 candidates
    const int tile_stride = (int)gridDim.x * kGatherWarps;
    int tile = (int)blockIdx.x * kGatherWarps + warp;
    if (tile >= n_tiles) return;
    stage(tile, 0);

    uint32_t phase_bits = 0;                               // bit i = parity to wait for on full_bar[warp][i]
    auto wait_staged = [&](int t, int b) {                 // the raw lists of tile t have landed in raw buffer b
        const int bar_i = kCompact ? 0 : b;
        if (tile_is_bulk(t)) {
            mbar_wait(&full_bar[warp][bar_i], (phase_bits >> bar_i) & 1u);
            phase_bits ^= 1u << bar_i;
        }
    };
    // the first tile is expanded up front; every later tile is expanded one candidate at a time BETWEEN the
    // candidates of the tile before it, while that tile's row loads are in flight
    wait_staged(tile, 0);
    {
        const int n0 = min(kWarpTile, (int)(p.n - (int64_t)tile * kWarpTile));
        for (int c = 0; c < n0; ++c) expand_one(raw_buf(0), exp_buf(0), c);
        __syncwarp();
    }
    bool have_prev = false;                                // a candidate's loads are in flight in ev[]
    float prev_wt[kIters];                                 // its weights, one per slot of this lane
    int prev_b = 0;                                        // its index in this launch
#pragma unroll
    for (int j = 0; j < kIters; ++j) prev_wt[j] = 0.f;
    const int xlane = g * kEmbDim + q * 8;                 // this lane's 8 halfs of field g within a row of X
    for (int it = 0; tile < n_tiles; tile += tile_stride, ++it) {
        const int bsel = it & 1;
        const int c0 = tile * kWarpTile;
        const int n_here = min(kWarpTile, (int)p.n - c0);
        const int next = tile + tile_stride;
        const int n_next = next < n_tiles ? min(kWarpTile, (int)p.n - next * kWarpTile) : 0;
        uint32_t cur_e = smem_u32(exp_buf(bsel)) + g * 8;
        for (int c = 0; c < kWarpTile; ++c, cur_e += F * 8) {
            if (c < n_here) {
                uint2 ent[kIters];
                lds_entries(cur_e, ent);
                if (have_prev) {
                    __half* xrow = kEmb ? p.X + ((size_t)prev_b * p.ldx + xlane) : nullptr;
                    reset_acc();
#pragma unroll
                    for (int j = 0; j < kIters; ++j) {
                        consume(j, prev_wt[j], xrow);
                        issue(j, ent[j].x);
                    }
                    finish(prev_b);
                } else {
#pragma unroll
                    for (int j = 0; j < kIters; ++j) issue(j, ent[j].x);
                }
#pragma unroll
                for (int j = 0; j < kIters; ++j) prev_wt[j] = __uint_as_float(ent[j].y);
                have_prev = true; prev_b = c0 + c;
            }
            if (n_next) {
                // c == 0: the previous tile's buffers have been read for the last time -> stage the next tile;
                // c >= 1: its lists have had a candidate's worth of time to land -> expand candidate c-1 (and the last)
                if (c == 0) stage(next, bsel ^ 1);
                else {
                    if (c == 1) wait_staged(next, bsel ^ 1);
                    if (c - 1 < n_next) expand_one(raw_buf(bsel ^ 1), exp_buf(bsel ^ 1), c - 1);
                    if (c == kWarpTile - 1) {
                        if (c < n_next) expand_one(raw_buf(bsel ^ 1), exp_buf(bsel ^ 1), c);
                        __syncwarp();
                    }
                }
            }
        }
    }
    if (have_prev) {
        __half* xrow = kEmb ? p.X + ((size_t)prev_b * p.ldx + xlane) : nullptr;
        reset_acc();
#pragma unroll
        for (int j = 0; j < kIters; ++j) consume(j, prev_wt[j], xrow);
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
