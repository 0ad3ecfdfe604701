// gather.cuh -- hashed sparse-feature embedding gather / pool kernel (north star (b)).
//
// Replaces the part of the served model that TF-Serving ran behind sendRequest
// (DCNClient.java:91-115): look up the row of every (candidate, field) id, scale it by the
// field weight, and pool.  Builder-defined math, SURVEY.md section 8(a-M):
//     x[b,f,:] = E[row] * wt          -> X[b, f*16 .. f*16+15]  (half, the GEMM's A operand)
//     lin[b]   = b0 + sum_f W1[row]*wt           (+ 0.5*sum_k[(sum_f x)^2 - sum_f x^2] for DeepFM)
//
// Data layout in HBM ("packed rows"): one table row holds the 16 embedding values AND the
// first-order weight, so a lookup is ONE place in DRAM (one 128-byte line, fetched once into L1):
//     fp32 row (128 B): [e0..e15 fp32 | w1 fp32 | pad]      f16 row (64 B): [e0..e15 half | w1 fp32 | pad]
// B200's DRAM moves 64-byte granules: an fp32 row (68 useful bytes) costs two of them, an f16 row one.
//
// Two kernels, one arithmetic.  Both map a candidate to a warp as lane = g*4 + q: the 8 lane groups g take 8 fields
// per round, lane q of a group loads embedding values 4q .. 4q+3 of the group's row (128-bit for fp32 rows, 64-bit
// for f16 rows) and lane 0 the first-order weight; per-lane sums run in field order and the warp reduces with a fixed
// shuffle tree.  Same lanes, same order, same tree => the same candidate gets bit-identical results at any
// position, in any chunk, on any GPU, from either kernel (tests/test_gpu_packed.py).
//   gather_tile_kernel   reference layout.  A CTA stages the lists of a 32-candidate tile with two bulk copies
//                        (cp.async.bulk, SASS UBLKCP, double buffered), a warp takes a candidate, issues all its row
//                        loads, then consumes them.
//   gather_pool_kernel   compact layouts.  A warp owns tiles of 4 candidates and never synchronises with the other
//                        warps; the tile's lists (bulk copy into the warp's own buffer on its own mbarrier) are first
//                        EXPANDED into one 8-byte entry {table row, weight} per (candidate, field) -- format, hash,
//                        range check and shared context handled once per id by one lane, one candidate of the next
//                        tile between two candidates of the current one -- then gathered with straight-line code
//                        (predicated loads and stores).  The loop is written as a rolling window (consume slot j of
//                        candidate c, issue slot j of candidate c+1), but every row load of a loop lands on the same
//                        hardware scoreboard (tools/sass_scoreboards.py), so waiting for the oldest load waits for the
//                        newest too and ptxas schedules it as "consume all, issue all": the overlap comes from the 32
//                        resident warps, not from the window.  DESIGN.md section 3.1 has the measurements.
//
// Input formats.  (1) the reference's: ids int64 [n, F], wts float [n, F] (DC:97-108).  (2) compact
// (kCompact): the first n_ctx fields of every candidate come from one shared list (kernel parameters),
// the per-candidate list holds the remaining Fk = F - n_ctx ids as int64 or as uint32 ROW indices
// (already hashed by the host packer), and weights only for the fields flagged in weighted_mask (the
// others are 1.0) -- what ctr_predict_shared_async / ctr_predict_packed_async ship over the host link.
//
// Footprint: 256 threads x 64 registers and <= 30 KB of shared memory per CTA: four CTAs per SM alone, or one beside
// the resident tower CTA (tower.cuh) when both kernels prefer the largest shared-memory carve-out (device_path.cuh).
//
// Roofline: HBM.  Algorithmic bytes per candidate (F=39: 13 dense + 26 sparse, fp32 table):
// 468 (ids+wts) + 26*(64+4) (rows; the 13 dense-field rows stay in L1/L2) + 4 = 2240 B.
#pragma once
#include "common.cuh"

namespace ctr {

constexpr int kWarpTile = 4;           // candidates per warp tile
constexpr int kGatherThreads = 256;    // 8 independent warps x 64 registers: the 16 K registers the tower leaves free on its SM
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
    uint64_t id_mask;          // bit f: per-candidate field f has an id column; clear: its id is ctx_ids[f] for every candidate
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
    int64_t ctx_ids[48];       // compact input, indexed by field: ids (or rows) of the context fields and of the fixed-id fields;
                               // ctx_wts: weights of the context fields 0 .. n_ctx-1
    float ctx_wts[48];
};

template <int kTableDtype>
struct RowLayout;
template <>
struct RowLayout<0> { static constexpr int shift = 7; static constexpr int w1_off = 64; };
template <>
struct RowLayout<1> { static constexpr int shift = 6; static constexpr int w1_off = 32; };

// ------------------------------------------------------------------ reference layout: the block-tile kernel
// ids int64 [n, F] / wts float [n, F] exactly as sendRequest ships them (DC:97-108).  A CTA stages the lists of a
// 32-candidate tile with two bulk copies (double buffered), a warp takes a candidate: 8 lane groups x 4 lanes, lane
// q of a group loads embedding values 4q .. 4q+3 of the group's row (128-bit) and lane 0 the first-order weight;
// all rounds' loads are issued before any is used.  Same lanes, same order of operations, same shuffle tree as the
// warp-tile kernel below (which serves the compact layouts): the two give bit-identical results.
constexpr int kGatherTile = 32;        // candidates per staged tile
constexpr int kTileThreads = 256;      // 8 warps -> 4 candidates per warp per tile
constexpr int kGatherRounds = 6;       // 8 fields per round -> F <= 48

__device__ __forceinline__ float4 ldg_nc_f4(const void* p) {
    float4 v;
    asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint2 ldg_nc_u2(const void* p) {
    uint2 v;
    asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ float ldg_nc_f1(const void* p) {
    float v;
    asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

template <int kTableDtype, bool kHashed, bool kEmb, bool kLin, bool kFM, bool kSharded>
__global__ void __launch_bounds__(kTileThreads, 4)
gather_tile_kernel(const __grid_constant__ GatherParams p) {
    extern __shared__ __align__(128) uint8_t gsm[];
    const int F = p.F;
    const int64_t* const gids = static_cast<const int64_t*>(p.ids);
    const uint32_t ids_bytes = kGatherTile * F * 8, wts_bytes = kGatherTile * F * 4;
    // [ids buf0 | ids buf1 | wts buf0 | wts buf1]; pointer arithmetic (not arrays) keeps this out of local memory
    auto ids_buf = [&](int buf) { return reinterpret_cast<int64_t*>(gsm + buf * ids_bytes); };
    auto wts_buf = [&](int buf) { return reinterpret_cast<float*>(gsm + 2 * ids_bytes + buf * wts_bytes); };
    __shared__ __align__(8) uint64_t full_bar[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const int64_t n_tiles = (p.n + kGatherTile - 1) / kGatherTile;

    if (tid == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();

    // Stage the id / weight lists of one tile.  Full, aligned tiles: one elected thread
    // issues two bulk copies that complete on full_bar[buf].  The ragged last tile (or
    // unaligned caller pointers) falls back to ordinary loads by the whole CTA.
    auto tile_is_bulk = [&](int64_t tile) { return p.bulk_ok && (tile + 1) * kGatherTile <= p.n; };
    auto stage = [&](int64_t tile, int buf) {
        const int64_t c0 = tile * kGatherTile;
        if (tile_is_bulk(tile)) {
            if (tid == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&full_bar[buf], ids_bytes + wts_bytes);
                bulk_g2s(ids_buf(buf), gids + c0 * F, ids_bytes, &full_bar[buf]);
                bulk_g2s(wts_buf(buf), p.wts + c0 * F, wts_bytes, &full_bar[buf]);
            }
        } else {
            const int64_t cnt = min((int64_t)kGatherTile, p.n - c0) * F;
            int64_t* di = ids_buf(buf);
            float* dw = wts_buf(buf);
            for (int64_t i = tid; i < cnt; i += kTileThreads) {
                di[i] = gids[c0 * F + i];
                dw[i] = p.wts[c0 * F + i];
            }
        }
    };

    uint32_t phase_bits = 0;                               // bit b = parity to wait for on full_bar[b]
    int64_t tile = blockIdx.x;
    if (tile < n_tiles) stage(tile, 0);
    __syncthreads();

    for (int it = 0; tile < n_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        const int64_t next = tile + gridDim.x;
        if (next < n_tiles) stage(next, buf ^ 1);          // buf^1 was released by the barrier below
        if (tile_is_bulk(tile)) {
            mbar_wait(&full_bar[buf], (phase_bits >> buf) & 1u);
            phase_bits ^= 1u << buf;
        }
        const int64_t c0 = tile * kGatherTile;
        const int n_here = (int)min((int64_t)kGatherTile, p.n - c0);

        for (int c = warp; c < n_here; c += kTileThreads / 32) {
            const int64_t b = c0 + c;
            const int64_t* sid = ids_buf(buf) + c * F;
            const float* swt = wts_buf(buf) + c * F;

            float4 ev[kGatherRounds];
            float wv[kGatherRounds], w1v[kGatherRounds];
#pragma unroll
            for (int r = 0; r < kGatherRounds; ++r) {
                ev[r] = make_float4(0.f, 0.f, 0.f, 0.f);
                wv[r] = 0.f;
                w1v[r] = 0.f;
                const int f = r * 8 + g;
                if (r * 8 < F && f < F) {
                    const int64_t id = sid[f];
                    wv[r] = swt[f];
                    uint64_t row;
                    if (kHashed) {
                        row = mix64((uint64_t)id) & (p.num_rows - 1);
                    } else {
                        row = (uint64_t)id;
                        if (row >= p.num_rows) { row = 0; wv[r] = 0.f; *p.error_flag = 1; }
                    }
                    const uint8_t* base = p.table;
                    if (kSharded) {
                        const uint32_t owner = (uint32_t)(row % (uint32_t)p.n_shards);
                        row = row / (uint32_t)p.n_shards;
                        base = p.shard_base[owner];
                    }
                    const uint8_t* rp = base + (row << RowLayout<kTableDtype>::shift);
                    if (kEmb || kFM) {
                        if (kTableDtype == 0) {
                            ev[r] = ldg_nc_f4(rp + q * 16);
                        } else {
                            const uint2 h = ldg_nc_u2(rp + q * 8);
                            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&h.x));
                            const float2 c2 = __half22float2(*reinterpret_cast<const __half2*>(&h.y));
                            ev[r] = make_float4(a.x, a.y, c2.x, c2.y);
                        }
                    }
                    if (kLin && q == 0) w1v[r] = ldg_nc_f1(rp + RowLayout<kTableDtype>::w1_off);
                }
            }

            float lin = 0.f, s2 = 0.f;
            float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < kGatherRounds; ++r) {
                const int f = r * 8 + g;
                if (r * 8 < F) {
                    const float wt = wv[r];
                    const float4 x = make_float4(ev[r].x * wt, ev[r].y * wt, ev[r].z * wt, ev[r].w * wt);
                    if (kLin) lin = fmaf(w1v[r], wt, lin);
                    if (kFM) {
                        s1.x += x.x; s1.y += x.y; s1.z += x.z; s1.w += x.w;
                        s2 = fmaf(x.x, x.x, s2); s2 = fmaf(x.y, x.y, s2); s2 = fmaf(x.z, x.z, s2); s2 = fmaf(x.w, x.w, s2);
                    }
                    if (kEmb && f < F) {
                        uint2 o;                                 // saturating: beyond the half range -> +-65504, never inf
                        o.x = pack_half2_sat(x.x, x.y);
                        o.y = pack_half2_sat(x.z, x.w);
                        *reinterpret_cast<uint2*>(p.X + b * p.ldx + f * kEmbDim + q * 4) = o;
                    }
                }
            }

            if (kLin || kFM) {
                float acc = lin;                              // only q==0 lanes hold first-order terms
                if (kFM) {
                    // field sums per embedding coordinate: reduce over the 8 lane groups
#pragma unroll
                    for (int off = 4; off <= 16; off <<= 1) {
                        s1.x += __shfl_xor_sync(0xffffffffu, s1.x, off);
                        s1.y += __shfl_xor_sync(0xffffffffu, s1.y, off);
                        s1.z += __shfl_xor_sync(0xffffffffu, s1.z, off);
                        s1.w += __shfl_xor_sync(0xffffffffu, s1.w, off);
                    }
                    const float sq = s1.x * s1.x + s1.y * s1.y + s1.z * s1.z + s1.w * s1.w;
                    acc += (g == 0 ? 0.5f * sq : 0.f) - 0.5f * s2;
                }
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
                if (lane == 0) {
                    const float z = acc + p.b0;
                    if (p.wide) p.wide[b] = z;
                    if (p.scores) p.scores[b] = sigmoidf_fast(z);
                }
            }
        }
        __syncthreads();   // every warp is done with buf before it is staged again
    }
}


// ------------------------------------------------------------------ compact layouts: the warp-tile kernel
constexpr int kLanesPerRow = 4;
constexpr int kRowsPerLoad = 32 / kLanesPerRow;    // 8 rows per load instruction

// One lane's piece of a packed row.  Loads are predicated (no branch: a lane that does not load keeps what it has)
// and do not allocate in L1: a row is used once.
//   fp32 rows: lane q of a group takes embedding values 4q .. 4q+3 (one 128-bit load; the four lanes cover the 64
//              bytes of the embedding) and lane 0 also takes the first-order weight (one 32-bit load): 5 registers a
//              slot, so that a window of 40 rows per warp fits the 64 registers that let 32 warps live on an SM.
//              (A variant with one 256-bit load per lane -- row and weight in one instruction -- needed 8 registers
//              a slot, ran at 24 warps per SM and was slower: the kernel lives on rows in flight.)
//   f16 rows:  the same mapping with 64-bit loads (4 halfs per lane).
template <int kTableDtype>
struct RowPiece;
template <>
struct RowPiece<0> {
    static constexpr int nv = 4;                   // embedding values per embedding lane
    static constexpr int emb_lanes = 4;
    static constexpr int w1_lane = 0;
    uint32_t r[4];
    uint32_t w;
    __device__ __forceinline__ void zero() { r[0] = r[1] = r[2] = r[3] = 0u; w = 0u; }
    __device__ __forceinline__ void load(const uint8_t* row, int q, bool emb_on, bool w1_on) {
        asm volatile("{\n\t.reg .pred P, Q;\n\tsetp.ne.u32 P, %7, 0;\n\tsetp.ne.u32 Q, %8, 0;\n\t"
                     "@P ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%5];\n\t"
                     "@Q ld.global.nc.u32 %4, [%6];\n\t}"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(w)
                     : "l"(row + q * 16), "l"(row + 64), "r"((uint32_t)emb_on), "r"((uint32_t)w1_on));
    }
    __device__ __forceinline__ float val(int k) const { return __uint_as_float(r[k]); }
    __device__ __forceinline__ float w1() const { return __uint_as_float(w); }
};
template <>
struct RowPiece<1> {
    static constexpr int nv = 4;
    static constexpr int emb_lanes = 4;
    static constexpr int w1_lane = 0;
    uint32_t r[2];
    uint32_t w;
    __device__ __forceinline__ void zero() { r[0] = r[1] = 0u; w = 0u; }
    __device__ __forceinline__ void load(const uint8_t* row, int q, bool emb_on, bool w1_on) {
        asm volatile("{\n\t.reg .pred P, Q;\n\tsetp.ne.u32 P, %5, 0;\n\tsetp.ne.u32 Q, %6, 0;\n\t"
                     "@P ld.global.nc.v2.u32 {%0,%1}, [%3];\n\t"
                     "@Q ld.global.nc.u32 %2, [%4];\n\t}"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(w) : "l"(row + q * 8), "l"(row + 32), "r"((uint32_t)emb_on), "r"((uint32_t)w1_on));
    }
    __device__ __forceinline__ float val(int k) const {
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&r[k >> 1]));
        return (k & 1) ? a.y : a.x;
    }
    __device__ __forceinline__ float w1() const { return __uint_as_float(w); }
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
// sum.  kFM: DeepFM second-order term.  kSharded: rows striped over the devices (row mod n_shards), read through peer pointers.
template <int kTableDtype, int kIters, bool kCompact, bool kEmb, bool kLin, bool kFM, bool kSharded>
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
        for (int f = tid; f < F; f += kGatherThreads) { s_ctx_id[f] = p.ctx_ids[f]; s_ctx_wt[f] = p.ctx_wts[f]; }
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
                    const int j = f - p.n_ctx;
                    const uint64_t im = p.id_mask >> p.n_ctx;
                    if ((im >> j) & 1ull) {
                        const int col = c * Fk + __popcll(im & ((1ull << j) - 1ull));
                        id = p.id_bytes == 4 ? (uint64_t)reinterpret_cast<const uint32_t*>(raw)[col] : (uint64_t)reinterpret_cast<const int64_t*>(raw)[col];
                    } else {
                        id = (uint64_t)s_ctx_id[f];                  // the same id for every candidate of the request
                    }
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
    constexpr int NV = Piece::nv;
    Piece ev[kIters];                                          // the rolling window: slot j = row of field j*8 + g
    const bool emb_lane = q < Piece::emb_lanes, w1_lane = q == Piece::w1_lane;
    const bool emb_on = emb_lane && (kEmb || kFM), w1_on = w1_lane && kLin;
    uint32_t slot_mask = 0;                                    // bit j: field j*8 + g exists
#pragma unroll
    for (int j = 0; j < kIters; ++j) slot_mask |= (j * kRowsPerLoad + g < F ? 1u : 0u) << j;
#pragma unroll
    for (int j = 0; j < kIters; ++j) ev[j].zero();             // lanes / slots that never load keep zeros for good

    // the entries {row, weight} of the candidate whose entry of field g is at shared address e0, one per slot of this
    // lane ({0, 0.0f} for a slot the lane does not have).  Fetched together in front of the candidate's loads: inside
    // the consume / issue sequence below the only long-latency instructions left are the row loads themselves, so
    // that each slot's loads can have a scoreboard of their own (tools/sass_scoreboards.py: with the shared-memory
    // reads and the shard arithmetic in between, ptxas put every row load on ONE scoreboard, and waiting for the
    // oldest load then also waited for the newest -- one full DRAM latency per slot)
    auto lds_entries = [&](uint32_t e0, uint2 (&ent)[kIters]) {
#pragma unroll
        for (int j = 0; j < kIters; ++j) {
            ent[j] = make_uint2(0u, 0u);
            asm volatile("{\n\t.reg .pred P;\n\tsetp.ne.u32 P, %3, 0;\n\t@P ld.shared.v2.u32 {%0,%1}, [%2];\n\t}"
                         : "+r"(ent[j].x), "+r"(ent[j].y) : "r"(e0 + j * (kRowsPerLoad * 8)), "r"((slot_mask >> j) & 1u));
        }
    };
    // issue the loads of slot j: table row `row`
    auto issue = [&](int j, uint32_t row) {
        const bool on = (slot_mask >> j) & 1u;
        const uint8_t* base = p.table;
        if (kSharded) {
            base = p.shard_base[row % (uint32_t)p.n_shards];
            row = row / (uint32_t)p.n_shards;
        }
        ev[j].load(base + ((uint64_t)row << RL::shift), q, emb_on && on, w1_on && on);
    };

    // per-candidate accumulators (of the candidate being consumed)
    float lin, s2, s1[NV];
    auto reset_acc = [&]() {
        lin = 0.f; s2 = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) s1[k] = 0.f;
    };
    // consume slot j (weight wt; 0 for a slot this lane does not have) of the candidate whose X piece starts at xrow
    auto consume = [&](int j, float wt, __half* xrow) {
        if (kEmb || kFM) {
            float x[NV];
#pragma unroll
            for (int k = 0; k < NV; ++k) x[k] = ev[j].val(k) * wt;     // every lane: the idle lanes' products go nowhere
            if (kFM) {
                if (emb_lane) {                                        // (in f16 rows the w1 lane holds pad bytes, not values)
#pragma unroll
                    for (int k = 0; k < NV; ++k) { s1[k] += x[k]; s2 = fmaf(x[k], x[k], s2); }
                }
            }
            if (kEmb) {
                uint32_t o[NV / 2];
#pragma unroll
                for (int k = 0; k < NV / 2; ++k) o[k] = pack_half2_sat(x[2 * k], x[2 * k + 1]);
                if (emb_lane && ((slot_mask >> j) & 1u)) {
                    __half* dst = xrow + j * (kRowsPerLoad * kEmbDim);
                    *reinterpret_cast<uint2*>(dst) = make_uint2(o[0], o[1]);
                }
            }
        }
        if (kLin) lin = fmaf(ev[j].w1(), w1_lane ? wt : 0.f, lin);     // other lanes: a finite value (or zero) times zero
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
                    for (int k = 0; k < NV; ++k) s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], off);
                }
                const float sq = s1[0] * s1[0] + s1[1] * s1[1] + s1[2] * s1[2] + s1[3] * s1[3];     // as the tile kernel writes it
                acc += (g == 0 && emb_lane ? 0.5f * sq : 0.f) - 0.5f * s2;
            }
            // first-order terms live in one lane of every group: without FM three steps over the lane groups are enough
#pragma unroll
            for (int off = 16; off >= (kFM ? 1 : kLanesPerRow); off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
            if (lane == (kFM ? 0 : Piece::w1_lane)) {
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
    const int xlane = g * kEmbDim + q * NV;                // this lane's halfs of field g within a row of X
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
