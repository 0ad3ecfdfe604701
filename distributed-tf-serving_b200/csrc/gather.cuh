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
// Mapping.  A CTA stages the id / weight lists of a 32-candidate tile with bulk copies (cp.async.bulk, SASS UBLKCP,
// on an mbarrier, double buffered); a warp takes a candidate as lane = g*4 + q: the 8 lane groups g take 8 fields per
// round, lane q of a group loads embedding values 4q .. 4q+3 of the group's row (128-bit for fp32 rows, 64-bit for f16
// rows) and lane 0 the first-order weight; all rounds' loads are issued before any is used; per-lane sums run in field
// order and the warp reduces with a fixed shuffle tree, so the same candidate gets bit-identical results at any
// position, in any chunk, on any GPU, from any input layout.  For the compact layouts (kCompact) every thread of the
// CTA first EXPANDS its share of the tile into one 8-byte entry {table row, weight} per (candidate, field) -- format,
// hash, range check, shared context and fixed ids handled once per id -- and the gather reads those entries.
//
// Round 2 also built a warp-autonomous variant (4-candidate tiles per warp, own mbarriers, a rolling window of row
// loads across candidates, 256-bit loads): every row load of its loop landed on the same hardware scoreboard
// (tools/sass_scoreboards.py), so waiting for the oldest load waited for the newest too, and it never beat this kernel
// (170-190 against 139-149 us per 151 552 candidates).  DESIGN.md section 3.1 has the measurements; the code is in the
// history (commit "Reference layout back on the block-tile gather kernel").
//
// Input formats.  (1) the reference's: ids int64 [n, F], wts float [n, F] (DC:97-108).  (2) compact
// (kCompact): the first n_ctx fields of every candidate come from one shared list (kernel parameters),
// the per-candidate list holds the ids of the remaining fields that have an id column (id_mask; the others have
// one id for the whole request, also in the kernel parameters) as int64 or as uint32 ROW indices
// (already hashed by the host packer), and weights only for the fields flagged in weighted_mask (the
// others are 1.0) -- what ctr_predict_shared_async / ctr_predict_packed_async ship over the host link.
//
// Footprint: 256 threads x 64 registers and <= 30 KB of shared memory per CTA for the reference layout: four CTAs per SM.
//
// Roofline: HBM.  Algorithmic bytes per candidate (F=39: 13 dense + 26 sparse, fp32 table):
// 468 (ids+wts) + 26*(64+4) (rows; the 13 dense-field rows stay in L1/L2) + 4 = 2240 B.
#pragma once
#include "common.cuh"

namespace ctr {


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

// kCompact: the staged lists are a compact layout (shared context / uint32 rows / sparse weights / fixed ids); every
// thread of the CTA first expands its share of the tile into {row, weight} entries, the gather then reads those.
template <int kTableDtype, bool kHashed, bool kEmb, bool kLin, bool kFM, bool kSharded, bool kCompact = false>
__global__ void __launch_bounds__(kTileThreads, 4)
gather_tile_kernel(const __grid_constant__ GatherParams p) {
    extern __shared__ __align__(128) uint8_t gsm[];
    const int F = p.F;
    const int Fk = kCompact ? p.Fk : F, Wk = kCompact ? p.Wk : F, id_bytes = kCompact ? p.id_bytes : 8;
    const uint8_t* const gids = static_cast<const uint8_t*>(p.ids);
    const uint32_t ids_bytes = kGatherTile * Fk * id_bytes, wts_bytes = kGatherTile * Wk * 4;
    // [ids buf0 | ids buf1 | wts buf0 | wts buf1 | entries]; pointer arithmetic (not arrays) keeps this out of local memory
    auto ids_buf = [&](int buf) { return reinterpret_cast<int64_t*>(gsm + buf * ids_bytes); };
    auto wts_buf = [&](int buf) { return reinterpret_cast<float*>(gsm + 2 * ids_bytes + buf * wts_bytes); };
    uint2* const ent = reinterpret_cast<uint2*>(gsm + 2 * (ids_bytes + wts_bytes));      // [32, F] {row, weight} (kCompact)
    __shared__ __align__(8) uint64_t full_bar[2];
    __shared__ int64_t s_ctx_id[kCompact ? 48 : 1];
    __shared__ float s_ctx_wt[kCompact ? 48 : 1];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const int64_t n_tiles = (p.n + kGatherTile - 1) / kGatherTile;

    if (tid == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        mbar_fence_init();
    }
    if (kCompact) {
        for (int f = tid; f < F; f += kTileThreads) { s_ctx_id[f] = p.ctx_ids[f]; s_ctx_wt[f] = p.ctx_wts[f]; }
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
                bulk_g2s(ids_buf(buf), gids + (size_t)c0 * Fk * id_bytes, ids_bytes, &full_bar[buf]);
                if (wts_bytes) bulk_g2s(wts_buf(buf), p.wts + (size_t)c0 * Wk, wts_bytes, &full_bar[buf]);
            }
        } else {
            const int cnt = (int)min((int64_t)kGatherTile, p.n - c0);
            const uint32_t* si = reinterpret_cast<const uint32_t*>(gids + (size_t)c0 * Fk * id_bytes);
            uint32_t* di = reinterpret_cast<uint32_t*>(ids_buf(buf));
            float* dw = wts_buf(buf);
            for (int i = tid; i < cnt * Fk * (id_bytes >> 2); i += kTileThreads) di[i] = si[i];
            for (int i = tid; i < cnt * Wk; i += kTileThreads) dw[i] = p.wts[(size_t)c0 * Wk + i];
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
        if (kCompact) {
            // expansion: one entry {table row, weight} per (candidate, field) -- format, hash, range check, shared context
            // and fixed ids handled once per id; an id outside a direct table becomes row 0 with weight 0 and raises the flag
            if (!tile_is_bulk(tile)) __syncthreads();      // plain-load staging by the whole CTA: visible to everyone first
            const uint8_t* raw = reinterpret_cast<const uint8_t*>(ids_buf(buf));
            const float* swt = wts_buf(buf);
            const uint64_t im = p.id_mask >> p.n_ctx, wm = p.weighted_mask >> p.n_ctx;
            for (int i = tid; i < n_here * F; i += kTileThreads) {
                const int c = i / F, f = i - c * F;
                uint64_t id;
                float wt;
                if (f < p.n_ctx) {
                    id = (uint64_t)s_ctx_id[f];
                    wt = s_ctx_wt[f];
                } else {
                    const int j = f - p.n_ctx;
                    if ((im >> j) & 1ull) {
                        const int col = c * Fk + __popcll(im & ((1ull << j) - 1ull));
                        id = id_bytes == 4 ? (uint64_t)reinterpret_cast<const uint32_t*>(raw)[col] : (uint64_t)reinterpret_cast<const int64_t*>(raw)[col];
                    } else {
                        id = (uint64_t)s_ctx_id[f];
                    }
                    wt = ((wm >> j) & 1ull) ? swt[c * Wk + __popcll(wm & ((1ull << j) - 1ull))] : 1.0f;
                }
                uint64_t row = p.hashed ? (mix64(id) & (p.num_rows - 1)) : id;
                if (row >= p.num_rows) { row = 0; wt = 0.f; *p.error_flag = 1; }
                ent[i] = make_uint2((uint32_t)row, __float_as_uint(wt));
            }
            __syncthreads();
        }

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
                    uint64_t row;
                    if (kCompact) {
                        const uint2 en = ent[c * F + f];
                        row = en.x;
                        wv[r] = __uint_as_float(en.y);
                    } else {
                        const int64_t id = sid[f];
                        wv[r] = swt[f];
                        if (kHashed) {
                            row = mix64((uint64_t)id) & (p.num_rows - 1);
                        } else {
                            row = (uint64_t)id;
                            if (row >= p.num_rows) { row = 0; wv[r] = 0.f; *p.error_flag = 1; }
                        }
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
