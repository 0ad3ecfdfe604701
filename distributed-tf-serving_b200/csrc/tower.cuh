// tower.cuh -- the dense tower as ONE persistent tcgen05 / TMEM / TMA kernel per chain of layers
// (north star (c)).  A chain is a sequence of layers in which each layer's output is the next
// layer's input: the MLP of Wide&Deep / DeepFM / DCN-v2, or DCN-v2's cross network.
//
//     layer:  OUT[m, n] = epi( sum_k A[m, k] * W[n, k] + bias[n] )        A, W: half, K-major
//       EPI_RELU      OUT = max(acc + b, 0)                       -> half, next layer's A
//       EPI_CROSS     OUT = X0[m,n] * (acc + b) + A[m,n]          -> half, next layer's A   (DCN-v2 cross)
//       EPI_HEAD      dot += sum_n max(acc + b, 0) * w_out[n]     (last MLP layer + the 1-wide output layer)
//       EPI_CROSS_DOT dot += sum_n (X0 * (acc + b) + A)[n] * w_out[n]   (last cross layer: the x_L part of the head)
//     after a dot layer:  scores[m] = sigmoid(dot + b_out + wide[m] + extra_in[m])   or   extra_out[m] = dot
//
// Why one kernel per chain.  Measured with cycle counters on B200: a single 18944 x 1024 x 640
// layer spends ~15 us in its MMA loop but ~29 us from event to event -- launch, prologue (barrier
// init, TMEM allocation, cluster sync, descriptor fetch), pipeline ramp and the last epilogue cost
// 11-14 us per launch, three times per chunk.  Here a CTA pair owns a 256-row tile of candidates
// and walks it through every layer of the chain; the intermediate activations of that tile go to
// a small per-pair scratch that never leaves L2, the accumulators stay double-buffered across
// layer boundaries, and the W tiles of the next layer are already in flight while the previous
// layer's epilogue drains.  A pair keeps TWO tiles in flight and alternates between them layer by
// layer: a layer can only start when the previous layer's last epilogue has stored its rows, and
// while one tile sits at that boundary the tensor cores work on the other one.
//
// Structure (CTAs paired into clusters of 2 = one TPC, persistent, 384 threads per CTA):
//   warp 0     TMA producer: cp.async.bulk.tensor 2-D loads (128-byte swizzle) of this CTA's 128
//              rows of A and its HALF of the W tile into a shared-memory ring (mbarrier paced);
//              both CTAs' loads complete on the leader's barrier
//   warp 1     MMA issuer (leader CTA): one thread issues tcgen05.mma.cta_group::2 (M=256, N=BN,
//              K=16, f16 -> f32), accumulators in the TMEM of both SMs; tcgen05.commit multicasts
//              "stage free" / "accumulator ready" to both CTAs
//   warp 2     TMEM allocator (2 accumulators x BN columns: tile i's epilogue overlaps tile i+1's MMAs)
//   warps 4-11 epilogue: tcgen05.ld -> registers -> bias / activation -> 16-byte stores straight from
//              registers into the TILE-MAJOR scratch (below)
//
// Tile-major scratch.  In TMEM a thread owns a ROW of the accumulator, so writing activations
// row-major costs 32 L1 wavefronts per store instruction (the first version was epilogue-bound: 7.5k
// cycles per tile against 5.1k of MMA), and DCN-v2's cross epilogue, which also READS x0 and x_l by
// row, measured ~12k.  Activations between layers are therefore kept in the layout the tensor core
// reads un-swizzled K-major operands in: per 128-row tile, per group of 8 columns, 128 rows x 16 B
// contiguous -- [tile][col / 8][row][8 halfs].  A warp's store (32 rows x 16 B of one column group)
// is one contiguous 512-byte segment; a 64-column K block of a tile is one contiguous 16 KB box for
// the TMA (3-D map, no swizzle) that lands in shared memory exactly as the MMA descriptor wants it
// (8 x 16 B core matrices, 128 B between 8-row groups, 2048 B between column groups); and the cross
// epilogue reads x0 / x_l back with the same coalesced accesses.  Only the first layer's A (the
// gathered X, written a row per warp by the gather kernel) is row-major with the 128-byte swizzle.
// Every mbarrier wait is bounded (2 s, then trap): a descriptor or phase bug becomes a CUDA error.
//
// Roofline: tensor pipe.  Per K block a CTA pulls 16 KB (A) + BN/2 * 128 B (W half) from L2
// -- with 1-CTA 128x256 tiles (48 KB per K block) the kernel measured ~10 TB/s of L2->SM
// traffic, i.e. it was L2-bandwidth bound.
#pragma once
#include "common.cuh"

namespace ctr {

constexpr int kGemmThreads = 384;       // 4 control warps + 8 epilogue warps
constexpr int kBM = 128;                // rows per CTA; a pair covers 2 * kBM
constexpr int kBK = 64;                 // 64 halfs = one 128-byte swizzle span
constexpr int kMaxChain = 4;
// Strides of the tile-major A operand in shared memory (bring-up switch: -DCTR_A_LBO=128 -DCTR_A_SBO=2048
// swaps the two descriptor fields).
#ifndef CTR_A_LBO
#define CTR_A_LBO 2048     // bytes from one 8-column (16-byte) group to the next
#define CTR_A_SBO 128      // bytes from one 8-row core matrix to the next
#endif

enum GemmEpilogue { EPI_RELU = 0, EPI_CROSS = 1, EPI_HEAD = 2, EPI_CROSS_DOT = 3 };

struct ChainLayer {
    int32_t N, K;              // padded output / reduction widths (N multiple of BN, K of 64)
    int32_t n_valid;           // columns below this exist in X0 / A (EPI_CROSS* must not read past their rows)
    int32_t epi;
    int32_t a_global;          // 1: A is row-major [M, lda] addressed by candidate row (first layer); 0: tile-major pair scratch
    const float* bias;         // [N]
    const __half* a_ptr;       // the A operand as a plain pointer (EPI_CROSS* reads x_l from it)
    int32_t lda;               // row stride (row-major) or column count (tile-major) of A
    __half* out;               // tile-major pair scratch, ldo columns (EPI_RELU / EPI_CROSS)
    int32_t ldo;
    int32_t out_global;        // test hook: store row-major by candidate row instead (slow path)
    const float* w_out;        // [N] (dot epilogues)
};

struct ChainParams {
    int32_t M;                 // candidates in this launch
    int32_t n_layers;
    ChainLayer layer[kMaxChain];
    const __half* x0;          // [M, ldx0] gathered activations, row-major (EPI_CROSS* of the first layer)
    int32_t ldx0;
    __half* x0e;               // tile-major copy of x0 in the pair scratch, written by the first cross layer's
    int32_t ldx0e;             //   epilogue and read by the later ones (nullptr: single-layer chain)
    const float* wide;         // [M] or nullptr
    const float* extra_in;     // [M] or nullptr
    float b_out;
    float* scores;             // [M]: sigmoid(dot + ...) after a dot layer, or nullptr
    float* extra_out;          // [M]: raw dot after a dot layer, or nullptr
    long long* prof;           // bring-up instrumentation (nullptr in production)
    // The pair scratch is ONE allocation per device, shared by every stream: a CTA pair takes a free slice from
    // this bitmap when it starts and gives it back when it leaves (at most num_sms / 2 pairs are ever resident,
    // one per TPC, whatever the number of streams).  nullptr: slice = pair index of this launch (the caller
    // guarantees that no other launch uses the same scratch at the same time).
    unsigned long long* slice_bitmap;
    int32_t n_slices;
    int32_t tiles_in_flight;   // 2 (default): a pair alternates between two tiles layer by layer; 1: one tile, half the scratch
};

struct ChainMaps {
    CUtensorMap a[kMaxChain];
    CUtensorMap w[kMaxChain];
};

template <int BN>
struct ChainCfg {
    static constexpr int kABytes = kBM * kBK * 2;                 // 16 KB: this CTA's 128 rows
    static constexpr int kBBytes = (BN / 2) * kBK * 2;            // 16 KB / 8 KB: this CTA's half of the W tile
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = BN == 256 ? 6 : 8;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

// Offset (in halfs) of the 16-byte unit (row r of 128, column group g) of tile `tile` in a tile-major
// buffer of `ld` columns.
__device__ __forceinline__ size_t tile_major_off(int tile, int ld, int g, int r) {
    return (size_t)tile * (size_t)(kBM * ld) + (size_t)g * (kBM * 8) + (size_t)r * 8;
}

// x0 and x_l pieces (4 column groups x 16 B each) of one 32-column chunk of a cross epilogue
struct CrossRegs { uint4 a[4], b[4]; };

// Issues the loads of one chunk's cross operands.  They do not depend on the accumulator, so the caller
// has them in flight one chunk ahead (also across N tiles): with 8 epilogue warps per SM nothing else
// hides their L2 round trip.
__device__ __forceinline__ void load_cross(const ChainParams& p, const ChainLayer& L, bool row_ok, int grow, int stile, int r,
                                           int col0, CrossRegs& x) {
    if (!row_ok || col0 >= L.n_valid) return;                  // same predicate as the consumer's
    if (L.a_global) {                                          // first cross layer: x_l = x0, row-major (gather output)
        const uint4* x0p = reinterpret_cast<const uint4*>(p.x0 + (size_t)grow * p.ldx0 + col0);
#pragma unroll
        for (int j = 0; j < 4; ++j) x.a[j] = __ldg(x0p + j);
    } else {                                                   // written by this very thread one layer ago: plain loads
        const __half* x0p = p.x0e + tile_major_off(stile, p.ldx0e, col0 >> 3, r);
        const __half* xlp = L.a_ptr + tile_major_off(stile, L.lda, col0 >> 3, r);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            x.a[j] = *reinterpret_cast<const uint4*>(x0p + j * (kBM * 8));
            x.b[j] = *reinterpret_cast<const uint4*>(xlp + j * (kBM * 8));
        }
    }
}

// Drains kChunks x 32 columns (from column n0) of a 128-row accumulator tile for the calling warp's 32 rows.
//   grow: candidate row of this thread (guards, row-major x0, wide); stile / r: its 128-row scratch
//   tile and its row in it.  Cross epilogues: x[0] holds the first chunk's operands on entry and, if
//   next_n0 >= 0, the first chunk's of the tile starting at column next_n0 on exit.
template <int kChunks>
__device__ __forceinline__ void epilogue_tile(const ChainParams& p, const ChainLayer& L, uint32_t taddr, int grow,
                                              int stile, int r, int n0, int next_n0, CrossRegs (&x)[2], float& dot) {
    static_assert(kChunks % 2 == 0, "the operand double buffer returns to slot 0 at a tile boundary");
    const int epi = L.epi;
    const bool cross = epi == EPI_CROSS || epi == EPI_CROSS_DOT;
    const bool dot_epi = epi == EPI_HEAD || epi == EPI_CROSS_DOT;
    const bool row_ok = grow < p.M;
    uint32_t rr[32];
    tmem_ld32(taddr, rr);
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
        const int col0 = n0 + c * 32;
        if (cross) {
            if (c + 1 < kChunks) load_cross(p, L, row_ok, grow, stile, r, col0 + 32, x[(c + 1) & 1]);
            else if (next_n0 >= 0) load_cross(p, L, row_ok, grow, stile, r, next_n0, x[0]);
        }
        tmem_ld_wait();                                        // chunk c has landed
        float v[32];
        const float4* bias4 = reinterpret_cast<const float4*>(L.bias + col0);   // same address in every lane: broadcast
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 bv = __ldg(bias4 + j);
            v[4 * j + 0] = __uint_as_float(rr[4 * j + 0]) + bv.x;
            v[4 * j + 1] = __uint_as_float(rr[4 * j + 1]) + bv.y;
            v[4 * j + 2] = __uint_as_float(rr[4 * j + 2]) + bv.z;
            v[4 * j + 3] = __uint_as_float(rr[4 * j + 3]) + bv.w;
        }
        if (c + 1 < kChunks) tmem_ld32(taddr + (c + 1) * 32, rr);   // rr is free again: the load overlaps the rest of this chunk
        if (!cross) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        } else if (row_ok && col0 < L.n_valid) {               // x0 * (acc + b) + x_l (padding columns stay acc + b = 0)
            const CrossRegs& xc = x[c & 1];
            if (L.a_global && p.x0e) {                         // keep x0 where the later layers read it coalesced
                __half* x0w = p.x0e + tile_major_off(stile, p.ldx0e, col0 >> 3, r);
#pragma unroll
                for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(x0w + j * (kBM * 8)) = xc.a[j];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 xl = L.a_global ? xc.a[j] : xc.b[j];
                const __half2* ah = reinterpret_cast<const __half2*>(&xc.a[j]);
                const __half2* bh = reinterpret_cast<const __half2*>(&xl);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 af = __half22float2(ah[e]), bf = __half22float2(bh[e]);
                    v[8 * j + 2 * e + 0] = fmaf(af.x, v[8 * j + 2 * e + 0], bf.x);
                    v[8 * j + 2 * e + 1] = fmaf(af.y, v[8 * j + 2 * e + 1], bf.y);
                }
            }
        }
        if (dot_epi) {
            const float4* w4 = reinterpret_cast<const float4*>(L.w_out + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 wv = __ldg(w4 + j);
                dot = fmaf(v[4 * j + 0], wv.x, dot);
                dot = fmaf(v[4 * j + 1], wv.y, dot);
                dot = fmaf(v[4 * j + 2], wv.z, dot);
                dot = fmaf(v[4 * j + 3], wv.w, dot);
            }
        } else if (row_ok) {
            // this thread's 32 columns = 4 column groups x 16 B; the warp's 32 rows of a group are contiguous
            __half* o = L.out_global ? L.out + (size_t)grow * L.ldo + col0 : L.out + tile_major_off(stile, L.ldo, col0 >> 3, r);
            const int step = L.out_global ? 8 : kBM * 8;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint4 h;                                             // saturating: a layer output beyond the half range is clamped, never inf
                h.x = pack_half2_sat(v[8 * j + 0], v[8 * j + 1]); h.y = pack_half2_sat(v[8 * j + 2], v[8 * j + 3]);
                h.z = pack_half2_sat(v[8 * j + 4], v[8 * j + 5]); h.w = pack_half2_sat(v[8 * j + 6], v[8 * j + 7]);
                *reinterpret_cast<uint4*>(o + j * step) = h;
            }
        }
    }
}

// Register budget.  The kernel is launched with 128 registers per thread (12 warps x 32 x 128 = 48 K of the SM's
// 64 K), then the four control warps shrink to 48 and the eight epilogue warps grow to 168 (setmaxnreg, one
// instruction per warp group): 4 x 32 x 48 + 8 x 32 x 168 = 48 K.  The 16 K registers and ~30 KB of shared
// memory this leaves on every SM are what one 256-thread CTA of the gather kernel (gather.cuh, 64 registers)
// needs, so the next chunk's gather runs on the same SMs UNDER this kernel instead of in front of it.
constexpr int kTowerLaunchRegThreads = 512;    // __launch_bounds__ max-threads that makes ptxas allocate 128 registers
template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTowerLaunchRegThreads, 1)
tower_chain_kernel(const __grid_constant__ ChainMaps maps, const __grid_constant__ ChainParams p) {
    using Cfg = ChainCfg<BN>;
    constexpr int kStages = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    // 128-byte swizzle needs 1024-byte aligned stage buffers
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kStages * Cfg::kABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
    uint64_t* full_bar = bars;                     // [kStages] used in the leader: both CTAs' TMA -> MMA
    uint64_t* empty_bar = bars + kStages;          // [kStages] in each CTA: MMA -> that CTA's producer
    uint64_t* acc_full = bars + 2 * kStages;       // [2] in each CTA: MMA -> that CTA's epilogue
    uint64_t* acc_empty = bars + 2 * kStages + 2;  // [2] used in the leader: both CTAs' epilogues -> MMA
    uint64_t* layer_done = bars + 2 * kStages + 4; // [2] in each CTA, one per interleaved tile: epilogue (layer l stored) -> producer (layer l+1 A loads)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 6);
    __shared__ float dot_part[2 * kBM];            // high-column-half partial dots, double-buffered
    __shared__ int slice_sh;                       // this pair's slice of the activation scratch

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int tiles_m = (p.M + 2 * kBM - 1) / (2 * kBM);            // 256-row tiles of candidates
    const int nl = p.n_layers;
    const int tif = p.tiles_in_flight == 1 ? 1 : 2;                  // 256-row tiles a pair alternates between

    if (warp == 0 && lane == 0) {
        for (int l = 0; l < nl; ++l) { tma_prefetch_desc(&maps.a[l]); tma_prefetch_desc(&maps.w[l]); }
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 2); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 16); }
        mbar_init(&layer_done[0], 8);
        mbar_init(&layer_done[1], 8);
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
        tmem_relinquish_pair();
    }
    if (warp == 3 && lane == 0 && leader) {
        // take a free scratch slice (the search starts at a different bit in every pair, so the pairs of a
        // launch rarely collide), publish it in both CTAs; the cluster barrier below makes it visible
        int slice = pair_id;
        if (p.slice_bitmap) {
            slice = -1;
            const uint64_t t0 = globaltimer_ns();
            for (int probe = pair_id % p.n_slices; slice < 0; probe = probe + 1 == p.n_slices ? 0 : probe + 1) {
                const unsigned long long bit = 1ull << (probe & 63);
                if (!(atomicOr(&p.slice_bitmap[probe >> 6], bit) & bit)) slice = probe;
                else if (probe == 0 && globaltimer_ns() - t0 > 2000000000ull) __trap();   // leaked slices: an earlier launch died
            }
        }
        slice_sh = slice;
        asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(map_to_cta(smem_u32(&slice_sh), 1)), "r"(slice) : "memory");
    }
    tc_fence_before();
    cluster_sync_all();                              // barriers of both CTAs are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int slice = slice_sh;

    if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");       // one instruction for the whole warp group
    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        // The whole warp walks the loop (so addresses and coordinates stay in uniform registers) and one
        // elected lane issues.
        {
            int s = 0; uint32_t ph = 0, ld_ph = 0;
            long long w_empty = 0, w_layer = 0;
            for (int t0 = pair_id; t0 < tiles_m; t0 += tif * n_pairs) {
                const int n_sub = (tif == 2 && t0 + n_pairs < tiles_m) ? 2 : 1;   // two tiles in flight, layer by layer
                for (int l = 0; l < nl; ++l) {
                  for (int sub = 0; sub < n_sub; ++sub) {
                    const ChainLayer& L = p.layer[l];
                    const int t = t0 + sub * n_pairs;
                    const int a_row = t * 2 * kBM + (int)rank * kBM;          // row-major A: candidate rows
                    const int a_tile = (slice * tif + sub) * 2 + (int)rank;     // tile-major A: this CTA's scratch tile
                    const int tiles_n = (L.N + BN - 1) / BN, k_blocks = L.K / kBK;
                    for (int nt = 0; nt < tiles_n; ++nt) {
                        // the last tile of a layer may be 128 wide: each CTA then feeds the first 64 rows of its box
                        const int w_row = nt * BN + (int)rank * (min(BN, L.N - nt * BN) / 2);
                        for (int kb = 0; kb < k_blocks; ++kb) {
                            long long c0 = p.prof ? clock64() : 0;
                            mbar_wait(&empty_bar[s], ph ^ 1);
                            if (p.prof) w_empty += clock64() - c0;
                            const uint32_t lead_full = map_to_cta(smem_u32(&full_bar[s]), 0);
                            const bool issuer = elect_one();
                            if (issuer) {
                                if (leader) mbar_expect_tx(&full_bar[s], 2 * Cfg::kStageBytes);  // bytes of both CTAs
                                tma_load_2d_pair(smem_b + s * Cfg::kBBytes, &maps.w[l], kb * kBK, w_row, lead_full);
                            }
                            if (l > 0 && nt == 0 && kb == 0) {
                                // A of this layer is what this CTA's epilogue stored for the previous layer:
                                // wait until all 8 epilogue warps have published it (the W tile above is
                                // already in flight)
                                c0 = p.prof ? clock64() : 0;
                                mbar_wait(&layer_done[sub], (ld_ph >> sub) & 1u);
                                ld_ph ^= 1u << sub;
                                if (p.prof) w_layer += clock64() - c0;
                            }
                            if (issuer) {
                                if (L.a_global) tma_load_2d_pair(smem_a + s * Cfg::kABytes, &maps.a[l], kb * kBK, a_row, lead_full);
                                else tma_load_3d_pair(smem_a + s * Cfg::kABytes, &maps.a[l], 0, kb * 32, a_tile, lead_full);   // 16 KB contiguous
                                if (!leader) mbar_arrive_cluster(lead_full);
                            }
                            __syncwarp();
                            if (++s == kStages) { s = 0; ph ^= 1; }
                        }
                    }
                  }
                }
            }
            if (p.prof && lane == 0) { p.prof[blockIdx.x * 8 + 0] = w_empty; p.prof[blockIdx.x * 8 + 7] = w_layer; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        // Warp-uniform loop, one elected lane issues: issued from inside a single-lane branch the
        // descriptors sit in per-thread registers and every tcgen05.mma pays a ~30-instruction
        // move-to-uniform-register loop, which leaves the issuer no run-ahead over the tensor pipe.
        if (leader) {
            int s = 0; uint32_t ph = 0;
            int lt = 0;
            long long w_acc = 0, w_full = 0, t_issue = 0;
            const long long k0 = p.prof ? clock64() : 0;
            const uint64_t ns0 = p.prof ? globaltimer_ns() : 0;
            for (int t0 = pair_id; t0 < tiles_m; t0 += tif * n_pairs) {
                const int n_sub = (tif == 2 && t0 + n_pairs < tiles_m) ? 2 : 1;
                for (int l = 0; l < nl; ++l) {
                    const int tiles_l = (p.layer[l].N + BN - 1) / BN, k_blocks = p.layer[l].K / kBK;
                    const int tiles_n = tiles_l * n_sub;                                            // same order as the producer
                    const bool a_rowmajor = p.layer[l].a_global != 0;
                    for (int nt = 0; nt < tiles_n; ++nt, ++lt) {
                        const int acc = lt & 1;
                        const uint32_t idesc = make_idesc_f16(2 * kBM, min(BN, p.layer[l].N - (nt % tiles_l) * BN), /*f16*/ 0);
                        long long c0 = p.prof ? clock64() : 0;
                        mbar_wait(&acc_empty[acc], ((lt >> 1) & 1) ^ 1);      // both CTAs' epilogues drained it
                        if (p.prof) w_acc += clock64() - c0;
                        tc_fence_after();
                        const uint32_t d_tmem = tmem_base + acc * BN;
                        for (int kb = 0; kb < k_blocks; ++kb) {
                            c0 = p.prof ? clock64() : 0;
                            mbar_wait(&full_bar[s], ph);
                            const long long c1 = p.prof ? clock64() : 0;
                            tc_fence_after();
                            const uint32_t a_addr = smem_u32(smem_a + s * Cfg::kABytes);
                            const uint32_t b_addr = smem_u32(smem_b + s * Cfg::kBBytes);
                            // advancing K by 16 halfs = 32 bytes inside the 128-byte swizzle span; in the
                            // tile-major A it is two column groups of 2048 B (start-address field: 16-byte units)
                            const uint64_t da0 = a_rowmajor ? make_sw128_kmajor_desc(a_addr) : make_interleaved_kmajor_desc(a_addr, CTR_A_LBO, CTR_A_SBO);
                            const uint64_t da_step = a_rowmajor ? (32 >> 4) : (4096 >> 4);
                            const uint64_t db0 = make_sw128_kmajor_desc(b_addr);
                            if (elect_one()) {
#pragma unroll
                                for (int k = 0; k < kBK / 16; ++k)
                                    umma_f16_pair(d_tmem, da0 + k * da_step, db0 + k * (32 >> 4), idesc, (kb | k) != 0 ? 1u : 0u);
                                umma_commit_pair(&empty_bar[s]);              // stage free in both CTAs
                            }
                            __syncwarp();
                            if (p.prof) { w_full += c1 - c0; t_issue += clock64() - c1; }
                            if (++s == kStages) { s = 0; ph ^= 1; }
                        }
                        if (elect_one()) umma_commit_pair(&acc_full[acc]);    // accumulator ready in both CTAs
                        __syncwarp();
                    }
                }
            }
            if (p.prof && lane == 0) {
                p.prof[blockIdx.x * 8 + 1] = w_acc; p.prof[blockIdx.x * 8 + 2] = w_full; p.prof[blockIdx.x * 8 + 3] = t_issue;
                p.prof[blockIdx.x * 8 + 4] = clock64() - k0;
                p.prof[(blockIdx.x + 1) * 8 + 4] = (long long)(globaltimer_ns() - ns0);     // peer's slot: wall time of the same loop
            }
        }
    }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 168;");
        // ===================== epilogue (8 warps, both CTAs, own 128 rows) =====================
        // A warp may only touch the TMEM lanes of its sub-partition (warp % 4); two warps share a
        // sub-partition and split the tile's columns in halves.
        const int sub = warp & 3;
        const int half = (warp - 4) >> 2;
        int lt = 0, dot_round = 0;
        CrossRegs x[2];
        long long w_accfull = 0, t_epi = 0;
        for (int t0 = pair_id; t0 < tiles_m; t0 += tif * n_pairs) {
          const int n_sub = (tif == 2 && t0 + n_pairs < tiles_m) ? 2 : 1;
          for (int l = 0; l < nl; ++l) {
            for (int st = 0; st < n_sub; ++st) {                                     // the two interleaved tiles of this pair
                const int t = t0 + st * n_pairs;
                const int grow = t * 2 * kBM + (int)rank * kBM + sub * 32 + lane;        // this thread's candidate row
                const int stile = (slice * tif + st) * 2 + (int)rank;                     // this CTA's 128-row scratch tile
                const ChainLayer& L = p.layer[l];
                const int tiles_n = (L.N + BN - 1) / BN;
                const bool dot_epi = L.epi == EPI_HEAD || L.epi == EPI_CROSS_DOT;
                const bool cross = L.epi == EPI_CROSS || L.epi == EPI_CROSS_DOT;
                const int r = sub * 32 + lane;
                float dot = 0.f;
                if (cross) load_cross(p, L, grow < p.M, grow, stile, r, half * (min(BN, L.N) / 2), x[0]);
                for (int nt = 0; nt < tiles_n; ++nt, ++lt) {
                    const int acc = lt & 1;
                    const int hc = min(BN, L.N - nt * BN) / 2;                 // columns per warp: 128, or 64 in a narrow last tile
                    const int next_n0 = nt + 1 < tiles_n ? (nt + 1) * BN + half * (min(BN, L.N - (nt + 1) * BN) / 2) : -1;
                    const long long c0 = p.prof ? clock64() : 0;
                    mbar_wait(&acc_full[acc], (lt >> 1) & 1);
                    const long long c1 = p.prof ? clock64() : 0;
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)(sub * 32) << 16) + acc * BN + half * hc;
                    if (BN == 256 && hc == 128) epilogue_tile<4>(p, L, taddr, grow, stile, r, nt * BN + half * hc, next_n0, x, dot);
                    else epilogue_tile<2>(p, L, taddr, grow, stile, r, nt * BN + half * hc, next_n0, x, dot);
                    // TMEM reads of this accumulator are done: hand it back to the MMA warp (in the leader)
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&acc_empty[acc]), 0));
                    if (p.prof) { w_accfull += c1 - c0; t_epi += clock64() - c1; }
                }
                if (dot_epi) {
                    // the two column halves of a row live in two warps: combine in a fixed order (low + high)
                    const int buf = dot_round & 1;
                    ++dot_round;
                    if (half == 1) dot_part[buf * kBM + sub * 32 + lane] = dot;
                    asm volatile("bar.sync %0, 64;" ::"r"(1 + sub) : "memory");
                    if (half == 0 && grow < p.M) {
                        float z = dot + dot_part[buf * kBM + sub * 32 + lane];
                        if (p.extra_out) p.extra_out[grow] = z;
                        if (p.scores) {
                            z += p.b_out;
                            if (p.wide) z += p.wide[grow];
                            if (p.extra_in) z += p.extra_in[grow];
                            p.scores[grow] = sigmoidf_fast(z);
                        }
                    }
                } else if (l + 1 < nl) {
                    // this layer's rows are stored: make them visible to the TMA (async proxy) reads of
                    // the next layer, then tell this CTA's producer
                    asm volatile("fence.proxy.async;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&layer_done[st]);
                }
            }
          }
        }
        if (p.prof && warp == 4 && lane == 0) { p.prof[blockIdx.x * 8 + 5] = w_accfull; p.prof[blockIdx.x * 8 + 6] = t_epi; }
    }

    // neither CTA may leave (or free TMEM) while its partner can still touch its shared memory / TMEM
    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
    }
    if (warp == 3 && lane == 0 && leader && p.slice_bitmap) {     // every access of this pair to its slice is behind the barrier above
        __threadfence();
        atomicAnd(&p.slice_bitmap[slice >> 6], ~(1ull << (slice & 63)));
    }
}

// ------------------------------------------------------------------ small companions
// fp32 [rows, cols] -> half [rows_pad, cols_pad], zero padded (tower weights at load time)
__global__ void convert_pad_f16_kernel(const float* src, int rows, int cols, __half* dst, int rows_pad, int cols_pad) {
    const int64_t total = (int64_t)rows_pad * cols_pad;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / cols_pad), c = (int)(i % cols_pad);
        dst[i] = __float2half_rn((r < rows && c < cols) ? src[(int64_t)r * cols + c] : 0.f);
    }
}

__global__ void half_to_float_kernel(const __half* src, int64_t rows, int cols, int ld, float* dst) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < rows * cols; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = __half2float(src[(i / cols) * ld + (i % cols)]);
}

}  // namespace ctr
