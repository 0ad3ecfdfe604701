// gemm_f16.cuh -- dense-tower layer as a tcgen05 / TMEM / TMA GEMM with a fused epilogue
// (north star (c)).  One layer of the served model's MLP / cross network:
//
//     OUT[m, n] = epi( sum_k A[m, k] * W[n, k] + bias[n] )          A, W: half, K-major
//
//   EPI_RELU : OUT = max(acc + bias, 0)                 -> half    (MLP hidden layer)
//   EPI_CROSS: OUT = X0[m,n] * (acc + bias) + XL[m,n]   -> half    (DCN-v2 cross layer)
//   EPI_HEAD : score[m] = sigmoid( wide[m] + extra[m] + b_out + sum_n max(acc+bias,0) * w_out[n] )
//              (last hidden layer with the 1-wide output layer, the wide/FM add and the
//               sigmoid fused in; needs the whole layer width inside one tile: N == BN)
//
// Structure (one CTA per SM, CTAs paired into clusters of 2, persistent over 256 x BN output
// tiles, 384 threads per CTA):
//   warp 0    TMA producer: cp.async.bulk.tensor 2-D loads (128-byte swizzle) of this CTA's 128
//             rows of A and its half of the W tile into a shared-memory ring, mbarrier paced
//   warp 1    MMA issuer (leader CTA): one thread issues tcgen05.mma.cta_group::2 (M=256, N=BN,
//             K=16, f16 -> f32) four times per stage, accumulating in the TMEM of both SMs;
//             tcgen05.commit frees the stage and, after the last K block, publishes the accumulator
//   warp 2    TMEM allocator (2 accumulators x BN columns, so the epilogue of tile i overlaps
//             the main loop of tile i+1)
//   warps 4-11 epilogue: tcgen05.ld -> registers -> bias/activation -> staged, coalesced stores;
//             two warps per TMEM sub-partition split the tile's columns
//
// Roofline: tensor pipe.  2*M*N*K flop per launch against the measured dense f16 rate.
#pragma once
#include "common.cuh"

namespace ctr {

constexpr int kGemmThreads = 384;             // 4 control warps + 8 epilogue warps
constexpr int kBM = 128;
constexpr int kBK = 64;                 // 64 halfs = one 128-byte swizzle span

enum GemmEpilogue { EPI_RELU = 0, EPI_CROSS = 1, EPI_HEAD = 2 };

struct GemmParams {
    int32_t M;                 // valid rows (candidates)
    int32_t N;                 // padded output width (multiple of BN)
    int32_t K;                 // padded reduction width (multiple of 64)
    const float* bias;         // [N]
    __half* out;               // [M_cap, ldo]
    int32_t ldo;
    // EPI_CROSS
    const __half* x0;          // [M_cap, ldx]
    const __half* xl;
    int32_t ldx;
    // EPI_HEAD
    const float* w_out;        // [N]
    const float* wide;         // [M] or nullptr
    const float* extra;        // [M] or nullptr (DCN-v2: w_out . x_L part, computed by the head kernel)
    float b_out;
    float* scores;             // [M]
    // bring-up instrumentation (nullptr in production): per-CTA cycle counters, 8 x int64 per CTA
    long long* prof;
};

// Drains one half (BN/2 columns) of a 128-row accumulator tile for the calling warp's 32 rows:
// tcgen05.ld 32 columns at a time (the next chunk's load in flight while this one is processed),
// bias + activation in fp32, then either the running dot product with the output layer's weights
// (EPI_HEAD) or half-precision stores.  In TMEM a thread owns a ROW, so storing straight from
// registers makes every 16-byte store instruction touch 32 different lines (32 L1 wavefronts; the
// first version of this kernel spent 7.5k cycles per tile in here against 5.1k of MMA).  Instead
// two chunks (64 columns = 128 bytes per row) go through a warp-private, XOR-swizzled 4 KB staging
// block in shared memory and are written back row-major: one store instruction = 4 rows x 128 B.
template <int BN, int EPI>
__device__ __forceinline__ void epilogue_half_tile(const GemmParams& p, uint32_t taddr, int row_base, int lane, int n0,
                                                   float& head_acc, uint4* stage) {
    constexpr int kChunks = (BN / 2) / 32;
    static_assert(kChunks % 2 == 0, "two 32-column chunks per staging round");
    const int row = row_base + lane;
    uint32_t r[2][32];
    tmem_ld32(taddr, r[0]);
#pragma unroll
    for (int c = 0; c < kChunks; ++c) {
        const int col0 = n0 + c * 32;
        // bias (and head weights) for this chunk: same address in every lane -> broadcast, L1-resident
        float4 bv[8];
        const float4* bias4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
        for (int j = 0; j < 8; ++j) bv[j] = __ldg(bias4 + j);
        tmem_ld_wait();                                        // chunk c has landed in r[c & 1]
        if (c + 1 < kChunks) tmem_ld32(taddr + (c + 1) * 32, r[(c + 1) & 1]);   // overlaps the math below
        const uint32_t* rc = r[c & 1];
        if (EPI == EPI_HEAD) {
            const float4* w4 = reinterpret_cast<const float4*>(p.w_out + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 wv = __ldg(w4 + j);
                head_acc = fmaf(fmaxf(__uint_as_float(rc[4 * j + 0]) + bv[j].x, 0.f), wv.x, head_acc);
                head_acc = fmaf(fmaxf(__uint_as_float(rc[4 * j + 1]) + bv[j].y, 0.f), wv.y, head_acc);
                head_acc = fmaf(fmaxf(__uint_as_float(rc[4 * j + 2]) + bv[j].z, 0.f), wv.z, head_acc);
                head_acc = fmaf(fmaxf(__uint_as_float(rc[4 * j + 3]) + bv[j].w, 0.f), wv.w, head_acc);
            }
        } else {
            float v[32];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[4 * j + 0] = __uint_as_float(rc[4 * j + 0]) + bv[j].x;
                v[4 * j + 1] = __uint_as_float(rc[4 * j + 1]) + bv[j].y;
                v[4 * j + 2] = __uint_as_float(rc[4 * j + 2]) + bv[j].z;
                v[4 * j + 3] = __uint_as_float(rc[4 * j + 3]) + bv[j].w;
            }
            if (EPI == EPI_RELU) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
            } else if (row < p.M) {   // EPI_CROSS: out = x0 * (acc + b) + x_l
                const uint4* x0p = reinterpret_cast<const uint4*>(p.x0 + (size_t)row * p.ldx + col0);
                const uint4* xlp = reinterpret_cast<const uint4*>(p.xl + (size_t)row * p.ldx + col0);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint4 a = __ldg(x0p + j), b = __ldg(xlp + j);
                    const __half2* ah = reinterpret_cast<const __half2*>(&a);
                    const __half2* bh = reinterpret_cast<const __half2*>(&b);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 af = __half22float2(ah[e]), bf = __half22float2(bh[e]);
                        v[8 * j + 2 * e + 0] = fmaf(af.x, v[8 * j + 2 * e + 0], bf.x);
                        v[8 * j + 2 * e + 1] = fmaf(af.y, v[8 * j + 2 * e + 1], bf.y);
                    }
                }
            }
            // this thread's 32 columns -> 4 x 16 B into its row of the staging block (swizzled: conflict-free)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                __half2 h[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[8 * j + 2 * e], v[8 * j + 2 * e + 1]);
                const int lc = (c & 1) * 4 + j;
                stage[lane * 8 + (lc ^ (lane & 7))] = *reinterpret_cast<const uint4*>(h);
            }
            if (c & 1) {
                __syncwarp();
                const int colbase = n0 + (c - 1) * 32;               // first of the 64 staged columns
                const int cc = lane & 7;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int rr = it * 4 + (lane >> 3);
                    const uint4 val = stage[rr * 8 + (cc ^ (rr & 7))];
                    if (row_base + rr < p.M)
                        *reinterpret_cast<uint4*>(p.out + (size_t)(row_base + rr) * p.ldo + colbase + cc * 8) = val;
                }
                __syncwarp();
            }
        }
    }
}

// ================================================================================================
// CTA-pair version (cluster of 2, tcgen05 cta_group::2): one MMA spans both SMs of a TPC.
// The pair computes a 256 x BN tile: CTA r holds rows [128r, 128r+128) of A and of the accumulator,
// and HALF of the W tile (BN/2 rows), which the tensor cores of both SMs read.  Per K block a CTA
// therefore pulls 16 KB (A) + BN/2*128 B (W half) from L2 instead of 16 KB + BN*128 B: the 1-CTA
// kernel measured ~10 TB/s of L2->SM traffic, i.e. it was bound by L2 bandwidth, not by the tensor
// pipe.  Smaller stages also make the ring deeper (6 x 32 KB for BN=256).
//   both CTAs   warp 0 TMA producer (own A rows, own W half; completes on the LEADER's full barrier)
//   leader CTA  warp 1 MMA issuer (tcgen05.mma.cta_group::2, M=256); tcgen05.commit multicasts to the
//               barriers of both CTAs (stage free / accumulator ready)
//   both CTAs   warps 4-11 epilogue for their own 128 rows; "accumulator drained" arrives on the
//               leader's barrier from both CTAs
template <int BN>
struct GemmPairCfg {
    static constexpr int kABytes = kBM * kBK * 2;                 // 16 KB: this CTA's 128 rows
    static constexpr int kBBytes = (BN / 2) * kBK * 2;            // 16 KB / 8 KB: this CTA's half of W
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = BN == 256 ? 5 : 6;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kStoreStageBytes = 8 * 4096;             // 8 epilogue warps x (32 rows x 128 B)
    static constexpr int kSmemBytes = kStages * kStageBytes + kStoreStageBytes + 1024 + 256;
};

template <int BN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_f16_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w, const GemmParams p) {
    using Cfg = GemmPairCfg<BN>;
    constexpr int kStages = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kStages * Cfg::kABytes;
    uint8_t* store_stage = smem + kStages * Cfg::kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(store_stage + Cfg::kStoreStageBytes);
    uint64_t* full_bar = bars;                     // [kStages] used in the leader: both CTAs' TMA -> MMA
    uint64_t* empty_bar = bars + kStages;          // [kStages] in each CTA: MMA -> that CTA's producer
    uint64_t* acc_full = bars + 2 * kStages;       // [2] in each CTA: MMA -> that CTA's epilogue
    uint64_t* acc_empty = bars + 2 * kStages + 2;  // [2] used in the leader: both CTAs' epilogues -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
    __shared__ float head_part[2 * kBM];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int tiles_m = (p.M + 2 * kBM - 1) / (2 * kBM);            // 256-row tiles
    const int tiles_n = p.N / BN;
    const int n_tiles = tiles_m * tiles_n;
    const int k_blocks = p.K / kBK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_w);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 2); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 16); }
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
        tmem_relinquish_pair();
    }
    tc_fence_before();
    cluster_sync_all();                              // barriers of both CTAs are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            long long w_empty = 0;
            for (int t = pair_id; t < n_tiles; t += n_pairs) {
                const int m0 = (t / tiles_n) * 2 * kBM + (int)rank * kBM;
                const int n0 = (t % tiles_n) * BN + (int)rank * (BN / 2);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    const long long c0 = p.prof ? clock64() : 0;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    if (p.prof) w_empty += clock64() - c0;
                    const uint32_t lead_full = map_to_cta(smem_u32(&full_bar[s]), 0);
                    if (leader) mbar_expect_tx(&full_bar[s], 2 * Cfg::kStageBytes);      // bytes of both CTAs
                    tma_load_2d_pair(smem_a + s * Cfg::kABytes, &map_a, kb * kBK, m0, lead_full);
                    tma_load_2d_pair(smem_b + s * Cfg::kBBytes, &map_w, kb * kBK, n0, lead_full);
                    if (!leader) mbar_arrive_cluster(lead_full);
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
            }
            if (p.prof) p.prof[blockIdx.x * 8 + 0] = w_empty;
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (leader && lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(2 * kBM, BN, /*f16*/ 0);
            int s = 0; uint32_t ph = 0;
            int lt = 0;
            long long w_acc = 0, w_full = 0, t_issue = 0;
            const long long k0 = p.prof ? clock64() : 0;
            for (int t = pair_id; t < n_tiles; t += n_pairs, ++lt) {
                const int acc = lt & 1;
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
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k) {
                        const uint64_t da = make_sw128_kmajor_desc(a_addr + k * 32);
                        const uint64_t db = make_sw128_kmajor_desc(b_addr + k * 32);
                        umma_f16_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit_pair(&empty_bar[s]);                  // stage free in both CTAs
                    if (p.prof) { w_full += c1 - c0; t_issue += clock64() - c1; }
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
                umma_commit_pair(&acc_full[acc]);                     // accumulator ready in both CTAs
            }
            if (p.prof) {
                p.prof[blockIdx.x * 8 + 1] = w_acc; p.prof[blockIdx.x * 8 + 2] = w_full; p.prof[blockIdx.x * 8 + 3] = t_issue;
                p.prof[blockIdx.x * 8 + 4] = clock64() - k0;
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue (8 warps, both CTAs, own 128 rows) =====================
        const int sub = warp & 3;
        const int half = (warp - 4) >> 2;
        constexpr int kHalfCols = BN / 2;
        int lt = 0;
        long long w_accfull = 0, t_epi = 0;
        for (int t = pair_id; t < n_tiles; t += n_pairs, ++lt) {
            const int acc = lt & 1;
            const int m0 = (t / tiles_n) * 2 * kBM + (int)rank * kBM, n0 = (t % tiles_n) * BN + half * kHalfCols;
            const int row_base = m0 + sub * 32, row = row_base + lane;
            const bool row_ok = row < p.M;
            const long long c0 = p.prof ? clock64() : 0;
            mbar_wait(&acc_full[acc], (lt >> 1) & 1);
            const long long c1 = p.prof ? clock64() : 0;
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(sub * 32) << 16) + acc * BN + half * kHalfCols;
            float head_acc = 0.f;
            epilogue_half_tile<BN, EPI>(p, taddr, row_base, lane, n0, head_acc, reinterpret_cast<uint4*>(store_stage + (warp - 4) * 4096));
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&acc_empty[acc]), 0));
            if (p.prof) { w_accfull += c1 - c0; t_epi += clock64() - c1; }
            if (EPI == EPI_HEAD) {
                if (half == 1) head_part[acc * kBM + sub * 32 + lane] = head_acc;
                asm volatile("bar.sync %0, 64;" ::"r"(1 + sub) : "memory");
                if (half == 0 && row_ok) {
                    float z = head_acc + head_part[acc * kBM + sub * 32 + lane] + p.b_out;
                    if (p.wide) z += p.wide[row];
                    if (p.extra) z += p.extra[row];
                    p.scores[row] = sigmoidf_fast(z);
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

// Generic head: one warp per candidate.
//   z = wide[b] + b_out + sum_j H[b,j]*w_h[j]  (+ sum_j XL[b,j]*w_x[j] for DCN-v2)
//   write_scores: scores[b] = sigmoid(z)   else: extra_out[b] = z   (partial sums for EPI_HEAD)
struct HeadParams {
    int64_t n;
    const __half* H; int32_t ldh; int32_t nh; const float* w_h;     // may be nullptr / 0
    const __half* XL; int32_t ldxl; int32_t nx; const float* w_x;   // may be nullptr / 0
    const float* wide;                                              // may be nullptr
    float b_out;
    float* scores;      // sigmoid output, or
    float* extra_out;   // raw partial logit
};
__global__ void __launch_bounds__(256) head_kernel(const HeadParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t b = warp_global; b < p.n; b += n_warps) {
        float acc = 0.f;
        for (int part = 0; part < 2; ++part) {
            const __half* v = part == 0 ? p.H : p.XL;
            const int len = part == 0 ? p.nh : p.nx;
            const float* w = part == 0 ? p.w_h : p.w_x;
            const int ld = part == 0 ? p.ldh : p.ldxl;
            if (!v || len == 0) continue;
            const __half* rowp = v + b * ld;
            for (int j = lane * 8; j < len; j += 256) {          // len is a multiple of 8
                const uint4 hv = __ldg(reinterpret_cast<const uint4*>(rowp + j));
                const __half2* hh = reinterpret_cast<const __half2*>(&hv);
                const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + j));
                const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + j + 4));
                float2 f;
                f = __half22float2(hh[0]); acc = fmaf(f.x, w0.x, acc); acc = fmaf(f.y, w0.y, acc);
                f = __half22float2(hh[1]); acc = fmaf(f.x, w0.z, acc); acc = fmaf(f.y, w0.w, acc);
                f = __half22float2(hh[2]); acc = fmaf(f.x, w1.x, acc); acc = fmaf(f.y, w1.y, acc);
                f = __half22float2(hh[3]); acc = fmaf(f.x, w1.z, acc); acc = fmaf(f.y, w1.w, acc);
            }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) {
            float z = acc + p.b_out + (p.wide ? p.wide[b] : 0.f);
            if (p.scores) p.scores[b] = sigmoidf_fast(z);
            if (p.extra_out) p.extra_out[b] = z;
        }
    }
}

}  // namespace ctr
