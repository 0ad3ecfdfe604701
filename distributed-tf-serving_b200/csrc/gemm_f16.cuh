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
// Structure (one CTA per SM, persistent over output tiles, 256 threads):
//   warp 0    TMA producer: cp.async.bulk.tensor 2-D loads of a 128x64 A box and a BNx64 W
//             box (128-byte swizzle) into a kStages-deep shared-memory ring, mbarrier paced
//   warp 1    MMA issuer: one thread issues tcgen05.mma (M=128, N=BN, K=16, f16 -> f32)
//             four times per stage, accumulating in TMEM; tcgen05.commit frees the stage
//             and, after the last K block, publishes the accumulator
//   warp 2    TMEM allocator (2 accumulators x BN columns, so the epilogue of tile i
//             overlaps the main loop of tile i+1)
//   warps 4-7 epilogue: tcgen05.ld 32 lanes x 32 columns -> registers, bias/activation,
//             pack to half, 16-byte stores (each thread owns one output row)
//
// Roofline: tensor pipe.  2*M*N*K flop per launch against the measured dense f16 rate.
#pragma once
#include "common.cuh"

namespace ctr {

constexpr int kGemmThreads = 256;
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
};

template <int BN>
struct GemmCfg {
    static constexpr int kStages = BN == 256 ? 4 : 6;
    static constexpr int kABytes = kBM * kBK * 2;            // 16 KB
    static constexpr int kBBytes = BN * kBK * 2;             // 32 KB / 16 KB
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kTmemCols = 2 * BN;                 // 512 / 256 (powers of two)
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/;
};

template <int BN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_f16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w, const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int kStages = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    // 128-byte swizzle needs 1024-byte aligned stage buffers
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kStages * Cfg::kABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
    uint64_t* full_bar = bars;                     // [kStages] TMA -> MMA
    uint64_t* empty_bar = bars + kStages;          // [kStages] MMA -> TMA
    uint64_t* acc_full = bars + 2 * kStages;       // [2] MMA -> epilogue
    uint64_t* acc_empty = bars + 2 * kStages + 2;  // [2] epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_m = (p.M + kBM - 1) / kBM;
    const int tiles_n = p.N / BN;
    const int n_tiles = tiles_m * tiles_n;
    const int k_blocks = p.K / kBK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_w);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 4); }
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_slot, Cfg::kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int m0 = (t / tiles_n) * kBM, n0 = (t % tiles_n) * BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    mbar_expect_tx(&full_bar[s], Cfg::kStageBytes);
                    tma_load_2d(smem_a + s * Cfg::kABytes, &map_a, kb * kBK, m0, &full_bar[s]);
                    tma_load_2d(smem_b + s * Cfg::kBBytes, &map_w, kb * kBK, n0, &full_bar[s]);
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(kBM, BN, /*f16*/ 0);
            int s = 0; uint32_t ph = 0;
            int lt = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++lt) {
                const int acc = lt & 1;
                mbar_wait(&acc_empty[acc], ((lt >> 1) & 1) ^ 1);      // epilogue drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + s * Cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smem_b + s * Cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k) {
                        // advancing K by 16 halfs = 32 bytes inside the 128-byte swizzle span
                        const uint64_t da = make_sw128_kmajor_desc(a_addr + k * 32);
                        const uint64_t db = make_sw128_kmajor_desc(b_addr + k * 32);
                        umma_f16(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[s]);                       // stage free once these MMAs retire
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
                umma_commit(&acc_full[acc]);                          // accumulator complete
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int sub = warp & 3;                                     // TMEM sub-partition = warp % 4
        int lt = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++lt) {
            const int acc = lt & 1;
            const int m0 = (t / tiles_n) * kBM, n0 = (t % tiles_n) * BN;
            const int row = m0 + sub * 32 + lane;
            const bool row_ok = row < p.M;
            mbar_wait(&acc_full[acc], (lt >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(sub * 32) << 16) + acc * BN;
            float head_acc = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t r[32];
                tmem_ld32(taddr + c * 32, r);
                tmem_ld_wait();
                const int col0 = n0 + c * 32;
                const float4* bias4 = reinterpret_cast<const float4*>(p.bias + col0);
                if (EPI == EPI_HEAD) {
                    const float4* w4 = reinterpret_cast<const float4*>(p.w_out + col0);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 bv = __ldg(bias4 + j), wv = __ldg(w4 + j);
                        head_acc = fmaf(fmaxf(__uint_as_float(r[4 * j + 0]) + bv.x, 0.f), wv.x, head_acc);
                        head_acc = fmaf(fmaxf(__uint_as_float(r[4 * j + 1]) + bv.y, 0.f), wv.y, head_acc);
                        head_acc = fmaf(fmaxf(__uint_as_float(r[4 * j + 2]) + bv.z, 0.f), wv.z, head_acc);
                        head_acc = fmaf(fmaxf(__uint_as_float(r[4 * j + 3]) + bv.w, 0.f), wv.w, head_acc);
                    }
                } else {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 bv = __ldg(bias4 + j);
                        v[4 * j + 0] = __uint_as_float(r[4 * j + 0]) + bv.x;
                        v[4 * j + 1] = __uint_as_float(r[4 * j + 1]) + bv.y;
                        v[4 * j + 2] = __uint_as_float(r[4 * j + 2]) + bv.z;
                        v[4 * j + 3] = __uint_as_float(r[4 * j + 3]) + bv.w;
                    }
                    if (row_ok) {
                        if (EPI == EPI_RELU) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
                        } else {   // EPI_CROSS
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
                        uint4* op = reinterpret_cast<uint4*>(p.out + (size_t)row * p.ldo + col0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            __half2 h[4];
#pragma unroll
                            for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[8 * j + 2 * e], v[8 * j + 2 * e + 1]);
                            op[j] = *reinterpret_cast<const uint4*>(h);
                        }
                    }
                }
            }
            // TMEM reads of this accumulator are done: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
            if (EPI == EPI_HEAD && row_ok) {
                float z = head_acc + p.b_out;
                if (p.wide) z += p.wide[row];
                if (p.extra) z += p.extra[row];
                p.scores[row] = sigmoidf_fast(z);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::kTmemCols);
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
