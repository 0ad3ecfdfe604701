// sort.cuh -- the reference's post-merge sort (Collections.sort(fullCtrList), DCNClient.java:195) on
// the device: ascending in Float.compareTo order (-0.0 < +0.0, every NaN last), and -- what the
// reference's sort throws away -- the candidate index of every sorted score.
//
// Requests of up to 32 768 candidates: the whole sort in one block and one launch (end of this file).
// Larger: LSD radix sort, 4 passes of 8 bits over (key, index) pairs, each pass = histogram per tile,
// one-block exclusive scan, stable scatter.  Integer work: results are bit-exact with the oracle
// (oracle/dcn_client.py::java_float_sort_key + a stable argsort).  HBM-bound, but tiny next to
// the scoring (8 bytes per candidate per pass); it exists for completeness of the timed region
// of processARequest, not as a hot kernel.
#pragma once
#include "common.cuh"
#ifdef CTR_SORT_PROF
#include <cstdio>
#endif

namespace ctr {

constexpr int kSortThreads = 256;
constexpr int kSortTile = 4096;        // elements per block per pass

// Float.compareTo as an unsigned key: canonical NaN, sign-magnitude -> biased
__device__ __forceinline__ uint32_t java_sort_key(float f) {
    uint32_t b = __float_as_uint(f);
    if (f != f) b = 0x7FC00000u;
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void sort_make_keys_kernel(const float* scores, int n, uint32_t* keys, int32_t* idx) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        keys[i] = java_sort_key(scores[i]);
        idx[i] = i;
    }
}

// counts[digit * n_blocks + block] = number of keys of this block's tile with that digit
__global__ void __launch_bounds__(kSortThreads) sort_hist_kernel(const uint32_t* keys, int n, int shift, uint32_t* counts) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int t0 = blockIdx.x * kSortTile, t1 = min(n, t0 + kSortTile);
    for (int i = t0 + threadIdx.x; i < t1; i += kSortThreads) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
    __syncthreads();
    counts[threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of `total` counters, in place, one block
__global__ void __launch_bounds__(1024) sort_scan_kernel(uint32_t* counts, int total) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < total; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < total ? counts[i] : 0;
        uint32_t incl = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_sums[lane], wi = w;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, wi, off);
                if (lane >= off) wi += t;
            }
            warp_sums[lane] = wi - w;                       // exclusive prefix of the warp totals
        }
        __syncthreads();
        const uint32_t excl = carry + warp_sums[warp] + incl - v;
        if (i < total) counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}

// stable scatter of this block's tile: position = scanned[digit][block] + rank among the tile's
// earlier keys with the same digit
__global__ void __launch_bounds__(kSortThreads) sort_scatter_kernel(const uint32_t* keys_in, const int32_t* idx_in, int n, int shift,
                                                                    const uint32_t* scanned, uint32_t* keys_out, int32_t* idx_out) {
    __shared__ uint32_t digit_base[256];
    __shared__ uint32_t wc[kSortThreads / 32][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    digit_base[threadIdx.x] = scanned[threadIdx.x * gridDim.x + blockIdx.x];
    const int t0 = blockIdx.x * kSortTile, t1 = min(n, t0 + kSortTile);
    for (int c0 = t0; c0 < t1; c0 += kSortThreads) {
        for (int w = 0; w < kSortThreads / 32; ++w) wc[w][threadIdx.x] = 0;
        __syncthreads();
        const int i = c0 + threadIdx.x;
        const bool ok = i < t1;
        const uint32_t key = ok ? keys_in[i] : 0xFFFFFFFFu;
        const uint32_t d = (key >> shift) & 255u;
        // lanes of this warp with the same digit (inactive tail lanes form their own group via ok)
        const uint32_t peers = __match_any_sync(0xffffffffu, ok ? d : 256u);
        const uint32_t rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (ok && rank_in_warp == 0) wc[warp][d] = __popc(peers);
        __syncthreads();
        if (ok) {
            uint32_t pos = digit_base[d] + rank_in_warp;
            for (int w = 0; w < warp; ++w) pos += wc[w][d];
            keys_out[pos] = key;
            idx_out[pos] = idx_in[i];
        }
        __syncthreads();
        uint32_t add = 0;
        for (int w = 0; w < kSortThreads / 32; ++w) add += wc[w][threadIdx.x];
        digit_base[threadIdx.x] += add;
        __syncthreads();
    }
}

__global__ void sort_gather_scores_kernel(const float* scores, const int32_t* idx, int n, float* sorted) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) sorted[i] = scores[idx[i]];
}

// ------------------------------------------------------------------ one request's worth: a single CTA
// Up to kBlockSortMax candidates (the reference's 1 500, config 2's 10 000) are sorted by ONE block in ONE
// launch: the same 4 x 8-bit LSD passes, with histogram, scan and stable scatter inside the block (the 14
// dependent launches of the multi-block path cost ~110 us of launch latency for 10 000 candidates).
// Each of the 32 warps owns a contiguous segment of the array, so warp order = element order: a pass is
//   count    every warp histograms its segment into its own row of a [warp][digit] table
//   prefix   per digit, exclusive prefix over (digit, warp): every warp's first output slot per digit
//   scatter  every warp walks its segment 32 elements at a time; ballots give an element its rank among
//            the equal digits of the step, the warp's own table row is its running offset (no block-wide
//            synchronisation inside the sweep)
// The (key, index) arrays ping-pong through global memory (L2-resident at this size).
constexpr int kBlockSortMax = 32768;
constexpr int kBlockSortThreads = 1024;

__global__ void __launch_bounds__(kBlockSortThreads) sort_block_kernel(const float* scores, int n, uint32_t* keys0, int32_t* idx0, uint32_t* keys1,
                                                                       int32_t* idx1, float* sorted) {
    __shared__ uint32_t table[32][256];         // count, then running output offset, per warp and digit
    __shared__ uint32_t part[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg = ((n + 31) / 32 + 31) & ~31;                 // elements per warp, a multiple of 32
    const int s0 = warp * seg, s1 = min(n, s0 + seg);
#ifdef CTR_SORT_PROF
    long long tp[8]; tp[0] = clock64();
#endif
    for (int i = tid; i < n; i += kBlockSortThreads) { keys0[i] = java_sort_key(scores[i]); idx0[i] = i; }
    __syncthreads();
#ifdef CTR_SORT_PROF
    tp[1] = clock64();
#endif
    for (int pass = 0; pass < 4; ++pass) {
        const uint32_t* ksrc = (pass & 1) ? keys1 : keys0;
        const int32_t* isrc = (pass & 1) ? idx1 : idx0;
        uint32_t* kdst = (pass & 1) ? keys0 : keys1;
        int32_t* idst = (pass & 1) ? idx0 : idx1;
        const int shift = pass * 8;
        // Segments of up to kRegSteps x 32 elements (n <= 16 384) are held in registers for the whole pass: all
        // their loads are in flight at once, and the scatter sweep's step-to-step chain has no memory access
        // but shared memory (with one L2 round trip per step a pass over 10 000 elements took 27k cycles)
        constexpr int kRegSteps = 16;
        const bool in_regs = seg <= kRegSteps * 32;
        uint32_t kreg[kRegSteps];
        int32_t ireg[kRegSteps];
        if (in_regs) {
#pragma unroll
            for (int q = 0; q < kRegSteps; ++q) {
                const int i = s0 + q * 32 + lane;
                kreg[q] = i < s1 ? ksrc[i] : 0u;
                ireg[q] = i < s1 ? isrc[i] : 0;
            }
        }
        // ---- count
#pragma unroll
        for (int q = 0; q < 8; ++q) table[warp][lane + q * 32] = 0;
        __syncwarp();
        if (in_regs) {
#pragma unroll
            for (int q = 0; q < kRegSteps; ++q)
                if (s0 + q * 32 + lane < s1) atomicAdd(&table[warp][(kreg[q] >> shift) & 255u], 1u);
        } else {
            for (int i = s0 + lane; i < s1; i += 32) atomicAdd(&table[warp][(ksrc[i] >> shift) & 255u], 1u);
        }
        __syncthreads();
        // ---- prefix over (digit, warp)
        uint32_t total = 0, incl = 0;
        if (tid < 256) {
#pragma unroll
            for (int w = 0; w < 32; ++w) total += table[w][tid];
            incl = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) part[warp] = incl;
        }
        __syncthreads();
        if (tid < 256) {
            uint32_t run = incl - total;
            for (int w = 0; w < warp; ++w) run += part[w];
#pragma unroll
            for (int w = 0; w < 32; ++w) { const uint32_t c = table[w][tid]; table[w][tid] = run; run += c; }
        }
        __syncthreads();
        // ---- scatter: this warp's segment, in order
        auto place = [&](bool valid, uint32_t key, int32_t src) {
            const uint32_t dgt = valid ? (key >> shift) & 255u : 256u;
            // lanes holding the same digit: 9 ballots (8 digit bits + validity); __match_any_sync costs one
            // round per DISTINCT value, ~32 here
            uint32_t peers = 0xFFFFFFFFu;
#pragma unroll
            for (int bit = 0; bit < 9; ++bit) {
                const uint32_t vote = __ballot_sync(0xFFFFFFFFu, (dgt >> bit) & 1u);
                peers &= ((dgt >> bit) & 1u) ? vote : ~vote;
            }
            const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
            uint32_t pos = 0;
            if (valid) pos = table[warp][dgt] + rank;
            __syncwarp();
            if (valid && rank == 0) table[warp][dgt] += __popc(peers);
            __syncwarp();
            if (valid) { kdst[pos] = key; idst[pos] = src; }
        };
        if (in_regs) {
#pragma unroll
            for (int q = 0; q < kRegSteps; ++q)
                if (s0 + q * 32 < s1) place(s0 + q * 32 + lane < s1, kreg[q], ireg[q]);
        } else {
            uint32_t key_n = s0 + lane < s1 ? ksrc[s0 + lane] : 0u;
            int32_t idx_n = s0 + lane < s1 ? isrc[s0 + lane] : 0;
            for (int i0 = s0; i0 < s1; i0 += 32) {
                const int i = i0 + lane;
                const uint32_t key = key_n;
                const int32_t src = idx_n;
                if (i + 32 < s1) { key_n = ksrc[i + 32]; idx_n = isrc[i + 32]; }
                place(i < s1, key, src);
            }
        }
        __syncthreads();                                        // this pass's stores are visible to the next pass's loads
#ifdef CTR_SORT_PROF
        tp[2 + pass] = clock64();
#endif
    }
    for (int i = tid; i < n; i += kBlockSortThreads) sorted[i] = scores[idx0[i]];       // 4 passes: the result is back in (keys0, idx0)
#ifdef CTR_SORT_PROF
    tp[6] = clock64();
    if (tid == 0) printf("[sort prof] n=%d keys %lld  pass0 %lld pass1 %lld pass2 %lld pass3 %lld  gather %lld cycles\n", n, tp[1] - tp[0],
                         tp[2] - tp[1], tp[3] - tp[2], tp[4] - tp[3], tp[5] - tp[4], tp[6] - tp[5]);
#endif
}

}  // namespace ctr
