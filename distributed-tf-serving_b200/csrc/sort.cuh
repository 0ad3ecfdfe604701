// sort.cuh -- the reference's post-merge sort (Collections.sort(fullCtrList), DCNClient.java:195) on
// the device: ascending in Float.compareTo order (-0.0 < +0.0, every NaN last), and -- what the
// reference's sort throws away -- the candidate index of every sorted score.
//
// LSD radix sort, 4 passes of 8 bits over (key, index) pairs, each pass = histogram per tile,
// one-block exclusive scan, stable scatter.  Integer work: results are bit-exact with the oracle
// (oracle/dcn_client.py::java_float_sort_key + a stable argsort).  HBM-bound, but tiny next to
// the scoring (8 bytes per candidate per pass); it exists for completeness of the timed region
// of processARequest, not as a hot kernel.
#pragma once
#include "common.cuh"

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

}  // namespace ctr
