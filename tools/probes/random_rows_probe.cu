// random_rows_probe.cu -- how many random table rows per second can a B200 fetch from HBM, whatever the kernel?
//
// The gather kernel's time does not follow the bytes it moves (f16 rows: 34 % fewer DRAM bytes, same time; embedding
// only or first-order weight only: same time as both).  This probe takes everything else away: a table of `rows`
// rows of ROW bytes, every warp fetches rows at hashed (uniformly random) positions with as many independent loads in
// flight as registers allow, and xors them into a sink.  Output: rows/s and bytes/s for ROW = 32 .. 512, for a table
// far larger than L2 and for one that fits it.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o random_rows_probe random_rows_probe.cu && ./random_rows_probe [table GB] [quick]
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__host__ __device__ inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// LANES lanes fetch one row of LANES * 16 bytes with one 128-bit load each; a warp covers 32 / LANES rows per
// instruction and keeps DEPTH instructions in flight.
template <int LANES, int DEPTH>
__global__ void __launch_bounds__(256) fetch_rows(const uint4* __restrict__ table, uint64_t row_mask, uint64_t lookups_per_group,
                                                  uint64_t seed, uint4* __restrict__ sink) {
    const uint32_t lane = threadIdx.x & 31, sub = lane % LANES;
    const uint64_t group = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES;      // one row per group and step
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (uint64_t i = 0; i < lookups_per_group; i += DEPTH) {
        uint4 v[DEPTH];
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            const uint64_t row = mix64(seed + group * lookups_per_group + i + d) & row_mask;
            v[d] = __ldg(table + row * LANES + sub);
        }
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) { acc.x ^= v[d].x; acc.y ^= v[d].y; acc.z ^= v[d].z; acc.w ^= v[d].w; }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x12345679u) sink[0] = acc;       // never true in practice: keeps the loads alive
}

template <int LANES, int DEPTH>
int run(const uint4* table, size_t table_bytes, uint4* sink, const char* what, int blocks_per_sm = 8) {
    const uint64_t rows = table_bytes / (LANES * 16);
    uint64_t pow2 = 1;
    while (pow2 * 2 <= rows) pow2 *= 2;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int blocks = sms * blocks_per_sm;
    const uint64_t groups = (uint64_t)blocks * 256 / LANES;
    const uint64_t per_group = (((uint64_t)64 << 20) / groups / DEPTH + 1) * DEPTH;           // ~64 M lookups per launch
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(a));
        fetch_rows<LANES, DEPTH><<<blocks, 256>>>(table, pow2 - 1, per_group, 1234 + rep * 7919, sink);
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (rep && ms < best) best = ms;
    }
    const double n = (double)groups * per_group;
    printf("%-18s %2d warps/SM, row %4d B, %2d loads in flight per lane: %6.2f G rows/s  %6.2f TB/s  (%.2f ms, %.0f M lookups, %.1f GB of rows)\n", what, blocks_per_sm * 8, LANES * 16, DEPTH,
           n / best / 1e6, n * LANES * 16 / best / 1e9, best, n / 1e6, (double)pow2 * LANES * 16 / 1e9);
    return 0;
}

int main(int argc, char** argv) {
    const size_t big = (size_t)(argc > 1 ? atof(argv[1]) : 16.0) * (1ull << 30), small = (size_t)48 << 20;
    uint4 *table, *sink;
    CK(cudaMalloc(&table, big));
    CK(cudaMalloc(&sink, 64));
    CK(cudaMemset(table, 1, big));
    if (argc > 2) {          // "quick": the one line bench.py reads (128-byte rows, 32 warps per SM, 8 loads in flight per lane)
        return run<8, 8>(table, big, sink, "table >> L2 (HBM)", 4);
    }
    int n_dev = 0;
    CK(cudaGetDeviceCount(&n_dev));
    if (n_dev > 1) {         // the same fetches from a PEER's memory (plain loads over NVLink, what the row-sharded gather does)
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, 0, 1));
        if (can) {
            uint4* remote;
            const size_t peer_bytes = (size_t)8 << 30;
            CK(cudaSetDevice(1));
            CK(cudaMalloc(&remote, peer_bytes));
            CK(cudaMemset(remote, 1, peer_bytes));
            CK(cudaDeviceSynchronize());
            CK(cudaSetDevice(0));
            CK(cudaDeviceEnablePeerAccess(1, 0));
            const char* what = "peer table (NVLink)";
            if (run<4, 8>(remote, peer_bytes, sink, what, 4)) return 1;
            if (run<8, 2>(remote, peer_bytes, sink, what, 4)) return 1;
            if (run<8, 8>(remote, peer_bytes, sink, what, 4)) return 1;
            if (run<8, 8>(remote, peer_bytes, sink, what, 8)) return 1;
            if (run<8, 16>(remote, peer_bytes, sink, what, 8)) return 1;
            if (run<16, 8>(remote, peer_bytes, sink, what, 8)) return 1;
            if (run<32, 8>(remote, peer_bytes, sink, what, 8)) return 1;
        }
    }
    for (int pass = 0; pass < 2; ++pass) {
        const size_t bytes = pass == 0 ? big : small;
        const char* what = pass == 0 ? "table >> L2 (HBM)" : "table 48 MB (L2)";
        if (run<2, 8>(table, bytes, sink, what)) return 1;
        if (run<2, 16>(table, bytes, sink, what)) return 1;
        if (run<4, 8>(table, bytes, sink, what)) return 1;
        if (run<4, 16>(table, bytes, sink, what)) return 1;
        if (run<8, 4>(table, bytes, sink, what)) return 1;
        if (run<8, 8>(table, bytes, sink, what)) return 1;
        if (run<8, 16>(table, bytes, sink, what)) return 1;
        if (pass == 0) {                     // how the rate depends on the warps per SM and on the loads each keeps in flight
            for (int bps = 1; bps <= 4; bps *= 2) {
                if (run<8, 2>(table, bytes, sink, what, bps)) return 1;
                if (run<8, 4>(table, bytes, sink, what, bps)) return 1;
                if (run<8, 8>(table, bytes, sink, what, bps)) return 1;
                if (run<8, 16>(table, bytes, sink, what, bps)) return 1;
                if (run<4, 8>(table, bytes, sink, what, bps)) return 1;
            }
        }
        if (run<16, 8>(table, bytes, sink, what)) return 1;
        if (run<32, 8>(table, bytes, sink, what)) return 1;
    }
    return 0;
}
