// coresidency_probe.cu -- can a CTA with the gather's footprint (B: 256 threads, 64 registers, ~30 KB shared memory)
// become resident on an SM that already holds a CTA with the tower's footprint (A: cluster of 2, 384 threads, 128
// registers, 197 888 B of dynamic shared memory)?  A spins for 2 ms; B is launched on a second stream 300 us later
// and every B block records when it started.  Result on B200 (profiles/r02_coresidency_probe.txt): never, unless
// BOTH kernels prefer the maximum shared-memory carve-out; then 128 + 64 registers and 194 + 32 KB fit, 33 KB do not.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o coresidency_probe coresidency_probe.cu
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <unistd.h>
#include <vector>

__device__ __forceinline__ uint64_t gtime() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

// keeps R floats live across the spin so that ptxas has to give the kernel about R + 8 registers
template <int R>
__device__ __forceinline__ float burn(const float* in, uint64_t ns, int tid) {
    float a[R];
#pragma unroll
    for (int i = 0; i < R; ++i) a[i] = in[i * 32 + (tid & 31)];
    const uint64_t t0 = gtime();
    while (gtime() - t0 < ns) {
#pragma unroll
        for (int i = 0; i < R; ++i) a[i] = a[i] * 1.0001f + 0.5f;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < R; ++i) s += a[i];
    return s;
}

template <int R>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(512, 1) kernel_a(const float* in, float* out, uint64_t ns, uint64_t* rec) {
    extern __shared__ uint8_t sm[];
    if (threadIdx.x == 0) { rec[blockIdx.x * 2] = gtime(); sm[0] = 1; }
    const float s = burn<R>(in, ns, threadIdx.x);
    if (s == 123.456f) out[0] = s;
    if (threadIdx.x == 0) rec[blockIdx.x * 2 + 1] = gtime();
}

template <int R>
__global__ void __launch_bounds__(256, 4) kernel_b(const float* in, float* out, uint64_t ns, uint64_t* rec) {
    extern __shared__ uint8_t sm[];
    if (threadIdx.x == 0) { rec[blockIdx.x * 2] = gtime(); sm[0] = 1; }
    const float s = burn<R>(in, ns, threadIdx.x);
    if (s == 123.456f) out[0] = s;
    if (threadIdx.x == 0) rec[blockIdx.x * 2 + 1] = gtime();
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

template <typename KA, typename KB>
void run(const char* name, KA ka, KB kb, int smem_a, int smem_b, bool carve_a, bool carve_b) {
    cudaFuncAttributes fa, fb;
    CK(cudaFuncGetAttributes(&fa, ka));
    CK(cudaFuncGetAttributes(&fb, kb));
    CK(cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_a));
    CK(cudaFuncSetAttribute(ka, cudaFuncAttributePreferredSharedMemoryCarveout, carve_a ? 100 : -1));
    CK(cudaFuncSetAttribute(kb, cudaFuncAttributePreferredSharedMemoryCarveout, carve_b ? 100 : -1));
    float *in, *out; uint64_t *ra, *rb;
    const int ga = 148, gb = 148 * 4;
    CK(cudaMalloc(&in, 1 << 20)); CK(cudaMemset(in, 0, 1 << 20)); CK(cudaMalloc(&out, 64));
    CK(cudaMalloc(&ra, ga * 16)); CK(cudaMalloc(&rb, gb * 16));
    cudaStream_t s1, s2;
    CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    for (int rep = 0; rep < 2; ++rep) {
        ka<<<ga, 384, smem_a, s1>>>(in, out, 2000000ull, ra);
        usleep(300);
        kb<<<gb, 256, smem_b, s2>>>(in, out, 100000ull, rb);
        CK(cudaDeviceSynchronize());
    }
    std::vector<uint64_t> ha(ga * 2), hb(gb * 2);
    CK(cudaMemcpy(ha.data(), ra, ga * 16, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hb.data(), rb, gb * 16, cudaMemcpyDeviceToHost));
    uint64_t a0 = ~0ull, a1 = 0, b0 = ~0ull;
    for (int i = 0; i < ga; ++i) { a0 = std::min(a0, ha[i * 2]); a1 = std::max(a1, ha[i * 2 + 1]); }
    int early = 0;
    for (int i = 0; i < gb; ++i) { if (hb[i * 2] < a1 - 100000) ++early; b0 = std::min(b0, hb[i * 2]); }
    printf("%-52s A %3d regs %6d B smem, B %2d regs %5d B smem: first B block %+5.0f us after A's start; %3d of %d B blocks started while A was resident\n",
           name, fa.numRegs, smem_a, fb.numRegs, smem_b, ((double)b0 - (double)a0) / 1e3, early, gb);
    cudaFree(in); cudaFree(out); cudaFree(ra); cudaFree(rb); cudaStreamDestroy(s1); cudaStreamDestroy(s2);
}

int main() {
    run("default carve-outs", kernel_a<120>, kernel_b<56>, 197888, 30080, false, false);
    run("max carve-out on B only", kernel_a<120>, kernel_b<56>, 197888, 30080, false, true);
    run("max carve-out on A only", kernel_a<120>, kernel_b<56>, 197888, 30080, true, false);
    run("max carve-out on both", kernel_a<120>, kernel_b<56>, 197888, 30080, true, true);
    run("max carve-out on both, B 32 KB", kernel_a<120>, kernel_b<56>, 197888, 32768, true, true);
    run("max carve-out on both, B 33 KB", kernel_a<120>, kernel_b<56>, 197888, 33792, true, true);
    run("max carve-out on both, A 64 regs, B 48 regs 16 KB", kernel_a<56>, kernel_b<40>, 197888, 16384, true, true);
    run("default carve-outs, A 64 regs, B 48 regs 16 KB", kernel_a<56>, kernel_b<40>, 197888, 16384, false, false);
    return 0;
}
