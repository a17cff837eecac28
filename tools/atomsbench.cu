// atomsbench.cu -- what shared-memory reductions cost on sm_100a, for the common-mode histogram design:
// red.shared.add.u32 / atom.shared.add.u32 / thread-private ld+add+st, under the address patterns a bank-median
// histogram produces (uniform bins, a 37-bin-sigma bell, distinct banks, one hot word) and with a fraction of the
// lanes predicated off.  Prints SM cycles per warp instruction and active lanes per clock per SM.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/atomsbench tools/atomsbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int NT = 352, NBIN = 2048;

// OP: 0 red, 1 atom (result used), 2 private ld/add/st (word [bin&31][tid]), 3 no memory op (address arithmetic only)
// PAT: 0 uniform over 2048 bins, 1 bell (sum of four 6-bit fields + 896), 2 one bank per lane, 3 one word
// ACT: lanes active out of 32 (lane < ACT)
template <int OP, int PAT, int ACT>
__global__ void __launch_bounds__(NT, 2) k(int iters, uint32_t* sink, long long* cyc)
{
    __shared__ uint32_t hist[NBIN];
    extern __shared__ uint32_t priv[];          // OP 2: 32 * NT words
    for (int i = threadIdx.x; i < NBIN; i += NT) hist[i] = 0;
    if (OP == 2) for (int i = threadIdx.x; i < 32 * NT; i += NT) priv[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    uint32_t s = (blockIdx.x * NT + threadIdx.x) * 2654435761u + 12345u, acc = 0;
    const uint32_t haddr = (uint32_t)__cvta_generic_to_shared(hist);
    const bool on = lane < ACT;
    const long long t0 = clock64();
#pragma unroll 8
    for (int it = 0; it < iters; ++it) {
        s = s * 1664525u + 1013904223u;
        uint32_t bin;
        if (PAT == 0) bin = s >> 21;
        else if (PAT == 1) bin = 896u + ((s >> 26) + ((s >> 20) & 63u) + ((s >> 14) & 63u) + ((s >> 8) & 63u));
        else if (PAT == 2) bin = (uint32_t)lane + ((s >> 21) & ~31u);
        else bin = 1000u;
        if (OP == 0) {
            asm volatile("{ .reg .pred q; setp.ne.u32 q, %1, 0; @q red.shared.add.u32 [%0], 1; }" :: "r"(haddr + bin * 4u), "r"((uint32_t)on) : "memory");
        } else if (OP == 1) {
            if (on) acc += atomicAdd(&hist[bin], 1u);
        } else if (OP == 2) {
            if (on) { uint32_t* p = priv + (bin & 31u) * NT + threadIdx.x; *(volatile uint32_t*)p = *(volatile uint32_t*)p + 1u; }
        } else acc += bin;
    }
    const long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    for (int i = threadIdx.x; i < NBIN; i += NT) acc += hist[i];
    if (acc == 0xdeadbeefu) sink[0] = acc;
}

template <int OP, int PAT, int ACT>
int run(const char* name, int nsm)
{
    const int iters = 4096, grid = nsm * 2;
    uint32_t* sink; long long* cyc;
    CK(cudaMalloc(&sink, 4)); CK(cudaMalloc(&cyc, grid * 8));
    const size_t dyn = OP == 2 ? (size_t)32 * NT * 4 : 0;
    CK(cudaFuncSetAttribute(k<OP, PAT, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k<OP, PAT, ACT><<<grid, NT, dyn>>>(iters, sink, cyc);
    CK(cudaEventRecord(e0));
    k<OP, PAT, ACT><<<grid, NT, dyn>>>(iters, sink, cyc);
    CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    long long* h = new long long[grid]; CK(cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost));
    double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid;
    // per SM: 2 CTAs x 11 warps x iters warp instructions in `avg` cycles
    const double winst = 2.0 * (NT / 32) * iters;
    printf("%-44s %8.1f us  %7.2f cyc/warp-op/SM  %6.2f active lanes/clk/SM\n", name, ms * 1e3, avg / winst, winst * ACT / avg);
    delete[] h; cudaFree(sink); cudaFree(cyc);
    return 0;
}

int main()
{
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    printf("%s, %d SMs; 2 CTAs x %d threads per SM\n", p.name, nsm, NT);
    run<3, 0, 32>("address arithmetic only, uniform", nsm);
    run<3, 1, 32>("address arithmetic only, bell", nsm);
    run<0, 0, 32>("red.shared uniform 2048 bins", nsm);
    run<0, 1, 32>("red.shared bell (sigma 37 bins)", nsm);
    run<0, 2, 32>("red.shared one bank per lane", nsm);
    run<0, 3, 32>("red.shared one word", nsm);
    run<0, 0, 4>("red.shared uniform, 4 of 32 lanes", nsm);
    run<0, 1, 4>("red.shared bell, 4 of 32 lanes", nsm);
    run<0, 1, 8>("red.shared bell, 8 of 32 lanes", nsm);
    run<0, 1, 16>("red.shared bell, 16 of 32 lanes", nsm);
    run<1, 0, 32>("atom.shared (result used) uniform", nsm);
    run<1, 1, 32>("atom.shared (result used) bell", nsm);
    run<1, 2, 32>("atom.shared one bank per lane", nsm);
    run<2, 0, 32>("private ld/add/st uniform", nsm);
    run<2, 1, 32>("private ld/add/st bell", nsm);
    return 0;
}
