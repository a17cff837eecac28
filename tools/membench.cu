// membench.cu -- what HBM gives for the calibration traffic mix (2 B in : 4 B out per pixel), independent of
// the calibration math: u16 -> f32 convert kernels in linear and frames-inner order.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/membench tools/membench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ float4 cvt(uint2 r) {
    return make_float4((float)(r.x & 0x3FFF), (float)((r.x >> 16) & 0x3FFF), (float)(r.y & 0x3FFF), (float)((r.y >> 16) & 0x3FFF));
}
// A: linear, one group (4 px) per thread per iteration, grid-stride, UN in flight
template <int UN>
__global__ void __launch_bounds__(256) lin(const uint2* __restrict__ in, float4* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * 256 * UN + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * 256 * UN;
    for (; i + 256 * (UN - 1) < n; i += stride) {
        uint2 r[UN];
#pragma unroll
        for (int u = 0; u < UN; ++u) r[u] = __ldcs(in + i + u * 256);
#pragma unroll
        for (int u = 0; u < UN; ++u) __stcs(out + i + u * 256, cvt(r[u]));
    }
}
// B: frames-inner like epix_calib_map_kernel: thread owns group g, walks frames f0..f1
template <int UN>
__global__ void __launch_bounds__(256) finner(const uint2* __restrict__ in, float4* __restrict__ out, int ngroups, int nframes, int fpb) {
    const int g = blockIdx.x * 256 + threadIdx.x;
    if (g >= ngroups) return;
    const int f0 = blockIdx.y * fpb, f1 = min(nframes, f0 + fpb);
    for (int f = f0; f + UN <= f1; f += UN) {
        uint2 r[UN];
#pragma unroll
        for (int u = 0; u < UN; ++u) r[u] = __ldcs(in + (size_t)(f + u) * ngroups + g);
#pragma unroll
        for (int u = 0; u < UN; ++u) __stcs(out + (size_t)(f + u) * ngroups + g, cvt(r[u]));
    }
}
// C: write only / read only
__global__ void __launch_bounds__(256) wonly(float4* __restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) __stcs(out + i, make_float4(1.f, 2.f, 3.f, 4.f));
}
__global__ void __launch_bounds__(256) ronly(const uint4* __restrict__ in, size_t n, unsigned* sink) {
    unsigned acc = 0;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) { uint4 v = __ldcs(in + i); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    if (acc == 0x12345678u) *sink = acc;
}

int main() {
    const int nframes = 64, npix = 16 * 352 * 384, ngroups = npix / 4;
    const size_t n = (size_t)nframes * ngroups;
    uint2* in; float4* out; unsigned* sink;
    CK(cudaMalloc(&in, n * 8)); CK(cudaMalloc(&out, n * 16)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(in, 0x11, n * 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double bytes = (double)n * 24;
    auto report = [&](const char* name, float ms, double b) { printf("%-44s %8.1f us  %7.0f GB/s\n", name, ms * 1e3, b / ms / 1e6); };
#define TIME(name, b, launch) { for (int i = 0; i < 5; ++i) { launch; } CK(cudaDeviceSynchronize()); cudaEventRecord(e0); \
    for (int i = 0; i < 40; ++i) { launch; } cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); float ms; cudaEventElapsedTime(&ms, e0, e1); report(name, ms / 40, b); }
    for (int blocks : {148 * 4, 148 * 8, 148 * 16, 148 * 32}) {
        char nm[64];
        snprintf(nm, 64, "linear UN=4 blocks=%d", blocks); TIME(nm, bytes, (lin<4><<<blocks, 256>>>(in, out, n)));
        snprintf(nm, 64, "linear UN=8 blocks=%d", blocks); TIME(nm, bytes, (lin<8><<<blocks, 256>>>(in, out, n)));
    }
    for (int fpb : {8, 16, 32, 64}) {
        char nm[64];
        dim3 grid((ngroups + 255) / 256, nframes / fpb);
        snprintf(nm, 64, "frames-inner UN=4 fpb=%d", fpb); TIME(nm, bytes, (finner<4><<<grid, 256>>>(in, out, ngroups, nframes, fpb)));
        snprintf(nm, 64, "frames-inner UN=8 fpb=%d", fpb); TIME(nm, bytes, (finner<8><<<grid, 256>>>(in, out, ngroups, nframes, fpb)));
    }
    TIME("write only (16 B/group)", (double)n * 16, (wonly<<<148 * 16, 256>>>(out, n)));
    TIME("read only (8 B/group)", (double)n * 8, (ronly<<<148 * 16, 256>>>((const uint4*)in, n / 2, sink)));
    return 0;
}
