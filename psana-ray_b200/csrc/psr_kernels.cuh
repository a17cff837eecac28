// psr_kernels.cuh -- device-side constants block and launcher prototypes shared by the
// kernel translation units and the C ABI (capi.cu).  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// Per-pixel flag byte (d_flags):
//   bit0  mask good (multiply by 1.0f) / bad (multiply by 0.0f)           calib_*: arrf*factor*mask
//   bit1  epix10ka: common-mode eligible when data bit 14 = 0 (already AND mask)
//   bit2  epix10ka: common-mode eligible when data bit 14 = 1 (already AND mask)
//   bit3  keep (producer-side np.where masks, calib layout); 1 when no keep mask is set
#define PSR_FL_MASK   1u
#define PSR_FL_ELIG0  2u
#define PSR_FL_ELIG1  4u
#define PSR_FL_KEEP   8u

struct PsrDev {
    int det_kind, nseg, rows, cols, npix;
    // epix10ka: 2-entry per-pixel table selected by data bit 14 (SoA, npix floats each)
    const float* ped0; const float* gf0; const float* ped1; const float* gf1;
    const uint8_t* modes;          // epix10ka: mode(d14=0) | mode(d14=1) << 4, 0xF = none
    // jungfrau: (3, npix) each; offs may be null (== zeros, exact to skip)
    const float* peds; const float* offs; const float* gfac;
    const uint8_t* flags;          // npix
    // image mode
    const int32_t* inv;            // H*W, source pixel or -1
    const uint8_t* keep_image;     // H*W or null
    int H, W;
    // image mode, scatter side (common mode fused with the assembly): per SOURCE pixel its image index if it is the
    // last writer of that target (>= 0), ~index if the image keep mask zeroes that target, INT32_MIN if another
    // pixel wins the target; gaps = image indices no pixel maps to (they are zero-filled)
    const int32_t* fwd; const int32_t* gaps; int ngaps;
    const int32_t* blk_order;      // scatter kernel: its 1024-pixel blocks sorted by where they land in the image (DRAM write locality), or null
    // common mode
    int cm_mode; float cormax; int npix_min; uint32_t elig_mask;
    int ignore_keep;               // 1: calib-layout keep mask not applied (image mode masks the 2-D image)
    int has_keep;                  // 1: a calib-layout keep mask is installed (else every flag byte has PSR_FL_KEEP set)
};

// launchers (each returns the launch's cudaError_t)
cudaError_t psr_launch_calib_map(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_out,
                                 int frames_per_block, cudaStream_t st);
cudaError_t psr_launch_decode(const PsrDev& c, const uint16_t* d_raw, int nframes, uint8_t* d_modes, cudaStream_t st);
cudaError_t psr_launch_image_fused(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_img,
                                   int frames_per_block, cudaStream_t st);
// raw -> image by scatter through c.fwd; writes every image pixel some detector pixel maps to -- the caller zeroes the
// rest with psr_launch_image_gap_fill (any order, any stream: the two write disjoint pixels)
cudaError_t psr_launch_image_scatter(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_img,
                                     int frames_per_block, cudaStream_t st);
cudaError_t psr_launch_image_gap_fill(const PsrDev& c, int nframes, float* d_img, cudaStream_t st);
cudaError_t psr_launch_image_gather(const PsrDev& c, const float* d_calib, int nframes, float* d_img, cudaStream_t st);
cudaError_t psr_launch_common_mode(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_out, cudaStream_t st);
// common mode + image assembly in one pass (bank medians only: cm_mode == 4; needs c.fwd); gaps as above
cudaError_t psr_launch_common_mode_image(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_img, cudaStream_t st);
cudaError_t psr_common_mode_setup(void);   // raises the dynamic shared-memory limit once per device

// ---- small device helpers -----------------------------------------------------------------------
#ifdef __CUDACC__
// order-preserving map float -> uint32 (a < b  <=>  key(a) < key(b)); inverse below
__device__ __forceinline__ uint32_t psr_f2key(float x) {
    uint32_t b = __float_as_uint(x);
    return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}
__device__ __forceinline__ float psr_key2f(uint32_t k) {
    uint32_t b = (k & 0x80000000u) ? (k ^ 0x80000000u) : ~k;
    return __uint_as_float(b);
}
// numpy median of an even count: (lo + hi) / 2 in float32, two roundings
__device__ __forceinline__ float psr_mid(float lo, float hi) { return __fmul_rn(__fadd_rn(lo, hi), 0.5f); }
#endif
