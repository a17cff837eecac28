// calib_map.cu -- streaming calibration kernels for sm_100a (no common mode):
//   K1 epix10ka  : gain-bit decode + pedestal + gain + mask          (calib_epix10ka_any)
//   K2 jungfrau  : gain-bit decode + pedestal + offset + gain + mask (calib_jungfrau,     BASELINE.md 2.3)
//   K5 masks     : np.where(keep, data, 0) of producer.py:92-95, folded into the epilogues
//   K4 assemble  : det.image.  Raw -> image directly by SCATTER through the forward map (image_scatter_kernel:
//                  the map kernels' ownership, every pixel stored where it is the last writer of its target; gaps zeroed by
//                  img_gap_fill_kernel); from an already calibrated batch by gather through the inverse map (image_gather_kernel);
//                  image_fused_kernel, the output-driven raw -> image gather, stays selectable as the cross-check
//   decode       : gain-mode index per pixel, for bit-exact checking of the decode
//
// Design: HBM-bound streaming map.  Work is pixel-tile-major, frames-inner: a thread owns a few
// pixels, loads their per-run constants ONCE into registers and then walks the frames of the batch,
// so HBM traffic per frame is raw-in (2 B/px) + float-out (4 B/px); the constants cost
// size/frames_per_block per frame and come from L2 when they fit (epix10ka 36.8 MB).  Raw loads and
// output stores carry streaming (evict-first) hints so they do not push the constants out of L2.
// Every float op is an explicit round-to-nearest intrinsic: NumPy does subtract, multiply, multiply
// as separate ufunc passes, so no FMA contraction and no re-association is allowed.
#include "psr_kernels.cuh"

namespace {

// tuning knobs (profiles/ holds the sweep that chose the defaults)
#ifndef PSR_MAP_THREADS
#define PSR_MAP_THREADS 256
#endif
#ifndef PSR_MAP_UNROLL
#define PSR_MAP_UNROLL 8
#endif
#ifndef PSR_MAP_MINB
#define PSR_MAP_MINB 4
#endif
#ifndef PSR_MAP_GROUPS      /* 4-pixel groups per thread in the epix10ka map kernel */
#define PSR_MAP_GROUPS 1
#endif
#ifndef PSR_MAP_LD          /* raw loads: 0 ld.global.nc, 1 .cs (evict-first), 2 .nc.L1::no_allocate, 3 .cv */
#define PSR_MAP_LD 1
#endif
#ifndef PSR_MAP_ST          /* out stores: 0 plain, 1 .cs, 2 .wt, 3 .L1::no_allocate */
#define PSR_MAP_ST 1
#endif
#ifndef PSR_JF_THREADS
#define PSR_JF_THREADS 128
#endif
#ifndef PSR_JF_UNROLL
#define PSR_JF_UNROLL 4
#endif
constexpr int MAP_THREADS = PSR_MAP_THREADS;
constexpr int JF_THREADS = PSR_JF_THREADS;
constexpr int MAP_UNROLL = PSR_MAP_UNROLL;       // frames in flight per thread

__device__ __forceinline__ uint2 ld_raw(const uint2* p) {
#if PSR_MAP_LD == 0
    return __ldg(p);
#elif PSR_MAP_LD == 1
    return __ldcs(p);
#elif PSR_MAP_LD == 2
    uint2 r; asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p)); return r;
#else
    return __ldcv(p);
#endif
}
__device__ __forceinline__ void st_out(float4* p, const float4& v) {
#if PSR_MAP_ST == 0
    *p = v;
#elif PSR_MAP_ST == 1
    __stcs(p, v);
#elif PSR_MAP_ST == 2
    __stwt(p, v);
#else
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
#endif
}
template <int N> struct IntC { static constexpr int value = N; };

__device__ __forceinline__ float maskf(uint32_t flbyte) { return (flbyte & PSR_FL_MASK) ? 1.0f : 0.0f; }

// ---- one epix10ka pixel -------------------------------------------------------------------------
__device__ __forceinline__ float epix_px(uint32_t v, float p0, float g0, float p1, float g1, uint32_t fl) {
    const bool sw = (v & 0x4000u) != 0;                         // data bit 14 -> table entry
    const float x = (float)(v & 0x3FFFu);
    float o = __fmul_rn(__fmul_rn(__fsub_rn(x, sw ? p1 : p0), sw ? g1 : g0), maskf(fl));
    return (fl & PSR_FL_KEEP) ? o : 0.0f;
}

// a if p != 0 else b, as one SELP
__device__ __forceinline__ float selp(float a, float b, uint32_t p) {
    float r;
    asm("{ .reg .pred q; setp.ne.u32 q, %3, 0; selp.f32 %0, %1, %2, q; }" : "=f"(r) : "f"(a), "f"(b), "r"(p));
    return r;
}

// ---- one jungfrau pixel -------------------------------------------------------------------------
template <bool HAS_OFFS>
__device__ __forceinline__ float jf_px(uint32_t v, const float* p, const float* o, const float* g, uint32_t fl) {
    const uint32_t gb = v >> 14;                                 // 00 g0, 01 g1, 11 g2, 10 none
    float x = (float)(v & 0x3FFFu);
    // branch-free 4-way select on the two gain bits (ptxas turns the plain ternaries into branches):
    // bit0 picks within {g0|g1} and {none|g2}, bit1 picks between the pairs
    const uint32_t b0 = gb & 1u, b1 = gb >> 1;
    const float ped = selp(selp(p[2], 0.0f, b0), selp(p[1], p[0], b0), b1);
    const float gf  = selp(selp(g[2], 1.0f, b0), selp(g[1], g[0], b0), b1);
    float off = 0.0f;
    if (HAS_OFFS) off = selp(selp(o[2], 0.0f, b0), selp(o[1], o[0], b0), b1);
    x = __fsub_rn(x, ped);                                       // arrf[gr] -= peds[i][gr]   (x - 0 is exact)
    if (HAS_OFFS) x = __fsub_rn(x, off);                         // arrf -= offset, separately rounded
    float r = __fmul_rn(__fmul_rn(x, gf), maskf(fl));
    return (fl & PSR_FL_KEEP) ? r : 0.0f;
}

__device__ __forceinline__ float f4get(const float4& v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }

// Work of a map kernel = ntiles pixel tiles x nframes frames.  Two schedules:
//   grid      (PERSIST = false): CTA (x, y) owns tile x and the frame chunk [y*fpb, (y+1)*fpb)
//   persistent (PERSIST = true): the grid is exactly the number of CTAs the GPU holds at once (SMs x
//     resident CTAs) and the linearised (tile, frame) space is cut into equal contiguous shares, so all
//     SMs start together and finish together (no partial last wave); a CTA reloads the tile constants
//     only when its share crosses a tile boundary.
template <bool PERSIST, typename F>
__device__ __forceinline__ void map_schedule(F&& do_tile, int nframes, int fpb, int ntiles) {
    if (!PERSIST) {
        const int f_begin = blockIdx.y * fpb;
        do_tile((int)blockIdx.x, f_begin, min(nframes, f_begin + fpb));
    } else {
        const long long total = (long long)ntiles * nframes;
        long long w = total * blockIdx.x / gridDim.x;
        const long long w_end = total * (blockIdx.x + 1) / gridDim.x;
        while (w < w_end) {
            const int tile = (int)(w / nframes), f = (int)(w - (long long)tile * nframes);
            const int f_end = (int)min((long long)nframes, f + (w_end - w));
            do_tile(tile, f, f_end);
            w += f_end - f;
        }
    }
}

// =================================================================================================
// K1: epix10ka calib map.  A warp covers 64 groups of 4 pixels; lane L owns groups L and L+32 so
// that both the 8-byte raw loads and the 16-byte output stores of a warp instruction are contiguous.
// =================================================================================================
template <bool PERSIST>
__global__ void __launch_bounds__(MAP_THREADS, PSR_MAP_MINB)
epix_calib_map_kernel(const uint2* __restrict__ raw, float4* __restrict__ out,
                      const float4* __restrict__ ped0, const float4* __restrict__ gf0,
                      const float4* __restrict__ ped1, const float4* __restrict__ gf1,
                      const uint32_t* __restrict__ flags, int ngroups, int nframes, int fpb, int ntiles)
{
    constexpr int NG = PSR_MAP_GROUPS;                 // 4-pixel groups per thread
    const int lane = threadIdx.x & 31;
    auto do_tile = [&](int tile, int f_begin, int f_end) {
    const int wbase = (tile * MAP_THREADS + (threadIdx.x & ~31)) * NG;
    if (wbase + lane >= ngroups) return;

    float4 p0[NG], q0[NG], p1[NG], q1[NG]; uint32_t fl[NG]; int gi[NG]; bool ok[NG];
#pragma unroll
    for (int h = 0; h < NG; ++h) {
        const int g = wbase + lane + 32 * h;
        ok[h] = g < ngroups;
        gi[h] = ok[h] ? g : wbase + lane;
        p0[h] = __ldg(ped0 + gi[h]); q0[h] = __ldg(gf0 + gi[h]); p1[h] = __ldg(ped1 + gi[h]); q1[h] = __ldg(gf1 + gi[h]);
        fl[h] = __ldg(flags + gi[h]);
    }

    auto frames = [&](int f, auto nconst) {           // nconst frames starting at f, loads first, then math + stores
        constexpr int NF = decltype(nconst)::value;
        uint2 r[NF][NG];
#pragma unroll
        for (int u = 0; u < NF; ++u) {
            const size_t base = (size_t)(f + u) * ngroups;
#pragma unroll
            for (int h = 0; h < NG; ++h) r[u][h] = ld_raw(raw + base + gi[h]);
        }
#pragma unroll
        for (int u = 0; u < NF; ++u) {
            const size_t base = (size_t)(f + u) * ngroups;
#pragma unroll
            for (int h = 0; h < NG; ++h) {
                if (h > 0 && !ok[h]) break;
                const uint32_t a = r[u][h].x, b = r[u][h].y;
                float4 o;
                o.x = epix_px(a & 0xFFFFu, p0[h].x, q0[h].x, p1[h].x, q1[h].x, fl[h]);
                o.y = epix_px(a >> 16,     p0[h].y, q0[h].y, p1[h].y, q1[h].y, fl[h] >> 8);
                o.z = epix_px(b & 0xFFFFu, p0[h].z, q0[h].z, p1[h].z, q1[h].z, fl[h] >> 16);
                o.w = epix_px(b >> 16,     p0[h].w, q0[h].w, p1[h].w, q1[h].w, fl[h] >> 24);
                st_out(out + base + gi[h], o);
            }
        }
    };
    int f = f_begin;
    for (; f + MAP_UNROLL <= f_end; f += MAP_UNROLL) frames(f, IntC<MAP_UNROLL>{});
    for (; f < f_end; ++f) frames(f, IntC<1>{});
    };
    map_schedule<PERSIST>(do_tile, nframes, fpb, ntiles);
}

// =================================================================================================
// K2: jungfrau calib map.  One group of 4 pixels per thread (9 float4 of constants = 36 registers).
// =================================================================================================
template <bool HAS_OFFS, bool PERSIST>
__global__ void __launch_bounds__(JF_THREADS)
jf_calib_map_kernel(const uint2* __restrict__ raw, float4* __restrict__ out,
                    const float4* __restrict__ peds, const float4* __restrict__ offs, const float4* __restrict__ gfac,
                    const uint32_t* __restrict__ flags, int ngroups, int nframes, int fpb, int ntiles)
{
    auto do_tile = [&](int tile, int f_begin, int f_end) {
    const int g = tile * JF_THREADS + threadIdx.x;
    if (g >= ngroups) return;
    float P[4][3], O[4][3], G[4][3];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        const float4 p = __ldg(peds + (size_t)m * ngroups + g);
        const float4 q = __ldg(gfac + (size_t)m * ngroups + g);
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
        if (HAS_OFFS) o = __ldg(offs + (size_t)m * ngroups + g);
#pragma unroll
        for (int i = 0; i < 4; ++i) { P[i][m] = f4get(p, i); G[i][m] = f4get(q, i); O[i][m] = f4get(o, i); }
    }
    const uint32_t fl = __ldg(flags + g);

    auto frames = [&](int f, auto nconst) {
        constexpr int NF = decltype(nconst)::value;
        uint2 r[NF];
#pragma unroll
        for (int u = 0; u < NF; ++u) r[u] = ld_raw(raw + (size_t)(f + u) * ngroups + g);
#pragma unroll
        for (int u = 0; u < NF; ++u) {
            float4 o;
            o.x = jf_px<HAS_OFFS>(r[u].x & 0xFFFFu, P[0], O[0], G[0], fl);
            o.y = jf_px<HAS_OFFS>(r[u].x >> 16,     P[1], O[1], G[1], fl >> 8);
            o.z = jf_px<HAS_OFFS>(r[u].y & 0xFFFFu, P[2], O[2], G[2], fl >> 16);
            o.w = jf_px<HAS_OFFS>(r[u].y >> 16,     P[3], O[3], G[3], fl >> 24);
            st_out(out + (size_t)(f + u) * ngroups + g, o);
        }
    };
    int f = f_begin;
    for (; f + PSR_JF_UNROLL <= f_end; f += PSR_JF_UNROLL) frames(f, IntC<PSR_JF_UNROLL>{});
    for (; f < f_end; ++f) frames(f, IntC<1>{});
    };
    map_schedule<PERSIST>(do_tile, nframes, fpb, ntiles);
}

// =================================================================================================
// decode: gain-mode index per pixel
// =================================================================================================
__global__ void __launch_bounds__(MAP_THREADS)
decode_kernel(const uint16_t* __restrict__ raw, uint8_t* __restrict__ modes_out, const uint8_t* __restrict__ modes,
              int det_kind, int npix, size_t total)
{
    const size_t i = (size_t)blockIdx.x * MAP_THREADS + threadIdx.x;
    if (i >= total) return;
    const uint32_t v = raw[i];
    uint32_t m;
    if (det_kind == 0) {
        const uint32_t mm = modes[i % npix];
        m = (v & 0x4000u) ? (mm >> 4) : (mm & 0xFu);
        if (m == 0xFu) m = 255u;
    } else {
        const uint32_t gb = v >> 14;
        m = gb == 0 ? 0u : gb == 1 ? 1u : gb == 3 ? 2u : 255u;
    }
    modes_out[i] = (uint8_t)m;
}

// =================================================================================================
// K4: image assembly as a gather.  Output tile = 64 px x 16 rows per CTA (thread = 4 px of one row)
// so that panels rotated by 90/270 degrees, whose image rows run along detector columns, still use
// whole 32-byte sectors of the source within one CTA.  VEC = W % 4 == 0 (16-byte stores).
// =================================================================================================
constexpr int IMG_TW = 64, IMG_TH = 16;

struct ImgGeom { int H, W; };

// VEC (W % 4 == 0): every thread's 4 pixels are one aligned 16-byte store.  Otherwise the widest store the actual
// address allows: image widths are arbitrary (1566 in the synthetic epix10k2M geometry: rows alternate between
// 16- and 8-byte alignment), and scalar 4-byte stores halve the kernel.
template <bool VEC>
__device__ __forceinline__ void img_store(float* __restrict__ img, size_t off, int ncol, const float (&v)[4]) {
    float* p = img + off;
    if (VEC) { __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3])); return; }
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    if (ncol == 4 && (a & 15) == 0) {
        __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
    } else if ((a & 7) == 0) {
        if (ncol >= 2) __stcs(reinterpret_cast<float2*>(p), make_float2(v[0], v[1])); else if (ncol == 1) __stcs(p, v[0]);
        if (ncol == 4) __stcs(reinterpret_cast<float2*>(p) + 1, make_float2(v[2], v[3])); else if (ncol == 3) __stcs(p + 2, v[2]);
    } else {
        if (ncol > 0) __stcs(p, v[0]);
        if (ncol > 1) __stcs(p + 1, v[1]);
        if (ncol > 2) __stcs(p + 2, v[2]);
        if (ncol > 3) __stcs(p + 3, v[3]);
    }
}

// +1: src[0..3] ascending, contiguous, src[0] % 4 == 0; -1: descending with src[3] % 4 == 0; 0: anything else.
// npix % 4 == 0, so a run that is aligned within a frame is aligned in every frame.
__device__ __forceinline__ int src_run(const int (&src)[4], int npix) {
    if (src[0] >= 0 && (src[0] & 3) == 0 && src[1] == src[0] + 1 && src[2] == src[0] + 2 && src[3] == src[0] + 3 && src[3] < npix) return 1;
    if (src[3] >= 0 && (src[3] & 3) == 0 && src[2] == src[3] + 1 && src[1] == src[3] + 2 && src[0] == src[3] + 3 && src[0] < npix) return -1;
    return 0;
}

// Tile bookkeeping shared by the two image kernels.  A CTA owns a 64 x 16 output tile; every thread STORES
// 4 consecutive pixels of one output row (coalesced 16-byte stores).  Which pixels a thread LOADS depends on
// the tile: on unrotated / 180-degree panels consecutive output x are consecutive detector pixels, so a thread
// loads what it stores; on panels rotated by 90 / 270 degrees consecutive output y are consecutive detector
// pixels, so the tile is walked column-wise for the loads (4 consecutive y per thread = one vector load) and
// transposed through shared memory.  The CTA votes once which way its tile goes.
constexpr int IMG_TS = IMG_TW + 1;                     // padded shared row stride

struct ImgTile {
    int x0, y, ncol; bool store_ok; size_t p0;         // store role
    int src[4]; int run; bool transposed; int tx, ty0; // load role
};

__device__ __forceinline__ int inv_at(const int32_t* __restrict__ inv, const uint8_t* __restrict__ keep, int H, int W, int y, int x) {
    if (y >= H || x >= W) return -1;
    const size_t p = (size_t)y * W + x;
    if (keep != nullptr && !__ldg(keep + p)) return -1;
    return __ldg(inv + p);
}

__device__ __forceinline__ void img_tile_setup(ImgTile& t, const int32_t* __restrict__ inv, const uint8_t* __restrict__ keep,
                                               int H, int W, int npix) {
    const int tid = threadIdx.x;
    const int bx = blockIdx.x * IMG_TW, by = blockIdx.y * IMG_TH;
    t.x0 = bx + (tid & 15) * 4; t.y = by + (tid >> 4);
    t.store_ok = t.y < H && t.x0 < W;
    t.ncol = t.store_ok ? min(4, W - t.x0) : 0;
    t.p0 = (size_t)t.y * W + t.x0;
    int rowsrc[4], colsrc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) rowsrc[i] = inv_at(inv, keep, H, W, t.y, t.x0 + i);
    t.tx = tid >> 2; t.ty0 = (tid & 3) * 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) colsrc[i] = inv_at(inv, keep, H, W, by + t.ty0 + i, bx + t.tx);
    const int rrun = src_run(rowsrc, npix), crun = src_run(colsrc, npix);
    const int nrow = __syncthreads_count(rrun != 0), ncolv = __syncthreads_count(crun != 0);
    t.transposed = ncolv > nrow;
#pragma unroll
    for (int i = 0; i < 4; ++i) t.src[i] = t.transposed ? colsrc[i] : rowsrc[i];
    t.run = t.transposed ? crun : rrun;
}

// v[u][0..3] = this thread's LOADED values of frames f0+u (u < nf) -> image, through shared memory when the
// tile is transposed (all frames of the group share two barriers)
#ifndef PSR_IMG_NU
#define PSR_IMG_NU 2
#endif
constexpr int IMG_NU = PSR_IMG_NU;                              // frames per group: loads of the whole group are in flight together

template <bool VEC>
__device__ __forceinline__ void img_tile_store(const ImgTile& t, float* __restrict__ tilebuf, float* __restrict__ img,
                                               size_t fimg, int f0, int nf, const float (&v)[IMG_NU][4]) {
    if (!t.transposed) {
        if (t.store_ok) {
#pragma unroll
            for (int u = 0; u < IMG_NU; ++u)
                if (u < nf) img_store<VEC>(img, (size_t)(f0 + u) * fimg + t.p0, t.ncol, v[u]);
        }
        return;
    }
#pragma unroll
    for (int u = 0; u < IMG_NU; ++u)
#pragma unroll
        for (int i = 0; i < 4; ++i) tilebuf[u * (IMG_TH * IMG_TS) + (t.ty0 + i) * IMG_TS + t.tx] = v[u][i];
    __syncthreads();
    if (t.store_ok) {
        const int r = threadIdx.x >> 4, cx = (threadIdx.x & 15) * 4;
#pragma unroll
        for (int u = 0; u < IMG_NU; ++u) {
            if (u < nf) {
                const float* tb = tilebuf + u * (IMG_TH * IMG_TS) + r * IMG_TS + cx;
                const float w[4] = {tb[0], tb[1], tb[2], tb[3]};
                img_store<VEC>(img, (size_t)(f0 + u) * fimg + t.p0, t.ncol, w);
            }
        }
    }
    __syncthreads();
}

// from an already calibrated batch (common mode on): out[f][p] = src < 0 ? 0 : calib[f][src]
template <bool VEC>
__global__ void __launch_bounds__(256)
image_gather_kernel(const float* __restrict__ calib, float* __restrict__ img, const int32_t* __restrict__ inv,
                    const uint8_t* __restrict__ keep, int H, int W, int npix, int nframes)
{
    __shared__ float tilebuf[IMG_NU * IMG_TH * IMG_TS];
    ImgTile t;
    img_tile_setup(t, inv, keep, H, W, npix);
    const size_t fimg = (size_t)H * W;
    // blockIdx.z owns a contiguous range of frames, walked in groups of IMG_NU
    const int per = (nframes + gridDim.z - 1) / gridDim.z;
    const int f_begin = blockIdx.z * per, f_end = min(nframes, f_begin + per);
    for (int f0 = f_begin; f0 < f_end; f0 += IMG_NU) {
        const int nf = min(IMG_NU, f_end - f0);
        float v[IMG_NU][4];
#pragma unroll
        for (int u = 0; u < IMG_NU; ++u) {
            const float* cf = calib + (size_t)(f0 + (u < nf ? u : 0)) * npix;
            if (t.run > 0)      { const float4 q = __ldg(reinterpret_cast<const float4*>(cf + t.src[0])); v[u][0] = q.x; v[u][1] = q.y; v[u][2] = q.z; v[u][3] = q.w; }
            else if (t.run < 0) { const float4 q = __ldg(reinterpret_cast<const float4*>(cf + t.src[3])); v[u][0] = q.w; v[u][1] = q.z; v[u][2] = q.y; v[u][3] = q.x; }
            else {
#pragma unroll
                for (int i = 0; i < 4; ++i) v[u][i] = t.src[i] >= 0 ? __ldg(cf + t.src[i]) : 0.0f;
            }
        }
        img_tile_store<VEC>(t, tilebuf, img, fimg, f0, nf, v);
    }
}

// fused raw -> image (no common mode): constants gathered once per loaded pixel, frames inner
template <int DET, bool HAS_OFFS, bool VEC>
__global__ void __launch_bounds__(256)
image_fused_kernel(const uint16_t* __restrict__ raw, float* __restrict__ img, PsrDev c, int nframes, int fpb)
{
    __shared__ float tilebuf[IMG_NU * IMG_TH * IMG_TS];
    ImgTile t;
    img_tile_setup(t, c.inv, c.keep_image, c.H, c.W, c.npix);
    const int f_begin = blockIdx.z * fpb;
    const int f_end = min(nframes, f_begin + fpb);
    uint32_t fl[4];
    float A[4][3], B[4][3], G[4][3];   // epix: A = {ped0, ped1}, G = {gf0, gf1}; jungfrau: A peds, B offs, G gfac
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = t.src[i] < 0 ? 0 : t.src[i];
        fl[i] = __ldg(c.flags + q) | PSR_FL_KEEP;     // calib-layout keep mask does not apply to images
        if (DET == 0) {
            A[i][0] = __ldg(c.ped0 + q); A[i][1] = __ldg(c.ped1 + q); A[i][2] = 0.f;
            G[i][0] = __ldg(c.gf0 + q);  G[i][1] = __ldg(c.gf1 + q);  G[i][2] = 0.f;
            B[i][0] = B[i][1] = B[i][2] = 0.f;
        } else {
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                A[i][m] = __ldg(c.peds + (size_t)m * c.npix + q);
                G[i][m] = __ldg(c.gfac + (size_t)m * c.npix + q);
                B[i][m] = HAS_OFFS ? __ldg(c.offs + (size_t)m * c.npix + q) : 0.f;
            }
        }
    }
    const size_t fimg = (size_t)c.H * c.W;
    for (int f0 = f_begin; f0 < f_end; f0 += IMG_NU) {
        const int nf = min(IMG_NU, f_end - f0);
        uint32_t r[IMG_NU][4];
#pragma unroll
        for (int u = 0; u < IMG_NU; ++u) {
            const uint16_t* rf = raw + (size_t)(f0 + (u < nf ? u : 0)) * c.npix;
            if (t.run > 0)      { const uint2 q = __ldg(reinterpret_cast<const uint2*>(rf + t.src[0])); r[u][0] = q.x & 0xFFFFu; r[u][1] = q.x >> 16; r[u][2] = q.y & 0xFFFFu; r[u][3] = q.y >> 16; }
            else if (t.run < 0) { const uint2 q = __ldg(reinterpret_cast<const uint2*>(rf + t.src[3])); r[u][3] = q.x & 0xFFFFu; r[u][2] = q.x >> 16; r[u][1] = q.y & 0xFFFFu; r[u][0] = q.y >> 16; }
            else {
#pragma unroll
                for (int i = 0; i < 4; ++i) r[u][i] = t.src[i] >= 0 ? (uint32_t)__ldg(rf + t.src[i]) : 0u;
            }
        }
        float v[IMG_NU][4];
#pragma unroll
        for (int u = 0; u < IMG_NU; ++u)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float o;
                if (DET == 0) o = epix_px(r[u][i], A[i][0], G[i][0], A[i][1], G[i][1], fl[i]);
                else          o = jf_px<HAS_OFFS>(r[u][i], A[i], B[i], G[i], fl[i]);
                v[u][i] = t.src[i] >= 0 ? o : 0.0f;
            }
        img_tile_store<VEC>(t, tilebuf, img, fimg, f0, nf, v);
    }
}


// =================================================================================================
// K4s: raw -> assembled image, SOURCE-driven (no common mode).  The map kernels' structure -- a thread
// owns 4 pixels, holds their constants in registers and walks the frames of the batch -- with the image
// index of every owned pixel (PsrDev::fwd, the scatter side of the geometry) held in registers too, so
// the calibrated values go straight to their places: 12 + ~3 thread-instructions per pixel against 36 for
// the output-driven gather.  Which pixels a CTA and its warps own depends on how the segment lies in the image (see `rot` below), so that stores cover whole image-row pieces either way.
// Neighbouring CTAs run concurrently on neighbouring tiles and the same frames, so sectors shared across a
// tile border are completed in L2.  Gaps are zeroed by psr_launch_image_gap_fill.
// =================================================================================================
#ifndef PSR_IMGS_UNROLL
#define PSR_IMGS_UNROLL 8
#endif
#ifndef PSR_IMGS_MINB
#define PSR_IMGS_MINB 3
#endif

#ifndef PSR_IMGS_ROT_ST      /* scattered stores: 0 = .cs (evict-first), 1 = plain write-back */
#define PSR_IMGS_ROT_ST 1
#endif
// Scattered stores leave partly written sectors at tile borders until the neighbouring CTA writes the rest; evicted
// early (evict-first) they cost a read-modify-write in DRAM, so these stores keep the normal eviction priority.
__device__ __forceinline__ void imgs_st_at(float* base, int32_t t, float v)
{
#if PSR_IMGS_ROT_ST == 0
    asm volatile("{ .reg .u64 a; mad.wide.u32 a, %1, 4, %0; st.global.cs.f32 [a], %2; }" :: "l"(base), "r"(t), "f"(v) : "memory");
#else
    asm volatile("{ .reg .u64 a; mad.wide.u32 a, %1, 4, %0; st.global.f32 [a], %2; }" :: "l"(base), "r"(t), "f"(v) : "memory");
#endif
}

// 4 consecutive floats at any 4-byte alignment
__device__ __forceinline__ void imgs_store_run4(float* p, float o0, float o1, float o2, float o3)
{
    const unsigned a = (unsigned)(reinterpret_cast<uintptr_t>(p) >> 2) & 3u;
    if (a == 0)      { __stcs(reinterpret_cast<float4*>(p), make_float4(o0, o1, o2, o3)); }
    else if (a == 2) { __stcs(reinterpret_cast<float2*>(p), make_float2(o0, o1)); __stcs(reinterpret_cast<float2*>(p + 2), make_float2(o2, o3)); }
    else if (a == 1) { __stcs(p, o0); __stcs(reinterpret_cast<float2*>(p + 1), make_float2(o1, o2)); __stcs(p + 3, o3); }
    else             { __stcs(p, o0); __stcs(reinterpret_cast<float2*>(p + 1), make_float2(o1, o2)); __stcs(p + 3, o3); }    // a == 3: p + 1 is 16-byte aligned
}

template <int DET, bool HAS_OFFS>
__global__ void __launch_bounds__(256, PSR_IMGS_MINB)
image_scatter_kernel(const uint2* __restrict__ raw, float* __restrict__ img, PsrDev c, int nframes, int fpb)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int ngroups = c.npix / 4, gps = ngroups / c.nseg;              // 4-pixel groups in the batch frame / in a segment
    const int bps = gps / 256;                                           // CTAs per segment (256 groups each)
    // blocks run in the order of their place in the IMAGE (PsrDev::blk_order): CTAs resident together then write neighbouring
    // image rows, whichever segments those come from, instead of half-filling DRAM pages that another segment completes later
    const int blk = c.blk_order != nullptr ? __ldg(c.blk_order + blockIdx.x) : (int)blockIdx.x;
    const int seg = blk / bps, lb = blk - seg * bps;
    // How the segment lies in the image decides who owns what (a speed matter only: any ownership is correct).
    // Unrotated / 180 degrees: the calib map kernel's ownership -- the CTA takes 1024 consecutive pixels, a warp 128 of them:
    // 256 contiguous bytes of raw and 512 contiguous bytes of image per warp instruction, whole detector rows per CTA.
    // Rotated by 90 / 270 degrees (detector rows run along image columns): four CTAs share a tile of 32 rows x 128 pixels, each
    // takes 32 columns of it and a warp 4 pixels of 32 consecutive ROWS, so each of its four scalar stores is 32 consecutive
    // floats of one image row; the 8-byte raw loads of a warp then touch 32 sectors a quarter each (the CTA's other warps use
    // the rest; loaded through L1).  Measured: 2.4x the L2 reads and 1.9x the DRAM reads of the unrotated pattern -- and still
    // faster than loading row-wise and exchanging through shared memory (319 vs 291 us on an all-rotated geometry).
    bool rot;
    {
        const int4 pt = __ldg(reinterpret_cast<const int4*>(c.fwd) + (size_t)seg * gps + (gps >> 1));     // a group in the middle of the segment
        const int d = pt.y - pt.x;
        rot = pt.x >= 0 && pt.y >= 0 && d != 1 && d != -1;
    }
    int g;
    if (rot) {
        const int tiles_x = c.cols / 128, t = lb >> 2, pass = lb & 3;
        const int ty = t / tiles_x, tx = t - ty * tiles_x;
        g = ((seg * c.rows + ty * 32 + lane) * c.cols + tx * 128 + pass * 32 + w * 4) >> 2;
    } else {
        g = seg * gps + lb * 256 + tid;
    }
    const size_t fimg = (size_t)c.H * c.W;

    // per-run constants of the 4 owned pixels, in registers for the whole batch
    float P[4][DET == 0 ? 2 : 3], G[4][DET == 0 ? 2 : 3], O[4][3];
    if (DET == 0) {
        const float4 p0 = __ldg(reinterpret_cast<const float4*>(c.ped0) + g), q0 = __ldg(reinterpret_cast<const float4*>(c.gf0) + g);
        const float4 p1 = __ldg(reinterpret_cast<const float4*>(c.ped1) + g), q1 = __ldg(reinterpret_cast<const float4*>(c.gf1) + g);
#pragma unroll
        for (int i = 0; i < 4; ++i) { P[i][0] = f4get(p0, i); P[i][1] = f4get(p1, i); G[i][0] = f4get(q0, i); G[i][1] = f4get(q1, i); }
    } else {
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            const float4 p = __ldg(reinterpret_cast<const float4*>(c.peds) + (size_t)m * ngroups + g);
            const float4 q = __ldg(reinterpret_cast<const float4*>(c.gfac) + (size_t)m * ngroups + g);
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (HAS_OFFS) o = __ldg(reinterpret_cast<const float4*>(c.offs) + (size_t)m * ngroups + g);
#pragma unroll
            for (int i = 0; i < 4; ++i) { P[i][m] = f4get(p, i); G[i][m] = f4get(q, i); O[i][m] = f4get(o, i); }
        }
    }
    // image mode: the calib-layout keep mask does not apply (ref: producer.py:92-97 masks the 2-D image)
    const uint32_t fl = __ldg(reinterpret_cast<const uint32_t*>(c.flags) + g) | 0x08080808u;
    const int4 t4 = __ldg(reinterpret_cast<const int4*>(c.fwd) + g);
    const int32_t t[4] = {t4.x, t4.y, t4.z, t4.w};
    const bool all = t[0] >= 0 && t[1] >= 0 && t[2] >= 0 && t[3] >= 0;
    const bool asc = all && t[1] == t[0] + 1 && t[2] == t[0] + 2 && t[3] == t[0] + 3;
    const bool desc = all && t[1] == t[0] - 1 && t[2] == t[0] - 2 && t[3] == t[0] - 3;
    const int kind = asc ? 1 : desc ? 2 : all ? 3 : 0;

    const int f_begin = blockIdx.y * fpb, f_end = min(nframes, f_begin + fpb);
    auto frames = [&](int f, auto nconst) {
        constexpr int NF = decltype(nconst)::value;
        uint2 r[NF];
#pragma unroll
        for (int u = 0; u < NF; ++u) r[u] = rot ? __ldg(raw + (size_t)(f + u) * ngroups + g) : ld_raw(raw + (size_t)(f + u) * ngroups + g);
#pragma unroll
        for (int u = 0; u < NF; ++u) {
            float o[4];
            if (DET == 0) {
                o[0] = epix_px(r[u].x & 0xFFFFu, P[0][0], G[0][0], P[0][1], G[0][1], fl);
                o[1] = epix_px(r[u].x >> 16,     P[1][0], G[1][0], P[1][1], G[1][1], fl >> 8);
                o[2] = epix_px(r[u].y & 0xFFFFu, P[2][0], G[2][0], P[2][1], G[2][1], fl >> 16);
                o[3] = epix_px(r[u].y >> 16,     P[3][0], G[3][0], P[3][1], G[3][1], fl >> 24);
            } else {
                o[0] = jf_px<HAS_OFFS>(r[u].x & 0xFFFFu, P[0], O[0], G[0], fl);
                o[1] = jf_px<HAS_OFFS>(r[u].x >> 16,     P[1], O[1], G[1], fl >> 8);
                o[2] = jf_px<HAS_OFFS>(r[u].y & 0xFFFFu, P[2], O[2], G[2], fl >> 16);
                o[3] = jf_px<HAS_OFFS>(r[u].y >> 16,     P[3], O[3], G[3], fl >> 24);
            }
            float* imgf = img + (size_t)(f + u) * fimg;
            if (kind == 1) imgs_store_run4(imgf + t[0], o[0], o[1], o[2], o[3]);
            else if (kind == 2) imgs_store_run4(imgf + t[3], o[3], o[2], o[1], o[0]);
            else if (kind == 3) {
#pragma unroll
                for (int i = 0; i < 4; ++i) imgs_st_at(imgf, t[i], o[i]);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (t[i] >= 0) imgs_st_at(imgf, t[i], o[i]);
                    else if (t[i] != INT32_MIN) imgs_st_at(imgf, ~t[i], 0.0f);      // the image keep mask drops this target
                }
            }
        }
    };
    int f = f_begin;
    for (; f + PSR_IMGS_UNROLL <= f_end; f += PSR_IMGS_UNROLL) frames(f, IntC<PSR_IMGS_UNROLL>{});
    for (; f < f_end; ++f) frames(f, IntC<1>{});
}

// image indices no detector pixel maps to: zero in every frame (ref: img_from_pixel_arrays starts from np.zeros)
__global__ void __launch_bounds__(256) img_gap_fill_kernel(float* __restrict__ img, const int32_t* __restrict__ gaps, int ngaps, size_t fimg, int nframes)
{
    for (int f = blockIdx.y; f < nframes; f += gridDim.y) {
        float* imgf = img + (size_t)f * fimg;
        for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < ngaps; g += gridDim.x * blockDim.x) __stcs(imgf + __ldg(gaps + g), 0.0f);
    }
}

}  // namespace

// ---- launchers --------------------------------------------------------------------------------
namespace {
template <typename K>
int resident_ctas(K kernel, int threads) {          // SMs x CTAs resident per SM for this kernel on the current device
    int dev = 0, sms = 0, per = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kernel, threads, 0);
    return sms * (per > 0 ? per : 1);
}
}  // namespace

// fpb > 0: grid schedule with that many frames per CTA; fpb == 0: grid schedule, whole batch (<= 64 frames) per CTA --
// the measured best (profiles/README.md); fpb < 0: persistent balanced schedule
cudaError_t psr_launch_calib_map(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_out,
                                 int fpb, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    const int ngroups = c.npix / 4;
    const bool persist = fpb < 0;
    if (fpb == 0) fpb = nframes < 64 ? nframes : 64;
    if (fpb < 0 || fpb > nframes) fpb = nframes;
    const int ysplit = (nframes + fpb - 1) / fpb;
    const uint2* raw = reinterpret_cast<const uint2*>(d_raw);
    float4* out = reinterpret_cast<float4*>(d_out);
    const uint32_t* fl = reinterpret_cast<const uint32_t*>(c.flags);
    if (c.det_kind == 0) {
        const int ntiles = (ngroups + PSR_MAP_GROUPS * MAP_THREADS - 1) / (PSR_MAP_GROUPS * MAP_THREADS);
        const float4 *a0 = reinterpret_cast<const float4*>(c.ped0), *a1 = reinterpret_cast<const float4*>(c.gf0);
        const float4 *a2 = reinterpret_cast<const float4*>(c.ped1), *a3 = reinterpret_cast<const float4*>(c.gf1);
        if (persist) {
            const int nres = resident_ctas(epix_calib_map_kernel<true>, MAP_THREADS);       // per launch: the current device's, no shared state
            const long long total = (long long)ntiles * nframes;
            const int grid = (int)(total < nres ? total : nres);
            epix_calib_map_kernel<true><<<grid, MAP_THREADS, 0, st>>>(raw, out, a0, a1, a2, a3, fl, ngroups, nframes, fpb, ntiles);
        } else {
            epix_calib_map_kernel<false><<<dim3(ntiles, ysplit), MAP_THREADS, 0, st>>>(raw, out, a0, a1, a2, a3, fl, ngroups, nframes, fpb, ntiles);
        }
    } else {
        const int ntiles = (ngroups + JF_THREADS - 1) / JF_THREADS;
        const float4 *p = reinterpret_cast<const float4*>(c.peds), *o = reinterpret_cast<const float4*>(c.offs);
        const float4 *g = reinterpret_cast<const float4*>(c.gfac);
        const long long total = (long long)ntiles * nframes;
#define PSR_JF_LAUNCH(OFFS)                                                                                          \
        if (persist) {                                                                                               \
            const int nres = resident_ctas(jf_calib_map_kernel<OFFS, true>, JF_THREADS);                            \
            const int grid = (int)(total < nres ? total : nres);                                                     \
            jf_calib_map_kernel<OFFS, true><<<grid, JF_THREADS, 0, st>>>(raw, out, p, o, g, fl, ngroups, nframes, fpb, ntiles); \
        } else {                                                                                                     \
            jf_calib_map_kernel<OFFS, false><<<dim3(ntiles, ysplit), JF_THREADS, 0, st>>>(raw, out, p, o, g, fl, ngroups, nframes, fpb, ntiles); \
        }
        if (c.offs != nullptr) { PSR_JF_LAUNCH(true) } else { PSR_JF_LAUNCH(false) }
#undef PSR_JF_LAUNCH
    }
    return cudaGetLastError();
}

cudaError_t psr_launch_decode(const PsrDev& c, const uint16_t* d_raw, int nframes, uint8_t* d_modes, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    const size_t total = (size_t)nframes * c.npix;
    const unsigned blocks = (unsigned)((total + MAP_THREADS - 1) / MAP_THREADS);
    decode_kernel<<<blocks, MAP_THREADS, 0, st>>>(d_raw, d_modes, c.modes, c.det_kind, c.npix, total);
    return cudaGetLastError();
}

cudaError_t psr_launch_image_gather(const PsrDev& c, const float* d_calib, int nframes, float* d_img, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    // the per-thread set-up (inverse-map lookups, run detection, tile vote) is amortised over the frames of one z slice
#ifndef PSR_IMG_GZ
#define PSR_IMG_GZ 8
#endif
    dim3 grid((c.W + IMG_TW - 1) / IMG_TW, (c.H + IMG_TH - 1) / IMG_TH, (nframes + PSR_IMG_GZ - 1) / PSR_IMG_GZ);
    if (c.W % 4 == 0)
        image_gather_kernel<true><<<grid, 256, 0, st>>>(d_calib, d_img, c.inv, c.keep_image, c.H, c.W, c.npix, nframes);
    else
        image_gather_kernel<false><<<grid, 256, 0, st>>>(d_calib, d_img, c.inv, c.keep_image, c.H, c.W, c.npix, nframes);
    return cudaGetLastError();
}

cudaError_t psr_launch_image_fused(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_img,
                                   int fpb, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    if (fpb <= 0 || fpb > nframes) fpb = nframes;
    dim3 grid((c.W + IMG_TW - 1) / IMG_TW, (c.H + IMG_TH - 1) / IMG_TH, (nframes + fpb - 1) / fpb);
    const bool vec = (c.W % 4 == 0);
#define PSR_IMG_LAUNCH(DET, OFFS, VEC) image_fused_kernel<DET, OFFS, VEC><<<grid, 256, 0, st>>>(d_raw, d_img, c, nframes, fpb)
    if (c.det_kind == 0) { if (vec) PSR_IMG_LAUNCH(0, false, true); else PSR_IMG_LAUNCH(0, false, false); }
    else if (c.offs != nullptr) { if (vec) PSR_IMG_LAUNCH(1, true, true); else PSR_IMG_LAUNCH(1, true, false); }
    else { if (vec) PSR_IMG_LAUNCH(1, false, true); else PSR_IMG_LAUNCH(1, false, false); }
#undef PSR_IMG_LAUNCH
    return cudaGetLastError();
}

cudaError_t psr_launch_image_gap_fill(const PsrDev& c, int nframes, float* d_img, cudaStream_t st)
{
    if (nframes <= 0 || c.ngaps <= 0) return cudaSuccess;
    dim3 g((unsigned)((c.ngaps + 256 * 8 - 1) / (256 * 8)), (unsigned)(nframes < 65535 ? nframes : 65535));
    img_gap_fill_kernel<<<g, 256, 0, st>>>(d_img, c.gaps, c.ngaps, (size_t)c.H * c.W, nframes);
    return cudaGetLastError();
}

cudaError_t psr_launch_image_scatter(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_img,
                                     int fpb, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    if (c.fwd == nullptr || c.rows % 32 != 0 || c.cols % 128 != 0) return cudaErrorInvalidValue;
    if (fpb <= 0 || fpb > nframes) fpb = nframes > 64 ? 64 : nframes;
    dim3 grid((unsigned)(c.npix / 1024), (unsigned)((nframes + fpb - 1) / fpb));       // rows % 32 == 0 and cols % 128 == 0: a segment is a whole number of CTAs
    const uint2* r = reinterpret_cast<const uint2*>(d_raw);
    if (c.det_kind == 0) image_scatter_kernel<0, false><<<grid, 256, 0, st>>>(r, d_img, c, nframes, fpb);
    else if (c.offs != nullptr) image_scatter_kernel<1, true><<<grid, 256, 0, st>>>(r, d_img, c, nframes, fpb);
    else image_scatter_kernel<1, false><<<grid, 256, 0, st>>>(r, d_img, c, nframes, fpb);
    return cudaGetLastError();
}

