// capi.cu -- the C ABI declared in include/psana_ray_cuda.h: context, per-run constant tables,
// launch dispatch, the host-fed chunked pipeline, pinned/IPC helpers.  No torch types, no exceptions
// across the boundary.
#include "../../include/psana_ray_cuda.h"
#include "psr_kernels.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <mutex>
#include <new>
#include <string>
#include <utility>
#include <vector>

// Scratch of psr_calib_image_noalloc: lives in the library image, so no call ever allocates it.
// Host code must not pass this name to a kernel -- on the host it is only a stand-in for the symbol;
// the device address comes from cudaGetSymbolAddress, looked up once per device (see the entry point).
static __device__ __align__(256) unsigned char g_noalloc_scratch[PSR_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_err = buf;
    return code;
}

#define PSR_CUDA(call)                                                                            \
    do { cudaError_t e_ = (call);                                                                 \
         if (e_ != cudaSuccess) return fail(PSR_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct DeviceGuard {            // binds the context's device for one call, restores the caller's
    int prev = -1; bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; }
        if (prev != dev) ok = (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// epix10ka decode of cbits = (cfg & 0o34) | d14 << 5, in psana's np.select order (first match wins)
inline int epix_mode_of_cbits(unsigned cb) {
    if ((cb & 28u) == 28u) return 0;   // FH
    if ((cb & 28u) == 12u) return 1;   // FM
    if ((cb & 12u) == 8u)  return 2;   // FL
    if ((cb & 60u) == 16u) return 3;   // AHL-H
    if ((cb & 60u) == 0u)  return 4;   // AML-M
    if ((cb & 60u) == 48u) return 5;   // AHL-L
    if ((cb & 60u) == 32u) return 6;   // AML-L
    return 15;                         // no map: ped 0, factor 1
}

inline float gfac_of(float g) { return g != 0.0f ? 1.0f / g : 0.0f; }   // divide_protected(1, gain)

struct HostPipe {
    int chunk = 8, depth = 4;
    cudaStream_t s_in = nullptr, s_comp = nullptr, s_out = nullptr;
    std::vector<uint16_t*> d_raw; std::vector<float*> d_out;
    std::vector<cudaEvent_t> ev_in, ev_comp, ev_out;
    size_t raw_elems = 0, out_elems = 0;      // per slot
    long long seq = 0;                        // chunks issued so far (across calls: a call may still be in flight when the next begins)
};

}  // namespace

struct psr_ctx {
    int device = 0;
    PsrDev dev{};
    bool have_constants = false, have_geometry = false;
    // host mirrors needed to rebuild the flag bytes
    std::vector<uint8_t> h_mode0, h_mode1, h_mask, h_keep_calib;
    uint32_t elig_mask = 0;
    // device allocations
    void* d_consts = nullptr; size_t consts_bytes = 0;
    uint8_t* d_flags = nullptr;
    int32_t* d_inv = nullptr;
    int32_t* d_blk_order = nullptr;
    int32_t* d_fwd = nullptr; int32_t* d_gaps = nullptr;     // scatter side of the geometry (common mode fused with assembly)
    std::vector<int32_t> h_fwd;                               // per source pixel: winning image index or INT32_MIN, before the image keep mask
    uint8_t* d_keep_image = nullptr;
    float* d_scratch = nullptr; size_t scratch_elems = 0;
    cudaStream_t s_gap = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;    // gap fill runs beside the scatter kernels
    int image_gather = 0;          // 1: raw -> image through the output-driven gather kernel instead of the scatter kernel
    int frames_per_block = 0;      // map kernels: 0 = whole batch (<= 64 frames) per CTA on a plain grid, > 0 = that many frames per CTA, < 0 = persistent balanced schedule
    HostPipe pipe;
};

namespace {

int rebuild_flags(psr_ctx* c) {
    const int n = c->dev.npix;
    std::vector<uint8_t> fl((size_t)n);
    const bool epix = c->dev.det_kind == PSR_DET_EPIX10KA;
    for (int i = 0; i < n; ++i) {
        const unsigned m = c->h_mask.empty() ? 1u : c->h_mask[i];
        const unsigned k = c->h_keep_calib.empty() ? 1u : (c->h_keep_calib[i] ? 1u : 0u);
        unsigned f = (m ? PSR_FL_MASK : 0u) | (k ? PSR_FL_KEEP : 0u);
        if (epix && !c->h_mode0.empty()) {
            const unsigned m0 = c->h_mode0[i], m1 = c->h_mode1[i];
            if (m && m0 < 7 && ((c->elig_mask >> m0) & 1u)) f |= PSR_FL_ELIG0;
            if (m && m1 < 7 && ((c->elig_mask >> m1) & 1u)) f |= PSR_FL_ELIG1;
        }
        fl[i] = (uint8_t)f;
    }
    PSR_CUDA(cudaMemcpy(c->d_flags, fl.data(), (size_t)n, cudaMemcpyHostToDevice));
    c->dev.has_keep = c->h_keep_calib.empty() ? 0 : 1;
    return PSR_OK;
}

// The mask is folded into the gain factors: a masked pixel's factor becomes factor * 0 = +-0, so that x * factor is already
// the (x * factor) * 0 psana computes (same sign of zero, NaN stays NaN) and kernels may skip the mask multiply; kernels that
// still multiply get the same bits.  The two differ only if x * factor overflows float32 on a masked pixel (inf * 0 = NaN
// there, +-0 here): such constants are refused instead of silently calibrated differently.
inline bool fold_overflows(float gf, float ped, float off) {
    if (!std::isfinite(gf) || !std::isfinite(ped) || !std::isfinite(off)) return false;     // non-finite inputs give NaN either way
    return (double)std::fabs(gf) * (16383.0 + std::fabs((double)ped) + std::fabs((double)off)) >= 1.0e38;
}

// device copy of the forward map with the image keep mask folded in (~index = "store 0 there")
int upload_fwd(psr_ctx* c, const uint8_t* h_keep_image) {
    const size_t n = c->h_fwd.size();
    if (n == 0) return PSR_OK;
    if (!c->d_fwd) PSR_CUDA(cudaMalloc(&c->d_fwd, n * sizeof(int32_t)));
    if (!h_keep_image) {
        PSR_CUDA(cudaMemcpy(c->d_fwd, c->h_fwd.data(), n * sizeof(int32_t), cudaMemcpyHostToDevice));
    } else {
        std::vector<int32_t> f(c->h_fwd);
        for (size_t i = 0; i < n; ++i) if (f[i] >= 0 && !h_keep_image[f[i]]) f[i] = ~f[i];
        PSR_CUDA(cudaMemcpy(c->d_fwd, f.data(), n * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    c->dev.fwd = c->d_fwd;
    return PSR_OK;
}

void free_pipe(HostPipe& p) {
    for (auto q : p.d_raw) if (q) cudaFree(q);
    for (auto q : p.d_out) if (q) cudaFree(q);
    for (auto e : p.ev_in) if (e) cudaEventDestroy(e);
    for (auto e : p.ev_comp) if (e) cudaEventDestroy(e);
    for (auto e : p.ev_out) if (e) cudaEventDestroy(e);
    p.d_raw.clear(); p.d_out.clear(); p.ev_in.clear(); p.ev_comp.clear(); p.ev_out.clear();
    p.raw_elems = p.out_elems = 0; p.seq = 0;
}

size_t out_frame_elems(const psr_ctx* c, int out_kind) {
    return out_kind == PSR_OUT_IMAGE ? (size_t)c->dev.H * c->dev.W : (size_t)c->dev.npix;
}

int ensure_scratch(psr_ctx* c, size_t elems) {
    if (c->scratch_elems >= elems) return PSR_OK;
    if (c->d_scratch) { PSR_CUDA(cudaFree(c->d_scratch)); c->d_scratch = nullptr; c->scratch_elems = 0; }
    PSR_CUDA(cudaMalloc(&c->d_scratch, elems * sizeof(float)));
    c->scratch_elems = elems;
    return PSR_OK;
}

// Zero the image pixels no detector pixel maps to, on a side stream forked from `st`: the scatter kernels write the
// other pixels, so the two need no order between them, only both after what `st` held before and both before what follows.
// fork marks the point on `st` BEFORE the scatter kernel; join, called after the scatter kernel is launched, launches the
// fill behind that point and makes `st` wait for it.  Launched second, the fill takes the SMs the scatter kernel leaves
// (its last, partly filled wave) instead of holding it up: measured 13-18 us per 64 epix10k2M frames against fill-first.
int gap_fill_fork(psr_ctx* c, cudaStream_t st) {
    if (c->dev.ngaps <= 0) return PSR_OK;
    PSR_CUDA(cudaEventRecord(c->ev_fork, st));
    return PSR_OK;
}
int gap_fill_join(psr_ctx* c, int nframes, float* d_img, cudaStream_t st) {
    if (c->dev.ngaps <= 0) return PSR_OK;
    PSR_CUDA(cudaStreamWaitEvent(c->s_gap, c->ev_fork, 0));
    PSR_CUDA(psr_launch_image_gap_fill(c->dev, nframes, d_img, c->s_gap));
    PSR_CUDA(cudaEventRecord(c->ev_join, c->s_gap));
    PSR_CUDA(cudaStreamWaitEvent(st, c->ev_join, 0));
    return PSR_OK;
}

int launch_batch(psr_ctx* c, const uint16_t* d_raw, int nframes, float* d_out, int out_kind, cudaStream_t st) {
    const PsrDev& d = c->dev;
    const int fpb = c->frames_per_block;      // see psr_ctx::frames_per_block
    if (out_kind == PSR_OUT_CALIB) {
        if (d.cm_mode == 0) PSR_CUDA(psr_launch_calib_map(d, d_raw, nframes, d_out, fpb, st));
        else                PSR_CUDA(psr_launch_common_mode(d, d_raw, nframes, d_out, st));
    } else {
        if (d.cm_mode == 0) {
            if (d.fwd != nullptr && !c->image_gather) {
                int rc = gap_fill_fork(c, st);
                if (rc != PSR_OK) return rc;
                PSR_CUDA(psr_launch_image_scatter(d, d_raw, nframes, d_out, fpb, st));
                rc = gap_fill_join(c, nframes, d_out, st);
                if (rc != PSR_OK) return rc;
            } else PSR_CUDA(psr_launch_image_fused(d, d_raw, nframes, d_out, fpb, st));
        } else if (d.cm_mode == 4 && d.fwd != nullptr) {
            // bank medians only: the common-mode kernel writes each tile straight into the images (no scratch, no second pass)
            PsrDev t = d; t.ignore_keep = 1;
            int rc = gap_fill_fork(c, st);
            if (rc != PSR_OK) return rc;
            PSR_CUDA(psr_launch_common_mode_image(t, d_raw, nframes, d_out, st));
            rc = gap_fill_join(c, nframes, d_out, st);
            if (rc != PSR_OK) return rc;
        } else {
            // no hidden allocation on the hot call: the scratch of this two-pass path is sized by psr_reserve
            if (c->scratch_elems < (size_t)nframes * d.npix)
                return fail(PSR_E_INVALID, "image output with common-mode word %d goes through a calibrated scratch of %d frames: "
                            "call psr_reserve(ctx, max_frames) once before the first batch", d.cm_mode, nframes);
            // the calib-layout keep mask does not apply to images (ref: producer.py:92-97 masks the 2-D image)
            PsrDev t = d; t.ignore_keep = 1;
            PSR_CUDA(psr_launch_common_mode(t, d_raw, nframes, c->d_scratch, st));
            PSR_CUDA(psr_launch_image_gather(d, c->d_scratch, nframes, d_out, st));
        }
    }
    return PSR_OK;
}

int check_batch_args(psr_ctx* c, const void* raw, int nframes, const void* out, int out_kind) {
    if (!c) return fail(PSR_E_INVALID, "null context");
    if (!c->have_constants) return fail(PSR_E_INVALID, "psr_set_constants has not been called");
    if (nframes < 0) return fail(PSR_E_INVALID, "nframes = %d", nframes);
    if (out_kind != PSR_OUT_CALIB && out_kind != PSR_OUT_IMAGE) return fail(PSR_E_INVALID, "out_kind = %d", out_kind);
    if (out_kind == PSR_OUT_IMAGE && !c->have_geometry) return fail(PSR_E_INVALID, "PSR_OUT_IMAGE needs psr_set_geometry");
    if (nframes > 0 && (!raw || !out)) return fail(PSR_E_INVALID, "null frame buffer");
    if (((uintptr_t)raw & 15u) || ((uintptr_t)out & 15u)) return fail(PSR_E_INVALID, "frame buffers must be 16-byte aligned");
    return PSR_OK;
}

}  // namespace

extern "C" {

int psr_version(void) { return PSR_VERSION; }
const char* psr_last_error(void) { return g_err.c_str(); }

int psr_device_count(int* count) {
    if (!count) return fail(PSR_E_INVALID, "null count");
    PSR_CUDA(cudaGetDeviceCount(count));
    return PSR_OK;
}

int psr_create(psr_ctx** out, int det_kind, int nseg, int rows, int cols, int device) {
    if (!out) return fail(PSR_E_INVALID, "null out");
    *out = nullptr;
    if (det_kind == PSR_DET_EPIX10KA) {
        if (rows != 352 || cols != 384) return fail(PSR_E_UNSUPPORTED, "epix10ka segments are 352x384, got %dx%d", rows, cols);
    } else if (det_kind == PSR_DET_JUNGFRAU) {
        if (rows != 512 || cols != 1024) return fail(PSR_E_UNSUPPORTED, "jungfrau segments are 512x1024, got %dx%d", rows, cols);
    } else return fail(PSR_E_INVALID, "det_kind = %d", det_kind);
    if (nseg <= 0 || nseg > 4096) return fail(PSR_E_INVALID, "nseg = %d", nseg);
    int ndev = 0;
    PSR_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(PSR_E_INVALID, "device %d of %d", device, ndev);
    DeviceGuard g(device);
    if (!g.ok) return fail(PSR_E_CUDA, "cudaSetDevice(%d) failed", device);
    cudaDeviceProp prop;
    PSR_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(PSR_E_UNSUPPORTED, "built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
    PSR_CUDA(psr_common_mode_setup());
    psr_ctx* c = new (std::nothrow) psr_ctx();
    if (!c) return fail(PSR_E_NOMEM, "out of host memory");
    c->device = device;
    c->dev.det_kind = det_kind; c->dev.nseg = nseg; c->dev.rows = rows; c->dev.cols = cols;
    c->dev.npix = nseg * rows * cols;
    c->dev.cm_mode = 0; c->dev.cormax = 0.f; c->dev.npix_min = 10;
    c->elig_mask = det_kind == PSR_DET_EPIX10KA ? 0x1Bu : 0x3u;
    c->dev.elig_mask = c->elig_mask;
    if (const char* e = getenv("PSR_FRAMES_PER_BLOCK")) { int v = atoi(e); if (v > 0) c->frames_per_block = v; }
    // everything a hot call would otherwise create lazily: the side stream and events of the gap fill, and the device address
    // of psr_calib_image_noalloc's built-in scratch (the lookup loads this library's device code -- and with it the array --
    // onto the device, so it happens here, at start-up, not inside a call that promised not to allocate)
    cudaError_t e1 = cudaStreamCreateWithFlags(&c->s_gap, cudaStreamNonBlocking);
    if (e1 == cudaSuccess) e1 = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    if (e1 == cudaSuccess) e1 = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
    if (e1 == cudaSuccess && device < NOALLOC_MAX_DEVICES) {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            void* p = nullptr;
            e1 = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e1 == cudaSuccess) g_noalloc_ptr[device] = static_cast<float*>(p);
        }
    }
    if (e1 != cudaSuccess) { psr_destroy(c); return fail(PSR_E_CUDA, "psr_create: %s", cudaGetErrorString(e1)); }
    *out = c;
    return PSR_OK;
}

void psr_destroy(psr_ctx* c) {
    if (!c) return;
    DeviceGuard g(c->device);
    if (c->pipe.s_in) {                              // a batch begun asynchronously may still be using the staging ring
        cudaStreamSynchronize(c->pipe.s_in); cudaStreamSynchronize(c->pipe.s_comp); cudaStreamSynchronize(c->pipe.s_out);
    }
    free_pipe(c->pipe);
    if (c->pipe.s_in) cudaStreamDestroy(c->pipe.s_in);
    if (c->pipe.s_comp) cudaStreamDestroy(c->pipe.s_comp);
    if (c->pipe.s_out) cudaStreamDestroy(c->pipe.s_out);
    if (c->s_gap) cudaStreamDestroy(c->s_gap);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->d_consts) cudaFree(c->d_consts);
    if (c->d_flags) cudaFree(c->d_flags);
    if (c->d_inv) cudaFree(c->d_inv);
    if (c->d_fwd) cudaFree(c->d_fwd);
    if (c->d_gaps) cudaFree(c->d_gaps);
    if (c->d_blk_order) cudaFree(c->d_blk_order);
    if (c->d_keep_image) cudaFree(c->d_keep_image);
    if (c->d_scratch) cudaFree(c->d_scratch);
    delete c;
}

int psr_set_constants(psr_ctx* c, const float* h_peds, const float* h_gain, const float* h_offs,
                      const uint8_t* h_cfg, const uint8_t* h_mask, int nmodes)
{
    if (!c) return fail(PSR_E_INVALID, "null context");
    if (!h_peds) return fail(PSR_E_INVALID, "peds is required");
    const bool epix = c->dev.det_kind == PSR_DET_EPIX10KA;
    if (nmodes != (epix ? 7 : 3)) return fail(PSR_E_INVALID, "nmodes = %d, expected %d", nmodes, epix ? 7 : 3);
    if (epix && (!h_cfg || !h_gain)) return fail(PSR_E_INVALID, "epix10ka needs cfg and gain");
    if (epix && h_offs) return fail(PSR_E_INVALID, "epix10ka has no offset constants");
    const size_t n = (size_t)c->dev.npix;
    if (h_mask) for (size_t i = 0; i < n; ++i)
        if (h_mask[i] > 1) return fail(PSR_E_INVALID, "mask must be 0/1; mask[%zu] = %u", i, (unsigned)h_mask[i]);
    DeviceGuard g(c->device);
    if (c->d_consts) { PSR_CUDA(cudaFree(c->d_consts)); c->d_consts = nullptr; }
    if (!c->d_flags) PSR_CUDA(cudaMalloc(&c->d_flags, n));
    c->have_constants = false;
    try {
        if (h_mask) c->h_mask.assign(h_mask, h_mask + n); else c->h_mask.clear();
        if (epix) {
            c->h_mode0.resize(n); c->h_mode1.resize(n);
            std::vector<float> tab(4 * n);
            std::vector<uint8_t> modes(n);
            float* p0 = tab.data(); float* g0 = p0 + n; float* p1 = g0 + n; float* g1 = p1 + n;
            for (size_t i = 0; i < n; ++i) {
                const unsigned cb = h_cfg[i] & 0x1Cu;              // bits 4..2; bits 1,0 ignored
                const int m0 = epix_mode_of_cbits(cb), m1 = epix_mode_of_cbits(cb | 32u);
                c->h_mode0[i] = (uint8_t)m0; c->h_mode1[i] = (uint8_t)m1;
                modes[i] = (uint8_t)(m0 | (m1 << 4));
                p0[i] = m0 < 7 ? h_peds[(size_t)m0 * n + i] : 0.0f;
                g0[i] = m0 < 7 ? gfac_of(h_gain[(size_t)m0 * n + i]) : 1.0f;
                p1[i] = m1 < 7 ? h_peds[(size_t)m1 * n + i] : 0.0f;
                g1[i] = m1 < 7 ? gfac_of(h_gain[(size_t)m1 * n + i]) : 1.0f;
                if (h_mask && !h_mask[i]) {
                    if (fold_overflows(g0[i], p0[i], 0.f) || fold_overflows(g1[i], p1[i], 0.f))
                        return fail(PSR_E_UNSUPPORTED, "masked pixel %zu: 1/gain times the pixel's value range overflows float32", i);
                    g0[i] = g0[i] * 0.0f; g1[i] = g1[i] * 0.0f;
                }
            }
            c->consts_bytes = 4 * n * sizeof(float) + n;
            PSR_CUDA(cudaMalloc(&c->d_consts, c->consts_bytes));
            PSR_CUDA(cudaMemcpy(c->d_consts, tab.data(), 4 * n * sizeof(float), cudaMemcpyHostToDevice));
            float* d = static_cast<float*>(c->d_consts);
            c->dev.ped0 = d; c->dev.gf0 = d + n; c->dev.ped1 = d + 2 * n; c->dev.gf1 = d + 3 * n;
            uint8_t* dm = reinterpret_cast<uint8_t*>(d + 4 * n);
            PSR_CUDA(cudaMemcpy(dm, modes.data(), n, cudaMemcpyHostToDevice));
            c->dev.modes = dm;
            c->dev.peds = c->dev.offs = c->dev.gfac = nullptr;
        } else {
            const size_t m = 3 * n;
            std::vector<float> gf(m);
            for (size_t i = 0; i < m; ++i) gf[i] = h_gain ? gfac_of(h_gain[i]) : 1.0f;
            if (h_mask) for (size_t i = 0; i < m; ++i) if (!h_mask[i % n]) {
                if (fold_overflows(gf[i], h_peds[i], h_offs ? h_offs[i] : 0.f))
                    return fail(PSR_E_UNSUPPORTED, "masked pixel %zu: 1/gain times the pixel's value range overflows float32", i % n);
                gf[i] = gf[i] * 0.0f;
            }
            c->consts_bytes = (h_offs ? 3 : 2) * m * sizeof(float);
            PSR_CUDA(cudaMalloc(&c->d_consts, c->consts_bytes));
            float* d = static_cast<float*>(c->d_consts);
            PSR_CUDA(cudaMemcpy(d, h_peds, m * sizeof(float), cudaMemcpyHostToDevice));
            PSR_CUDA(cudaMemcpy(d + m, gf.data(), m * sizeof(float), cudaMemcpyHostToDevice));
            c->dev.peds = d; c->dev.gfac = d + m; c->dev.offs = nullptr;
            if (h_offs) { PSR_CUDA(cudaMemcpy(d + 2 * m, h_offs, m * sizeof(float), cudaMemcpyHostToDevice)); c->dev.offs = d + 2 * m; }
            c->dev.ped0 = c->dev.gf0 = c->dev.ped1 = c->dev.gf1 = nullptr; c->dev.modes = nullptr;
        }
    } catch (const std::bad_alloc&) { return fail(PSR_E_NOMEM, "out of host memory"); }
    c->dev.flags = c->d_flags;
    int rc = rebuild_flags(c);
    if (rc != PSR_OK) return rc;
    c->have_constants = true;
    return PSR_OK;
}

int psr_set_common_mode(psr_ctx* c, int mode_bits, float cormax, int npix_min, uint32_t eligible_modes_mask) {
    if (!c) return fail(PSR_E_INVALID, "null context");
    if (mode_bits < 0 || mode_bits > 7) return fail(PSR_E_INVALID, "mode_bits = %d (bit word of 4|1|2)", mode_bits);
    const uint32_t dflt = c->dev.det_kind == PSR_DET_EPIX10KA ? 0x1Bu : 0x3u;
    const uint32_t em = eligible_modes_mask ? eligible_modes_mask : dflt;
    c->dev.cm_mode = mode_bits; c->dev.cormax = cormax; c->dev.npix_min = npix_min;
    const bool changed = em != c->elig_mask;
    c->elig_mask = em; c->dev.elig_mask = em;
    if (changed && c->have_constants) { DeviceGuard g(c->device); return rebuild_flags(c); }
    return PSR_OK;
}

int psr_set_geometry(psr_ctx* c, const int32_t* h_iX, const int32_t* h_iY) {
    if (!c || !h_iX || !h_iY) return fail(PSR_E_INVALID, "null argument");
    const size_t n = (size_t)c->dev.npix;
    int32_t mx = 0, my = 0;
    for (size_t i = 0; i < n; ++i) {
        if (h_iX[i] < 0 || h_iY[i] < 0) return fail(PSR_E_INVALID, "negative image index at pixel %zu", i);
        if (h_iX[i] > mx) mx = h_iX[i];
        if (h_iY[i] > my) my = h_iY[i];
    }
    const int H = mx + 1, W = my + 1;
    if ((size_t)H * W > ((size_t)1 << 31) - 1) return fail(PSR_E_INVALID, "image %dx%d too large", H, W);
    DeviceGuard g(c->device);
    c->have_geometry = false;            // a failure below must not leave the old map's pointer live
    try {
        std::vector<int32_t> inv((size_t)H * W, -1);
        for (size_t i = 0; i < n; ++i) inv[(size_t)h_iX[i] * W + h_iY[i]] = (int32_t)i;   // flat C order: last writer wins
        if (c->d_inv) { PSR_CUDA(cudaFree(c->d_inv)); c->d_inv = nullptr; }
        if (c->d_keep_image) { PSR_CUDA(cudaFree(c->d_keep_image)); c->d_keep_image = nullptr; c->dev.keep_image = nullptr; }
        PSR_CUDA(cudaMalloc(&c->d_inv, inv.size() * sizeof(int32_t)));
        PSR_CUDA(cudaMemcpy(c->d_inv, inv.data(), inv.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
        // scatter side: every source pixel that is the last writer of its target stores there, nobody else does
        c->dev.fwd = nullptr; c->dev.gaps = nullptr; c->dev.ngaps = 0; c->dev.blk_order = nullptr;
        if (c->d_blk_order) { PSR_CUDA(cudaFree(c->d_blk_order)); c->d_blk_order = nullptr; }
        if (c->d_fwd) { PSR_CUDA(cudaFree(c->d_fwd)); c->d_fwd = nullptr; }
        if (c->d_gaps) { PSR_CUDA(cudaFree(c->d_gaps)); c->d_gaps = nullptr; }
        c->h_fwd.assign(n, INT32_MIN);
        for (size_t i = 0; i < n; ++i) {
            const size_t t = (size_t)h_iX[i] * W + h_iY[i];
            if (inv[t] == (int32_t)i) c->h_fwd[i] = (int32_t)t;
        }
        std::vector<int32_t> gaps;
        for (size_t t = 0; t < inv.size(); ++t) if (inv[t] < 0) gaps.push_back((int32_t)t);
        if (!gaps.empty()) {
            PSR_CUDA(cudaMalloc(&c->d_gaps, gaps.size() * sizeof(int32_t)));
            PSR_CUDA(cudaMemcpy(c->d_gaps, gaps.data(), gaps.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
        }
        int rc = upload_fwd(c, nullptr);
        if (rc != PSR_OK) return rc;
        c->dev.gaps = c->d_gaps; c->dev.ngaps = (int)gaps.size();
        if (c->dev.rows % 32 == 0 && c->dev.cols % 128 == 0) {
            // launch order of the scatter kernel's 1024-pixel blocks: by image position of the block's first pixel that has a
            // target.  Mirrors the kernel's block -> pixel rule (image_scatter_kernel): linear blocks on unrotated segments,
            // quarter tiles of 32 rows x 128 pixels on segments rotated by 90/270 degrees.
            const int rows = c->dev.rows, cols = c->dev.cols, pps = rows * cols, bps = pps / 1024, nblk = c->dev.nseg * bps;
            std::vector<std::pair<int64_t, int32_t>> key((size_t)nblk);
            for (int seg = 0; seg < c->dev.nseg; ++seg) {
                const size_t probe = (size_t)seg * pps + (size_t)(pps / 8) * 4;          // group in the middle of the segment
                const int32_t p0 = c->h_fwd[probe], p1 = c->h_fwd[probe + 1];
                const bool rot = p0 >= 0 && p1 >= 0 && p1 - p0 != 1 && p1 - p0 != -1;
                for (int lb = 0; lb < bps; ++lb) {
                    int64_t best = INT64_MAX;
                    for (int k = 0; k < 1024 && best == INT64_MAX; ++k) {                 // usually the very first pixel
                        size_t px;
                        if (!rot) px = (size_t)seg * pps + (size_t)lb * 1024 + k;
                        else {
                            const int tiles_x = cols / 128, t = lb >> 2, pass = lb & 3, ty = t / tiles_x, tx = t - ty * tiles_x;
                            px = (size_t)seg * pps + (size_t)(ty * 32 + k / 32) * cols + tx * 128 + pass * 32 + k % 32;
                        }
                        if (c->h_fwd[px] >= 0) best = c->h_fwd[px];
                    }
                    key[(size_t)seg * bps + lb] = {best, (int32_t)(seg * bps + lb)};
                }
            }
            std::sort(key.begin(), key.end());
            std::vector<int32_t> order((size_t)nblk);
            for (int i = 0; i < nblk; ++i) order[i] = key[i].second;
            PSR_CUDA(cudaMalloc(&c->d_blk_order, order.size() * sizeof(int32_t)));
            PSR_CUDA(cudaMemcpy(c->d_blk_order, order.data(), order.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
            c->dev.blk_order = c->d_blk_order;
        }
    } catch (const std::bad_alloc&) { return fail(PSR_E_NOMEM, "out of host memory"); }
    c->dev.inv = c->d_inv; c->dev.H = H; c->dev.W = W;
    c->have_geometry = true;
    free_pipe(c->pipe);
    return PSR_OK;
}

int psr_image_shape(const psr_ctx* c, int* H, int* W) {
    if (!c || !H || !W) return fail(PSR_E_INVALID, "null argument");
    if (!c->have_geometry) return fail(PSR_E_INVALID, "psr_set_geometry has not been called");
    *H = c->dev.H; *W = c->dev.W;
    return PSR_OK;
}

int psr_set_keep_mask(psr_ctx* c, const uint8_t* h_keep, size_t n, int out_kind) {
    if (!c) return fail(PSR_E_INVALID, "null context");
    DeviceGuard g(c->device);
    if (out_kind == PSR_OUT_CALIB) {
        if (h_keep && n != (size_t)c->dev.npix) return fail(PSR_E_INVALID, "keep mask has %zu elements, frame has %d", n, c->dev.npix);
        try { if (h_keep) c->h_keep_calib.assign(h_keep, h_keep + n); else c->h_keep_calib.clear(); }
        catch (const std::bad_alloc&) { return fail(PSR_E_NOMEM, "out of host memory"); }
        if (c->have_constants) return rebuild_flags(c);
        return PSR_OK;
    }
    if (out_kind != PSR_OUT_IMAGE) return fail(PSR_E_INVALID, "out_kind = %d", out_kind);
    if (!c->have_geometry) return fail(PSR_E_INVALID, "image keep mask needs psr_set_geometry first");
    const size_t hw = (size_t)c->dev.H * c->dev.W;
    if (!h_keep) { if (c->d_keep_image) { PSR_CUDA(cudaFree(c->d_keep_image)); c->d_keep_image = nullptr; } c->dev.keep_image = nullptr; return upload_fwd(c, nullptr); }
    if (n != hw) return fail(PSR_E_INVALID, "keep mask has %zu elements, image has %zu", n, hw);
    if (!c->d_keep_image) PSR_CUDA(cudaMalloc(&c->d_keep_image, hw));
    PSR_CUDA(cudaMemcpy(c->d_keep_image, h_keep, hw, cudaMemcpyHostToDevice));
    c->dev.keep_image = c->d_keep_image;
    try { return upload_fwd(c, h_keep); }
    catch (const std::bad_alloc&) { return fail(PSR_E_NOMEM, "out of host memory"); }
}

int psr_reserve(psr_ctx* c, int max_frames) {
    if (!c || max_frames < 0) return fail(PSR_E_INVALID, "bad argument");
    DeviceGuard g(c->device);
    return ensure_scratch(c, (size_t)max_frames * c->dev.npix);
}

int psr_set_option(psr_ctx* c, const char* name, int value) {
    if (!c || !name) return fail(PSR_E_INVALID, "null argument");
    if (!strcmp(name, "frames_per_block")) { c->frames_per_block = value; return PSR_OK; }
    if (!strcmp(name, "image_gather")) { c->image_gather = value != 0; return PSR_OK; }
    return fail(PSR_E_INVALID, "unknown option %s", name);
}

int psr_calib_batch(psr_ctx* c, const uint16_t* d_raw, int nframes, float* d_out, int out_kind, void* stream) {
    int rc = check_batch_args(c, d_raw, nframes, d_out, out_kind);
    if (rc != PSR_OK) return rc;
    if (nframes == 0) return PSR_OK;
    DeviceGuard g(c->device);
    return launch_batch(c, d_raw, nframes, d_out, out_kind, static_cast<cudaStream_t>(stream));
}

int psr_calib_image_noalloc(psr_ctx* c, const uint16_t* d_raw, int nframes, float* d_img, void* stream) {
    int rc = check_batch_args(c, d_raw, nframes, d_img, PSR_OUT_IMAGE);
    if (rc != PSR_OK) return rc;
    if (nframes == 0) return PSR_OK;
    const size_t frame_bytes = (size_t)c->dev.npix * sizeof(float);
    const size_t need = (size_t)nframes * frame_bytes;
    if (need > (size_t)PSR_NOALLOC_SCRATCH_BYTES)
        return fail(PSR_E_INVALID, "psr_calib_image_noalloc: %d frames need %zu bytes of scratch, the built-in array holds %u "
                    "(at most %d frames of this detector per call); slice the batch", nframes, need,
                    (unsigned)PSR_NOALLOC_SCRATCH_BYTES, (int)((size_t)PSR_NOALLOC_SCRATCH_BYTES / frame_bytes));
    if (c->device < 0 || c->device >= NOALLOC_MAX_DEVICES) return fail(PSR_E_INVALID, "device index %d not supported here", c->device);
    DeviceGuard g(c->device);
    if (!g.ok) return fail(PSR_E_CUDA, "cudaSetDevice(%d) failed", c->device);
    float* scratch = nullptr;
    {   // the symbol's address on THIS device, looked up once per device by psr_create with that device current
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        scratch = g_noalloc_ptr[c->device];
    }
    if (!scratch) return fail(PSR_E_CUDA, "psr_calib_image_noalloc: no scratch address for device %d", c->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    PsrDev t = c->dev; t.ignore_keep = 1;       // image mode: producer masks act on the 2-D image only
    if (t.fwd != nullptr && (t.cm_mode == 0 || t.cm_mode == 4)) {
        // no common mode, or bank medians only: the kernels that write each pixel straight to its place in the image
        // need no scratch at all (the gap fill's side stream and events exist since psr_create)
        rc = gap_fill_fork(c, st);
        if (rc != PSR_OK) return rc;
        if (t.cm_mode == 0) PSR_CUDA(psr_launch_image_scatter(c->dev, d_raw, nframes, d_img, c->frames_per_block, st));
        else PSR_CUDA(psr_launch_common_mode_image(t, d_raw, nframes, d_img, st));
        return gap_fill_join(c, nframes, d_img, st);
    }
    // pass 1 writes every pixel of frames [0, nframes) of the scratch, pass 2 reads exactly those: stale
    // contents from an earlier call are never seen, so nothing is cleared
    PSR_CUDA(psr_launch_common_mode(t, d_raw, nframes, scratch, st));
    PSR_CUDA(psr_launch_image_gather(c->dev, scratch, nframes, d_img, st));
    return PSR_OK;
}

int psr_decode_gain_modes(psr_ctx* c, const uint16_t* d_raw, int nframes, uint8_t* d_modes, void* stream) {
    if (!c || !c->have_constants) return fail(PSR_E_INVALID, "context without constants");
    if (nframes < 0 || (nframes > 0 && (!d_raw || !d_modes))) return fail(PSR_E_INVALID, "bad argument");
    DeviceGuard g(c->device);
    PSR_CUDA(psr_launch_decode(c->dev, d_raw, nframes, d_modes, static_cast<cudaStream_t>(stream)));
    return PSR_OK;
}

int psr_set_host_chunk(psr_ctx* c, int chunk_frames, int depth) {
    if (!c || chunk_frames <= 0 || depth < 2 || depth > 16) return fail(PSR_E_INVALID, "chunk_frames > 0, 2 <= depth <= 16");
    DeviceGuard g(c->device);
    free_pipe(c->pipe);
    c->pipe.chunk = chunk_frames; c->pipe.depth = depth;
    return PSR_OK;
}

int psr_calib_batch_host_begin(psr_ctx* c, const uint16_t* h_raw, int nframes, float* h_out, float* d_out, int out_kind, void* done_event) {
    if (!h_out && !d_out) return fail(PSR_E_INVALID, "need h_out and/or d_out");
    int rc = check_batch_args(c, h_raw, nframes, h_out ? (const void*)h_out : (const void*)d_out, out_kind);
    if (rc != PSR_OK) return rc;
    if (d_out && ((uintptr_t)d_out & 15u)) return fail(PSR_E_INVALID, "d_out must be 16-byte aligned");
    DeviceGuard g(c->device);
    HostPipe& p = c->pipe;
    const size_t npix = (size_t)c->dev.npix, oelems = out_frame_elems(c, out_kind);
    if (!p.s_in) {
        PSR_CUDA(cudaStreamCreateWithFlags(&p.s_in, cudaStreamNonBlocking));
        PSR_CUDA(cudaStreamCreateWithFlags(&p.s_comp, cudaStreamNonBlocking));
        PSR_CUDA(cudaStreamCreateWithFlags(&p.s_out, cudaStreamNonBlocking));
    }
    const bool need_out_slots = (d_out == nullptr);
    if (p.d_raw.empty() || p.raw_elems != (size_t)p.chunk * npix ||
        (need_out_slots && (p.d_out.empty() || p.out_elems < (size_t)p.chunk * oelems))) {
        // (re)building the ring: nothing of an earlier call may still be using the old one
        if (p.s_in) { cudaStreamSynchronize(p.s_in); cudaStreamSynchronize(p.s_comp); cudaStreamSynchronize(p.s_out); }
        free_pipe(p);
        p.raw_elems = (size_t)p.chunk * npix;
        for (int i = 0; i < p.depth; ++i) {
            uint16_t* r = nullptr; PSR_CUDA(cudaMalloc(&r, p.raw_elems * sizeof(uint16_t))); p.d_raw.push_back(r);
            cudaEvent_t e;
            PSR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); p.ev_in.push_back(e);
            PSR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); p.ev_comp.push_back(e);
            PSR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); p.ev_out.push_back(e);
        }
        if (need_out_slots) {
            p.out_elems = (size_t)p.chunk * oelems;
            for (int i = 0; i < p.depth; ++i) { float* o = nullptr; PSR_CUDA(cudaMalloc(&o, p.out_elems * sizeof(float))); p.d_out.push_back(o); }
        }
    }
    if (out_kind == PSR_OUT_IMAGE && c->dev.cm_mode != 0 && c->dev.cm_mode != 4) { rc = ensure_scratch(c, (size_t)p.chunk * npix); if (rc != PSR_OK) return rc; }

    // Chunk schedule: the copy in one direction is the bottleneck (H2D when the batch stays on the GPU, D2H when it is
    // copied back), and what the pipeline adds on top is its fill -- the first chunk's H2D + kernel before the first D2H
    // can start -- and its drain.  So the chunks ramp up from one frame (1, 1, 2, 4, ... up to the configured size).
    // Slots are numbered by a counter that runs on across calls, so a call that begins while the previous one is still
    // in flight waits for exactly the slots that one has not finished with.
    int f0 = 0, ramp = 1, it = 0;
    while (f0 < nframes) {
        int nf = ramp < p.chunk ? ramp : p.chunk;
        if (it >= 1 && ramp < p.chunk) ramp *= 2;
        if (nf > nframes - f0) nf = nframes - f0;
        const int s = (int)(p.seq % p.depth);
        if (p.seq >= p.depth) {
            PSR_CUDA(cudaStreamWaitEvent(p.s_in, p.ev_comp[s], 0));                 // raw slot consumed
            if (need_out_slots) PSR_CUDA(cudaStreamWaitEvent(p.s_comp, p.ev_out[s], 0));   // out slot drained
        }
        PSR_CUDA(cudaMemcpyAsync(p.d_raw[s], h_raw + (size_t)f0 * npix, (size_t)nf * npix * sizeof(uint16_t), cudaMemcpyHostToDevice, p.s_in));
        PSR_CUDA(cudaEventRecord(p.ev_in[s], p.s_in));
        PSR_CUDA(cudaStreamWaitEvent(p.s_comp, p.ev_in[s], 0));
        float* o = d_out ? d_out + (size_t)f0 * oelems : p.d_out[s];
        rc = launch_batch(c, p.d_raw[s], nf, o, out_kind, p.s_comp);
        if (rc != PSR_OK) {              // drain what is already queued on the caller's buffers before reporting
            cudaStreamSynchronize(p.s_in); cudaStreamSynchronize(p.s_comp); cudaStreamSynchronize(p.s_out);
            return rc;
        }
        PSR_CUDA(cudaEventRecord(p.ev_comp[s], p.s_comp));
        if (h_out) {
            PSR_CUDA(cudaStreamWaitEvent(p.s_out, p.ev_comp[s], 0));
            PSR_CUDA(cudaMemcpyAsync(h_out + (size_t)f0 * oelems, o, (size_t)nf * oelems * sizeof(float), cudaMemcpyDeviceToHost, p.s_out));
            PSR_CUDA(cudaEventRecord(p.ev_out[s], p.s_out));
        }
        f0 += nf; ++it; ++p.seq;
    }
    if (done_event) {
        // the batch is final once the compute stream (d_out) and, with h_out, the copy-out stream are: record on the
        // one that finishes last, after making it wait for the other
        cudaStream_t last = p.s_comp;
        if (h_out) { last = p.s_out; if (nframes > 0) PSR_CUDA(cudaStreamWaitEvent(p.s_out, p.ev_comp[(int)((p.seq - 1) % p.depth)], 0)); }
        PSR_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(done_event), last));
    }
    return PSR_OK;
}

int psr_calib_batch_host_wait(psr_ctx* c, int what) {
    if (!c) return fail(PSR_E_INVALID, "null context");
    if (what != PSR_WAIT_INPUT && what != PSR_WAIT_ALL) return fail(PSR_E_INVALID, "what = %d", what);
    DeviceGuard g(c->device);
    HostPipe& p = c->pipe;
    if (!p.s_in) return PSR_OK;
    PSR_CUDA(cudaStreamSynchronize(p.s_in));                     // every H2D copy issued so far has read its host source
    if (what == PSR_WAIT_ALL) { PSR_CUDA(cudaStreamSynchronize(p.s_comp)); PSR_CUDA(cudaStreamSynchronize(p.s_out)); }
    return PSR_OK;
}

int psr_calib_batch_host(psr_ctx* c, const uint16_t* h_raw, int nframes, float* h_out, float* d_out, int out_kind) {
    int rc = psr_calib_batch_host_begin(c, h_raw, nframes, h_out, d_out, out_kind, nullptr);
    if (rc != PSR_OK) return rc;
    if (nframes == 0) return PSR_OK;
    DeviceGuard g(c->device);
    PSR_CUDA(cudaStreamSynchronize(c->pipe.s_comp));
    if (h_out) PSR_CUDA(cudaStreamSynchronize(c->pipe.s_out));
    return PSR_OK;
}

// ---- memory helpers -----------------------------------------------------------------------------
int psr_host_alloc(void** h_ptr, size_t bytes) {
    if (!h_ptr) return fail(PSR_E_INVALID, "null h_ptr");
    PSR_CUDA(cudaHostAlloc(h_ptr, bytes, cudaHostAllocPortable));
    return PSR_OK;
}
int psr_host_alloc_wc(void** h_ptr, size_t bytes) {
    if (!h_ptr) return fail(PSR_E_INVALID, "null h_ptr");
    PSR_CUDA(cudaHostAlloc(h_ptr, bytes, cudaHostAllocPortable | cudaHostAllocWriteCombined));
    return PSR_OK;
}
int psr_host_free(void* h_ptr) { if (h_ptr) PSR_CUDA(cudaFreeHost(h_ptr)); return PSR_OK; }
int psr_host_register(void* h_ptr, size_t bytes) {
    if (!h_ptr) return fail(PSR_E_INVALID, "null h_ptr");
    PSR_CUDA(cudaHostRegister(h_ptr, bytes, cudaHostRegisterPortable));
    return PSR_OK;
}
int psr_host_unregister(void* h_ptr) { if (h_ptr) PSR_CUDA(cudaHostUnregister(h_ptr)); return PSR_OK; }
int psr_device_alloc(void** d_ptr, size_t bytes, int device) {
    if (!d_ptr) return fail(PSR_E_INVALID, "null d_ptr");
    DeviceGuard g(device);
    if (!g.ok) return fail(PSR_E_CUDA, "cudaSetDevice(%d) failed", device);
    PSR_CUDA(cudaMalloc(d_ptr, bytes));
    return PSR_OK;
}
int psr_device_free(void* d_ptr, int device) {
    if (!d_ptr) return PSR_OK;
    DeviceGuard g(device);
    PSR_CUDA(cudaFree(d_ptr));
    return PSR_OK;
}
int psr_memcpy_h2d(void* d_dst, const void* h_src, size_t bytes, void* stream) {
    PSR_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
    return PSR_OK;
}
int psr_memcpy_d2h(void* h_dst, const void* d_src, size_t bytes, void* stream) {
    PSR_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
    return PSR_OK;
}
int psr_stream_synchronize(void* stream) { PSR_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream))); return PSR_OK; }

// ---- CUDA IPC -----------------------------------------------------------------------------------
static_assert(sizeof(cudaIpcMemHandle_t) == PSR_IPC_HANDLE_BYTES, "handle size");
static_assert(sizeof(cudaIpcEventHandle_t) == PSR_IPC_HANDLE_BYTES, "handle size");

int psr_ipc_export_mem(const void* d_ptr, uint8_t handle[PSR_IPC_HANDLE_BYTES]) {
    if (!d_ptr || !handle) return fail(PSR_E_INVALID, "null argument");
    cudaIpcMemHandle_t h;
    PSR_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
    memcpy(handle, &h, sizeof h);
    return PSR_OK;
}
int psr_ipc_open_mem(const uint8_t handle[PSR_IPC_HANDLE_BYTES], int device, void** d_ptr) {
    if (!d_ptr || !handle) return fail(PSR_E_INVALID, "null argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(PSR_E_CUDA, "cudaSetDevice(%d) failed", device);
    cudaIpcMemHandle_t h; memcpy(&h, handle, sizeof h);
    PSR_CUDA(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return PSR_OK;
}
int psr_ipc_close_mem(void* d_ptr, int device) {
    if (!d_ptr) return PSR_OK;
    DeviceGuard g(device);
    PSR_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return PSR_OK;
}
int psr_ipc_event_create(void** event, uint8_t handle[PSR_IPC_HANDLE_BYTES], int device) {
    if (!event || !handle) return fail(PSR_E_INVALID, "null argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(PSR_E_CUDA, "cudaSetDevice(%d) failed", device);
    cudaEvent_t e;
    PSR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming | cudaEventInterprocess));
    cudaIpcEventHandle_t h;
    cudaError_t er = cudaIpcGetEventHandle(&h, e);
    if (er != cudaSuccess) { cudaEventDestroy(e); return fail(PSR_E_CUDA, "cudaIpcGetEventHandle: %s", cudaGetErrorString(er)); }
    memcpy(handle, &h, sizeof h);
    *event = e;
    return PSR_OK;
}
int psr_ipc_event_open(const uint8_t handle[PSR_IPC_HANDLE_BYTES], int device, void** event) {
    if (!event || !handle) return fail(PSR_E_INVALID, "null argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(PSR_E_CUDA, "cudaSetDevice(%d) failed", device);
    cudaIpcEventHandle_t h; memcpy(&h, handle, sizeof h);
    cudaEvent_t e;
    PSR_CUDA(cudaIpcOpenEventHandle(&e, h));
    *event = e;
    return PSR_OK;
}
int psr_event_record(void* event, void* stream) { PSR_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(event), static_cast<cudaStream_t>(stream))); return PSR_OK; }
int psr_event_wait(void* event, void* stream) { PSR_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), static_cast<cudaEvent_t>(event), 0)); return PSR_OK; }
int psr_event_synchronize(void* event) { PSR_CUDA(cudaEventSynchronize(static_cast<cudaEvent_t>(event))); return PSR_OK; }
int psr_event_destroy(void* event) { if (event) PSR_CUDA(cudaEventDestroy(static_cast<cudaEvent_t>(event))); return PSR_OK; }

}  // extern "C"
