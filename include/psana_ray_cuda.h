/*
 * psana_ray_cuda.h -- C ABI of libpsana_ray_cuda.so: the B200 (sm_100a) calibration hot path
 * of psana-ray.
 *
 * The reference (carbonscott/psana-ray) is pure Python and has no FFI of its own; the work this
 * library does is what the reference obtains, per event and on the CPU, from
 *     psana_wrapper.PsanaWrapperSmd.iter_events(mode=...)      ref: psana_ray/producer.py:88
 *       -> psana Detector.calib(evt) / Detector.image(evt)     ref: psana_ray/producer.py:156-159
 * followed by the producer-side masks                          ref: psana_ray/producer.py:92-97
 * and what it ships through the Ray object store               ref: psana_ray/producer.py:101,
 *                                                                    psana_ray/data_reader.py:35.
 * Each entry point below names the reference interface it replaces.
 *
 * Conventions
 *   - every function returns PSR_OK (0) or a negative PSR_E_* code; the message for the calling
 *     thread's last failure is psr_last_error().  Nothing throws across the ABI or calls exit().
 *   - "h_" pointers are host memory, "d_" pointers are device memory on the context's device.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Launches
 *     are asynchronous on it.  One context is used from one host thread at a time.
 *   - the caller owns every frame buffer; a context owns only its device copies of the per-run
 *     constants, the scratch psr_reserve() sized, and the staging ring of psr_calib_batch_host().
 *   - pixel order everywhere is the detector's natural C order (nseg, rows, cols); a batch is
 *     (nframes, nseg, rows, cols).
 */
#ifndef PSANA_RAY_CUDA_H
#define PSANA_RAY_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PSR_VERSION 1

/* error codes */
#define PSR_OK            0
#define PSR_E_INVALID    -1   /* bad argument / call order */
#define PSR_E_CUDA       -2   /* a CUDA runtime call failed (message has the CUDA error string) */
#define PSR_E_NOMEM      -3
#define PSR_E_UNSUPPORTED -4  /* detector shape not compiled in */

/* detector kinds: which psana calib routine is restated */
#define PSR_DET_EPIX10KA  0   /* Detector/UtilsEpix10ka.py::calib_epix10ka_any, segments 352x384, 8 banks */
#define PSR_DET_JUNGFRAU  1   /* Detector/UtilsJungfrau.py::calib_jungfrau,    segments 512x1024, 16 banks */

/* output kinds: ImageRetrievalMode.calib / .image, ref: psana_ray/producer.py:156-159 */
#define PSR_OUT_CALIB     0   /* (nframes, nseg, rows, cols) float32 */
#define PSR_OUT_IMAGE     1   /* (nframes, H, W) float32, geometry-assembled */

/* common-mode bit word, cmpars[1]: applied in the order 4, 1, 2 (UtilsCommonMode.py) */
#define PSR_CM_ROWS       1
#define PSR_CM_COLS       2
#define PSR_CM_BANKS      4

typedef struct psr_ctx psr_ctx;

int         psr_version(void);
const char* psr_last_error(void);
int         psr_device_count(int* count);

/* Replaces PsanaWrapperSmd(exp, run, detector_name) + psana.Detector(detector_name) as far as the
 * arithmetic is concerned (ref: psana_ray/producer.py:150-154): one context per rank, bound to
 * `device` (rank -> GPU binding is the caller's; ref: producer.py:138-140 gives rank/size). */
int  psr_create(psr_ctx** out, int det_kind, int nseg, int rows, int cols, int device);
void psr_destroy(psr_ctx* ctx);

/* Per-run constants, host pointers, copied to the device and re-laid-out for the kernels
 * (replaces det.pedestals/gain/offset/mask_total fetched inside det.calib).
 *   h_peds, h_gain, h_offs : (nmodes, nseg, rows, cols) float32.  nmodes = 7 (epix10ka, order FH,
 *                            FM, FL, AHL-H, AML-M, AHL-L, AML-L) or 3 (jungfrau g0, g1, g2).
 *                            h_gain is the gain itself; gfac = divide_protected(1, gain) is formed
 *                            here.  h_gain NULL -> ones, h_offs NULL -> zeros (jungfrau only).
 *   h_cfg  : (nseg, rows, cols) uint8 per-pixel config byte, epix10ka only: bits 3,2 = pixel gain
 *            config, bit 4 = ASIC trbit broadcast over the ASIC, bits 1,0 ignored.
 *   h_mask : (nseg, rows, cols) uint8, 1 = good, 0 = bad, or NULL for no mask.  Values other than
 *            0/1 are rejected. */
int psr_set_constants(psr_ctx* ctx, const float* h_peds, const float* h_gain, const float* h_offs,
                      const uint8_t* h_cfg, const uint8_t* h_mask, int nmodes);

/* cmpars = (alg, mode, cormax, npix_min) of det.calib(evt, cmpars=...); mode_bits = 0 disables.
 * eligible_modes_mask: bit i set = gain mode i takes part in the median (0 = detector default:
 * epix10ka {FH, FM, AHL-H, AML-M} = 0x1B, jungfrau {g0, g1} = 0x3). */
int psr_set_common_mode(psr_ctx* ctx, int mode_bits, float cormax, int npix_min, uint32_t eligible_modes_mask);

/* Geometry for PSR_OUT_IMAGE: per-pixel image indices as det.indexes_x/y give them
 * (PSCalib/GeometryAccess.py::img_from_pixel_arrays).  Builds the inverse gather map with
 * last-writer-wins for duplicate targets; image shape = (max(iX)+1, max(iY)+1).
 * A new geometry is a new image: an installed PSR_OUT_IMAGE keep mask is CLEARED by this call (its
 * element count belongs to the old image shape) -- install it again afterwards. */
int psr_set_geometry(psr_ctx* ctx, const int32_t* h_iX, const int32_t* h_iY);
int psr_image_shape(const psr_ctx* ctx, int* H, int* W);

/* Producer-side masks, ref: psana_ray/producer.py:92-95 -- np.where(mask, data, 0) for the bad-pixel
 * mask then the manual mask; pass their logical AND.  `n` must equal the element count of one
 * output frame of `out_kind`.  NULL clears. */
int psr_set_keep_mask(psr_ctx* ctx, const uint8_t* h_keep, size_t n, int out_kind);

/* Size the context's scratch.  PSR_OUT_IMAGE with a common-mode word other than 4 goes through a
 * calibrated-frame scratch of max_frames (bank medians alone -- word 4 -- and no common mode write
 * the image directly and need none).  The hot call never allocates: psr_calib_batch returns
 * PSR_E_INVALID if such a batch is larger than what was reserved.  psr_calib_batch_host sizes the
 * scratch for its chunks itself, together with its staging ring, on its first call. */
int psr_reserve(psr_ctx* ctx, int max_frames);

/* Tuning knobs that change speed, never results.  "frames_per_block": frames one CTA of the
 * streaming kernels walks with its constants held in registers: 0 = auto (whole batch, at most 64),
 * > 0 = that many, < 0 = persistent schedule (one equal share of the batch per resident CTA).
 * "image_gather": 1 = PSR_OUT_IMAGE without common mode runs the output-driven gather kernel instead of
 * the source-driven scatter kernel (0, default). */
int psr_set_option(psr_ctx* ctx, const char* name, int value);

/* THE HOT CALL.  Replaces, for a whole batch at once, det.calib(evt) / det.image(evt) inside
 * iter_events (ref: producer.py:88) plus the masks of producer.py:92-95.
 *   d_raw : (nframes, nseg, rows, cols) uint16 on the device, 16-byte aligned
 *   d_out : PSR_OUT_CALIB (nframes, nseg, rows, cols) / PSR_OUT_IMAGE (nframes, H, W) float32 */
int psr_calib_batch(psr_ctx* ctx, const uint16_t* d_raw, int nframes, float* d_out, int out_kind, void* stream);

/* No counterpart in the reference: an image-mode hot call for callers that cannot allocate device
 * memory.  Two kinds of caller are in that position: a monitoring or consumer process attached to a
 * GPU whose memory another job (typically a PyTorch caching allocator) has already claimed, where
 * cudaMalloc simply fails; and a latency-critical acquisition loop that must never take the
 * device-wide synchronisation that cudaMalloc / cudaFree imply.  psr_calib_batch(PSR_OUT_IMAGE) with
 * a row / column common-mode word parks the calibrated frames in a scratch buffer the context
 * allocates (psr_reserve); this variant runs the same two kernels but parks them in a fixed array
 * that is part of the library image, PSR_NOALLOC_SCRATCH_BYTES long, which every GPU that loads the
 * library carries until the process exits.  It calls no cudaMalloc, cudaFree or any other allocating
 * routine.  (With common mode off, or bank medians only, the kernels write the image directly and
 * the array is not touched; the limit below is the call's contract all the same.)
 *   - limit: nframes * nseg*rows*cols * 4 bytes must fit the array -- 3 frames of epix10k2M, 2 frames
 *     of Jungfrau 4M per call; larger requests are rejected with PSR_E_INVALID and a message naming
 *     the limit.  Loop over the batch in slices.
 *   - d_raw and d_img are the caller's, exactly as for psr_calib_batch; results are identical to it,
 *     with common mode on or off.
 *   - set-up happens in psr_create, not here: with the context's device current it looks the array's
 *     device address up once per device, which is also the moment the driver loads this library's
 *     device code -- the 32 MiB array included -- onto that GPU.  Create the context at start-up,
 *     before other jobs claim the GPU's memory and before any stream capture; the calls themselves
 *     then neither load nor allocate anything.
 *   - the array is shared by every call and every context on the device and keeps the previous
 *     call's contents; nothing in it has to start at zero, because each call overwrites the part it
 *     later reads.  Two calls whose kernels overlap on the GPU (different streams, or different
 *     host threads) race on it: keep calls on one device in one stream. */
#define PSR_NOALLOC_SCRATCH_BYTES (32u << 20)
int psr_calib_image_noalloc(psr_ctx* ctx, const uint16_t* d_raw, int nframes, float* d_img, void* stream);

/* Gain-mode index per pixel (0..nmodes-1, 255 = no mode matches): the decode of
 * gain_maps_epix10ka_any / the jungfrau gain-bit ranges, exposed so it can be checked bit-exactly. */
int psr_decode_gain_modes(psr_ctx* ctx, const uint16_t* d_raw, int nframes, uint8_t* d_modes, void* stream);

/* Host-fed form of the hot call: raw frames in (preferably pinned) host memory are copied H2D in
 * chunks on a side stream while earlier chunks are calibrated; results go to d_out (device, may be
 * NULL) and/or h_out (host, may be NULL).  Replaces one producer loop iteration block
 * (ref: producer.py:88-101) for `nframes` events.  Synchronous: returns when h_out/d_out are final. */
int psr_calib_batch_host(psr_ctx* ctx, const uint16_t* h_raw, int nframes, float* h_out, float* d_out, int out_kind);

/* The same, asynchronously: _begin queues the copies and kernels of the batch on the context's own
 * streams and returns; `done_event` (nullable; a cudaEvent_t as void*, e.g. from
 * psr_ipc_event_create) is recorded where the batch is final, so a producer can publish the batch's
 * descriptor at once -- ref: producer.py:99-103 puts only after the data exists -- and a consumer in
 * another process waits on the event in its own stream (psr_event_wait) instead of the producer
 * blocking.  Calls may overlap: a second _begin may be issued while the first is in flight (the
 * staging slots are handed on in order).  _wait(PSR_WAIT_INPUT) returns when every h_raw passed so
 * far has been read (the host buffer may be refilled); _wait(PSR_WAIT_ALL) when everything is done. */
#define PSR_WAIT_INPUT 0
#define PSR_WAIT_ALL   1
int psr_calib_batch_host_begin(psr_ctx* ctx, const uint16_t* h_raw, int nframes, float* h_out, float* d_out, int out_kind, void* done_event);
int psr_calib_batch_host_wait(psr_ctx* ctx, int what);
int psr_set_host_chunk(psr_ctx* ctx, int chunk_frames, int depth);   /* staging ring geometry; default 8 x 4; the chunks of a call
                                                                      * ramp up 1, 1, 2, 4 ... to chunk_frames so the pipeline fills fast */

/* memory helpers for callers without their own CUDA allocator */
int psr_host_alloc(void** h_ptr, size_t bytes);          /* pinned */
int psr_host_alloc_wc(void** h_ptr, size_t bytes);       /* pinned, write-combined: for staging buffers the CPU only writes (raw frames on their way
                                                           * to the GPU); CPU reads of it are slow.  Free with psr_host_free */
int psr_host_free(void* h_ptr);
int psr_host_register(void* h_ptr, size_t bytes);        /* pin an existing buffer */
int psr_host_unregister(void* h_ptr);
int psr_device_alloc(void** d_ptr, size_t bytes, int device);
int psr_device_free(void* d_ptr, int device);
int psr_memcpy_h2d(void* d_dst, const void* h_src, size_t bytes, void* stream);
int psr_memcpy_d2h(void* h_dst, const void* d_src, size_t bytes, void* stream);
int psr_stream_synchronize(void* stream);

/* Same-box hand-off of calibrated batches to consumers as GPU-resident memory: replaces pickling the
 * frame into the Ray object store on put (ref: producer.py:101) and out of it on get
 * (ref: data_reader.py:35).  Handles are the 64-byte cudaIpcMemHandle_t / cudaIpcEventHandle_t. */
#define PSR_IPC_HANDLE_BYTES 64
int psr_ipc_export_mem(const void* d_ptr, uint8_t handle[PSR_IPC_HANDLE_BYTES]);
int psr_ipc_open_mem(const uint8_t handle[PSR_IPC_HANDLE_BYTES], int device, void** d_ptr);
int psr_ipc_close_mem(void* d_ptr, int device);
int psr_ipc_event_create(void** event, uint8_t handle[PSR_IPC_HANDLE_BYTES], int device);
int psr_ipc_event_open(const uint8_t handle[PSR_IPC_HANDLE_BYTES], int device, void** event);
int psr_event_record(void* event, void* stream);
int psr_event_wait(void* event, void* stream);           /* stream waits, host does not */
int psr_event_synchronize(void* event);
int psr_event_destroy(void* event);

#ifdef __cplusplus
}
#endif
#endif /* PSANA_RAY_CUDA_H */
