// common_mode.cu -- K3: calibration with common-mode correction in ONE pass over HBM (sm_100a).
//
// psana applies the common-mode bit word in the order 4 (2-d median per bank per half), 1 (row
// medians per bank), 2 (column medians per half) on arrf = float(raw & 0x3fff) - ped [- offset],
// over common-mode-eligible pixels only, before gain and mask (UtilsCommonMode.py; BASELINE.md
// 2.2-2.4).  All three sub-modes are confined to one "bank-half" tile -- (rows/2) x (cols/nbanks)
// pixels: 176 x 48 for epix10ka, 256 x 64 for jungfrau -- so one CTA owns one tile of one frame,
// keeps it in shared memory, runs the enabled sub-modes there and streams the result out.  HBM
// traffic stays at raw-in + float-out; the constants of the tile come from L2.
//
// The median must be EXACT (NumPy: mean of the two middle order statistics, in float32).  Values are
// mapped to order-preserving uint32 keys; nothing here is approximate and nothing depends on the data:
//   * bank (2-d) median, 8 448 / 16 384 keys per CTA: keys stay in registers, 3-pass radix select
//     (11 + 11 + 10 bits) on a 2048-bin shared histogram with predicated red.shared, refined on a
//     compacted candidate list and finished by one warp ranking the last <= 32 candidates;
//   * row medians, 48 / 64 keys: one thread per row, keys in registers through a generated Batcher
//     sorting network (sortnet.cuh);
//   * column medians, 176 / 256 keys: one warp per column, 8-bit radix select on a warp-private
//     histogram over a linear integer image of the values (their float keys share the high bits and
//     would all hit the same bins), ties decided by the float keys.
// Guards (npix_min, cormax) and the eligible-pixels-only subtraction follow UtilsCommonMode.py.
#include <type_traits>
#include "psr_kernels.cuh"
#include "sortnet.cuh"

// experiment switches (profiles/README.md records what each one measured)
#ifndef PSR_CMB_JF_NT          /* jungfrau banks-only kernel: threads, pixels per chunk, min CTAs/SM */
#define PSR_CMB_JF_NT 512
#endif
#ifndef PSR_CMB_JF_CW
#define PSR_CMB_JF_CW 4   /* 4-pixel chunks: 1 348 -> 1 173 us per 64 frames against 8 (fewer values in flight per load, fewer spills) */
#endif
#ifndef PSR_CMB_JF_KS          /* keys parked in shared memory instead of registers: a second resident CTA per SM */
#define PSR_CMB_JF_KS 1
#endif
#ifndef PSR_CMB_JF_MINB
#define PSR_CMB_JF_MINB (PSR_CMB_JF_KS ? 2 : 1)
#endif
#ifndef PSR_CMF_MAGIC           /* frames-inner kernel: int -> float through 2^23 + v (FMA pipe) instead of a conversion */
#define PSR_CMF_MAGIC 1
#endif
#ifndef PSR_CMF_CHAIN           /* frames-inner kernel: chunk j + 1's loads wait for chunk j's values (bounds registers) */
#define PSR_CMF_CHAIN 1
#endif
#ifndef PSR_CMF_SELASM          /* selects as predicated second operations (FMA pipe) */
#define PSR_CMF_SELASM 1
#endif
#ifndef PSR_LQ_ATOM             /* histogram update: 1 = unconditional add of 0/1, 0 = increment under a branch */
#define PSR_LQ_ATOM 1
#endif
#ifndef PSR_CMF_OPAQUE
#define PSR_CMF_OPAQUE 1
#endif
#ifndef PSR_LQ_NARROW_M          /* linear median: halve the window when the median's bin holds more values than this (and the tails are empty) */
#define PSR_LQ_NARROW_M 24
#endif
#ifndef PSR_LQ_MIN_WINDOW        /* linear median: narrowest window, ADU */
#define PSR_LQ_MIN_WINDOW 16.0f
#endif
#ifndef PSR_CMW_COL_LQ       /* cm_calib_kernel column medians: one-pass quantised histogram per warp, radix select as the fallback */
#define PSR_CMW_COL_LQ 1
#endif
#ifndef PSR_CMW_JF_MINB
#define PSR_CMW_JF_MINB 1
#endif
#ifndef PSR_CM_WORDS_GROUPS       /* epix10ka batches, words with row / column medians: 1 = the two-frame-group kernel, 0 = cm_calib_kernel */
#define PSR_CM_WORDS_GROUPS 1
#endif
#ifndef PSR_CM_WORDS_NG           /* that kernel's frame groups per CTA: 1 = all 704 threads on one frame (12 values per thread), 2 = two groups of 352 */
#define PSR_CM_WORDS_NG 1
#endif
#ifndef PSR_CM_MATCH
#define PSR_CM_MATCH 0
#endif
#ifndef PSR_CM_UMULHI
#define PSR_CM_UMULHI 0
#endif

namespace {

constexpr uint32_t KEY_INELIGIBLE = 0xFFFFFFFFu;   // key of no finite float

// eight consecutive floats (32-byte aligned source) into registers
__device__ __forceinline__ void ld8(const float* __restrict__ p, float (&d)[8]) {
    const float4 u = __ldg(reinterpret_cast<const float4*>(p));
    const float4 v = __ldg(reinterpret_cast<const float4*>(p) + 1);
    d[0] = u.x; d[1] = u.y; d[2] = u.z; d[3] = u.w; d[4] = v.x; d[5] = v.y; d[6] = v.z; d[7] = v.w;
}

// CW (4 or 8) consecutive floats / raw pixels / flag bytes into registers
template <int CW>
__device__ __forceinline__ void ldN(const float* __restrict__ p, float (&d)[8]) {
    const float4 u = __ldg(reinterpret_cast<const float4*>(p));
    d[0] = u.x; d[1] = u.y; d[2] = u.z; d[3] = u.w;
    if (CW == 8) { const float4 v = __ldg(reinterpret_cast<const float4*>(p) + 1); d[4] = v.x; d[5] = v.y; d[6] = v.z; d[7] = v.w; }
}
template <int CW>
__device__ __forceinline__ void ld_flags(const uint8_t* __restrict__ fl, uint32_t (&fw)[2]) {
    if (CW == 8) { const uint2 f = __ldg(reinterpret_cast<const uint2*>(fl)); fw[0] = f.x; fw[1] = f.y; }
    else { fw[0] = __ldg(reinterpret_cast<const uint32_t*>(fl)); fw[1] = 0; }
}
template <int CW>
__device__ __forceinline__ void ld_raw_flags(const uint16_t* __restrict__ raw, const uint8_t* __restrict__ fl, uint32_t (&w)[4], uint32_t (&fw)[2]) {
    if (CW == 8) { const uint4 r = __ldcs(reinterpret_cast<const uint4*>(raw)); w[0] = r.x; w[1] = r.y; w[2] = r.z; w[3] = r.w; }
    else { const uint2 r = __ldcs(reinterpret_cast<const uint2*>(raw)); w[0] = r.x; w[1] = r.y; w[2] = w[3] = 0; }
    ld_flags<CW>(fl, fw);
}

template <int NW>
struct BlockRed {
    int* buf;      // [2][32]
    int phase;
    __device__ __forceinline__ int sum(int v) {
        v = __reduce_add_sync(0xFFFFFFFFu, v);
        if ((threadIdx.x & 31) == 0) buf[phase * 32 + (threadIdx.x >> 5)] = v;
        __syncthreads();
        int t = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) t += buf[phase * 32 + w];
        phase ^= 1;
        return t;
    }
    __device__ __forceinline__ uint32_t minu(uint32_t v) {
        v = __reduce_min_sync(0xFFFFFFFFu, v);
        if ((threadIdx.x & 31) == 0) buf[phase * 32 + (threadIdx.x >> 5)] = (int)v;
        __syncthreads();
        uint32_t t = 0xFFFFFFFFu;
#pragma unroll
        for (int w = 0; w < NW; ++w) t = min(t, (uint32_t)buf[phase * 32 + w]);
        phase ^= 1;
        return t;
    }
};

// Medians of C lines (columns) held by one warp, NK keys per lane per line, by a 4 x 8-bit radix select on
// warp-private 256-bin histograms in shared memory (wh: C * 256 words).  The C selections are independent
// instruction streams, interleaved to hide the shared-memory round trips of each pass.  Ineligible keys are
// KEY_INELIGIBLE; a line with n == 0 returns garbage that the caller ignores.
template <int NK, int C>
__device__ __forceinline__ void warp_kth_radix_multi(const uint32_t (&key)[C][NK], const int (&kin)[C], uint32_t* wh, uint32_t (&res)[C]) {
    const int lane = threadIdx.x & 31;
    const uint32_t waddr = (uint32_t)__cvta_generic_to_shared(wh);
    uint32_t prefix[C]; int k[C];
#pragma unroll
    for (int q = 0; q < C; ++q) { prefix[q] = 0; k[q] = kin[q]; }
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
#pragma unroll
        for (int q = 0; q < C; ++q) {
            reinterpret_cast<uint4*>(wh + q * 256)[2 * lane] = make_uint4(0, 0, 0, 0);
            reinterpret_cast<uint4*>(wh + q * 256)[2 * lane + 1] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
#pragma unroll
        for (int q = 0; q < C; ++q)
#pragma unroll
            for (int i = 0; i < NK; ++i) {
                uint32_t m = key[q][i] != KEY_INELIGIBLE ? 1u : 0u;
                if (pass > 0) m &= ((key[q][i] >> (shift + 8)) == prefix[q]) ? 1u : 0u;
#if PSR_CM_MATCH
                // a column's values share their high bits, so plain atomics would hit one address 32 times:
                // lanes with the same digit elect a leader that adds their count once
                const uint32_t digit = m ? ((key[q][i] >> shift) & 0xFFu) : 256u;
                const uint32_t peers = __match_any_sync(0xFFFFFFFFu, digit);
                const uint32_t lead = (peers & (0u - peers)) == (1u << lane) ? m : 0u;
                asm volatile("{ .reg .pred p; setp.ne.u32 p, %1, 0; @p red.shared.add.u32 [%0], %2; }"
                             :: "r"(waddr + q * 1024 + ((digit & 0xFFu) << 2)), "r"(lead), "r"(__popc(peers)) : "memory");
#else
                asm volatile("{ .reg .pred p; setp.ne.u32 p, %1, 0; @p red.shared.add.u32 [%0], 1; }"
                             :: "r"(waddr + q * 1024 + (((key[q][i] >> shift) & 0xFFu) << 2)), "r"(m) : "memory");
#endif
            }
        __syncwarp();
#pragma unroll
        for (int q = 0; q < C; ++q) {
            const uint4 b0 = reinterpret_cast<const uint4*>(wh + q * 256)[2 * lane], b1 = reinterpret_cast<const uint4*>(wh + q * 256)[2 * lane + 1];
            const uint32_t b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            uint32_t s = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) s += b[j];
            uint32_t incl = s;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
            const uint32_t excl = incl - s;
            const bool mine = (uint32_t)k[q] >= excl && (uint32_t)k[q] < incl;
            uint32_t bin = 0, r = 0;
            if (mine) {
                r = (uint32_t)k[q] - excl; bool done = false;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (!done) { if (r < b[j]) { bin = 8 * lane + j; done = true; } else r -= b[j]; }
            }
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, mine);
            const int src = bal ? __ffs(bal) - 1 : 0;
            bin = __shfl_sync(0xFFFFFFFFu, bin, src);
            k[q] = (int)__shfl_sync(0xFFFFFFFFu, r, src);
            prefix[q] = (prefix[q] << 8) | bin;
        }
        __syncwarp();
    }
#pragma unroll
    for (int q = 0; q < C; ++q) res[q] = prefix[q];
}


// k-th smallest (0-based) float key of one warp's line, selecting on a LINEAR integer image of the values.
// A detector column's values sit in a narrow range, so their float keys share the sign/exponent bits and a
// radix pass on those bits sends all 32 lanes to the same few histogram bins.  Instead: p = floor(x * 2^16)
// (monotone in x), u = p - min(p) over the line, and the radix passes start at the highest bit of
// max(u): the first digit spreads the values linearly over the 256 bins.  p is not injective; if several
// values share the selected p the float keys of those values decide (rare: a 2^-16 ADU cell), so the
// result is still the exact order statistic.
template <int NK>
__device__ __forceinline__ uint32_t warp_kth_linear(const uint32_t (&key)[NK], int k, uint32_t* wh) {
    const int lane = threadIdx.x & 31;
    const uint32_t waddr = (uint32_t)__cvta_generic_to_shared(wh);
    int p[NK]; int pmin = 0x7FFFFFFF, pmax = (int)0x80000000;
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        p[i] = __float2int_rd(__fmul_rn(psr_key2f(key[i]), 65536.0f));       // saturates, stays monotone
        if (key[i] != KEY_INELIGIBLE) { pmin = min(pmin, p[i]); pmax = max(pmax, p[i]); }
    }
    pmin = __reduce_min_sync(0xFFFFFFFFu, pmin);
    pmax = __reduce_max_sync(0xFFFFFFFFu, pmax);
    uint32_t u[NK];
#pragma unroll
    for (int i = 0; i < NK; ++i) u[i] = key[i] != KEY_INELIGIBLE ? (uint32_t)p[i] - (uint32_t)pmin : 0xFFFFFFFFu;
    const uint32_t range = (uint32_t)pmax - (uint32_t)pmin;
    int s = 32 - __clz(range);                       // bits in play (0 when all values share one p); warp-uniform
    uint32_t prefix = 0;                             // value of the bits of u above s, once s < bits
    bool first = true;
    while (s > 0) {
        const int w = s > 8 ? 8 : s;
        s -= w;
        reinterpret_cast<uint4*>(wh)[2 * lane] = make_uint4(0, 0, 0, 0);
        reinterpret_cast<uint4*>(wh)[2 * lane + 1] = make_uint4(0, 0, 0, 0);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < NK; ++i) {
            uint32_t m = key[i] != KEY_INELIGIBLE ? 1u : 0u;
            if (!first) m &= ((u[i] >> (s + w)) == prefix) ? 1u : 0u;
            asm volatile("{ .reg .pred q; setp.ne.u32 q, %1, 0; @q red.shared.add.u32 [%0], 1; }"
                         :: "r"(waddr + (((u[i] >> s) & ((1u << w) - 1u)) << 2)), "r"(m) : "memory");
        }
        __syncwarp();
        const uint4 b0 = reinterpret_cast<const uint4*>(wh)[2 * lane], b1 = reinterpret_cast<const uint4*>(wh)[2 * lane + 1];
        const uint32_t b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        uint32_t sum = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += b[j];
        uint32_t incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        const uint32_t excl = incl - sum;
        const bool mine = (uint32_t)k >= excl && (uint32_t)k < incl;
        uint32_t bin = 0, r = 0;
        if (mine) {
            r = (uint32_t)k - excl; bool done = false;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (!done) { if (r < b[j]) { bin = 8 * lane + j; done = true; } else r -= b[j]; }
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, mine);
        const int src = bal ? __ffs(bal) - 1 : 0;
        bin = __shfl_sync(0xFFFFFFFFu, bin, src);
        k = (int)__shfl_sync(0xFFFFFFFFu, r, src);
        prefix = first ? bin : ((prefix << w) | bin);
        first = false;
        __syncwarp();
    }
    // every bit of u is decided: the values with u == target tie in p; k is now the rank among them
    const uint32_t target = first ? 0u : prefix;
    int nt = 0; uint32_t tk = KEY_INELIGIBLE;
#pragma unroll
    for (int i = 0; i < NK; ++i) if (u[i] == target) { ++nt; tk = key[i]; }
    const int ties = __reduce_add_sync(0xFFFFFFFFu, nt);
    if (ties == 1) return __reduce_min_sync(0xFFFFFFFFu, tk);         // the usual case
    uint32_t pre = 0;                                                  // rare: exact counting search on the float keys of the ties
    for (int bit = 31; bit >= 0; --bit) {
        const uint32_t cand = pre | (1u << bit);
        int c = 0;
#pragma unroll
        for (int i = 0; i < NK; ++i) c += (u[i] == target && key[i] < cand) ? 1 : 0;
        c = __reduce_add_sync(0xFFFFFFFFu, c);
        if (c <= k) pre = cand;
    }
    return pre;
}

template <int NB> __device__ __forceinline__ uint32_t lq_bin8(float x, float a, float b);
// Median of one warp's line (a bank-tile column: lane = row mod 32, NK values per lane) in ONE histogram pass: the values
// go linearly into 256 bins of a quarter ADU around `centre` (lq_bin8: monotone in x, so the bin of the wanted rank holds
// the wanted value), the lanes scan the bins, the members of that bin -- a handful -- are ranked against each other as
// floats.  lo / hi = the two middle order statistics (lo == hi for an odd count).  Returns false, with nothing decided,
// when the rank falls into an end bin (values outside the window) or the bin holds more than 32 values: the caller then
// takes the radix select (warp_kth_linear).  n = eligible values of the line (> 0); wh: 256 words of this warp's own.
template <int NK>
__device__ __forceinline__ bool warp_median_lq(const float (&x)[NK], uint32_t elig, int n, float centre, uint32_t* wh, float& lo, float& hi)
{
    const int lane = threadIdx.x & 31;
    const uint32_t waddr = (uint32_t)__cvta_generic_to_shared(wh);
    const float a = 1.0f / 64.0f, b = __fsub_rn(0.5f, __fmul_rn(centre, a));
    reinterpret_cast<uint4*>(wh)[2 * lane] = make_uint4(0, 0, 0, 0);
    reinterpret_cast<uint4*>(wh)[2 * lane + 1] = make_uint4(0, 0, 0, 0);
    __syncwarp();
    uint32_t bin[NK];
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        bin[i] = lq_bin8<256>(x[i], a, b);
        asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(waddr + bin[i] * 4u), "r"((elig >> i) & 1u) : "memory");
    }
    __syncwarp();
    const uint4 b0 = reinterpret_cast<const uint4*>(wh)[2 * lane], b1 = reinterpret_cast<const uint4*>(wh)[2 * lane + 1];
    const uint32_t cnt[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) sum += cnt[j];
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
    const uint32_t excl = incl - sum, k1 = (uint32_t)(n - 1) >> 1;
    const bool mine = k1 >= excl && k1 < incl;                       // exactly one lane
    uint32_t B = 0, r = 0, M = 0;
    if (mine) {
        r = k1 - excl; bool done = false;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (!done) { if (r < cnt[j]) { B = 8 * lane + j; M = cnt[j]; done = true; } else r -= cnt[j]; }
    }
    const int src = __ffs(__ballot_sync(0xFFFFFFFFu, mine)) - 1;
    B = __shfl_sync(0xFFFFFFFFu, B, src); r = __shfl_sync(0xFFFFFFFFu, r, src); M = __shfl_sync(0xFFFFFFFFu, M, src);
    if (B == 0u || B == 255u || M > 32u) return false;               // warp-uniform
    __syncwarp();                                                     // every lane has read its bins: wh becomes the candidate list
    float* list = reinterpret_cast<float*>(wh);
    uint32_t base = 0;
    uint32_t above = KEY_INELIGIBLE;                                  // smallest key beyond the bin (for hi when the bin ends at rank k1)
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        const bool e = (elig >> i) & 1u;
        const bool hit = e && bin[i] == B;
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, hit);
        if (hit) list[base + __popc(bal & ((1u << lane) - 1u))] = x[i];
        base += __popc(bal);
        if (e && bin[i] > B) above = min(above, psr_f2key(x[i]));
    }
    __syncwarp();
    // rank the M candidates: lane j takes candidate j; ties are ordered by position
    const float cj = lane < (int)M ? list[lane] : 0.0f;
    uint32_t rank = 0;
    for (uint32_t i = 0; i < M; ++i) { const float ci = list[i]; rank += (ci < cj || (ci == cj && (int)i < lane)) ? 1u : 0u; }
    const uint32_t blo = __ballot_sync(0xFFFFFFFFu, lane < (int)M && rank == r);
    lo = __shfl_sync(0xFFFFFFFFu, cj, __ffs(blo) - 1);
    hi = lo;
    if ((n & 1) == 0) {
        if (r + 1u < M) {
            const uint32_t bhi = __ballot_sync(0xFFFFFFFFu, lane < (int)M && rank == r + 1u);
            hi = __shfl_sync(0xFFFFFFFFu, cj, __ffs(bhi) - 1);
        } else hi = psr_key2f(__reduce_min_sync(0xFFFFFFFFu, above));
    }
    __syncwarp();
    return true;
}

// Values of one thread parked in shared memory instead of registers: slot i of thread t at base[i * NT + t]
// (conflict-free), same [] syntax as a register array so the selection code serves both.
template <int NT>
struct SmemVals {
    const float* base;                   // already offset by threadIdx.x
    __device__ __forceinline__ float operator[](int i) const { return base[i * NT]; }
};

// ---- CTA-wide exact selection: 3-pass radix select (11 + 11 + 10 bits) on a 2048-bin shared histogram.
// Keys stay in registers; bit i of `elig` says key[i] takes part.  Returns the k-th smallest (0-based)
// eligible key; k must be < the number of eligible keys.  Every thread of the CTA must call it.
// hist: 2048 words of shared memory; sh: 40 ints of shared memory (sh[34] = population of the chosen bin).
template <int NT, int NKEY, int PASS, typename K>
__device__ __forceinline__ void block_kth_pass(const K& key, uint32_t elig, uint32_t& prefix, int& k,
                                               uint32_t* hist, int* sh)
{
    constexpr int NB = 2048, BPT = (NB + NT - 1) / NT, NW = NT / 32;
    constexpr int SHIFT = PASS == 0 ? 21 : PASS == 1 ? 10 : 0;          // digit position
    constexpr int HSHIFT = PASS == 1 ? 21 : 10;                         // bits above the digit (PASS > 0)
    constexpr uint32_t DMASK = PASS == 2 ? 0x3FFu : 0x7FFu;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t haddr = (uint32_t)__cvta_generic_to_shared(hist);
#pragma unroll
    for (int i = 0; i < NKEY; ++i) {
        uint32_t m = (elig >> i) & 1u;
        // shifts as multiply-high / multiply-add: those issue on the FMA pipe, the ALU pipe (LOP3/SHF/ISETP) is the busy one
#if PSR_CM_UMULHI
        if (PASS > 0) m &= (__umulhi(key[i], 1u << (32 - HSHIFT)) == prefix) ? 1u : 0u;
        const uint32_t digit = PASS == 0 ? __umulhi(key[i], 1u << 11) : ((key[i] >> SHIFT) & DMASK);
#else
        if (PASS > 0) m &= ((key[i] >> HSHIFT) == prefix) ? 1u : 0u;
        const uint32_t digit = (key[i] >> SHIFT) & DMASK;
#endif
        // predicated shared-memory reduction: no branch around the atomic
        asm volatile("{ .reg .pred q; setp.ne.u32 q, %1, 0; @q red.shared.add.u32 [%0], 1; }"
                     :: "r"(digit * 4u + haddr), "r"(m) : "memory");
    }
    __syncthreads();
    uint32_t b[BPT]; uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < BPT; ++j) {
        const int idx = tid * BPT + j;
        b[j] = idx < NB ? hist[idx] : 0u;
        if (idx < NB) hist[idx] = 0u;                // owner re-zeroes its bins for the next pass
        s += b[j];
    }
    uint32_t incl = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) sh[warp] = (int)incl;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) base += (w < warp) ? (uint32_t)sh[w] : 0u;
    const uint32_t excl = base + incl - s;
    if ((uint32_t)k >= excl && (uint32_t)k < excl + s) {
        uint32_t r = (uint32_t)k - excl; int bin = tid * BPT; bool done = false;
#pragma unroll
        for (int j = 0; j < BPT; ++j) {
            if (!done) { if (r < b[j]) { bin = tid * BPT + j; done = true; sh[34] = (int)b[j]; } else r -= b[j]; }
        }
        sh[32] = bin; sh[33] = (int)r;
    }
    __syncthreads();
    prefix = (prefix << (PASS == 2 ? 10 : 11)) | (uint32_t)sh[32];
    k = sh[33];
}

template <int NT, int NKEY, typename K>
__device__ __forceinline__ uint32_t block_kth(const K& key, uint32_t elig, int k, uint32_t* hist, int* sh)
{
    for (int i = threadIdx.x; i < 2048; i += NT) hist[i] = 0u;
    __syncthreads();
    uint32_t prefix = 0;
    block_kth_pass<NT, NKEY, 0>(key, elig, prefix, k, hist, sh);
    block_kth_pass<NT, NKEY, 1>(key, elig, prefix, k, hist, sh);
    block_kth_pass<NT, NKEY, 2>(key, elig, prefix, k, hist, sh);
    return prefix;
}


// Selects written as a predicated second operation: the arithmetic runs on the FMA pipe, which has room, instead
// of a select on the ALU pipe, which is the one these kernels fill.  Same roundings as select-then-operate.
__device__ __forceinline__ float psr_sub_sel(float f, float pa, float pb, uint32_t pick_b) {        // f - (pick_b ? pb : pa)
#if !PSR_CMF_SELASM
    return __fsub_rn(f, pick_b ? pb : pa);
#endif
    float r;
    asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %4, 0;\n\tsub.rn.f32 %0, %1, %2;\n\t@p sub.rn.f32 %0, %1, %3;\n\t}" : "=&f"(r) : "f"(f), "f"(pa), "f"(pb), "r"(pick_b));
    return r;
}
__device__ __forceinline__ float psr_mul_sel(float f, float ga, float gb, uint32_t pick_b) {        // f * (pick_b ? gb : ga)
#if !PSR_CMF_SELASM
    return __fmul_rn(f, pick_b ? gb : ga);
#endif
    float r;
    asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %4, 0;\n\tmul.rn.f32 %0, %1, %2;\n\t@p mul.rn.f32 %0, %1, %3;\n\t}" : "=&f"(r) : "f"(f), "f"(ga), "f"(gb), "r"(pick_b));
    return r;
}
__device__ __forceinline__ float psr_sub_if(float f, float d, uint32_t on) {                        // on ? f - d : f
#if !PSR_CMF_SELASM
    return on ? __fsub_rn(f, d) : f;
#endif
    asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p sub.rn.f32 %0, %0, %1;\n\t}" : "+f"(f) : "f"(d), "r"(on));
    return f;
}

// ---- median of the CTA's eligible values on a LINEAR-quantised histogram ---------------------------
// The values of a bank tile sit within a few ADU of its common-mode offset, so their float keys share the
// sign/exponent bits and a radix pass on those spends its 2048 bins on a handful of them.  Instead:
//     y = sat(x * a + b)  in [0, 1]   (one saturating FMA: a = 1/64 per ADU, b = 0.5 - centre * a)
//     bin(x) = floor(y * 2047)        (one round-down FMA onto 2^23: the low mantissa bits are the integer)
// -- about 1/32 ADU per bin in a 64-ADU window around the centre, everything outside in bins 0 / 2047.
// bin() is monotone in x (a composition of monotone roundings), so the bin in which the cumulative count
// crosses rank k holds the k-th smallest value, and with sigma ~ 3 ADU that bin holds a few dozen of the
// tile's values: they are collected (the same two FMAs decide membership, so the set is exactly the bin's)
// and ranked directly by as many threads.  One histogram pass, one collection pass, no float <-> key
// transform, and the arithmetic is on the FMA pipe rather than the ALU pipe the rest of the kernel loads.
// The centre follows the previous frame's median.  If the rank lands in an end bin the window is relocated:
// one pass at 32 ADU per bin over +-32768 ADU, then the fine pass again around the bin that pass found.
// Whatever is still unusual (values beyond that, more than CAP values in one bin) takes the bitwise
// search -- same exact answer.
constexpr int   LQ_NB = 2048;
constexpr float LQ_A = 1.0f / 64.0f;          // fine window: 64 ADU
constexpr float LQ_A_COARSE = 1.0f / 65536.0f;

// bits of 2^23 + bin(x): the histogram address is bits * 4 + (base - bits(2^23) * 4), all in 32-bit wrap-around
__device__ __forceinline__ uint32_t lq_bits(float x, float a, float b) {
    return __float_as_uint(__fmaf_rd(__saturatef(__fmaf_rn(x, a, b)), (float)(LQ_NB - 1), 8388608.0f));
}

struct BitsLinear { static __device__ __forceinline__ constexpr int pos(int i) { return i; } };
// value i = pixel (2k + h) of chunk j (i = 8j + 2k + h: word k of the chunk's four raw words, low / high half):
// bit 4j + k + 16h, the place where ONE shift of the raw word drops its two bit-14 flags
struct BitsEpixWords { static __device__ __forceinline__ constexpr int pos(int i) { return 4 * (i >> 3) + ((i & 7) >> 1) + 16 * (i & 1); } };
// the same for 4-pixel chunks (two raw words): value i = pixel (2k + h) of chunk j (i = 4j + 2k + h) -> bit 2j + k + 16h
struct BitsEpixWords4 { static __device__ __forceinline__ constexpr int pos(int i) { return 2 * (i >> 2) + ((i & 3) >> 1) + 16 * (i & 1); } };

// Who synchronises with whom: the whole CTA, or one of several equal groups of warps inside it, each on a hardware
// barrier of its own (the groups then run their phases independently, like CTAs that share the tile's constants).
struct CtaGroup {
    __device__ __forceinline__ int tid() const { return threadIdx.x; }
    __device__ __forceinline__ void sync() const { __syncthreads(); }
};
template <int N>
struct WarpGroup {
    int t, id;                                       // thread within the group; barrier 1 + group index
    __device__ __forceinline__ int tid() const { return t; }
    __device__ __forceinline__ void sync() const { asm volatile("bar.sync %0, %1;" :: "r"(id), "n"(N) : "memory"); }
};
// BlockRed over a group
template <int NW, typename G>
struct GroupRed {
    int* buf; int phase; G g;
    __device__ __forceinline__ int sum(int v) {
        v = __reduce_add_sync(0xFFFFFFFFu, v);
        if ((g.tid() & 31) == 0) buf[phase * 32 + (g.tid() >> 5)] = v;
        g.sync();
        int t = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) t += buf[phase * 32 + w];
        phase ^= 1;
        return t;
    }
    __device__ __forceinline__ uint32_t minu(uint32_t v) {
        v = __reduce_min_sync(0xFFFFFFFFu, v);
        if ((g.tid() & 31) == 0) buf[phase * 32 + (g.tid() >> 5)] = (int)v;
        g.sync();
        uint32_t t = 0xFFFFFFFFu;
#pragma unroll
        for (int w = 0; w < NW; ++w) t = min(t, (uint32_t)buf[phase * 32 + w]);
        phase ^= 1;
        return t;
    }
};

struct LocalVals { const float* p; __device__ __forceinline__ float operator[](int i) const { return p[i]; } };

// One histogram pass + scan: on return sh[32] = bin holding rank k1 = (n - 1) / 2, sh[33] = rank inside it, sh[34] = its
// population, sh[36] = n (eligible values), sh[35] = 0; hist is zero again.  Every thread of the group must call it.
template <int NT, int NK, typename BP, typename X, typename G>
__device__ __forceinline__ void lq_pass(const X& x, uint32_t elig, float a, float b, uint32_t* hist, int* sh, const G& g)
{
    constexpr int NSCAN = LQ_NB / 8;                 // threads that scan 8 bins each
    static_assert(NT >= NSCAN && NSCAN % 32 == 0, "scan layout");
    const int tid = g.tid(), lane = tid & 31, warp = tid >> 5;
    const uint32_t cadd = (uint32_t)__cvta_generic_to_shared(hist) - 0x2C000000u;        // bits(2^23) * 4 mod 2^32
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        const uint32_t ad = lq_bits(x[i], a, b) * 4u + cadd;
#if PSR_LQ_ATOM
        const uint32_t m = (elig << (31 - BP::pos(i))) >> 31;                             // ineligible: add 0 (no branch around the atomic)
        asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(ad), "r"(m) : "memory");
#else
        asm volatile("{ .reg .pred q; setp.ne.u32 q, %1, 0; @q red.shared.add.u32 [%0], 1; }" :: "r"(ad), "r"(elig & (1u << BP::pos(i))) : "memory");
#endif
    }
    g.sync();
    uint32_t bn[8]; uint32_t s = 0, incl = 0;
    if (tid < NSCAN) {
        uint4* h4 = reinterpret_cast<uint4*>(hist) + 2 * tid;
        const uint4 u = h4[0], v = h4[1];
        h4[0] = make_uint4(0, 0, 0, 0); h4[1] = make_uint4(0, 0, 0, 0);                   // the owner leaves its bins zero
        bn[0] = u.x; bn[1] = u.y; bn[2] = u.z; bn[3] = u.w; bn[4] = v.x; bn[5] = v.y; bn[6] = v.z; bn[7] = v.w;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += bn[j];
        incl = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        if (lane == 31) sh[warp] = (int)incl;
    }
    g.sync();
    int n = 0; uint32_t base = 0;
    {
        const int4 t0 = reinterpret_cast<const int4*>(sh)[0], t1 = reinterpret_cast<const int4*>(sh)[1];
        const int tw[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
        for (int w = 0; w < NSCAN / 32; ++w) { n += tw[w]; base += (w < warp) ? (uint32_t)tw[w] : 0u; }
    }
    const int k1 = (n - 1) >> 1;
    if (tid < NSCAN && n > 0) {
        const uint32_t excl = base + incl - s;
        if ((uint32_t)k1 >= excl && (uint32_t)k1 < excl + s) {
            uint32_t rr = (uint32_t)k1 - excl; bool done = false;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (!done) { if (rr < bn[j]) { sh[32] = 8 * tid + j; sh[33] = (int)rr; sh[34] = (int)bn[j]; done = true; } else rr -= bn[j]; }
        }
    }
    if (tid == 0) sh[39] = (int)bn[0];                 // the two end bins: everything outside the window
    if (tid == NSCAN - 1) sh[40] = (int)bn[7];
    if (tid == NT - 1) { sh[35] = 0; sh[36] = n; }
    g.sync();
}

// The members of bin B (fine window b) are collected and ranked by one thread each: lo = rank r, hi = rank r + 1 when
// the bin holds it (have_hi).  sh[35] must be 0 on entry (lq_pass leaves it so); M <= CAP.  Every thread must call it.
template <int NT, int NK, int CAP, typename BP, typename X, typename G>
__device__ __forceinline__ void lq_collect_rank(const X& x, uint32_t elig, float a, float b, int B, int r, int M, int* sh, float* list,
                                                float& lo, float& hi, bool& have_hi, const G& g)
{
    static_assert(CAP <= NT && CAP % 4 == 0, "rank layout");
    const int tid = g.tid();
    const uint32_t target = 0x4B000000u + (uint32_t)B;
    uint32_t hits = 0; float cand = 0.f;
    asm volatile("" : "+f"(b));                          // the bins are recomputed, not kept from the histogram pass (NK registers)
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        const float xi = x[i];
        if (lq_bits(xi, a, b) == target) { hits |= 1u << BP::pos(i); cand = xi; }
    }
    // a thread almost never holds two members of the bin: one candidate register, one branch per thread
    if (hits != 0u) {
        if ((hits & (hits - 1u)) == 0u) { if (hits & elig) list[atomicAdd(&sh[35], 1)] = cand; }
        else {
            hits &= elig;
#pragma unroll
            for (int i = 0; i < NK; ++i) if (hits & (1u << BP::pos(i))) list[atomicAdd(&sh[35], 1)] = x[i];
        }
    }
    if (tid < 4) list[M + tid] = __int_as_float(0x7f800000);                              // pad to a multiple of four
    g.sync();
    if (tid < M) {
        const float v = list[tid];
        int lt = 0, le = 0;
        for (int j = 0; j < M; j += 4) {
            const float4 u = *reinterpret_cast<const float4*>(list + j);
            lt += (u.x < v) + (u.y < v) + (u.z < v) + (u.w < v);
            le += (u.x <= v) + (u.y <= v) + (u.z <= v) + (u.w <= v);
        }
        if (lt <= r && r < le) sh[37] = __float_as_int(v);
        if (lt <= r + 1 && r + 1 < le) sh[38] = __float_as_int(v);
    }
    g.sync();
    lo = __int_as_float(sh[37]);
    have_hi = r + 1 < M;
    hi = have_hi ? __int_as_float(sh[38]) : lo;
}

// Everything unusual about a tile's median, in one place and not inlined, so that it costs the usual path neither
// registers nor code: works on a copy of the values in local memory.
//   have_lo: the fine pass found lo but the bin does not hold the next value (lo is the last of its bin).
//   else: the rank fell into an end bin of the fine window, or its bin holds more than CAP values.  End bin: one pass at
//   32 ADU per bin over +-32768 ADU, the fine window moved to the middle of the bin that pass finds, fine pass + ranking
//   again.  What even that does not settle goes to a 32-step binary search on the float keys, one block-wide count per bit.
// out[0] = lo, out[1] = hi, out[2] = the fine window's b as used (unchanged unless relocated).
template <int NT, int NK, int CAP, typename BP, typename G>
__device__ __noinline__ void lq_median_slow(const float* xs, uint32_t elig, int n, bool have_lo, float lo, float a, float b,
                                            uint32_t* hist, int* sh, float* list, int* redbuf, float* out, G g)
{
    const LocalVals x{xs};
    const int k1 = (n - 1) >> 1, k2 = n >> 1;
    float hi = lo; bool have_hi = false;
    g.sync();
    if (!have_lo) {
        int B = sh[32];
        g.sync();
        if (B <= 0 || B >= LQ_NB - 1) {
            lq_pass<NT, NK, BP>(x, elig, LQ_A_COARSE, 0.5f, hist, sh, g);
            B = sh[32];
            g.sync();
            if (B > 0 && B < LQ_NB - 1) {
                const float xc = __fmul_rn(__fsub_rn(__fmul_rn((float)B + 0.5f, 1.0f / (float)(LQ_NB - 1)), 0.5f), 1.0f / LQ_A_COARSE);
                b = __fsub_rn(0.5f, __fmul_rn(xc, a));
                lq_pass<NT, NK, BP>(x, elig, a, b, hist, sh, g);
                B = sh[32];
                const int r = sh[33], M = sh[34];
                if (B > 0 && B < LQ_NB - 1 && M <= CAP) {
                    lq_collect_rank<NT, NK, CAP, BP>(x, elig, a, b, B, r, M, sh, list, lo, hi, have_hi, g);
                    have_lo = true;
                }
            }
        }
        if (!have_lo) {
            GroupRed<NT / 32, G> red{redbuf, 0, g};
            uint32_t pre = 0;
#pragma unroll 1
            for (int bit = 31; bit >= 0; --bit) {
                const uint32_t cand = pre | (1u << bit);
                int cnt = 0;
#pragma unroll 1
                for (int i = 0; i < NK; ++i) cnt += ((elig & (1u << BP::pos(i))) && psr_f2key(xs[i]) < cand) ? 1 : 0;
                cnt = red.sum(cnt);
                if (cnt <= k1) pre = cand;
            }
            g.sync();
            lo = psr_key2f(pre); hi = lo;
        }
    }
    if (k2 != k1 && !have_hi) {
        // the upper middle value: the smallest eligible value above lo, or lo itself when its duplicates reach rank k2
        GroupRed<NT / 32, G> red{redbuf, 0, g};
        int cle = 0; uint32_t mg = 0xFFFFFFFFu;
#pragma unroll 1
        for (int i = 0; i < NK; ++i)
            if (elig & (1u << BP::pos(i))) { cle += (xs[i] <= lo); if (xs[i] > lo) mg = min(mg, psr_f2key(xs[i])); }
        cle = red.sum(cle);
        mg = red.minu(mg);
        g.sync();
        hi = cle <= k2 ? psr_key2f(mg) : lo;
    }
    if (k2 == k1) hi = lo;
    out[0] = lo; out[1] = hi; out[2] = b;
}

// hist: LQ_NB words, all zero on entry and on exit; sh: 48 ints (16-byte aligned); list: CAP + 4 floats.
// win = the window carried from frame to frame: win.a = 1 / (window width in ADU), win.b = 0.5 - centre * a.  It is
// re-centred on the median found and re-scaled to the distribution seen: twice as wide when more than an eighth of the
// values fell outside it (they all hit the two end bins: atomics on one word, and a rank out there costs a relocation),
// half as wide when the median's bin was crowded and the tails are empty.  Returns false (uniformly) when fewer than
// max(npix_min, 1) values are eligible; else lo/hi = the two middle order statistics.  Every thread of the group must call it.
struct LqWindow { float a, b; };
constexpr float LQ_A_MIN = LQ_A_COARSE * 2.0f, LQ_A_MAX = 1.0f / PSR_LQ_MIN_WINDOW;

template <int NT, int NK, int CAP, typename BP, typename X, typename G>
__device__ __forceinline__ bool block_median_linear(const X& x, uint32_t elig, int npix_min, LqWindow& win,
                                                    uint32_t* hist, int* sh, float* list, int* redbuf,
                                                    float& lo_out, float& hi_out, const G& g)
{
    const float a = win.a;
    float b = win.b;
    lq_pass<NT, NK, BP>(x, elig, a, b, hist, sh, g);
    const int n = sh[36];
    if (n < npix_min || n < 1) return false;                                              // uniform
    const int B = sh[32], r = sh[33], M = sh[34], tails = sh[39] + sh[40];
    float lo = 0.f, hi = 0.f; bool have_lo = false, have_hi = false;
    if (B > 0 && B < LQ_NB - 1 && M <= CAP) {                                             // uniform, the usual case
        lq_collect_rank<NT, NK, CAP, BP>(x, elig, a, b, B, r, M, sh, list, lo, hi, have_hi, g);
        have_lo = true;
    }
    if (!have_lo || (!have_hi && (n & 1) == 0)) {                                         // uniform, rare
        float xs[NK], res[3];
#pragma unroll
        for (int i = 0; i < NK; ++i) xs[i] = x[i];
        lq_median_slow<NT, NK, CAP, BP>(xs, elig, n, have_lo, lo, a, b, hist, sh, list, redbuf, res, g);
        lo = res[0]; hi = res[1]; b = res[2];
    }
    if (n & 1) hi = lo;
    lo_out = lo; hi_out = hi;
    float na = a;
    if (tails * 2 > n) na = a * 0.25f;
    else if (tails * 8 > n) na = a * 0.5f;
    else if (M > PSR_LQ_NARROW_M && tails * 32 < n) na = a * 2.0f;
    na = fminf(fmaxf(na, LQ_A_MIN), LQ_A_MAX);
    const float nb = __fsub_rn(0.5f, __fmul_rn(psr_mid(lo, hi), na));                     // a wild value just costs that frame a relocation
    win.a = na; win.b = fabsf(nb) < 1.0e6f ? nb : 0.5f;
    return true;
}

// smem flag byte of a tile pixel: bit0 mask, bit1 eligible (this frame), bit3 keep, bits 4-5 gain selector
//   epix10ka selector: data bit 14;  jungfrau selector: 0,1,2 = g0,g1,g2, 3 = none
// a if p != 0 else b, as one SELP (ptxas turns nested ternaries on the gain bits into branches)
__device__ __forceinline__ float cm_selp(float a, float b, uint32_t p) {
    float r;
    asm("{ .reg .pred q; setp.ne.u32 q, %3, 0; selp.f32 %0, %1, %2, q; }" : "=f"(r) : "f"(a), "f"(b), "r"(p));
    return r;
}

template <int DET, int ROWS, int COLS, int NB, int NT>
__global__ void __launch_bounds__(NT, DET == 0 ? 2 : PSR_CMW_JF_MINB)      // epix10ka: two CTAs/SM (measured: 1.3x on the row medians)
cm_calib_kernel(const uint16_t* __restrict__ raw, float* __restrict__ out, PsrDev c)
{
    constexpr int TR = ROWS / 2, TC = COLS / NB, CH = TC / 8, NCH = TR * CH, CPT = NCH / NT;
    constexpr int ST = TC + 1;                       // padded row stride: column walks hit 32 banks
    constexpr int NW = NT / 32;
    static_assert(NCH % NT == 0 && NT % 32 == 0, "tile must split evenly over the CTA");
    extern __shared__ float smem[];
    float* a = smem;                                           // TR * ST
    uint8_t* fl = reinterpret_cast<uint8_t*>(a + TR * ST);     // TR * ST
    // four 256-bin histograms per warp for the column medians, 16-byte aligned behind the tile
    uint32_t* whist = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(smem) + ((TR * ST * 5 + 15) / 16) * 16);
    __shared__ int redbuf[64];
    __shared__ uint32_t hist[2048];
    __shared__ int selsh[40];

    const int tid = threadIdx.x, lane = tid & 31;
    // Block order: bank fastest, then FRAME, then (half, segment).  The CTAs resident at one time then work on all frames of
    // a few half-segments, so the tiles' constants are fetched from DRAM once per batch and come from L2 for the other frames
    // (Jungfrau: 155 MB of constants do not fit L2; with the frame slowest they were re-read for every frame -- ncu: 10.5 GB of
    // DRAM reads per 64 frames), while the banks of one frame's row band still run side by side (whole rows of raw and output).
    int bid = blockIdx.x;
    const int nframes = (int)(gridDim.x / (unsigned)(c.nseg * 2 * NB));
    const int bank = bid % NB; bid /= NB;
    const int frame = bid % nframes; bid /= nframes;
    const int half = bid & 1;
    const int seg = bid >> 1;
    const uint16_t* rawf = raw + (size_t)frame * c.npix;
    float* outf = out + (size_t)frame * c.npix;
    const int mode = c.cm_mode;
    const bool has_offs = DET != 0 && c.offs != nullptr;

    // ---- load: raw + constants -> arrf in smem, keys in registers -------------------------------
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const int ch = tid + j * NT;
        const int r = ch / CH, cc = ch - r * CH;
        const int gp = ((seg * ROWS + half * TR + r) * COLS) + bank * TC + cc * 8;
        const uint4 rv = __ldcs(reinterpret_cast<const uint4*>(rawf + gp));
        const uint2 fv = __ldg(reinterpret_cast<const uint2*>(c.flags + gp));
        const uint32_t w[4] = {rv.x, rv.y, rv.z, rv.w};
        float pa[8], pb[8], pc[8], oa[8], ob[8], oc[8];
        if (DET == 0) {
            ld8(c.ped0 + gp, pa);
            ld8(c.ped1 + gp, pb);
        } else {
            ld8(c.peds + gp, pa);
            ld8(c.peds + c.npix + gp, pb);
            ld8(c.peds + 2 * (size_t)c.npix + gp, pc);
            if (has_offs) {
                ld8(c.offs + gp, oa);
                ld8(c.offs + c.npix + gp, ob);
                ld8(c.offs + 2 * (size_t)c.npix + gp, oc);
            }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t v = (i & 1) ? (w[i >> 1] >> 16) : (w[i >> 1] & 0xFFFFu);
            const uint32_t f = (((i < 4 ? fv.x : fv.y) >> (8 * (i & 3))) & 0xFFu) | (c.ignore_keep ? PSR_FL_KEEP : 0u);
            float x = (float)(v & 0x3FFFu);
            uint32_t sel, elig;
            if (DET == 0) {
                sel = (v >> 14) & 1u;
                x = __fsub_rn(x, sel ? pb[i] : pa[i]);
                elig = sel ? (f & PSR_FL_ELIG1) : (f & PSR_FL_ELIG0);
            } else {
                const uint32_t gb = v >> 14, b0 = gb & 1u, b1 = gb >> 1;     // 00 g0, 01 g1, 11 g2, 10 none
                sel = gb ^ b1;                                             // 0, 1, 2 = g0, g1, g2; 3 = none
                x = __fsub_rn(x, cm_selp(cm_selp(pc[i], 0.0f, b0), cm_selp(pb[i], pa[i], b0), b1));
                if (has_offs) x = __fsub_rn(x, cm_selp(cm_selp(oc[i], 0.0f, b0), cm_selp(ob[i], oa[i], b0), b1));
                elig = ((c.elig_mask & 7u) >> sel) & ((f & PSR_FL_MASK) ? 1u : 0u);
            }
            const int si = r * ST + cc * 8 + i;
            a[si] = x;
            fl[si] = (uint8_t)((f & (PSR_FL_MASK | PSR_FL_KEEP)) | (elig ? 2u : 0u) | (sel << 4));
        }
    }

    // ---- mode 4: 2-d median over the whole tile (common_mode_2d: skip if n < npix_min, skip if |cm| > cormax)
    if (mode & 4) {
        uint32_t key[CPT * 8];                             // the thread's own pixels back from the tile (its own writes: no barrier needed)
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            const int ch = tid + j * NT;
            const int r = ch / CH, cc = ch - r * CH;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int si = r * ST + cc * 8 + i;
                key[j * 8 + i] = (fl[si] & 2u) ? psr_f2key(a[si]) : KEY_INELIGIBLE;
            }
        }
        BlockRed<NW> red{redbuf, 0};
        int cnt = 0;
#pragma unroll
        for (int i = 0; i < CPT * 8; ++i) cnt += (key[i] != KEY_INELIGIBLE);
        const int n = red.sum(cnt);
        if (n >= c.npix_min && n > 0) {                    // uniform across the CTA
            const int k1 = (n - 1) >> 1, k2 = n >> 1;
            uint32_t elig = 0;
#pragma unroll
            for (int i = 0; i < CPT * 8; ++i) elig |= (key[i] != KEY_INELIGIBLE ? 1u : 0u) << i;
            const uint32_t prefix = block_kth<NT, CPT * 8>(key, elig, k1, hist, selsh);
            const uint32_t lo = prefix;
            uint32_t hi = lo;
            if (k2 != k1) {
                int cle = 0; uint32_t mg = 0xFFFFFFFFu;
#pragma unroll
                for (int i = 0; i < CPT * 8; ++i) { cle += (key[i] <= lo); if (key[i] > lo) mg = min(mg, key[i]); }
                cle = red.sum(cle);
                mg = red.minu(mg);
                if (cle <= k2) hi = mg;
            }
            const float cm = psr_mid(psr_key2f(lo), psr_key2f(hi));
            if (!(fabsf(cm) > c.cormax)) {
#pragma unroll
                for (int j = 0; j < CPT; ++j) {
                    const int ch = tid + j * NT;
                    const int r = ch / CH, cc = ch - r * CH;
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        if (key[j * 8 + i] != KEY_INELIGIBLE) {
                            const int si = r * ST + cc * 8 + i;
                            a[si] = __fsub_rn(a[si], cm);
                        }
                }
            }
        }
    }
    __syncthreads();

    // ---- mode 1: per-row medians within the bank (common_mode_rows: n > npix_min and |cm| < cormax)
    // One THREAD per row: the row's 48 / 64 keys go into registers and through a Batcher sorting network
    // (sortnet.cuh; ineligible pixels are +inf keys that sink to the end), ~8x fewer instructions than a
    // warp-wide counting search per row.  Row r of the tile is read with stride ST = TC + 1: conflict-free.
    if (mode & 1) {
        for (int r = tid; r < TR; r += NT) {
            uint32_t k[TC]; int n = 0;
#pragma unroll
            for (int col = 0; col < TC; ++col) {
                const bool e = (fl[r * ST + col] & 2u) != 0;
                k[col] = e ? psr_f2key(a[r * ST + col]) : KEY_INELIGIBLE;
                n += e ? 1 : 0;
            }
            if (n > c.npix_min && n > 0) {
                if (TC == 48) psr_sort48(reinterpret_cast<uint32_t (&)[48]>(k)); else psr_sort64(reinterpret_cast<uint32_t (&)[64]>(k));
                const int k1 = (n - 1) >> 1, k2 = n >> 1;           // both <= TC / 2
                uint32_t lo = k[0], hi = k[0];
#pragma unroll
                for (int i = 1; i <= TC / 2; ++i) { if (i == k1) lo = k[i]; if (i == k2) hi = k[i]; }
                const float cm = psr_mid(psr_key2f(lo), psr_key2f(hi));
                if (fabsf(cm) < c.cormax) {
#pragma unroll
                    for (int col = 0; col < TC; ++col)
                        if (fl[r * ST + col] & 2u) a[r * ST + col] = __fsub_rn(a[r * ST + col], cm);
                }
            }
        }
        __syncthreads();
    }

    // ---- mode 2: per-column medians within the half (common_mode_cols: same guards)
    // One warp per column (lane = row mod 32); radix select on a linear image of the values (warp_kth_linear).
    if (mode & 2) {
        constexpr int NK = (TR + 31) / 32;
        const int warp = tid >> 5;
        uint32_t* wh = whist + warp * 256;
        float centre = 0.0f;                                       // of the one-pass window: the warp's previous column's median
        for (int col = warp; col < TC; col += NW) {
            float xv[NK]; uint32_t el = 0; int cnt = 0;
#pragma unroll
            for (int i = 0; i < NK; ++i) {
                const int r = lane + 32 * i;
                xv[i] = 0.0f;
                if (r < TR) {
                    xv[i] = a[r * ST + col];
                    if (fl[r * ST + col] & 2u) { el |= 1u << i; ++cnt; }
                }
            }
            const int n = __reduce_add_sync(0xFFFFFFFFu, cnt);
            if (n > c.npix_min && n > 0) {                         // warp-uniform
                float flo, fhi;
                if (!PSR_CMW_COL_LQ || !warp_median_lq<NK>(xv, el, n, centre, wh, flo, fhi)) {   // rare: values outside the window, or a crowded bin
                    uint32_t k[NK];
#pragma unroll
                    for (int i = 0; i < NK; ++i) k[i] = ((el >> i) & 1u) ? psr_f2key(xv[i]) : KEY_INELIGIBLE;
                    const uint32_t lo = warp_kth_linear<NK>(k, (n - 1) >> 1, wh);
                    uint32_t hi = lo;
                    if ((n & 1) == 0) {                                // even count: the upper middle value
                        int cle = 0; uint32_t mg = 0xFFFFFFFFu;
#pragma unroll
                        for (int i = 0; i < NK; ++i) { cle += (k[i] <= lo); if (k[i] > lo) mg = min(mg, k[i]); }
                        cle = __reduce_add_sync(0xFFFFFFFFu, cle);
                        mg = __reduce_min_sync(0xFFFFFFFFu, mg);
                        if (cle <= (n >> 1)) hi = mg;
                    }
                    flo = psr_key2f(lo); fhi = psr_key2f(hi);
                }
                const float cm = psr_mid(flo, fhi);
                if (fabsf(cm) < 1.0e6f) centre = cm;
                if (fabsf(cm) < c.cormax) {
#pragma unroll
                    for (int i = 0; i < NK; ++i) {
                        const int r = lane + 32 * i;
                        if ((el >> i) & 1u) a[r * ST + col] = __fsub_rn(xv[i], cm);
                    }
                }
            }
        }
        __syncthreads();
    }

    // ---- epilogue: arrf * factor * mask, keep-select, stream out ---------------------------------
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const int ch = tid + j * NT;
        const int r = ch / CH, cc = ch - r * CH;
        const int gp = ((seg * ROWS + half * TR + r) * COLS) + bank * TC + cc * 8;
        float ga[8], gb_[8], gc[8];
        if (DET == 0) {
            ld8(c.gf0 + gp, ga);
            ld8(c.gf1 + gp, gb_);
        } else {
            ld8(c.gfac + gp, ga);
            ld8(c.gfac + c.npix + gp, gb_);
            ld8(c.gfac + 2 * (size_t)c.npix + gp, gc);
        }
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int si = r * ST + cc * 8 + i;
            const uint32_t f = fl[si];
            const uint32_t sel = f >> 4;
            float gf;
            if (DET == 0) gf = sel ? gb_[i] : ga[i];
            else gf = cm_selp(cm_selp(1.0f, gc[i], sel & 1u), cm_selp(gb_[i], ga[i], sel & 1u), sel >> 1);
            const float v = __fmul_rn(__fmul_rn(a[si], gf), (f & PSR_FL_MASK) ? 1.0f : 0.0f);
            o[i] = (f & PSR_FL_KEEP) ? v : 0.0f;
        }
        __stcs(reinterpret_cast<float4*>(outf + gp), make_float4(o[0], o[1], o[2], o[3]));
        __stcs(reinterpret_cast<float4*>(outf + gp + 4), make_float4(o[4], o[5], o[6], o[7]));
    }
}


// ---- image-mode epilogue: the tile's pixels go straight to their places in the assembled image -----------------
// A run of CW consecutive floats at any 4-byte alignment (image rows need not be a multiple of 4 wide), with the
// widest stores the address allows.
template <int CW>
__device__ __forceinline__ void img_store_run(float* p, const float (&o)[8])
{
    const unsigned a = (unsigned)(reinterpret_cast<uintptr_t>(p) >> 2) & 3u;
    if (CW == 8) {
        if (a == 0)      { __stcs(reinterpret_cast<float4*>(p), make_float4(o[0], o[1], o[2], o[3])); __stcs(reinterpret_cast<float4*>(p + 4), make_float4(o[4], o[5], o[6], o[7])); }
        else if (a == 2) { __stcs(reinterpret_cast<float2*>(p), make_float2(o[0], o[1])); __stcs(reinterpret_cast<float4*>(p + 2), make_float4(o[2], o[3], o[4], o[5])); __stcs(reinterpret_cast<float2*>(p + 6), make_float2(o[6], o[7])); }
        else if (a == 1) { __stcs(p, o[0]); __stcs(reinterpret_cast<float2*>(p + 1), make_float2(o[1], o[2])); __stcs(reinterpret_cast<float4*>(p + 3), make_float4(o[3], o[4], o[5], o[6])); __stcs(p + 7, o[7]); }
        else             { __stcs(p, o[0]); __stcs(reinterpret_cast<float4*>(p + 1), make_float4(o[1], o[2], o[3], o[4])); __stcs(reinterpret_cast<float2*>(p + 5), make_float2(o[5], o[6])); __stcs(p + 7, o[7]); }
    } else {
        if (a == 0)      { __stcs(reinterpret_cast<float4*>(p), make_float4(o[0], o[1], o[2], o[3])); }
        else if (a == 2) { __stcs(reinterpret_cast<float2*>(p), make_float2(o[0], o[1])); __stcs(reinterpret_cast<float2*>(p + 2), make_float2(o[2], o[3])); }
        else if (a == 1) { __stcs(p, o[0]); __stcs(reinterpret_cast<float2*>(p + 1), make_float2(o[1], o[2])); __stcs(p + 3, o[3]); }
        else             { __stcs(p, o[0]); __stcs(reinterpret_cast<float2*>(p + 1), make_float2(o[1], o[2])); __stcs(p + 3, o[3]); }
    }
}

// one streaming store at base[t]: a single multiply-add forms the address (the compiler otherwise rebuilds the 64-bit
// frame offset for every pixel)
__device__ __forceinline__ void img_st_at(float* base, int32_t t, float v)
{
    asm volatile("{ .reg .u64 a; mad.wide.u32 a, %1, 4, %0; st.global.cs.f32 [a], %2; }" :: "l"(base), "r"(t), "f"(v) : "memory");
}

// Targets of one chunk (PsrDev::fwd): kind 1 = CW consecutive ascending image indices from t0, 2 = descending
// down to t0 (panel rotated by 180 degrees), 3 = every pixel has a target of its own, anywhere (rotated by 90/270:
// plain stores, nothing to test), 0 = some pixel lost a duplicate target or is masked in the image: tested one by one.
template <int CW>
__device__ __forceinline__ uint32_t img_chunk_kind(const int32_t* __restrict__ fwd_gp, int& t0)
{
    int32_t t[8];
    const int4 a = __ldg(reinterpret_cast<const int4*>(fwd_gp));
    t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w;
    if (CW == 8) { const int4 b = __ldg(reinterpret_cast<const int4*>(fwd_gp) + 1); t[4] = b.x; t[5] = b.y; t[6] = b.z; t[7] = b.w; }
    bool asc = true, desc = true, all = t[0] >= 0;
#pragma unroll
    for (int i = 1; i < CW; ++i) { asc = asc && t[i] == t[0] + i; desc = desc && t[i] == t[0] - i; all = all && t[i] >= 0; }
    t0 = asc ? t[0] : t[CW - 1];
    return !all ? 0u : asc ? 1u : desc ? 2u : 3u;
}

template <int CW>
__device__ __forceinline__ void img_store_chunk(float* __restrict__ imgf, const int32_t* __restrict__ fwd_gp, uint32_t kind, int t0,
                                                const float (&o)[8])
{
    if (kind == 1u) img_store_run<CW>(imgf + t0, o);
    else if (kind == 2u) {
        float r[8];
#pragma unroll
        for (int i = 0; i < CW; ++i) r[i] = o[CW - 1 - i];
        img_store_run<CW>(imgf + t0, r);
    } else {
        int32_t t[8];
        const int4 a = __ldg(reinterpret_cast<const int4*>(fwd_gp));
        t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w;
        if (CW == 8) { const int4 b = __ldg(reinterpret_cast<const int4*>(fwd_gp) + 1); t[4] = b.x; t[5] = b.y; t[6] = b.z; t[7] = b.w; }
        if (kind == 3u) {
#pragma unroll
            for (int i = 0; i < CW; ++i) img_st_at(imgf, t[i], o[i]);
        } else {
#pragma unroll
            for (int i = 0; i < CW; ++i) {
                if (t[i] >= 0) img_st_at(imgf, t[i], o[i]);
                else if (t[i] != INT32_MIN) img_st_at(imgf, ~t[i], 0.0f);     // the image keep mask drops this target
            }
        }
    }
}

// =================================================================================================
// mode 4 only (the per-bank 2-d median -- the usual setting): no shared-memory tile at all.  Each
// thread keeps the keys of its CPT*CW pixels in registers from load to store; eligibility, gain
// selector, mask and keep bits are packed in four 32-bit registers.
// =================================================================================================
template <int DET, int ROWS, int COLS, int NB, int NT, int MINB, int CW, bool KS, bool IMG>
__global__ void __launch_bounds__(NT, MINB)
cm_banks_kernel(const uint16_t* __restrict__ raw, float* __restrict__ out, PsrDev c)
{
    constexpr int TR = ROWS / 2, TC = COLS / NB, CH = TC / CW, NCH = TR * CH, CPT = NCH / NT, NK = CPT * CW;   // CW = pixels per chunk (4 or 8)
    constexpr int NW = NT / 32;
    static_assert(NCH % NT == 0 && NT % 32 == 0 && NK <= 32, "tile must split evenly over the CTA");
    constexpr int CAP = NT > 512 ? 512 : NT;
    __shared__ int redbuf[64];
    __shared__ __align__(16) uint32_t hist[LQ_NB];
    __shared__ __align__(16) float list[CAP + 4];
    __shared__ __align__(16) int selsh[48];

    const int tid = threadIdx.x;
    // Block order: bank fastest, then FRAME, then (half, segment).  The CTAs resident at one time then work on all frames of
    // a few half-segments, so the tiles' constants are fetched from DRAM once per batch and come from L2 for the other frames
    // (Jungfrau: 155 MB of constants do not fit L2; with the frame slowest they were re-read for every frame -- ncu: 10.5 GB of
    // DRAM reads per 64 frames), while the banks of one frame's row band still run side by side (whole rows of raw and output).
    int bid = blockIdx.x;
    const int nframes = (int)(gridDim.x / (unsigned)(c.nseg * 2 * NB));
    const int bank = bid % NB; bid /= NB;
    const int frame = bid % nframes; bid /= nframes;
    const int half = bid & 1;
    const int seg = bid >> 1;
    const uint16_t* rawf = raw + (size_t)frame * c.npix;
    float* outf = out + (size_t)frame * (IMG ? (size_t)c.H * c.W : (size_t)c.npix);    // IMG: out is the batch of assembled images

    // KS: the values are parked in shared memory (slot q of thread t at sx[q * NT + t]) instead of registers, which
    // buys another resident CTA per SM
    extern __shared__ float sx[];
    float xr[KS ? 1 : NK];
    const SmemVals<NT> xs{sx + tid};
    for (int i = tid; i < LQ_NB; i += NT) hist[i] = 0u;
    __syncthreads();
    uint32_t elig = 0, sel0 = 0, sel1 = 0;
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const int ch = tid + j * NT;
        const int r = ch / CH, cc = ch - r * CH;
        const int gp = ((seg * ROWS + half * TR + r) * COLS) + bank * TC + cc * CW;
        uint32_t w[4], fw[2];
        ld_raw_flags<CW>(rawf + gp, c.flags + gp, w, fw);
        float pa[8], pb[8], pc[8], oa[8], ob[8], oc[8];
        if (DET == 0) { ldN<CW>(c.ped0 + gp, pa); ldN<CW>(c.ped1 + gp, pb); }
        else {
            ldN<CW>(c.peds + gp, pa);
            if (c.offs != nullptr) ldN<CW>(c.offs + gp, oa);
            ldN<CW>(c.peds + c.npix + gp, pb); ldN<CW>(c.peds + 2 * (size_t)c.npix + gp, pc);
            if (c.offs != nullptr) { ldN<CW>(c.offs + c.npix + gp, ob); ldN<CW>(c.offs + 2 * (size_t)c.npix + gp, oc); }
        }
#pragma unroll
        for (int i = 0; i < CW; ++i) {
            const int q = j * CW + i;
            const uint32_t v = (i & 1) ? (w[i >> 1] >> 16) : (w[i >> 1] & 0xFFFFu);
            const uint32_t f = ((fw[i >> 2] >> (8 * (i & 3))) & 0xFFu) | (c.ignore_keep ? PSR_FL_KEEP : 0u);
            float x = (float)(v & 0x3FFFu);
            uint32_t sel, el;
            if (DET == 0) {
                sel = (v >> 14) & 1u;
                x = __fsub_rn(x, sel ? pb[i] : pa[i]);
                el = sel ? (f & PSR_FL_ELIG1) : (f & PSR_FL_ELIG0);
            } else {
                const uint32_t gb = v >> 14;
                sel = gb == 0 ? 0u : gb == 1 ? 1u : gb == 3 ? 2u : 3u;
                float ped = 0.f, off = 0.f;
                if (sel == 0) { ped = pa[i]; if (c.offs) off = oa[i]; }
                else if (sel == 1) { ped = pb[i]; if (c.offs) off = ob[i]; }
                else if (sel == 2) { ped = pc[i]; if (c.offs) off = oc[i]; }
                x = __fsub_rn(x, ped);
                if (c.offs != nullptr) x = __fsub_rn(x, off);
                el = (sel < 3 && ((c.elig_mask >> sel) & 1u) && (f & PSR_FL_MASK)) ? 1u : 0u;
            }
            if constexpr (KS) sx[q * NT + tid] = x; else xr[q] = x;
            elig |= (el ? 1u : 0u) << q;
            sel0 |= (sel & 1u) << q;
            if (DET != 0) sel1 |= (sel >> 1) << q;
        }
    }

    // common_mode_2d: skip if n < npix_min, skip if |cm| > cormax
    float cm = 0.f, lo, hi; bool apply = false;
    LqWindow win{LQ_A, 0.5f};
    bool have;
    if constexpr (KS) have = block_median_linear<NT, NK, CAP, BitsLinear>(xs, elig, c.npix_min, win, hist, selsh, list, redbuf, lo, hi, CtaGroup{});
    else have = block_median_linear<NT, NK, CAP, BitsLinear>(xr, elig, c.npix_min, win, hist, selsh, list, redbuf, lo, hi, CtaGroup{});
    if (have) {                                            // uniform across the CTA
        cm = psr_mid(lo, hi);
        apply = !(fabsf(cm) > c.cormax);
    }

#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const int ch = tid + j * NT;
        const int r = ch / CH, cc = ch - r * CH;
        const int gp = ((seg * ROWS + half * TR + r) * COLS) + bank * TC + cc * CW;
        float ga[8], gb_[8], gc[8];
        if (DET == 0) { ldN<CW>(c.gf0 + gp, ga); ldN<CW>(c.gf1 + gp, gb_); }
        else {
            ldN<CW>(c.gfac + gp, ga);
            ldN<CW>(c.gfac + c.npix + gp, gb_); ldN<CW>(c.gfac + 2 * (size_t)c.npix + gp, gc);
        }
        uint32_t fw[2];
        ld_flags<CW>(c.flags + gp, fw);                                              // L1/L2 hit: read in the load phase
        float o[8];
#pragma unroll
        for (int i = 0; i < CW; ++i) {
            const int q = j * CW + i;
            const uint32_t f = (fw[i >> 2] >> (8 * (i & 3))) | (c.ignore_keep ? PSR_FL_KEEP : 0u);
            float x;
            if constexpr (KS) x = xs[q]; else x = xr[q];
            if (apply && ((elig >> q) & 1u)) x = __fsub_rn(x, cm);
            const uint32_t sel = ((sel0 >> q) & 1u) | (((sel1 >> q) & 1u) << 1);
            float gf;
            if (DET == 0) gf = sel ? gb_[i] : ga[i];
            else gf = sel == 0 ? ga[i] : sel == 1 ? gb_[i] : sel == 2 ? gc[i] : 1.0f;
            const float v = __fmul_rn(__fmul_rn(x, gf), (f & PSR_FL_MASK) ? 1.0f : 0.0f);
            o[i] = (f & PSR_FL_KEEP) ? v : 0.0f;
        }
        if constexpr (IMG) {
            int t0; const uint32_t kind = img_chunk_kind<CW>(c.fwd + gp, t0);
            img_store_chunk<CW>(outf, c.fwd + gp, kind, t0, o);
        } else {
            __stcs(reinterpret_cast<float4*>(outf + gp), make_float4(o[0], o[1], o[2], o[3]));
            if (CW == 8) __stcs(reinterpret_cast<float4*>(outf + gp + 4), make_float4(o[4], o[5], o[6], o[7]));
        }
    }
}

// =================================================================================================
// Many small medians at once (the row and column sub-modes): one linear histogram PER SEGMENT -- a row of a bank
// (48 values) or a column of a half (176 values) -- with byte counters (a segment holds at most 176 values), all
// segments of the tile filled in one pass over the values the threads hold in registers.  As for the bank median
// the bin function is monotone, so the bin where a segment's cumulative count crosses rank (n - 1) / 2 holds that
// order statistic; its members -- and those of the next occupied bin when the upper middle value lies there -- are
// collected per segment and ranked.  A segment whose rank falls into an end bin, or whose bins hold more than CAP
// values, is redone exactly, one at a time, from a list of all its values (seg_slow).
//
// Shared memory per group (SegBuf): `buf` holds the histograms (NSEG x NB bytes) and, once they are scanned, the
// candidate lists (NSEG x CAP floats) in the same place.
// =================================================================================================
constexpr int SEG_MAX = 176;                       // most segments of one phase (rows of a bank tile)
constexpr int SEG_BUF_BYTES = 13056;               // rows: 176 x 64 bins / 176 x 16 floats; columns: 48 x (256 bins + 16 bytes of skew) / 48 x 64 floats
constexpr int SEG_SMEM_BYTES = SEG_BUF_BYTES + 4 * (3 * SEG_MAX + 8 + SEG_MAX);

struct SegBuf {
    uint32_t* buf;       // histograms, then candidate lists
    uint32_t* info;      // per segment: B | B2 << 8 | r << 16 | n << 24 (n = 0: nothing to collect)
    int* cnt;            // per segment: candidates collected
    float* cmv;          // per segment: the correction its eligible pixels subtract (0 = none)
    int* misc;           // 0: failed segments, 1: values in end bins, 2: values in all, 3: slow-path counter, 4: sum of medians (float bits)
    int* fails;          // ids of the failed segments
    __device__ __forceinline__ explicit SegBuf(uint8_t* base) {
        buf = reinterpret_cast<uint32_t*>(base);
        info = buf + SEG_BUF_BYTES / 4; cnt = reinterpret_cast<int*>(info + SEG_MAX);
        cmv = reinterpret_cast<float*>(cnt + SEG_MAX); misc = reinterpret_cast<int*>(cmv + SEG_MAX); fails = misc + 8;
    }
};

template <int NB>
__device__ __forceinline__ uint32_t lq_bin8(float x, float a, float b) {            // bin of x in an NB-bin window (NB <= 256)
    return __float_as_uint(__fmaf_rd(__saturatef(__fmaf_rn(x, a, b)), (float)(NB - 1), 8388608.0f)) & 0xFFu;
}

// psana's guards for the row / column corrections: apply only where n > npix_min and |cm| < cormax (both strict)
__device__ __forceinline__ float seg_guard(float lo, float hi, int n, int npix_min, float cormax) {
    const float cm = psr_mid(lo, (n & 1) ? lo : hi);
    return (n > npix_min && fabsf(cm) < cormax) ? cm : 0.0f;
}

// Exact redo of one segment from the list of ALL its eligible values (in sb.buf, n of them): warp 0 ranks them by counting.
template <typename G>
__device__ __noinline__ void seg_slow_rank(const SegBuf sb, int seg, int npix_min, float cormax, G g)
{
    const int tid = g.tid(), lane = tid & 31;
    const int n = sb.misc[3];
    if (tid < 32) {
        const float* v = reinterpret_cast<const float*>(sb.buf);
        const int k1 = (n - 1) >> 1, k2 = n >> 1;
        float lo = 0.f, hi = 0.f; bool flo = false, fhi = false;
        for (int i = lane; i < n; i += 32) {
            const float vi = v[i];
            int lt = 0, le = 0;
            for (int j = 0; j < n; ++j) { const float u = v[j]; lt += (u < vi); le += (u <= vi); }
            if (lt <= k1 && k1 < le) { lo = vi; flo = true; }
            if (lt <= k2 && k2 < le) { hi = vi; fhi = true; }
        }
        const uint32_t mlo = __ballot_sync(0xFFFFFFFFu, flo), mhi = __ballot_sync(0xFFFFFFFFu, fhi);
        if (n > 0) {
            lo = __shfl_sync(0xFFFFFFFFu, lo, __ffs(mlo) - 1);
            hi = __shfl_sync(0xFFFFFFFFu, hi, __ffs(mhi) - 1);
            if (lane == 0) sb.cmv[seg] = seg_guard(lo, hi, n, npix_min, cormax);
        } else if (lane == 0) sb.cmv[seg] = 0.0f;
    }
    g.sync();
}

// x: the thread's NK values; elig: bit BP::pos(i) = value i takes part; SEG::of(i) = the segment of value i (SEG::base(j) may
// hoist per-chunk work).  On return sb.cmv[seg] holds every segment's correction (guards applied); win is updated.
// Every thread of the group must call it.
template <int GT, int NK, int NSEG, int NB, int TPS, int CAP, typename BP, typename SEG, typename G>
__device__ __forceinline__ void seg_medians(const float (&x)[NK], uint32_t elig, const SEG& sg, int npix_min, float cormax, LqWindow& win,
                                            const SegBuf sb, const G& g)
{
    // a segment's histogram starts STRIDE bytes after its neighbour's: with NB a multiple of 128 the plain layout would put
    // bin b of EVERY segment into the same bank, and a warp's lanes -- different segments, similar bins -- would collide
    constexpr int STRIDE = (NB % 128 == 0) ? NB + 16 : NB;
    static_assert(NSEG * STRIDE <= SEG_BUF_BYTES && NSEG * CAP * 4 <= SEG_BUF_BYTES && NSEG <= SEG_MAX && NSEG * TPS <= GT, "segment layout");
    static_assert((NB / 4) % TPS == 0 && (NB / 4 / TPS) % 4 == 0 && (TPS & (TPS - 1)) == 0 && NB <= 256 && CAP % 4 == 0, "scan layout");
    constexpr int W = NB / 4 / TPS;                  // histogram words one scanning thread owns
    const int tid = g.tid(), lane = tid & 31;
    const float a = win.a, b = win.b;
    const uint32_t baddr = (uint32_t)__cvta_generic_to_shared(sb.buf);

    // ---- clear ---------------------------------------------------------------------------------------------------
    for (int i = tid; i < NSEG * STRIDE / 16; i += GT) reinterpret_cast<uint4*>(sb.buf)[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < NSEG; i += GT) sb.cnt[i] = 0;
    if (tid < 8) sb.misc[tid] = 0;
    g.sync();
    // ---- histograms: byte counters, one 32-bit add per value (ineligible: add 0) -----------------------------------
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        const uint32_t bin = lq_bin8<NB>(x[i], a, b);
        const uint32_t m = (elig << (31 - BP::pos(i))) >> 31;
        asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(baddr + (uint32_t)sg.of(i) * STRIDE + (bin & ~3u)), "r"(m << ((bin & 3u) * 8u)) : "memory");
    }
    g.sync();
    // ---- scan: TPS threads per segment ------------------------------------------------------------------------------
    const int seg = tid / TPS, part = tid % TPS;
    uint32_t myinfo = 0; bool writer = false, failed = false;
    if (tid < NSEG * TPS) {
        uint32_t w[W];
        const uint4* h4 = reinterpret_cast<const uint4*>(sb.buf + seg * (STRIDE / 4) + part * W);
#pragma unroll
        for (int i = 0; i < W / 4; ++i) { const uint4 u = h4[i]; w[4 * i] = u.x; w[4 * i + 1] = u.y; w[4 * i + 2] = u.z; w[4 * i + 3] = u.w; }
        uint32_t ps = 0; int fne = 255;              // bytewise sums (no byte exceeds the segment's size < 256); first occupied bin of this part
#pragma unroll
        for (int i = W - 1; i >= 0; --i) { ps += w[i]; if (w[i]) fne = 4 * (part * W + i) + ((__ffs(w[i]) - 1) >> 3); }
        const uint32_t ptot = (ps * 0x01010101u) >> 24;
        uint32_t n = 0, carry = 0; int fnext = 255;
#pragma unroll
        for (int p = TPS - 1; p >= 0; --p) {
            const uint32_t tp = __shfl_sync(0xFFFFFFFFu, ptot, (lane & ~(TPS - 1)) + p);
            const int fp = __shfl_sync(0xFFFFFFFFu, fne, (lane & ~(TPS - 1)) + p);
            n += tp; if (p < part) carry += tp; if (p > part && fp != 255) fnext = fp;          // descending p: ends at the nearest higher part
        }
        if (part == 0) {                              // window statistics: what fell into the two end bins
            int tails = (int)(w[0] & 0xFFu);
            if (TPS == 1) tails += (int)(w[W - 1] >> 24);
            if (tails) atomicAdd(&sb.misc[1], tails);
            if (n) atomicAdd(&sb.misc[2], (int)n);
        } else if (part == TPS - 1) { const int tails = (int)(w[W - 1] >> 24); if (tails) atomicAdd(&sb.misc[1], tails); }
        const uint32_t k1 = n ? (n - 1) >> 1 : 0;
        if (n == 0) { writer = part == 0; myinfo = 0; }
        else if (k1 >= carry && k1 < carry + ptot) {
            writer = true;
            // the word in which the cumulative count crosses k1, then its four bytes once; the next occupied bin after it
            uint32_t cum = carry, wsel = 0, cumsel = 0; int isel = -1, B = -1, B2 = 255; uint32_t r = 0, M = 0;
#pragma unroll
            for (int i = 0; i < W; ++i) {
                const uint32_t tw = (w[i] * 0x01010101u) >> 24;
                if (isel < 0 && k1 < cum + tw) { isel = i; wsel = w[i]; cumsel = cum; }
                else if (isel >= 0 && B2 == 255 && w[i]) B2 = 4 * (part * W + i) + ((__ffs(w[i]) - 1) >> 3);
                cum += tw;
            }
            {
                const uint32_t pre = wsel * 0x01010101u;                          // byte k = the word's bytes 0..k summed
#pragma unroll
                for (int k = 3; k >= 0; --k) {
                    const uint32_t ck = cumsel + ((pre >> (8 * k)) & 0xFFu), bk = (wsel >> (8 * k)) & 0xFFu;
                    if (ck > k1 && bk != 0u && ck - bk <= k1) { B = 4 * (part * W + isel) + k; M = bk; r = k1 - (ck - bk); }
                }
                const int kb = B & 3;
                const uint32_t rest = kb < 3 ? (wsel >> (8 * (kb + 1))) : 0u;
                if (rest) B2 = B + 1 + ((__ffs(rest) - 1) >> 3);
            }
            if (B2 == 255) B2 = fnext;
            const bool need_hi = (n & 1u) == 0u && r + 1 >= M;                       // the upper middle value lies in the next occupied bin
            uint32_t M2 = 0;
            if (need_hi && B2 != 255) M2 = reinterpret_cast<const uint8_t*>(sb.buf)[seg * STRIDE + B2];
            failed = B <= 0 || B >= NB - 1 || (need_hi && (B2 == 255 || B2 >= NB - 1)) || M + M2 > (uint32_t)CAP;
            myinfo = failed ? 0u : ((uint32_t)B | ((need_hi ? (uint32_t)B2 : (uint32_t)B) << 8) | (r << 16) | (n << 24));
        }
    }
    g.sync();                                         // every histogram word has been read: buf becomes the candidate lists
    if (writer) {
        sb.info[seg] = myinfo;
        if (failed) sb.fails[atomicAdd(&sb.misc[0], 1)] = seg;
    }
    g.sync();
    // ---- collect the members of each segment's one or two bins ------------------------------------------------------
    float* lists = reinterpret_cast<float*>(sb.buf);
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        const int s = sg.of(i);
        const uint32_t inf = sb.info[s], bin = lq_bin8<NB>(x[i], a, b);
        if ((inf >> 24) != 0u && (elig & (1u << BP::pos(i))) && (bin == (inf & 0xFFu) || bin == ((inf >> 8) & 0xFFu))) {
            const int slot = atomicAdd(&sb.cnt[s], 1);
            if (slot < CAP) lists[s * CAP + slot] = x[i];
        }
    }
    g.sync();
    // ---- rank: the TPS threads of a segment share its candidates ----------------------------------------------------
    if (tid < NSEG * TPS) {
        const uint32_t inf = sb.info[seg];
        const int n = (int)(inf >> 24), r = (int)((inf >> 16) & 0xFFu), M = min(sb.cnt[seg], CAP);
        const float* L = lists + seg * CAP;
        uint32_t lo = 0, hi = 0; bool flo = false, fhi = false;
        if (n > 0) {
            for (int i = part; i < M; i += TPS) {
                const float v = L[i];
                int lt = 0, le = 0;
                for (int j = 0; j < M; ++j) { const float u = L[j]; lt += (u < v); le += (u <= v); }
                if (lt <= r && r < le) { lo = __float_as_uint(v); flo = true; }
                if (lt <= r + 1 && r + 1 < le) { hi = __float_as_uint(v); fhi = true; }
            }
        }
#pragma unroll
        for (int d = 1; d < TPS; d <<= 1) {           // any finder's value (equal values rank equal)
            const uint32_t olo = __shfl_xor_sync(0xFFFFFFFFu, lo, d), ohi = __shfl_xor_sync(0xFFFFFFFFu, hi, d);
            const bool oflo = __shfl_xor_sync(0xFFFFFFFFu, (int)flo, d) != 0, ofhi = __shfl_xor_sync(0xFFFFFFFFu, (int)fhi, d) != 0;
            if (!flo && oflo) { lo = olo; flo = true; }
            if (!fhi && ofhi) { hi = ohi; fhi = true; }
        }
        if (part == 0) sb.cmv[seg] = n > 0 ? seg_guard(__uint_as_float(lo), __uint_as_float(fhi ? hi : lo), n, npix_min, cormax) : 0.0f;
    }
    g.sync();
    // ---- the segments that did not fit the window, exactly, one at a time (uniform; none on ordinary data) ---------
    const int nfail = sb.misc[0];
    for (int f = 0; f < nfail; ++f) {
        const int fs = sb.fails[f];
        if (tid == 0) sb.misc[3] = 0;
        g.sync();
#pragma unroll
        for (int i = 0; i < NK; ++i)
            if (sg.of(i) == fs && (elig & (1u << BP::pos(i)))) lists[atomicAdd(&sb.misc[3], 1)] = x[i];
        g.sync();
        seg_slow_rank(sb, fs, npix_min, cormax, g);
    }
    // ---- next frame's window: twice as wide when an eighth of the values fell outside, half when nothing did ------
    const int tails = sb.misc[1], ntot = sb.misc[2];
    float na = a;
    if (tails * 2 > ntot) na = a * 0.25f;
    else if (tails * 8 > ntot || nfail * 16 > NSEG) na = a * 0.5f;
    else if (tails * 64 < ntot && nfail == 0) na = a * 2.0f;
    win.a = fminf(fmaxf(na, LQ_A_MIN), 1.0f / 32.0f);
    win.b = 0.5f;                                     // rows and columns are corrected around the bank level: centred on 0
    g.sync();                                         // cmv is final and misc may be reused
}

// the segment of value i of thread t of a group of GT threads (chunk j = i / 4 of the thread is chunk t + GT j of the tile, 12 chunks per row)
template <int GT>
struct SegRows {            // segment = row of the tile (48 values)
    int t;
    __device__ __forceinline__ int of(int i) const { return (t + GT * (i >> 2)) / 12; }
};
template <int GT>
struct SegCols {            // segment = column of the tile (176 values)
    int t;
    __device__ __forceinline__ int of(int i) const { const int ch = t + GT * (i >> 2); return 4 * (ch - 12 * (ch / 12)) + (i & 3); }
};

// Row medians of one bank tile by sorting: the tile's float keys sit in shared memory (row stride 49: conflict-free),
// one thread per row takes its 48 keys through the generated Batcher network (ineligible pixels are +inf keys that sink
// to the end).  Cheaper than per-row histograms for segments this small (measured: 580 -> ~200 us per 64 frames for the
// row phase).  Not inlined: the 48 keys live in this function's registers only.  cmv[r] = the row's correction (guards
// applied).  Every thread of the group must call it.
template <typename G>
__device__ __noinline__ void rows_sort_phase(const uint32_t* tile, float* cmv, int npix_min, float cormax, G g)
{
    const int r = g.tid();
    if (r < 176) {
        uint32_t k[48]; int n = 0;
#pragma unroll
        for (int col = 0; col < 48; ++col) { k[col] = tile[r * 49 + col]; n += k[col] != KEY_INELIGIBLE ? 1 : 0; }
        float cm = 0.0f;
        if (n > npix_min && n > 0) {
            psr_sort48(k);
            const int k1 = (n - 1) >> 1, k2 = n >> 1;               // both <= 24
            uint32_t lo = k[0], hi = k[0];
#pragma unroll
            for (int i = 1; i <= 24; ++i) { if (i == k1) lo = k[i]; if (i == k2) hi = k[i]; }
            const float med = psr_mid(psr_key2f(lo), psr_key2f(hi));
            if (fabsf(med) < cormax) cm = med;
        }
        cmv[r] = cm;
    }
    g.sync();
}

// =================================================================================================
// epix10ka, mode 4, TWO FRAME GROUPS: one 704-thread CTA per SM owns one bank tile and walks `fpc` frames of the
// batch, with ALL of the tile's constants (both pedestal and both gain-factor entries, 135 KB) staged once in
// shared memory, so the frame loop's global traffic is raw in and floats out, nothing else.  The CTA is two groups of 352 threads that each take
// every other frame of the CTA's share and synchronise only among themselves (hardware barriers 1 and 2):
// the groups drift apart and fill each other's barrier and latency gaps, like two CTAs that share their
// 135 KB of constants.  A thread holds 24 pixels (six 4-pixel chunks) of its group's frame.
// =================================================================================================
#ifndef PSR_IMG_RUNS     /* ... and no per-chunk tests on tiles that are all aligned ascending / descending runs */
#define PSR_IMG_RUNS 1
#endif
#ifndef PSR_IMG_ROWOWN   /* image output of the two-group kernel: row ownership on panels rotated by 90 / 270 degrees */
#define PSR_IMG_ROWOWN 1
#endif
template <bool IMG, bool KEEP, bool WORDS, int NG>
__global__ void __launch_bounds__(704, 1)
cm_banks_epix_groups_kernel(const uint16_t* __restrict__ raw, float* __restrict__ out, PsrDev c, int nframes, int fpc)
{
    constexpr int NT = 704, GT = NT / NG, ROWS = 352, COLS = 384, NB = 8, TR = ROWS / 2, TC = COLS / NB, CH = TC / 4, NCH = TR * CH;
    constexpr int CPT = NCH / GT, NK = CPT * 4, CAP = GT > 512 ? 512 : GT;
    using BP = BitsEpixWords4;
    static_assert(NCH % NT == 0 && NCH % GT == 0 && (NG == 1 || NG == 2), "tile layout");
    extern __shared__ float4 sm4[];                               // [4][NCH]: ped0, ped1, gf0, gf1 of the 4 pixels of a chunk
    int* sgp = reinterpret_cast<int*>(sm4 + 4 * NCH);             // [CPT][GT] first pixel of a thread's chunk; IMG: + [CPT][GT] its image index
    __shared__ int redbuf[NG][64];
    static_assert(!(IMG && WORDS), "row / column words write the calib layout (images go through the scratch + gather)");
    __shared__ __align__(16) uint32_t hist[NG][LQ_NB];
    __shared__ __align__(16) float list[NG][CAP + 4];
    __shared__ __align__(16) int selsh[NG][48];

    const int tid = threadIdx.x, grp = tid / GT, t = tid - grp * GT;
    int bid = blockIdx.x;
    const int bank = bid % NB; bid /= NB;
    const int half = bid & 1;  bid >>= 1;
    const int seg = bid;
    const int f_begin = blockIdx.y * fpc, f_end = min(nframes, f_begin + fpc);
    const int tile0 = ((seg * ROWS + half * TR) * COLS) + bank * TC;

    for (int i = tid; i < NG * LQ_NB; i += NT) (&hist[0][0])[i] = 0u;
    // IMG: is the tile's place in the image t00 + r * dr + c * dc with dr = +-1 for every pixel -- detector rows running
    // along image rows' x, the panel rotated by 90 / 270 degrees?  Then the threads own the tile differently (below).
    int rdc = 0;                                                  // dc if so, else 0
    if constexpr (IMG && PSR_IMG_ROWOWN) {
        const int t00 = __ldg(c.fwd + tile0), dr = __ldg(c.fwd + tile0 + COLS) - t00, dc = __ldg(c.fwd + tile0 + 1) - t00;
        bool affine = t00 >= 0 && (dr == 1 || dr == -1);
#pragma unroll
        for (int j = 0; j < NCH / NT; ++j) {
            const int ch = tid + j * NT;
            const int r = ch / CH, cc = ch - r * CH;
            const int4 a = __ldg(reinterpret_cast<const int4*>(c.fwd + tile0 + r * COLS + cc * 4));
            const int e = t00 + r * dr + cc * 4 * dc;
            affine = affine && a.x == e && a.y == e + dc && a.z == e + 2 * dc && a.w == e + 3 * dc;
        }
        if (__syncthreads_and(affine ? 1 : 0)) rdc = dc;
    }
    // Which chunk is thread tt's j-th.  Normally chunk tt + j * GT: a warp instruction covers 32 consecutive chunks.  On a
    // rotated tile a warp owns 16 whole tile rows, lane l row l & 15: its chunks 2m, 2m + 1 are the 16 bytes of raw at chunk
    // 4m + 2 (l >> 4) of that row (one 16-byte load; two lanes share a 32-byte sector), and pixel i of chunk j of lanes 0-15
    // / 16-31 is 16 consecutive floats of one image row each: stores of 64-byte runs instead of 12-byte ones.
    auto chunk_of = [&](int tt, int j) {
        if (IMG && PSR_IMG_ROWOWN && rdc != 0) return ((tt >> 5) * 16 + (tt & 15)) * CH + 4 * (j >> 1) + 2 * ((tt >> 4) & 1) + (j & 1);
        return tt + j * GT;
    };
    // the tile's constants, by all 704 threads, thread-major: [table][j][thread] (whatever the ownership, a warp reads 32
    // consecutive float4); the mask is folded into the gain factors (psr_set_constants)
#pragma unroll
    for (int m = 0; m < NCH / NT; ++m) {
        const int slot = tid + m * NT;                            // = j * GT + tt
        const int j = slot / GT, tt = slot - j * GT;
        const int ch = chunk_of(tt, j);
        const int r = ch / CH, cc = ch - r * CH;
        const int gp = tile0 + r * COLS + cc * 4;
        sm4[0 * NCH + slot] = __ldg(reinterpret_cast<const float4*>(c.ped0 + gp));
        sm4[1 * NCH + slot] = __ldg(reinterpret_cast<const float4*>(c.ped1 + gp));
        sm4[2 * NCH + slot] = __ldg(reinterpret_cast<const float4*>(c.gf0 + gp));
        sm4[3 * NCH + slot] = __ldg(reinterpret_cast<const float4*>(c.gf1 + gp));
    }
    uint32_t kbits = 0, e0bits = 0, e1bits = 0;                   // bit BP::pos(i) = value i of this thread
    uint32_t kinds = 0;                                           // IMG: where each chunk goes in the assembled image is the same for every frame
    bool run_asc = true, run_desc = true, run_a16 = true;
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const int ch = chunk_of(t, j);
        const int r = ch / CH, cc = ch - r * CH;
        const int gp = tile0 + r * COLS + cc * 4;
        if (grp == 0) sgp[j * GT + t] = gp;
        const uint32_t f = __ldg(reinterpret_cast<const uint32_t*>(c.flags + gp));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t fb = (f >> (8 * i)) & 0xFFu;
            const int q = BP::pos(j * 4 + i);
            if (KEEP) kbits |= ((fb & PSR_FL_KEEP) ? 1u : 0u) << q;
            e0bits |= ((fb & PSR_FL_ELIG0) ? 1u : 0u) << q;
            e1bits |= ((fb & PSR_FL_ELIG1) ? 1u : 0u) << q;
        }
        if constexpr (IMG) {
            int t0; const uint32_t kind = img_chunk_kind<4>(c.fwd + gp, t0);
            kinds |= kind << (2 * j);
            run_asc = run_asc && kind == 1u && (t0 & 1) == 0; run_desc = run_desc && kind == 2u && (t0 & 1) == 0; run_a16 = run_a16 && (t0 & 3) == 0;
            if (PSR_IMG_ROWOWN && rdc != 0) t0 = __ldg(c.fwd + gp);       // rotated tile: pixel i of the chunk goes to t0 + i * dc
            if (grp == 0) sgp[(CPT + j) * GT + t] = t0;
        }
    }
    int runcls = 0; bool al16 = false;                            // see OWN in the frame loop
    if constexpr (IMG && PSR_IMG_RUNS) {
        const int na = __syncthreads_and(run_asc ? 1 : 0), nd = __syncthreads_and(run_desc ? 1 : 0);
        al16 = __syncthreads_and(run_a16 ? 1 : 0) != 0;
        runcls = na ? 2 : nd ? 3 : 0;
    }
    const size_t fout = IMG ? (size_t)c.H * c.W : (size_t)c.npix;
    // always 0 (this kernel is epix10ka's), but not to the compiler: lets a value be made to depend on another, which
    // keeps ptxas from hoisting every chunk's loads to the top and spilling the values it has already computed
    const uint32_t zero = (uint32_t)c.det_kind;
    __syncthreads();

    const WarpGroup<GT> g{t, 1 + grp};
    LqWindow win{LQ_A, 0.5f};                                     // 64 ADU centred on 0, then following the group's medians and their spread
    LqWindow winr{LQ_A, 0.5f}, winc{LQ_A, 0.5f};                  // WORDS: the windows of the row and column histograms
    const SegBuf sb(reinterpret_cast<uint8_t*>(sgp + 2 * CPT * GT) + (WORDS ? grp * SEG_SMEM_BYTES : 0));
    // the frame loop, compiled twice for images: ROT = the rotated tile's ownership (a tile runs one of the two)
    auto frames = [&](auto own_c) {
    // OWN (images): 0 = each chunk by its kind (any geometry), 1 = rotated tile, 2 / 3 = every chunk of the tile is an
    // ascending / descending run at an 8-byte aligned image index (panel flat or turned by 180 degrees): two 8-byte stores
    // per chunk, or one of 16 bytes when all of them are 16-byte aligned -- no per-chunk tests in the frame loop
    constexpr int OWN = decltype(own_c)::value;
    constexpr bool ROT = OWN == 1;
    for (int frame = f_begin + grp; frame < f_end; frame += NG) {
        const uint16_t* rawf = raw + (size_t)frame * c.npix;
        float* outf = out + (size_t)frame * fout;
        float x[NK];
        uint32_t sbits = 0;                                       // bit BP::pos(i) = data bit 14 of value i
        if (KEEP) asm volatile("" : "+r"(kbits));                 // not loop-invariant to the compiler: it would hoist (and spill) the single-bit tests
        uint2 rv[CPT];
        if constexpr (ROT) {                                      // chunks 2m and 2m + 1 are neighbours: one 16-byte load
#pragma unroll
            for (int m = 0; m < CPT / 2; ++m) {
                const uint4 v = __ldcs(reinterpret_cast<const uint4*>(rawf + sgp[2 * m * GT + t]));
                rv[2 * m] = make_uint2(v.x, v.y); rv[2 * m + 1] = make_uint2(v.z, v.w);
            }
        } else {
#pragma unroll
            for (int j = 0; j < CPT; ++j) rv[j] = __ldcs(reinterpret_cast<const uint2*>(rawf + sgp[j * GT + t]));
        }
        if (frame + NG < f_end) {                                 // the group's next raw tile on its way into L2 meanwhile
#pragma unroll
            for (int j = 0; j < CPT; ++j) asm volatile("prefetch.global.L2 [%0];" :: "l"(rawf + (size_t)NG * c.npix + sgp[j * GT + t]));
        }
        uint32_t chain = 0;
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            const int ch = t + j * GT + (int)chain;               // (slot in the thread-major tables) chunk j's pedestals are fetched once chunk j - 1 is done
            const uint32_t w[2] = {rv[j].x, rv[j].y};
            const float4 p0 = sm4[0 * NCH + ch], p1 = sm4[1 * NCH + ch];
            const float pa[4] = {p0.x, p0.y, p0.z, p0.w}, pb[4] = {p1.x, p1.y, p1.z, p1.w};
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                // both bit-14 flags of the word with one shift: bit 14 -> 2j + k, bit 30 -> 16 + 2j + k
                sbits |= (w[k] >> (14 - 2 * j - k)) & (0x00010001u << (2 * j + k));
#if PSR_CMF_MAGIC
                const uint32_t wm = w[k] & 0x3FFF3FFFu;
                // float(v) without a conversion: 0x4B000000 | v is 2^23 + v; both subtractions are exact or correctly rounded
                const float fe = __fsub_rn(__uint_as_float(__byte_perm(wm, 0x4B000000u, 0x7610)), 8388608.0f);
                const float fo = __fsub_rn(__uint_as_float(__byte_perm(wm, 0x4B000000u, 0x7632)), 8388608.0f);
#else
                const float fe = (float)(w[k] & 0x3FFFu), fo = (float)((w[k] >> 16) & 0x3FFFu);
#endif
                x[j * 4 + 2 * k] = psr_sub_sel(fe, pa[2 * k], pb[2 * k], w[k] & 0x4000u);
                x[j * 4 + 2 * k + 1] = psr_sub_sel(fo, pa[2 * k + 1], pb[2 * k + 1], w[k] & 0x40000000u);
            }
#if PSR_CMF_CHAIN
            chain = __float_as_uint(x[j * 4 + 3]) & zero;
#endif
        }
        const uint32_t elig = (sbits & e1bits) | (~sbits & e0bits);

        // common_mode_2d: skip if n < npix_min, skip if |cm| > cormax.  An unapplied correction is cm = +0: x - 0 == x for every x.
        float lo, hi, cm = 0.f;
        if (!WORDS || (c.cm_mode & 4)) {
            if (block_median_linear<GT, NK, CAP, BP>(x, elig, c.npix_min, win, &hist[grp][0], &selsh[grp][0], &list[grp][0], &redbuf[grp][0], lo, hi, g)) {
                const float med = psr_mid(lo, hi);
                if (!(fabsf(med) > c.cormax)) cm = med;
            }
        }
        if constexpr (WORDS) {
            // psana's order: bank medians (4), then row medians (1), then column medians (2), each on what the previous left
            if (c.cm_mode & 4) {
#pragma unroll
                for (int i = 0; i < NK; ++i) x[i] = psr_sub_if(x[i], cm, elig & (1u << BP::pos(i)));
                cm = 0.f;
            }
            if (c.cm_mode & 1) {
                const SegRows<GT> sr{t};
                if constexpr (NG == 1) {                              // one group: room for the tile's keys in shared memory -> sort each row
                    uint32_t* tile = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(sgp + 2 * CPT * GT) + SEG_SMEM_BYTES);
                    const SegCols<GT> sc{t};
#pragma unroll
                    for (int i = 0; i < NK; ++i)
                        tile[sr.of(i) * 49 + sc.of(i)] = (elig & (1u << BP::pos(i))) ? psr_f2key(x[i]) : KEY_INELIGIBLE;
                    g.sync();
                    rows_sort_phase(tile, sb.cmv, c.npix_min, c.cormax, g);
                } else seg_medians<GT, NK, 176, 64, 2 * (2 / NG), 16, BP>(x, elig, sr, c.npix_min, c.cormax, winr, sb, g);
#pragma unroll
                for (int i = 0; i < NK; ++i) x[i] = psr_sub_if(x[i], sb.cmv[sr.of(i)], elig & (1u << BP::pos(i)));
                g.sync();                                             // cmv is read before the column phase rewrites it
            }
            if (c.cm_mode & 2) {
                const SegCols<GT> sc{t};
                seg_medians<GT, NK, 48, 256, 4 * (2 / NG), 64, BP>(x, elig, sc, c.npix_min, c.cormax, winc, sb, g);
#pragma unroll
                for (int i = 0; i < NK; ++i) x[i] = psr_sub_if(x[i], sb.cmv[sc.of(i)], elig & (1u << BP::pos(i)));
                g.sync();
            }
        }

        chain = 0;
        uint32_t eligq = elig, sbitsq = sbits;
#if PSR_CMF_OPAQUE
        asm volatile("" : "+r"(eligq), "+r"(sbitsq));             // single-bit tests are made here, not carried (spilled) from the load phase
#endif
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            const int ch = t + j * GT + (int)chain;               // one chunk's gain factors in registers at a time
            const float4 ga = sm4[2 * NCH + ch], gb = sm4[3 * NCH + ch];
            const float g0[4] = {ga.x, ga.y, ga.z, ga.w}, g1[4] = {gb.x, gb.y, gb.z, gb.w};
            float o[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int q = BP::pos(j * 4 + i);
                const float xv = psr_sub_if(x[j * 4 + i], cm, eligq & (1u << q));
                o[i] = psr_mul_sel(xv, g0[i], g1[i], sbitsq & (1u << q));
            }
            if (KEEP) {                                               // np.where(keep, v, 0): only the variant launched when a keep mask is installed
#pragma unroll
                for (int i = 0; i < 4; ++i) if (!(kbits & (1u << BP::pos(j * 4 + i)))) o[i] = 0.0f;
            }
#if PSR_CMF_CHAIN
            chain = __float_as_uint(o[3]) & zero;
#endif
            const int gp = sgp[j * GT + t];
            if constexpr (IMG) {
                if constexpr (ROT) {
                    int idx = sgp[(CPT + j) * GT + t];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { img_st_at(outf, idx, o[i]); idx += rdc; }
                } else if constexpr (OWN == 2 || OWN == 3) {
                    float* p = outf + sgp[(CPT + j) * GT + t];
                    const float q0 = OWN == 2 ? o[0] : o[3], q1 = OWN == 2 ? o[1] : o[2], q2 = OWN == 2 ? o[2] : o[1], q3 = OWN == 2 ? o[3] : o[0];
                    if (al16) __stcs(reinterpret_cast<float4*>(p), make_float4(q0, q1, q2, q3));
                    else { __stcs(reinterpret_cast<float2*>(p), make_float2(q0, q1)); __stcs(reinterpret_cast<float2*>(p + 2), make_float2(q2, q3)); }
                } else img_store_chunk<4>(outf, c.fwd + gp, (kinds >> (2 * j)) & 3u, sgp[(CPT + j) * GT + t], o);
            }
            else __stcs(reinterpret_cast<float4*>(outf + gp), make_float4(o[0], o[1], o[2], o[3]));
        }
    }
    };
    if constexpr (IMG && PSR_IMG_ROWOWN) {
        if (rdc != 0) { frames(std::integral_constant<int, 1>{}); return; }
    }
    if constexpr (IMG && PSR_IMG_RUNS) {
        if (runcls == 2) { frames(std::integral_constant<int, 2>{}); return; }
        if (runcls == 3) { frames(std::integral_constant<int, 3>{}); return; }
    }
    frames(std::integral_constant<int, 0>{});
}

constexpr int EPIX_NT = 352, JF_NT = 512;
#ifndef PSR_CMB_EPIX_NT          /* banks-only kernel: threads, min CTAs/SM, pixels per chunk */
#define PSR_CMB_EPIX_NT 352
#endif
#ifndef PSR_CMB_EPIX_KS        /* keys in shared memory: a third resident CTA per SM */
#define PSR_CMB_EPIX_KS 0      /* single frames: 256 CTAs do not fill the GPU, values in registers are faster (20.5 -> 16.4 us per frame with 4-pixel chunks) */
#endif
#ifndef PSR_CMB_EPIX_MINB
#define PSR_CMB_EPIX_MINB (PSR_CMB_EPIX_KS ? 3 : 2)
#endif
#ifndef PSR_CMB_EPIX_CW
#define PSR_CMB_EPIX_CW 4
#endif
constexpr size_t EPIX_SMEM = ((size_t)(176 * 49) * 5 + 15) / 16 * 16 + (EPIX_NT / 32) * 4 * 256 * 4;   // padded tile (float + flag byte) + warp histograms
constexpr size_t JF_SMEM = ((size_t)(256 * 65) * 5 + 15) / 16 * 16 + (JF_NT / 32) * 4 * 256 * 4;
#ifndef PSR_CM_GROUPS_FPC
#define PSR_CM_GROUPS_FPC 16   /* frames per CTA of the two-frame-group kernel (8 per group) */
#endif
constexpr size_t EPIX_GROUPS_SMEM = (size_t)2112 * (4 * 16) + (size_t)12 * 352 * 4;
constexpr size_t EPIX_WORDS_SMEM = EPIX_GROUPS_SMEM + (PSR_CM_WORDS_NG == 1 ? (size_t)SEG_SMEM_BYTES + 176 * 49 * 4 : 2 * (size_t)SEG_SMEM_BYTES);   // + the column histograms and the tile's keys for the row sort (one group), or the histograms of two groups   // all four constant tables of one bank tile + per-thread chunk offsets

}  // namespace

cudaError_t psr_common_mode_setup(void)
{
    cudaError_t e = cudaFuncSetAttribute(cm_calib_kernel<0, 352, 384, 8, EPIX_NT>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EPIX_SMEM);
    if (e != cudaSuccess) return e;
    if (PSR_CMB_JF_KS) {
        e = cudaFuncSetAttribute(cm_banks_kernel<1, 512, 1024, 16, PSR_CMB_JF_NT, PSR_CMB_JF_MINB, PSR_CMB_JF_CW, PSR_CMB_JF_KS != 0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 64 * 4);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(cm_banks_kernel<1, 512, 1024, 16, PSR_CMB_JF_NT, PSR_CMB_JF_MINB, PSR_CMB_JF_CW, PSR_CMB_JF_KS != 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 64 * 4);
        if (e != cudaSuccess) return e;
    }
    e = cudaFuncSetAttribute(cm_banks_epix_groups_kernel<false, false, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EPIX_GROUPS_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(cm_banks_epix_groups_kernel<false, true, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EPIX_GROUPS_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(cm_banks_epix_groups_kernel<true, false, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EPIX_GROUPS_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(cm_banks_epix_groups_kernel<false, false, true, PSR_CM_WORDS_NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EPIX_WORDS_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(cm_banks_epix_groups_kernel<false, true, true, PSR_CM_WORDS_NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EPIX_WORDS_SMEM);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(cm_calib_kernel<1, 512, 1024, 16, JF_NT>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JF_SMEM);
}

namespace {
template <bool IMG>
cudaError_t launch_banks(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_out, cudaStream_t st)
{
    if (c.det_kind == 0) {
        if (nframes > 1) {
            const int fpc = PSR_CM_GROUPS_FPC;
            dim3 g2((unsigned)c.nseg * 2 * 8, (unsigned)((nframes + fpc - 1) / fpc));
            // the calib-layout keep mask never applies to images (ref: producer.py:92-97 masks the 2-D image), and without an
            // installed mask every keep bit is set: only calib output with a mask runs the variant that tests them
            if (!IMG && c.has_keep && !c.ignore_keep) cm_banks_epix_groups_kernel<false, true, false, 2><<<g2, 704, EPIX_GROUPS_SMEM, st>>>(d_raw, d_out, c, nframes, fpc);
            else cm_banks_epix_groups_kernel<IMG, false, false, 2><<<g2, 704, EPIX_GROUPS_SMEM, st>>>(d_raw, d_out, c, nframes, fpc);
        } else {
            const unsigned grid = (unsigned)nframes * c.nseg * 2 * 8;
            cm_banks_kernel<0, 352, 384, 8, PSR_CMB_EPIX_NT, PSR_CMB_EPIX_MINB, PSR_CMB_EPIX_CW, PSR_CMB_EPIX_KS != 0, IMG>
                <<<grid, PSR_CMB_EPIX_NT, PSR_CMB_EPIX_KS ? 176 * 48 * 4 : 0, st>>>(d_raw, d_out, c);
        }
    } else {
        const unsigned grid = (unsigned)nframes * c.nseg * 2 * 16;
        cm_banks_kernel<1, 512, 1024, 16, PSR_CMB_JF_NT, PSR_CMB_JF_MINB, PSR_CMB_JF_CW, PSR_CMB_JF_KS != 0, IMG>
            <<<grid, PSR_CMB_JF_NT, PSR_CMB_JF_KS ? 256 * 64 * 4 : 0, st>>>(d_raw, d_out, c);
    }
    return cudaGetLastError();
}
}  // namespace

cudaError_t psr_launch_common_mode(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_out, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    if (c.cm_mode == 4) return launch_banks<false>(c, d_raw, nframes, d_out, st);
    if (c.det_kind == 0 && nframes > 1 && PSR_CM_WORDS_GROUPS) {
        // any word with row / column medians, epix10ka batches: the two-frame-group kernel with its row and column phases
        const int fpc = PSR_CM_GROUPS_FPC;
        dim3 g2((unsigned)c.nseg * 2 * 8, (unsigned)((nframes + fpc - 1) / fpc));
        // the tile's keys for the row sort only when the word has row medians: what shared memory a CTA leaves is L1
        const size_t smem = (PSR_CM_WORDS_NG == 1 && !(c.cm_mode & 1)) ? EPIX_GROUPS_SMEM + (size_t)SEG_SMEM_BYTES : EPIX_WORDS_SMEM;
        if (c.has_keep && !c.ignore_keep) cm_banks_epix_groups_kernel<false, true, true, PSR_CM_WORDS_NG><<<g2, 704, smem, st>>>(d_raw, d_out, c, nframes, fpc);
        else cm_banks_epix_groups_kernel<false, false, true, PSR_CM_WORDS_NG><<<g2, 704, smem, st>>>(d_raw, d_out, c, nframes, fpc);
        return cudaGetLastError();
    }
    if (c.det_kind == 0) {
        const unsigned grid = (unsigned)nframes * c.nseg * 2 * 8;
        cm_calib_kernel<0, 352, 384, 8, EPIX_NT><<<grid, EPIX_NT, EPIX_SMEM, st>>>(d_raw, d_out, c);
    } else {
        const unsigned grid = (unsigned)nframes * c.nseg * 2 * 16;
        cm_calib_kernel<1, 512, 1024, 16, JF_NT><<<grid, JF_NT, JF_SMEM, st>>>(d_raw, d_out, c);
    }
    return cudaGetLastError();
}

cudaError_t psr_launch_common_mode_image(const PsrDev& c, const uint16_t* d_raw, int nframes, float* d_img, cudaStream_t st)
{
    if (nframes <= 0) return cudaSuccess;
    if (c.cm_mode != 4 || c.fwd == nullptr) return cudaErrorInvalidValue;
    return launch_banks<true>(c, d_raw, nframes, d_img, st);
}
