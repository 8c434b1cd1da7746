// K1 -- fused per-candidate importance score (row-wise log-softmax + NLL, reduction='none').
// Replaces F.cross_entropy at reference pytorch_collab.py:102 / util.py:129.
//
// The logits are read from HBM exactly once; everything else happens in registers / shared
// memory.  Three layouts, chosen by row size (DESIGN.md "K1"):
//   small rows (<= 256 B, e.g. 10 classes)   : tiles of rows staged by 1-D bulk (TMA) copies into
//                                              a shared-memory ring, one thread per row;
//   medium rows (<= 8 KB, e.g. 1000 x bf16)  : one warp per row, 128-bit streaming loads, the row
//                                              lives in registers, warp-shuffle max / sum;
//   wide rows (e.g. 32768 x bf16 = 64 KB)    : one CTA per row, the row streams through a
//                                              bulk-copy ring in 16 KB chunks, online softmax.
// A scalar warp-per-row kernel covers everything unaligned or strided.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "internal.cuh"
#include "score_small.cuh"

namespace mercury {

namespace {


// ------------------------------------------------------------------------------------------------
// generic: one warp per row, scalar loads, two passes (second pass served by L1/L2)
template <typename T>
__global__ void __launch_bounds__(256) score_generic_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                            int64_t n, int64_t c, int64_t row_stride, int64_t ignore_index,
                                                            float* __restrict__ scores, int32_t* status) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < n; row += nwarps) {
        const T* rp = logits + row * row_stride;
        float m = -INFINITY;
        for (int64_t j = lane; j < c; j += 32) m = fmaxf(m, Elem<T>::ld(rp + j));
        m = warp_max(m);
        float l = 0.f;
        for (int64_t j = lane; j < c; j += 32) l += ex2_approx((Elem<T>::ld(rp + j) - m) * kLog2e);
        l = warp_sum(l);
        if (lane == 0) {
            const int64_t y = labels[row];
            float out = 0.f;
            if (label_ok(y, c, ignore_index, status)) out = logf(l) - (Elem<T>::ld(rp + y) - m);
            scores[row] = out;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// 2^t on the FMA pipe (t <= ~0): Cody-Waite split by the 1.5*2^23 trick, degree-5 polynomial on [-0.5, 0.5] (max relative
// error 1.9e-7, the same as MUFU.EX2's), exponent added as an integer.  With one MUFU.EX2 per logit the special-function
// unit is the co-limiter of the bf16 1000-class shape (ncu: XU 68 % beside DRAM 75 %); POLY of every 8 logits go this way.
__device__ __forceinline__ float ex2_poly(float t) {
    t = fmaxf(t, -125.f);
    const float z = __fadd_rn(t, 12582912.f);
    const float f = __fadd_rn(t, -__fadd_rn(z, -12582912.f));
    float g = fmaf(f, 0.0013264815788716078f, 0.009671512059867382f);
    g = fmaf(f, g, 0.05550733581185341f);
    g = fmaf(f, g, 0.24022242426872253f);
    g = fmaf(f, g, 0.6931470036506653f);
    return __int_as_float(__float_as_int(fmaf(f, g, 1.f)) + (__float_as_int(z) << 23));
}

// medium rows: one warp per row, the row held in registers as VPL 128-bit vectors per lane
constexpr int kDefaultPoly = -1;
// Two logits per instruction where the ISA has it (sm_100: fma.rn.f32x2 / add.f32x2 on register pairs): the bf16 1000-class
// shape is limited by instruction issue beside DRAM (profiles/r2_k1_poly_ab.jsonl), so the scale-and-shift and the
// accumulation of a bf16x2 word are one instruction each instead of two.
__device__ __forceinline__ unsigned long long pack2(float a, float b) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void term2_acc(uint32_t w, unsigned long long kk, unsigned long long nh, unsigned long long& acc) {
    const unsigned long long xx = pack2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
    unsigned long long t;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t) : "l"(xx), "l"(kk), "l"(nh));
    float t0, t1;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(t0), "=f"(t1) : "l"(t));
    const unsigned long long ee = pack2(ex2_approx(t0), ex2_approx(t1));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(acc) : "l"(acc), "l"(ee));
}
template <typename T, int VPL, int POLY = 0>
__global__ void __launch_bounds__(256) score_warp_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                         int64_t n, int64_t c, int64_t row_stride, int64_t ignore_index,
                                                         float* __restrict__ scores, int32_t* status) {
    const int lane = threadIdx.x & 31;
    const int nvec = (int)((c * (int64_t)sizeof(T)) >> 4);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    // (a software pipeline over rows -- next row's loads in flight while this one is reduced -- was measured
    // SLOWER at 1.28M x 1000 bf16: 489 us against 448 us; 32 resident warps x 2 KB already cover the latency)
    constexpr bool kPrefetch = false;
    auto load_row = [&](int64_t row, uint4 (&dst)[VPL]) {
        const uint4* rp4 = reinterpret_cast<const uint4*>(logits + row * row_stride);
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int k = lane + 32 * i;
            dst[i] = k < nvec ? ldg_stream(rp4 + k)
                              : make_uint4(Elem<T>::kNegInfWord, Elem<T>::kNegInfWord, Elem<T>::kNegInfWord, Elem<T>::kNegInfWord);
        }
    };
    const int64_t row0 = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    uint4 v[VPL], nxt[kPrefetch ? VPL : 1];
    if (kPrefetch && row0 < n) load_row(row0, v);
    for (int64_t row = row0; row < n; row += nwarps) {
        const T* rp = logits + row * row_stride;
        if constexpr (kPrefetch) {
            if (row + nwarps < n) load_row(row + nwarps, nxt);
        } else {
            load_row(row, v);
        }
        const int64_t y = labels[row];
        const bool ok = label_in_range(y, c, ignore_index);
        if (!ok && lane == 0) label_ok(y, c, ignore_index, status);
        const float xy = ok ? Elem<T>::ld(rp + y) : 0.f;

        float m;
        if constexpr (sizeof(T) == 4) {
            m = -INFINITY;
#pragma unroll
            for (int i = 0; i < VPL; ++i) {
                m = fmaxf(m, fmaxf(fmaxf(__uint_as_float(v[i].x), __uint_as_float(v[i].y)),
                                   fmaxf(__uint_as_float(v[i].z), __uint_as_float(v[i].w))));
            }
        } else {
            uint32_t mx = Elem<T>::kNegInfWord;
#pragma unroll
            for (int i = 0; i < VPL; ++i) mx = bf16x2_max(bf16x2_max(mx, bf16x2_max(v[i].x, v[i].y)), bf16x2_max(v[i].z, v[i].w));
            m = fmaxf(bf16_lo(mx), bf16_hi(mx));
        }
        m = warp_max(m);
        const ExpScale es(m);
        float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;  // independent partial sums (ILP)
        if constexpr (sizeof(T) == 2 && POLY < 0) {   // packed pairs
            const unsigned long long kk = pack2(kLog2e, kLog2e), nh = pack2(-es.hi, -es.hi);
            unsigned long long a0 = 0ull, a1 = 0ull, a2 = 0ull, a3 = 0ull;
#pragma unroll
            for (int i = 0; i < VPL; ++i) {
                term2_acc(v[i].x, kk, nh, a0); term2_acc(v[i].y, kk, nh, a1);
                term2_acc(v[i].z, kk, nh, a2); term2_acc(v[i].w, kk, nh, a3);
            }
            asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a0) : "l"(a1));
            asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a2) : "l"(a3));
            asm("mov.b64 {%0, %1}, %2;" : "=f"(l0), "=f"(l1) : "l"(a0));
            asm("mov.b64 {%0, %1}, %2;" : "=f"(l2), "=f"(l3) : "l"(a2));
        } else
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            if constexpr (sizeof(T) == 4) {
                l0 += es.term(__uint_as_float(v[i].x));
                l1 += es.term(__uint_as_float(v[i].y));
                l2 += es.term(__uint_as_float(v[i].z));
                l3 += es.term(__uint_as_float(v[i].w));
            } else {
                auto term = [&](float x, int slot) { return slot < POLY ? ex2_poly(fmaf(x, kLog2e, -es.hi)) : es.term(x); };
                l0 += term(bf16_lo(v[i].x), 0); l1 += term(bf16_hi(v[i].x), 4);
                l2 += term(bf16_lo(v[i].y), 1); l3 += term(bf16_hi(v[i].y), 5);
                l0 += term(bf16_lo(v[i].z), 2); l1 += term(bf16_hi(v[i].z), 6);
                l2 += term(bf16_lo(v[i].w), 3); l3 += term(bf16_hi(v[i].w), 7);
            }
        }
        const float l = warp_sum((l0 + l1) + (l2 + l3));
        if (lane == 0) scores[row] = ok ? (logf(l) - es.log_fix()) - (xy - m) : 0.f;
        if constexpr (kPrefetch) {
#pragma unroll
            for (int i = 0; i < VPL; ++i) v[i] = nxt[i];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// small rows: the body lives in score_small.cuh (shared with the fused one-kernel step)
template <typename T, int CT, int STAGES, int CONSUMERS>
__global__ void __launch_bounds__(CONSUMERS + kProducerThreads, 3)
score_small_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels, int64_t n, int c_rt, int tile_rows,
                   int64_t ignore_index, float* __restrict__ scores, int32_t* status, double* __restrict__ tile_sums, int64_t ntiles) {
    extern __shared__ __align__(128) unsigned char smem[];
    // (ntiles comes from the host: a 64-bit division by every thread in front of the first load costs the kernel a microsecond)
    score_small_body<T, CT, STAGES, CONSUMERS, false, true>(smem, logits, labels, n, c_rt, tile_rows, ignore_index, scores, status,
                                                            tile_sums, blockIdx.x, ntiles, gridDim.x, nullptr);
}


// wide rows: one CTA per row, CHUNK-byte pieces through the ring, per-thread online softmax
template <typename T, int STAGES, int CHUNK, int CONSUMERS>
__global__ void __launch_bounds__(CONSUMERS + kProducerThreads)
score_wide_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels, int64_t n, int64_t c, int64_t row_stride,
                  int64_t ignore_index, float* __restrict__ scores, int32_t* status) {
    extern __shared__ __align__(128) unsigned char smem[];
    Ring<STAGES>* ring = reinterpret_cast<Ring<STAGES>*>(smem);
    float* red = reinterpret_cast<float*>(smem + 128);  // [2][CONSUMERS/32] max, sum
    unsigned char* stage0 = smem + 256;
    ring_init(ring, CONSUMERS / 32);
    constexpr int VPT = CHUNK / 16 / CONSUMERS;  // vectors per thread per chunk
    static_assert(VPT >= 1 && VPT * 16 * CONSUMERS == CHUNK, "chunk/consumer geometry");
    constexpr int NW = CONSUMERS / 32;

    const uint32_t row_bytes = (uint32_t)(c * sizeof(T));
    const int chunks = (int)((row_bytes + CHUNK - 1) / CHUNK);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
                const unsigned char* src = reinterpret_cast<const unsigned char*>(logits + row * row_stride);
                for (int k = 0; k < chunks; ++k, ++it) {
                    const int s = it % STAGES, round = it / STAGES;
                    if (round > 0) mbar_wait(&ring->empty[s], (round - 1) & 1);
                    const uint32_t bytes = min((uint32_t)CHUNK, row_bytes - (uint32_t)k * CHUNK);
                    mbar_arrive_expect_tx(&ring->full[s], bytes);
                    bulk_g2s(stage0 + (size_t)s * CHUNK, src + (size_t)k * CHUNK, bytes, &ring->full[s]);
                }
            }
        }
    } else {
        const int t = threadIdx.x - kProducerThreads;
        const int cw = t >> 5;
        int it = 0;
        for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
            const int64_t y = labels[row];
            const bool ok = (t == 0) ? label_ok(y, c, ignore_index, status) : false;
            const float xy = ok ? Elem<T>::ld(logits + row * row_stride + y) : 0.f;
            float m = -INFINITY, l = 0.f;
            for (int k = 0; k < chunks; ++k, ++it) {
                const int s = it % STAGES, round = it / STAGES;
                mbar_wait(&ring->full[s], round & 1);
                const uint32_t bytes = min((uint32_t)CHUNK, row_bytes - (uint32_t)k * CHUNK);
                const int nv = (int)(bytes >> 4);
                const uint4* sv = reinterpret_cast<const uint4*>(stage0 + (size_t)s * CHUNK);
                uint4 v[VPT];
#pragma unroll
                for (int i = 0; i < VPT; ++i) {
                    const int q = t + i * CONSUMERS;
                    v[i] = q < nv ? sv[q]
                                  : make_uint4(Elem<T>::kNegInfWord, Elem<T>::kNegInfWord, Elem<T>::kNegInfWord, Elem<T>::kNegInfWord);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&ring->empty[s]);  // the chunk now lives in registers
                float cm;
                if constexpr (sizeof(T) == 4) {
                    cm = -INFINITY;
#pragma unroll
                    for (int i = 0; i < VPT; ++i)
                        cm = fmaxf(cm, fmaxf(fmaxf(__uint_as_float(v[i].x), __uint_as_float(v[i].y)),
                                             fmaxf(__uint_as_float(v[i].z), __uint_as_float(v[i].w))));
                } else {
                    uint32_t mx = Elem<T>::kNegInfWord;
#pragma unroll
                    for (int i = 0; i < VPT; ++i) mx = bf16x2_max(bf16x2_max(mx, bf16x2_max(v[i].x, v[i].y)), bf16x2_max(v[i].z, v[i].w));
                    cm = fmaxf(bf16_lo(mx), bf16_hi(mx));
                }
                if (cm > -INFINITY) {
                    const float mn = fmaxf(m, cm);
                    const ExpScale es(mn);
                    float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
#pragma unroll
                    for (int i = 0; i < VPT; ++i) {
                        if constexpr (sizeof(T) == 4) {
                            p0 += es.term(__uint_as_float(v[i].x));
                            p1 += es.term(__uint_as_float(v[i].y));
                            p2 += es.term(__uint_as_float(v[i].z));
                            p3 += es.term(__uint_as_float(v[i].w));
                        } else {
                            p0 += es.term(bf16_lo(v[i].x)); p1 += es.term(bf16_hi(v[i].x));
                            p2 += es.term(bf16_lo(v[i].y)); p3 += es.term(bf16_hi(v[i].y));
                            p0 += es.term(bf16_lo(v[i].z)); p1 += es.term(bf16_hi(v[i].z));
                            p2 += es.term(bf16_lo(v[i].w)); p3 += es.term(bf16_hi(v[i].w));
                        }
                    }
                    const float ps = (p0 + p1) + (p2 + p3);
                    // rescale the running sum to the new max (m == -inf -> factor 0, l is 0 anyway) and
                    // take the common 2^lo factor out of this chunk's terms (|lo*ln2| ~ 1e-6: first order)
                    l = fmaf(l, ex2_approx((m - mn) * kLog2e), fmaf(ps, -es.log_fix(), ps));
                    m = mn;
                }
            }
            // combine the CONSUMERS partial (m, l) pairs: consumers-only named barrier (id 1)
            const float wm = warp_max(m);
            if (lane == 0) red[cw] = wm;
            asm volatile("bar.sync 1, %0;" ::"n"(CONSUMERS) : "memory");
            float M = red[0];
#pragma unroll
            for (int i = 1; i < NW; ++i) M = fmaxf(M, red[i]);
            const float part = (m > -INFINITY) ? l * ex2_approx((m - M) * kLog2e) : 0.f;
            const float ws = warp_sum(part);
            if (lane == 0) red[NW + cw] = ws;
            asm volatile("bar.sync 1, %0;" ::"n"(CONSUMERS) : "memory");
            if (t == 0) {
                float L = 0.f;
#pragma unroll
                for (int i = 0; i < NW; ++i) L += red[NW + i];
                scores[row] = ok ? logf(L) - (xy - M) : 0.f;
            }
            // red[] is rewritten only after the next row's first barrier, which every consumer
            // reaches after reading it here
            asm volatile("bar.sync 1, %0;" ::"n"(CONSUMERS) : "memory");
        }
    }
}

template <typename T>
int launch_score(const T* logits, const int64_t* labels, int64_t n, int64_t c, int64_t row_stride, int64_t ignore_index,
                 float* scores, int32_t* status, cudaStream_t st, double* part_sums, int* per_out, bool chained) {
    if (per_out) *per_out = 0;
    const int64_t row_bytes = c * (int64_t)sizeof(T);
    const bool base16 = (reinterpret_cast<uintptr_t>(logits) & 15) == 0;
    const bool stride16 = ((row_stride * (int64_t)sizeof(T)) & 15) == 0;
    const bool lab16 = (reinterpret_cast<uintptr_t>(labels) & 15) == 0;

    if (row_stride == c && base16 && lab16 && row_bytes <= 256) {
        constexpr int STAGES = 3, CONSUMERS = 256;
        // one row per consumer thread and tile (~12 KB per stage at 10 classes): three CTAs (27 warps) share an
        // SM and cover each other's stalls; measured best of {128..1024} rows x {1..8} CTAs/SM at 1M x 10
        // developer knobs, read once and validated: rows per tile must divide the CDF chunk (the fused statistics
        // sum whole tiles per chunk) and keep the ring inside shared memory
        static const int knob_tile = env_knob("MERCURY_K1_TILE", 0, 32, 2048);
        static const int knob_ctas = env_knob("MERCURY_K1_CTAS", 3, 1, 8);
        int tile_rows = small_tile_rows(row_bytes);
        if (knob_tile > 0 && (knob_tile & (knob_tile - 1)) == 0 && kCdfChunk % knob_tile == 0 &&
            small_smem_bytes(STAGES, knob_tile, row_bytes) <= 200 * 1024)
            tile_rows = knob_tile;
        const int ctas_per_sm = knob_ctas;
        const int64_t ntiles = (n + tile_rows - 1) / tile_rows;
        // every CTA gets the same number of tiles (the last ones one fewer at most): size the grid for that
        const int64_t slots = (int64_t)ctas_per_sm * kSMs;
        const int64_t per_cta = (ntiles + slots - 1) / slots;
        const int grid = (int)min_i64(ntiles, (ntiles + per_cta - 1) / per_cta);
        const size_t smem = small_smem_bytes(STAGES, tile_rows, row_bytes);
        auto go = [&](auto kern) -> int {
            int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), smem, "score_small smem attribute");
            if (rc) return rc;
            cudaError_t e = launch_chain(chained, kern, dim3(grid), dim3(CONSUMERS + kProducerThreads), smem, st, logits, labels, n, (int)c, tile_rows,
                             ignore_index, scores, status, part_sums, ntiles);
            if (e != cudaSuccess) return cuda_fail(e, "score_small_kernel launch");
            if (per_out && part_sums) *per_out = tile_rows;
            return MERCURY_OK;
        };
        if (c == 10) return go(score_small_kernel<T, 10, STAGES, CONSUMERS>);
        return go(score_small_kernel<T, 0, STAGES, CONSUMERS>);
    }
    if (base16 && stride16 && (row_bytes & 15) == 0 && row_bytes <= 8192) {
        const int nvec = (int)(row_bytes >> 4);
        const int64_t warps_needed = n;
        // Grid: 96 CTAs per SM, far more than the 6 that are resident.  With 8 per SM (1.33 waves of resident CTAs, long
        // grid-stride loops) the last third of the rows ran at a third of the occupancy: 0.445 ms at 1.28 M x 1000 bf16; short
        // CTAs that the hardware hands out as slots free up give 0.378-0.390 ms (profiles/r2_k1_grid_sweep.jsonl).
        static const int ctas_per_sm = env_knob("MERCURY_K1_WARP_CTAS", 96, 1, 4096);
        const int grid = (int)min_i64((warps_needed + 7) / 8, (int64_t)kSMs * ctas_per_sm);
#define MB_LAUNCH_WARP(VPL)                                                                                         \
    score_warp_kernel<T, VPL><<<grid, 256, 0, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status)
        static const int poly = env_knob("MERCURY_K1_POLY", kDefaultPoly, -1, 3);   // -1: packed f32x2 arithmetic   // exponentials per 8 logits on the FMA pipe
        if (nvec <= 32) MB_LAUNCH_WARP(1);
        else if (nvec <= 64) MB_LAUNCH_WARP(2);
        else if (nvec <= 128 && sizeof(T) == 2 && poly == -1) score_warp_kernel<T, 4, -1><<<grid, 256, 0, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status);
        else if (nvec <= 128 && sizeof(T) == 2 && poly == 1) score_warp_kernel<T, 4, 1><<<grid, 256, 0, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status);
        else if (nvec <= 128 && sizeof(T) == 2 && poly == 2) score_warp_kernel<T, 4, 2><<<grid, 256, 0, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status);
        else if (nvec <= 128 && sizeof(T) == 2 && poly == 3) score_warp_kernel<T, 4, 3><<<grid, 256, 0, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status);
        else if (nvec <= 128) MB_LAUNCH_WARP(4);
        else if (nvec <= 256) MB_LAUNCH_WARP(8);
        else MB_LAUNCH_WARP(16);
#undef MB_LAUNCH_WARP
        MB_CHECK_LAUNCH("score_warp_kernel");
        return MERCURY_OK;
    }
    if (base16 && stride16 && (row_bytes & 15) == 0) {
        constexpr int STAGES = 4, CHUNK = 16384, CONSUMERS = 256;
        const size_t smem = 256 + (size_t)STAGES * CHUNK;
        auto kern = score_wide_kernel<T, STAGES, CHUNK, CONSUMERS>;
        const int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), smem, "score_wide smem attribute");
        if (rc) return rc;
        static const int wide_ctas = env_knob("MERCURY_K1_WIDE_CTAS", 64, 1, 4096);
        const int grid = (int)min_i64(n, (int64_t)kSMs * wide_ctas);
        kern<<<grid, CONSUMERS + kProducerThreads, smem, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status);
        MB_CHECK_LAUNCH("score_wide_kernel");
        return MERCURY_OK;
    }
    const int grid = (int)min_i64((n + 7) / 8, (int64_t)kSMs * 8);
    score_generic_kernel<T><<<grid, 256, 0, st>>>(logits, labels, n, c, row_stride, ignore_index, scores, status);
    MB_CHECK_LAUNCH("score_generic_kernel");
    return MERCURY_OK;
}

}  // namespace

int score_dispatch(const void* logits, int dtype, const int64_t* labels, int64_t n, int64_t c, int64_t row_stride,
                   int64_t ignore_index, float* scores, int32_t* status, cudaStream_t st, double* part_sums, int* per_out,
                   bool chained) {
    if (per_out) *per_out = 0;
    if (n == 0) return MERCURY_OK;
    if (dtype == MERCURY_F32)
        return launch_score(static_cast<const float*>(logits), labels, n, c, row_stride, ignore_index, scores, status, st,
                            part_sums, per_out, chained);
    return launch_score(static_cast<const __nv_bfloat16*>(logits), labels, n, c, row_stride, ignore_index, scores, status, st,
                        part_sums, per_out, chained);
}

}  // namespace mercury

extern "C" int mercury_score(const void* logits, int dtype, const int64_t* labels, int64_t n, int64_t c, int64_t row_stride,
                             int64_t ignore_index, float* scores, int32_t* status, void* stream) {
    using namespace mercury;
    MB_CHECK_ARG(n >= 0 && c >= 1, "mercury_score: need n >= 0 and c >= 1 (n=%lld c=%lld)", (long long)n, (long long)c);
    MB_CHECK_ARG(dtype == MERCURY_F32 || dtype == MERCURY_BF16, "mercury_score: dtype must be MERCURY_F32 or MERCURY_BF16");
    MB_CHECK_ARG(row_stride >= c, "mercury_score: row_stride (%lld) < c (%lld)", (long long)row_stride, (long long)c);
    MB_CHECK_ARG(n == 0 || (logits && labels && scores && status), "mercury_score: null pointer");
    // (plain stream order.  Launching it with the programmatic-serialization attribute was measured: back-to-back launches got
    // SLOWER, 10.2 -> 11.7 us at 1 M x 10 -- the early CTAs of the next launch sit on shared memory while the previous one drains)
    return score_dispatch(logits, dtype, labels, n, c, row_stride, ignore_index, scores, status, static_cast<cudaStream_t>(stream));
}
