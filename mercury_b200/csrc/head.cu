// Classifier head fused into the scoring kernel: scores straight from the
// penultimate features, logits never written to memory.
//
//   reference: out = self.linear(out.view(N, -1))            pytorch_model.py:94-96 (ResNet: 512 -> C)
//              F.cross_entropy(out, label, reduction='none') pytorch_collab.py:102
//
// Shape of the work at the reference's sizes (D = 512, C = 10): 5120 multiply-adds per candidate for
// 2 KB (fp32) or 1 KB (bf16) of features read once -- HBM-bound, not a tensor-core job (an MMA tile
// needs N >= 8..16 columns of which 10 would be real, and the stream of features is the bill either
// way).  So: plain FFMA, features streamed with coalesced 128-bit loads, the weight matrix resident in
// shared memory for the life of the CTA.
//
// One warp scores kRows candidates at a time.  Lane l owns the feature columns {128 j + 4 l .. +3}: a
// warp-wide load instruction reads 512 contiguous bytes of one row.  Per 128-column slab a lane loads
// one float4 per row, and for every class one float4 of weights from shared memory (consecutive lanes,
// consecutive 16-byte words: conflict-free) that is used for all kRows rows.  The kRows x C partial
// dot products are then summed across the warp and every lane finishes the cross-entropy redundantly.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/mercury_b200.h"
#include "common.cuh"
#include "internal.cuh"

namespace mercury {
namespace {

constexpr int kHeadNT = 256;   // 8 warps
constexpr int kRows = 4;       // candidates per warp iteration

typedef unsigned long long u64;
struct X4 { u64 lo, hi; };  // four consecutive fp32 values as two packed pairs (x, y) and (z, w)

// d += a * b on both halves (SASS FFMA2: two fp32 multiply-adds per issue slot)
__device__ __forceinline__ void fma2(u64& d, u64 a, u64 b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b)); }

template <typename T>
__device__ __forceinline__ X4 load4(const T* row, int col);
template <>
__device__ __forceinline__ X4 load4<float>(const float* row, int col) {
    const uint4 v = ldg_stream(row + col);
    return X4{(u64)v.x | ((u64)v.y << 32), (u64)v.z | ((u64)v.w << 32)};
}
template <>
__device__ __forceinline__ X4 load4<uint16_t>(const uint16_t* row, int col) {  // bf16 -> fp32 is a shift
    const uint2 v = *reinterpret_cast<const uint2*>(row + col);
    return X4{(u64)(v.x << 16) | ((u64)(v.x & 0xffff0000u) << 32), (u64)(v.y << 16) | ((u64)(v.y & 0xffff0000u) << 32)};
}

constexpr int kSlots = 16;  // class slots per row in the cross-warp reduction (c <= 16)

// Sums the kRows x CP packed partial dot products across the warp and finishes the cross-entropy of the group's rows.
// Reduce-scatter: 64 slots (row r, class k at r * 16 + k) halved five times -- after the step with offset o a lane keeps
// the half selected by its bit o -- so that lane l ends with slots 2l and 2l + 1, i.e. row l >> 3, classes 2 (l & 7) and
// the next (62 shuffles instead of 5 per accumulator); 8 lanes per row then do the log-sum-exp.
template <int CP>
__device__ __forceinline__ void head_finish(const u64 (&acc2)[kRows][CP], int lane, int c, float bias0, float bias1, int64_t y,
                                            int64_t my_row, int64_t n, int64_t ignore_index, float* __restrict__ scores, int& bad) {
    float v[kRows * kSlots];
#pragma unroll
    for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int k = 0; k < kSlots; ++k)
            v[r * kSlots + k] = k < CP ? __uint_as_float((unsigned)acc2[r][k < CP ? k : 0]) + __uint_as_float((unsigned)(acc2[r][k < CP ? k : 0] >> 32)) : 0.f;
#pragma unroll
    for (int half = (kRows * kSlots) / 2, o = 16; o >= 1; half >>= 1, o >>= 1) {
        const bool upper = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float send = upper ? v[i] : v[i + half], keep = upper ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    // F.cross_entropy(reduction='none'): log sum exp(x - m) - (x_y - m)
    const int k0 = 2 * (lane & 7);
    const float a0 = v[0] + bias0, a1 = v[1] + bias1;
    const bool ok0 = k0 < c, ok1 = k0 + 1 < c;
    float m = fmaxf(ok0 ? a0 : -INFINITY, ok1 ? a1 : -INFINITY);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float sum = (ok0 ? expf(a0 - m) : 0.f) + (ok1 ? expf(a1 - m) : 0.f);
    float xy = ((int64_t)k0 == y ? a0 : 0.f) + ((int64_t)(k0 + 1) == y ? a1 : 0.f);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        xy += __shfl_xor_sync(0xffffffffu, xy, o);
    }
    float loss = logf(sum) - (xy - m);
    if (y == ignore_index) loss = 0.f;
    else if (y < 0 || y >= c) { loss = 0.f; if (my_row < n) bad = MERCURY_ST_BAD_LABEL; }
    if ((lane & 7) == 0 && my_row < n) scores[my_row] = loss;
}

// CP: classes padded to a compile-time count (accumulators live in registers); c <= CP are real.
// SLABS: d / 128 when known at compile time, 0 = run-time loop.  With SLABS known a group's 4 x SLABS feature vectors live
// in registers and the NEXT group's are requested slab by slab as soon as the current group's slab has been consumed, so
// that the memory system stays busy through the arithmetic, the reduction and the cross-entropy (same registers, no
// extra buffer).
template <typename T, int CP, int SLABS>
__global__ void __launch_bounds__(kHeadNT) score_head_kernel(const T* __restrict__ feats, int64_t n, int d, int64_t row_stride,
                                                             const float* __restrict__ weight, const float* __restrict__ bias, int c,
                                                             const int64_t* __restrict__ labels, int64_t ignore_index,
                                                             float* __restrict__ scores, int32_t* __restrict__ status) {
    extern __shared__ __align__(16) float w_s[];  // [CP][d], rows >= c are zero
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < CP * d; i += kHeadNT) w_s[i] = i < c * d ? weight[i] : 0.f;
    __syncthreads();
    // after the reduction lane l holds row (l >> 3)'s logits for classes k0 = 2 (l & 7) and k0 + 1
    const int k0 = 2 * (lane & 7), my_r = lane >> 3;
    const float bias0 = (bias && k0 < c) ? bias[k0] : 0.f, bias1 = (bias && k0 + 1 < c) ? bias[k0 + 1] : 0.f;
    const int64_t groups = (n + kRows - 1) / kRows;
    const int64_t gstride = (int64_t)gridDim.x * (kHeadNT / 32);
    const int col0 = lane << 2;
    int bad = 0;
    int64_t g = (int64_t)blockIdx.x * (kHeadNT / 32) + warp;
    if constexpr (SLABS > 0) {
        if (g >= groups) return;
        X4 x[SLABS][kRows];
        const T* rows[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) rows[r] = feats + (g * kRows + r < n ? g * kRows + r : n - 1) * row_stride;  // (tail rows re-read the last one)
#pragma unroll
        for (int j = 0; j < SLABS; ++j)
#pragma unroll
            for (int r = 0; r < kRows; ++r) x[j][r] = load4<T>(rows[r], (j << 7) + col0);
        int64_t y = g * kRows + my_r < n ? labels[g * kRows + my_r] : 0;
        for (; g < groups; g += gstride) {
            const int64_t gn = g + gstride < groups ? g + gstride : g;   // (the last group is simply requested again)
#pragma unroll
            for (int r = 0; r < kRows; ++r) rows[r] = feats + (gn * kRows + r < n ? gn * kRows + r : n - 1) * row_stride;
            const int64_t y_next = gn * kRows + my_r < n ? labels[gn * kRows + my_r] : 0;
            u64 acc2[kRows][CP];
#pragma unroll
            for (int r = 0; r < kRows; ++r)
#pragma unroll
                for (int k = 0; k < CP; ++k) acc2[r][k] = 0ull;
#pragma unroll
            for (int j = 0; j < SLABS; ++j) {
#pragma unroll
                for (int k = 0; k < CP; ++k) {
                    const ulonglong2 w = *reinterpret_cast<const ulonglong2*>(w_s + k * d + (j << 7) + col0);
#pragma unroll
                    for (int r = 0; r < kRows; ++r) { fma2(acc2[r][k], x[j][r].lo, w.x); fma2(acc2[r][k], x[j][r].hi, w.y); }
                }
#pragma unroll
                for (int r = 0; r < kRows; ++r) x[j][r] = load4<T>(rows[r], (j << 7) + col0);   // next group's slab j
            }
            head_finish<CP>(acc2, lane, c, bias0, bias1, y, g * kRows + my_r, n, ignore_index, scores, bad);
            y = y_next;
        }
    } else {
        const int slabs = d >> 7;
        for (; g < groups; g += gstride) {
            const int64_t r0 = g * kRows;
            u64 acc2[kRows][CP];
#pragma unroll
            for (int r = 0; r < kRows; ++r)
#pragma unroll
                for (int k = 0; k < CP; ++k) acc2[r][k] = 0ull;
            const T* rows[kRows];
#pragma unroll
            for (int r = 0; r < kRows; ++r) rows[r] = feats + (r0 + r < n ? r0 + r : n - 1) * row_stride;
            const int64_t my_row = r0 + my_r;
            const int64_t y = my_row < n ? labels[my_row] : 0;   // (early: used after the reduction)
#pragma unroll 2
            for (int j = 0; j < slabs; ++j) {
                const int col = (j << 7) + col0;
                X4 x[kRows];
#pragma unroll
                for (int r = 0; r < kRows; ++r) x[r] = load4<T>(rows[r], col);
#pragma unroll
                for (int k = 0; k < CP; ++k) {
                    const ulonglong2 w = *reinterpret_cast<const ulonglong2*>(w_s + k * d + col);
#pragma unroll
                    for (int r = 0; r < kRows; ++r) { fma2(acc2[r][k], x[r].lo, w.x); fma2(acc2[r][k], x[r].hi, w.y); }
                }
            }
            head_finish<CP>(acc2, lane, c, bias0, bias1, y, my_row, n, ignore_index, scores, bad);
        }
    }
    if (bad) atomicOr(status, bad);
}

template <typename T>
int launch_head(const T* feats, int64_t n, int d, int64_t row_stride, const float* weight, const float* bias, int c,
                const int64_t* labels, int64_t ignore_index, float* scores, int32_t* status, cudaStream_t st) {
    const int cp = c <= 2 ? 2 : c <= 4 ? 4 : c <= 8 ? 8 : c <= 10 ? 10 : 16;
    const size_t smem = (size_t)cp * d * sizeof(float);
    const int64_t groups = (n + kRows - 1) / kRows;
    const int grid = (int)min_i64((groups + kHeadNT / 32 - 1) / (kHeadNT / 32), (int64_t)148 * 2);
    cudaError_t e = cudaSuccess;
#define MB_HEAD_LAUNCH(CPV, SL)                                                                                                 \
    if (ensure_dynamic_smem(reinterpret_cast<const void*>(score_head_kernel<T, CPV, SL>), smem, "score_head smem attribute"))   \
        return MERCURY_ERR_CUDA;                                                                                                \
    if (e == cudaSuccess)                                                                                                       \
        score_head_kernel<T, CPV, SL><<<grid, kHeadNT, smem, st>>>(feats, n, d, row_stride, weight, bias, c, labels,           \
                                                                   ignore_index, scores, status);
#define MB_HEAD_CASE(CPV)                                                                                                       \
    case CPV:                                                                                                                   \
        if (d == 512) { MB_HEAD_LAUNCH(CPV, 4) } else { MB_HEAD_LAUNCH(CPV, 0) }                                                \
        break;
    switch (cp) {
        MB_HEAD_CASE(2)
        MB_HEAD_CASE(4)
        MB_HEAD_CASE(8)
        MB_HEAD_CASE(10)
        MB_HEAD_CASE(16)
    }
#undef MB_HEAD_LAUNCH
#undef MB_HEAD_CASE
    if (e != cudaSuccess) return cuda_fail(e, "score_head_kernel attribute");
    MB_CHECK_LAUNCH("score_head_kernel");
    return MERCURY_OK;
}

}  // namespace
}  // namespace mercury

using namespace mercury;

extern "C" int mercury_score_head(const void* features, int dtype, int64_t n, int64_t d, int64_t row_stride, const float* weight,
                                  const float* bias, int64_t c, const int64_t* labels, int64_t ignore_index, float* scores,
                                  int32_t* status, void* stream) {
    MB_CHECK_ARG(n >= 0 && d >= 128 && d % 128 == 0 && row_stride >= d, "mercury_score_head: need n >= 0, d a multiple of 128, row_stride >= d");
    MB_CHECK_ARG(c >= 1 && c <= 16, "mercury_score_head: 1..16 classes (wider heads: score the logits with mercury_score)");
    {
        const int64_t cp = c <= 2 ? 2 : c <= 4 ? 4 : c <= 8 ? 8 : c <= 10 ? 10 : 16;   // (class count as padded by launch_head)
        MB_CHECK_ARG(c * d <= 51200 && cp * d * 4 <= 224 * 1024,
                     "mercury_score_head: the weight matrix must fit shared memory (c * d <= 51200, padded classes * d * 4 <= 224 KiB)");
    }
    MB_CHECK_ARG(dtype == MERCURY_F32 || dtype == MERCURY_BF16, "mercury_score_head: bad dtype");
    if (n == 0) return MERCURY_OK;
    MB_CHECK_ARG(features && weight && labels && scores && status, "mercury_score_head: null pointer");
    const size_t esz = dtype == MERCURY_F32 ? 4 : 2;
    MB_CHECK_ARG((reinterpret_cast<uintptr_t>(features) & 15) == 0 && (row_stride * esz) % 16 == 0,
                 "mercury_score_head: feature rows must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == MERCURY_F32)
        return launch_head(static_cast<const float*>(features), n, (int)d, row_stride, weight, bias, (int)c, labels, ignore_index, scores, status, st);
    return launch_head(static_cast<const uint16_t*>(features), n, (int)d, row_stride, weight, bias, (int)c, labels, ignore_index, scores, status, st);
}
