// K1, small-row layout (rows <= 256 B): the device code shared by the stand-alone score kernel (score.cu) and
// the fused one-kernel step (step_fused.cuh).  See score.cu for the description of the layouts.
#pragma once
#include <math.h>

#include "common.cuh"

namespace mercury {
namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr int kProducerThreads = 32;
constexpr int kSmallHeader = 512;  // ring barriers + per-warp partial sums in front of the stages

template <typename T> struct Elem;
template <> struct Elem<float> {
    static __device__ __forceinline__ float ld(const float* p) { return *p; }
    static constexpr uint32_t kNegInfWord = 0xff800000u;
};
template <> struct Elem<__nv_bfloat16> {
    static __device__ __forceinline__ float ld(const __nv_bfloat16* p) {
        return __uint_as_float(static_cast<uint32_t>(*reinterpret_cast<const unsigned short*>(p)) << 16);
    }
    static constexpr uint32_t kNegInfWord = 0xff80ff80u;
};

constexpr float kLn2 = 0.6931471805599453f;

// packed bf16x2 max (one ALU op per two logits): NaN-free operand wins, like fmaxf
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// exp(x - m) summed as 2^(x*k - hi) with hi = fl(m*k): one FFMA + one MUFU per logit.  Every
// term is off by the same factor 2^lo, lo = m*k - hi (exact, from an FMA); `log_fix` removes it
// from log(sum) afterwards, so nothing is lost against computing (x - m) first.
struct ExpScale {
    float hi, lo;
    __device__ __forceinline__ explicit ExpScale(float m) { hi = m * kLog2e; lo = fmaf(m, kLog2e, -hi); }
    __device__ __forceinline__ float term(float x) const { return ex2_approx(fmaf(x, kLog2e, -hi)); }
    __device__ __forceinline__ float log_fix() const { return lo * kLn2; }  // log(sum_true) = log(sum) - log_fix
};

// label handling shared by all layouts: returns true if the row contributes a loss.
// Call it from ONE thread per row (it raises the status bit with an atomic).
__device__ __forceinline__ bool label_ok(int64_t y, int64_t c, int64_t ignore_index, int32_t* status) {
    if (y == ignore_index) return false;
    if (y < 0 || y >= c) {
        atomicOr(status, MERCURY_ST_BAD_LABEL);
        return false;
    }
    return true;
}
__device__ __forceinline__ bool label_in_range(int64_t y, int64_t c, int64_t ignore_index) {
    return y != ignore_index && y >= 0 && y < c;
}

// ------------------------------------------------------------------------------------------------
// bulk-copy ring shared by the small-row and wide-row kernels
template <int STAGES> struct Ring {
    uint64_t full[STAGES];
    uint64_t empty[STAGES];
};

template <int STAGES>
__device__ __forceinline__ void ring_init(Ring<STAGES>* ring, int consumer_warps) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&ring->full[s], 1);
            mbar_init(&ring->empty[s], consumer_warps);
        }
        mbar_fence_init();
    }
    __syncthreads();
}

// small rows: tile = TR rows of logits + their labels per stage; consumers own rows.
// CT > 0: compile-time class count (rows cached in registers); CT == 0: run-time class count.
// PRODUCER_LAST: the producer warp is the CTA's last warp and the consumers are threads 0..CONSUMERS-1 (the fused step
// wants its worker threads numbered from zero); otherwise warp 0 produces and the consumers follow it.
// PDL: first kernel of a programmatic-dependent-launch chain (the stand-alone kernel).
// This CTA scores tiles tile_begin, tile_begin + tile_step, ... < tile_end.  cta_sum_out (optional): the fp64 sum of this
// CTA's tile sums, added in the order the tiles were taken (needs tile_sums).
// Every thread of the CTA must call this; it returns when the CTA's tiles are done (no barrier at the end).
template <typename T, int CT, int STAGES, int CONSUMERS, bool PRODUCER_LAST, bool PDL>
__device__ __forceinline__ void score_small_body(unsigned char* smem, const T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                 int64_t n, int c_rt, int tile_rows, int64_t ignore_index, float* __restrict__ scores,
                                                 int32_t* status, double* __restrict__ tile_sums, int64_t tile_begin, int64_t tile_end,
                                                 int64_t tile_step, double* __restrict__ cta_sum_out,
                                                 const unsigned int* range_smem = nullptr) {
    const int c = CT > 0 ? CT : c_rt;
    Ring<STAGES>* ring = reinterpret_cast<Ring<STAGES>*>(smem);
    const uint32_t logit_bytes = (uint32_t)tile_rows * c * sizeof(T);
    const uint32_t logit_slot = (logit_bytes + 127u) & ~127u;
    const uint32_t stage_bytes = logit_slot + (uint32_t)tile_rows * 8u;
    double* tred = reinterpret_cast<double*>(smem + 128);  // [2][CONSUMERS / 32] per-warp score sums
    unsigned char* stage0 = smem + kSmallHeader;
    static_assert(sizeof(Ring<STAGES>) <= 128 && 128 + 2 * (CONSUMERS / 32) * 8 <= kSmallHeader, "header layout");
    ring_init(ring, CONSUMERS / 32);
    if (range_smem) { tile_begin = range_smem[0]; tile_end = range_smem[1]; }   // (written by one thread in front of ring_init's barrier)
    if constexpr (PDL) {
        pdl_wait();                // (the previous step's draw kernel may still be reading the scores we overwrite)
        pdl_launch_dependents();
    }

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool producer = PRODUCER_LAST ? warp == CONSUMERS / 32 : warp == 0;

    if (producer) {
        // ---- producer warp
        int it = 0;
        for (int64_t tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
            const int s = it % STAGES, round = it / STAGES;
            if (round > 0) mbar_wait(&ring->empty[s], (round - 1) & 1);
            const int64_t row0 = tile * tile_rows;
            const int rows = (int)min((int64_t)tile_rows, n - row0);
            unsigned char* dst = stage0 + (size_t)s * stage_bytes;
            const uint32_t lb = (uint32_t)rows * c * sizeof(T), yb = (uint32_t)rows * 8u;
            if (((lb | yb) & 15u) == 0) {
                if (lane == 0) {
                    mbar_arrive_expect_tx(&ring->full[s], lb + yb);
                    bulk_g2s(dst, logits + row0 * c, lb, &ring->full[s]);
                    bulk_g2s(dst + logit_slot, labels + row0, yb, &ring->full[s]);
                }
            } else {  // ragged last tile: plain loads by the producer warp
                const T* src = logits + row0 * c;
                T* d = reinterpret_cast<T*>(dst);
                for (int k = lane; k < rows * c; k += 32) d[k] = src[k];
                int64_t* dy = reinterpret_cast<int64_t*>(dst + logit_slot);
                for (int k = lane; k < rows; k += 32) dy[k] = labels[row0 + k];
                __syncwarp();
                if (lane == 0) mbar_arrive(&ring->full[s]);
            }
            __syncwarp();
        }
    } else {
        // ---- consumers
        const int t = PRODUCER_LAST ? (int)threadIdx.x : (int)threadIdx.x - kProducerThreads;
        double cta_acc = 0.0;
        int it = 0;
        for (int64_t tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
            const int s = it % STAGES, round = it / STAGES;
            mbar_wait(&ring->full[s], round & 1);
            const int64_t row0 = tile * tile_rows;
            const int rows = (int)min((int64_t)tile_rows, n - row0);
            const unsigned char* base = stage0 + (size_t)s * stage_bytes;
            const T* tl = reinterpret_cast<const T*>(base);
            const int64_t* ty = reinterpret_cast<const int64_t*>(base + logit_slot);
            double tile_acc = 0.0;
            for (int r = t; r < rows; r += CONSUMERS) {
                const T* rp = tl + (size_t)r * c;
                const int64_t y = ty[r];
                float m = -INFINITY, l = 0.f, fix = 0.f;
                if constexpr (CT > 0) {
                    float x[CT > 0 ? CT : 1];
                    if constexpr (sizeof(T) == 4 && (CT % 2) == 0) {
#pragma unroll
                        for (int j = 0; j < CT / 2; ++j) {
                            const float2 p2 = reinterpret_cast<const float2*>(rp)[j];
                            x[2 * j] = p2.x;
                            x[2 * j + 1] = p2.y;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < CT; ++j) x[j] = Elem<T>::ld(rp + j);
                    }
#pragma unroll
                    for (int j = 0; j < CT; ++j) m = fmaxf(m, x[j]);
                    const ExpScale es(m);
                    float la = 0.f, lb = 0.f;
#pragma unroll
                    for (int j = 0; j + 1 < CT; j += 2) { la += es.term(x[j]); lb += es.term(x[j + 1]); }
                    if (CT & 1) la += es.term(x[CT - 1]);
                    l = la + lb;
                    fix = es.log_fix();
                } else {
                    // rotate the start column by the row so that lanes hit different banks
                    int j0 = r % c;
                    int j = j0;
                    for (int k = 0; k < c; ++k) {
                        m = fmaxf(m, Elem<T>::ld(rp + j));
                        if (++j == c) j = 0;
                    }
                    j = j0;
                    for (int k = 0; k < c; ++k) {
                        l += ex2_approx((Elem<T>::ld(rp + j) - m) * kLog2e);
                        if (++j == c) j = 0;
                    }
                }
                float out = 0.f;
                if (label_ok(y, c, ignore_index, status)) out = (logf(l) - fix) - (Elem<T>::ld(rp + y) - m);
                scores[row0 + r] = out;
                tile_acc += (double)out;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&ring->empty[s]);
            if (tile_sums) {  // sum of this tile's scores, fixed order: feeds the pool statistics
                tile_acc = warp_sum(tile_acc);
                double* buf = tred + (it & 1) * (CONSUMERS / 32);
                if (lane == 0) buf[(t >> 5)] = tile_acc;
                asm volatile("bar.sync 2, %0;" ::"n"(CONSUMERS) : "memory");
                if (t == 0) {
                    double sum = 0.0;
#pragma unroll
                    for (int i = 0; i < CONSUMERS / 32; ++i) sum += buf[i];
                    tile_sums[tile] = sum;
                    cta_acc += sum;
                }
            }
        }
        if (cta_sum_out && t == 0) *cta_sum_out = cta_acc;
    }
}

// tile geometry shared by both users
__host__ __device__ inline int small_tile_rows(int64_t row_bytes) { return row_bytes <= 88 ? 256 : row_bytes <= 184 ? 128 : 64; }
constexpr int kSmallMinTileRows = 64;   // (the whole-step workspace holds one sum per tile: cdf.cu carve() sizes it by this)
__host__ __device__ inline size_t small_smem_bytes(int stages, int tile_rows, int64_t row_bytes) {
    const size_t logit_slot = ((size_t)tile_rows * row_bytes + 127) & ~(size_t)127;
    return kSmallHeader + (size_t)stages * (logit_slot + (size_t)tile_rows * 8);
}

}  // namespace
}  // namespace mercury
