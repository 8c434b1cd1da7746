// K2 -- pool statistics, the cumulative distribution, and the inverse-CDF draw.
//
// Replaces torch.mean / EMAverage.update / `+ alpha*ema` / torch.sum (reference
// pytorch_collab.py:109-112) and torch.multinomial(p, B, replacement=True) (:114).
//
// The exact mode reproduces ATen's serial fp32 `sum += val` bit for bit, in parallel
// (chain.cuh explains the arithmetic).  Work is cut into chunks of kChunk candidates:
//   1. chunk sums (fp64) -> an approximate running total per chunk, used only to GUESS the
//      floating-point phase (binade) each chunk will be summed in;
//   2. `summarize`: every chunk, in parallel, reduces its addends to the 2-state map they apply
//      to the running sum, for the guessed phase and the one above it;
//   3. `spine` (one CTA): walks the chunk maps in order with block-wide scans, verifying every
//      guess; a chunk whose guess was wrong, or in which the running sum changes phase, is
//      "resolved" exactly (phase by phase, inside the CTA).  Output: the exact running sum at
//      every chunk boundary;
//   4. `materialize` (or the fused draw kernel) rebuilds any chunk's exact prefix from its
//      boundary value.
// A wrong guess costs time, never correctness.
#include <math.h>

#include <mutex>

#include "chain.cuh"
#include "common.cuh"
#include "internal.cuh"

namespace mercury {

namespace {

constexpr int kChunk = kCdfChunk;        // candidates per chunk
constexpr int kMaxBounds = 64;

__device__ __forceinline__ uint32_t f2b(float f) { return __float_as_uint(f); }
__device__ __forceinline__ float b2f(uint32_t b) { return __uint_as_float(b); }

struct Bounds { int64_t v[kMaxBounds]; int n; };

// ---------------------------------------------------------------------------------------------
// element access: either the weights themselves, or p_i = (score_i + shift) / total on the fly
struct WeightSrc {
    const float* w;      // weights, or scores when stats != nullptr
    const float* stats;  // nullptr -> raw weights
    __device__ __forceinline__ void prepare(float& shift, float& total) const {
        if (stats) { shift = stats[MERCURY_STAT_SHIFT]; total = stats[MERCURY_STAT_TOTAL]; }
        else { shift = 0.f; total = 1.f; }
    }
    __device__ __forceinline__ float get(int64_t i, float shift, float total) const {
        const float x = w[i];
        return stats ? __fdiv_rn(__fadd_rn(x, shift), total) : x;
    }
};

// ---------------------------------------------------------------------------------------------
// block-wide helpers
template <int NT>
__device__ __forceinline__ double block_sum(double v, double* red) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < NT / 32; ++i) t += red[i];  // fixed order: deterministic
    return t;
}

__device__ __forceinline__ Pair shfl_up_pair(Pair p, int off) {
    Pair r;
    r.d0 = __shfl_up_sync(0xffffffffu, p.d0, off);
    r.d1 = __shfl_up_sync(0xffffffffu, p.d1, off);
    return r;
}

// exclusive scan of the maps held one per thread, in thread order; also returns the block total.
// `tot` is shared scratch of NT/32 Pairs.  Contains two __syncthreads().
template <int NT>
__device__ __forceinline__ Pair block_exclusive_scan(Pair mine, Pair* tot, Pair* block_total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Pair inc = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        Pair o = shfl_up_pair(inc, off);
        if (lane >= off) inc = compose(o, inc);
    }
    __syncthreads();  // protects tot[] against the previous use
    if (lane == 31) tot[warp] = inc;
    __syncthreads();
    Pair ex = shfl_up_pair(inc, 1);
    if (lane == 0) ex = identity_pair();
    Pair pre = identity_pair();
    Pair all = identity_pair();
#pragma unroll
    for (int i = 0; i < NT / 32; ++i) {
        if (i == warp) pre = all;
        all = compose(all, tot[i]);
    }
    if (block_total) *block_total = all;
    return compose(pre, ex);
}

// ---------------------------------------------------------------------------------------------
// exact resolution of one chunk from its exact start state, by a whole CTA.
// pb[k] = bits of element tid*IPT+k (blocked layout); len = valid elements in the chunk.
// If out_norm != nullptr, writes the normalised cdf (chain / total) of every element.
struct ResolveSmem {
    Pair tot[32];
    int cross;
    uint32_t next_bits;
    uint32_t end_bits;
};

template <int NT, int IPT>
__device__ uint32_t resolve_chunk(const uint32_t (&pb)[IPT], int len, uint32_t bits, float* out_norm, float total,
                                  ResolveSmem* sm) {
    const int base = threadIdx.x * IPT;
    int pos = 0;
    while (true) {
        const uint32_t E = phase_of(bits), S = bits - (E << 23);
        Pair loc = identity_pair();
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const int i = base + k;
            if (i >= pos && i < len) loc = compose(loc, addend_pair(pb[k], E));
        }
        if (threadIdx.x == 0) sm->cross = INT_MAX;
        const Pair ex = block_exclusive_scan<NT>(loc, sm->tot, nullptr);  // (its barriers publish cross)
        uint32_t s = apply(ex, S);
        int my_cross = INT_MAX;
        uint32_t s_before = 0;
        int k_cross = 0;
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const int i = base + k;
            if (i >= pos && i < len && my_cross == INT_MAX) {
                const Pair d = addend_pair(pb[k], E);
                const uint32_t sn = s + ((s & 1u) ? d.d1 : d.d0);
                if (sn >= kPhaseLimit) {
                    my_cross = i; s_before = s; k_cross = k;
                } else {
                    if (out_norm) out_norm[i] = __fdiv_rn(b2f(bits_of(E, sn)), total);
                    s = sn;
                }
            }
        }
        if (my_cross != INT_MAX) atomicMin(&sm->cross, my_cross);
        // the thread owning the last valid element publishes the end state (valid if no crossing)
        if (len - 1 >= base && len - 1 < base + IPT && my_cross == INT_MAX) sm->end_bits = bits_of(E, s);
        __syncthreads();
        const int cross = sm->cross;
        if (cross == INT_MAX) {
            const uint32_t r = sm->end_bits;
            __syncthreads();
            return r;
        }
        if (my_cross == cross) {  // the one real fp32 add of this phase
            uint32_t pk = 0;
#pragma unroll
            for (int k = 0; k < IPT; ++k) if (k == k_cross) pk = pb[k];
            const uint32_t nb = f2b(__fadd_rn(b2f(bits_of(E, s_before)), b2f(pk)));
            sm->next_bits = nb;
            if (out_norm) out_norm[cross] = __fdiv_rn(b2f(nb), total);
        }
        __syncthreads();
        bits = sm->next_bits;
        pos = cross + 1;
        __syncthreads();
        if (pos >= len) return bits;
    }
}

template <int NT, int IPT>
__device__ __forceinline__ void load_chunk(const WeightSrc& src, int64_t start, int len, uint32_t (&pb)[IPT]) {
    float shift, total;
    src.prepare(shift, total);
    const int base = threadIdx.x * IPT;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
        const int i = base + k;
        pb[k] = i < len ? f2b(src.get(start + i, shift, total)) : 0u;
    }
}

// ---------------------------------------------------------------------------------------------
// 1. chunk sums (fp64) + ATen's validity checks (every element is visited exactly once here)
__global__ void __launch_bounds__(256) chunk_sums_kernel(const float* __restrict__ w, int64_t n, double* __restrict__ sums,
                                                         int32_t* status, int check) {
    __shared__ double red[8];
    const int64_t start = (int64_t)blockIdx.x * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    double acc = 0.0;
    int bad = 0;
    for (int i = threadIdx.x; i < len; i += 256) {
        const float x = w[start + i];
        if (check) {
            if (!(x >= 0.f)) bad |= MERCURY_ST_NEGATIVE;
            else if (isinf(x)) bad |= MERCURY_ST_NONFINITE;
        }
        acc += (double)x;
    }
    if (bad) atomicOr(status, bad);
    const double t = block_sum<256>(acc, red);
    if (threadIdx.x == 0) sums[blockIdx.x] = t;
}

// 2a. statistics of a pool of scores (single CTA): mean(s), EMA, shift, total, approximate
//     running total of p at every chunk boundary.
__global__ void __launch_bounds__(1024) stats_finalize_kernel(const float* __restrict__ scores, int64_t n,
                                                              const double* __restrict__ sums, int nchunks, Bounds bounds,
                                                              float* ema_state, int ema_first, double ema_alpha,
                                                              double smooth_alpha, float* __restrict__ stats,
                                                              double* __restrict__ approx) {
    __shared__ double red[32];
    __shared__ double carry;
    float ema = *ema_state;
    float mean = 0.f;
    double pool_sum = 0.0;
    const float a = (float)ema_alpha, one_minus_a = (float)(1.0 - ema_alpha);
    // prefix means at the requested pool sizes (one EMA update each, reference :95-110)
    for (int j = 0; j < bounds.n; ++j) {
        const int64_t b = bounds.v[j];
        const int64_t full = b / kChunk;
        double acc = 0.0;
        for (int64_t c = threadIdx.x; c < full; c += 1024) acc += sums[c];
        for (int64_t i = full * kChunk + threadIdx.x; i < b; i += 1024) acc += (double)scores[i];
        pool_sum = block_sum<1024>(acc, red);
        mean = (float)(pool_sum / (double)b);
        if (ema_first && j == 0) ema = mean;
        else ema = __fadd_rn(__fmul_rn(a, ema), __fmul_rn(one_minus_a, mean));
    }
    const float shift = __fmul_rn((float)smooth_alpha, ema);
    const double total_d = pool_sum + (double)n * (double)shift;
    const float total = (float)total_d;
    if (threadIdx.x == 0) {
        stats[MERCURY_STAT_MEAN] = mean;
        stats[MERCURY_STAT_EMA] = ema;
        stats[MERCURY_STAT_SHIFT] = shift;
        stats[MERCURY_STAT_TOTAL] = total;
        stats[4] = stats[5] = stats[6] = stats[7] = 0.f;
        *ema_state = ema;
        carry = 0.0;
    }
    __syncthreads();
    // exclusive scan of the per-chunk totals of p ~ (sum_scores + len*shift) / total
    for (int c0 = 0; c0 < nchunks; c0 += 1024) {
        const int c = c0 + threadIdx.x;
        double v = 0.0;
        if (c < nchunks) {
            const int len = (int)min((int64_t)kChunk, n - (int64_t)c * kChunk);
            v = (sums[c] + (double)len * (double)shift) / (double)total;
        }
        // inclusive warp scan, then across warps
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        double inc = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const double o = __shfl_up_sync(0xffffffffu, inc, off);
            if (lane >= off) inc += o;
        }
        __syncthreads();
        if (lane == 31) red[warp] = inc;
        __syncthreads();
        double pre = carry;
        for (int i = 0; i < warp; ++i) pre += red[i];
        if (c < nchunks) approx[c] = pre + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = pre + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) approx[nchunks] = carry;
}

// 2b. the same scan for raw weights
__global__ void __launch_bounds__(1024) prefix_finalize_kernel(const double* __restrict__ sums, int nchunks,
                                                               double* __restrict__ approx) {
    __shared__ double red[32];
    __shared__ double carry;
    if (threadIdx.x == 0) carry = 0.0;
    __syncthreads();
    for (int c0 = 0; c0 < nchunks; c0 += 1024) {
        const int c = c0 + threadIdx.x;
        const double v = c < nchunks ? sums[c] : 0.0;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        double inc = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const double o = __shfl_up_sync(0xffffffffu, inc, off);
            if (lane >= off) inc += o;
        }
        __syncthreads();
        if (lane == 31) red[warp] = inc;
        __syncthreads();
        double pre = carry;
        for (int i = 0; i < warp; ++i) pre += red[i];
        if (c < nchunks) approx[c] = pre + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = pre + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) approx[nchunks] = carry;
}

__global__ void __launch_bounds__(256) normalize_kernel(const float* __restrict__ scores, int64_t n,
                                                        const float* __restrict__ stats, float* __restrict__ p) {
    const float shift = stats[MERCURY_STAT_SHIFT], total = stats[MERCURY_STAT_TOTAL];
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
        p[i] = __fdiv_rn(__fadd_rn(scores[i], shift), total);
}

// ---------------------------------------------------------------------------------------------
// 3. per-chunk maps for the guessed phase and the next one
struct __align__(32) Summary { uint32_t e; Pair q[2]; uint32_t pad[3]; };

constexpr int kSumNT = 256, kSumIPT = kChunk / kSumNT;

__global__ void __launch_bounds__(kSumNT) summarize_kernel(WeightSrc src, int64_t n, const double* __restrict__ approx,
                                                           Summary* __restrict__ out, int32_t* status, int check) {
    __shared__ Pair wt[2][kSumNT / 32];
    const int64_t start = (int64_t)blockIdx.x * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    uint32_t pb[kSumIPT];
    load_chunk<kSumNT, kSumIPT>(src, start, len, pb);
    // guess low: the walk tolerates "one phase too low", not "one too high"
    const uint32_t e = phase_of(f2b((float)(approx[blockIdx.x] * (1.0 - 1e-4))));
    Pair a0 = identity_pair(), a1 = identity_pair();
    int bad = 0;
#pragma unroll
    for (int k = 0; k < kSumIPT; ++k) {
        if (check) {
            const float x = b2f(pb[k]);
            if (!(x >= 0.f)) bad |= MERCURY_ST_NEGATIVE;
            else if (isinf(x)) bad |= MERCURY_ST_NONFINITE;
        }
        a0 = compose(a0, addend_pair(pb[k], e));
        a1 = compose(a1, addend_pair(pb[k], e + 1u));
    }
    if (bad) atomicOr(status, bad);
    // ordered reduction: lane 0 ends with the composition of the whole warp
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        Pair o0, o1;
        o0.d0 = __shfl_down_sync(0xffffffffu, a0.d0, off); o0.d1 = __shfl_down_sync(0xffffffffu, a0.d1, off);
        o1.d0 = __shfl_down_sync(0xffffffffu, a1.d0, off); o1.d1 = __shfl_down_sync(0xffffffffu, a1.d1, off);
        if (lane + off < 32) { a0 = compose(a0, o0); a1 = compose(a1, o1); }
    }
    if (lane == 0) { wt[0][warp] = a0; wt[1][warp] = a1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        Pair t0 = identity_pair(), t1 = identity_pair();
#pragma unroll
        for (int i = 0; i < kSumNT / 32; ++i) { t0 = compose(t0, wt[0][i]); t1 = compose(t1, wt[1][i]); }
        Summary s;
        s.e = e; s.q[0] = t0; s.q[1] = t1; s.pad[0] = s.pad[1] = s.pad[2] = 0;
        out[blockIdx.x] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// 4. the spine: exact running sum at every chunk boundary
constexpr int kSpineNT = 512, kSpineIPT = kChunk / kSpineNT;

__global__ void __launch_bounds__(kSpineNT) spine_kernel(WeightSrc src, int64_t n, int nchunks,
                                                         const Summary* __restrict__ sums, uint32_t* __restrict__ start_bits,
                                                         float* __restrict__ total_out, int32_t* status) {
    __shared__ ResolveSmem rs;
    __shared__ int first_bad;
    uint32_t bits = 0u;  // the running sum starts at +0
    int c = 0;
    while (c < nchunks) {
        const uint32_t E = phase_of(bits), S = bits - (E << 23);
        const int W = min(nchunks - c, kSpineNT);
        Pair mine = identity_pair();
        if ((int)threadIdx.x < W) {
            const Summary s = sums[c + threadIdx.x];
            if (E == s.e) mine = s.q[0];
            else if (E == s.e + 1u) mine = s.q[1];
            else mine = Pair{kSatCap, kSatCap};  // wrong guess: force a resolve
        }
        if (threadIdx.x == 0) first_bad = INT_MAX;
        Pair all;
        const Pair ex = block_exclusive_scan<kSpineNT>(mine, rs.tot, &all);
        const uint32_t s_in = apply(ex, S);
        const uint32_t s_out = s_in + ((s_in & 1u) ? mine.d1 : mine.d0);
        const bool in_win = (int)threadIdx.x < W;
        if (in_win && s_in < kPhaseLimit && s_out >= kPhaseLimit) atomicMin(&first_bad, (int)threadIdx.x);
        __syncthreads();
        const int f = first_bad;
        if (in_win && (int)threadIdx.x <= min(f, W - 1)) start_bits[c + threadIdx.x] = bits_of(E, s_in);
        if (f == INT_MAX) {
            bits = bits_of(E, apply(all, S));
            c += W;
            __syncthreads();
            continue;
        }
        // chunk c+f leaves the phase (or was guessed wrong): resolve it exactly
        if ((int)threadIdx.x == f) rs.next_bits = bits_of(E, s_in);
        __syncthreads();
        const uint32_t start = rs.next_bits;
        __syncthreads();
        const int64_t e0 = (int64_t)(c + f) * kChunk;
        const int len = (int)min((int64_t)kChunk, n - e0);
        uint32_t pb[kSpineIPT];
        load_chunk<kSpineNT, kSpineIPT>(src, e0, len, pb);
        bits = resolve_chunk<kSpineNT, kSpineIPT>(pb, len, start, nullptr, 1.f, &rs);
        c += f + 1;
    }
    if (threadIdx.x == 0) {
        start_bits[nchunks] = bits;
        const float total = b2f(bits);
        *total_out = total;
        if (!(total > 0.f)) atomicOr(status, MERCURY_ST_ZERO_SUM);
        else if (isinf(total)) atomicOr(status, MERCURY_ST_NONFINITE);
    }
}

// 5. rebuild every chunk's prefix and write the normalised CDF
constexpr int kMatNT = 256, kMatIPT = kChunk / kMatNT;

__global__ void __launch_bounds__(kMatNT) materialize_kernel(WeightSrc src, int64_t n, const uint32_t* __restrict__ start_bits,
                                                             const float* __restrict__ total_ptr, float* __restrict__ cdf) {
    __shared__ ResolveSmem rs;
    const int64_t start = (int64_t)blockIdx.x * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    uint32_t pb[kMatIPT];
    load_chunk<kMatNT, kMatIPT>(src, start, len, pb);
    resolve_chunk<kMatNT, kMatIPT>(pb, len, start_bits[blockIdx.x], cdf + start, *total_ptr, &rs);
}

// ---------------------------------------------------------------------------------------------
// fp64 mode: deterministic reduce-then-scan
__global__ void __launch_bounds__(256) materialize_f64_kernel(const float* __restrict__ w, int64_t n,
                                                              const double* __restrict__ approx, int nchunks,
                                                              float* __restrict__ cdf, float* __restrict__ total_out,
                                                              int32_t* status) {
    __shared__ double red[8];
    const int64_t start = (int64_t)blockIdx.x * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    const double total = approx[nchunks];
    constexpr int IPT = kChunk / 256;
    const int base = threadIdx.x * IPT;
    double v[IPT];
    double loc = 0.0;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
        v[k] = base + k < len ? (double)w[start + base + k] : 0.0;
        loc += v[k];
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double inc = loc;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const double o = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += o;
    }
    if (lane == 31) red[warp] = inc;
    __syncthreads();
    double run = approx[blockIdx.x] + inc - loc;
    for (int i = 0; i < warp; ++i) run += red[i];
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
        run += v[k];
        if (base + k < len) cdf[start + base + k] = (float)(run / total);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *total_out = (float)total;
        if (!(total > 0.0)) atomicOr(status, MERCURY_ST_ZERO_SUM);
        else if (isinf((float)total)) atomicOr(status, MERCURY_ST_NONFINITE);
    }
}

// ---------------------------------------------------------------------------------------------
// inverse-CDF search on a materialised CDF: one warp per draw, 32-way narrowing
__global__ void __launch_bounds__(256) search_kernel(const float* __restrict__ cdf, int64_t n, const double* __restrict__ u,
                                                     int64_t b, int64_t* __restrict__ idx) {
    const int lane = threadIdx.x & 31;
    const int64_t j = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (j >= b) return;
    const double uj = u[j];
    int64_t lo = 0, hi = n;  // answer in [lo, hi]
    while (hi - lo > 32) {
        const int64_t step = (hi - lo + 31) / 32;
        int64_t pos = lo + (int64_t)(lane + 1) * step;
        if (pos > hi) pos = hi;
        pos -= 1;  // last slot of this lane's segment
        const bool less = pos >= lo + (int64_t)lane * step ? ((double)cdf[pos] < uj) : false;
        // lanes whose segment is empty (beyond hi) report "not less"
        const unsigned m = __ballot_sync(0xffffffffu, less);
        const int cnt = __popc(m);  // monotone: the first cnt segments lie entirely below u
        const int64_t nlo = lo + (int64_t)cnt * step;
        if (nlo >= hi) { lo = hi; break; }
        int64_t nhi = nlo + step;
        if (nhi > hi) nhi = hi;
        lo = nlo;
        hi = nhi - 1;  // that segment's last slot is >= u, so the answer is at most there
    }
    {
        const int64_t pos = lo + lane;
        const bool less = pos < hi ? ((double)cdf[pos] < uj) : false;
        const int cnt = __popc(__ballot_sync(0xffffffffu, less));
        int64_t r = lo + cnt;
        if (r > n - 1) r = n - 1;  // u >= 1 is outside the contract; keep the index valid
        if (lane == 0) idx[j] = r;
    }
}

// ---------------------------------------------------------------------------------------------
// fused draw: one CTA per draw; finds the chunk from the boundary values, rebuilds only that
// chunk's exact CDF in shared memory, finishes the search there, then gathers the candidate.
constexpr int kDrawNT = 256, kDrawIPT = kChunk / kDrawNT;

struct DrawArgs {
    WeightSrc src; int64_t n; int nchunks;
    const uint32_t* start_bits; const float* total_ptr;
    const double* u; int64_t b;
    const unsigned char* data; int64_t row_bytes; const int64_t* labels; const int64_t* index;
    int64_t* idx; float* nprob; float* seed; unsigned char* out_data; int64_t* out_labels; int64_t* out_index;
};

__global__ void __launch_bounds__(kDrawNT) draw_kernel(DrawArgs a) {
    __shared__ ResolveSmem rs;
    __shared__ float cdf_s[kChunk];
    __shared__ int wcnt[kDrawNT / 32];
    __shared__ int64_t chosen;
    const int64_t j = blockIdx.x;
    const double uj = a.u[j];
    const float total = *a.total_ptr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // (a) chunks whose last cdf value is < u
    int cnt = 0;
    for (int c = threadIdx.x; c < a.nchunks; c += kDrawNT) {
        const float endv = __fdiv_rn(b2f(a.start_bits[c + 1]), total);
        cnt += ((double)endv < uj) ? 1 : 0;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) wcnt[warp] = cnt;
    __syncthreads();
    int chunk = 0;
#pragma unroll
    for (int i = 0; i < kDrawNT / 32; ++i) chunk += wcnt[i];
    if (chunk > a.nchunks - 1) chunk = a.nchunks - 1;
    __syncthreads();
    // (b) exact CDF of that chunk
    const int64_t start = (int64_t)chunk * kChunk;
    const int len = (int)min((int64_t)kChunk, a.n - start);
    uint32_t pb[kDrawIPT];
    load_chunk<kDrawNT, kDrawIPT>(a.src, start, len, pb);
    resolve_chunk<kDrawNT, kDrawIPT>(pb, len, a.start_bits[chunk], cdf_s, total, &rs);
    __syncthreads();
    // (c) slots of the chunk below u
    cnt = 0;
    for (int i = threadIdx.x; i < len; i += kDrawNT) cnt += ((double)cdf_s[i] < uj) ? 1 : 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) wcnt[warp] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int i = 0; i < kDrawNT / 32; ++i) t += wcnt[i];
        if (t > len - 1) t = len - 1;
        const int64_t pick = start + t;
        chosen = pick;
        a.idx[j] = pick;
        float shift, tot2;
        a.src.prepare(shift, tot2);
        const float p = a.src.get(pick, shift, tot2);
        // pytorch_collab.py:116: losses_pros[idx] * N ; :137,:145: grad = (1/B) / probs
        const float np_ = __fmul_rn(p, (float)a.n);
        a.nprob[j] = np_;
        a.seed[j] = __fdiv_rn(__fdiv_rn(1.f, (float)a.b), np_);
        if (a.labels && a.out_labels) a.out_labels[j] = a.labels[pick];
        if (a.index && a.out_index) a.out_index[j] = a.index[pick];
    }
    __syncthreads();
    // (d) payload row
    if (a.data && a.out_data) {
        const unsigned char* srow = a.data + chosen * a.row_bytes;
        unsigned char* drow = a.out_data + j * a.row_bytes;
        if (((reinterpret_cast<uintptr_t>(srow) | reinterpret_cast<uintptr_t>(drow) | (uintptr_t)a.row_bytes) & 15) == 0) {
            const uint4* s4 = reinterpret_cast<const uint4*>(srow);
            uint4* d4 = reinterpret_cast<uint4*>(drow);
            for (int64_t i = threadIdx.x; i < (a.row_bytes >> 4); i += kDrawNT) d4[i] = s4[i];
        } else {
            for (int64_t i = threadIdx.x; i < a.row_bytes; i += kDrawNT) drow[i] = srow[i];
        }
    }
}

// workspace carving (all offsets 256-byte aligned)
struct CdfWs {
    double* sums; double* approx; Summary* summaries; uint32_t* start_bits; float* total; size_t bytes;
};
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
inline CdfWs carve(void* ws, int64_t n) {
    const size_t nchunks = (size_t)((n + kChunk - 1) / kChunk);
    unsigned char* p = static_cast<unsigned char*>(ws);
    CdfWs w;
    size_t off = 0;
    w.sums = reinterpret_cast<double*>(p + off); off += align256(nchunks * 8);
    w.approx = reinterpret_cast<double*>(p + off); off += align256((nchunks + 1) * 8);
    w.summaries = reinterpret_cast<Summary*>(p + off); off += align256(nchunks * sizeof(Summary));
    w.start_bits = reinterpret_cast<uint32_t*>(p + off); off += align256((nchunks + 1) * 4);
    w.total = reinterpret_cast<float*>(p + off); off += 256;
    w.bytes = off;
    return w;
}

}  // namespace

int check_ws(void* ws, size_t have, size_t need, const char* who) {
    if (have < need) { set_last_error("%s: workspace too small (%zu < %zu bytes)", who, have, need); return MERCURY_ERR_WORKSPACE; }
    if (need && (!ws || (reinterpret_cast<uintptr_t>(ws) & 255))) {
        set_last_error("%s: workspace must be non-null and 256-byte aligned", who);
        return MERCURY_ERR_WORKSPACE;
    }
    return MERCURY_OK;
}

namespace {

// exact boundaries of the chain over `src` (weights or scores+stats); approx must be filled
int run_exact_boundaries(const WeightSrc& src, int64_t n, const CdfWs& w, int32_t* status, int check, cudaStream_t st) {
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    summarize_kernel<<<nchunks, kSumNT, 0, st>>>(src, n, w.approx, w.summaries, status, check);
    MB_CHECK_LAUNCH("summarize_kernel");
    spine_kernel<<<1, kSpineNT, 0, st>>>(src, n, nchunks, w.summaries, w.start_bits, w.total, status);
    MB_CHECK_LAUNCH("spine_kernel");
    return MERCURY_OK;
}

}  // namespace

size_t cdf_ws_bytes(int64_t n) { return carve(nullptr, n < 1 ? 1 : n).bytes; }

int pool_stats_launch(const float* scores, int64_t n, const int64_t* h_bounds, int n_bounds, float* ema_state, int ema_first,
                      double ema_alpha, double smooth_alpha, float* stats, void* ws, cudaStream_t st) {
    const CdfWs w = carve(ws, n);
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    Bounds bd;
    if (n_bounds > 0) {
        bd.n = n_bounds;
        for (int i = 0; i < n_bounds; ++i) bd.v[i] = h_bounds[i];
    } else {
        bd.n = 1;
        bd.v[0] = n;
    }
    chunk_sums_kernel<<<nchunks, 256, 0, st>>>(scores, n, w.sums, nullptr, 0);
    MB_CHECK_LAUNCH("chunk_sums_kernel");
    stats_finalize_kernel<<<1, 1024, 0, st>>>(scores, n, w.sums, nchunks, bd, ema_state, ema_first, ema_alpha, smooth_alpha,
                                              stats, w.approx);
    MB_CHECK_LAUNCH("stats_finalize_kernel");
    return MERCURY_OK;
}

int update_draw_launch(const mercury_update_args& a, const float* scores, void* ws, cudaStream_t st) {
    const CdfWs w = carve(ws, a.n);
    const int nchunks = (int)((a.n + kChunk - 1) / kChunk);
    WeightSrc src{scores, a.stats};
    int rc = run_exact_boundaries(src, a.n, w, a.status, 1, st);
    if (rc) return rc;
    if (a.b > 0) {
        DrawArgs d;
        d.src = src; d.n = a.n; d.nchunks = nchunks; d.start_bits = w.start_bits; d.total_ptr = w.total;
        d.u = a.uniforms; d.b = a.b;
        d.data = static_cast<const unsigned char*>(a.data); d.row_bytes = a.row_bytes; d.labels = a.labels; d.index = a.index;
        d.idx = a.idx; d.nprob = a.nprob; d.seed = a.seed; d.out_data = static_cast<unsigned char*>(a.out_data);
        d.out_labels = a.out_labels; d.out_index = a.out_index;
        draw_kernel<<<(unsigned)a.b, kDrawNT, 0, st>>>(d);
        MB_CHECK_LAUNCH("draw_kernel");
    }
    return MERCURY_OK;
}

}  // namespace mercury

using namespace mercury;

extern "C" size_t mercury_pool_stats_workspace_bytes(int64_t n) { return cdf_ws_bytes(n); }

extern "C" int mercury_pool_stats(const float* scores, int64_t n, const int64_t* h_bounds, int n_bounds, float* ema_state,
                                  int ema_first, double ema_alpha, double smooth_alpha, float* stats, void* ws,
                                  size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(n >= 1, "mercury_pool_stats: empty pool");
    MB_CHECK_ARG(n <= ((int64_t)1 << 24), "number of categories cannot exceed 2^24");
    MB_CHECK_ARG(scores && ema_state && stats, "mercury_pool_stats: null pointer");
    MB_CHECK_ARG(n_bounds >= 0 && n_bounds <= kMaxBounds, "mercury_pool_stats: at most %d pool sizes", kMaxBounds);
    for (int i = 0; i < n_bounds; ++i) {
        MB_CHECK_ARG(h_bounds[i] >= 1 && h_bounds[i] <= n && (i == 0 || h_bounds[i] > h_bounds[i - 1]),
                     "mercury_pool_stats: pool sizes must be ascending in [1, n]");
    }
    MB_CHECK_ARG(n_bounds == 0 || h_bounds[n_bounds - 1] == n, "mercury_pool_stats: last pool size must equal n");
    int rc = check_ws(ws, ws_bytes, cdf_ws_bytes(n), "mercury_pool_stats");
    if (rc) return rc;
    return pool_stats_launch(scores, n, h_bounds, n_bounds, ema_state, ema_first, ema_alpha, smooth_alpha, stats, ws,
                             static_cast<cudaStream_t>(stream));
}

extern "C" int mercury_normalize(const float* scores, int64_t n, const float* stats, float* p, void* stream) {
    MB_CHECK_ARG(n >= 0 && (n == 0 || (scores && stats && p)), "mercury_normalize: bad arguments");
    if (n == 0) return MERCURY_OK;
    const int grid = (int)min_i64((n + 255) / 256, (int64_t)kSMs * 8);
    normalize_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(scores, n, stats, p);
    MB_CHECK_LAUNCH("normalize_kernel");
    return MERCURY_OK;
}

extern "C" size_t mercury_cdf_workspace_bytes(int64_t n) { return cdf_ws_bytes(n); }

extern "C" int mercury_cdf(const float* weights, int64_t n, int mode, float* cdf, float* total, int32_t* status, void* ws,
                           size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(n >= 1, "mercury_cdf: empty distribution");
    MB_CHECK_ARG(n <= ((int64_t)1 << 24), "number of categories cannot exceed 2^24");
    MB_CHECK_ARG(weights && cdf && total && status, "mercury_cdf: null pointer");
    MB_CHECK_ARG(mode == MERCURY_CDF_EXACT_SERIAL_F32 || mode == MERCURY_CDF_F64, "mercury_cdf: unknown mode %d", mode);
    int rc = check_ws(ws, ws_bytes, cdf_ws_bytes(n), "mercury_cdf");
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const CdfWs w = carve(ws, n);
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    chunk_sums_kernel<<<nchunks, 256, 0, st>>>(weights, n, w.sums, status, 1);
    MB_CHECK_LAUNCH("chunk_sums_kernel");
    prefix_finalize_kernel<<<1, 1024, 0, st>>>(w.sums, nchunks, w.approx);
    MB_CHECK_LAUNCH("prefix_finalize_kernel");
    if (mode == MERCURY_CDF_F64) {
        materialize_f64_kernel<<<nchunks, 256, 0, st>>>(weights, n, w.approx, nchunks, cdf, total, status);
        MB_CHECK_LAUNCH("materialize_f64_kernel");
        return MERCURY_OK;
    }
    WeightSrc src{weights, nullptr};
    rc = run_exact_boundaries(src, n, w, status, 0, st);
    if (rc) return rc;
    materialize_kernel<<<nchunks, kMatNT, 0, st>>>(src, n, w.start_bits, w.total, cdf);
    MB_CHECK_LAUNCH("materialize_kernel");
    cudaError_t e = cudaMemcpyAsync(total, w.total, sizeof(float), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "mercury_cdf: total copy");
    return MERCURY_OK;
}

extern "C" int mercury_sample_with_replacement(const float* cdf, int64_t n, const double* uniforms, int64_t b, int64_t* idx,
                                               void* stream) {
    MB_CHECK_ARG(n >= 1 && b >= 0, "mercury_sample_with_replacement: need n >= 1, b >= 0");
    MB_CHECK_ARG(b == 0 || (cdf && uniforms && idx), "mercury_sample_with_replacement: null pointer");
    if (b == 0) return MERCURY_OK;
    search_kernel<<<(unsigned)((b + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(cdf, n, uniforms, b, idx);
    MB_CHECK_LAUNCH("search_kernel");
    return MERCURY_OK;
}

// ---------------------------------------------------------------------------------------------
// mercury_cdf_static: the same cumulative distribution, for callers that cannot hand us scratch.
//
// Who needs this: a host that has no device allocator to call on -- a plugin running inside
// another runtime that reserved the whole GPU for its own pool at start-up (a cudaMalloc there
// fails or fragments that pool), or a small C/Go/Java host binding this ABI that would otherwise
// have to size a workspace, keep it alive and free it around every call.  They pass only the
// input and the outputs; the scratch lives in the array below, inside the library image, so
// every GPU that loads the library carries these 112 KiB until the process exits.
//
// The name `g_static_scratch` is only a stand-in on the host side: its device address has to be
// asked for (cudaGetSymbolAddress) with the right device current, once per device.  The first
// call on a device does that, so make that first call outside any stream capture.
//
// Limits and sharing: n <= MERCURY_STATIC_MAX_N (1 << 22); every call on a device uses the same
// array, so two calls whose kernels overlap on the GPU (different streams, different host
// threads) race on it -- keep them on one stream or order them with events.  The array keeps
// whatever the previous call left there; that is harmless because each stage writes a field
// before any later stage reads it, so nothing in it has to start at zero and nothing is cleared.
constexpr int64_t kStaticMaxN = (int64_t)MERCURY_STATIC_MAX_N;
constexpr size_t kStaticScratchBytes = 112 * 1024;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {};

int static_scratch_for_current_device(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "mercury_cdf_static: cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) {
        set_last_error("mercury_cdf_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
        return MERCURY_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);  // device `dev` is current here
        if (e != cudaSuccess) return cuda_fail(e, "mercury_cdf_static: cudaGetSymbolAddress");
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255)) {
            set_last_error("mercury_cdf_static: scratch symbol resolved to a bad address");
            return MERCURY_ERR_CUDA;
        }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return MERCURY_OK;
}
}  // namespace

extern "C" int mercury_cdf_static(const float* weights, int64_t n, int mode, float* cdf, float* total, int32_t* status,
                                  void* stream) {
    MB_CHECK_ARG(n >= 1, "mercury_cdf_static: empty distribution");
    MB_CHECK_ARG(n <= kStaticMaxN,
                 "mercury_cdf_static: n = %lld exceeds the built-in scratch (at most %lld candidates); "
                 "use mercury_cdf with a caller-provided workspace",
                 (long long)n, (long long)kStaticMaxN);
    if (cdf_ws_bytes(n) > kStaticScratchBytes) {  // cannot happen for n <= kStaticMaxN; kept as a guard
        set_last_error("mercury_cdf_static: internal scratch too small for n = %lld", (long long)n);
        return MERCURY_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    const int rc = static_scratch_for_current_device(&scratch);
    if (rc) return rc;
    return mercury_cdf(weights, n, mode, cdf, total, status, scratch, kStaticScratchBytes, stream);
}
