// K2' -- exponential race (Gumbel top-k in its p/Exp(1) form): draws without replacement.
// Replaces torch.multinomial(p, k, replacement=False) and every n_sample == 1 draw of the
// reference's Groupwise_Sampler (util.py:150-151).  ATen's fast path computes
//   q_i = float(-log1p(-u_i)),  keys_i = p_i / q_i,  topk(keys, k)
// with u_i fp64 uniforms; the kernels below compute the same keys and select the k largest by a
// radix select on the key bits (keys are non-negative, so their bit patterns are ordered),
// followed by a bitonic sort of the few survivors.  Ties are broken towards the lower index.
#include <math.h>

#include "common.cuh"
#include "internal.cuh"

namespace mercury {
namespace {

constexpr int kBins = 4096;
constexpr int kMaxK = 2048;
constexpr int kBinCap = 2048;            // candidates kept from the threshold bin
constexpr int kCandCap = kMaxK + kBinCap;

struct SelState {
    uint32_t prefix;   // resolved high bits of the k-th key
    uint32_t mask;     // which bits are resolved
    uint32_t k_rem;    // how many still to take among keys matching the prefix
    uint32_t done;     // threshold bin small enough to collect wholesale
    uint32_t count;    // candidates appended so far
    uint32_t pad[3];
};

struct TopkWs { float* keys; uint32_t* hist; SelState* st; unsigned long long* cand; size_t bytes; };
inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
inline TopkWs carve(void* ws, int64_t n) {
    unsigned char* p = static_cast<unsigned char*>(ws);
    TopkWs w; size_t off = 0;
    w.hist = reinterpret_cast<uint32_t*>(p + off); off += al256(kBins * 4);
    w.st = reinterpret_cast<SelState*>(p + off); off += 256;
    w.cand = reinterpret_cast<unsigned long long*>(p + off); off += al256((size_t)kCandCap * 8);
    w.keys = reinterpret_cast<float*>(p + off); off += al256((size_t)(n < 1 ? 1 : n) * 4);
    w.bytes = off;
    return w;
}

__device__ __forceinline__ int digit_of(uint32_t bits, int pass) {
    return pass == 0 ? (int)(bits >> 20) : pass == 1 ? (int)((bits >> 8) & 0xfffu) : (int)(bits & 0xffu);
}

// keys + first histogram
__global__ void __launch_bounds__(256) race_keys_kernel(const float* __restrict__ w, const double* __restrict__ u, int64_t n,
                                                        float* __restrict__ keys, uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[kBins];
    for (int i = threadIdx.x; i < kBins; i += 256) sh[i] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const float q = (float)(-log1p(-u[i]));
        const float key = __fdiv_rn(w[i], q);
        keys[i] = key;
        atomicAdd(&sh[__float_as_uint(key) >> 20], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kBins; i += 256)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// histogram of the next digit among keys that match the resolved prefix
__global__ void __launch_bounds__(256) hist_kernel(const float* __restrict__ keys, int64_t n, int pass,
                                                   const SelState* __restrict__ st, uint32_t* __restrict__ hist) {
    if (st->done) return;
    __shared__ uint32_t sh[kBins];
    for (int i = threadIdx.x; i < kBins; i += 256) sh[i] = 0;
    __syncthreads();
    const uint32_t prefix = st->prefix, mask = st->mask;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const uint32_t b = __float_as_uint(keys[i]);
        if ((b & mask) == prefix) atomicAdd(&sh[digit_of(b, pass)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kBins; i += 256)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// pick the digit that holds the k_rem-th largest key; clears the histogram for the next pass
__global__ void __launch_bounds__(1024) select_kernel(int pass, uint32_t k, SelState* st, uint32_t* hist) {
    if (st->done) return;
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t found_digit, found_above, found_cnt;
    const int t = threadIdx.x;
    // thread t owns bins 4*(1023-t) .. +3, so thread order = descending key order
    const int b0 = 4 * (1023 - t);
    uint32_t h[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = hist[b0 + i];
    const uint32_t local = h[0] + h[1] + h[2] + h[3];
    uint32_t inc = local;
    const int lane = t & 31, warp = t >> 5;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    uint32_t pre = 0;
    for (int i = 0; i < warp; ++i) pre += wsum[i];
    uint32_t above = pre + inc - local;  // keys in strictly higher bins than this thread's
    const uint32_t k_rem = pass == 0 ? k : st->k_rem;
#pragma unroll
    for (int i = 3; i >= 0; --i) {
        if (above < k_rem && above + h[i] >= k_rem) { found_digit = (uint32_t)(b0 + i); found_above = above; found_cnt = h[i]; }
        above += h[i];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) hist[b0 + i] = 0;
    __syncthreads();
    if (t == 0) {
        const int shift = pass == 0 ? 20 : pass == 1 ? 8 : 0;
        const uint32_t dmask = pass == 0 ? 0xfff00000u : pass == 1 ? 0x000fff00u : 0x000000ffu;
        st->prefix |= found_digit << shift;
        st->mask |= dmask;
        st->k_rem = k_rem - found_above;
        if (found_cnt <= (uint32_t)kBinCap || pass == 2) st->done = 1;
    }
}

// every key above the threshold bin, plus the bin itself, as (key bits, ~index) composites
__global__ void __launch_bounds__(256) collect_kernel(const float* __restrict__ keys, int64_t n, SelState* st,
                                                      unsigned long long* __restrict__ cand, int32_t* status) {
    const uint32_t prefix = st->prefix, mask = st->mask;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const uint32_t b = __float_as_uint(keys[i]);
        if ((b & mask) >= prefix) {
            const uint32_t slot = atomicAdd(&st->count, 1u);
            if (slot < (uint32_t)kCandCap) cand[slot] = ((unsigned long long)b << 32) | (unsigned long long)(0xffffffffu - (uint32_t)i);
            else atomicOr(status, MERCURY_ST_TOPK_TIES);
        }
    }
}

// bitonic sort (descending) of the candidates in shared memory; emits the first k indices
__global__ void __launch_bounds__(1024) sort_kernel(const SelState* __restrict__ st, const unsigned long long* __restrict__ cand,
                                                    int64_t k, int64_t* __restrict__ idx) {
    __shared__ unsigned long long s[kCandCap];
    const uint32_t cnt = min(st->count, (uint32_t)kCandCap);
    for (int i = threadIdx.x; i < kCandCap; i += 1024) s[i] = i < (int)cnt ? cand[i] : 0ull;
    __syncthreads();
    for (int size = 2; size <= kCandCap; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < kCandCap / 2; i += 1024) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const unsigned long long a = s[lo], b = s[hi];
                if ((a < b) == desc) { s[lo] = b; s[hi] = a; }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < k; i += 1024) idx[i] = (int64_t)(0xffffffffu - (uint32_t)(s[i] & 0xffffffffull));
}

// ---------------------------------------------------------------------------------------------------------------
// Group-masked draws (Groupwise_Sampler.__iter__, reference util.py:140-157): the members of the current group are the
// positions whose tag equals `group`; p = (imp + mean(imp)) / sum(...) over the members in float64, rounded once to
// fp32 (util.py:144-150: numpy float64, then torch.Tensor); the m-th member (in position order) races with the m-th
// uniform of the draw, exactly the stream torch.multinomial consumes on the compacted vector.  No compaction happens:
// a count + scan gives every member its rank, non-members never enter the race.
constexpr int kGwChunk = 2048;
struct GwState { double mean, denom; unsigned int count; unsigned int pad; };

__global__ void __launch_bounds__(256) gw_count_kernel(const float* __restrict__ imp, const int32_t* __restrict__ tags, int32_t group,
                                                       int64_t n, unsigned int* __restrict__ cnt, double* __restrict__ sum) {
    __shared__ double rs[8];
    __shared__ unsigned int rc[8];
    const int64_t start = (int64_t)blockIdx.x * kGwChunk;
    unsigned int c = 0;
    double s = 0.0;
    for (int i = threadIdx.x; i < kGwChunk && start + i < n; i += 256)
        if (tags[start + i] == group) { ++c; s += (double)imp[start + i]; }
    s = warp_sum(s);
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) { rs[threadIdx.x >> 5] = s; rc[threadIdx.x >> 5] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        unsigned int k = 0;
        for (int i = 0; i < 8; ++i) { t += rs[i]; k += rc[i]; }
        cnt[blockIdx.x] = k; sum[blockIdx.x] = t;
    }
}

// exclusive prefix of the chunk counts (in place), the group's mean and normaliser
__global__ void __launch_bounds__(1024) gw_scan_kernel(unsigned int* __restrict__ cnt, const double* __restrict__ sum, int nchunks,
                                                       unsigned int expect_count, GwState* st, int32_t* status) {
    __shared__ unsigned int wtot[32];
    __shared__ double dsum[32];
    __shared__ unsigned int carry;
    if (threadIdx.x == 0) carry = 0u;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double acc = 0.0;
    for (int c0 = 0; c0 < nchunks; c0 += 1024) {
        const int c = c0 + threadIdx.x;
        const unsigned int v = c < nchunks ? cnt[c] : 0u;
        if (c < nchunks) acc += sum[c];
        unsigned int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) wtot[warp] = inc;
        __syncthreads();
        unsigned int pre = carry;
        for (int i = 0; i < warp; ++i) pre += wtot[i];
        if (c < nchunks) cnt[c] = pre + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = pre + inc;
        __syncthreads();
    }
    acc = warp_sum(acc);
    if (lane == 0) dsum[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 32; ++i) t += dsum[i];
        const unsigned int count = carry;
        const double mean = count ? t / (double)count : 0.0;
        st->mean = mean; st->denom = t + (double)count * mean; st->count = count;   // sum(imp + mean) over the members
        if (count != expect_count) atomicOr(status, MERCURY_ST_GROUP_COUNT);
        if (!(st->denom > 0.0)) atomicOr(status, MERCURY_ST_ZERO_SUM);
    }
}

// the race.  grid (chunks, draws).  RACE1: every draw is a single draw -> block maximum of (key bits << 32 | ~position)
// into best[draw]; else (one draw, k > 1): keys + first histogram for the radix select above (non-members: key 0).
template <bool RACE1>
__global__ void __launch_bounds__(256) gw_race_kernel(const float* __restrict__ imp, const int32_t* __restrict__ tags, int32_t group,
                                                      int64_t n, const unsigned int* __restrict__ base, const GwState* __restrict__ st,
                                                      const double* __restrict__ u, unsigned long long* __restrict__ best,
                                                      float* __restrict__ keys, uint32_t* __restrict__ hist, float* __restrict__ p_out) {
    __shared__ unsigned int wtot[8];
    __shared__ unsigned long long wbest[8];
    __shared__ uint32_t sh[RACE1 ? 1 : kBins];
    if (!RACE1) { for (int i = threadIdx.x; i < kBins; i += 256) sh[i] = 0; }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t start = (int64_t)blockIdx.x * kGwChunk + (int64_t)threadIdx.x * 8;
    const double mean = st->mean, denom = st->denom;
    const unsigned int count = st->count;
    bool mem[8];
    unsigned int local = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { mem[k] = start + k < n && tags[start + k] == group; local += mem[k]; }
    unsigned int inc = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    unsigned int rank = base[blockIdx.x] + inc - local;
    for (int i = 0; i < warp; ++i) rank += wtot[i];
    const double* ud = u + (size_t)blockIdx.y * count;
    unsigned long long mine = 0ull;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float key = 0.f;
        if (mem[k]) {
            const float p = (float)(((double)imp[start + k] + mean) / denom);
            const float q = (float)(-log1p(-ud[rank]));
            key = __fdiv_rn(p, q);
            ++rank;
            if (p_out && blockIdx.y == 0) p_out[start + k] = p;
            const unsigned long long comp = ((unsigned long long)__float_as_uint(key) << 32) | (0xffffffffu - (uint32_t)(start + k));
            mine = comp > mine ? comp : mine;
        } else if (p_out && blockIdx.y == 0 && start + k < n) p_out[start + k] = 0.f;
        if (!RACE1 && start + k < n) { keys[start + k] = key; atomicAdd(&sh[__float_as_uint(key) >> 20], 1u); }
    }
    if (RACE1) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, mine, o); mine = t > mine ? t : mine; }
        if (lane == 0) wbest[warp] = mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long b = 0ull;
            for (int i = 0; i < 8; ++i) b = wbest[i] > b ? wbest[i] : b;
            if (b) atomicMax(&best[blockIdx.y], b);
        }
    } else {
        __syncthreads();
        for (int i = threadIdx.x; i < kBins; i += 256)
            if (sh[i]) atomicAdd(&hist[i], sh[i]);
    }
}

__global__ void gw_decode_kernel(const unsigned long long* __restrict__ best, int64_t ndraws, int64_t* __restrict__ idx) {
    const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (j < ndraws) idx[j] = (int64_t)(0xffffffffu - (uint32_t)best[j]);
}

struct GwWs { unsigned int* cnt; double* sum; GwState* st; unsigned long long* best; size_t topk_off; size_t bytes; };
inline GwWs gw_carve(void* ws, int64_t n, int64_t ndraws) {
    unsigned char* p = static_cast<unsigned char*>(ws);
    const size_t nch = (size_t)((n + kGwChunk - 1) / kGwChunk);
    GwWs w; size_t off = 0;
    w.cnt = reinterpret_cast<unsigned int*>(p + off); off += al256(nch * 4);
    w.sum = reinterpret_cast<double*>(p + off); off += al256(nch * 8);
    w.st = reinterpret_cast<GwState*>(p + off); off += 256;
    w.best = reinterpret_cast<unsigned long long*>(p + off); off += al256((size_t)(ndraws < 1 ? 1 : ndraws) * 8);
    w.topk_off = off; off += carve(nullptr, n).bytes;
    w.bytes = off;
    return w;
}

}  // namespace
}  // namespace mercury

using namespace mercury;

extern "C" size_t mercury_sample_topk_workspace_bytes(int64_t n, int64_t k) {
    (void)k;
    return carve(nullptr, n).bytes;
}

extern "C" int mercury_sample_topk(const float* weights, int64_t n, const double* uniforms, int64_t k, int64_t* idx,
                                   int32_t* status, void* ws, size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(n >= 1 && k >= 1, "mercury_sample_topk: need n >= 1 and k >= 1");
    MB_CHECK_ARG(k <= n, "cannot sample n_sample > prob_dist.size(-1) samples without replacement");
    MB_CHECK_ARG(k <= kMaxK, "mercury_sample_topk: k (%lld) exceeds %d", (long long)k, kMaxK);
    MB_CHECK_ARG(n <= ((int64_t)1 << 24), "number of categories cannot exceed 2^24");
    MB_CHECK_ARG(weights && uniforms && idx && status, "mercury_sample_topk: null pointer");
    const TopkWs w = carve(ws, n);
    int rc = check_ws(ws, ws_bytes, w.bytes, "mercury_sample_topk");
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // hist + state are contiguous at the head of the workspace
    cudaError_t e = cudaMemsetAsync(w.hist, 0, al256(kBins * 4) + 256, st);
    if (e != cudaSuccess) return cuda_fail(e, "mercury_sample_topk: memset");
    const int grid = (int)min_i64((n + 255) / 256, (int64_t)kSMs * 4);
    race_keys_kernel<<<grid, 256, 0, st>>>(weights, uniforms, n, w.keys, w.hist);
    MB_CHECK_LAUNCH("race_keys_kernel");
    for (int pass = 0; pass < 3; ++pass) {
        if (pass > 0) {
            hist_kernel<<<grid, 256, 0, st>>>(w.keys, n, pass, w.st, w.hist);
            MB_CHECK_LAUNCH("hist_kernel");
        }
        select_kernel<<<1, 1024, 0, st>>>(pass, (uint32_t)k, w.st, w.hist);
        MB_CHECK_LAUNCH("select_kernel");
    }
    collect_kernel<<<grid, 256, 0, st>>>(w.keys, n, w.st, w.cand, status);
    MB_CHECK_LAUNCH("collect_kernel");
    sort_kernel<<<1, 1024, 0, st>>>(w.st, w.cand, k, idx);
    MB_CHECK_LAUNCH("sort_kernel");
    return MERCURY_OK;
}

extern "C" size_t mercury_groupwise_draw_workspace_bytes(int64_t n, int64_t ndraws) { return gw_carve(nullptr, n, ndraws).bytes; }

extern "C" int mercury_groupwise_draw(const float* importance, const int32_t* tags, int32_t group, int64_t n, int64_t count,
                                      const double* uniforms, int64_t ndraws, int64_t k, int64_t* idx, float* p_out,
                                      int32_t* status, void* ws, size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(n >= 1 && n <= ((int64_t)1 << 24), "mercury_groupwise_draw: 1 <= n <= 2^24");
    MB_CHECK_ARG(count >= 1 && count <= n, "mercury_groupwise_draw: the group must have between 1 and n members");
    MB_CHECK_ARG(ndraws >= 1 && k >= 1 && (k == 1 || ndraws == 1), "mercury_groupwise_draw: either single draws (k == 1) or one draw of k");
    MB_CHECK_ARG(k <= count, "cannot sample n_sample > prob_dist.size(-1) samples without replacement");
    MB_CHECK_ARG(k <= kMaxK, "mercury_groupwise_draw: k (%lld) exceeds %d", (long long)k, kMaxK);
    MB_CHECK_ARG(ndraws <= 65535, "mercury_groupwise_draw: at most 65535 draws per call");
    MB_CHECK_ARG(importance && tags && uniforms && idx && status, "mercury_groupwise_draw: null pointer");
    const GwWs w = gw_carve(ws, n, ndraws);
    int rc = check_ws(ws, ws_bytes, w.bytes, "mercury_groupwise_draw");
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nch = (int)((n + kGwChunk - 1) / kGwChunk);
    gw_count_kernel<<<nch, 256, 0, st>>>(importance, tags, group, n, w.cnt, w.sum);
    MB_CHECK_LAUNCH("gw_count_kernel");
    gw_scan_kernel<<<1, 1024, 0, st>>>(w.cnt, w.sum, nch, (unsigned int)count, w.st, status);
    MB_CHECK_LAUNCH("gw_scan_kernel");
    if (k == 1) {
        cudaError_t e = cudaMemsetAsync(w.best, 0, (size_t)ndraws * 8, st);
        if (e != cudaSuccess) return cuda_fail(e, "mercury_groupwise_draw: memset");
        gw_race_kernel<true><<<dim3(nch, (unsigned)ndraws), 256, 0, st>>>(importance, tags, group, n, w.cnt, w.st, uniforms, w.best, nullptr,
                                                                          nullptr, p_out);
        MB_CHECK_LAUNCH("gw_race_kernel");
        gw_decode_kernel<<<(unsigned)((ndraws + 255) / 256), 256, 0, st>>>(w.best, ndraws, idx);
        MB_CHECK_LAUNCH("gw_decode_kernel");
        return MERCURY_OK;
    }
    const TopkWs t = carve(static_cast<unsigned char*>(ws) + w.topk_off, n);
    cudaError_t e = cudaMemsetAsync(t.hist, 0, al256(kBins * 4) + 256, st);
    if (e != cudaSuccess) return cuda_fail(e, "mercury_groupwise_draw: memset");
    gw_race_kernel<false><<<dim3(nch, 1), 256, 0, st>>>(importance, tags, group, n, w.cnt, w.st, uniforms, nullptr, t.keys, t.hist, p_out);
    MB_CHECK_LAUNCH("gw_race_kernel");
    const int grid = (int)min_i64((n + 255) / 256, (int64_t)kSMs * 4);
    for (int pass = 0; pass < 3; ++pass) {
        if (pass > 0) {
            hist_kernel<<<grid, 256, 0, st>>>(t.keys, n, pass, t.st, t.hist);
            MB_CHECK_LAUNCH("hist_kernel");
        }
        select_kernel<<<1, 1024, 0, st>>>(pass, (uint32_t)k, t.st, t.hist);
        MB_CHECK_LAUNCH("select_kernel");
    }
    collect_kernel<<<grid, 256, 0, st>>>(t.keys, n, t.st, t.cand, status);
    MB_CHECK_LAUNCH("collect_kernel");
    sort_kernel<<<1, 1024, 0, st>>>(t.st, t.cand, k, idx);
    MB_CHECK_LAUNCH("sort_kernel");
    return MERCURY_OK;
}
